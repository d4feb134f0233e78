// mg.cu -- the multi-GPU host of the shading path in C: ONE process drives the N GPUs of a box (north_star:
// "screen tiles, cubemap faces/mips and batched camera views are partitioned across the 8 GPUs of one box, and NCCL over
// NVLink is used only to gather the finished tiles and faces").  The reference renders on one device from one host thread
// (src/pbr_engine/core/pbr/core/GpuHandle.cpp:77-81, src/gltf_viewer/App.cpp:444-483) and has no counterpart; a C++ client of
// pbr::PbrRenderSystem drives this through include/pbr/DeviceGroup.hpp.
//
// A group = one pbrcu_ctx + one stream per member device, peer access enabled between every pair, so memory from
// cudaMalloc on one member is addressable from kernels on another.  Work is ENQUEUED on the members' streams:
//  * pbrcu_mg_shade_frame: every member clusters (K3c) and shades (K3) its row band of the frame.  The gather of the
//    finished bands is either fused into K3's stores (every texel also stored into the peers' frames over NVLink:
//    pbrcu_shade_ggx_ibl_replicated) or one grouped NCCL broadcast per band, or left out;
//  * pbrcu_mg_shade_views: contiguous view blocks, no exchange;
//  * pbrcu_mg_prefilter: row blocks of every level from prepared sources, finished texels stored into every chain.
// After the launches an event barrier makes every member's stream wait for all the others', so whatever the caller enqueues
// next on ANY member's stream sees the whole frame / chain.  No kernel waits for another kernel: the ordering is events only.
// The same device may be listed more than once (two members on one GPU): the bands, the replica stores and the barrier
// then run on a single-GPU box exactly as they do across GPUs -- that is how the tests cover N > 1 on one device.
#include <dlfcn.h>

#include <algorithm>
#include <cstdlib>
#include <new>
#include <vector>

#include "common.cuh"

namespace {
constexpr int kMaxMembers = 16;

// the few NCCL entry points the gather arm needs, resolved at run time (libnccl is not a link dependency: the library must load
// on a box without it, and a host that already carries NCCL -- e.g. through PyTorch -- keeps its own copy)
struct Nccl {
  void* lib = nullptr;
  int (*CommInitAll)(void**, int, const int*) = nullptr;
  int (*CommDestroy)(void*) = nullptr;
  int (*GroupStart)() = nullptr;
  int (*GroupEnd)() = nullptr;
  int (*Broadcast)(const void*, void*, size_t, int, int, void*, cudaStream_t) = nullptr;
  const char* (*GetErrorString)(int) = nullptr;
  bool load() {
    if (lib) return true;
    // Which NCCL: (1) the file PBRCU_NCCL_LIB names; (2) the copy the process ALREADY carries (a host with PyTorch loaded has its
    // own, newer libnccl.so.2 -- loading a second one under the same SONAME would make later loads of the host's libraries bind
    // to the wrong version); (3) the system's.  A host that imports its NCCL user later should set PBRCU_NCCL_LIB to that copy.
    if (const char* path = getenv("PBRCU_NCCL_LIB")) lib = dlopen(path, RTLD_NOW | RTLD_LOCAL);
    if (!lib) lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_LOCAL | RTLD_NOLOAD);
    for (const char* name : {"libnccl.so.2", "libnccl.so"}) {
      if (lib) break;
      lib = dlopen(name, RTLD_NOW | RTLD_LOCAL);
    }
    if (!lib) return false;
    CommInitAll = reinterpret_cast<decltype(CommInitAll)>(dlsym(lib, "ncclCommInitAll"));
    CommDestroy = reinterpret_cast<decltype(CommDestroy)>(dlsym(lib, "ncclCommDestroy"));
    GroupStart = reinterpret_cast<decltype(GroupStart)>(dlsym(lib, "ncclGroupStart"));
    GroupEnd = reinterpret_cast<decltype(GroupEnd)>(dlsym(lib, "ncclGroupEnd"));
    Broadcast = reinterpret_cast<decltype(Broadcast)>(dlsym(lib, "ncclBroadcast"));
    GetErrorString = reinterpret_cast<decltype(GetErrorString)>(dlsym(lib, "ncclGetErrorString"));
    return CommInitAll && CommDestroy && GroupStart && GroupEnd && Broadcast && GetErrorString;
  }
};
constexpr int kNcclChar = 0;  // ncclInt8 / ncclChar
}  // namespace

struct pbrcu_mg {
  int n;
  int device[kMaxMembers];
  pbrcu_ctx* ctx[kMaxMembers];
  cudaStream_t stream[kMaxMembers];
  cudaEvent_t done[kMaxMembers];
  pbrcu_light* lights[kMaxMembers];  // per-member copy of the frame's light list (tile lists index it), grown on demand
  uint32_t lights_cap[kMaxMembers];
  bool distinct;                     // no device listed twice (NCCL needs that)
  Nccl nccl;
  void* comm[kMaxMembers];
  bool comms_ready;
};

using namespace pbrcu;

#define PBRCU_MG_ON(mg, i) PBRCU_CUDA(cudaSetDevice((mg)->device[i]))

extern "C" void pbrcu_mg_destroy(pbrcu_mg* mg) {
  if (!mg) return;
  for (int i = 0; i < mg->n; ++i) {
    cudaSetDevice(mg->device[i]);
    if (mg->comms_ready && mg->comm[i]) mg->nccl.CommDestroy(mg->comm[i]);
    if (mg->stream[i]) { cudaStreamSynchronize(mg->stream[i]); cudaStreamDestroy(mg->stream[i]); }
    if (mg->done[i]) cudaEventDestroy(mg->done[i]);
    if (mg->lights[i]) cudaFree(mg->lights[i]);
    if (mg->ctx[i]) pbrcu_ctx_destroy(mg->ctx[i]);
  }
  delete mg;
}

extern "C" int pbrcu_mg_create(const int* devices, int n, pbrcu_mg** out) {
  PBRCU_REQUIRE(devices && out, "NULL argument");
  PBRCU_REQUIRE(n >= 1 && n <= kMaxMembers, "a device group has 1 to 16 members");
  *out = nullptr;
  pbrcu_mg* mg = new (std::nothrow) pbrcu_mg();
  if (!mg) return PBRCU_ERR_NOMEM;
  *mg = pbrcu_mg{};
  mg->n = n;
  mg->distinct = true;
  for (int i = 0; i < n; ++i) {
    mg->device[i] = devices[i];
    for (int j = 0; j < i; ++j) mg->distinct = mg->distinct && devices[j] != devices[i];
  }
  int rc = PBRCU_OK;
  for (int i = 0; i < n && rc == PBRCU_OK; ++i) {
    rc = pbrcu_ctx_create(devices[i], &mg->ctx[i]);  // (makes the device current)
    if (rc) break;
    cudaError_t e = cudaStreamCreateWithFlags(&mg->stream[i], cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&mg->done[i], cudaEventDisableTiming);
    if (e != cudaSuccess) { rc = cuda_fail(e, "stream / event of a group member"); break; }
    for (int j = 0; j < n; ++j) {  // every other member's memory becomes addressable from this device
      if (devices[j] == devices[i]) continue;
      int can = 0;
      e = cudaDeviceCanAccessPeer(&can, devices[i], devices[j]);
      if (e != cudaSuccess || !can) {
        set_error("device %d cannot address the memory of device %d: a device group needs peer access (NVLink / NVSwitch)", devices[i], devices[j]);
        rc = PBRCU_ERR_CUDA;
        break;
      }
      e = cudaDeviceEnablePeerAccess(devices[j], 0);
      if (e == cudaErrorPeerAccessAlreadyEnabled) { cudaGetLastError(); e = cudaSuccess; }
      if (e != cudaSuccess) { rc = cuda_fail(e, "cudaDeviceEnablePeerAccess"); break; }
    }
  }
  if (rc) { pbrcu_mg_destroy(mg); return rc; }
  *out = mg;
  return PBRCU_OK;
}

extern "C" int pbrcu_mg_size(const pbrcu_mg* mg) { return mg ? mg->n : 0; }
extern "C" pbrcu_ctx* pbrcu_mg_ctx(pbrcu_mg* mg, int member) { return mg && member >= 0 && member < mg->n ? mg->ctx[member] : nullptr; }
extern "C" pbrcu_stream pbrcu_mg_stream(pbrcu_mg* mg, int member) { return mg && member >= 0 && member < mg->n ? mg->stream[member] : nullptr; }
extern "C" int pbrcu_mg_device(const pbrcu_mg* mg, int member) { return mg && member >= 0 && member < mg->n ? mg->device[member] : -1; }

extern "C" int pbrcu_mg_set_device(pbrcu_mg* mg, int member) {
  PBRCU_REQUIRE(mg && member >= 0 && member < mg->n, "member out of range");
  PBRCU_MG_ON(mg, member);
  return PBRCU_OK;
}

extern "C" int pbrcu_mg_synchronize(pbrcu_mg* mg) {
  PBRCU_REQUIRE(mg, "mg is NULL");
  for (int i = 0; i < mg->n; ++i) {
    PBRCU_MG_ON(mg, i);
    PBRCU_CUDA(cudaStreamSynchronize(mg->stream[i]));
  }
  return PBRCU_OK;
}

// every member's stream waits for the work enqueued so far on all the others' (events only)
static int mg_barrier(pbrcu_mg* mg) {
  for (int i = 0; i < mg->n; ++i) {
    PBRCU_MG_ON(mg, i);
    PBRCU_CUDA(cudaEventRecord(mg->done[i], mg->stream[i]));
  }
  for (int i = 0; i < mg->n; ++i) {
    PBRCU_MG_ON(mg, i);
    for (int j = 0; j < mg->n; ++j)
      if (j != i) PBRCU_CUDA(cudaStreamWaitEvent(mg->stream[i], mg->done[j], 0));
  }
  return PBRCU_OK;
}
extern "C" int pbrcu_mg_barrier(pbrcu_mg* mg) {
  PBRCU_REQUIRE(mg, "mg is NULL");
  return mg_barrier(mg);
}

// Row bands [y0, y1) for n members, cut on light-tile boundaries (PBRCU_TILE rows) so that per-band light clustering equals
// whole-frame clustering.  tile_row_costs (one per tile row, e.g. the previous frame's sum over the row's tiles of 9 + lights in
// the tile) balances the bands by COST; NULL cuts equal numbers of tile rows.  Contiguous, cover [0, height), may be empty.
extern "C" int pbrcu_mg_row_bands(uint32_t height, int n, const float* tile_row_costs, uint32_t* bands_y0y1) {
  PBRCU_REQUIRE(bands_y0y1 && n >= 1 && n <= kMaxMembers && height > 0, "bad argument");
  const uint32_t rows = (height + PBRCU_TILE - 1) / PBRCU_TILE;
  std::vector<uint32_t> cuts(size_t(n) + 1, 0u);
  double total = 0.0;
  if (tile_row_costs)
    for (uint32_t r = 0; r < rows; ++r) total += std::max(double(tile_row_costs[r]), 0.0);
  if (total > 0.0) {
    double acc = 0.0;
    uint32_t r = 0;
    for (int k = 1; k < n; ++k) {
      const double target = total * k / n;
      while (r < rows && acc + std::max(double(tile_row_costs[r]), 0.0) * 0.5 <= target) {  // a row goes to the side of the cut most of it lies on
        acc += std::max(double(tile_row_costs[r]), 0.0);
        ++r;
      }
      cuts[k] = r;
    }
  } else {
    for (int k = 1; k < n; ++k) cuts[k] = uint32_t(uint64_t(rows) * k / n);
  }
  cuts[n] = rows;
  for (int k = 0; k < n; ++k) {
    bands_y0y1[2 * k] = std::min(cuts[k] * PBRCU_TILE, height);
    bands_y0y1[2 * k + 1] = std::min(cuts[k + 1] * PBRCU_TILE, height);
  }
  return PBRCU_OK;
}

// rows [y0, y1) of a G-buffer / of its tile lists as sub-views: pointer arithmetic only (y0 on a tile boundary)
static pbrcu_gbuffer band_of(const pbrcu_gbuffer& g, uint32_t y0, uint32_t y1) {
  pbrcu_gbuffer b = g;
  b.height = y1 - y0;
  auto adv = [&](void* p, size_t pitch) -> void* { return p ? static_cast<uint8_t*>(p) + size_t(y0) * pitch : nullptr; };
  b.positions = adv(g.positions, g.pitch8); b.normals = adv(g.normals, g.pitch8); b.albedo = adv(g.albedo, g.pitch8);
  b.emissive = adv(g.emissive, g.pitch8); b.orm = adv(g.orm, g.pitch4);
  b.depth = static_cast<float*>(adv(g.depth, g.pitch4));
  return b;
}
static pbrcu_light_tiles band_of(const pbrcu_light_tiles& t, uint32_t y0, uint32_t y1) {
  pbrcu_light_tiles b = t;
  const uint32_t ty0 = y0 / PBRCU_TILE;
  b.tiles_y = (y1 - y0 + PBRCU_TILE - 1) / PBRCU_TILE;
  b.counts = t.counts + size_t(ty0) * t.tiles_x;
  b.indices = t.indices + size_t(ty0) * t.tiles_x * t.max_per_tile;
  return b;
}

static int ensure_comms(pbrcu_mg* mg) {
  if (mg->comms_ready) return PBRCU_OK;
  if (!mg->distinct) { set_error("the NCCL gather needs every member on its own device"); return PBRCU_ERR_INVALID; }
  if (!mg->nccl.load()) { set_error("NCCL is not available on this host (dlopen of libnccl.so.2 failed): use the fused gather"); return PBRCU_ERR_CUDA; }
  const int r = mg->nccl.CommInitAll(mg->comm, mg->n, mg->device);
  if (r != 0) { set_error("ncclCommInitAll failed: %s", mg->nccl.GetErrorString(r)); return PBRCU_ERR_CUDA; }
  mg->comms_ready = true;
  return PBRCU_OK;
}

// One frame over the group.  g[i], tiles[i], ibl[i], frames[i]: member i's full-frame G-buffer (only the rows of ITS band need to
// hold data), its tile lists (tiles == NULL: every pixel loops the whole list), its IBL set and its frame of `pitch` bytes per
// row, all in member i's memory.  bands = 2 n rows from pbrcu_mg_row_bands.  lights: HOST array.
// gather: PBRCU_MG_GATHER_NONE (member i's frame holds its band), _P2P (K3 stores every texel into every member's frame),
// _NCCL (one grouped ncclBroadcast per band after the shading).  Returns after enqueueing; ends with pbrcu_mg_barrier.
static int mg_shade_frame_impl(pbrcu_mg* mg, const pbrcu_gbuffer* g, const float cam_pos[3], const pbrcu_light* lights_host, uint32_t n_lights,
                               const pbrcu_light_tiles* tiles, const pbrcu_ibl* ibl, const uint32_t* bands_y0y1, void* const* frames, size_t pitch,
                               int gather, bool ldr) {
  PBRCU_REQUIRE(mg && g && cam_pos && bands_y0y1 && frames, "NULL argument");
  PBRCU_REQUIRE(n_lights == 0 || lights_host, "lights_host is NULL");
  PBRCU_REQUIRE(gather == PBRCU_MG_GATHER_NONE || gather == PBRCU_MG_GATHER_P2P || gather == PBRCU_MG_GATHER_NCCL, "unknown gather mode");
  PBRCU_REQUIRE(gather != PBRCU_MG_GATHER_P2P || mg->n - 1 <= PBRCU_MAX_REPLICAS, "too many members for the fused gather");
  for (int i = 0; i < mg->n; ++i) {
    PBRCU_REQUIRE(frames[i], "frame is NULL");
    PBRCU_REQUIRE(bands_y0y1[2 * i] <= bands_y0y1[2 * i + 1] && bands_y0y1[2 * i + 1] <= g[i].height && bands_y0y1[2 * i] % PBRCU_TILE == 0, "bad band");
  }
  if (gather == PBRCU_MG_GATHER_NCCL)
    if (int rc = ensure_comms(mg)) return rc;
  for (int i = 0; i < mg->n; ++i) {
    const uint32_t y0 = bands_y0y1[2 * i], y1 = bands_y0y1[2 * i + 1];
    if (y1 == y0) continue;
    PBRCU_MG_ON(mg, i);
    const pbrcu_gbuffer gb = band_of(g[i], y0, y1);
    uint8_t* out = static_cast<uint8_t*>(frames[i]) + size_t(y0) * pitch;
    void* reps[PBRCU_MAX_REPLICAS];
    uint32_t nrep = 0;
    if (gather == PBRCU_MG_GATHER_P2P)
      for (int j = 0; j < mg->n; ++j)
        if (j != i && frames[j] != frames[i]) reps[nrep++] = static_cast<uint8_t*>(frames[j]) + size_t(y0) * pitch;
    int rc;
    if (tiles) {
      if (mg->lights_cap[i] < n_lights) {
        if (mg->lights[i]) PBRCU_CUDA(cudaFree(mg->lights[i]));
        mg->lights[i] = nullptr; mg->lights_cap[i] = 0;
        PBRCU_CUDA(cudaMalloc(&mg->lights[i], size_t(n_lights) * sizeof(pbrcu_light)));
        mg->lights_cap[i] = n_lights;
      }
      if (n_lights) PBRCU_CUDA(cudaMemcpyAsync(mg->lights[i], lights_host, size_t(n_lights) * sizeof(pbrcu_light), cudaMemcpyHostToDevice, mg->stream[i]));
      const pbrcu_light_tiles tb = band_of(tiles[i], y0, y1);
      if ((rc = pbrcu_cluster_lights(mg->ctx[i], &gb, mg->lights[i], n_lights, &tb, mg->stream[i]))) return rc;
      rc = ldr    ? pbrcu_shade_ggx_ibl_ldr(mg->ctx[i], &gb, cam_pos, mg->lights[i], n_lights, &tb, ibl ? &ibl[i] : nullptr, pbrcu_rect{0, 0, 0, 0}, out, pitch,
                                            nrep ? reps : nullptr, nrep, mg->stream[i])
           : nrep ? pbrcu_shade_ggx_ibl_replicated(mg->ctx[i], &gb, cam_pos, mg->lights[i], n_lights, &tb, ibl ? &ibl[i] : nullptr, pbrcu_rect{0, 0, 0, 0}, out, pitch,
                                                   reps, nrep, mg->stream[i])
                  : pbrcu_shade_ggx_ibl(mg->ctx[i], &gb, cam_pos, mg->lights[i], n_lights, &tb, ibl ? &ibl[i] : nullptr, pbrcu_rect{0, 0, 0, 0}, out, pitch, mg->stream[i]);
    } else {
      rc = ldr    ? pbrcu_shade_ggx_ibl_ldr(mg->ctx[i], &gb, cam_pos, lights_host, n_lights, nullptr, ibl ? &ibl[i] : nullptr, pbrcu_rect{0, 0, 0, 0}, out, pitch,
                                            nrep ? reps : nullptr, nrep, mg->stream[i])
           : nrep ? pbrcu_shade_ggx_ibl_replicated(mg->ctx[i], &gb, cam_pos, lights_host, n_lights, nullptr, ibl ? &ibl[i] : nullptr, pbrcu_rect{0, 0, 0, 0}, out, pitch, reps,
                                                   nrep, mg->stream[i])
                  : pbrcu_shade_ggx_ibl(mg->ctx[i], &gb, cam_pos, lights_host, n_lights, nullptr, ibl ? &ibl[i] : nullptr, pbrcu_rect{0, 0, 0, 0}, out, pitch, mg->stream[i]);
    }
    if (rc) return rc;
  }
  if (gather == PBRCU_MG_GATHER_NCCL && mg->n > 1) {
    // the exchange north_star names: the finished bands travel over NVLink by NCCL, all broadcasts of the frame in ONE group
    int r = mg->nccl.GroupStart();
    for (int root = 0; root < mg->n && r == 0; ++root) {
      const uint32_t y0 = bands_y0y1[2 * root], y1 = bands_y0y1[2 * root + 1];
      if (y1 == y0) continue;
      for (int i = 0; i < mg->n && r == 0; ++i) {
        uint8_t* p = static_cast<uint8_t*>(frames[i]) + size_t(y0) * pitch;
        r = mg->nccl.Broadcast(p, p, size_t(y1 - y0) * pitch, kNcclChar, root, mg->comm[i], mg->stream[i]);
      }
    }
    const int r2 = mg->nccl.GroupEnd();
    if (r != 0 || r2 != 0) { set_error("NCCL gather failed: %s", mg->nccl.GetErrorString(r ? r : r2)); return PBRCU_ERR_CUDA; }
  }
  return mg_barrier(mg);
}

extern "C" int pbrcu_mg_shade_frame(pbrcu_mg* mg, const pbrcu_gbuffer* g, const float cam_pos[3], const pbrcu_light* lights_host, uint32_t n_lights,
                                    const pbrcu_light_tiles* tiles, const pbrcu_ibl* ibl, const uint32_t* bands_y0y1, void* const* frames, size_t pitch,
                                    int gather) {
  return mg_shade_frame_impl(mg, g, cam_pos, lights_host, n_lights, tiles, ibl, bands_y0y1, frames, pitch, gather, false);
}
// The same frame delivered display-ready: RGBA8 frames (pitch in bytes of 4-byte texels), the reference's ACES + gamma tonemap fused into K3's stores
// (pbrcu_shade_ggx_ibl_ldr) -- half the bytes of the HDR gather on NVLink, no separate tonemap pass.
extern "C" int pbrcu_mg_shade_frame_ldr(pbrcu_mg* mg, const pbrcu_gbuffer* g, const float cam_pos[3], const pbrcu_light* lights_host, uint32_t n_lights,
                                        const pbrcu_light_tiles* tiles, const pbrcu_ibl* ibl, const uint32_t* bands_y0y1, void* const* frames, size_t pitch,
                                        int gather) {
  return mg_shade_frame_impl(mg, g, cam_pos, lights_host, n_lights, tiles, ibl, bands_y0y1, frames, pitch, gather, true);
}

// Batched multi-view over the group (BASELINE config 5): member i shades ITS block of views -- n_views[i] G-buffers of one shape laid
// out view_stride8 / view_stride4 bytes apart from g0[i]'s planes, cameras cam_pos_xyz[i] (host, 3 n_views[i] floats) -- into
// outs[i], view v at out_view_stride bytes after view v - 1.  Views are independent: no exchange.
extern "C" int pbrcu_mg_shade_views(pbrcu_mg* mg, const pbrcu_gbuffer* g0, size_t view_stride8, size_t view_stride4, const uint32_t* n_views,
                                    const float* const* cam_pos_xyz, const pbrcu_light* lights_host, uint32_t n_lights, const pbrcu_ibl* ibl,
                                    void* const* outs, size_t out_pitch, size_t out_view_stride) {
  PBRCU_REQUIRE(mg && g0 && n_views && cam_pos_xyz && outs, "NULL argument");
  for (int i = 0; i < mg->n; ++i) {
    if (n_views[i] == 0) continue;
    PBRCU_MG_ON(mg, i);
    if (int rc = pbrcu_shade_ggx_ibl_batched(mg->ctx[i], &g0[i], view_stride8, view_stride4, n_views[i], cam_pos_xyz[i], lights_host, n_lights, ibl ? &ibl[i] : nullptr,
                                             outs[i], out_pitch, out_view_stride, mg->stream[i]))
      return rc;
  }
  return mg_barrier(mg);
}

// The geometry pass of a block of views per member (BASELINE configs[4]: "256 camera views ... of the glTF scene, views partitioned across the
// GPUs"): member i rasterises n_views[i] cameras (views16[i], projs16[i]: 16 floats each, host) over ITS copy of the scene -- draws[i], whose
// vertex / index / texture pointers live in member i's memory -- into the G-buffers that start at g0[i] (pbrcu_geometry_pass_batched per member).
// Follow it with pbrcu_mg_shade_views on the same blocks.  No exchange.
extern "C" int pbrcu_mg_geometry_views(pbrcu_mg* mg, const pbrcu_gbuffer* g0, size_t view_stride8, size_t view_stride4, const uint32_t* n_views,
                                       const float* const* views16, const float* const* projs16, const pbrcu_draw* const* draws, const uint32_t* n_draws) {
  PBRCU_REQUIRE(mg && g0 && n_views && views16 && projs16 && draws && n_draws, "NULL argument");
  for (int i = 0; i < mg->n; ++i) {
    if (n_views[i] == 0) continue;
    PBRCU_MG_ON(mg, i);
    if (int rc = pbrcu_geometry_pass_batched(mg->ctx[i], &g0[i], view_stride8, view_stride4, n_views[i], views16[i], projs16[i], draws[i], n_draws[i], mg->stream[i]))
      return rc;
  }
  return mg_barrier(mg);
}

// The GGX prefilter over the group: member i convolves row block i of n of every level 1.. from ITS prepared source (every member
// holds the environment: K4 / K5 are replicated) and copies level 0 whole.  gather != 0 stores every finished texel into every
// member's chain (pbrcu_ibl_prefilter_ggx_prepared's replicas), so all chains are complete after the barrier.
extern "C" int pbrcu_mg_prefilter(pbrcu_mg* mg, const pbrcu_ibl_source* const* sources, uint32_t nmips, uint32_t nsamples, void* const* dst_chains, int gather) {
  PBRCU_REQUIRE(mg && sources && dst_chains, "NULL argument");
  PBRCU_REQUIRE(!gather || mg->n - 1 <= PBRCU_MAX_REPLICAS, "too many members for the fused gather");
  for (int i = 0; i < mg->n; ++i) {
    PBRCU_MG_ON(mg, i);
    void* reps[PBRCU_MAX_REPLICAS];
    uint32_t nrep = 0;
    if (gather)
      for (int j = 0; j < mg->n; ++j)
        if (j != i && dst_chains[j] != dst_chains[i]) reps[nrep++] = dst_chains[j];
    int rc = pbrcu_ibl_prefilter_ggx_prepared(mg->ctx[i], sources[i], nsamples, dst_chains[i], 0, 0, 0, 1, nullptr, 0, mg->stream[i]);
    if (!rc && nmips > 1)
      rc = pbrcu_ibl_prefilter_ggx_prepared(mg->ctx[i], sources[i], nsamples, dst_chains[i], 1, nmips - 1, uint32_t(i), uint32_t(mg->n), nrep ? reps : nullptr, nrep,
                                            mg->stream[i]);
    if (rc) return rc;
  }
  return mg_barrier(mg);
}
