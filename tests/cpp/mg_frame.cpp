// mg_frame -- one frame over a pbr::DeviceGroup (include/pbr/DeviceGroup.hpp over pbrcu_mg_*) from C++, the way a multi-GPU client of
// the facade drives it: per member a G-buffer (synthetic, generated on the device from one seed so every member holds the same frame),
// tile lists and a frame; cost-balanced row bands from a first clustering; K3c + K3 per member with the gather fused into the stores.
// Every member's frame must equal, bit for bit, the frame one device shades alone.  Groups tried: {0, 0} (two members on one GPU: runs
// on a one-GPU box) and, with two or more GPUs, all of them (also the NCCL gather).  Prints "mg_frame ok".
#include "pbr/DeviceGroup.hpp"

#include <cstdio>
#include <cstring>
#include <vector>

namespace {
constexpr std::uint32_t W = 320, H = 208, NL = 48;

auto lightsOf(std::uint32_t n) -> std::vector<pbr::PointLight> {
  std::vector<pbr::PointLight> lights(n);
  std::uint32_t state = 12345u;
  auto rnd = [&] { state = state * 1664525u + 1013904223u; return float(state >> 8) / 16777216.0f; };
  for (auto& l : lights) {
    l.position = {rnd() * 16.0f - 8.0f, rnd() * 16.0f - 8.0f, rnd() * 16.0f - 8.0f};
    l.range = 2.0f + rnd() * 4.0f;
    l.color = {0.2f + 0.8f * rnd(), 0.2f + 0.8f * rnd(), 0.2f + 0.8f * rnd()};
    l.intensity = 50.0f + 150.0f * rnd();
  }
  return lights;
}

auto download(void* dev, std::size_t bytes) -> std::vector<unsigned char> {
  std::vector<unsigned char> host(bytes);
  pbr::core::check(pbrcu_memcpy_d2h(host.data(), dev, bytes, nullptr), "download");
  pbr::core::check(pbrcu_stream_synchronize(nullptr), "sync");
  return host;
}

auto runGroup(std::vector<int> const& devices, std::vector<unsigned char> const& want, std::vector<float> const& costs, bool nccl) -> void {
  pbr::DeviceGroup group {std::span<int const>(devices)};
  std::size_t const n = group.size();
  auto const lights = lightsOf(NL);
  float const cam[3] = {0.0f, -1.0f, -12.0f}, target[3] = {0.0f, 0.0f, 0.0f};
  std::vector<pbrcu_gbuffer> g(n);
  std::vector<pbrcu_light_tiles> tiles(n);
  std::vector<void*> frames(n);
  std::size_t const pitch = std::size_t(W) * 8;
  for (std::size_t i = 0; i < n; ++i) {
    group.makeCurrent(i);
    auto* ctx = group.getContext(i);
    pbr::core::check(pbrcu_gbuffer_alloc(ctx, W, H, 1, &g[i]), "gbuffer");
    pbr::core::check(pbrcu_synth_gbuffer(ctx, &g[i], 1, 0x5EED0007ull, cam, target, nullptr), "synth");
    pbr::core::check(pbrcu_light_tiles_alloc(ctx, W, H, NL, &tiles[i]), "tiles");
    pbr::core::check(pbrcu_malloc(ctx, pitch * H, &frames[i]), "frame");
    pbr::core::check(pbrcu_stream_synchronize(nullptr), "sync");
  }
  if (g[0].pitch8 != pitch) throw std::runtime_error("unexpected G-buffer pitch");
  auto const bands = group.rowBands(H, costs);
  if (bands.front().y0 != 0 || bands.back().y1 != H) throw std::runtime_error("bands do not cover the frame");
  for (auto gather : {pbr::Gather::P2P, pbr::Gather::Nccl}) {
    if (gather == pbr::Gather::Nccl && !nccl) continue;
    for (std::size_t i = 0; i < n; ++i) {
      group.makeCurrent(i);
      pbr::core::check(pbrcu_memset(frames[i], 0x5A, pitch * H, nullptr), "memset");
      pbr::core::check(pbrcu_stream_synchronize(nullptr), "sync");
    }
    group.shadeFrame(g, {cam[0], cam[1], cam[2]}, lights, tiles, {}, bands, frames, pitch, gather);
    group.synchronize();
    for (std::size_t i = 0; i < n; ++i) {
      group.makeCurrent(i);
      if (download(frames[i], pitch * H) != want)
        throw std::runtime_error("member " + std::to_string(i) + " of a " + std::to_string(n) + "-member group: the gathered frame differs from the single-device frame");
    }
  }
  for (std::size_t i = 0; i < n; ++i) {
    group.makeCurrent(i);
    auto* ctx = group.getContext(i);
    pbrcu_light_tiles_free(ctx, &tiles[i]);
    pbrcu_gbuffer_free(ctx, &g[i]);
    pbrcu_free(ctx, frames[i]);
  }
}
}  // namespace

int main() try {
  int count = 0;
  pbr::core::check(pbrcu_device_count(&count), "device count");
  if (count < 1) { std::fprintf(stderr, "mg_frame: no CUDA device\n"); return 1; }
  // the frame on ONE device, through the plain entry points; its tile-row costs are "the previous frame's light lists"
  std::vector<unsigned char> want;
  std::vector<float> costs((H + PBRCU_TILE - 1) / PBRCU_TILE, 0.0f);
  {
    auto gpu = pbr::core::makeGpuHandle({});
    auto* ctx = gpu->getContext();
    float const cam[3] = {0.0f, -1.0f, -12.0f}, target[3] = {0.0f, 0.0f, 0.0f};
    pbrcu_gbuffer g {};
    pbrcu_light_tiles tiles {};
    pbr::core::check(pbrcu_gbuffer_alloc(ctx, W, H, 1, &g), "gbuffer");
    pbr::core::check(pbrcu_synth_gbuffer(ctx, &g, 1, 0x5EED0007ull, cam, target, nullptr), "synth");
    pbr::core::check(pbrcu_light_tiles_alloc(ctx, W, H, NL, &tiles), "tiles");
    auto const lights = lightsOf(NL);
    void *lightsDev = nullptr, *frame = nullptr;
    pbr::core::check(pbrcu_malloc(ctx, lights.size() * sizeof(pbrcu_light), &lightsDev), "lights");
    pbr::core::check(pbrcu_memcpy_h2d(lightsDev, lights.data(), lights.size() * sizeof(pbrcu_light), nullptr), "lights upload");
    pbr::core::check(pbrcu_malloc(ctx, std::size_t(W) * H * 8, &frame), "frame");
    pbr::core::check(pbrcu_cluster_lights(ctx, &g, static_cast<pbrcu_light const*>(lightsDev), NL, &tiles, nullptr), "cluster");
    pbr::core::check(pbrcu_shade_ggx_ibl(ctx, &g, cam, static_cast<pbrcu_light const*>(lightsDev), NL, &tiles, nullptr, pbrcu_rect {0, 0, 0, 0}, frame, std::size_t(W) * 8, nullptr),
                     "shade");
    want = download(frame, std::size_t(W) * H * 8);
    std::vector<std::uint32_t> counts(std::size_t(tiles.tiles_x) * tiles.tiles_y);
    pbr::core::check(pbrcu_memcpy_d2h(counts.data(), tiles.counts, counts.size() * 4, nullptr), "counts");
    pbr::core::check(pbrcu_stream_synchronize(nullptr), "sync");
    for (std::uint32_t r = 0; r < tiles.tiles_y; ++r)
      for (std::uint32_t c = 0; c < tiles.tiles_x; ++c) costs[r] += 9.0f + float(counts[std::size_t(r) * tiles.tiles_x + c]);
    pbrcu_light_tiles_free(ctx, &tiles);
    pbrcu_gbuffer_free(ctx, &g);
    pbrcu_free(ctx, lightsDev);
    pbrcu_free(ctx, frame);
  }
  bool lit = false;
  for (auto b : want) lit = lit || b != 0;
  if (!lit) { std::fprintf(stderr, "mg_frame: the single-device frame is empty\n"); return 1; }
  runGroup({0, 0}, want, costs, false);
  runGroup({0, 0, 0}, want, {}, false);  // equal tile rows
  if (count >= 2) {
    std::vector<int> all;
    for (int d = 0; d < count && d < 8; ++d) all.push_back(d);
    runGroup(all, want, costs, true);
  }
  std::printf("mg_frame ok (%d device%s)\n", count, count == 1 ? "" : "s");
  return 0;
} catch (std::exception const& e) {
  std::fprintf(stderr, "mg_frame: %s\n", e.what());
  return 1;
}
