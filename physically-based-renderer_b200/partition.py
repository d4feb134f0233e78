"""Host-side partitioning of the hot path across the GPUs of one box.

Every unit of the path is independent (pixels, views, cube texels), so ranks own disjoint, contiguous
blocks and there is NO data-path collective.  When a caller wants the whole image (or prefiltered chain) on
every rank there are two ways: NCCL through torch.distributed (`gather_rows`, `gather_chain`), or -- faster,
no exchange pass -- the kernels' own stores into the peers' buffers (`PeerFrames` + the C ABI's
`pbrcu_shade_ggx_ibl_replicated` / `_multicast` / `pbrcu_ibl_prefilter_ggx_replicated`)."""
from __future__ import annotations


def split_range(n: int, parts: int, index: int) -> tuple[int, int]:
    """[begin, end) of block `index` when `n` units are cut into `parts` near-equal contiguous blocks
    (the same rule the C ABI uses for prefilter shards: begin = n*index // parts)."""
    if parts < 1 or not (0 <= index < parts):
        raise ValueError(f"bad partition {index}/{parts}")
    return n * index // parts, n * (index + 1) // parts


def row_band(height: int, world: int, rank: int, align: int = 1) -> tuple[int, int]:
    """Rows [y0, y1) of a frame owned by `rank` (single-frame configs 2 and 4).  With align=16 the bands
    start on light-tile boundaries so per-band light clustering equals whole-frame clustering."""
    units = (height + align - 1) // align
    b, e = split_range(units, world, rank)
    return min(b * align, height), min(e * align, height)


def view_block(n_views: int, world: int, rank: int) -> tuple[int, int]:
    """Views [v0, v1) owned by `rank` (batched multi-view, config 5)."""
    return split_range(n_views, world, rank)


def gather_rows(tensor, y0: int, y1: int, world: int, height: int, dist=None):
    """All-gathers row bands of a (H, ...) tensor so every rank holds the full frame.  `tensor` is the
    full-size buffer with only rows [y0, y1) valid on this rank.  Uses all_gather when the bands are
    equal-sized, broadcast per band otherwise.  Works with any torch.distributed backend (NCCL on the
    GPUs, gloo in the CPU tests)."""
    if dist is None or world == 1:
        return tensor
    bands = [row_band(height, world, r) for r in range(world)]
    sizes = {e - b for b, e in bands}
    if len(sizes) == 1:
        dist.all_gather_into_tensor(tensor, tensor[y0:y1].contiguous()) if hasattr(dist, "all_gather_into_tensor") and tensor.is_cuda \
            else _gather_list(tensor, bands, y0, y1, dist)
    else:
        for r, (b, e) in enumerate(bands):
            if e > b:
                dist.broadcast(tensor[b:e], src=r)
    return tensor


def _gather_list(tensor, bands, y0, y1, dist):
    parts = [tensor[b:e].clone() for b, e in bands]
    dist.all_gather(parts, tensor[y0:y1].contiguous())
    for (b, e), p in zip(bands, parts):
        tensor[b:e] = p


def chain_blocks(base_res: int, nmips: int, world: int, rank: int, texel_bytes: int = 8):
    """Per-level (byte_offset, byte_length) of the rows of a cube mip chain that `rank` owns when the
    prefilter is sharded (pbrcu_ibl_prefilter_ggx shard/nshards: level l has 6*res_l rows of res_l texels,
    cut with split_range).  Level order and offsets are those of pbrcu_cube_chain_texels."""
    out, off = [], 0
    for m in range(nmips):
        res = max(base_res >> m, 1)
        b, e = split_range(6 * res, world, rank)
        out.append((off + b * res * texel_bytes, (e - b) * res * texel_bytes))
        off += 6 * res * res * texel_bytes
    return out


def gather_chain(chain, base_res: int, nmips: int, world: int, rank: int, dist=None, texel_bytes: int = 8, first_level: int = 0):
    """All-gathers a sharded prefilter result: `chain` is the full-size flat uint8 tensor in which this rank
    has written only its chain_blocks; on return every rank holds the whole chain.  Levels whose rows divide
    evenly go through ONE in-place all-gather each (NCCL: all_gather_into_tensor), the tiny tail levels
    through per-owner broadcasts.  `first_level` = 1 leaves level 0 out of the exchange: the prefiltered
    level 0 is a copy of the source level every rank already holds (K4/K5 are replicated), so a rank copies it
    whole (shard 0 of 1 for that level) and three quarters of the chain's bytes never touch NVLink."""
    if dist is None or world == 1:
        return chain
    blocks = [chain_blocks(base_res, nmips, world, r, texel_bytes) for r in range(world)]
    for m in range(first_level, nmips):
        lens = {blocks[r][m][1] for r in range(world)}
        off0 = blocks[0][m][0]
        if len(lens) == 1 and chain.is_cuda and hasattr(dist, "all_gather_into_tensor"):
            n = blocks[0][m][1]
            o, _ = blocks[rank][m]
            dist.all_gather_into_tensor(chain[off0:off0 + n * world], chain[o:o + n])
        else:
            for r in range(world):
                o, n = blocks[r][m]
                if n:
                    dist.broadcast(chain[o:o + n], src=r)
    return chain


class BandViews:
    """Sub-views of a G-buffer and of its light-tile lists restricted to rows [y0, y1) (y0 a multiple of 16):
    plain pointer arithmetic on the C structs, so a rank can cluster and shade ONLY its band with the ordinary
    entry points (pbrcu_cluster_lights / pbrcu_shade_ggx_ibl) and still write the whole-frame buffers in place."""

    def __init__(self, gbuffer, tiles, out_offset_rows):
        self.gbuffer, self.tiles, self.out_offset_rows = gbuffer, tiles, out_offset_rows


def band_views(pbr, g, tiles, y0: int, y1: int, tile: int = 16) -> BandViews:
    if y0 % tile != 0:
        raise ValueError("a band must start on a light-tile boundary")
    gv = pbr.GBuffer()
    gv.width, gv.height, gv.pitch8, gv.pitch4 = g.width, max(y1 - y0, 0), g.pitch8, g.pitch4
    for name, pitch in (("positions", g.pitch8), ("normals", g.pitch8), ("albedo", g.pitch8), ("emissive", g.pitch8), ("orm", g.pitch4), ("depth", g.pitch4)):
        p = getattr(g, name)
        setattr(gv, name, (p + y0 * pitch) if p else None)
    tv = None
    if tiles is not None:
        tv = pbr.LightTiles()
        ty0 = y0 // tile
        tv.tiles_x, tv.tiles_y, tv.max_per_tile = tiles.tiles_x, (max(y1 - y0, 0) + tile - 1) // tile, tiles.max_per_tile
        tv.counts = tiles.counts + ty0 * tiles.tiles_x * 4
        tv.indices = tiles.indices + ty0 * tiles.tiles_x * tiles.max_per_tile * 2
    return BandViews(gv, tv, y0)


def cost_balanced_bands(tile_row_costs, height: int, world: int, tile: int = 16) -> list[tuple[int, int]]:
    """Row bands [y0, y1) for every rank, cut on light-tile boundaries so that each band carries about the same COST
    instead of the same number of rows.  `tile_row_costs[r]` is the estimated shading cost of tile row r -- e.g. from the
    previous frame's light lists, sum over the row's tiles of (fixed + lights in the tile).  Every rank computes the same
    cuts from the same numbers, so no exchange is needed.  Bands are contiguous, cover [0, height) and may be empty."""
    costs = [max(float(c), 0.0) for c in tile_row_costs]
    n = len(costs)
    if n != (height + tile - 1) // tile:
        raise ValueError("one cost per tile row expected")
    total = sum(costs)
    if total <= 0.0:
        return [row_band(height, world, r, align=tile) for r in range(world)]
    cuts, acc, r = [0], 0.0, 0
    for k in range(1, world):
        target = total * k / world
        while r < n and acc + costs[r] * 0.5 <= target:  # a row goes to the side of the cut most of it lies on
            acc += costs[r]
            r += 1
        cuts.append(r)
    cuts.append(n)
    return [(min(cuts[k] * tile, height), min(cuts[k + 1] * tile, height)) for k in range(world)]


def _cpulist(text: str) -> set[int]:
    """'0-15,32-47' -> {0..15, 32..47} (the sysfs cpulist syntax)."""
    cpus: set[int] = set()
    for part in text.strip().split(","):
        if not part:
            continue
        lo, _, hi = part.partition("-")
        cpus.update(range(int(lo), int(hi or lo) + 1))
    return cpus


def bind_host_near_device(pci_bus_id: str, sysfs: str = "/sys") -> dict:
    """Pins the calling process to the CPUs of the NUMA node its GPU hangs off, so that the pinned
    staging buffers allocated afterwards (first touch) and the copy-issuing thread sit on the socket whose
    PCIe root complex serves that GPU.  One rank per GPU on a two-socket box otherwise has half its
    host<->device copies cross the socket interconnect.  `pci_bus_id` is 'dddd:bb:dd.f' as CUDA reports it.
    Does nothing (and says so) on a single-node host or when sysfs has no answer.  The reference runs one
    host thread on one GPU (`core/GpuHandle.cpp:77-81`) and has no counterpart."""
    import os
    info = {"pci_bus_id": pci_bus_id, "node": None, "bound": False}
    try:
        nodes = [d for d in os.listdir(f"{sysfs}/devices/system/node") if d.startswith("node") and d[4:].isdigit()]
        info["nodes"] = len(nodes)
        with open(f"{sysfs}/bus/pci/devices/{pci_bus_id.lower()}/numa_node") as f:
            node = int(f.read().strip())
        info["node"] = node
        if node < 0 or len(nodes) < 2:
            return info
        with open(f"{sysfs}/devices/system/node/node{node}/cpulist") as f:
            cpus = _cpulist(f.read()) & os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
            info["bound"], info["cpus"] = True, len(cpus)
    except (OSError, ValueError) as e:
        info["why_not"] = f"{type(e).__name__}: {e}"
    return info


class PeerFrames:
    """This rank's frame buffer in symmetric memory plus the same buffer of every other rank mapped into this
    process: what `pbrcu_shade_ggx_ibl_replicated` needs to store a finished row band straight into the peers'
    frames over NVLink.  `local` is a uint8 tensor; `peers` the mapped device addresses of the other ranks' buffers
    (rank order, self left out); `barrier()` = every rank's launches so far have completed (device-side, on the
    current stream) -- call it after the replicated launch and before reading or overwriting a frame."""

    def __init__(self, nbytes: int, device, dist):
        import torch
        import torch.distributed._symmetric_memory as symm

        self.local = symm.empty(nbytes, dtype=torch.uint8, device=device)
        self._handle = symm.rendezvous(self.local, dist.group.WORLD)
        h = self._handle
        off = int(getattr(h, "offset", 0))
        self.rank, self.world = int(h.rank), int(h.world_size)
        ptrs = [int(p) + off for p in h.buffer_ptrs]
        if ptrs[self.rank] != self.local.data_ptr():
            raise RuntimeError("symmetric memory: the mapped address of this rank's buffer is not the tensor's")
        self.peers = [p for r, p in enumerate(ptrs) if r != self.rank]
        # NVSwitch multicast address of the same buffers (0 when the box has no NVLS multicast)
        mc = int(getattr(h, "multicast_ptr", 0) or 0)
        self.multicast = mc + off if mc else 0

    def replica_array(self, byte_offset: int = 0):
        """(ctypes array of void*, count) of the peers' buffers advanced by byte_offset -- the replica_outs argument."""
        import ctypes as C
        arr = (C.c_void_p * max(len(self.peers), 1))(*[p + byte_offset for p in self.peers])
        return arr, len(self.peers)

    def barrier(self):
        self._handle.barrier()
