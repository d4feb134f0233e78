"""CPU tests of the multi-GPU host logic: partitions are disjoint and complete, and a world_size-2 gloo
run of the gather reassembles a frame from per-rank row bands (the N>1 path of bench.py, minus the GPU)."""
import os
import socket

import numpy as np
import pytest


def test_partitions_cover_exactly(pbr):
    P = pbr.partition
    for n in (1, 7, 256, 2160, 4320, 12288):
        for world in (1, 2, 3, 4, 8):
            blocks = [P.split_range(n, world, r) for r in range(world)]
            assert blocks[0][0] == 0 and blocks[-1][1] == n
            assert all(blocks[i][1] == blocks[i + 1][0] for i in range(world - 1))
            assert max(e - b for b, e in blocks) - min(e - b for b, e in blocks) <= 1
    assert P.row_band(2160, 8, 3) == (810, 1080)
    bands = [P.row_band(2160, 8, r, align=16) for r in range(8)]
    assert bands[0][0] == 0 and bands[-1][1] == 2160 and all(b % 16 == 0 for b, _ in bands)
    assert P.view_block(256, 8, 7) == (224, 256)
    with pytest.raises(ValueError):
        P.split_range(10, 2, 2)
    # prefilter shards: per level, the ranks' row blocks tile the level exactly
    for world in (1, 2, 3, 8):
        per_rank = [P.chain_blocks(2048, 12, world, r) for r in range(world)]
        off = 0
        for m in range(12):
            res = 2048 >> m
            assert per_rank[0][m][0] == off
            for r in range(world - 1):
                assert per_rank[r][m][0] + per_rank[r][m][1] == per_rank[r + 1][m][0]
            off += 6 * res * res * 8
            assert per_rank[-1][m][0] + per_rank[-1][m][1] == off


def test_cost_balanced_bands(pbr):
    P = pbr.partition
    rng = np.random.default_rng(1)
    for height, world in ((4320, 8), (4320, 2), (1080, 4), (100, 3), (16, 8)):
        n = (height + 15) // 16
        costs = rng.uniform(1, 10, n) * (1 + 5 * np.exp(-((np.arange(n) - n / 2) / (n / 6 + 1)) ** 2))  # heavy middle
        bands = P.cost_balanced_bands(costs, height, world)
        assert len(bands) == world and bands[0][0] == 0 and bands[-1][1] == height
        assert all(bands[i][1] == bands[i + 1][0] for i in range(world - 1)) and all(b % 16 == 0 for b, _ in bands)
        if n >= 8 * world:
            per = [costs[b // 16:(e + 15) // 16].sum() for b, e in bands]
            equal = [costs[b // 16:(e + 15) // 16].sum() for b, e in (P.row_band(height, world, r, align=16) for r in range(world))]
            assert max(per) <= max(equal) + 1e-9 and max(per) < 1.25 * sum(per) / world
    assert P.cost_balanced_bands([0, 0, 0], 48, 2) == [P.row_band(48, 2, 0, align=16), P.row_band(48, 2, 1, align=16)]
    with pytest.raises(ValueError):
        P.cost_balanced_bands([1.0], 48, 2)


def _worker(rank, world, port, height, q):
    import torch
    import torch.distributed as dist
    import pbr_b200 as pbr
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    y0, y1 = pbr.partition.row_band(height, world, rank)
    frame = torch.zeros(height, 8, 4, dtype=torch.int32)
    full = torch.arange(height * 32, dtype=torch.int32).reshape(height, 8, 4)
    frame[y0:y1] = full[y0:y1]  # "shade" only the rows this rank owns
    pbr.partition.gather_rows(frame, y0, y1, world, height, dist)
    # the sharded prefilter's chain gather: every rank fills only its row blocks of every level
    res, mips = 8, 4
    nbytes = sum(6 * (res >> m) ** 2 * 8 for m in range(mips))
    want = (torch.arange(nbytes, dtype=torch.int64) * 7 % 251).to(torch.uint8)
    chain = torch.zeros(nbytes, dtype=torch.uint8)
    for o, n in pbr.partition.chain_blocks(res, mips, world, rank):
        chain[o:o + n] = want[o:o + n]
    pbr.partition.gather_chain(chain, res, mips, world, rank, dist)
    ok_chain = bool(torch.equal(chain, want))
    # level 0 held whole by every rank (it is a copy of the replicated source level) and left out of the exchange
    chain1 = torch.zeros(nbytes, dtype=torch.uint8)
    n0 = 6 * res * res * 8
    chain1[:n0] = want[:n0]
    for o, n in pbr.partition.chain_blocks(res, mips, world, rank)[1:]:
        chain1[o:o + n] = want[o:o + n]
    pbr.partition.gather_chain(chain1, res, mips, world, rank, dist, first_level=1)
    ok_chain = ok_chain and bool(torch.equal(chain1, want))
    t = torch.tensor([float(rank + 1)])
    dist.all_reduce(t, op=dist.ReduceOp.MAX)  # the max-over-ranks timing reduction of bench.py
    q.put((rank, bool(torch.equal(frame, full)) and ok_chain, float(t.item())))
    dist.destroy_process_group()


@pytest.mark.parametrize("height", [64, 37])
def test_gloo_world2_gather(height):
    import torch.multiprocessing as mp
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, height, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
    assert res == [(0, True, 2.0), (1, True, 2.0)]


def test_bind_host_near_device_on_a_fake_sysfs(pbr, tmp_path):
    """Host placement for one rank per GPU: the rank is pinned to its GPU's NUMA node on a two-node host,
    left alone on a one-node host, and a missing sysfs entry is reported instead of raised."""
    P = pbr.partition
    assert P._cpulist("0-3,8,10-11\n") == {0, 1, 2, 3, 8, 10, 11}
    mine = sorted(os.sched_getaffinity(0))
    try:
        def fake(nodes, gpu_node):
            root = tmp_path / f"sys{nodes}{gpu_node}"
            for n in range(nodes):
                d = root / "devices/system/node" / f"node{n}"
                d.mkdir(parents=True)
                cpus = mine[n::nodes] if nodes > 1 else mine
                (d / "cpulist").write_text(",".join(map(str, cpus)) + "\n")
            (root / "devices/system/node/possible").write_text("0\n")  # a non-node entry, as in the real sysfs
            dev = root / "bus/pci/devices/0000:1b:00.0"
            dev.mkdir(parents=True)
            (dev / "numa_node").write_text(f"{gpu_node}\n")
            return str(root)

        one = P.bind_host_near_device("0000:1B:00.0", sysfs=fake(1, 0))
        assert one["bound"] is False and one["node"] == 0 and one["nodes"] == 1 and os.sched_getaffinity(0) == set(mine)
        unknown = P.bind_host_near_device("0000:1B:00.0", sysfs=fake(2, -1))
        assert unknown["bound"] is False and unknown["node"] == -1
        absent = P.bind_host_near_device("0000:ff:00.0", sysfs=fake(3, 0))
        assert absent["bound"] is False and "why_not" in absent
        if len(mine) >= 2:
            two = P.bind_host_near_device("0000:1B:00.0", sysfs=fake(2, 1))
            assert two["bound"] is True and two["node"] == 1 and os.sched_getaffinity(0) == set(mine[1::2])
    finally:
        os.sched_setaffinity(0, mine)


def test_c_row_bands_equal_the_python_partition():
    """pbrcu_mg_row_bands (the C multi-GPU host's band cuts; touches no CUDA state) == partition.cost_balanced_bands /
    row_band on random costs, heights that are not tile multiples, more members than tile rows, all-zero costs."""
    import ctypes as C

    import numpy as np
    import pbr_b200 as pbr
    from pbr_b200.partition import cost_balanced_bands, row_band
    lib = pbr.load()
    rng = np.random.default_rng(5)
    for height, n in [(4320, 8), (2160, 4), (1080, 2), (100, 3), (33, 16), (16, 5), (7, 2), (4320, 1)]:
        rows = (height + 15) // 16
        out = (C.c_uint32 * (2 * n))()
        lib.mg_row_bands(height, n, None, out)
        assert [(out[2 * k], out[2 * k + 1]) for k in range(n)] == [row_band(height, n, r, align=16) for r in range(n)]
        for costs in (rng.uniform(0, 50, rows).astype(np.float32), np.zeros(rows, np.float32), (rng.uniform(0, 1, rows) ** 8 * 1000).astype(np.float32)):
            lib.mg_row_bands(height, n, costs.ctypes.data_as(C.POINTER(C.c_float)), out)
            got = [(out[2 * k], out[2 * k + 1]) for k in range(n)]
            assert got == cost_balanced_bands(costs.astype(np.float64), height, n), (height, n)
            assert got[0][0] == 0 and got[-1][1] == height and all(got[k][1] == got[k + 1][0] for k in range(n - 1))
