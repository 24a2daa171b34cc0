"""Host logic of tiled whole-field inference: tile plan coverage, band sharding, and the N>1 path on CPU with a
world_size-2 gloo group (each rank blends its band with the oracle blend; the gathered frame equals the 1-rank one)."""
import os

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def test_plan_covers_frame():
    from superresolution_b200 import tiling
    from oracle import tiling as otil
    for size, tile, stride in ((4096, 128, 96), (160, 32, 24), (128, 128, 96), (100, 32, 32)):
        xs = tiling.tile_origins(size, tile, stride)
        assert xs == otil.tile_origins(size, tile, stride)
        cov = np.zeros(size, bool)
        for x in xs:
            assert 0 <= x and x + tile <= max(size, tile)
            cov[x:x + tile] = True
        assert cov.all()
    assert len(tiling.plan(4096, 4096)) == 43 * 43  # BASELINE config 3 / SURVEY.md §8d


def test_bands_partition_rows():
    from superresolution_b200 import tiling
    for world in (1, 2, 3, 8):
        rows = [tiling.band(r, world, 16384) for r in range(world)]
        assert rows[0][0] == 0 and rows[-1][1] == 16384
        assert all(rows[i][1] == rows[i + 1][0] for i in range(world - 1))


def _worker(rank, world, port, q):
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    from superresolution_b200 import tiling
    from oracle import tiling as otil
    H = W = 80
    tile, stride, scale = 32, 24, 4
    rng = np.random.default_rng(0)
    origins = tiling.plan(H, W, tile, stride)
    outs = [rng.random((scale * tile, scale * tile), dtype=np.float32) for _ in origins]  # stand-in tile outputs
    FH = FW = scale * H
    r0, r1 = tiling.band(rank, world, FH)
    mine = [(o, t) for o, t in zip(origins, outs) if scale * o[0] < r1 and scale * o[0] + scale * tile > r0]
    full = otil.blend([t for _, t in mine], [(scale * o[0], scale * o[1]) for o, _ in mine], FH, FW,
                      scale * (tile - stride))
    band = torch.from_numpy(full[r0:r1].copy())
    gathered = [torch.empty(tiling.band(r, world, FH)[1] - tiling.band(r, world, FH)[0], FW) for r in range(world)]
    dist.all_gather(gathered, band) if len({g.shape for g in gathered}) == 1 else None
    if rank == 0:
        ref = otil.blend(outs, [(scale * o[0], scale * o[1]) for o in origins], FH, FW, scale * (tile - stride))
        q.put(float(np.abs(torch.cat(gathered).numpy() - ref).max()))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_band_sharding_matches_single_rank():
    ctx = mp.get_context('spawn')
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    err = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert err < 1e-6


def _allreduce_worker(rank, world, port, q):
    import os
    import torch
    import torch.distributed as dist
    from superresolution_b200.optim import allreduce_gradients
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port))
    dist.init_process_group('gloo', rank=rank, world_size=world)
    g = torch.Generator().manual_seed(100 + rank)
    flat = torch.randn(100_003, generator=g)
    allreduce_gradients(flat, bucket_bytes=64 << 10)  # 7 buckets, the last one ragged
    # the overlapped form of train_step: the tail range (stage 2) first, other work in between, then the head range
    from superresolution_b200.optim import allreduce_gradients_start, allreduce_gradients_wait
    flat2 = torch.randn(100_003, generator=torch.Generator().manual_seed(100 + rank))
    works = allreduce_gradients_start(flat2, bucket_bytes=64 << 10, lo=60_001)
    head = flat2[:10].clone()
    flat2[:10] = 0.0   # "stage-1 backward" still writes the head range while the tail is in flight
    flat2[:10] = head
    works += allreduce_gradients_start(flat2, bucket_bytes=64 << 10, hi=60_001)
    allreduce_gradients_wait(works)
    assert torch.equal(flat2, flat), (flat2 - flat).abs().max()
    q.put((rank, flat[:5].tolist(), float(flat.double().sum())))
    dist.destroy_process_group()


def test_bucketed_gradient_allreduce_world2():
    """N>1 path of the training step's exchange (replaces DataParallel's reduce-add): gloo, world 2, on CPU."""
    import torch
    import torch.multiprocessing as mp
    from superresolution_b200.optim import bucket_ranges
    assert bucket_ranges(10, 4) == [(0, 4), (4, 8), (8, 10)]
    ctx = mp.get_context('spawn')
    q = ctx.Queue()
    port = 29500 + (os.getpid() % 2000)
    ps = [ctx.Process(target=_allreduce_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in ps:
        p.start()
    res = sorted(q.get(timeout=120) for _ in ps)
    for p in ps:
        p.join(60)
    want = sum(torch.randn(100_003, generator=torch.Generator().manual_seed(100 + r)) for r in range(2))
    for _, head, tot in res:
        assert head == want[:5].tolist()
        assert abs(tot - float(want.double().sum())) < 1e-6


def _seam_worker(rank, world, port, q):
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    from superresolution_b200 import tiling
    from oracle import tiling as otil
    H, W = 128, 80
    tile, stride, scale = 32, 24, 4
    ov = scale * (tile - stride)
    rng = np.random.default_rng(0)
    origins = tiling.plan(H, W, tile, stride)
    outs = {o: rng.random((scale * tile, scale * tile), dtype=np.float32) for o in origins}  # stand-in tile outputs
    FH, FW = scale * H, scale * W
    plan_all = tiling.seam_plan(H, tile, stride, scale, world)
    rows, own0, own1, end = plan_all[rank]
    sums = np.zeros((2, end - own0, FW), np.float32)
    for (y, x), t in outs.items():
        if y not in rows:
            continue
        w = np.outer(otil.ramp(scale * tile, ov), otil.ramp(scale * tile, ov)).astype(np.float32)
        oy, ox = scale * y - own0, scale * x
        sums[0, oy:oy + scale * tile, ox:ox + scale * tile] += w * t
        sums[1, oy:oy + scale * tile, ox:ox + scale * tile] += w
    sums = tiling.exchange_seams_planned(torch.from_numpy(sums), plan_all, rank).numpy()
    mine = sums[0, :own1 - own0] / np.maximum(sums[1, :own1 - own0], 1e-30)
    parts = [None] * world
    dist.all_gather_object(parts, (own0, own1, mine))
    if rank == 0:
        ref = otil.blend(list(outs.values()), [(scale * y, scale * x) for (y, x) in outs], FH, FW, ov)
        got = np.zeros_like(ref)
        covered = np.zeros(FH, bool)
        for a, b, m in parts:
            got[a:b] = m
            assert not covered[a:b].any()
            covered[a:b] = True
        q.put((bool(covered.all()), float(np.abs(got - ref).max()), [len(p[0]) for p in plan_all]))
    dist.barrier()
    dist.destroy_process_group()


def test_three_rank_seam_exchange_matches_single_rank():
    """seams='p2p': tile rows dealt out by count, every tile computed once, overhanging partial sums shipped to the next
    rank point-to-point — the stitched frame equals the single-rank blend (fp32 addition order differs: 1e-6)."""
    ctx = mp.get_context('spawn')
    q = ctx.Queue()
    port = 31500 + os.getpid() % 2000
    procs = [ctx.Process(target=_seam_worker, args=(r, 3, port, q)) for r in range(3)]
    for p in procs:
        p.start()
    covered, err, counts = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert covered and err < 1e-6
    assert sum(counts) == 5 and max(counts) - min(counts) <= 1  # 5 tile rows over 3 ranks, none computed twice


def test_seam_plan_partitions_rows_and_tiles():
    from superresolution_b200 import tiling
    for world in (1, 2, 3, 8, 50):
        pl = tiling.seam_plan(4096, 128, 96, 4, world)
        rows = [y for p in pl for y in p[0]]
        assert rows == tiling.tile_origins(4096, 128, 96)           # every tile row exactly once
        edges = [(p[1], p[2]) for p in pl if p[0]]
        assert edges[0][0] == 0 and edges[-1][1] == 16384
        assert all(edges[i][1] == edges[i + 1][0] for i in range(len(edges) - 1))
        assert all(p[3] - p[2] <= 4 * 128 for p in pl)               # a strip is at most one tile high


def _tile_worker(rank, world, port, q):
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    from superresolution_b200 import tiling
    from oracle import tiling as otil
    H, W = 128, 104
    tile, stride, scale = 32, 24, 4
    ov = scale * (tile - stride)
    rng = np.random.default_rng(0)
    origins = tiling.plan(H, W, tile, stride)
    outs = {o: rng.random((scale * tile, scale * tile), dtype=np.float32) for o in origins}  # stand-in tile outputs
    FH, FW = scale * H, scale * W
    plan_all = tiling.tile_plan(H, W, tile, stride, scale, world)
    me = plan_all[rank]
    b0, b1 = me['buf']
    sums = np.zeros((2, b1 - b0, FW), np.float32)
    w = np.outer(otil.ramp(scale * tile, ov), otil.ramp(scale * tile, ov)).astype(np.float32)
    for (y, x) in me['tiles']:
        oy, ox = scale * y - b0, scale * x
        sums[0, oy:oy + scale * tile, ox:ox + scale * tile] += w * outs[(y, x)]
        sums[1, oy:oy + scale * tile, ox:ox + scale * tile] += w
    own = tiling.exchange_tiles_planned(torch.from_numpy(sums), plan_all, rank).numpy()
    mine = own[0] / np.maximum(own[1], 1e-30)
    parts = [None] * world
    dist.all_gather_object(parts, (me['own'][0], me['own'][1], mine))
    if rank == 0:
        ref = otil.blend(list(outs.values()), [(scale * y, scale * x) for (y, x) in outs], FH, FW, ov)
        got = np.zeros_like(ref)
        covered = np.zeros(FH, bool)
        for a, b, m in parts:
            got[a:b] = m
            assert not covered[a:b].any()
            covered[a:b] = True
        q.put((bool(covered.all()), float(np.abs(got - ref).max()), [len(p['tiles']) for p in plan_all]))
    dist.barrier()
    dist.destroy_process_group()


def test_three_rank_tile_count_sharding_matches_single_rank():
    """seams='tiles': the tiles (not the tile rows) are dealt out by count; tile rows shared by two ranks are stitched
    through the two-sided point-to-point exchange — the frame equals the single-rank blend."""
    ctx = mp.get_context('spawn')
    q = ctx.Queue()
    port = 33500 + os.getpid() % 2000
    procs = [ctx.Process(target=_tile_worker, args=(r, 3, port, q)) for r in range(3)]
    for p in procs:
        p.start()
    covered, err, counts = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert covered and err < 1e-6
    assert sum(counts) == 5 * 4 and max(counts) - min(counts) <= 1  # 20 tiles over 3 ranks: 6 / 7 / 7, rows are split


def test_tile_plan_is_balanced_and_partitions_the_frame():
    from superresolution_b200 import tiling
    every = tiling.plan(4096, 4096)
    for world in (1, 2, 3, 4, 8, 50):
        pl = tiling.tile_plan(4096, 4096, 128, 96, 4, world)
        assert [t for p in pl for t in p['tiles']] == every           # every tile exactly once, raster order
        cnt = [len(p['tiles']) for p in pl]
        assert max(cnt) - min(cnt) <= 1                               # balance by tile COUNT
        own = [p['own'] for p in pl]
        assert own[0][0] == 0 and own[-1][1] == 16384 and all(own[i][1] == own[i + 1][0] for i in range(world - 1))
        for p in pl:  # the local accumulator covers the rank's tiles and its own range
            assert p['buf'][0] <= min(4 * p['tiles'][0][0], p['own'][0]) and p['buf'][1] >= max(4 * (p['tiles'][-1][0] + 128), p['own'][1])
    assert [len(p['tiles']) for p in tiling.tile_plan(4096, 4096, 128, 96, 4, 8)].count(231) == 7
