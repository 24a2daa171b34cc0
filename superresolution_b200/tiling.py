"""Whole-field tiled inference (BASELINE config 3).  Not in the reference (Modello_5 iterates
pre-cut patches): a frame is cut into overlapping tile x tile inputs on a regular stride (last row/column clamped to
the frame edge), every tile is an independent model forward, and the 4x outputs are stitched with a separable ramp
weight (libsrb blend kernels).  Three shardings, none with a collective:
  seams='recompute'  rank r owns a band of OUTPUT rows and runs every tile that intersects it — tiles straddling a band
                     edge are recomputed by both neighbours (one extra tile row per boundary: 16 % more forwards at 8
                     ranks on the 43-row plan of config 3); no communication at all.
  seams='p2p'        the tile ROWS are dealt out by count; rank r owns the output rows from the top of its first tile row
                     to the top of rank r+1's, blends its own tiles only, and ships the strip of partial sums that hangs
                     over into rank r+1's rows (sum w*y and sum w, <= one overlap height) to that neighbour with one
                     point-to-point send / receive pair (torch.distributed isend / irecv: NCCL P2P over NVLink).
  seams='tiles'      the TILES are dealt out by count in raster order (config 3: 1849 tiles -> 231 or 232 per rank at 8
                     ranks instead of 5 or 6 whole tile rows): perfectly balanced, every tile computed once.  A tile row
                     shared by two ranks belongs to the rank that starts it; every rank ships the rows of its partial
                     sums that fall into a neighbour's range (up and down) point-to-point.  Default for world > 1."""
import torch

from . import _lib as L


def tile_origins(size, tile, stride):
    """1-D tile origins covering [0,size): regular stride, last origin clamped to size - tile."""
    if size <= tile:
        return [0]
    xs = list(range(0, size - tile, stride))
    xs.append(size - tile)
    return xs


def plan(H, W, tile=128, stride=96):
    return [(y, x) for y in tile_origins(H, tile, stride) for x in tile_origins(W, tile, stride)]


def band(rank, world, out_rows):
    r0 = (out_rows * rank) // world
    r1 = (out_rows * (rank + 1)) // world
    return r0, r1


def seam_plan(H, tile, stride, scale, world):
    """seams='p2p' geometry.  Returns per rank (tile_row_origins, own0, own1, end): the rank runs the tile rows whose
    input origins are listed, owns output rows [own0, own1) and accumulates rows [own0, end) locally; rows [own1, end)
    are the strip it sends to rank + 1.  Ranks beyond the number of tile rows get no tiles and an empty range."""
    ys = tile_origins(H, tile, stride)
    n = len(ys)
    used = min(world, n)
    cut = [(n * r) // used for r in range(used + 1)]
    FH = scale * H
    out = []
    for r in range(world):
        if r >= used:
            out.append(([], FH, FH, FH))
            continue
        mine = ys[cut[r]:cut[r + 1]]
        own0 = 0 if r == 0 else scale * mine[0]
        own1 = FH if r == used - 1 else scale * ys[cut[r + 1]]
        end = scale * (mine[-1] + tile)
        if r < used - 1 and end > (FH if r + 1 == used - 1 else scale * ys[cut[r + 2]]):
            raise ValueError('seam_plan: a tile of one rank overhangs beyond its neighbour (stride too small for this world)')
        out.append((mine, own0, own1, max(end, own1)))
    return out


def tile_plan(H, W, tile, stride, scale, world):
    """seams='tiles' geometry.  Returns per rank a dict: ``tiles`` (its (y, x) input origins, consecutive in raster order,
    counts differ by at most one), ``own`` = (own0, own1) the output rows it finalises, ``buf`` = (b0, b1) the output rows
    its local accumulator covers (its tiles' rows and its own range).  Own ranges partition [0, scale*H)."""
    ys, xs = tile_origins(H, tile, stride), tile_origins(W, tile, stride)
    every = [(y, x) for y in ys for x in xs]
    n, nx, FH = len(every), len(xs), scale * H
    used = min(world, n)
    cut = [(n * r) // used for r in range(used + 1)]
    edge = [0]
    for r in range(1, used):
        a, col = divmod(cut[r], nx)
        if col:  # the tile row is started by an earlier rank: it belongs to that rank
            a += 1
        edge.append(scale * ys[a] if a < len(ys) else FH)
    edge.append(FH)
    for r in range(used):
        edge[r + 1] = max(edge[r + 1], edge[r])
    out = []
    for r in range(world):
        if r >= used:
            out.append({'tiles': [], 'own': (FH, FH), 'buf': (FH, FH)})
            continue
        mine = every[cut[r]:cut[r + 1]]
        top, bot = scale * mine[0][0], scale * (mine[-1][0] + tile)
        out.append({'tiles': mine, 'own': (edge[r], edge[r + 1]), 'buf': (min(top, edge[r]), max(bot, edge[r + 1]))})
    return out


def exchange_tiles_planned(sums, plan_all, rank, group=None):
    """seams='tiles' exchange: sums [2, b1 - b0, FW] fp32 holds this rank's partial sums over its buffer rows.  Rows that
    lie in another rank's own range are sent there; rows other ranks hold for this rank's own range are received and
    added in rank order (deterministic).  Returns the [2, own1 - own0, FW] sums of the own range."""
    import torch.distributed as dist
    b0, b1 = plan_all[rank]['buf']
    o0, o1 = plan_all[rank]['own']
    ops, recv = [], []
    for q, pq in enumerate(plan_all):
        if q == rank:
            continue
        lo, hi = max(b0, pq['own'][0]), min(b1, pq['own'][1])      # my rows inside q's range
        if hi > lo and plan_all[rank]['tiles']:
            ops.append(dist.P2POp(dist.isend, sums[:, lo - b0:hi - b0].contiguous(), q, group))
        lo, hi = max(pq['buf'][0], o0), min(pq['buf'][1], o1)      # q's rows inside my range
        if hi > lo and pq['tiles']:
            buf = torch.empty(2, hi - lo, sums.shape[2], dtype=sums.dtype, device=sums.device)
            recv.append((lo, hi, buf))
            ops.append(dist.P2POp(dist.irecv, buf, q, group))
    if ops:
        for w in dist.batch_isend_irecv(ops):
            w.wait()
    own = sums[:, o0 - b0:o1 - b0].contiguous()
    for lo, hi, buf in recv:
        own[:, lo - o0:hi - o0] += buf
    return own


def exchange_seams_planned(sums, plan_all, rank, group=None):
    """seams='p2p' exchange for the geometry ``plan_all`` = seam_plan(...): rank r sends its rows [own1, end) (relative
    to own0) to r + 1 and adds what r - 1 sends onto its first rows.  sums: [2, end - own0, FW] fp32, any device."""
    import torch.distributed as dist
    world = len(plan_all)
    _, own0, own1, end = plan_all[rank]
    reqs, strip = [], None
    if rank + 1 < world and plan_all[rank + 1][0] and end > own1:
        reqs.append(dist.P2POp(dist.isend, sums[:, own1 - own0:].contiguous(), rank + 1, group))
    if rank > 0 and plan_all[rank][0]:
        _, p0, p1, pend = plan_all[rank - 1]
        if pend > p1:
            strip = torch.empty(2, pend - p1, sums.shape[2], dtype=sums.dtype, device=sums.device)
            reqs.append(dist.P2POp(dist.irecv, strip, rank - 1, group))
    if reqs:
        for w in dist.batch_isend_irecv(reqs):
            w.wait()
    if strip is not None:
        sums[:, :strip.shape[1]] += strip
    return sums


def infer_frame(model, frame, tile=128, stride=96, batch=16, rank=0, world=1, out_dtype=torch.uint16, scale=4,
                seams=None, group=None, _exchange=None):
    """frame: [H,W] uint16 or float32 CUDA tensor.  Returns this rank's rows of the stitched output
    ([r1-r0, scale*W], out_dtype) and (r0, r1).  ``seams``: see the module docstring (default 'tiles' when world > 1)."""
    if frame.device.type != 'cuda':
        raise RuntimeError('tiled inference runs on CUDA only')
    if seams is None:
        seams = 'tiles' if world > 1 else 'recompute'
    H, W = frame.shape
    FH, FW = scale * H, scale * W
    th = scale * tile
    ov = scale * (tile - stride)
    dev = frame.device
    lib = L.lib()
    plan_all = None
    if seams == 'tiles' and world > 1:
        plan_all = tile_plan(H, W, tile, stride, scale, world)
        tiles = plan_all[rank]['tiles']
        r0, r1 = plan_all[rank]['buf']
    elif seams == 'p2p' and world > 1:
        plan_all = seam_plan(H, tile, stride, scale, world)
        rows, r0, own1, r1 = plan_all[rank]          # accumulate over [r0, r1) = own rows + the strip for rank + 1
        xs = tile_origins(W, tile, stride)
        tiles = [(y, x) for y in rows for x in xs]
    elif seams in ('recompute', 'p2p', 'tiles'):
        r0, r1 = band(rank, world, FH)
        tiles = [(y, x) for (y, x) in plan(H, W, tile, stride) if scale * y < r1 and scale * y + th > r0]
    else:
        raise ValueError(f"seams must be 'recompute', 'p2p' or 'tiles', got {seams!r}")
    if frame.dtype != torch.uint16:
        frame = frame.float()
    frame = frame.contiguous()
    fdt = L.SRB_U16 if frame.dtype == torch.uint16 else L.SRB_F32
    sums = torch.zeros(2, r1 - r0, FW, dtype=torch.float32, device=dev)
    acc, wsum = sums[0], sums[1]
    nt = len(tiles)
    with torch.no_grad(), torch.cuda.device(dev):
        if nt:
            # tile origins live on the device (one upload per frame); input and output origins per batch are slices
            org = torch.tensor([[t[0] for t in tiles], [t[1] for t in tiles]], dtype=torch.int32).to(dev, non_blocking=True)
            org4 = org * scale
            xb = torch.empty(batch, 1, tile, tile, dtype=frame.dtype, device=dev)
        for i in range(0, nt, batch):
            n = min(batch, nt - i)
            L.check(lib.srb_gather_tiles(L.ptr(frame), fdt, H, W, L.ptr(org[0, i:]), L.ptr(org[1, i:]), n, tile,
                                         L.ptr(xb), L.current_stream()), 'gather_tiles')
            y_t = model.run(xb[:n], 3)  # [n,1,th,th] fp32
            L.check(lib.srb_blend_tiles(L.ptr(y_t), L.ptr(org4[0, i:]), L.ptr(org4[1, i:]), n, th, th, max(ov, 1),
                                        L.ptr(acc), L.ptr(wsum), FH, FW, r0, r1, L.current_stream()), 'blend_tiles')
        if plan_all is not None and seams == 'tiles':
            own = (_exchange or exchange_tiles_planned)(sums, plan_all, rank, group)  # _exchange: test hook
            r0, r1 = plan_all[rank]['own']
            acc, wsum = own[0], own[1]
        elif plan_all is not None:
            (_exchange or exchange_seams_planned)(sums, plan_all, rank, group)
            r1 = own1
            acc, wsum = sums[0, :r1 - r0].contiguous(), sums[1, :r1 - r0].contiguous()
        out = torch.empty(r1 - r0, FW, dtype=out_dtype, device=dev)
        if out.numel() == 0:
            return out, (r0, r1)
        L.check(lib.srb_blend_finalize(L.ptr(acc), L.ptr(wsum), L.ptr(out),
                                       L.SRB_U16 if out_dtype == torch.uint16 else L.SRB_F32, acc.numel(),
                                       L.current_stream()), 'blend_finalize')
    return out, (r0, r1)
