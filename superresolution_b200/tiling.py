"""Whole-field tiled inference (BASELINE config 3).  Not in the reference (Modello_5 iterates
pre-cut patches): a frame is cut into overlapping tile x tile inputs on a regular stride (last row/column clamped to
the frame edge), every tile is an independent model forward, and the 4x outputs are stitched with a separable ramp
weight (libsrb blend kernels).  Two shardings, neither with a collective:
  seams='recompute'  rank r owns a band of OUTPUT rows and runs every tile that intersects it — tiles straddling a band
                     edge are recomputed by both neighbours (one extra tile row per boundary: 16 % more forwards at 8
                     ranks on the 43-row plan of config 3); no communication at all.
  seams='p2p'        the tile ROWS are dealt out by count; rank r owns the output rows from the top of its first tile row
                     to the top of rank r+1's, blends its own tiles only, and ships the strip of partial sums that hangs
                     over into rank r+1's rows (sum w*y and sum w, <= one overlap height) to that neighbour with one
                     point-to-point send / receive pair (torch.distributed isend / irecv: NCCL P2P over NVLink)."""
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
                seams='recompute', group=None, _exchange=None):
    """frame: [H,W] uint16 or float32 CUDA tensor.  Returns this rank's band of the stitched output
    ([r1-r0, scale*W], out_dtype) and (r0, r1).  ``seams``: see the module docstring."""
    if frame.device.type != 'cuda':
        raise RuntimeError('tiled inference runs on CUDA only')
    H, W = frame.shape
    FH, FW = scale * H, scale * W
    th = scale * tile
    ov = scale * (tile - stride)
    dev = frame.device
    lib = L.lib()
    if seams == 'p2p' and world > 1:
        plan_all = seam_plan(H, tile, stride, scale, world)
        rows, r0, own1, r1 = plan_all[rank]          # accumulate over [r0, r1) = own rows + the strip for rank + 1
        xs = tile_origins(W, tile, stride)
        tiles = [(y, x) for y in rows for x in xs]
    elif seams in ('recompute', 'p2p'):
        plan_all = None
        r0, r1 = band(rank, world, FH)
        tiles = [(y, x) for (y, x) in plan(H, W, tile, stride) if scale * y < r1 and scale * y + th > r0]
    else:
        raise ValueError(f"seams must be 'recompute' or 'p2p', got {seams!r}")
    sums = torch.zeros(2, r1 - r0, FW, dtype=torch.float32, device=dev)
    acc, wsum = sums[0], sums[1]
    with torch.no_grad():
        for i in range(0, len(tiles), batch):
            chunk = tiles[i:i + batch]
            x = torch.stack([frame[y:y + tile, x:x + tile] for (y, x) in chunk]).unsqueeze(1)
            if x.dtype != torch.uint16:
                x = x.float()
            y_t = model.run(x.contiguous(), 3)  # [T,1,th,th] fp32
            oy = torch.tensor([scale * t[0] for t in chunk], dtype=torch.int32, device=dev)
            ox = torch.tensor([scale * t[1] for t in chunk], dtype=torch.int32, device=dev)
            with torch.cuda.device(dev):
                L.check(lib.srb_blend_tiles(L.ptr(y_t), L.ptr(oy), L.ptr(ox), len(chunk), th, th, max(ov, 1),
                                            L.ptr(acc), L.ptr(wsum), FH, FW, r0, r1, L.current_stream()), 'blend_tiles')
        if plan_all is not None:
            (_exchange or exchange_seams_planned)(sums, plan_all, rank, group)  # _exchange: test hook
            r1 = own1
            acc, wsum = sums[0, :r1 - r0].contiguous(), sums[1, :r1 - r0].contiguous()
        out = torch.empty(r1 - r0, FW, dtype=out_dtype, device=dev)
        if out.numel() == 0:
            return out, (r0, r1)
        with torch.cuda.device(dev):
            L.check(lib.srb_blend_finalize(L.ptr(acc), L.ptr(wsum), L.ptr(out),
                                           L.SRB_U16 if out_dtype == torch.uint16 else L.SRB_F32, acc.numel(),
                                           L.current_stream()), 'blend_finalize')
    return out, (r0, r1)
