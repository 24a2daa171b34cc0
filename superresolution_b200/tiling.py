"""Whole-field tiled inference (BASELINE config 3).  Not in the reference (Modello_5 iterates
pre-cut patches): a frame is cut into overlapping tile x tile inputs on a regular stride (last row/column clamped to
the frame edge), every tile is an independent model forward, and the 4x outputs are stitched with a separable ramp
weight (libsrb blend kernels).  Sharding: rank r owns a band of OUTPUT rows and runs every tile that intersects it —
tiles straddling a band edge are recomputed by both neighbours, so there is no collective and no peer copy."""
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


def infer_frame(model, frame, tile=128, stride=96, batch=16, rank=0, world=1, out_dtype=torch.uint16, scale=4):
    """frame: [H,W] uint16 or float32 CUDA tensor.  Returns this rank's band of the stitched output
    ([r1-r0, scale*W], out_dtype) and (r0, r1)."""
    if frame.device.type != 'cuda':
        raise RuntimeError('tiled inference runs on CUDA only')
    H, W = frame.shape
    FH, FW = scale * H, scale * W
    r0, r1 = band(rank, world, FH)
    th = scale * tile
    ov = scale * (tile - stride)
    tiles = [(y, x) for (y, x) in plan(H, W, tile, stride) if scale * y < r1 and scale * y + th > r0]
    dev = frame.device
    lib = L.lib()
    acc = torch.zeros(r1 - r0, FW, dtype=torch.float32, device=dev)
    wsum = torch.zeros_like(acc)
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
        out = torch.empty(r1 - r0, FW, dtype=out_dtype, device=dev)
        with torch.cuda.device(dev):
            L.check(lib.srb_blend_finalize(L.ptr(acc), L.ptr(wsum), L.ptr(out),
                                           L.SRB_U16 if out_dtype == torch.uint16 else L.SRB_F32, acc.numel(),
                                           L.current_stream()), 'blend_finalize')
    return out, (r0, r1)
