"""ORACLE (test infrastructure): CPU restatement of the tile plan + ramp blend used for whole-field inference.
The reference has no tiled inference (scripts/Modello_5.py:72-85 iterates pre-cut patches) so this defines the
semantics the CUDA blend kernels are checked against: w(i) = min(i+1, n-i, ov)/ov, out = sum(w*y)/sum(w)."""
import numpy as np


def tile_origins(size, tile, stride):
    if size <= tile:
        return [0]
    xs = list(range(0, size - tile, stride))
    xs.append(size - tile)
    return xs


def ramp(n, ov):
    i = np.arange(n)
    return (np.minimum(np.minimum(i + 1, n - i), ov) / np.float32(ov)).astype(np.float32)


def blend(tile_outputs, origins, FH, FW, ov):
    """tile_outputs: list of [th,tw] float32 arrays; origins: list of (oy, ox) in output pixels."""
    acc = np.zeros((FH, FW), np.float64)
    ws = np.zeros((FH, FW), np.float64)
    for t, (oy, ox) in zip(tile_outputs, origins):
        th, tw = t.shape
        w = np.outer(ramp(th, ov), ramp(tw, ov)).astype(np.float32)
        acc[oy:oy + th, ox:ox + tw] += (w * t).astype(np.float64)
        ws[oy:oy + th, ox:ox + tw] += w
    return (acc / np.maximum(ws, 1e-30)).astype(np.float32)
