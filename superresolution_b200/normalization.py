"""LOG / 16-bit normalisation of reference ``scripts/Dataset_step4_normalization.py:161-179`` and the u16 <-> [0,1]
edges (``src/dataset.py:31-34``, ``scripts/Modello_5.py:29-33``) on the libsrb kernels."""
import ctypes as C

import numpy as np
import torch

from . import _lib as L


def log_normalize_u16(data, gmin, gmax):
    """data: float32 CUDA tensor (any shape) -> uint16 tensor of the same shape.
    u16(clip((log1p(max(nan_to_num(x), 0)) - gmin) / (gmax - gmin + 1e-8), 0, 1) * 65535), float32 arithmetic."""
    if data.device.type != 'cuda':
        raise RuntimeError('log_normalize_u16 (libsrb) runs on CUDA only; no CPU fallback')
    x = data.float().contiguous()
    out = torch.empty(x.shape, dtype=torch.uint16, device=x.device)
    gmin32 = np.float32(gmin)
    denom = np.float32(np.float32(gmax) - gmin32 + 1e-8)  # same float32 rounding as the numpy expression
    with torch.cuda.device(x.device):
        L.check(L.lib().srb_log_normalize_u16(L.ptr(x), L.ptr(out), x.numel(), float(gmin32), float(denom),
                                              L.current_stream()), 'log_normalize_u16')
    return out


# ---------------------------------------------------------------------------------------------------------------------
# Robust black / white points: calculate_robust_stats of scripts/Dataset_step4_normalization.py:55-91
_VALID_LOG = 1e-5            # `valid_mask = batch_flat > 1e-5` (:70)
_VALID_THRESHOLD = None


def valid_threshold():
    """Smallest float32 x with torch.log1p(x) > 1e-5 in the reference's own (CPU, float32) arithmetic: the valid-pixel
    mask of :68-70 restated on the raw value, so that the device pass needs no log1p (log1p is monotone)."""
    global _VALID_THRESHOLD
    if _VALID_THRESHOLD is None:
        lo = int(np.float32(1e-6).view(np.int32))   # log1p(lo) <= 1e-5
        hi = int(np.float32(1e-4).view(np.int32))   # log1p(hi) >  1e-5
        ok = lambda bits: bool(torch.log1p(torch.tensor(np.int32(bits).view(np.float32))) > _VALID_LOG)
        assert not ok(lo) and ok(hi)
        while hi - lo > 1:
            mid = (lo + hi) // 2
            lo, hi = (lo, mid) if ok(mid) else (mid, hi)
        _VALID_THRESHOLD = float(np.int32(hi).view(np.float32))
    return _VALID_THRESHOLD


def _percentile_from_order_statistics(n, q, fetch):
    """np.percentile(sample, q) (default 'linear' method, float32 sample of n values) given access to single order
    statistics: fetch(k) = k-th smallest as np.float32.  Follows numpy's own arithmetic step by step — q / float32(100),
    float32 virtual index (n - 1) * q, float32 gamma, _lerp's two-sided form — so the result is bit-identical."""
    q32 = np.true_divide(q, np.float32(100))
    v = np.asanyarray((n - 1) * q32)
    lo = int(np.floor(v))
    g = np.asanyarray(np.asanyarray(v - lo), dtype=v.dtype)
    a = fetch(min(max(lo, 0), n - 1))
    b = fetch(min(max(lo + 1, 0), n - 1))
    d = np.subtract(b, a)
    r = np.add(a, d * g)
    if g >= 0.5:
        r = np.subtract(b, d * (1 - g))
    return r, (min(max(lo, 0), n - 1), min(max(lo + 1, 0), n - 1))


def _needed_ranks(n, q):
    return _percentile_from_order_statistics(n, q, lambda k: np.float32(0))[1]


def robust_stats(batches, black_clip_percentile=3.0, white_percentile=99.99, use_log_stretch=True):
    """(global_min, global_max) of calculate_robust_stats (:55-91): the ``black_clip_percentile``-th and 99.99th
    percentiles of log1p(max(x, 0)) over the valid pixels (log1p > 1e-5; NaN / inf count as 0 like :47).

    The reference estimates them from an unseeded ``torch.randperm`` subsample (4000 pixels per image); this computes the
    exact population percentiles — three HBM-bound histogram passes of a radix select on the float bit pattern
    (``srb_robust_hist``), deterministic.  ``batches``: a list / tuple of float32 CUDA tensors, or a zero-argument
    callable returning a fresh iterator over such tensors (it is called three times, for data that is streamed).
    Returns two np.float32; (0.0, 1.0) when no pixel is valid (the reference's fallback, :79-81)."""
    if not use_log_stretch:
        raise NotImplementedError('libsrb implements USE_LOG_STRETCH = True, the reference setting (:21)')
    it = batches if callable(batches) else (lambda: iter(batches))
    lib = L.lib()
    thresh = valid_threshold()
    dev = None

    def run_level(level, prefixes):
        nonlocal dev
        hist, pre = None, None
        for t in it():
            if t.device.type != 'cuda':
                raise RuntimeError('robust_stats (libsrb) runs on CUDA only; no CPU fallback')
            if t.dtype != torch.float32:
                raise TypeError('robust_stats: float32 data expected (RawFitsDataset casts to float32, :48)')
            t = t.contiguous()
            if t.numel() == 0:
                continue
            with torch.cuda.device(t.device):
                if hist is None:
                    dev = t.device
                    hist = torch.zeros(max(1, len(prefixes)), 2048, dtype=torch.int64, device=dev)
                    pre = (C.c_uint32 * 4)(*prefixes)  # read by the host side of the call
                L.check(lib.srb_robust_hist(L.ptr(t), t.numel(), thresh, level, pre, len(prefixes), L.ptr(hist),
                                            L.current_stream()), 'robust_hist')
        return None if hist is None else hist.cpu().numpy()

    h0 = run_level(0, [])
    n = 0 if h0 is None else int(h0.sum())
    if n == 0:
        return np.float32(0.0), np.float32(1.0)
    ranks = sorted(set(_needed_ranks(n, black_clip_percentile) + _needed_ranks(n, white_percentile)))
    state = {r: (0, r) for r in ranks}  # rank -> (bit prefix so far, rank within that prefix)

    def descend(hists, prefixes, width):
        for r, (pre, rr) in state.items():
            h = hists[prefixes.index(pre)] if prefixes else hists[0]
            c = np.cumsum(h)
            b = int(np.searchsorted(c, rr, side='right'))
            state[r] = ((pre << width) | b, rr - (int(c[b - 1]) if b else 0))

    descend(h0, [], 11)                                   # bits[31:21] (bit 31 is 0 for valid pixels)
    for level, width in ((1, 11), (2, 10)):
        prefixes = sorted({p for p, _ in state.values()})
        descend(run_level(level, prefixes), prefixes, width)
    value = {}
    for r, (bits, _) in state.items():
        x = torch.tensor(np.uint32(bits).view(np.float32))
        value[r] = torch.log1p(torch.maximum(x, torch.tensor(0.0))).numpy()[()]   # the reference's expression (:66)
    gmin, _ = _percentile_from_order_statistics(n, black_clip_percentile, lambda k: value[k])
    gmax, _ = _percentile_from_order_statistics(n, white_percentile, lambda k: value[k])
    return np.float32(gmin), np.float32(gmax)
