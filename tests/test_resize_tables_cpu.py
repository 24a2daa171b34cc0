"""Host-side tap tables of the bicubic resize branch (src/architecture.py:80-81): the (index, weight) tables that
srb_resample_separable consumes, expanded to dense matrices here, must reproduce ATen's upsample_bicubic2d, and the
transposed tables must be the exact transpose (they drive the backward).  No kernel is called."""
import pytest
import torch
import torch.nn.functional as F

from superresolution_b200.ops import _cubic_taps, _transpose_taps


def _dense(idx, w, n_cols):
    m = torch.zeros(idx.shape[0], n_cols, dtype=torch.float64)
    for k in range(idx.shape[1]):
        m.index_put_((torch.arange(idx.shape[0]), idx[:, k].long()), w[:, k].double(), accumulate=True)
    return m


@pytest.mark.parametrize('n_in,n_out', [(128, 192), (64, 512), (96, 37), (512, 512), (8, 5)])
def test_cubic_tables_reproduce_aten(n_in, n_out):
    idx, w = _cubic_taps(n_in, n_out)
    assert idx.shape == (n_out, 4) and idx.min() >= 0 and idx.max() < n_in
    assert torch.allclose(w.sum(1), torch.ones(n_out), atol=1e-6)
    M = _dense(idx, w, n_in)
    g = torch.Generator().manual_seed(n_in * 1000 + n_out)
    x = torch.rand(1, 1, n_in, n_in, generator=g)
    ref = F.interpolate(x, size=(n_out, n_out), mode='bicubic', align_corners=False)[0, 0].double()
    got = M @ x[0, 0].double() @ M.T
    assert (got - ref).abs().max().item() <= 2e-6


@pytest.mark.parametrize('n_in,n_out', [(128, 192), (64, 512), (96, 37)])
def test_transposed_tables_are_the_transpose(n_in, n_out):
    idx, w = _cubic_taps(n_in, n_out)
    it, wt = _transpose_taps(idx, w, n_in)
    assert it.shape[0] == n_in
    assert torch.equal(_dense(it, wt, n_out), _dense(idx, w, n_in).T)


# ---- host arithmetic of normalization.robust_stats: no kernel call ---------------------------------------
def test_percentile_from_order_statistics_is_bit_identical_to_numpy():
    import numpy as np
    from superresolution_b200.normalization import _percentile_from_order_statistics
    rng = np.random.default_rng(0)
    for _ in range(400):
        n = int(rng.integers(1, 4000))
        arr = np.exp(rng.normal(size=n)).astype(np.float32)
        q = float(rng.choice([3.0, 99.99, 5.0, 15.0, 50.0, 0.0, 100.0]))
        s = np.sort(arr)
        got, ranks = _percentile_from_order_statistics(n, q, lambda k: s[k])
        ref = np.percentile(arr, q)
        assert got == ref and got.dtype == ref.dtype, (n, q, got, ref)
        assert 0 <= ranks[0] <= ranks[1] < n


def test_valid_threshold_is_the_edge_of_the_reference_mask():
    import numpy as np
    from superresolution_b200.normalization import valid_threshold
    t = np.float32(valid_threshold())
    below = np.nextafter(t, np.float32(0))
    assert bool(torch.log1p(torch.tensor(t)) > 1e-5) and not bool(torch.log1p(torch.tensor(below)) > 1e-5)


def test_oracle_robust_stats_matches_reference_formula_on_a_tiny_case():
    import numpy as np
    from oracle import aux
    x = np.array([[0.0, np.nan, -1.0, np.inf], [1.0, 2.0, 3.0, 4.0]], dtype=np.float32)
    lo, hi = aux.robust_stats([x], 3.0)
    v = np.log1p(np.array([1, 2, 3, 4], dtype=np.float32))
    assert lo == np.percentile(v, 3.0) and hi == np.percentile(v, 99.99)
    assert aux.robust_stats([np.zeros((2, 2), np.float32)]) == (0.0, 1.0)


# ---- 16-bit TIFF container (host code, no kernel) -----------------------------------------------------------
@pytest.mark.parametrize('shape', [(37, 53), (128, 128), (1, 7)])
def test_tiff16_codec_round_trips_with_pil(tmp_path, shape):
    import numpy as np
    from PIL import Image
    from superresolution_b200 import io16
    a = np.random.default_rng(shape[0]).integers(0, 65536, shape, dtype=np.uint16)
    io16.write_tiff16(a, tmp_path / 'ours.tiff')
    pil = Image.open(tmp_path / 'ours.tiff')
    assert pil.mode == 'I;16' and np.array_equal(np.array(pil), a)          # what src/dataset.py:28-31 would read
    Image.fromarray(a).save(tmp_path / 'pil.tiff')                          # what Modello_5.save_as_tiff16 writes
    assert np.array_equal(io16.read_tiff16(tmp_path / 'pil.tiff'), a)
    assert np.array_equal(io16.read_tiff16(tmp_path / 'ours.tiff'), a)


def test_antialiased_bicubic_tables_match_aten():
    """the stale variant resizes with antialias=True (Backup/src/architecture.py:91-92); the tap tables fed to
    srb_resample_separable reproduce ATen's anti-aliased bicubic per axis (checked densely on the CPU, 1e-5)."""
    import torch
    import torch.nn.functional as F
    from superresolution_b200 import ops
    torch.manual_seed(0)
    for n_in, n_out in ((128, 48), (512, 192), (96, 37), (64, 512), (100, 100), (53, 128)):
        idx, w = ops._cubic_taps_aa(n_in, n_out)
        assert idx.min() >= 0 and idx.max() < n_in and torch.allclose(w.sum(1), torch.ones(n_out), atol=1e-6)
        x = torch.rand(1, 1, 1, n_in)
        ref = F.interpolate(x, size=(1, n_out), mode='bicubic', align_corners=False, antialias=True)[0, 0, 0]
        got = (x[0, 0, 0][idx.long()] * w).sum(1)
        assert (got - ref).abs().max().item() <= 1e-5, (n_in, n_out)
