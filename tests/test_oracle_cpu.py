"""CPU suite (-m "not gpu"): the oracle against the committed golden fixtures and the structural pins of the
reference (scripts/Modello_4.py:46); drop-in key parity of the product module; C-ABI exports."""
import hashlib
import os
import re

import numpy as np
import pytest
import torch

from _inputs import aux_inputs, model_inputs

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)


@pytest.fixture(scope='module')
def oracle():
    from oracle.hybrid import HybridSuperResolutionModel as Oracle
    torch.manual_seed(0)
    return Oracle(output_size=128, smoothing='balanced').eval()


def test_state_dict_schema(oracle):
    sd = oracle.state_dict()
    assert 'stage1.conv_first.weight' in sd  # the only key the reference itself pins (scripts/Modello_4.py:46)
    assert len(sd) == 1285 and len(oracle.stage1.state_dict()) == 702 and len(oracle.stage2.state_dict()) == 580
    assert sum(p.numel() for p in oracle.stage1.parameters()) == 16697409
    assert sum(p.numel() for p in oracle.stage2.parameters()) == 4101585
    assert tuple(sd['stage1.conv_first.weight'].shape) == (64, 4, 3, 3)
    assert tuple(sd['stage2.layers.0.residual_group.blocks.0.attn.relative_position_bias_table'].shape) == (225, 6)
    assert tuple(sd['stage2.layers.3.residual_group.overlap_attn.relative_position_bias_table'].shape) == (361, 6)
    sa, oca = sd['stage2.relative_position_index_SA'], sd['stage2.relative_position_index_OCA']
    assert sa.dtype == torch.int64 and tuple(sa.shape) == (64, 64) and sa.min() == 0 and sa.max() == 224
    assert tuple(oca.shape) == (64, 144) and oca.min() == -200 and oca.max() == 160
    assert len(torch.unique(oca % 361)) == 361
    assert torch.allclose(sd['s1.weight'].sum(), torch.tensor(1.0)) and tuple(sd['sf.weight'].shape) == (1, 1, 3, 3)


def test_smoothing_none_has_no_blur_keys():
    from oracle.hybrid import HybridSuperResolutionModel as Oracle
    o = Oracle(smoothing='none')
    assert len(o.state_dict()) == 1282 and not any(k.startswith(('s1.', 's2.', 'sf.')) for k in o.state_dict())


def test_oracle_matches_golden(oracle):
    g = np.load(os.path.join(HERE, 'golden', 'model.npz'))
    h = hashlib.sha256()
    sd = oracle.state_dict()
    for k in sorted(sd):
        h.update(sd[k].numpy().tobytes())
    assert np.array_equal(np.frombuffer(h.digest(), dtype=np.uint8), g['weights_sha256']), 'seeded init drifted'
    x32, x64 = model_inputs()
    with torch.no_grad():
        assert np.allclose(oracle.stage1(x32).numpy(), g['stage1'], atol=2e-5)
        assert np.allclose(oracle(x32).numpy(), g['full'], atol=2e-5)


def test_blur_impulse_is_kernel():
    from oracle.hybrid import AntiCheckerboardLayer
    for mode, k, s in (('light', 3, 16.0), ('balanced', 5, 256.0), ('strong', 7, 1600.0)):
        layer = AntiCheckerboardLayer(mode)
        x = torch.zeros(1, 1, 15, 15)
        x[0, 0, 7, 7] = 1.0
        y = layer(x)
        r = k // 2
        assert torch.equal(y[0, 0, 7 - r:8 + r, 7 - r:8 + r], layer.weight[0, 0].flip(0, 1))
        assert abs(float(layer.weight.sum()) - {3: 1.0, 5: 1.0, 7: 4096.0 / 1600.0}[k]) < 1e-6


def test_aux_oracle_matches_reference_goldens():
    from oracle import aux
    g = np.load(os.path.join(HERE, 'golden', 'aux.npz'))
    p, t, raw = aux_inputs()
    assert float(aux.combined_loss(p, t)) == float(g['loss'])
    m = aux.Metrics()
    m.update(p, t)
    r = m.compute()
    assert abs(r['psnr'] - float(g['psnr'])) < 1e-5 and abs(r['ssim'] - float(g['ssim'])) < 1e-6
    assert np.array_equal(aux.log_normalize_u16(raw, g['gmin'], g['gmax']), g['lognorm_u16'])


def test_aux_known_answers():
    from oracle import aux
    x = torch.rand(2, 1, 32, 32)
    m = aux.Metrics()
    m.update(x, x)
    r = m.compute()
    assert abs(r['ssim'] - 1.0) < 1e-6 and abs(r['psnr'] - 80.0) < 1e-3  # mse = 0 -> 10*log10(1/1e-8)
    m = aux.Metrics()
    m.update(torch.full((1, 1, 16, 16), 0.5), torch.full((1, 1, 16, 16), 0.6))
    assert abs(m.compute()['psnr'] - 20.0) < 1e-3  # -20*log10(0.1)
    t = torch.zeros(1, 1, 8, 8)
    p = torch.full((1, 1, 8, 8), 0.25)
    assert float(aux.combined_loss(p, t)) == 0.25  # all-dark target: plain mean |p - t|
    t[0, 0, 0, 0] = 1.0
    assert abs(float(aux.combined_loss(p, t)) - (63 * 0.25 + 500 * 0.75) / 64) < 1e-5
    v = aux.log_normalize_u16(np.array([-5.0, 0.0, np.nan, np.e - 1, 1e9], np.float32), 0.0, 1.0)
    assert list(v) == [0, 0, 0, 65535, 65535] or list(v) == [0, 0, 0, 65534, 65535]
    assert aux.log_normalize_u16(np.zeros((0,), np.float32), 0.0, 1.0).shape == (0,)


def test_product_module_is_a_drop_in():
    """same keys / shapes / dtypes / order as the oracle (= reference layout), strict load both ways."""
    from oracle.hybrid import HybridSuperResolutionModel as Oracle
    from superresolution_b200.architecture import HybridSuperResolutionModel as Model
    for smoothing in ('balanced', 'none', 'strong'):
        o, m = Oracle(smoothing=smoothing), Model(smoothing=smoothing)
        a, b = o.state_dict(), m.state_dict()
        assert list(a) == list(b)
        assert all(a[k].shape == b[k].shape and a[k].dtype == b[k].dtype for k in a)
        m.load_state_dict(a, strict=True)
        o.load_state_dict(m.state_dict(), strict=True)
        assert m.output_size == 512 and m.has_stage2
    with pytest.raises(RuntimeError):
        Model()(torch.zeros(1, 1, 32, 32))  # CPU tensor: no fallback path, must fail loudly


def test_abi_exports_every_declared_symbol():
    import ctypes
    hdr = open(os.path.join(ROOT, 'include', 'srb.h')).read()
    hdr = re.sub(r'/\*.*?\*/', '', hdr, flags=re.S)
    declared = set(re.findall(r'\b(srb_[a-z0-9_]+)\s*\(', hdr))
    assert len(declared) >= 15
    path = os.path.join(ROOT, 'superresolution_b200', 'libsrb.so')
    if not os.path.exists(path):
        import __graft_entry__ as g
        g.build()
    lib = ctypes.CDLL(path)
    for name in sorted(declared):
        assert hasattr(lib, name), f'{name} declared in include/srb.h but not exported'
    lib.srb_param_name.restype = ctypes.c_char_p
    names = [lib.srb_param_name(i).decode() for i in range(lib.srb_param_count())]
    from oracle.hybrid import HybridSuperResolutionModel as Oracle
    sd = Oracle(smoothing='none').state_dict()
    assert names == list(sd)  # the C parameter table is the state-dict, in order
    lib.srb_forward_launch_count.restype = ctypes.c_int
    assert lib.srb_forward_launch_count(1, 30, 32, 2, 3) == -1  # SRB_E_SHAPE, host-side check only


def test_abi_argument_errors_need_no_gpu():
    """Argument validation happens before any CUDA call: negative code + message, also on a machine without a GPU."""
    import ctypes as C
    from superresolution_b200 import _lib as L
    lib = L.lib()
    assert lib.srb_forward_launch_count(16, 128, 128, 2, 3) > 100
    assert lib.srb_forward_launch_count(16, 100, 128, 2, 3) > 100                              # H % 4 == 0: accepted (ragged tiles)
    assert lib.srb_forward_launch_count(16, 102, 128, 2, 3) < 0 and b'multiples of 4' in lib.srb_last_error()
    assert lib.srb_forward_launch_count(16, 128, 128, 2, 5) < 0 and b'stage_mask' in lib.srb_last_error()
    assert lib.srb_forward_launch_count(16, 128, 128, 9, 3) < 0 and b'smoothing' in lib.srb_last_error()
    rc = lib.srb_forward(None, L.SRB_F32, None, L.SRB_F32, 1, 32, 32, None, None, 0, 2, 3, None)
    assert rc < 0 and b'null' in lib.srb_last_error()
    assert lib.srb_blur(None, None, L.SRB_F32, 2, 1, 8, 8, None) < 0
    assert lib.srb_robust_hist(None, 0, C.c_float(1e-5), 0, None, 0, None, None) < 0
    assert lib.srb_workspace_bytes(16, 128, 128) > (1 << 30)


def test_backup_loss_terms_match_reference_golden():
    """the oracle restatement of the stale Charbonnier / astro terms equals the values produced by the
    reference's own Backup/src/losses.py (tests/golden/make_golden.py) on the seeded aux inputs."""
    import numpy as np
    import torch
    from _inputs import aux_inputs
    from oracle import aux
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden', 'backup_loss.npz'))
    p, t, _ = aux_inputs()
    char, l1_raw, astro = aux.astro_loss_terms(p, t)
    assert char.item() == float(g['char']) and astro.item() == float(g['astro']) and l1_raw.item() == float(g['l1_raw'])
