"""The `src` shim (shim/src): `from src.architecture import HybridSuperResolutionModel` (reference
scripts/Modello_5.py:23, scripts/Modello_supporto.py:21) resolves to the drop-in without BasicSR / HAT installed."""
import importlib
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_src_shim_resolves_to_dropin():
    sys.path.insert(0, os.path.join(ROOT, 'shim'))
    try:
        for k in [k for k in sys.modules if k == 'src' or k.startswith('src.')]:
            del sys.modules[k]
        arch = importlib.import_module('src.architecture')
        losses = importlib.import_module('src.losses')
        metrics = importlib.import_module('src.metrics')
        import superresolution_b200.architecture as A
        assert arch.HybridSuperResolutionModel is A.HybridSuperResolutionModel
        assert arch.AntiCheckerboardLayer is A.AntiCheckerboardLayer
        assert hasattr(losses, 'CombinedLoss') and hasattr(metrics, 'Metrics') and hasattr(metrics, 'ssim_torch')
        m = arch.HybridSuperResolutionModel(output_size=512, smoothing='none', device='cpu')  # the call of Modello_supporto.py:101
        assert 'stage1.conv_first.weight' in m.state_dict() and len(m.state_dict()) == 1282
    finally:
        sys.path.remove(os.path.join(ROOT, 'shim'))
        for k in [k for k in sys.modules if k == 'src' or k.startswith('src.')]:
            del sys.modules[k]
