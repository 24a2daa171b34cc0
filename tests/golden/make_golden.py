"""Mints the committed golden fixtures (run in the build container, where /root/reference exists):

  python tests/golden/make_golden.py

* model.npz  — outputs of the fp32 CPU oracle (oracle/) for seeded weights + inputs: RRDBNet (stage 1), HAT (stage 2),
  full hybrid ('balanced' and 'none').  The reference model itself cannot be imported here (BasicSR / HAT are not
  vendored), so these pin the ORACLE against drift; parity with upstream stays "unpinned".
* aux.npz    — outputs of the REFERENCE's own src/losses.py, src/metrics.py (imported from /root/reference) and of
  the normalisation lines of scripts/Dataset_step4_normalization.py:167-179 (exec'd from the reference file), on
  seeded inputs.  The oracle restatements (oracle/aux.py) are asserted equal to them here and again in the CPU tests.
* backup_loss.npz — the Charbonnier and astro terms of the stale Backup/src/losses.py on the same seeded inputs (f4).
"""
import hashlib
import importlib.util
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
REF = '/root/reference'


def load_ref(rel, name):
    spec = importlib.util.spec_from_file_location(name, os.path.join(REF, rel))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def model_inputs():
    g = torch.Generator().manual_seed(1)
    x32 = torch.randint(0, 65536, (1, 1, 32, 32), generator=g).float() / 65535.0
    g = torch.Generator().manual_seed(2)
    x64 = torch.randint(0, 65536, (1, 1, 64, 64), generator=g).float() / 65535.0
    return x32, x64


def aux_inputs():
    g = torch.Generator().manual_seed(11)
    t = torch.rand(2, 1, 96, 80, generator=g) * 0.03           # mostly dark, some pixels above the 0.02 star threshold
    t[0, 0, 10:14, 20:24] = 0.9
    p = (t + 0.05 * torch.randn(2, 1, 96, 80, generator=g))
    rng = np.random.default_rng(3)
    raw = np.expm1(rng.normal(1.0, 1.2, size=(64, 48))).astype(np.float32)  # heavy-tailed positive "counts"
    raw[rng.random(raw.shape) < 0.01] *= -1.0
    raw[3, 5] = np.nan
    raw[7, 9] = np.inf
    raw[8, 9] = -np.inf
    raw[0, 0] = 0.0
    return p, t, raw


def main():
    from oracle.hybrid import HybridSuperResolutionModel as Oracle
    from oracle import aux
    torch.manual_seed(0)
    o = Oracle(output_size=128, smoothing='balanced').eval()
    sd = o.state_dict()
    h = hashlib.sha256()
    for k in sorted(sd):
        h.update(sd[k].numpy().tobytes())
    x32, x64 = model_inputs()
    with torch.no_grad():
        s1 = o.stage1(x32)
        s2 = o.stage2(x64)
        full = o(x32)
        o.s1 = o.s2 = o.sf = torch.nn.Identity()
        full_none = o(x32)
    np.savez_compressed(os.path.join(HERE, 'model.npz'), stage1=s1.numpy(), stage2=s2.numpy(), full=full.numpy(),
                        full_none=full_none.numpy(), weights_sha256=np.frombuffer(h.digest(), dtype=np.uint8))

    # ---- reference's own loss / metrics ----
    ref_losses = load_ref('src/losses.py', 'ref_losses')
    ref_metrics = load_ref('src/metrics.py', 'ref_metrics')
    p, t, raw = aux_inputs()
    loss, _ = ref_losses.CombinedLoss()(p, t)
    m = ref_metrics.Metrics()
    m.update(p, t)
    mm = m.compute()
    assert torch.equal(loss, aux.combined_loss(p, t))
    om = aux.Metrics()
    om.update(p, t)
    assert om.compute() == mm
    # ---- reference normalisation lines, exec'd verbatim from the script ----
    lines = open(os.path.join(REF, 'scripts/Dataset_step4_normalization.py')).read().split('\n')
    body = '\n'.join(l.strip() for l in lines[166:179] if l.strip() and not l.strip().startswith('#'))
    body = body.replace('if USE_LOG_STRETCH:', 'if True:\n    pass')
    gmin, gmax = np.float32(0.08), np.float32(5.5)
    env = {'np': np, 'd_h': np.nan_to_num(raw.copy()), 'd_o': np.nan_to_num(raw.copy()[:32]), 'h_min': gmin,
           'h_max': gmax, 'o_min': gmin, 'o_max': gmax}
    exec(body, env)
    ref_u16 = env['h_u16']
    assert np.array_equal(ref_u16, aux.log_normalize_u16(raw, gmin, gmax)), 'oracle log_normalize != reference'
    np.savez_compressed(os.path.join(HERE, 'aux.npz'), loss=loss.numpy(), psnr=np.float64(mm['psnr']),
                        ssim=np.float64(mm['ssim']), lognorm_u16=ref_u16, gmin=gmin, gmax=gmax)
    # ---- stale-variant loss terms: Backup/src/losses.py CharbonnierLoss imported, the astro lines exec'd ----
    bl = load_ref('Backup/src/losses.py', 'ref_backup_losses')   # (needs torchvision importable; no weights are fetched)
    char = bl.CharbonnierLoss()(p, t)
    lines = open(os.path.join(REF, 'Backup/src/losses.py')).read().split('\n')
    body = '\n'.join(l.strip() for l in lines[45:48])            # diff / weight_map / astro_loss (:46-48)
    env = {'torch': torch, 'pred': p, 'target': t}
    exec(body, env)
    oc, ol1, oa = aux.astro_loss_terms(p, t)
    assert torch.equal(char, oc) and torch.equal(env['astro_loss'], oa)
    np.savez_compressed(os.path.join(HERE, 'backup_loss.npz'), char=char.numpy(), astro=env['astro_loss'].numpy(),
                        l1_raw=ol1.numpy())
    print('golden written: stage1', s1.shape, 'stage2', s2.shape, 'full', full.shape, 'loss', float(loss), mm)


if __name__ == '__main__':
    main()
