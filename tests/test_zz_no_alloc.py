"""srb_psnr_ssim_noalloc (include/srb.h): the allocation-free variant of srb_psnr_ssim, whose scratch is a static
__device__ array inside libsrb.so.  No reference counterpart (added for callers that may not call cudaMalloc); the
checker is the existing entry point srb_psnr_ssim on the same inputs.  Both run the same kernels in the same order on
a scratch of the same layout, so the results are compared bit for bit (tighter than the 1e-4 PSNR/SSIM bar of
tests/test_aux_gpu.py, which is also asserted against the oracle once).  Collected last on purpose (file name)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _err(L):
    return L.lib().srb_last_error().decode('utf-8', 'replace')


def _pair(B, H, W, seed, dev):
    g = torch.Generator().manual_seed(seed)
    t = torch.rand(B, 1, H, W, generator=g)
    p = t + 0.02 * torch.randn(B, 1, H, W, generator=g)  # leaves [0,1] here and there: exercises the clamp
    return p.to(dev).contiguous(), t.to(dev).contiguous()


def _with_scratch(L, p, t):
    B, _, H, W = p.shape
    lib = L.lib()
    out = torch.full((2,), float('nan'), dtype=torch.float32, device=p.device)
    scratch = torch.empty(lib.srb_psnr_ssim_scratch_bytes(B, H, W), dtype=torch.uint8, device=p.device)
    L.check(lib.srb_psnr_ssim(L.ptr(p), L.ptr(t), B, H, W, L.ptr(out), L.ptr(scratch), L.current_stream()), 'psnr_ssim')
    return out


def _no_alloc(L, p, t, out):
    B, _, H, W = p.shape
    return L.lib().srb_psnr_ssim_noalloc(L.ptr(p), L.ptr(t), B, H, W, L.ptr(out), L.current_stream())


def test_no_alloc_entry_point_matches_the_scratch_one_twice_in_a_row(cuda):
    from oracle import aux
    from superresolution_b200 import _lib as L
    # a large call first, then a smaller one of another shape: the second finds the first one's partials in the array
    cases = [_pair(6, 512, 512, 71, cuda), _pair(2, 75, 139, 72, cuda), _pair(6, 512, 512, 73, cuda)]
    refs = [_with_scratch(L, p, t) for p, t in cases]
    outs = [torch.empty(2, dtype=torch.float32, device=cuda) for _ in cases]
    for rnd in range(2):  # round 0 also pays the one-time symbol look-up; round 1 must leave the device's free memory alone
        torch.cuda.synchronize()
        free0, alloc0 = torch.cuda.mem_get_info(cuda)[0], torch.cuda.memory_allocated(cuda)
        for (p, t), out in zip(cases, outs):  # back to back on one stream, no sync in between
            out.fill_(float('nan'))
            assert _no_alloc(L, p, t, out) == 0, _err(L)
        torch.cuda.synchronize()
        assert torch.cuda.memory_allocated(cuda) == alloc0
        if rnd:
            assert free0 - torch.cuda.mem_get_info(cuda)[0] < (1 << 20)
    for out, ref in zip(outs, refs):
        assert torch.isfinite(out).all()
        assert torch.equal(out, ref), (out.tolist(), ref.tolist())
    # and against the oracle (src/metrics.py:5-32 restated), at the bar of tests/test_aux_gpu.py
    p, t = cases[1]
    m = aux.Metrics()
    m.update(p.cpu(), t.cpu())
    r = m.compute()
    B = p.shape[0]
    assert abs(outs[1][0].item() / B - r['psnr']) < 1e-4 and abs(outs[1][1].item() / B - r['ssim']) < 1e-4


def test_no_alloc_entry_point_rejects_sizes_beyond_the_static_array(cuda):
    from superresolution_b200 import _lib as L
    lib = L.lib()
    limit = 256 * 1024  # SRB_PSNR_SSIM_NOALLOC_MAX_BYTES
    assert lib.srb_psnr_ssim_scratch_bytes(252, 512, 512) <= limit < lib.srb_psnr_ssim_scratch_bytes(253, 512, 512)
    p, t = _pair(1, 64, 64, 74, cuda)
    out = torch.zeros(2, dtype=torch.float32, device=cuda)
    # shape check happens before any launch: the (too small) buffers are never touched
    rc = lib.srb_psnr_ssim_noalloc(L.ptr(p), L.ptr(t), 253, 512, 512, L.ptr(out), L.current_stream())
    assert rc == -3, rc  # SRB_E_WORKSPACE
    assert 'static array' in _err(L)
    assert lib.srb_psnr_ssim_noalloc(L.ptr(p), L.ptr(t), 0, 64, 64, L.ptr(out), L.current_stream()) == -5
    assert lib.srb_psnr_ssim_noalloc(None, L.ptr(t), 1, 64, 64, L.ptr(out), L.current_stream()) == -5
    torch.cuda.synchronize()
    assert out.tolist() == [0.0, 0.0]
    # still usable afterwards
    assert _no_alloc(L, p, t, out) == 0
    assert torch.equal(out, _with_scratch(L, p, t))
