"""superresolution_b200.pipeline.infer_host_batches: host batches in, host results out, copies overlapped with the
forwards on side streams (the loop of scripts/Modello_5.py:35-91).  Same results as one forward per batch."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def test_host_batches_match_plain_forwards(cuda):
    from superresolution_b200.architecture import HybridSuperResolutionModel
    from superresolution_b200.pipeline import infer_host_batches
    torch.manual_seed(0)
    m = HybridSuperResolutionModel(output_size=128, smoothing='balanced').to(cuda).eval()
    g = torch.Generator().manual_seed(3)
    xs = [torch.rand(2, 1, 32, 32, generator=g).pin_memory() for _ in range(5)]
    with torch.no_grad():
        want = [m(x.to(cuda)).cpu() for x in xs]
    got = infer_host_batches(m, xs)
    assert len(got) == 5 and all(t.is_pinned() and t.device.type == 'cpu' for t in got)
    for a, b in zip(got, want):
        assert torch.equal(a, b)
    # caller-provided result buffers, one of them reused for every batch: the last batch's result stays
    buf = torch.empty(2, 1, 128, 128).pin_memory()
    out = infer_host_batches(m, xs, out=[buf] * 5)
    assert out[-1] is buf and torch.equal(buf, want[-1])
    assert infer_host_batches(m, []) == []
    # eager forwards instead of graph replays, and a weight update between calls (the cached graphs must not go stale)
    eager = infer_host_batches(m, xs[:2], use_graphs=False)
    assert torch.equal(eager[0], want[0]) and torch.equal(eager[1], want[1])
    with torch.no_grad():
        m.stage1.conv_first.weight.mul_(1.5)
        want2 = m(xs[0].to(cuda)).cpu()
    assert not torch.equal(want2, want[0])
    assert torch.equal(infer_host_batches(m, xs[:1])[0], want2)
    with pytest.raises(ValueError, match='len'):
        infer_host_batches(m, xs, out=[buf])
