"""Error convention of the drop-in boundary: every C-ABI entry point returns a negative code and leaves a
message in srb_last_error(); the Python wrappers raise; nothing falls back to the CPU or silently changes the request."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def test_model_refuses_cpu_tensors_bad_ranks_and_bad_geometry(cuda):
    from superresolution_b200.architecture import HybridSuperResolutionModel
    from superresolution_b200.training import hybrid_forward
    m = HybridSuperResolutionModel(output_size=128, smoothing='balanced').to(cuda).eval()
    with pytest.raises(RuntimeError, match='CUDA only'):
        m(torch.rand(1, 1, 32, 32))
    with pytest.raises(ValueError, match=r'\[B,1,H,W\]'):
        m(torch.rand(1, 3, 32, 32, device=cuda))
    with pytest.raises(RuntimeError, match='libsrb forward failed'):
        m(torch.rand(1, 1, 42, 40, device=cuda))            # H, W must be multiples of 4 (pixel_unshuffle(2) + 8x8 windows)
    m.train()
    with pytest.raises(ValueError, match='multiple of 32'):
        hybrid_forward(m, torch.rand(1, 1, 32, 32, device=cuda))   # training path: W % 64
    with pytest.raises(RuntimeError, match='CUDA only'):
        hybrid_forward(m, torch.rand(1, 1, 32, 64))


def test_operator_wrappers_raise_with_the_library_message(cuda):
    from superresolution_b200 import _lib as L
    from superresolution_b200 import ops
    x = torch.zeros(1, 16, 32, 8, dtype=torch.bfloat16, device=cuda)
    w = ops.pack_conv_weight(torch.zeros(32, 8, 3, 3, device=cuda))
    with pytest.raises(RuntimeError, match='Cin=8'):
        ops.conv_nhwc(x, 8, w, None, out=torch.empty(1, 16, 32, 32, dtype=torch.bfloat16, device=cuda))
    assert 'Cin=8' in L.lib().srb_last_error().decode()
    qkv = torch.zeros(1, 12, 16, 288, dtype=torch.bfloat16, device=cuda)
    with pytest.raises(RuntimeError, match='multiples of 8'):
        ops.window_attention(qkv, torch.zeros(6, 64, 64, device=cuda), 0)
    with pytest.raises(RuntimeError, match='shift'):
        ops.window_attention(torch.zeros(1, 16, 16, 288, dtype=torch.bfloat16, device=cuda),
                             torch.zeros(6, 64, 64, device=cuda), 3)
    with pytest.raises(TypeError, match='fp32'):
        ops.blur(torch.zeros(1, 1, 8, 8, dtype=torch.float16, device=cuda), 'light')
    # training epilogue exists for the shapes that use it only: anything else is refused, not silently un-fused
    x96 = torch.zeros(1, 16, 32, 96, dtype=torch.bfloat16, device=cuda)
    w64 = ops.pack_conv_weight(torch.zeros(64, 96, 1, 1, device=cuda))
    with pytest.raises(RuntimeError, match='pre_out'):
        ops.conv_nhwc_train(x96, 96, w64, None, pre_out=torch.empty(1, 16, 32, 64, dtype=torch.bfloat16, device=cuda))
    with pytest.raises(RuntimeError, match='pitches'):  # a 32-channel buffer cannot hold 64 input channels
        ops.conv_wgrad(torch.zeros(1, 16, 32, 32, dtype=torch.bfloat16, device=cuda), 64,
                       torch.zeros(1, 16, 32, 32, dtype=torch.bfloat16, device=cuda), 32, 9)
    with pytest.raises(RuntimeError, match='wgrad'):
        ops.conv_wgrad(torch.zeros(1, 16, 24, 32, dtype=torch.bfloat16, device=cuda), 32,
                       torch.zeros(1, 16, 24, 32, dtype=torch.bfloat16, device=cuda), 32, 9)   # W % 32


def test_aux_paths_refuse_cpu_and_wrong_dtypes(cuda):
    from superresolution_b200 import io16, normalization, tiling
    from superresolution_b200.architecture import HybridSuperResolutionModel
    with pytest.raises(RuntimeError, match='CUDA only'):
        normalization.robust_stats([torch.rand(4, 4)])
    with pytest.raises(TypeError, match='float32'):
        normalization.robust_stats([torch.rand(4, 4, dtype=torch.float64, device=cuda)])
    with pytest.raises(RuntimeError, match='CUDA only'):
        normalization.log_normalize_u16(torch.rand(4, 4), 0.0, 1.0)
    with pytest.raises(RuntimeError, match='CUDA only'):
        io16.to_u16(torch.rand(4, 4))
    m = HybridSuperResolutionModel(output_size=128).to(cuda).eval()
    with pytest.raises(ValueError, match='seams'):
        tiling.infer_frame(m, torch.zeros(64, 64, dtype=torch.uint16, device=cuda), 32, 24, seams='allgather')
    with pytest.raises(RuntimeError, match='CUDA only'):
        tiling.infer_frame(m, torch.zeros(64, 64), 32, 24)
