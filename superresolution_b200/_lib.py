"""ctypes binding of libsrb.so (include/srb.h).  Fails loudly: there is no CPU / PyTorch fallback."""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, 'libsrb.so')

SRB_F32, SRB_BF16, SRB_U16 = 0, 1, 2
SMOOTHING = {'none': 0, 'light': 1, 'balanced': 2, 'strong': 3}
ACT_NONE, ACT_LRELU, ACT_GELU = 0, 1, 2

_lib = None

_vp, _i, _f, _sz = C.c_void_p, C.c_int, C.c_float, C.c_size_t

_PROTOS = {
    'srb_version': (C.c_int, []),
    'srb_last_error': (C.c_char_p, []),
    'srb_pack_conv_weight': (_i, [_vp, _vp, _i, _i, _i, _vp]),
    'srb_param_count': (_i, []),
    'srb_param_name': (C.c_char_p, [_i]),
    'srb_param_numel': (C.c_int64, [_i]),
    'srb_packed_bytes': (_sz, []),
    'srb_pack_weights': (_i, [_vp, _i, _vp, _sz, _vp]),
    'srb_workspace_bytes': (_sz, [_i, _i, _i]),
    'srb_forward': (_i, [_vp, _i, _vp, _i, _i, _i, _i, _vp, _vp, _sz, _i, _i, _vp]),
    'srb_forward_profile': (_i, [_vp, _i, _vp, _i, _i, _i, _i, _vp, _vp, _sz, _i, _i, _vp, _vp, _vp, _vp]),
    'srb_forward_launch_count': (_i, [_i, _i, _i, _i, _i]),
    'srb_loss_scratch_bytes': (_sz, []),
    'srb_weighted_l1': (_i, [_vp, _vp, _sz, _vp, _vp, _vp, _vp]),
    'srb_charbonnier_astro': (_i, [_vp, _vp, _sz, C.c_float, C.c_float, _vp, _vp, _vp, _vp]),
    'srb_psnr_ssim_scratch_bytes': (_sz, [_i, _i, _i]),
    'srb_psnr_ssim': (_i, [_vp, _vp, _i, _i, _i, _vp, _vp, _vp]),
    'srb_psnr_ssim_noalloc': (_i, [_vp, _vp, _i, _i, _i, _vp, _vp]),
    'srb_log_normalize_u16': (_i, [_vp, _vp, _sz, _f, _f, _vp]),
    'srb_pack_conv_weight_dgrad': (_i, [_vp, _vp, _i, _i, _i, _vp]),
    'srb_conv_wgrad_scratch_bytes': (_sz, [_i] * 6),
    'srb_conv_wgrad': (_i, [_vp, _i, _vp, _i, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _vp]),
    'srb_upsample2x_nearest': (_i, [_vp, _i, _vp, _i, _i, _i, _vp]),
    'srb_conv3x3_cout1': (_i, [_vp, _i, _vp, _vp, _vp, _i, _i, _i, _vp]),
    'srb_to_u16': (_i, [_vp, _vp, _sz, _vp]),
    'srb_preview_strip': (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _vp]),
    'srb_robust_hist': (_i, [_vp, _sz, _f, _i, _vp, _i, _vp, _vp]),
    'srb_blur': (_i, [_vp, _vp, _i, _i, _i, _i, _i, _vp]),
    'srb_resample_separable': (_i, [_vp, _vp, _i, _i, _i, _i, _i, _vp, _vp, _vp, _vp, _i, _vp]),
    'srb_lrelu_bwd': (_i, [_vp, _i, _vp, _i, _sz, _i, _f, _vp]),
    'srb_axpby': (_i, [_vp, _i, _i, _f, _vp, _i, _i, _f, _vp, _i, _i, _sz, _i, _vp]),
    'srb_axpby_samples': (_i, [_vp, _i, _i, _vp, _sz, _vp, _i, _i, _f, _vp, _i, _i, _sz, _i, _vp]),
    'srb_upsample2x_bwd': (_i, [_vp, _i, _vp, _i, _i, _i, _i, _i, _vp]),
    'srb_layernorm96_bwd_scratch_bytes': (_sz, []),
    'srb_layernorm96': (_i, [_vp, _vp, _vp, _vp, _sz, _vp]),
    'srb_layernorm96_bwd': (_i, [_vp, _vp, _vp, _vp, _i, _vp, _vp, _vp, _vp, _sz, _i, _vp]),
    'srb_gelu': (_i, [_vp, _i, _vp, _i, _vp, _i, _sz, _i, _vp]),
    'srb_window_attention': (_i, [_vp, _vp, _vp, _i, _i, _i, _i, _vp]),
    'srb_window_attention_bwd': (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _vp]),
    'srb_overlap_attention_bwd_scratch_bytes': (_sz, [_i, _i, _i]),
    'srb_overlap_attention': (_i, [_vp, _vp, _vp, _i, _i, _i, _vp]),
    'srb_overlap_attention_infer': (_i, [_vp, _vp, _vp, _i, _i, _i, _vp]),
    'srb_overlap_attention_bwd': (_i, [_vp] * 7 + [_i, _i, _i, _vp]),
    'srb_channel_attention_scratch_bytes': (_sz, [_i, _i]),
    'srb_channel_attention': (_i, [_vp] * 10 + [_i, _i, _vp]),
    'srb_channel_attention_bwd': (_i, [_vp] * 14 + [_i, _i, _i, _f, _vp]),
    'srb_grad_norm_scratch_bytes': (_sz, []),
    'srb_grad_sumsq': (_i, [_vp, _sz, _vp, _vp, _vp]),
    'srb_adamw_step': (_i, [_vp, _vp, _vp, _vp, _sz, _f, _f, _f, _f, _f, _i, _f, _f, _vp, _vp, _vp]),
    'srb_blend_tiles': (_i, [_vp, _vp, _vp, _i, _i, _i, _i, _vp, _vp, _i, _i, _i, _i, _vp]),
    'srb_blend_finalize': (_i, [_vp, _vp, _vp, _i, _sz, _vp]),
    'srb_gather_tiles': (_i, [_vp, _i, _i, _i, _vp, _vp, _i, _i, _vp, _vp]),
    'srb_expand_rel_bias': (_i, [_vp, _i, _vp, _i, _vp, _f, _vp]),
    'srb_hab_attention': (_i, [_vp] * 10 + [_i, _i, _i, _i, _vp]),
    'srb_mlp_branch': (_i, [_vp] * 8 + [_sz, _sz, _vp]),
    'srb_mlp_branch_ln': (_i, [_vp] * 11 + [_sz, _sz, _vp]),
    'srb_fold_ln_linear': (_i, [_vp] * 6 + [_i, _i, _vp]),
    'srb_conv_nhwc_bf16_train': (_i, [_vp, _i, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _i, _f, _vp, _i, _i, _vp, _i,
                                      _vp, _i, _vp, _vp, _i, _vp, _vp]),
    'srb_conv_nhwc_bf16': (_i, [_vp, _i, _vp, _vp, _vp, _i, _vp, _i, _i, _i, _i, _i, _i, _i, _i, _f, _f, _vp, _i, _i,
                                _f, _vp, _i, _i, _vp, _i, _vp]),
}


def register(name, restype, argtypes):
    _PROTOS[name] = (restype, argtypes)
    if _lib is not None:
        fn = getattr(_lib, name)
        fn.restype, fn.argtypes = restype, argtypes


def lib():
    """Load libsrb.so (built by __graft_entry__.build()).  Raises if it is missing: no fallback path exists."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(f'{LIB_PATH} not built: run `python -c "import __graft_entry__ as g; g.build()"` '
                               f'(or `make -C superresolution_b200/csrc`). There is no CPU fallback.')
        _lib = C.CDLL(LIB_PATH)
        for name, (restype, argtypes) in _PROTOS.items():
            fn = getattr(_lib, name)
            fn.restype, fn.argtypes = restype, argtypes
    return _lib


def check(rc, what=''):
    if rc != 0:
        msg = lib().srb_last_error().decode('utf-8', 'replace')
        raise RuntimeError(f'libsrb {what} failed (code {rc}): {msg}')


def ptr(t):
    """device pointer of a torch tensor (or None)."""
    return None if t is None else C.c_void_p(t.data_ptr())


def current_stream():
    import torch
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)
