"""ctypes binding of libpn_b200.so (include/pn_b200.h).

There is no CPU fallback: ``lib()`` raises if the shared library is missing, and ``require_gpu()``
raises unless the current device is a B200 (sm_100).  PyTorch is used for device memory and streams
only; every compute call below goes through the C-ABI.
"""
import ctypes as C
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
# PN_B200_LIB: another build of the same library (A/B measurements of a build-time option on one box)
LIB_PATH = os.environ.get("PN_B200_LIB") or os.path.join(_HERE, "libpn_b200.so")

c_i32, c_i64, c_u64, c_f32, c_f64, c_vp, c_int = (C.c_int32, C.c_int64, C.c_uint64, C.c_float,
                                                 C.c_double, C.c_void_p, C.c_int)

# name -> (restype, argtypes); mirrors include/pn_b200.h one to one
SIGNATURES = {
    "pn_version": (c_int, []),
    "pn_last_error": (C.c_char_p, []),
    "pn_device_check": (c_int, []),
    "pn_launch_count": (c_i64, []),
    "pn_weights_fwd": (c_int, [c_vp, c_i64, c_vp, c_f32, c_int, c_vp, c_vp, c_vp, c_vp]),
    "pn_weights_bwd": (c_int, [c_vp, c_i64, c_vp, c_f32, c_int, c_vp, c_vp, c_vp, c_vp]),
    "pn_weights_table_update": (c_int, [c_vp, c_vp, c_vp, c_i64, c_f32, c_f32, c_vp]),
    "pn_token_ce_fwd": (c_int, [c_vp, c_int, c_i64, c_i32, c_vp, c_i64, c_vp, c_vp, c_vp]),
    "pn_token_ce_bwd": (c_int, [c_vp, c_int, c_i64, c_i32, c_vp, c_i64, c_vp, c_vp, c_vp, c_vp]),
    "pn_group_loss_fwd": (c_int, [c_vp, c_vp, c_i64, c_i32, c_int, c_vp, c_vp, c_vp]),
    "pn_group_loss_bwd": (c_int, [c_vp, c_i64, c_i32, c_int, c_vp, c_vp, c_vp]),
    "pn_lce_restricted_fwd": (c_int, [c_vp, c_i64, c_i32, c_vp, c_vp, c_vp]),
    "pn_lce_restricted_bwd": (c_int, [c_vp, c_i64, c_i32, c_vp, c_vp, c_vp]),
    "pn_check_success": (c_int, [c_vp, c_vp, c_i64, c_i64, c_i64, c_i32, c_i32, c_vp, c_vp]),
    "pn_pack_prompts": (c_int, [c_vp, c_i32, c_vp, c_i32, c_vp, c_i32, c_vp, c_vp, c_i64, c_vp, c_vp, c_i64, c_vp, c_vp,
                                c_i64, c_i32, c_i32, c_i32, c_vp, c_vp, c_vp, c_vp]),
    "pn_select_level_scalar": (c_int, [c_vp, c_i64, c_i32, c_int, c_f64, c_int, c_vp, c_vp, c_vp,
                                       c_vp]),
    "pn_select_level_items": (c_int, [c_vp, c_i64, c_i32, c_int, c_vp, c_int, c_vp, c_vp, c_vp,
                                      c_vp]),
    "pn_binomial_table": (c_int, [c_f64, c_f64, c_f64, c_f64, c_i32, c_vp, c_vp, c_vp, c_vp]),
    "pn_select_binomial": (c_int, [c_vp, c_i64, c_i32, c_int, c_vp, c_i32, c_vp, c_i64, c_int,
                                   c_u64, c_vp, c_vp, c_vp, c_vp, c_vp]),
    "pn_gemm_bf16": (c_int, [c_vp, c_i64, c_vp, c_i64, c_vp, c_i64, c_i64, c_i64, c_i64, c_int, c_int,
                             c_vp, c_int, c_vp]),
    "pn_gemm_bf16_wgrad": (c_int, [c_vp, c_i64, c_vp, c_i64, c_vp, c_i64, c_i64, c_i64, c_i64, c_int, c_vp]),
    "pn_gemm_bf16_relu_bwd": (c_int, [c_vp, c_i64, c_vp, c_i64, c_vp, c_i64, c_vp, c_vp, c_i64, c_vp, c_i64, c_i64,
                                      c_i64, c_vp]),
    "pn_gemm_bf16_norm_bwd": (c_int, [c_vp, c_i64, c_vp, c_i64, c_vp, c_i64, c_vp, c_vp, c_i64, c_vp, c_i64, c_i64,
                                      c_i64, c_i64, c_vp]),
    "pn_t5_norm_coef": (c_int, [c_vp, c_i32, c_vp, c_i64, c_i32, c_vp, c_vp]),
    "pn_t5_self_attention_bwd_dot": (c_int, [c_vp, c_i32, c_i32, c_i32, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp,
                                             c_vp, c_vp, c_vp]),
    "pn_t5_self_attention": (c_int, [c_vp, c_i32, c_i32, c_i32, c_vp, c_vp, c_vp]),
    "pn_t5_self_attention_lse": (c_int, [c_vp, c_i32, c_i32, c_i32, c_vp, c_vp, c_vp, c_vp]),
    "pn_t5_self_attention_bwd": (c_int, [c_vp, c_i32, c_i32, c_i32, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp,
                                         c_vp, c_vp]),
    "pn_t5_self_attention_bwd_static": (c_int, [c_vp, c_i32, c_i32, c_i32, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp,
                                                c_vp]),
    "pn_t5_train_workspace_bytes": (c_i64, [c_vp, c_i64, c_i32, c_i32]),
    "pn_t5_set_train_buffers": (c_int, [c_vp, c_vp, c_vp, c_vp]),
    "pn_t5_train_forward": (c_int, [c_vp, c_vp, c_vp, c_i64, c_i32, c_i32, c_vp, c_vp, c_i64, c_vp]),
    "pn_t5_train_backward": (c_int, [c_vp, c_vp, c_vp, c_i64, c_i32, c_i32, c_vp, c_i64, c_vp]),
    "pn_t5_train_forward_z2": (c_int, [c_vp, c_vp, c_vp, c_i64, c_i32, c_i32, c_i32, c_i32, c_vp, c_vp, c_i64, c_vp]),
    "pn_t5_train_backward_z2": (c_int, [c_vp, c_vp, c_vp, c_i32, c_i32, c_i64, c_i32, c_i32, c_vp, c_i64, c_vp]),
    "pn_t5_set_dropout": (c_int, [c_vp, c_f32, c_u64, c_i64]),
    "pn_t5_zero_grad": (c_int, [c_vp, c_vp]),
    "pn_t5_set_grad_hook": (c_int, [c_vp, c_vp, c_vp]),
    "pn_t5_adamw_step": (c_int, [c_vp, c_f32, c_f32, c_f32, c_f32, c_f32, c_i64, c_f32, c_vp]),
    "pn_t5_sgd_step": (c_int, [c_vp, c_f32, c_i64, c_vp]),
    "pn_adamw": (c_int, [c_vp, c_vp, c_vp, c_vp, c_i64, c_f32, c_f32, c_f32, c_f32, c_f32, c_i64, c_f32, c_vp]),
    "pn_t5_num_params": (c_i64, [c_vp]),
    "pn_t5_num_tensors": (c_i32, [c_vp]),
    "pn_t5_tensor_info": (c_int, [c_vp, c_i32, c_vp, c_i32, c_vp, c_vp]),
    "pn_t5_params": (c_vp, [c_vp]),
    "pn_t5_create": (c_int, [c_vp, c_vp]),
    "pn_t5_destroy": (c_int, [c_vp]),
    "pn_t5_set_option": (c_int, [c_vp, C.c_char_p, c_i64]),
    "pn_t5_load_weight": (c_int, [c_vp, C.c_char_p, c_vp, c_i64, c_vp]),
    "pn_t5_finalize": (c_int, [c_vp, c_vp]),
    "pn_t5_workspace_bytes": (c_i64, [c_vp, c_i64, c_i32, c_i32, c_int]),
    "pn_t5_vocab_pad": (c_i32, [c_vp]),
    "pn_t5_score": (c_int, [c_vp, c_vp, c_i64, c_i32, c_i32, c_i32, c_vp, c_vp, c_vp, c_i64, c_vp]),
    "pn_t5_score_lens": (c_int, [c_vp, c_vp, c_vp, c_i64, c_i32, c_i32, c_i32, c_vp, c_vp, c_vp, c_i64, c_vp]),
    "pn_t5_logits": (c_int, [c_vp, c_vp, c_vp, c_i64, c_i32, c_i32, c_i32, c_i32, c_vp, c_vp, c_vp,
                             c_i64, c_vp]),
    "pn_prof_enable": (c_int, [c_int]),
    "pn_prof_read": (c_int, [c_int, c_vp, c_vp, c_vp]),
    "pn_topk_metrics": (c_int, [c_vp, c_vp, c_vp, c_i64, c_i32, c_i32, c_vp, c_vp, c_vp, c_vp]),
    "pn_topk_metrics_qrels": (c_int, [c_vp, c_vp, c_vp, c_vp, c_vp, c_i64, c_i32, c_i32, c_vp, c_vp, c_vp, c_vp]),
}

_lib = None


def lib():
    """The loaded library; raises (never falls back) if it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; "
                "g.build()'` (or `make -C pacednegatives_b200/csrc`).  pacednegatives_b200 has no "
                "CPU fallback.")
        l = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(l, name)
            fn.restype, fn.argtypes = res, args
        _lib = l
    return _lib


_gpu_ok = False


def require_gpu():
    global _gpu_ok
    if _gpu_ok:
        return
    if not torch.cuda.is_available():
        raise RuntimeError("pacednegatives_b200 needs a CUDA device (B200, sm_100a); none is visible "
                           "and there is no CPU fallback")
    check(lib().pn_device_check())
    _gpu_ok = True


def check(rc):
    if rc != 0:
        msg = lib().pn_last_error().decode("utf-8", "replace")
        if rc == -1:
            raise ValueError(msg)
        raise RuntimeError(f"libpn_b200 error {rc}: {msg}")


def ptr(t):
    """Device pointer of a tensor (None -> NULL).  The tensor must be contiguous and on CUDA."""
    if t is None:
        return None
    if not t.is_cuda:
        raise ValueError("expected a CUDA tensor")
    if not t.is_contiguous():
        raise ValueError("expected a contiguous tensor")
    return C.c_void_p(t.data_ptr())


def stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def launch_count():
    return int(lib().pn_launch_count())
