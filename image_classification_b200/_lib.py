"""ctypes binding of libcnx_b200.so — the only way Python reaches the CUDA kernels.

There is deliberately no fallback: if the shared library is missing or the device is not sm_100,
every op raises.  (The reference loop swallows per-batch exceptions,
/root/reference/train_advanced_v4.py:277-279, so callers that want loud failures should not rely
on that loop to surface them; see `train_step.py`.)
"""
from __future__ import annotations

import ctypes
import os
from ctypes import POINTER, c_char_p, c_float, c_int, c_int32, c_int64, c_size_t, c_void_p

HERE = os.path.dirname(os.path.abspath(__file__))
# CNX_LIB_PATH: load another build of the library (same-box A/B of two kernel versions)
LIB_PATH = os.environ.get("CNX_LIB_PATH") or os.path.join(HERE, "libcnx_b200.so")


class CnxError(RuntimeError):
    """A libcnx_b200 entry point returned a non-zero status."""


class MixPlanEntry(ctypes.Structure):
    """Mirror of `cnx_mix_plan_entry` (include/cnx_b200.h)."""

    _fields_ = [
        ("mode", c_int32),
        ("partner", c_int32),
        ("lam", c_float),
        ("one_minus_lam", c_float),
        ("bbx1", c_int32),
        ("bby1", c_int32),
        ("bbx2", c_int32),
        ("bby2", c_int32),
    ]


# cnx_op values (include/cnx_b200.h)
(CNX_OP_DWCONV_LN_FWD, CNX_OP_DWCONV_LN_BWD, CNX_OP_MLP_FWD, CNX_OP_MLP_BWD, CNX_OP_MIX_BATCH,
 CNX_OP_SOFTCE, CNX_OP_EMA) = range(7)

P = c_void_p
I = c_int
I64 = c_int64
F = c_float

# name -> (restype, argtypes); must list every symbol include/cnx_b200.h declares
SIGNATURES = {
    "cnx_version": (I, []),
    "cnx_last_error_string": (c_char_p, []),
    "cnx_check_device": (I, []),
    "cnx_debug_timeout_flag": (I, []),
    "cnx_debug_set_timeout_flag": (None, [I]),
    "cnx_debug_gemm_tuning": (None, [I, I, I, I]),
    "cnx_debug_gemm_epilogue_warps": (None, [I]),
    "cnx_debug_gemm_staging_chunk": (None, [I]),
    "cnx_debug_max_clusters": (I, [I]),
    "cnx_workspace_bytes": (c_size_t, [I, I, I, I, I]),
    "cnx_prep_block_weights": (I, [P] * 8 + [P] * 8 + [I, P]),
    "cnx_unfold_block_grads": (I, [P] * 11 + [P] * 7 + [I, P]),
    "cnx_dwconv_ln_fwd": (I, [P, P, P, P, P, P, I, I, I, I, F, P]),
    "cnx_dwconv_ln_bwd": (I, [P] * 10 + [I, I, I, I, P, c_size_t, P]),
    "cnx_dwconv_ln_bwd_acc": (I, [P] * 10 + [I, I, I, I, P, c_size_t, P]),
    "cnx_layernorm_fwd": (I, [P, P, P, P, P, P, I64, I, F, P]),
    "cnx_layernorm_bwd": (I, [P, P, P, P, P, P, P, P, I64, I, P]),
    "cnx_colsum_bf16": (I, [P, P, I64, I, P]),
    "cnx_dwconv_kxk_fwd": (I, [P, P, P, I, I, I, I, I, I, P]),
    "cnx_dwconv_kxk_bwd": (I, [P, P, P, P, P, I, I, I, I, I, I, I, P]),
    "cnx_avgpool_nhwc_fwd": (I, [P, P, I, I, I, P]),
    "cnx_avgpool_nhwc_bwd": (I, [P, P, I, I, I, P]),
    "cnx_bn_stats": (I, [P, P, I64, I, P]),
    "cnx_bn_act_fwd": (I, [P, P, P, P, P, P, F, F, I, P, P, P, P, I, I, I, P]),
    "cnx_bn_act_bwd": (I, [P, P, P, P, P, P, I, P, P, P, I64, I, P]),
    "cnx_se_scale_fwd": (I, [P, P, P, I, I, I, P]),
    "cnx_se_scale_bwd_reduce": (I, [P, P, P, I, I, I, P]),
    "cnx_se_scale_bwd_apply": (I, [P, P, P, P, I, I, I, P]),
    "cnx_se_gate_fwd": (I, [P, F, P, P, P, P, P, P, I, I, I, P]),
    "cnx_se_gate_bwd": (I, [P, P, P, P, F, P, P, P, P, P, P, P, P, P, I, I, I, P]),
    "cnx_layernorm_patch_fwd": (I, [P, P, P, P, P, P, I, I, I, I, I, F, P]),
    "cnx_layernorm_patch_bwd": (I, [P, P, P, P, P, P, P, P, I, I, I, I, I, P]),
    "cnx_patchify_nchw_f32": (I, [P, P, I, I, I, I, I, P]),
    "cnx_gemm_bf16": (I, [P, I64, I, P, I64, I, I, I, I, I, P, I64, P, I64, P, P, I64, P, P]),
    "cnx_fc1_gelu_fwd": (I, [P, P, P, P, P, I, I, P]),
    "cnx_fc2_residual_fwd": (I, [P, P, P, P, P, I, I, P]),
    "cnx_mlp_bwd_dz": (I, [P, P, P, P, P, I, I, P]),
    "cnx_mlp_bwd_dxhat": (I, [P, P, P, I, I, P]),
    "cnx_mlp_bwd_dz_stats": (I, [P, P, P, P, P, P, P, I, I, P]),
    "cnx_mlp_bwd_dconv": (I, [P, P, P, P, P, P, I, I, P]),
    "cnx_dwconv_bwd": (I, [P] * 8 + [I, I, I, I, I, P]),
    "cnx_mlp_bwd_wgrad": (I, [P, P, P, P, P, P, I, I, P]),
    "cnx_mix_batch": (I, [P, P, P, P, P, I, I, I, I, I, P]),
    "cnx_mix_batch_smooth": (I, [P, P, P, P, P, I, I, I, I, I, F, F, P]),
    "cnx_softce_fwd_bwd": (I, [P, P, P, P, I, I, F, P]),
    "cnx_ema_lerp_multi": (I, [P, P, P, P, I, I64, F, F, P]),
    "cnx_ema_chunk_elems": (I, []),
    "cnx_optim_grad_norm": (I, [P, P, P, I, I64, P, F, P, P, P]),
    "cnx_optim_grad_norm_noalloc": (I, [P, P, P, I, I64, P, F, P, P]),
    "cnx_optim_grad_norm_noalloc_max_chunks": (I, []),
    "cnx_optim_adamw_ema": (I, [P, P, P, P, P, P, P, I, I64, F, F, F, F, F, F, F, P, P]),
    "cnx_nchw_f32_to_nhwc_bf16": (I, [P, P, I, I, I, I, P]),
    "cnx_nhwc_bf16_to_nchw_f32": (I, [P, P, I, I, I, I, P]),
}

_lib = None


def load() -> ctypes.CDLL:
    """Load libcnx_b200.so (built in-tree by `image_classification_b200.build`)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise CnxError(
            f"{LIB_PATH} is missing: build it with `python -m image_classification_b200.build` "
            "(there is no PyTorch/CPU fallback for these ops)")
    lib = ctypes.CDLL(LIB_PATH)
    for name, (restype, argtypes) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError here = header/library mismatch
        fn.restype = restype
        fn.argtypes = argtypes
    _lib = lib
    return lib


def last_error() -> str:
    msg = load().cnx_last_error_string()
    return msg.decode("utf-8", "replace") if msg else ""


def check(rc: int, what: str) -> None:
    if rc != 0:
        raise CnxError(f"{what} failed with status {rc}: {last_error()}")


# kernels launched by one successful call of each entry point (bench.py's `gpu_launches` claim;
# cudaMemsetAsync nodes are not counted)
LAUNCHES_PER_CALL = {
    "cnx_prep_block_weights": 1, "cnx_unfold_block_grads": 2, "cnx_dwconv_ln_fwd": 2,
    "cnx_dwconv_ln_bwd": 2, "cnx_dwconv_ln_bwd_acc": 2, "cnx_gemm_bf16": 1, "cnx_layernorm_fwd": 1, "cnx_layernorm_bwd": 1, "cnx_colsum_bf16": 1, "cnx_dwconv_kxk_fwd": 1, "cnx_dwconv_kxk_bwd": 2, "cnx_avgpool_nhwc_fwd": 1, "cnx_avgpool_nhwc_bwd": 1, "cnx_layernorm_patch_fwd": 1, "cnx_layernorm_patch_bwd": 1, "cnx_patchify_nchw_f32": 1, "cnx_fc1_gelu_fwd": 1, "cnx_fc2_residual_fwd": 1,
    "cnx_mlp_bwd_dz": 1, "cnx_mlp_bwd_dxhat": 1, "cnx_mlp_bwd_dz_stats": 1, "cnx_mlp_bwd_dconv": 1, "cnx_dwconv_bwd": 1, "cnx_mlp_bwd_wgrad": 2, "cnx_mix_batch": 1, "cnx_mix_batch_smooth": 1,
    "cnx_softce_fwd_bwd": 1, "cnx_ema_lerp_multi": 1, "cnx_optim_grad_norm": 1, "cnx_optim_grad_norm_noalloc": 1, "cnx_optim_adamw_ema": 1, "cnx_nchw_f32_to_nhwc_bf16": 1,
    "cnx_nhwc_bf16_to_nchw_f32": 1,
    "cnx_bn_stats": 1, "cnx_bn_act_fwd": 1, "cnx_bn_act_bwd": 2, "cnx_se_scale_fwd": 1,
    "cnx_se_scale_bwd_reduce": 1, "cnx_se_scale_bwd_apply": 1, "cnx_se_gate_fwd": 2,
    "cnx_se_gate_bwd": 3,
}
launch_count = 0          # total kernels launched through this binding
call_hook = None          # optional callable(name, phase, args), phase in {"pre", "post"} (profiling)


def call(name: str, *args) -> None:
    """Invoke an int-returning entry point and raise CnxError on failure."""
    global launch_count
    fn = getattr(load(), name)
    if call_hook is not None:
        call_hook(name, "pre", args)
    rc = fn(*args)
    if call_hook is not None:
        call_hook(name, "post", args)
    check(rc, name)
    launch_count += LAUNCHES_PER_CALL.get(name, 0)
