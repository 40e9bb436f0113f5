"""ctypes binding of ``crackfl/_lib/libcrackfl.so`` (C ABI in ``include/crackfl.h``).

``get_lib()`` loads the in-tree library (building it first if the sources changed and nvcc is
available).  On a box with a GPU a missing library is a hard error: there is no eager/PyTorch
fallback for the sm_100a kernels.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

from . import build as _build

_lib: Optional[C.CDLL] = None

vp, i32, i64, f32 = C.c_void_p, C.c_int, C.c_long, C.c_float

MAX_CLIENTS, MAX_OUT, MAX_RANKS = 128, 16, 16


class PrepDesc(C.Structure):
    _fields_ = [("src_off", i64), ("dst_off", i64), ("N", i32), ("T", i32), ("K", i32), ("pad_", i32),
                ("s_n", i64), ("s_t", i64), ("s_k", i64)]


class BnFin(C.Structure):
    _fields_ = [("stats", vp), ("gamma", vp), ("beta", vp), ("run_mean", vp), ("run_var", vp), ("bn_out", vp),
                ("count", f32), ("eps", f32), ("momentum", f32)]


class BnEvalDesc(C.Structure):
    _fields_ = [("gamma_off", i64), ("beta_off", i64), ("mean_off", i64), ("var_off", i64), ("bn_off", i64), ("C", i32),
                ("pad_", i32)]


class DwDiagDesc(C.Structure):
    _fields_ = [("src_off", i64), ("dst_off", i64), ("C", i32), ("kc", i32), ("flip", i32), ("pad_", i32)]


class WgradRedDesc(C.Structure):
    _fields_ = [("ws_off", i64), ("dw_off", i64), ("s_tap", i64), ("s_ci", i64), ("s_co", i64), ("need", i64),
                ("groups", i32), ("ntn", i32), ("splits", i32), ("BN", i32), ("cw", i32), ("mblk", i32),
                ("total_blocks", i32), ("cchunks", i32), ("cin_valid", i32), ("zl", i32)]


class FedAggParams(C.Structure):
    _fields_ = [("client", vp * MAX_CLIENTS), ("coef", f32 * MAX_CLIENTS), ("out", vp * MAX_OUT),
                ("num_clients", i32), ("num_out", i32), ("payload_bf16", i32), ("out_bf16", i32),
                ("delta", i32), ("server_opt", i32), ("eta", f32), ("beta1", f32), ("beta2", f32), ("tau", f32),
                ("master", vp), ("m", vp), ("v", vp), ("trainable", vp), ("mc_out", vp),
                ("begin", i64), ("end", i64), ("pads", vp * MAX_RANKS), ("rank", i32), ("world", i32),
                ("epoch", C.c_uint32), ("error_flag", vp), ("local_sync", vp), ("mc_in", vp), ("unroll", i32), ("load_mode", i32),
                ("timeout_cycles", C.c_longlong)]


_SIGS = {
    "crackfl_entry_conv_fwd": [vp, vp, vp, vp, i32, i32, i32, i32, vp],
    "crackfl_entry_conv_wgrad": [vp, vp, vp, i32, i32, i32, i32, vp],
    "crackfl_entry_im2col": [vp, vp, i32, i32, i32, vp],
    "crackfl_bn_stats": [vp, vp, i64, i32, vp],
    "crackfl_bn_finalize": [vp, vp, vp, vp, vp, vp, i32, f32, f32, f32, vp],
    "crackfl_bn_eval_prepare": [vp, vp, vp, vp, vp, i32, f32, vp],
    "crackfl_bn_eval_prepare_all": [vp, vp, vp, i32, f32, vp],
    "crackfl_bn_act": [vp, vp, vp, vp, i32, i32, i32, i32, i32, vp],
    "crackfl_bn_bwd_reduce": [vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, vp],
    "crackfl_bn_bwd_apply": [vp, vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, vp],
    "crackfl_bn_train_fwd_noalloc": [vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, f32, f32, vp],
    "crackfl_noalloc_max_channels": [],
    "crackfl_dw3x3_fwd": [vp, vp, vp, vp, i32, i32, i32, i32, i32, vp],
    "crackfl_dw3x3_bwd": [vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, vp],
    "crackfl_bn_act_fin": [vp, C.POINTER(BnFin), vp, vp, i32, i32, i32, i32, i32, vp],
    "crackfl_dw3x3_fwd_fin": [vp, vp, C.POINTER(BnFin), vp, i32, i32, i32, i32, vp],
    "crackfl_pool_add_fwd_fin": [vp, C.POINTER(BnFin), vp, vp, vp, vp, vp, i32, i32, i32, i32, vp],
    "crackfl_dec_combine_fwd_fin": [vp, C.POINTER(BnFin), vp, vp, vp, i32, i32, i32, i32, i32, vp],
    "crackfl_dw3x3_tma": [vp, vp, vp, C.POINTER(BnFin), vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, i32, vp],
    "crackfl_dw3x3_wgrad_tma": [vp, vp, vp, vp, i32, i32, i32, i32, i32, vp],
    "crackfl_dw3x3_bwd_tma": [vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, vp],
    "crackfl_dw3x3_bwd_tma_bn": [vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, vp, vp, vp],
    "crackfl_dw3x3_dgrad": [vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, vp],
    "crackfl_dw3x3_wgrad": [vp, vp, vp, vp, i32, i32, i32, i32, i32, vp],
    "crackfl_pool_add_fwd": [vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, vp],
    "crackfl_dec_combine_fwd": [vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, vp],
    "crackfl_pool_bwd": [vp, vp, vp, i32, i32, i32, i32, vp],
    "crackfl_pool_bwd_bn": [vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, vp],
    "crackfl_down2sum": [vp, vp, vp, vp, i32, i32, i32, i32, vp],
    "crackfl_down2sum_bn": [vp, vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, vp],
    "crackfl_relu_bwd_add": [vp, vp, vp, vp, i64, vp],
    "crackfl_colsum": [vp, vp, i64, i32, vp],
    "crackfl_head_fwd": [vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, vp],
    "crackfl_head_bwd": [vp, vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, f32, f32, i32, vp],
    "crackfl_head_fwd_bwd": [vp, vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, f32, i32, vp],
    "crackfl_conv_simt_fwd": [vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, vp],
    "crackfl_wgrad_simt": [vp, vp, vp, i32, i32, i32, i32, i32, i32, i64, i64, i64, vp],
    "crackfl_conv_tc_fwd": [vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, vp],
    "crackfl_wgrad_tc": [vp, vp, vp, i32, i32, i32, i32, i32, i32, i64, i64, i64, vp],
    "crackfl_wgrad_tc_ws": [vp, vp, vp, i32, i32, i32, i32, i32, i32, i64, i64, i64, vp, i64, vp],
    "crackfl_wgrad_tc_ws_cv": [vp, vp, vp, i32, i32, i32, i32, i32, i32, i64, i64, i64, vp, i64, i32, vp],
    "crackfl_wgrad_plan": [i32, i32, i32, i32, i32, i32, i64, i64, i64, i32, C.POINTER(WgradRedDesc)],
    "crackfl_wgrad_tc_deferred": [vp, vp, i32, i32, i32, i32, i32, i32, vp, i64, vp],
    "crackfl_wgrad_reduce_all": [vp, i32, vp, vp, vp],
    "crackfl_tc_selftest_supported": [],
    "crackfl_weight_prep": [vp, vp, vp, i32, vp],
    "crackfl_weight_prep_flat": [vp, vp, vp, vp, i32, i32, vp],
    "crackfl_adam_step": [vp, vp, vp, vp, vp, i64, f32, f32, f32, f32, vp],
    "crackfl_adam_part": [vp, vp, vp, vp, vp, i64, f32, f32, f32, f32, vp],
    "crackfl_f32_to_bf16": [vp, vp, i64, vp],
    "crackfl_adam_step_tail": [vp, vp, vp, vp, vp, i64, f32, f32, f32, f32, vp, vp, vp, i32, vp, vp, i32, vp],
    "crackfl_accum_small": [vp, vp, i32, vp],
    "crackfl_publish_delta": [vp, vp, vp, i64, vp],
    "crackfl_synth_cracks": [vp, vp, i32, i32, i32, C.c_uint, vp],
    "crackfl_measure_masks": [vp, vp, i32, i32, i32, i32, vp],
    "crackfl_conv_tc_xf": [vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, vp, C.POINTER(BnFin), vp, i32, vp],
    "crackfl_conv_tc_bnbwd": [vp, vp, vp, vp, i32, i32, i32, i32, i32, vp, vp, i32, vp, vp, vp, vp],
    "crackfl_dw_diag_prep": [vp, vp, vp, i32, vp],
    "crackfl_fedagg": [C.POINTER(FedAggParams), i32, vp],
    "crackfl_ipc_export": [vp, vp],
    "crackfl_ipc_open": [vp, C.POINTER(vp)],
    "crackfl_ipc_close": [vp],
    "crackfl_enable_peer_access": [i32],
    "crackfl_dev_malloc": [C.POINTER(vp), C.c_size_t],
    "crackfl_dev_free": [vp],
}


class KernelError(RuntimeError):
    pass


def lib_path() -> str:
    return _build.LIBPATH


def get_lib(build_if_needed: bool = True) -> C.CDLL:
    global _lib
    if _lib is not None:
        return _lib
    path = _build.LIBPATH
    # Rebuild only when the library is missing, or when it is stale AND we are a single process (developer loop).
    # Ranks of a multi-process job never race to rebuild a library that is already there.
    multi = int(os.environ.get("WORLD_SIZE", "1")) > 1
    if build_if_needed and (not os.path.exists(path) or (not multi and not _build.is_current())):
        try:
            _build.build()
        except Exception as e:
            if not os.path.exists(path):
                raise KernelError(f"libcrackfl.so is missing and could not be built: {e}") from e
    if not os.path.exists(path):
        raise KernelError(f"{path} not found: run `python -m crackfl.ops.build` (no PyTorch fallback exists)")
    lib = C.CDLL(path)
    for name, sig in _SIGS.items():
        try:
            fn = getattr(lib, name)
        except AttributeError:
            continue
        fn.argtypes = sig
        fn.restype = i32
    _lib = lib
    return lib


def get_ieee_lib() -> C.CDLL:
    """The optimiser / elementwise kernels built without --use_fast_math (tests only, see build.IEEE_LIBPATH)."""
    get_lib()
    lib = C.CDLL(_build.IEEE_LIBPATH)
    for name, sig in _SIGS.items():
        try:
            fn = getattr(lib, name)
        except AttributeError:
            continue
        fn.argtypes = sig
        fn.restype = i32
    return lib


def check(code: int, what: str = ""):
    if code != 0:
        raise KernelError(f"{what or 'kernel launch'} failed with cudaError {code}")


def ptr(t) -> int:
    """Device pointer of a torch tensor (or None -> NULL)."""
    return 0 if t is None else t.data_ptr()
