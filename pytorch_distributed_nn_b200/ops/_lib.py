"""ctypes binding of libpdnn_b200.so (the hand-written sm_100a kernel library, C ABI in include/pdnn_b200.h).

The library is built in-tree by ``csrc/build.py`` (nvcc, ``-gencode arch=compute_100a,code=sm_100a``).
Nothing here falls back to PyTorch: if the shared object is missing, or a kernel returns an error, the
call raises. CPU-only code paths (gloo backend, oracles) never import this module's ``lib()``.
"""
from __future__ import annotations

import ctypes as C
import os
import threading

import torch

# PDNN_B200_LIB points at another build of the same library (A/B runs of a kernel variant); default: the in-tree build
_LIB_PATH = os.environ.get("PDNN_B200_LIB") or os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "lib", "libpdnn_b200.so")
_lib = None
_lock = threading.Lock()

MAX_WORKERS = 16


class Epilogue(C.Structure):
    """Mirror of pdnn_epilogue_t."""
    _fields_ = [
        ("out", C.c_void_p), ("ldo", C.c_longlong), ("out_mode", C.c_int), ("relu", C.c_int),
        ("alpha", C.c_float), ("bias", C.c_void_p), ("residual", C.c_void_p),
        ("guard_closed", C.c_void_p), ("guard_step", C.c_void_p), ("max_ctas", C.c_int),
    ]


MAX_RECEIVERS = 24
MAX_BUCKETS = 32


class PSMaster(C.Structure):
    """Mirror of pdnn_ps_master_t."""
    _fields_ = [
        ("slots", C.c_void_p * MAX_WORKERS), ("arrival_flags", C.c_void_p), ("arrival_stride", C.c_int),
        ("recv_weights", (C.c_void_p * MAX_RECEIVERS) * 2), ("recv_ready", C.c_void_p * MAX_RECEIVERS),
        ("worker_closed", C.c_void_p * MAX_WORKERS), ("multicast_weights", C.c_void_p * 2), ("multicast_slots", C.c_void_p),
        ("master_w", C.c_void_p), ("momentum", C.c_void_p), ("control", C.c_void_p),
        ("n", C.c_longlong), ("n_workers", C.c_int), ("n_receivers", C.c_int), ("k", C.c_int), ("slot_is_bf16", C.c_int),
        ("weights_are_bf16", C.c_int), ("avg_mode", C.c_int), ("lr", C.c_float), ("mu", C.c_float),
        ("step", C.c_uint32), ("deadline_ns", C.c_longlong), ("max_ctas", C.c_int), ("step_ptr", C.c_void_p),
        ("poison_slots", C.c_int), ("n_buckets", C.c_int),
        ("bucket_lo", C.c_longlong * MAX_BUCKETS), ("bucket_hi", C.c_longlong * MAX_BUCKETS), ("hard_timeout_ns", C.c_longlong),
        ("poll_sleep_ns", C.c_int),
    ]


class DelaySpec(C.Structure):
    """Mirror of pdnn_delay_spec_t."""
    _fields_ = [("kind", C.c_int), ("a", C.c_float), ("b", C.c_float), ("scale", C.c_float), ("prob", C.c_float),
                ("rank_mask", C.c_ulonglong), ("seed", C.c_ulonglong)]


class FcPsConfig(C.Structure):
    """Mirror of pdnn_fc_ps_config_t."""
    _fields_ = [
        ("n_dims", C.c_int), ("dims", C.c_int * 64), ("batch", C.c_int), ("lr", C.c_float), ("n_workers", C.c_int),
        ("n_to_collect", C.c_int), ("n_train_iters", C.c_int), ("shortcircuit", C.c_int), ("generate_timeline", C.c_int),
        ("with_evaluator", C.c_int), ("eval_batches", C.c_int), ("delay_worker_mask", C.c_ulonglong),
        ("delay_us_per_layer", C.c_longlong), ("seed", C.c_uint), ("synthetic_examples", C.c_int), ("max_devices", C.c_int),
        ("data_dir", C.c_char * 256), ("out_dir", C.c_char * 256), ("device_protocol", C.c_int),
    ]


class FcPsResult(C.Structure):
    """Mirror of pdnn_fc_ps_result_t."""
    _fields_ = [
        ("steps", C.c_int), ("total_ms", C.c_double), ("mean_step_ms", C.c_double), ("gradients_accepted", C.c_longlong),
        ("gradients_dropped_stale", C.c_longlong), ("steps_abandoned", C.c_longlong), ("avg_gradients_per_step", C.c_double),
        ("first_loss", C.c_float), ("final_loss", C.c_float), ("first_err", C.c_float), ("final_err", C.c_float),
        ("scheme", C.c_char * 128),
    ]


class FcPsIO(C.Structure):
    """Mirror of pdnn_fc_ps_io_t (host pointers)."""
    _fields_ = [
        ("train_images", C.c_void_p), ("train_labels", C.c_void_p), ("n_train", C.c_int),
        ("test_images", C.c_void_p), ("test_labels", C.c_void_p), ("n_test", C.c_int),
        ("init_weights", C.c_void_p), ("final_weights", C.c_void_p),
        ("accept_log", C.c_void_p), ("accept_log_rows", C.c_longlong), ("accept_log_used", C.c_longlong),
    ]


_P, _I, _L, _F, _U = C.c_void_p, C.c_int, C.c_longlong, C.c_float, C.c_uint32

# name -> argtypes (stream pointer included where the C function takes one)
_SIGS = {
    "pdnn_gemm_bf16": [_P, _I, _L, _P, _I, _L, _I, _I, _I, C.POINTER(Epilogue), _I, _P],
    "pdnn_conv_fprop_bf16": [_P, _P, _I, _I, _I, _I, _I, _I, _I, _I, C.POINTER(Epilogue), _P],
    "pdnn_conv_fprop_strided_bf16": [_P, _P, _I, _I, _I, _I, _I, _I, _I, _I, _I, C.POINTER(Epilogue), _P],
    "pdnn_conv_wgrad_strided_bf16": [_P, _P, _I, _I, _I, _I, _I, _I, _I, _I, _I, C.POINTER(Epilogue), _I, _P],
    "pdnn_conv_dgrad_s2_bf16": [_P, _P, _I, _I, _I, _I, _I, _I, _I, _I, C.POINTER(Epilogue), _P],
    "pdnn_conv_halo_supported": [_I] * 8,
    "pdnn_conv_stem_supported": [_I] * 9,
    "pdnn_conv_halo_fprop_bf16": [_P, _P, _I, _I, _I, _I, _I, C.POINTER(Epilogue), _P],
    "pdnn_conv_halo_fprop_stats_bf16": [_P, _P, _I, _I, _I, _I, _I, C.POINTER(Epilogue), _P, _P],
    "pdnn_conv_halo_wgrad_supported": [_I] * 8,
    "pdnn_conv_halo_wgrad_bf16": [_P, _P, _I, _I, _I, _I, _I, _P, _P, _P, _I, _P],
    "pdnn_conv_wgrad_bf16": [_P, _P, _I, _I, _I, _I, _I, _I, _I, _I, C.POINTER(Epilogue), _I, _P],
    "pdnn_cast_f32_to_bf16": [_P, _P, _L, _P],
    "pdnn_cast_bf16_to_f32": [_P, _P, _L, _P],
    "pdnn_nchw_f32_to_nhwc_bf16": [_P, _P, _I, _I, _I, _I, _I, _P],
    "pdnn_nchw_f32_to_nhwc_f32": [_P, _P, _I, _I, _I, _I, _I, _P],
    "pdnn_wgt_flip_transpose": [_P, _P, _I, _I, _I, _I, _P],
    "pdnn_conv_stem_bf16": [_P, _P, _I, _I, _I, _I, _I, _P, _P, _P, _P],
    "pdnn_set_cluster_mode": [_I],
    "pdnn_wgt_flip_transpose_multi": [_P, _P, _I, _I, _P],
    "pdnn_oihw_to_krsc_f32": [_P, _P, _I, _I, _I, _I, _I, _P],
    "pdnn_im2col_bf16": [_P, _P] + [_I] * 11 + [_P],
    "pdnn_im2col_f32": [_P, _P] + [_I] * 11 + [_P],
    "pdnn_col2im_bf16": [_P, _P] + [_I] * 11 + [_P],
    "pdnn_col2im_f32": [_P, _P] + [_I] * 11 + [_P],
    "pdnn_bn_fwd_stats": [_P, _P, _L, _I, _P],
    "pdnn_bn_fwd_apply": [_P, _P, _P, _P, _P, _P, _P, _P, _P, _L, _I, _F, _F, _I, _P],
    "pdnn_bn_bwd_reduce": [_P, _P, _P, _P, _P, _L, _I, _I, _P, _P, _P],
    "pdnn_bn_bwd_apply": [_P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _L, _I, _I, _P, _P],
    "pdnn_relu_fwd": [_P, _P, _L, _P],
    "pdnn_relu_bwd": [_P, _P, _P, _L, _P],
    "pdnn_relu_fwd_f32": [_P, _P, _L, _P],
    "pdnn_relu_bwd_f32": [_P, _P, _P, _L, _P],
    "pdnn_add_bf16": [_P, _P, _P, _L, _P],
    "pdnn_maxpool2_fwd": [_P, _P, _I, _I, _I, _I, _P],
    "pdnn_maxpool2_bwd": [_P, _P, _P, _P, _I, _I, _I, _I, _P],
    "pdnn_maxpool2_fwd_f32": [_P, _P, _I, _I, _I, _I, _P],
    "pdnn_maxpool2_bwd_f32": [_P, _P, _P, _P, _I, _I, _I, _I, _P],
    "pdnn_avgpool_fwd": [_P, _P, _I, _I, _I, _P],
    "pdnn_avgpool_bwd": [_P, _P, _I, _I, _I, _P],
    "pdnn_col_sum": [_P, _L, _P, _L, _I, _I, _P],
    "pdnn_col_sum_f32": [_P, _L, _P, _L, _I, _I, _P],
    "pdnn_cross_entropy_fwd_bwd": [_P, _P, _P, _P, _P, _I, _I, _I, _P],
    "pdnn_cross_entropy_fwd_bwd_f32": [_P, _P, _P, _P, _P, _I, _I, _I, _P],
    "pdnn_topk_correct": [_P, _P, _I, _I, _I, C.POINTER(C.c_int), _I, _P, _P],
    "pdnn_sgd_momentum": [_P, _P, _P, _P, _L, _F, _F, _F, _I, _P],
    "pdnn_sigmoid_fwd": [_P, _P, _L, _P],
    "pdnn_sigmoid_bwd": [_P, _P, _P, _L, _P],
    "pdnn_sigmoid_fwd_f32": [_P, _P, _L, _P],
    "pdnn_sigmoid_bwd_f32": [_P, _P, _P, _L, _P],
    "pdnn_gemm_simt_f32": [_P, _I, _L, _P, _I, _L, _P, _L, _I, _I, _I, _F, _P, _I, _I, _P],
    "pdnn_ps_push": [_P, _P, _L, _I, _P, _U, _P, _P, _I, _P],
    "pdnn_ps_push_dev": [_P, _P, _L, _I, _P, _P, _P, _P, _I, _P],
    "pdnn_ps_wait_next": [_P, _P, _L, _P],
    "pdnn_ps_publish": [_P, _U, _P, _P, _P, _P],
    "pdnn_ps_master_step": [C.POINTER(PSMaster), _P, _P],
    "pdnn_ps_master_bucket": [C.POINTER(PSMaster), _P, _I, _P],
    "pdnn_ps_fetch_weights": [_P, _P, _P, _L, _P, _U, _P],
    "pdnn_ps_wait_flag": [_P, _U, _P, _L, _P],
    "pdnn_delay": [_L, _P, _P, _P],
    "pdnn_delay_sampled": [C.POINTER(DelaySpec), _I, _I, _P, _P, _P, _P],
    "pdnn_set_word": [_P, _U, _P],
    "pdnn_set_abort_guard": [_P, _P],
    "pdnn_p2p_copy": [_P, _P, _L, _I, _P],
    "pdnn_mc_store": [_P, _P, _L, _I, _P],
    "pdnn_mc_ld_reduce_f32": [_P, _P, _L, _I, _P],
    "pdnn_flag_pingpong": [_P, _P, _I, _I, _U, _P, _P],
    "pdnn_malloc": [C.POINTER(C.c_void_p), _L],
    "pdnn_free": [_P],
    "pdnn_ipc_get_handle": [_P, _P],
    "pdnn_ipc_open_handle": [_P, C.POINTER(C.c_void_p)],
    "pdnn_ipc_close": [_P],
    "pdnn_enable_peer_access": [_I],
    "pdnn_bn_fwd_noalloc": [_P, _P, _P, _P, _P, _P, _P, _P, _L, _I, _F, _F, _I, _P],
    "pdnn_fc_ps_run": [C.POINTER(FcPsConfig), C.POINTER(FcPsResult)],
    "pdnn_fc_ps_run_ex": [C.POINTER(FcPsConfig), C.POINTER(FcPsResult), C.POINTER(FcPsIO)],
    "pdnn_fc_ps_selftest": [_I, C.POINTER(C.c_float), C.POINTER(C.c_float)],
}


class PdnnError(RuntimeError):
    pass


def lib_path() -> str:
    return _LIB_PATH


def available() -> bool:
    return os.path.exists(_LIB_PATH)


def lib():
    """Load (once) and return the ctypes handle. Raises if the library has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is not None:
            return _lib
        if not os.path.exists(_LIB_PATH):
            raise PdnnError(
                f"{_LIB_PATH} is missing: run `python csrc/build.py` (or __graft_entry__.build()). "
                "The B200 path has no PyTorch fallback.")
        h = C.CDLL(_LIB_PATH, mode=C.RTLD_GLOBAL)
        h.pdnn_last_error.restype = C.c_char_p
        h.pdnn_launch_count.restype = C.c_longlong
        h.pdnn_reset_launch_count.restype = None
        h.pdnn_add_launch_count.restype = None
        h.pdnn_add_launch_count.argtypes = [C.c_longlong]
        for name, sig in _SIGS.items():
            fn = getattr(h, name, None)
            if fn is None:
                continue
            fn.argtypes = sig
            fn.restype = C.c_int
        _lib = h
    return _lib


def ptr(t):
    """Device (or host) pointer of a tensor, None -> NULL, ints pass through."""
    if t is None:
        return None
    if isinstance(t, int):
        return t
    return t.data_ptr()


def stream_ptr():
    return torch.cuda.current_stream().cuda_stream


def call(name: str, *args):
    """Invoke an entry point; tensors are converted to pointers; raises PdnnError on failure."""
    h = lib()
    fn = getattr(h, name)
    conv = [ptr(a) if isinstance(a, torch.Tensor) else a for a in args]
    rc = fn(*conv)
    if rc != 0:
        raise PdnnError(f"{name} failed (rc={rc}): {h.pdnn_last_error().decode(errors='replace')}")
    return rc


def launch_count() -> int:
    return int(lib().pdnn_launch_count())


def reset_launch_count() -> None:
    lib().pdnn_reset_launch_count()


def add_launch_count(n: int) -> None:
    """Account for library kernels that ran through a CUDA-graph replay (see pdnn_add_launch_count)."""
    lib().pdnn_add_launch_count(int(n))
