"""ctypes bindings to the in-tree native libraries (C ABI in ``include/crossscale.h``).

Plain C ABI + ctypes (no torch C++ extension): the kernels take raw device pointers and a
``cudaStream_t``, so a launch costs ~1 us of Python and is capturable in CUDA graphs.  On a machine
with a GPU a missing/unbuildable CUDA library is a hard error — there is no silent eager fallback.
"""
from __future__ import annotations

import ctypes as C
import os
import threading
from typing import Optional

import torch

from ..utils import build as _build

_lock = threading.Lock()
_cuda: Optional[C.CDLL] = None
_cpu: Optional[C.CDLL] = None
_rt: Optional[C.CDLL] = None

vp, i32, i64, f32, sz = C.c_void_p, C.c_int, C.c_int64, C.c_float, C.c_size_t

_CUDA_SIGS = {
    "cs_version": (i32, []),
    "cs_device_is_sm100": (i32, []),
    "cs_gather_rows": (i32, [vp, vp, vp, vp, vp, i32, i32, i32, vp]),
    "cs_tiny_ecg_smem_bytes": (sz, [i32, i32, i32]),
    "cs_tiny_ecg_scratch_floats": (sz, [i32]),
    "cs_tiny_ecg_forward": (i32, [vp, vp, i32, i32, i32, i32, vp, vp, i32, vp]),
    "cs_tiny_ecg_backward": (i32, [vp, vp, i32, i32, i32, i32, vp, vp, vp, i32, vp]),
    "cs_tiny_ecg_train_step": (i32, [vp, vp, vp, i32, i32, i32, i32, vp, vp, vp, vp, vp, f32, f32, f32,
                                     i32, i32, f32, f32, i32, vp]),
    "cs_tiny_ecg_round_scratch_floats": (sz, [i32]),
    "cs_perm_index": (C.c_uint64, [C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint64]),
    "cs_tiny_ecg_train_round": (i32, [vp, vp, vp, i32, i64, i32, i32, i32, i32, i32, C.c_uint64, vp, vp, vp, vp, vp,
                                      f32, f32, f32, i32, i32, f32, f32, i32, vp, vp, f32, i32, f32, i32, sz, i32, vp]),
    "cs_tiny_ecg_train_round_local": (i32, [C.POINTER(vp), i32, C.POINTER(vp), C.POINTER(vp), C.POINTER(vp), i32,
                                            C.POINTER(i64), i32, i32, i32, i32, i32, C.POINTER(C.c_uint64),
                                            C.POINTER(vp), C.POINTER(vp), C.POINTER(vp), C.POINTER(vp), C.POINTER(vp),
                                            f32, f32, f32, i32, i32, f32, f32, i32, C.POINTER(vp), C.POINTER(f32),
                                            i32, f32, i32, sz, i32, vp]),
    "cs_cross_entropy_fwd_bwd": (i32, [vp, vp, i32, i32, vp, vp, vp, vp]),
    "cs_sgd_momentum": (i32, [vp, vp, vp, i64, f32, f32, f32, i32, vp, vp, f32, f32, i32, vp, vp]),
    "cs_conv1d_fwd": (i32, [vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, i32, vp]),
    "cs_conv1d_dgrad": (i32, [vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, vp]),
    "cs_conv1d_fwd_bn": (i32, [vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, i32, vp, vp, vp, vp, vp, vp, f32, f32, vp, vp, vp, vp]),
    "cs_conv1d_plan": (i32, [i32, i32, i32, i32, i32, i32, i32, vp]),
    "cs_conv1d_dgrad_direct": (i32, [vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, vp]),
    "cs_conv1d_prepare_dgrad_weight": (i32, [vp, vp, i32, i32, i32, vp]),
    "cs_conv1d_wgrad": (i32, [vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, vp]),
    "cs_conv1d_cin1_fwd": (i32, [vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, vp]),
    "cs_conv1d_cin1_wgrad": (i32, [vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, vp]),
    "cs_bn_stats": (i32, [vp, vp, i64, i32, vp]),
    "cs_bn_apply": (i32, [vp, vp, vp, vp, vp, vp, vp, i64, i32, i32, vp]),
    "cs_bn_bwd_reduce": (i32, [vp, vp, vp, vp, vp, vp, vp, i64, i32, i32, vp]),
    "cs_bn_bwd_apply": (i32, [vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, i64, i32, i32, vp]),
    "cs_bn_finalize": (i32, [vp, vp, vp, vp, vp, i32, i64, f32, f32, vp]),
    "cs_bn_stats_finalize": (i32, [vp, vp, vp, vp, vp, vp, vp, i64, i32, f32, f32, vp, vp, vp, vp]),
    "cs_bn_bwd_reduce_ws": (i32, [vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, i64, i32, i32, vp]),
    "cs_bn_apply_coef": (i32, [vp, vp, vp, vp, i64, i32, i32, vp]),
    "cs_bn_bwd_apply_coef": (i32, [vp, vp, vp, vp, vp, vp, i64, i32, i32, vp]),
    "cs_avgpool2_fwd": (i32, [vp, vp, i32, i32, i32, vp]),
    "cs_avgpool2_bwd": (i32, [vp, vp, i32, i32, i32, vp]),
    "cs_avgpool_fwd": (i32, [vp, vp, i32, i32, i32, vp]),
    "cs_avgpool_bwd": (i32, [vp, vp, i32, i32, i32, vp]),
    "cs_ncl_f32_to_nlc_bf16": (i32, [vp, vp, i32, i32, i32, vp]),
    "cs_nlc_bf16_to_ncl_f32": (i32, [vp, vp, i32, i32, i32, vp]),
    "cs_comm_create": (i32, [i32, i32, sz, C.POINTER(vp), vp, vp]),
    "cs_comm_connect": (i32, [vp, vp, vp]),
    "cs_comm_create_local": (i32, [i32, sz, C.POINTER(vp)]),
    "cs_comm_create_external": (i32, [i32, i32, sz, C.POINTER(vp), C.POINTER(vp), C.POINTER(vp)]),
    "cs_comm_signal_bytes": (sz, []),
    "cs_vmm_support": (i32, []),
    "cs_vmm_alloc": (i32, [sz, i32, i32, C.POINTER(vp), C.POINTER(sz), C.POINTER(C.c_ulonglong), C.POINTER(i32)]),
    "cs_vmm_import": (i32, [i32, sz, C.POINTER(vp), C.POINTER(C.c_ulonglong)]),
    "cs_vmm_free": (i32, [vp, sz, C.c_ulonglong]),
    "cs_vmm_mc_create": (i32, [i32, sz, C.POINTER(C.c_ulonglong), C.POINTER(i32)]),
    "cs_vmm_mc_import": (i32, [i32, C.POINTER(C.c_ulonglong)]),
    "cs_vmm_mc_add_device": (i32, [C.c_ulonglong]),
    "cs_vmm_mc_bind_map": (i32, [C.c_ulonglong, C.c_ulonglong, sz, C.POINTER(vp)]),
    "cs_vmm_close_fd": (i32, [i32]),
    "cs_comm_arena": (vp, [vp]),
    "cs_comm_peer_arena": (vp, [vp, i32]),
    "cs_comm_arena_bytes": (sz, [vp]),
    "cs_comm_destroy": (i32, [vp]),
    "cs_comm_status": (i32, [vp, C.POINTER(i32)]),
    "cs_comm_set_timeout_ms": (i32, [vp, C.c_longlong]),
    "cs_comm_clear_status": (i32, [vp]),
    "cs_comm_set_max_blocks": (i32, [vp, i32]),
    "cs_debug_fedavg_aggregate_alone": (i32, [vp, vp, vp, i64, f32, vp, vp]),
    "cs_comm_barrier": (i32, [vp, i32, vp]),
    "cs_fedavg_arena_bytes": (sz, [i64, i32, i32]),
    "cs_fedavg_aggregate": (i32, [vp, vp, vp, i64, f32, i32, f32, vp, sz, i32, vp]),
    "cs_fedavg_aggregate_local": (i32, [C.POINTER(vp), i32, C.POINTER(vp), C.POINTER(vp), i64,
                                        C.POINTER(f32), i32, f32, C.POINTER(vp), sz, i32, vp]),
    "cs_fedavg_aggregate_inplace": (i32, [vp, sz, vp, i64, C.POINTER(f32), f32, vp, i32, vp]),
    "cs_fedavg_aggregate_inplace_local": (i32, [C.POINTER(vp), i32, sz, C.POINTER(vp), i64, C.POINTER(f32), f32,
                                                C.POINTER(vp), i32, vp]),
    "cs_fedavg_aggregate_nvls": (i32, [vp, vp, vp, vp, i64, f32, f32, f32, vp, i32, vp]),
    "cs_fedavg_bucket": (i32, [vp, sz, vp, vp, vp, vp, i64, C.POINTER(f32), f32, f32, f32, f32, f32, i32, i32, i32, vp]),
    "cs_fedavg_bucket_local": (i32, [C.POINTER(vp), i32, sz, C.POINTER(vp), C.POINTER(vp), C.POINTER(vp), C.POINTER(vp), i64,
                                     C.POINTER(f32), f32, f32, f32, f32, f32, i32, i32, vp]),
    "cs_allreduce_bucket": (i32, [vp, vp, i64, f32, i32, sz, i32, vp]),
    "cs_allreduce_bucket_local": (i32, [C.POINTER(vp), i32, C.POINTER(vp), i64, f32, i32, sz, i32, vp]),
    "cs_stage_h2d": (i32, [vp, vp, sz, vp]),
    "cs_stage_d2h": (i32, [vp, vp, sz, vp]),
    "cs_conv1d_batch_valid": (i32, [vp, vp, vp, i32, i32, i32, vp]),
    "cs_zscore_rows": (i32, [vp, vp, i32, i32, i32, vp]),
}


class NativeError(RuntimeError):
    pass


def _load(path: str, builder) -> C.CDLL:
    if not os.path.exists(path) or os.environ.get("CROSSSCALE_REBUILD") == "1":
        builder()
    return C.CDLL(path, mode=C.RTLD_GLOBAL)


def cuda_lib() -> C.CDLL:
    """The sm_100a kernel library; builds it in-tree on first use when nvcc is present."""
    global _cuda
    if _cuda is None:
        with _lock:
            if _cuda is None:
                lib = _load(_build.cuda_lib_path(), _build.build_cuda)
                lib.cs_last_error.restype = C.c_char_p
                for name, (res, args) in _CUDA_SIGS.items():
                    fn = getattr(lib, name, None)
                    if fn is None:  # optional entry points added later keep older .so files loadable
                        continue
                    fn.restype, fn.argtypes = res, args
                _cuda = lib
    return _cuda


def cpu_lib() -> C.CDLL:
    global _cpu
    if _cpu is None:
        with _lock:
            if _cpu is None:
                lib = _load(_build.cpu_lib_path(), _build.build_cpu)
                lib.conv1d_batch_omp_simd.restype = None
                lib.conv1d_batch_omp_simd.argtypes = [vp, vp, vp, i32, i32, i32, i32]
                lib.cs_cpu_conv1d_isa.restype = i32
                _cpu = lib
    return _cpu


def rt_lib() -> C.CDLL:
    global _rt
    if _rt is None:
        with _lock:
            if _rt is None:
                lib = _load(_build.rt_lib_path(), _build.build_rt)
                lib.cs_rt_last_error.restype = C.c_char_p
                lib.cs_prefetcher_create.restype = vp
                lib.cs_prefetcher_create.argtypes = [C.POINTER(C.c_char_p), i32, i32, i32, C.POINTER(vp),
                                                     C.POINTER(vp), i32, i32, i32]
                lib.cs_prefetcher_window_len.restype = i32
                lib.cs_prefetcher_window_len.argtypes = [vp]
                lib.cs_prefetcher_start.restype = i32
                lib.cs_prefetcher_start.argtypes = [vp]
                lib.cs_prefetcher_next.restype = i32
                lib.cs_prefetcher_next.argtypes = [vp, C.POINTER(i32), C.POINTER(i32), C.POINTER(C.c_double)]
                lib.cs_prefetcher_recycle.restype = i32
                lib.cs_prefetcher_recycle.argtypes = [vp, i32]
                lib.cs_prefetcher_destroy.restype = None
                lib.cs_prefetcher_destroy.argtypes = [vp]
                lib.cs_rendezvous_allgather.restype = i32
                lib.cs_rendezvous_allgather.argtypes = [C.c_char_p, C.c_char_p, i32, i32, vp, sz, vp, C.c_double]
                lib.cs_rendezvous_cleanup.restype = i32
                lib.cs_rendezvous_cleanup.argtypes = [C.c_char_p, C.c_char_p, i32]
                lib.cs_rendezvous_last_error.restype = C.c_char_p
                _rt = lib
    return _rt


CS_OK, CS_ERR_INVALID, CS_ERR_CUDA, CS_ERR_UNSUPPORTED = 0, 1, 2, 3      # include/crossscale.h


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = cuda_lib().cs_last_error()
        raise NativeError(f"{what or 'native call'} failed (code {rc}): {msg.decode() if msg else ''}")


def have_gpu() -> bool:
    return torch.cuda.is_available()


def stream_ptr(stream: Optional[torch.cuda.Stream] = None) -> int:
    return (stream or torch.cuda.current_stream()).cuda_stream


def ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


def is_sm100() -> bool:
    return have_gpu() and bool(cuda_lib().cs_device_is_sm100())
