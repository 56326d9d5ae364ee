"""ctypes binding of libkafka_b200.so (declarations: include/kafka_b200.h).

There is NO fallback: if the library is missing the import of any compute path
raises, and every non-zero status from the library raises ``KafkaCudaError``.
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("KAFKA_B200_LIB") or os.path.join(HERE, "csrc", "libkafka_b200.so")

MAX_PARAMS = 16
MAX_BANDS = 16
MAX_DIM = 16
BAND_ARRAYS = 0
BAND_LINEAR = 1
BAND_UNIFORM = 2
BAND_WCM = 3
RAW_I16, RAW_U16, RAW_F32, RAW_F64 = 0, 1, 2, 3
ROUND_F32 = 1
PRIOR_BROADCAST = 2
SOLVE_PAIR = 4
SOLVE_COUNT_LU = 8
COMM_ID_BYTES = 128
STATIC_SCRATCH_DOUBLES = 524288
STATIC_MAX_PIXELS = 16758144


class KafkaCudaError(RuntimeError):
    pass


class GpBand(C.Structure):
    _fields_ = [("table", C.c_void_p), ("n_train", C.c_int32), ("n_dim", C.c_int32),
                ("mapper", C.c_int32 * MAX_DIM)]


class SolveBand(C.Structure):
    _fields_ = [("kind", C.c_int32), ("n_dim", C.c_int32), ("mapper", C.c_int32 * MAX_DIM),
                ("row", C.c_double * MAX_PARAMS), ("offset", C.c_double),
                ("h0", C.c_void_p), ("jac", C.c_void_p), ("y", C.c_void_p), ("w", C.c_void_p)]


_p, _i64, _i32, _u32 = C.c_void_p, C.c_int64, C.c_int, C.c_uint32

# name -> (restype, argtypes).  Every symbol declared in include/kafka_b200.h.
SIGNATURES = {
    "kafka_last_error": (C.c_char_p, []),
    "kafka_version": (C.c_int, []),
    "kafka_compact_scratch_elems": (_i64, [_i64]),
    "kafka_compact_mask": (C.c_int, [_p, _i64, _p, _i64, _p, C.POINTER(C.c_int64), _p]),
    "kafka_gather_band": (C.c_int, [_p, _p, _p, _p, _i64, _p, _p, _p, _p]),
    "kafka_gather_bands": (C.c_int, [_p, _p, _p, _i64, _i32, _p, _i64, _i64, _p, _p, _p, _p]),
    "kafka_gather_bhr_band": (C.c_int, [_p, _i32, _i64, _p, _p, _p, _i64, _p, _p, _p, _p]),
    "kafka_gather_backscatter_band": (C.c_int, [_p, _i32, _p, _i64, C.c_double, C.c_double, _p, _p, _p, _p]),
    "kafka_gather_scaled_band": (C.c_int, [_p, _i32, _p, _p, _i64, C.c_double, C.c_double, _p, _p, _p, _p]),
    "kafka_scatter_output": (C.c_int, [_p, _p, _i64, _i64, _i32, _p, _i64, _p, _p, _p]),
    "kafka_output_rasters": (C.c_int, [_p, _p, _i64, _i64, _i32, _p, _i64, _p, _p, _i32, _p, _p]),
    "kafka_state_uniform": (C.c_int, [_p, _i64, _i64, _i32, _p, _p]),
    "kafka_read_small": (C.c_int, [_p, _p, _i64, _p]),
    "kafka_interleaved_to_planar": (C.c_int, [_p, _i64, _i32, _p, _i64, _p]),
    "kafka_planar_to_interleaved": (C.c_int, [_p, _i64, _i64, _i32, _p, _p]),
    "kafka_blocks_to_packed": (C.c_int, [_p, _i64, _i32, _p, _i64, _p]),
    "kafka_packed_to_blocks": (C.c_int, [_p, _i64, _i64, _i32, _p, _p]),
    "kafka_packed_diagonal": (C.c_int, [_p, _i64, _i64, _i32, _p, _p]),
    "kafka_gp_table_doubles": (_i64, [_i32, _i32]),
    "kafka_gp_pack_host": (C.c_int, [_p, _p, _p, _i32, _i32, _p]),
    "kafka_gp_linearize": (C.c_int, [_p, _i64, _i64, C.POINTER(GpBand), _i32, _p, _p, _p, _u32, _p]),
    "kafka_gp_predict": (C.c_int, [_p, _i64, _i64, _p, _i32, _i32, _p, _p, _p, _p, _p]),
    "kafka_gp_hessian_correction": (C.c_int, [_p, _i64, _i64, _i32, C.POINTER(GpBand), _i32, _p, _p, _p, _p, _p]),
    "kafka_gp_hessian": (C.c_int, [_p, _i64, _i64, _p, _i32, _i32, _p, _p]),
    "kafka_solve_scratch_doubles": (_i64, [_i64]),
    "kafka_solve_bands": (C.c_int, [_i64, _i64, _i32, C.POINTER(SolveBand), _i32, _p, _p, _p, _p, _p, _p, _p,
                                    _p, _p, _p, _u32, _p]),
    "kafka_solve_bands_static": (C.c_int, [_i64, _i64, _i32, C.POINTER(SolveBand), _i32, _p, _p, _p, _p, _p, _p, _p,
                                           _p, _p, _u32, _p]),
    "kafka_assimilate_iter": (C.c_int, [_i64, _i64, _i32, C.POINTER(GpBand), _i32, _p, _p, _p, _p, _p, _p, _p, _p,
                                        _p, _p, _p, _p, _p, _p, _p, _u32, _p]),
    "kafka_assimilate_date": (C.c_int, [_i64, _i64, _i32, C.POINTER(GpBand), _i32, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p,
                                        _p, _p, _p, _p, C.c_double, _i32, _u32, _p, _p, _p, _p, _p, _p, _p, _p, _p]),
    "kafka_upload_rasters": (C.c_int, [_i32, _p, _p, _p, _p]),
    "kafka_wcm_forward": (C.c_int, [_p, _i64, _i64, _p, _p, _p, _p]),
    "kafka_advance_carry_slot": (C.c_int, [_p, _p, _p, _i64, _i64, _i32, _i32, _p, _p, _p, _p, _p]),
    "kafka_advance_approx": (C.c_int, [_p, _p, _i64, _i64, _i32, _p, _p]),
    "kafka_advance_exact": (C.c_int, [_p, _p, _i64, _i64, _i32, _p, _p, _p]),
    "kafka_advance_standard": (C.c_int, [_p, _p, _p, _p, _i64, _i64, _i32, _p, _p, _p]),
    "kafka_blend_prior": (C.c_int, [_p, _p, _p, _p, _i64, _i64, _i32, _p, _p, _p, _u32, _p]),
    "kafka_comm_unique_id": (C.c_int, [_p]),
    "kafka_ctx_create": (C.c_int, [_i32, _i32, _p, C.POINTER(C.c_void_p)]),
    "kafka_ctx_destroy": (C.c_int, [_p]),
    "kafka_ctx_info": (C.c_int, [_p, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "kafka_allreduce_norm": (C.c_int, [_p, _p, _i32, _p]),
    "kafka_gather_outputs": (C.c_int, [_p, _p, C.POINTER(C.c_int64), _p, _p]),
}

_lib = None


def get_lib():
    """Load the CUDA library or raise (no CPU path exists)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            "libkafka_b200.so is not built (%s). Run `python -m kafka_inferenceengine_b200.build` "
            "(needs nvcc); this package has no CPU fallback." % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)          # AttributeError if the symbol is missing
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


# kernels launched by one successful call of each entry point (for bench.py's
# gpu_launches count; the set of kernels is in csrc/*.cu)
KERNELS_PER_CALL = {"kafka_compact_mask": 3, "kafka_solve_bands": 2, "kafka_solve_bands_static": 3,
                    "kafka_assimilate_iter": 3, "kafka_assimilate_date": 0, "kafka_upload_rasters": 0,     # counted by the caller: 4 per solve
                    # NCCL's kernels, not this library's
                    "kafka_comm_unique_id": 0, "kafka_ctx_create": 0, "kafka_ctx_destroy": 0, "kafka_ctx_info": 0,
                    "kafka_allreduce_norm": 0, "kafka_gather_outputs": 0}
launch_count = 0


def check(status, what):
    global launch_count
    if status == 0 and not what.endswith("_host"):
        name = what.split("(")[0]
        launch_count += 2 if name == "kafka_compact_mask" and what.endswith("(count)") else \
            KERNELS_PER_CALL.get(name, 1)
    if status != 0:
        msg = get_lib().kafka_last_error().decode("utf-8", "replace")
        raise KafkaCudaError("%s failed with status %d: %s" % (what, status, msg))


def ptr(t):
    """Device pointer of a torch tensor (or None)."""
    if t is None:
        return None
    return C.c_void_p(t.data_ptr())


def np_ptr(a):
    return C.c_void_p(a.ctypes.data)


HOST_HOOK = C.CFUNCTYPE(None, C.c_void_p)      # kafka_host_hook


def stream_ptr():
    """The current CUDA stream of the current device as a void*."""
    import torch
    try:
        return C.c_void_p(torch._C._cuda_getCurrentRawStream(torch._C._cuda_getDevice()))
    except AttributeError:                      # a torch without these helpers
        return C.c_void_p(torch.cuda.current_stream().cuda_stream)


_small = {}          # device index -> (pinned uint8 tensor, event)


def read_small(t):
    """Host copy (numpy, same dtype) of a SMALL contiguous CUDA tensor whose byte size is a
    multiple of 8, through kafka_read_small: one tiny kernel on the current stream + an event
    wait, no device->host memcpy (which would queue behind large transfers in flight)."""
    import torch
    nbytes = t.numel() * t.element_size()
    if not (t.is_cuda and t.is_contiguous() and nbytes % 8 == 0 and 0 < nbytes <= 4096):
        raise ValueError("read_small needs a contiguous CUDA tensor of 8..4096 bytes (multiple of 8)")
    key = t.device.index
    if key not in _small:
        _small[key] = (torch.empty(4096, dtype=torch.uint8).pin_memory(), torch.cuda.Event())
    host, event = _small[key]
    with torch.cuda.device(t.device):
        check(get_lib().kafka_read_small(ptr(t), C.c_void_p(host.data_ptr()), nbytes, stream_ptr()), "kafka_read_small")
        event.record(torch.cuda.current_stream())
    event.synchronize()
    return host[:nbytes].numpy().view(np.dtype(str(t.dtype).replace("torch.", ""))).copy().reshape(tuple(t.shape))


def gp_pack(inputs, theta, invqt):
    """Host-side table for one emulator (kafka_gp_pack_host)."""
    lib = get_lib()
    inputs = np.ascontiguousarray(inputs, dtype=np.float64)
    theta = np.ascontiguousarray(theta, dtype=np.float64)
    invqt = np.ascontiguousarray(invqt, dtype=np.float64)
    T, D = inputs.shape
    if theta.shape != (D + 2,) or invqt.shape != (T,):
        raise ValueError("emulator fields have inconsistent shapes")
    n = lib.kafka_gp_table_doubles(T, D)
    if n < 0:
        raise ValueError("emulator with T=%d, D=%d is not supported (D <= %d)" % (T, D, MAX_DIM))
    table = np.empty(n, dtype=np.float64)
    check(lib.kafka_gp_pack_host(np_ptr(inputs), np_ptr(theta), np_ptr(invqt), T, D, np_ptr(table)),
          "kafka_gp_pack_host")
    return table
