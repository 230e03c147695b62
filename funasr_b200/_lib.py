"""ctypes binding of libfunasr_b200.so (the C-ABI in include/funasr_b200.h).

No torch types cross the boundary: tensors are passed as raw device pointers, sizes as ints
and the stream as an opaque handle.  There is NO CPU or PyTorch fallback: if the library is
missing or a call fails, a RuntimeError is raised.
"""
import ctypes
import os
from ctypes import c_float, c_int, c_int64, c_uint32, c_void_p

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libfunasr_b200.so")

_P = c_void_p
_I = c_int
_F = c_float

# name -> argument types (all return int unless noted); mirrors include/funasr_b200.h
SIGNATURES = {
    "fb200_fbank": [_P, _P, _P, _I, _I, _P, _I, _I, _I, _P, _I, _F, _F, _P, _P],
    "fb200_lfr_cmvn": [_P, _P, _P, _I, _I, _I, _I, _I, _P, _P, _P, _F, _I, _P, _I, _P],
    "fb200_scale_add_pe": [_P, _I, _P, _I, _I, _I, _P, _F, _P, _I, _P],
    "fb200_layernorm": [_P, _I, _I, _I, _P, _P, _F, _P, _I, _P, _I, _I, _P, _I, _P],
    "fb200_split_rows": [_P, _I, _I, _I, _P, _I, _P],
    "fb200_split_rows_halo": [_P, _I, _I, _I, _P, _I, _P, _I, _P],
    "fb200_pack_rows": [_P, ctypes.c_longlong, _I, _P, _I, _I, _I, _P, _I, _P],
    "fb200_unpack_rows": [_P, _I, _P, _I, _I, _I, _F, _P, _P],
    "fb200_fsmn": [_P, _I, _P, _I, _I, _I, _P, _I, _P, _I, _P, _I, _P],
    "fb200_gemm": [_P, _I, _I, _P, _I, _I, _I, _I, _I, _P, _I, _P, _I, _P, _I, _I, _P, _I, _I, _P],
    "fb200_gemm_ex": [_P, _I, _I, _I, _I, _I, _P, _I, _I, _I, _I, _I, _P, _P, _P, _I, _P, _I, _P, _I, _I, _P, _I, _I, _P],
    "fb200_gemm_heads": [_P, _I, _I, _I, _I, _I, _I, _P, _I, _I, _I, _I, _I, _I, _P, _I, _I, _P],
    "fb200_lstm_workspace_bytes": [_I, _I, _P],
    "fb200_lstm_parts": [_I, _P],
    "fb200_lstm_bidir": [_P, _I, _P, _I, _I, _I, _I, _I, _P, _P, _P, ctypes.c_longlong, _P, _P, _I, _P],
    "fb200_lstm_bidir_static": [_P, _I, _P, _I, _I, _I, _I, _I, _P, _P, _P, _P, _I, _P],
    "fb200_us_fires": [_P, _I, _I, _F, _F, _F, _F, _P, _I, _I, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P],
    "fb200_asf_scores": [_P, _I, _P, _I, _P, _P, _I, _I, _I, _I, _F, _P, _P],
    "fb200_merge_hotword_ids": [_P, _P, _P, _P, _I, _I, _P, _P, _P],
    "fb200_conv2d_sub1": [_P, _I, _P, _P, _I, _I, _I, _P, _P, _I, _P, _I, _P],
    "fb200_gather_rows": [_P, _I, _P, _P, _I, _P, _I, _I, _I, _P],
    "fb200_glu_dwconv": [_P, _I, _P, _I, _I, _I, _I, _P, _I, _P, _P, _P, _P, _I, _P],
    "fb200_attention_relpos": [_P, _I, _I, _I, _P, _I, _I, _P, _I, _I, _I, _I, _P, _I, ctypes.c_longlong, _P, _P, _I, _I,
                               _I, _I, _I, _F, _P, _I, _P, _I, _I, _P],
    "fb200_split_weight": [_P, _I, _I, _P, _I, _P],
    "fb200_fill_head_rows": [_P, _I, _I, _P, _I, _P, _F, _P, _I, _P],
    "fb200_attention_plan": [_P, _P, _I, _P, _I, _P],
    "fb200_attention": [_P, _I, _I, _I, _P, _I, _I, _P, _I, _I, _I, _I, _P, _P, _P, _I, _I, _I, _I, _I, _I, _F, _P, _I,
                        _P, _I, _I, _P],
    "fb200_cif_alpha": [_P, _I, _P, _I, _I, _I, _P, _P, _F, _F, _P, _P],
    "fb200_cif_scan": [_P, _P, _I, _F, _P, _P, _P, _P, _P],
    "fb200_cif_integrate": [_P, _I, _P, _P, _P, _I, _I, _I, _F, _P, _P, _P, _I, _P],
    "fb200_argmax_logprob": [_P, _I, _I, _I, _P, _P, _P],
    "fb200_ctc_collapse": [_P, _P, _I, c_uint32, _I, _P, _P, _P],
    "fb200_gemm_argmax_workspace": [_I, _I, _P],
    "fb200_gemm_argmax": [_P, _I, _I, _P, _I, _I, _I, _I, _P, _P, ctypes.c_longlong, _P, _P, _P],
    "fb200_stack_begin": [_P, _P, _I],
    "fb200_stack_end": [_P],
    "fb200_stack_info": [_P, _P, _P],
    "fb200_stack_run": [_P, _P, _I, _P],
    "fb200_stack_destroy": [_P],
    "fb200_stack_profile": [_P, _I],
}
# native offline runtime (csrc/runtime.cu): name -> (argument types, return type)
RUNTIME_SIGNATURES = {
    "fb200_offline_init": ([ctypes.c_char_p, _I, _I, _I], _P),
    "fb200_offline_infer": ([_P, ctypes.POINTER(_P), ctypes.POINTER(ctypes.c_int32), _I], _P),
    "fb200_offline_infer_buffer": ([_P, ctypes.c_char_p, _I, _I], _P),
    "fb200_result_count": ([_P], _I),
    "fb200_result_text": ([_P, _I], ctypes.c_char_p),
    "fb200_result_ids": ([_P, _I, ctypes.POINTER(_I)], ctypes.POINTER(ctypes.c_int32)),
    "fb200_result_stamps": ([_P, _I, ctypes.POINTER(_I)], ctypes.POINTER(ctypes.c_int32)),
    "fb200_free_result": ([_P], None),
    "fb200_offline_uninit": ([_P], None),
}
EXPORTS = sorted(list(SIGNATURES) + list(RUNTIME_SIGNATURES) + ["fb200_version", "fb200_lstm_static_max_batch", "fb200_last_error", "fb200_launch_count",
                  "fb200_profile_enable", "fb200_profile_read", "fb200_stack_profile_read"])
PROFILE_KINDS = ["fbank", "lfr_cmvn", "scale_add_pe", "layernorm", "split_rows", "fsmn", "gemm", "attention",
                 "cif_alpha", "cif_scan", "cif_integrate", "argmax_logprob", "ctc_collapse", "conv2d_sub1", "glu_dwconv",
                 "attention_relpos", "lstm_bidir", "us_fires"]

_lib = None


def load():
    """Load the shared library (building it is `__graft_entry__.build()` / `python -m funasr_b200.build`)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            "funasr_b200: %s not found -- the CUDA extension is not built and there is no fallback path. "
            "Run `python -m funasr_b200.build`." % LIB_PATH)
    lib = ctypes.CDLL(LIB_PATH)
    for name, args in SIGNATURES.items():
        fn = getattr(lib, name)
        fn.argtypes = args
        fn.restype = c_int
    for name, (args, ret) in RUNTIME_SIGNATURES.items():
        fn = getattr(lib, name)
        fn.argtypes = args
        fn.restype = ret
    lib.fb200_version.restype = c_int
    lib.fb200_stack_profile_read.argtypes = [_P, _P, _P, _P, _I, _P]
    lib.fb200_stack_profile_read.restype = c_int  # number of ops (>= 0) or a negative error code
    lib.fb200_lstm_static_max_batch.argtypes = []
    lib.fb200_lstm_static_max_batch.restype = c_int
    lib.fb200_last_error.restype = ctypes.c_char_p
    lib.fb200_launch_count.argtypes = [c_int]
    lib.fb200_launch_count.restype = ctypes.c_longlong
    lib.fb200_profile_enable.argtypes = [c_int]
    lib.fb200_profile_enable.restype = c_int
    lib.fb200_profile_read.argtypes = [c_int, ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_longlong),
                                       ctypes.POINTER(ctypes.c_double), c_int]
    lib.fb200_profile_read.restype = c_int
    _lib = lib
    return lib


def call(name, *args):
    lib = load()
    rc = getattr(lib, name)(*args)
    if rc != 0:
        raise RuntimeError("%s failed (%d): %s" % (name, rc, lib.fb200_last_error().decode()))


def launch_count(reset: bool = False) -> int:
    return int(load().fb200_launch_count(1 if reset else 0))


def profile_enable(on: bool) -> None:
    load().fb200_profile_enable(1 if on else 0)


def profile_read(reset: bool = True) -> dict:
    """{kind: {"ms": summed device ms, "calls": n, "work": algorithmic bytes or FLOPs}} for this thread."""
    lib = load()
    out = {}
    for i, name in enumerate(PROFILE_KINDS):
        ms, calls, work = ctypes.c_double(), ctypes.c_longlong(), ctypes.c_double()
        rc = lib.fb200_profile_read(i, ctypes.byref(ms), ctypes.byref(calls), ctypes.byref(work), 1 if reset else 0)
        if rc != 0:
            raise RuntimeError("fb200_profile_read failed (%d): %s" % (rc, lib.fb200_last_error().decode()))
        if calls.value:
            out[name] = {"ms": ms.value, "calls": int(calls.value), "work": work.value}
    return out
