"""ctypes binding of libper_b200.so (C ABI declared in include/per_b200.h).

The reference has no native layer to bind (it is pure Python, memory.py), so this file
plays the part the cgo/JNI stub would play elsewhere: it loads the shared library that
sits next to it and declares every entry point's signature.  There is NO fallback: if
the library has not been built (python -c "import __graft_entry__ as g; g.build()" at
the repo root) importing this module raises, and every product class fails with it.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("PER_B200_LIB") or os.path.join(_HERE, "libper_b200.so")  # override: A/B runs of two builds

PER_OK, PER_ERR_ASSERT, PER_ERR_CUDA, PER_ERR_ARG, PER_ERR_NOMEM = 0, -1, -2, -3, -4
STATUS_BAD_PRIORITY, STATUS_BAD_INDEX, STATUS_BAD_PREFIXSUM, STATUS_NSTEP_EMPTY = 1, 2, 4, 8
STATUS_EXCHANGE_TIMEOUT = 16
TREE_SUM, TREE_MIN = 0, 1
SEG_NOALLOC_MAX_CAPACITY = 1 << 20
PRIO_F64, PRIO_F32 = 0, 1

if not os.path.exists(LIB_PATH):
    raise ImportError(
        "libper_b200.so is missing at %s — build it first (repo root: "
        "python -c 'import __graft_entry__ as g; g.build()').  This package has no CPU fallback." % LIB_PATH)

lib = C.CDLL(LIB_PATH)

_p = C.c_void_p
_i64 = C.c_int64
_f64 = C.c_double

# name -> (restype, argtypes); mirrors include/per_b200.h one to one
SIGNATURES = {
    "per_last_error": (C.c_char_p, []),
    "per_version": (C.c_int, []),
    "seg_create": (C.c_int, [C.POINTER(_p), C.c_int, _i64, C.c_int]),
    "seg_destroy": (C.c_int, [_p]),
    "seg_set": (C.c_int, [_p, _p, _p, _i64, _p]),
    "seg_reduce": (C.c_int, [_p, _i64, _i64, _p, _p]),
    "seg_find_prefixsum": (C.c_int, [_p, _p, _i64, _p, _p]),
    "seg_read": (C.c_int, [_p, _p, _p]),
    "seg_get": (C.c_int, [_p, _i64, C.POINTER(_f64), _p]),
    "per_tree_get": (C.c_int, [_p, C.c_int, _i64, C.POINTER(_f64), _p]),
    "per_tree_reduce": (C.c_int, [_p, C.c_int, _i64, _i64, _p, _p]),
    "seg_capacity": (_i64, [_p]),
    "seg_status": (C.c_int, [_p, C.POINTER(_i64), _p]),
    "aql_create": (C.c_int, [C.POINTER(_p), C.c_int, C.c_int, C.c_int]),
    "aql_destroy": (C.c_int, [_p]),
    "aql_set_weights": (C.c_int, [_p, _p, _p, _p, _p, _p, _p, _p, _p, _p]),
    "aql_trace": (C.c_int, [_p, C.POINTER(_i64), _p]),
    "per_timeline": (C.c_int, [_p, C.POINTER(_i64), _i64, C.POINTER(_i64), _p]),
    "per_sample_host": (C.c_int, [_p, _p, _i64, C.c_double, _p, _p, _p, _p, _p, _p, _p, _p, _p]),
    "per_update_host": (C.c_int, [_p, _p, _p, _i64, C.POINTER(C.c_double), _p]),
    "aql_candidate_q": (C.c_int, [_p, _p, _p, _p, _p, _i64, _i64, _p, _p, _p]),
    "nstep_assemble": (C.c_int, [C.c_int, _i64, _p, _p, _p, _p, _p, _p, _i64, C.c_int, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p]),
    "seg_set_noalloc": (C.c_int, [C.c_int, _p, _p, _i64, _p, _p, _i64, _p, _p]),
    "per_create": (C.c_int, [C.POINTER(_p), C.c_int, _i64, _f64, _i64, _i64]),
    "per_destroy": (C.c_int, [_p]),
    "per_create_shard": (C.c_int, [C.POINTER(_p), C.c_int, _i64, C.c_int, C.c_int, _f64, _i64, _i64]),
    "per_shard_record": (C.c_int, [_p, _i64, _p, _p]),
    "per_shard_attach_record": (C.c_int, [_p, _p, _i64]),
    "per_ipc_alloc": (C.c_int, [C.c_int, _i64, C.POINTER(_p), _p]),
    "per_ipc_open": (C.c_int, [C.c_int, _p, C.POINTER(_p)]),
    "per_ipc_close": (C.c_int, [C.c_int, _p]),
    "per_ipc_free": (C.c_int, [C.c_int, _p]),
    "per_shard_attach_exchange": (C.c_int, [_p, _p, _p]),
    "per_shard_publish": (C.c_int, [_p, _i64, _p]),
    "per_shard_deliver": (C.c_int, [_p, _p, _p, _i64, _p, _p, _p]),
    "per_shard_exchange_stats": (C.c_int, [_p, C.POINTER(_i64), _p]),
    "per_sample_sharded": (C.c_int, [_p, _p, _p, _i64, _f64, _i64, _p, _p, _p, _p]),
    "per_capacity": (_i64, [_p]),
    "per_len": (_i64, [_p]),
    "per_cursor": (_i64, [_p]),
    "per_set_cursor": (C.c_int, [_p, _i64, _f64, _p]),
    "per_max_priority": (C.c_int, [_p, C.POINTER(_f64), _p]),
    "per_status": (C.c_int, [_p, C.POINTER(_i64), _p]),
    "per_add_batch": (C.c_int, [_p, _p, _p, _p, _p, _p, _p, _p, _p, _f64, _i64, _p]),
    "per_sample_indices": (C.c_int, [_p, _p, _i64, _f64, _p, _p, _p, _p]),
    "per_gather": (C.c_int, [_p, _p, _i64, _p, _p, _p, _p, _p, _p, _p]),
    "per_gather_cast": (C.c_int, [_p, _p, _i64, C.c_int, C.c_float, _p, _p, _p, _p, _p, _p]),
    "per_gather_cast_nhwc": (C.c_int, [_p, _p, _i64, C.c_int, C.c_float, C.c_int, _p, _p, _p, _p, _p, _p]),
    "per_update_priorities": (C.c_int, [_p, _p, _p, C.c_int, _i64, _p]),
    "per_update_leaves": (C.c_int, [_p, _p, _p, _f64, _i64, _p]),
    "per_pow_alpha_host": (C.c_int, [_p, _f64, _p, _i64]),
    "per_td_priority_update": (C.c_int, [_p, _p, _p, _p, _i64, _p, _p, _p, _p, C.c_float, _p, _i64, _p, _p, _p, _p]),
    "per_tree_read": (C.c_int, [_p, C.c_int, _p, _p]),
    "per_tune": (C.c_int, [_p, C.c_char_p, _i64]),
    "per_update_trace": (C.c_int, [_p, C.POINTER(_i64), _p]),
}

for _name, (_res, _args) in SIGNATURES.items():
    _fn = getattr(lib, _name)  # AttributeError here = header and library disagree
    _fn.restype = _res
    _fn.argtypes = _args


class NativeError(RuntimeError):
    pass


def check(rc):
    """0 passes; PER_ERR_ASSERT becomes AssertionError (what the reference raises at
    memory.py:34,224,284,313); anything else is a NativeError."""
    if rc == PER_OK:
        return
    msg = (lib.per_last_error() or b"").decode("utf-8", "replace")
    if rc == PER_ERR_ASSERT:
        raise AssertionError(msg)
    if rc == PER_ERR_NOMEM:
        raise MemoryError(msg)
    raise NativeError("libper_b200 error %d: %s" % (rc, msg))


def ptr(t):
    """Device (or pinned-host) address of a torch tensor, or None."""
    if t is None:
        return None
    assert t.is_contiguous(), "the C ABI takes dense buffers"
    return C.c_void_p(t.data_ptr())


def stream_ptr(stream=None):
    import torch
    s = stream if stream is not None else torch.cuda.current_stream()
    return C.c_void_p(s.cuda_stream)
