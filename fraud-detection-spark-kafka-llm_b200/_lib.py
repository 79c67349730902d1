"""ctypes binding of csrc/libfdb200.so (declared in include/fdb200.h).

There is NO CPU fallback: if the library is missing or a call fails the error is raised."""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("FDB200_LIB") or os.path.join(_HERE, "csrc", "libfdb200.so")  # FDB200_LIB: another build of the same ABI (A/B timing)

FD_OK, FD_ERR_INVALID, FD_ERR_CAPACITY, FD_ERR_CUDA, FD_ERR_STATE, FD_ERR_LIMIT = 0, -1, -2, -3, -4, -5
FD_FLAG_CLEAN = 1
FD_KIND_DT, FD_KIND_RF, FD_KIND_GBT = 1, 2, 3

_p = C.c_void_p
_i32, _i64, _u32, _f64 = C.c_int32, C.c_int64, C.c_uint32, C.c_double

# name -> (restype, argtypes); kept in the order of include/fdb200.h
SIGNATURES = {
    "fd_create": (C.c_int, [C.c_int, C.POINTER(_p)]),
    "fd_destroy": (C.c_int, [_p]),
    "fd_last_error": (C.c_char_p, [_p]),
    "fd_version": (C.c_int, []),
    "fd_set_stopwords": (C.c_int, [_p, _p, _p, _i32, _i32]),
    "fd_set_hashing": (C.c_int, [_p, _i32, _i32]),
    "fd_set_vocabulary": (C.c_int, [_p, _p, _p, _i32, _i32]),
    "fd_vocab_count": (C.c_int, [_p, _p, _p, _i64, _u32, _p, _p, _i32, _p, _i64, _p, _p]),
    "fd_hash_terms": (C.c_int, [_p, _p, _i32, _i32, _p]),
    "fd_doc_freq": (C.c_int, [_p, _p, _p, _i64, _i32, _p, _p]),
    "fd_set_idf": (C.c_int, [_p, _p, _i32]),
    "fd_set_linear": (C.c_int, [_p, _p, _i32, _f64, _f64]),
    "fd_set_forest": (C.c_int, [_p, _i32, _i32, _p, _p, _p, _p, _p, _p, _p, _i32]),
    "fd_featurize_begin": (C.c_int, [_p, _p, _p, _i64, _u32, C.POINTER(_i64), _p]),
    "fd_featurize_finish": (C.c_int, [_p, _p, _p, _p, _i64, _p]),
    "fd_featurize": (C.c_int, [_p, _p, _p, _i64, _u32, _p, _p, _p, _i64, C.POINTER(_i64), _p]),
    "fd_score": (C.c_int, [_p, _p, _p, _p, _i64, _p, _p, _p, _p]),
    "fd_transform": (C.c_int, [_p, _p, _p, _i64, _u32, _p, _p, _p, _p]),
    "fd_status": (C.c_int, [_p, _p]),
    "fd_transform_host": (C.c_int, [_p, _p, _p, _i64, _u32, _p, _p, _p]),
    "fd_transform_noalloc": (C.c_int, [_p, _p, _p, _i64, _u32, _p, _p, _p, _p, _p]),
    "fd_eval_auc_roc": (C.c_int, [_p, _p, _i64, _p, _i64, _i32, _p, _p]),
    "fd_eval_confusion2": (C.c_int, [_p, _p, _p, _i64, _p, _p]),
    "fd_idf_scale": (C.c_int, [_p, _p, _p, _i64, _p, _i32, _p]),
    "fd_workspace_bytes": (_i64, [_p, _i64]),
    "fd_csc_scratch_bytes": (_i64, [_p, _i64, _i32]),
    "fd_reserve": (C.c_int, [_p, _i64]),
    "fd_launch_count": (_i64, [_p]),
    "fd_bin_csr": (C.c_int, [_p, _p, _p, _i64, _p, _p, _p, _p]),
    "fd_csc_build": (C.c_int, [_p, _p, _p, _p, _p, _i64, _i64, _i32, _p, _p, _p, _i64, _p]),
    "fd_rf_histogram": (C.c_int, [_p, _p, _p, _p, _p, _i64, _i32, _p, _p, _p, _p, _i32, _p, _p, _i32, _p, _p]),
    "fd_rf_best_split": (C.c_int, [_p, _p, _i32, _p, _p, _p, _p, _p, _i32, _f64, _p, _p, _p, _p, _p, _p]),
    "fd_rf_bag_poisson": (C.c_int, [_p, C.c_uint64, _f64, _i32, _i64, _p, _p, _p, _p]),
    "fd_rf_advance": (C.c_int, [_p, _p, _p, _p, _p, _i64, _i32, _i32, _i32, _p, _p, _p, _p, _p, _i32, _p, _p, _i32, _p, _p]),
}

_lib = None


class FdError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"libfdb200 error {code}: {msg}")
        self.code = code


def load_library():
    """dlopen the in-tree library and declare every entry point; raises if it is not built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(f"{LIB_PATH} is missing: build it with `python __graft_entry__.py` "
                           "(nvcc, sm_100a). There is no CPU fallback for this path.")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        if not hasattr(lib, name):
            raise RuntimeError(f"{LIB_PATH} does not export {name}: it was built from another version of the sources; "
                               "rebuild it with `python __graft_entry__.py`")
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def ptr(t):
    """data pointer of a torch tensor / numpy array / None as c_void_p."""
    if t is None:
        return None
    if hasattr(t, "data_ptr"):
        return C.c_void_p(t.data_ptr())
    return C.c_void_p(t.ctypes.data)


class Handle:
    """Owns one fd_handle (one device)."""

    def __init__(self, device: int = 0):
        self.lib = load_library()
        h = _p()
        rc = self.lib.fd_create(int(device), C.byref(h))
        if rc != FD_OK:
            raise FdError(rc, f"fd_create(device={device}) failed (is a CUDA device visible?)")
        self.h = h
        self.device = int(device)

    def check(self, rc):
        if rc != FD_OK:
            raise FdError(rc, (self.lib.fd_last_error(self.h) or b"").decode("utf-8", "replace"))

    def call(self, name, *args):
        self.check(getattr(self.lib, name)(self.h, *args))

    def launch_count(self) -> int:
        return int(self.lib.fd_launch_count(self.h))

    def close(self):
        if getattr(self, "h", None):
            self.lib.fd_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
