"""ctypes binding of libsphb200.so (the C ABI declared in include/sphb200.h).

There is NO CPU fallback: if the library is missing, cannot be loaded, or no B200 is visible, every
entry point raises.  PyTorch is used only to own device buffers and streams.
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libsphb200.so")

c_i64 = ctypes.c_int64
c_int = ctypes.c_int
c_dbl = ctypes.c_double
c_ptr = ctypes.c_void_p

# name -> argtypes; every function returns int (0 = ok) unless listed in _RESTYPES
SIGNATURES = {
    "sphb200_create": [ctypes.POINTER(c_ptr), c_int],
    "sphb200_destroy": [c_ptr],
    "sphb200_last_error": [],
    "sphb200_version": [],
    "sphb200_set_constants": [c_ptr, c_dbl, c_dbl, c_dbl, c_dbl],
    "sphb200_set_option": [c_ptr, ctypes.c_char_p, c_dbl],
    "sphb200_get_counter": [c_ptr, ctypes.c_char_p, ctypes.POINTER(c_i64), c_ptr],
    "sphb200_fp64_peak": [c_ptr, ctypes.POINTER(c_dbl), c_ptr],
    "sphb200_neighbors": [c_ptr, c_ptr, c_i64, c_int, c_dbl, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr],
    "sphb200_hydro_update": [c_ptr, c_i64, c_int] + [c_ptr] * 5 + [c_int] + [c_ptr] * 5 + [c_ptr] * 7 + [c_ptr],
    "sphb200_crossing_time": [c_ptr, c_i64, c_int, c_ptr, c_ptr, c_ptr, c_ptr, c_dbl, ctypes.POINTER(c_dbl), c_ptr],
    "sphb200_loop_density": [c_ptr, c_int, c_i64, c_int] + [c_ptr] * 6 + [c_dbl, c_ptr, c_ptr],
    "sphb200_del_pressure": [c_ptr, c_i64, c_int] + [c_ptr] * 6 + [c_dbl, c_ptr, c_ptr],
    "sphb200_artificial_viscosity": [c_ptr, c_i64, c_int] + [c_ptr] * 10 + [c_dbl, c_ptr, c_ptr, c_ptr],
    "sphb200_gravity": [c_ptr, c_i64, c_ptr, c_ptr, c_dbl, c_dbl, c_ptr, c_ptr],
    "sphb200_crossing_time_noalloc": [c_int, c_i64, c_int, c_ptr, c_ptr, c_ptr, c_ptr, c_dbl, c_ptr, c_ptr, c_ptr],
    "sphb200_session_set_gravity": [c_ptr, c_dbl, c_dbl],
    "sphb200_net_impulse": [c_ptr, c_i64, c_int] + [c_ptr] * 7 + [c_int, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr],
    "sphb200_session_create": [c_ptr, c_i64, c_int, c_dbl],
    "sphb200_session_destroy": [c_ptr],
    "sphb200_session_resize": [c_ptr, c_i64, c_i64, c_ptr],
    "sphb200_session_append_rows": [c_ptr, c_ptr, c_i64, c_int, c_ptr],
    "sphb200_session_download_rows_tagged": [c_ptr, c_ptr, c_i64, c_int, c_ptr, c_ptr],
    "sphb200_session_upload_rows_at": [c_ptr, c_ptr, c_i64, c_i64, c_int, c_ptr],
    "sphb200_session_cover_check": [c_ptr, c_dbl, c_dbl, c_ptr],
    "sphb200_session_buffer": [c_ptr, ctypes.c_char_p, c_ptr, c_ptr],
    "sphb200_session_set_drag": [c_ptr, c_ptr, c_int, c_ptr, c_ptr, c_ptr],
    "sphb200_session_upload": [c_ptr, ctypes.c_char_p, c_ptr, c_i64, c_ptr],
    "sphb200_session_download": [c_ptr, ctypes.c_char_p, c_ptr, c_i64, c_ptr],
    "sphb200_session_gather": [c_ptr, ctypes.c_char_p, c_ptr, c_i64, c_ptr, c_ptr],
    "sphb200_session_scatter": [c_ptr, ctypes.c_char_p, c_ptr, c_i64, c_ptr, c_ptr],
    "sphb200_session_download_rows": [c_ptr, c_ptr, c_i64, c_int, c_ptr],
    "sphb200_session_upload_rows": [c_ptr, c_ptr, c_i64, c_int, c_ptr],
    "sphb200_session_startup": [c_ptr, c_ptr],
    "sphb200_session_step": [c_ptr, c_ptr],
    "sphb200_session_enable_gid": [c_ptr],
    "sphb200_session_migrate_pack": [c_ptr, c_int, c_ptr, c_int, c_int, c_ptr, c_i64, c_int, c_ptr, c_ptr],
    "sphb200_session_set_count": [c_ptr, c_i64, c_i64],
    "sphb200_session_receive_rows": [c_ptr, c_ptr, c_i64, c_int, c_int, c_ptr],
    "sphb200_session_halo_pack": [c_ptr, c_int, c_ptr, c_int, c_int, c_dbl, c_ptr, c_ptr, c_i64, c_int, c_ptr, c_ptr],
    "sphb200_session_perturb": [c_ptr, c_dbl, ctypes.c_uint64, c_ptr],
    "sphb200_session_scalar": [c_ptr, ctypes.c_char_p, ctypes.POINTER(c_dbl), c_ptr],
    "sphb200_session_set_scalar": [c_ptr, ctypes.c_char_p, c_dbl],
    "sphb200_session_phase": [c_ptr, ctypes.c_char_p, c_ptr],
}
_RESTYPES = {"sphb200_last_error": ctypes.c_char_p}

_lib = None


class Sphb200Error(RuntimeError):
    pass


def load():
    """Load libsphb200.so; raises (never falls back) when it is missing."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.isfile(LIB_PATH):
        raise Sphb200Error("libsphb200.so is not built (%s); run `python sph-code_b200/build.py` — "
                           "there is no CPU fallback" % LIB_PATH)
    lib = ctypes.CDLL(LIB_PATH)
    for name, argtypes in SIGNATURES.items():
        fn = getattr(lib, name)          # AttributeError if the symbol is not exported
        fn.argtypes = argtypes
        fn.restype = _RESTYPES.get(name, c_int)
    _lib = lib
    return lib


def check(rc):
    if rc != 0:
        msg = load().sphb200_last_error()
        raise Sphb200Error(msg.decode("utf-8", "replace") if msg else "libsphb200 call failed")


class Context(object):
    """One libsphb200 context on one GPU."""

    def __init__(self, device=0):
        from . import constants as C
        lib = load()
        self._lib = lib
        self.device = int(device)
        h = c_ptr()
        check(lib.sphb200_create(ctypes.byref(h), self.device))
        self.handle = h
        check(lib.sphb200_set_constants(h, C.k, C.amu, C.m_h, C.m_0))

    def set_option(self, name, value):
        check(self._lib.sphb200_set_option(self.handle, name.encode(), float(value)))

    def counter(self, name, stream=None):
        v = c_i64(0)
        check(self._lib.sphb200_get_counter(self.handle, name.encode(), ctypes.byref(v), stream))
        return int(v.value)

    def close(self):
        if getattr(self, "handle", None):
            self._lib.sphb200_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


_contexts = {}


def context(device=None):
    """Process-wide context for `device` (default: torch's current CUDA device)."""
    import torch
    if not torch.cuda.is_available():
        raise Sphb200Error("no CUDA device visible: sphb200 runs on a B200 only (no CPU fallback)")
    if device is None:
        device = torch.cuda.current_device()
    device = int(device)
    if device not in _contexts:
        _contexts[device] = Context(device)
    return _contexts[device]
