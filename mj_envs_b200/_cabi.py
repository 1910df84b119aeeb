"""ctypes declarations for the C ABI of include/adroit_b200.h.

The CUDA library exports `adroit_*`; the test-only CPU oracle exports the same signatures as `oracle_*`.
Pointers are passed as integers (device pointers for the CUDA library, host pointers for the oracle).
"""
import ctypes as C

_VP = C.c_void_p
SYMBOLS = {
    "create": (C.c_int, [_VP, C.c_size_t, C.c_int, C.c_int, C.c_uint64, C.c_int64, C.c_uint32, C.POINTER(_VP)]),
    "destroy": (None, [_VP]),
    "last_error": (C.c_char_p, []),
    "sizes": (C.c_int, [_VP, _VP]),
    "reset": (C.c_int, [_VP, _VP, _VP, _VP]),
    "step": (C.c_int, [_VP, _VP, _VP, _VP, _VP, _VP, _VP]),
    "step_host": (C.c_int, [_VP, _VP, _VP, _VP, _VP, _VP, _VP]),
    "get_obs": (C.c_int, [_VP, _VP, _VP]),
    "forward": (C.c_int, [_VP, _VP]),
    "get_state": (C.c_int, [_VP, _VP, _VP, _VP, _VP]),
    "set_state": (C.c_int, [_VP, _VP, _VP, _VP, _VP]),
    "set_elapsed_steps": (C.c_int, [_VP, _VP, _VP]),
    "get_warmstart": (C.c_int, [_VP, _VP, _VP]),
    "set_warmstart": (C.c_int, [_VP, _VP, _VP]),
    "get_frames": (C.c_int, [_VP, _VP, _VP, _VP, _VP]),
    "get_solver": (C.c_int, [_VP, _VP, _VP, _VP, _VP, _VP]),
    "reduce_episode": (C.c_int, [_VP, C.c_int, _VP, _VP]),
    "reduce_episode_noalloc": (C.c_int, [_VP, C.c_int, _VP, _VP]),
    "launch_count": (C.c_int64, [_VP]),
    "overflow": (C.c_int, [_VP, _VP, _VP]),
    "render": (C.c_int, [_VP, _VP, _VP]),
    "profile": (C.c_int, [_VP, C.c_int]),
    "phase_times": (C.c_int, [_VP, _VP, _VP, _VP]),
}

F64, AUTO_RESET, RESET_ON_DONE = 1, 2, 4


class CABI:
    def __init__(self, path: str, prefix: str):
        self.lib = C.CDLL(path)
        self.prefix = prefix
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(self.lib, f"{prefix}_{name}")
            fn.restype = res
            fn.argtypes = args
            setattr(self, name, fn)

    def check(self, rc: int):
        if rc != 0:
            msg = self.last_error()
            raise RuntimeError(f"{self.prefix}: {msg.decode() if msg else 'error'} (rc={rc})")
