"""ctypes loader for libstm_b200.so (the C ABI declared in include/stm_b200.h).

There is no CPU fallback: if the library is missing or no CUDA device is present, the product path
raises.  Nothing under oracle/ is ever imported from here.
"""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("STM_B200_LIB", os.path.join(HERE, "libstm_b200.so"))   # override: A/B builds of the same ABI
# where stm_ctx_create looks for the translated step kernel of a cartridge image the library was not built with (translate.py)
os.environ.setdefault("STM_XL_PLUGIN_DIR", os.path.join(HERE, "xl_plugins"))

STATE_BYTES = 256
OBS = 84
STACK = 4
STACK_BYTES = STACK * OBS * OBS


class GameDesc(C.Structure):
    """Mirror of stm_game_desc (include/stm_state.h)."""
    _fields_ = [("rom_size", C.c_uint32), ("bank_scheme", C.c_uint8), ("controller", C.c_uint8),
                ("n_actions", C.c_uint8), ("actions", C.c_uint8 * 18), ("score_addr", C.c_uint8 * 3),
                ("score2_addr", C.c_uint8 * 3), ("lives_addr", C.c_uint8), ("lives_shift", C.c_uint8),
                ("lives_mask", C.c_uint8), ("term_addr", C.c_uint8), ("term_mask", C.c_uint8),
                ("term_value", C.c_uint8), ("pad", C.c_uint8 * 3)]


class StmError(RuntimeError):
    pass


_lib = None

_SIGS = {
    "stm_ctx_create": (C.c_int, [C.POINTER(C.c_void_p), C.POINTER(GameDesc), C.c_void_p, C.c_int]),
    "stm_ctx_destroy": (C.c_int, [C.c_void_p]),
    "stm_last_error": (C.c_char_p, [C.c_void_p]),
    "stm_state_bytes": (C.c_size_t, [C.c_void_p]),
    "stm_num_actions": (C.c_int, [C.c_void_p]),
    "stm_set_palette": (C.c_int, [C.c_void_p, C.c_void_p]),
    "stm_reset_states": (C.c_int, [C.c_void_p] * 5 + [C.c_int, C.c_void_p]),
    "stm_expand": (C.c_int, [C.c_void_p] * 9 + [C.c_int, C.c_int, C.c_void_p]),
    "stm_step": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_float,
                           C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]),
    "stm_step_frame": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_float,
                                 C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "stm_frame_obs": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]),
    "stm_max_step_envs": (C.c_int, [C.c_void_p]),
    "stm_set_scratch_envs": (C.c_int, [C.c_void_p, C.c_int]),
    "stm_frames": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p]),
    "stm_bfs": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_float, C.c_int, C.c_int,
                          C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "stm_bfs_after": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_float, C.c_int, C.c_int,
                                C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "stm_bfs_resume": (C.c_int, [C.c_void_p] * 6 + [C.c_int, C.c_int, C.c_int, C.c_float, C.c_int, C.c_int, C.c_int]
                       + [C.c_void_p] * 6),
    "stm_bfs_host": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_float, C.c_int, C.c_int,
                               C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "stm_timing_enable": (C.c_int, [C.c_void_p, C.c_int]),
    "stm_timing_read": (C.c_int, [C.c_void_p, C.POINTER(C.c_float), C.POINTER(C.c_int)]),
    "stm_launch_count": (C.c_longlong, [C.c_void_p]),
    "stm_translated_core": (C.c_int, [C.c_void_p]),
    "stm_debug_shift_probe": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p]),
    "stm_step_frames_noalloc": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_float,
                                          C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]),
    "stm_cnn_set_weights": (C.c_int, [C.c_void_p] * 13 + [C.c_int, C.c_void_p]),
    "stm_cnn_forward": (C.c_int, [C.c_void_p, C.c_void_p, C.c_longlong, C.c_void_p, C.c_void_p]),
    "stm_cnn_forward_train": (C.c_int, [C.c_void_p, C.c_void_p, C.c_longlong, C.c_void_p, C.c_void_p]),
    "stm_cnn_backward": (C.c_int, [C.c_void_p, C.c_void_p, C.c_longlong, C.c_void_p] + [C.c_void_p] * 12 + [C.c_void_p]),
    "stm_logits_fwd": (C.c_int, [C.c_void_p] * 6 + [C.c_int] * 3 + [C.c_float] * 4 + [C.c_int] + [C.c_void_p] * 4),
    "stm_logits_bwd": (C.c_int, [C.c_void_p] * 5 + [C.c_int] * 3 + [C.c_float] * 4 + [C.c_int] + [C.c_void_p] * 6),
}


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise StmError(f"{LIB_PATH} is missing: build it with `python -m softtreemax_b200.build` "
                           "(there is no CPU fallback)")
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in _SIGS.items():
            fn = getattr(L, name)
            fn.restype, fn.argtypes = res, args
        _lib = L
    return _lib


def exported_symbols():
    return sorted(_SIGS)


def check(rc, ctx=None):
    if rc != 0:
        msg = lib().stm_last_error(ctx)
        raise StmError(f"stm error {rc}: {msg.decode() if msg else '?'}")
