"""ctypes binding of libragen_b200.so (the C ABI declared in include/ragen_b200.h).

There is no CPU fallback: if the library is missing or no B200 is visible, calls raise.
"""
import ctypes
import os
import threading

from . import build as _build

_c = ctypes
_P = _c.c_void_p
_I = _c.c_int
_L = _c.c_int64
_F = _c.c_float
_D = _c.c_double

# name -> argument ctypes (return type is int unless listed in _RESTYPE)
SIGNATURES = {
    "ragen_create": [_I, _c.POINTER(_P)],
    "ragen_destroy": [_P],
    "ragen_last_error": [],
    "ragen_version": [],
    "ragen_launch_count": [_P],
    "ragen_sm_count": [_P],
    "ragen_sokoban_step": [_P, _L, _I, _I, _I, _I, _P, _I, _P, _P, _P, _P, _P, _P, _P, _P, _P],
    "ragen_sokoban_step_packed": [_P, _L, _I, _I, _I, _I, _P, _I, _P, _P, _P, _P, _P, _P, _P, _P, _P],
    "ragen_sokoban_render": [_P, _L, _I, _I, _P, _P, _P, _P],
    "ragen_sokoban_render_packed": [_P, _L, _I, _I, _P, _P, _P, _P],
    "ragen_sokoban_step_packed3": [_P, _L, _I, _I, _I, _I, _P, _I, _P, _P, _P, _P, _P, _P, _P, _P, _P],
    "ragen_sokoban_step_compact": [_P, _L, _I, _I, _I, _I, _P, _I, _P, _P, _P, _P, _P, _P, _P, _P, _P],
    "ragen_sokoban_step_delta": [_P, _L, _I, _I, _I, _I, _P, _I, _P, _P, _P, _P, _P],
    "ragen_sokoban_render_packed3": [_P, _L, _I, _I, _P, _P, _P, _P],
    "ragen_sokoban_render_rgb": [_P, _L, _I, _I, _I, _P, _P, _P, _P, _L, _P, _P],
    "ragen_sokoban_adopt": [_P, _L, _I, _I, _I, _P, _P, _P, _P, _P],
    "ragen_sokoban_adopt_slots": [_P, _L, _I, _I, _I, _P, _P, _P, _P, _P, _P, _P, _P],
    "ragen_sokoban_step_glyph3": [_P, _L, _I, _I, _I, _I, _P, _I, _P, _P, _P, _P, _P, _P, _P, _P, _P],
    "ragen_sokoban_step_delta1": [_P, _L, _I, _I, _I, _I, _P, _I, _P, _I, _P, _P, _P, _P],
    "ragen_sokoban_generate": [_P, _L, _I, _I, _I, _I, _I, _P, _P, _P, _P, _P, _P],
    "ragen_sokoban_solve": [_P, _L, _I, _I, _P, _P, _I, _L, _P, _P, _P, _P, _P],
    "ragen_frozenlake_reset": [_P, _L, _I, _D, _P, _P, _P, _P, _P, _P, _P],
    "ragen_frozenlake_step": [_P, _L, _I, _I, _P, _I, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P],
    "ragen_frozenlake_render": [_P, _L, _I, _P, _P, _P, _P],
    "ragen_bandit_reset": [_P, _L, _P, _P, _P, _P],
    "ragen_bandit_step": [_P, _L, _I, _D, _D, _D, _D, _P, _P, _P, _P, _P, _P, _P],
    "ragen_parse_actions": [_P, _L, _P, _P, _I, _c.c_char_p, _I, _P, _P, _P, _P, _I, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P],
    "ragen_countdown_score": [_P, _L, _P, _P, _P, _P, _P, _P, _P, _P, _P],
    "ragen_masks_scores": [_P, _I, _I, _P, _L, _L, _I, _I, _I, _P, _P, _I, _P, _P, _P, _P],
    "ragen_traj_assemble": [_P, _I, _I, _P, _L, _P, _P, _P, _P, _P, _L, _P, _P, _P, _P],
    "ragen_glyph_tokens": [_P, _L, _I, _I, _P, _L, _P, _P, _I, _P, _P, _P],
    "ragen_glyph_utf8": [_P, _L, _I, _I, _I, _P, _L, _P, _P, _I, _P, _P, _P],
    "ragen_group_norm": [_P, _I, _P, _P, _P, _I, _I, _P, _P],
    "ragen_group_norm_stats": [_P, _I, _P, _P, _P, _I, _P, _P, _P],
    "ragen_group_norm_apply": [_P, _I, _P, _P, _I, _P, _I, _P],
    "ragen_group_norm_noalloc": [_P, _I, _P, _P, _P, _I, _I, _P, _P],
    "ragen_gae": [_P, _I, _I, _P, _P, _P, _D, _D, _P, _P, _P, _P],
    "ragen_gae_scan": [_P, _I, _I, _P, _P, _P, _D, _D, _P, _P, _P, _P],
    "ragen_bilevel_gae": [_P, _I, _I, _P, _P, _P, _D, _D, _D, _P, _P, _P, _P],
    "ragen_bilevel_gae_scan": [_P, _I, _I, _P, _P, _P, _D, _D, _D, _P, _P, _P, _P],
    "ragen_masked_moments": [_P, _L, _P, _P, _P, _P],
    "ragen_whiten": [_P, _L, _P, _P, _P],
    "ragen_grpo_outcome": [_P, _I, _I, _P, _P, _P, _I, _D, _I, _P, _P, _P],
    "ragen_outcome_advantage": [_P, _I, _I, _P, _P, _P, _I, _I, _D, _P, _P, _P, _P],
    "ragen_discounted_returns": [_P, _I, _I, _P, _P, _D, _P, _P, _P],
    "ragen_remax": [_P, _I, _I, _P, _P, _P, _P, _P, _P],
    "ragen_remax_scan": [_P, _I, _I, _P, _P, _P, _P, _P, _P],
    "ragen_discounted_returns_scan": [_P, _I, _I, _P, _P, _D, _P, _P, _P],
    "ragen_whiten_masked": [_P, _L, _P, _P, _P, _P, _P],
    "ragen_kl_rewards": [_P, _I, _I, _P, _P, _P, _P, _I, _D, _P, _P, _P],
    "ragen_text_score": [_P, _L, _I, _P, _P, _P, _P, _P, _P],
    "ragen_group_filter_stats": [_P, _I, _I, _I, _P, _P, _P, _P, _P, _P],
}
_RESTYPE = {"ragen_last_error": _c.c_char_p, "ragen_launch_count": _c.c_uint64}

_lock = threading.Lock()
_lib = None


class RagenError(RuntimeError):
    pass


def lib_path():
    return _build.LIB_PATH


def load():
    """Load (building first if the sources changed and nvcc is present) and type the library."""
    global _lib
    with _lock:
        if _lib is None:
            path = _build.build()
            if not os.path.exists(path):
                raise RagenError(f"{path} is missing: run `python -m ragen_b200.build` (no CPU fallback exists)")
            lib = _c.CDLL(path)
            for name, argtypes in SIGNATURES.items():
                fn = getattr(lib, name)  # AttributeError if the symbol is not exported
                fn.argtypes = argtypes
                fn.restype = _RESTYPE.get(name, _I)
            _lib = lib
    return _lib


def check(rc):
    if rc != 0:
        raise RagenError(f"libragen_b200 error {rc}: {load().ragen_last_error().decode()}")


def ptr(t):
    """Device (or host) pointer of a torch tensor / None."""
    if t is None:
        return None
    return _P(t.data_ptr())


class Context:
    """Owns a ragen_ctx for one CUDA device."""

    _per_device = {}

    def __init__(self, device=0):
        import torch

        if not torch.cuda.is_available():
            raise RagenError("no CUDA device visible: ragen_b200 has no CPU fallback")
        self.lib = load()
        self.device = int(device)
        h = _P()
        check(self.lib.ragen_create(self.device, _c.byref(h)))
        self.handle = h

    @classmethod
    def get(cls, device=0):
        device = int(device)
        if device not in cls._per_device:
            cls._per_device[device] = cls(device)
        return cls._per_device[device]

    @property
    def launches(self):
        return int(self.lib.ragen_launch_count(self.handle))

    @property
    def sm_count(self):
        return int(self.lib.ragen_sm_count(self.handle))

    def stream(self):
        import torch

        return _P(torch.cuda.current_stream(self.device).cuda_stream)

    def close(self):
        if self.handle:
            self.lib.ragen_destroy(self.handle)
            self.handle = None
