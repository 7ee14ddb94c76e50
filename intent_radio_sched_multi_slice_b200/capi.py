"""ctypes binding of include/irs_b200.h (the C-ABI shared library libirs_b200.so).

There is no CPU fallback: if the library is missing, `load()` raises.  PyTorch is used by the
callers only to own device memory and streams; nothing here takes torch types.
"""
from __future__ import annotations

import ctypes as C
import os

from . import build as _build

MAX_UES, MAX_SLICES, NUM_RBS, WINDOW, OBS0, OBSI, NUM_METRICS = 25, 5, 135, 10, 50, 19, 16

IRS_OK, IRS_ERR_ARG, IRS_ERR_UNBOUND, IRS_ERR_CUDA, IRS_ERR_ENV = 0, -1, -2, -3, -4
IRS_NOALLOC_MAX_ENVS = 1048576  # irs_check_errors_noalloc: size of the library's static scratch (include/irs_b200.h)
ERR_INTRA_ACTION, ERR_EPISODE_RANGE, ERR_RB_SUM, ERR_TRACE_EPISODE, ERR_RB_SHARED = 1, 2, 4, 8, 16

METRIC_NAMES = [
    "reward_player_0", "reward_player_1", "reward_player_2", "reward_player_3", "reward_player_4",
    "reward_player_5", "sent_pkts", "dropped_pkts", "incoming_pkts", "throughput_mbit", "violations",
    "violations_priority", "intent_distance", "intent_distance_priority", "env_steps", "episodes",
]


class Config(C.Structure):
    _fields_ = [(n, C.c_int32) for n in (
        "num_envs", "depth_cap", "channel_mode", "traffic_mode", "sort_slices", "fixed_intra", "policy",
        "assoc_mode", "scen_mod", "scen_div", "n_scen", "max_steps", "max_episodes", "initial_episode",
        "auto_reset", "env_id_base", "surface", "random_episodes")] + [("seed", C.c_uint64), ("bandwidth", C.c_double)]


_LAYOUT_ARRAYS = [
    "buf", "ue_i", "ue_d", "ring", "rb", "step", "episode", "scen", "win_len",
    "win_head", "err", "visit", "ep_pos", "ep_gen", "ep_base", "wsum", "sorted", "acc", "bank", "cdf", "in_a0", "in_intra", "out_obs0", "out_obsi",
    "out_mask0", "out_maski", "out_reward", "out_done", "metrics_total", "scratch", "out32",
]
UI_FIELDS = ["tot", "sum_sent", "sum_drop", "last_sent", "last_drop", "last_inc", "occ1z", "path"]
UD_FIELDS = ["occ", "lat", "sebar", "thr"]


class Layout(C.Structure):
    _fields_ = [("total_bytes", C.c_size_t), ("row_stride", C.c_int32), ("reserved", C.c_int32)] + \
               [(n, C.c_size_t) for n in _LAYOUT_ARRAYS]


class Actions(C.Structure):
    _fields_ = [("a0", C.c_void_p), ("intra", C.c_void_p), ("sched_bits", C.c_void_p)]


class Inject(C.Structure):
    _fields_ = [("se", C.c_void_p), ("traffic", C.c_void_p)]


RAW_FIELDS = ["pkt_incoming", "pkt_throughputs", "sent", "dropped", "occ", "lat", "rb_start", "rb_count",
              "sorted_slices", "a0_used", "intra_used", "se_used", "traffic_used"]


class RawOutputs(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in RAW_FIELDS]


SURFACE_FIELDS = ["obs0", "obsi", "mask0", "maski", "reward", "done"]
SURFACE_F32_FIELDS = ["obs0_f32", "obsi_f32", "reward_f32"]  # optional binary32 copies (converted on the device)


class HostSurface(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in SURFACE_FIELDS + SURFACE_F32_FIELDS]


# every symbol include/irs_b200.h declares: (restype, argtypes)
_P = C.c_void_p
SYMBOLS = {
    "irs_version": (C.c_int, []),
    "irs_create": (C.c_int, [C.POINTER(Config), C.c_int32, C.POINTER(_P)]),
    "irs_destroy": (C.c_int, [_P]),
    "irs_state_layout": (C.c_int, [_P, C.POINTER(Layout)]),
    "irs_bind": (C.c_int, [_P, _P, C.c_size_t, _P]),
    "irs_load_scenarios": (C.c_int, [_P, _P, C.c_int32, _P, C.c_int32, _P]),
    "irs_reset": (C.c_int, [_P, _P, C.c_int32, _P, C.POINTER(Inject), C.POINTER(RawOutputs), _P]),
    "irs_step": (C.c_int, [_P, C.POINTER(Actions), C.POINTER(Inject), C.POINTER(RawOutputs), _P]),
    "irs_step_profiled": (C.c_int, [_P, C.POINTER(Actions), C.POINTER(Inject), C.POINTER(RawOutputs), _P,
                                    C.POINTER(C.c_float)]),
    "irs_set_seed": (C.c_int, [_P, C.c_uint64]),
    "irs_set_se_trace": (C.c_int, [_P, _P, C.c_int32, C.c_int32, C.c_int32, C.c_int32]),
    "irs_power_to_se": (C.c_int, [_P, _P, C.c_int64, C.c_int32, C.c_double, C.c_double, _P]),
    "irs_run": (C.c_int, [_P, C.c_int32, _P]),
    "irs_step_host": (C.c_int, [_P, C.POINTER(Actions), C.POINTER(HostSurface), _P]),
    "irs_reset_host": (C.c_int, [_P, C.c_int32, C.POINTER(HostSurface), _P]),
    "irs_episode_metrics": (C.c_int, [_P, _P, _P]),
    "irs_check_errors": (C.c_int, [_P, C.POINTER(C.c_int32), C.POINTER(C.c_int32), _P]),
    "irs_set_episode_order": (C.c_int, [_P, _P, C.c_int32, _P]),
    "irs_read_buffers": (C.c_int, [_P, _P, _P]),
    "irs_check_errors_noalloc": (C.c_int, [_P, C.c_int32, _P, _P, C.c_char_p, C.c_size_t]),
    "irs_last_error": (C.c_char_p, [_P]),
}

_lib = None


def lib_path() -> str:
    return _build.LIB


def load(build_if_missing: bool = True):
    """dlopen libirs_b200.so and bind every exported symbol; raises if it cannot."""
    global _lib
    if _lib is not None:
        return _lib
    path = lib_path()
    if not os.path.exists(path) or (build_if_missing and not os.environ.get("IRS_LIB_PATH") and _build.stale()):
        if not build_if_missing:
            raise RuntimeError(f"{path} is missing: run `python -m intent_radio_sched_multi_slice_b200.build`")
        try:
            _build.build()
        except Exception as exc:  # no nvcc on the box: use the prebuilt library if there is one
            if not os.path.exists(path):
                raise RuntimeError(f"cannot build {path}: {exc}") from exc
    lib = C.CDLL(path)
    for name, (res, args) in SYMBOLS.items():
        fn = getattr(lib, name)  # AttributeError if the export is missing
        fn.restype, fn.argtypes = res, args
    _lib = lib
    return lib


class IrsError(RuntimeError):
    def __init__(self, code: int, message: str):
        super().__init__(f"irs_b200 error {code}: {message}")
        self.code = code


def check(handle, rc: int):
    if rc != IRS_OK:
        msg = load().irs_last_error(handle)
        raise IrsError(rc, msg.decode() if msg else "")
