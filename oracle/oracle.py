"""ctypes wrapper around oracle/irs_oracle.c (the CPU restatement of the reference path).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and the cpu_baseline /
`--impl reference` legs of bench.py - never by the product package.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "_build", "libirs_oracle.so")
U, S, R = 25, 5, 135


def build(force: bool = False) -> str:
    src = os.path.join(HERE, "irs_oracle.c")
    if force or not os.path.exists(LIB) or os.path.getmtime(LIB) < os.path.getmtime(src):
        os.makedirs(os.path.dirname(LIB), exist_ok=True)
        subprocess.check_call(
            ["gcc", "-O2", "-fPIC", "-shared", "-ffp-contract=off", "-fopenmp", "-mfma",
             "-o", LIB, src, "-lm"]
        )
    return LIB


class Config(C.Structure):
    _fields_ = [(n, C.c_int32) for n in (
        "num_envs", "depth_cap", "channel_mode", "traffic_mode", "sort_slices", "fixed_intra", "policy",
        "assoc_mode", "scen_mod", "scen_div", "n_scen", "max_steps", "max_episodes", "initial_episode",
        "auto_reset", "env_id_base", "surface", "random_episodes")] + [("seed", C.c_uint64), ("bandwidth", C.c_double)]


_IO_FIELDS = ["pkt_incoming", "pkt_throughputs", "sent", "dropped", "occ", "lat", "rb_start", "rb_count",
              "obs0", "mask0", "obs_i", "mask_i", "sorted_slices", "reward", "done", "a0_used", "intra_used",
              "se_used", "traffic_used"]


class IO(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in _IO_FIELDS]


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(LIB)
        _lib.orc_create.restype = C.c_void_p
        _lib.orc_create.argtypes = [C.POINTER(Config), C.c_void_p, C.c_void_p]
        _lib.orc_destroy.argtypes = [C.c_void_p]
        _lib.orc_get_scenario.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
        _lib.orc_set_se_trace.argtypes = [C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_int32]
        _lib.orc_set_episode_order.argtypes = [C.c_void_p, C.c_void_p, C.c_int32]
        _lib.orc_power_to_se.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_double, C.c_double]
        _lib.orc_reset.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.POINTER(IO)]
        _lib.orc_step.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(IO)]
        _lib.orc_buffers.restype = C.POINTER(C.c_int32)
        _lib.orc_buffers.argtypes = [C.c_void_p]
        _lib.orc_get_state.argtypes = [C.c_void_p] * 4
        _lib.orc_episode_metrics.argtypes = [C.c_void_p] * 3
        _lib.orc_buffer_main.restype = C.c_double
        _lib.orc_buffer_main.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
        _lib.orc_np_sum.restype = C.c_double
        _lib.orc_np_sum.argtypes = [C.c_void_p, C.c_int]
        _lib.orc_round_int_equal_sum.argtypes = [C.c_void_p, C.c_int, C.c_int64, C.c_void_p]
        _lib.orc_normal16.restype = C.c_float
        _lib.orc_normal16.argtypes = [C.c_uint32]
        _lib.orc_feistel_perm.restype = C.c_uint32
        _lib.orc_feistel_perm.argtypes = [C.c_uint32, C.c_uint32, C.c_void_p]
        _lib.orc_philox.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32]
        assert _lib.orc_sizeof_config() == C.sizeof(Config)
        assert _lib.orc_sizeof_scenario() == 432
    return _lib


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def buffer_main(max_pkts, max_age, arrivals, send_per_step):
    arr = np.asarray(arrivals, np.int32)
    n, d = len(arr), max_age + 1
    rows = np.zeros((n, 2, d), np.int32)
    dropped = np.zeros(n, np.int64)
    sent = np.zeros(n, np.int64)
    delay = lib().orc_buffer_main(max_pkts, max_age, _ptr(arr), n, send_per_step, _ptr(rows), _ptr(dropped), _ptr(sent))
    return rows, dropped, sent, delay


def power_to_se(power, layout=0, tx_power=100.0, noise=10e-14):
    """channels/quadriga.py:56-74 over frames [..., 135, 25] (layout 0) or [..., 25, 135] (layout 1)."""
    p = np.ascontiguousarray(power, np.float64)
    out = np.empty(p.shape[:-2] + (U, R))
    lib().orc_power_to_se(_ptr(p), _ptr(out), p.size // (U * R), layout, tx_power, noise)
    return out


def set_threads(n: int) -> int:
    """OpenMP threads of orc_step/orc_reset (torchrun exports OMP_NUM_THREADS=1); returns the count in force."""
    lib().orc_set_threads(int(n))
    return int(lib().orc_max_threads())


def np_sum(a):
    a = np.ascontiguousarray(a, np.float64)
    return lib().orc_np_sum(_ptr(a), a.size)


def round_int_equal_sum(x, target):
    x = np.ascontiguousarray(x, np.float64)
    out = np.zeros(x.size, np.int64)
    lib().orc_round_int_equal_sum(_ptr(x), x.size, int(target), _ptr(out))
    return out


def normal16(h):
    """One standard normal from a 16-bit word (the native generators' inverse-CDF table transform)."""
    return float(lib().orc_normal16(int(h)))


def feistel_perm(i, n, keys):
    k = np.asarray(keys, np.uint32)
    return int(lib().orc_feistel_perm(int(i), int(n), _ptr(k)))


def philox(counter, k0, k1):
    c = np.array(counter, np.uint32)
    lib().orc_philox(_ptr(c), int(k0), int(k1))
    return c


class OracleEnv:
    """Batch of E reference-equivalent envs stepped on the CPU."""

    def __init__(self, bank_array, poisson_cdf, num_envs=1, depth_cap=None, channel_mode=0, traffic_mode=0,
                 sort_slices=1, fixed_intra=-1, policy=0, assoc_mode=0, scen_mod=200, scen_div=100,
                 max_steps=1000, max_episodes=10000, initial_episode=0, auto_reset=0, env_id_base=0, seed=10, surface=0,
                 bandwidth=1e8, random_episodes=0):
        self.bank = np.ascontiguousarray(bank_array)
        assert self.bank.dtype.itemsize == 432
        self.cdf = np.ascontiguousarray(poisson_cdf, np.float64)
        if depth_cap is None:
            depth_cap = int(self.bank["sl"]["buffer_latency"].max()) + 1
        self.E = num_envs
        self.depth_cap = depth_cap
        self.cfg = Config(num_envs, depth_cap, channel_mode, traffic_mode, sort_slices, fixed_intra, policy,
                          assoc_mode, scen_mod, scen_div, len(self.bank), max_steps, max_episodes,
                          initial_episode, auto_reset, env_id_base, surface, int(random_episodes), seed, bandwidth)
        self.h = lib().orc_create(C.byref(self.cfg), _ptr(self.bank), _ptr(self.cdf))
        E = num_envs
        self.out = {
            "pkt_incoming": np.zeros((E, U)), "pkt_throughputs": np.zeros((E, U)),
            "sent": np.zeros((E, U), np.int32), "dropped": np.zeros((E, U), np.int32),
            "occ": np.zeros((E, U)), "lat": np.zeros((E, U)),
            "rb_start": np.zeros((E, U), np.int16), "rb_count": np.zeros((E, U), np.int16),
            "obs0": np.zeros((E, 50)), "mask0": np.zeros((E, S), np.int8),
            "obs_i": np.zeros((E, S, 19)), "mask_i": np.zeros((E, S, 5), np.int8),
            "sorted_slices": np.zeros((E, S), np.int8), "reward": np.zeros((E, 6)),
            "done": np.zeros(E, np.uint8), "a0_used": np.zeros((E, S)), "intra_used": np.zeros((E, S), np.int32),
            "se_used": None, "traffic_used": np.zeros((E, U)),
        }
        self._io = None

    def set_se_trace(self, se_bank, first_episode=0, wrap=False):
        """channel_mode 3: float64 [n_episodes, n_steps, 25, 135] (kept alive by this object); slot i = env episode
        first_episode + i (modulo n_episodes with `wrap`)."""
        self._se_bank = np.ascontiguousarray(se_bank, np.float64)
        assert self._se_bank.ndim == 4 and self._se_bank.shape[2:] == (U, R)
        lib().orc_set_se_trace(self.h, _ptr(self._se_bank), self._se_bank.shape[0], self._se_bank.shape[1],
                               int(first_episode), int(bool(wrap)))

    def set_episode_order(self, order):
        """random_episodes with a given order: int32 [E, n] (kept alive by this object); None = built-in permutation."""
        if order is None:
            self._order = None
            lib().orc_set_episode_order(self.h, None, 0)
            return
        self._order = np.ascontiguousarray(order, np.int32)
        assert self._order.ndim == 2 and self._order.shape[0] == self.E
        lib().orc_set_episode_order(self.h, _ptr(self._order), self._order.shape[1])

    def scenario_rows(self):
        """The scenario row every env is playing (432-byte records, dtype of the bank)."""
        out = np.zeros(self.E, self.bank.dtype)
        for e in range(self.E):
            lib().orc_get_scenario(self.h, e, out[e:e + 1].ctypes.data)
        return out

    def want_se(self):
        self.out["se_used"] = np.zeros((self.E, U, R))
        self._io = None

    def _get_io(self):
        if self._io is None:
            self._io = IO(*[_ptr(self.out[n]) for n in _IO_FIELDS])
        return self._io

    def reset(self, mask=None, initial_episode=-1, se=None):
        m = None if mask is None else np.ascontiguousarray(mask, np.uint8)
        se = None if se is None else np.ascontiguousarray(se, np.float64)
        rc = lib().orc_reset(self.h, _ptr(m), int(initial_episode), _ptr(se), C.byref(self._get_io()))
        if rc:
            raise RuntimeError(f"oracle reset failed rc={rc}")
        return self.out

    def step(self, a0=None, intra=None, se=None, traffic=None, sched_bits=None):
        """sched_bits: uint32 [E, 25, 5] - a scheduling matrix given directly (bit r of the 135-bit row = RB r)."""
        sched_bits = None if sched_bits is None else np.ascontiguousarray(sched_bits, np.uint32)
        a0 = None if a0 is None else np.ascontiguousarray(a0, np.float64)
        intra = None if intra is None else np.ascontiguousarray(intra, np.int32)
        se = None if se is None else np.ascontiguousarray(se, np.float64)
        traffic = None if traffic is None else np.ascontiguousarray(traffic, np.float64)
        rc = lib().orc_step(self.h, _ptr(a0), _ptr(intra), _ptr(se), _ptr(traffic), _ptr(sched_bits), C.byref(self._get_io()))
        if rc:
            raise RuntimeError(f"oracle step failed rc={rc}")
        return self.out

    def buffers(self):
        p = lib().orc_buffers(self.h)
        return np.ctypeslib.as_array(p, shape=(self.E, U, self.depth_cap))

    def state(self):
        st = [np.zeros(self.E, np.int32) for _ in range(3)]
        lib().orc_get_state(self.h, *[_ptr(a) for a in st])
        return dict(step=st[0], episode=st[1], scen=st[2])

    def episode_metrics(self):
        per = np.zeros((self.E, 16))
        tot = np.zeros(16)
        lib().orc_episode_metrics(self.h, _ptr(per), _ptr(tot))
        return per, tot

    def close(self):
        if self.h:
            lib().orc_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
