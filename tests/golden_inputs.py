"""Rebuild the inputs of a golden trace (SE frames) from its recipe.

The golden fixtures under tests/golden/ were recorded from the unmodified reference by
oracle/refharness/gen_golden.py.  Storing every SE frame (25 x 135 float64 per step) would
make the fixtures hundreds of MB, so a fixture stores the *recipe* (seed, channel kind) plus a
per-frame checksum, and this module replays the reference's RNG consumption order
(`channels/mimic_quadriga.py:37-56`, `traffics/mult_slice.py:24-32`,
`comm_env.py:449-473`) with NumPy's own PCG64 generator.  The checksum is verified, so a NumPy
whose streams differ is detected rather than silently accepted.  Traffic frames and actions are
stored in full (they are small).
"""
from __future__ import annotations

import os

import numpy as np

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
U, S, R = 25, 5, 135

# traffic lambda per slice type id (associations/mult_slice.py:58-347)
_LAMBDA = [5, 10, 30, 15, 10, 100, 20, 100, 50, 30]


def load_scenarios():
    f = np.load(os.path.join(GOLDEN, "scenarios_seed10.npz"))
    return f["slice_type"], f["ue_slice"]


class FrameSource:
    """Yields SE frames in the order the reference env draws them."""

    def __init__(self, seed: int, channel: str, slice_type, ue_slice):
        self.rng = np.random.Generator(np.random.PCG64(np.random.SeedSequence(int(seed))))
        self.channel = channel
        self.slice_type, self.ue_slice = slice_type, ue_slice
        self.mu = None

    def _se(self):
        if self.channel == "fixed":
            return 2.0 * np.ones((U, R))
        se = np.empty((U, R))
        for u in range(U):
            se[u] = np.abs(self.rng.normal(self.mu[u], 1.5, size=(R,)))
        return se

    def _traffic(self, scen):
        out = np.zeros(U)
        for s in range(S):
            t = self.slice_type[scen, s]
            if t >= 0:
                idx = np.flatnonzero(self.ue_slice[scen] == s)
                out[idx] = self.rng.poisson(_LAMBDA[t], len(idx)) * 1e6
        return out

    def reset_frame(self, scen):
        if self.channel == "mimic":  # a fresh channel object per reset -> means are redrawn
            self.mu = np.abs(self.rng.normal(10, 7.5, size=(U,)))
        se = self._se()
        self._traffic(scen)  # drawn and discarded (comm_env.py:468-473)
        return se

    def step_frame(self, scen):
        se = self._se()
        return se, self._traffic(scen)


def bits_checksum(frames):
    """Order-independent exact checksum: sum of the IEEE bit patterns mod 2^64, per frame."""
    a = np.ascontiguousarray(frames, np.float64).view(np.uint64)
    return a.reshape(a.shape[0], -1).sum(axis=1, dtype=np.uint64)


def scen_of_episode(meta, episode):
    return episode % 200 if meta["assoc"] == "mult_slice" else episode // 100


def load_trace(name):
    f = np.load(os.path.join(GOLDEN, f"trace_{name}.npz"), allow_pickle=False)
    d = {k: f[k] for k in f.files}
    meta = {k[5:]: (v.item() if v.shape == () else v) for k, v in d.items() if k.startswith("meta_")}
    for k in ("agent", "channel", "assoc"):
        meta[k] = str(meta[k])
    return meta, d


def regenerate(meta, d):
    """Returns (reset_se [n_resets,U,R], step_se [n,U,R], step_traffic [n,U]) with the
    checksums verified against the fixture."""
    st, us = load_scenarios()
    src = FrameSource(meta["seed"], meta["channel"], st, us)
    n = int(meta["n_steps"])
    reset_se, step_se, step_tr = [], np.empty((n, U, R)), np.empty((n, U))
    ri = 0
    reset_se.append(src.reset_frame(scen_of_episode(meta, int(d["reset_episode"][0]))))
    for k in range(n):
        scen = scen_of_episode(meta, int(d["episode"][k]))
        step_se[k], step_tr[k] = src.step_frame(scen)
        if d["done"][k]:
            ri += 1
            reset_se.append(src.reset_frame(scen_of_episode(meta, int(d["reset_episode"][ri]))))
    reset_se = np.stack(reset_se)
    if not np.array_equal(bits_checksum(step_se), d["se_checksum"]):
        raise AssertionError("regenerated SE frames do not match the fixture checksum (NumPy RNG stream differs)")
    if not np.array_equal(bits_checksum(reset_se), d["reset_se_checksum"]):
        raise AssertionError("regenerated reset SE frames do not match the fixture checksum")
    if not np.array_equal(step_tr, d["traffic"]):
        raise AssertionError("regenerated traffic frames do not match the fixture")
    return reset_se, step_se, step_tr
