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


QUADRIGA_TX_POWER, QUADRIGA_NOISE = 100.0, 10e-14  # channels/quadriga.py:33-34


def quadriga_power(assoc: int, ep: int, n_steps: int):
    """Synthetic stand-in for mult_slice_channel_generation/.../assoc_A/ep_E/target_cell_power.mat
    (the real files are an external QuaDRiGa/MATLAB product): a fixed recipe, in the order h5py
    shows the MATLAB array of simu.m:103 - [step][rb][sector=1][1][ue]."""
    rng = np.random.default_rng(7_000_000 + 1000 * int(assoc) + int(ep))
    return 10.0 ** rng.uniform(-13.5, -9.5, size=(n_steps, R, 1, 1, U))


def quadriga_file_of_episode(channel: str, episode: int):
    """choose_episode of channels/quadriga.py:79-87 / channels/quadriga_seq.py:28-39."""
    return (episode, 0) if channel == "quadriga" else (episode // 100, episode % 100)


def quadriga_se_numpy(power_frame):
    """channels/quadriga.py:56-77 on one frame [R,1,1,U] -> [U,R], NumPy restatement."""
    se = np.log2(1 + np.divide((QUADRIGA_TX_POWER / np.array([R])[0]) * power_frame,
                               (np.zeros_like(power_frame) + QUADRIGA_NOISE)))
    return np.squeeze(se.transpose())


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
        self.n_steps = None  # quadriga channels: frames per file
        self._files = {}

    def _quadriga(self, episode, k):
        key = quadriga_file_of_episode(self.channel, episode)
        if key not in self._files:
            self._files = {key: quadriga_power(key[0], key[1], self.n_steps)}
        return quadriga_se_numpy(self._files[key][k])

    def _se(self, episode=None, k=None):
        if self.channel.startswith("quadriga"):
            return self._quadriga(episode, k)
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

    def reset_frame(self, scen, episode=None):
        if self.channel == "mimic":  # a fresh channel object per reset -> means are redrawn
            self.mu = np.abs(self.rng.normal(10, 7.5, size=(U,)))
        se = self._se(episode, 0)
        self._traffic(scen)  # drawn and discarded (comm_env.py:468-473)
        return se

    def step_frame(self, scen, episode=None, k=None):
        se = self._se(episode, k)
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
    src.n_steps = int(meta["max_steps"])
    reset_se, step_se, step_tr = [], np.empty((n, U, R)), np.empty((n, U))
    ri = 0
    ep0 = int(d["reset_episode"][0])
    n_perm = [0]

    def permutation_draw(ri):
        """enable_random_episodes: a reset with options["initial_episode"] != -1 draws a new permutation from the
        env's generator BEFORE the scenario is created (comm_env.py:435-442) - the recorded order must be what the
        stream gives at that point."""
        if not int(meta.get("random_episodes", 0)) or int(d["reset_initial"][ri]) == -1:
            return
        want = d[f"reset_order_{n_perm[0]}"]
        got = src.rng.permutation(np.arange(int(d["reset_initial"][ri]), int(meta["max_episodes"])))
        if not np.array_equal(got, want):
            raise AssertionError("regenerated episode permutation does not match the fixture")
        n_perm[0] += 1

    permutation_draw(0)
    reset_se.append(src.reset_frame(scen_of_episode(meta, ep0), ep0))
    k_in_ep = 0  # the env's step counter when channel.step is called (comm_env.py:303-305)
    for k in range(n):
        ep = int(d["episode"][k])
        step_se[k], step_tr[k] = src.step_frame(scen_of_episode(meta, ep), ep, k_in_ep)
        k_in_ep += 1
        if d["done"][k]:
            ri += 1
            epr = int(d["reset_episode"][ri])
            permutation_draw(ri)
            reset_se.append(src.reset_frame(scen_of_episode(meta, epr), epr))
            k_in_ep = 0
    reset_se = np.stack(reset_se)
    if meta["channel"].startswith("quadriga"):
        # np.log2 may round differently on another CPU (SIMD dispatch): frames are pinned by their
        # sums to 1e-13 instead of by bit patterns
        np.testing.assert_allclose(step_se.reshape(n, -1).sum(1), d["se_sum"], rtol=1e-13)
        np.testing.assert_allclose(reset_se.reshape(len(reset_se), -1).sum(1), d["reset_se_sum"], rtol=1e-13)
        if not np.array_equal(step_tr, d["traffic"]):
            raise AssertionError("regenerated traffic frames do not match the fixture")
        return reset_se, step_se, step_tr
    if not np.array_equal(bits_checksum(step_se), d["se_checksum"]):
        raise AssertionError("regenerated SE frames do not match the fixture checksum (NumPy RNG stream differs)")
    if not np.array_equal(bits_checksum(reset_se), d["reset_se_checksum"]):
        raise AssertionError("regenerated reset SE frames do not match the fixture checksum")
    if not np.array_equal(step_tr, d["traffic"]):
        raise AssertionError("regenerated traffic frames do not match the fixture")
    return reset_se, step_se, step_tr


def quadriga_bank_powers(meta, d, n_eps=8):
    """Target-cell powers of the files a quadriga trace visits, laid out as an SE-trace bank wants
    them: [n_eps][T][135][25] (file order per frame), slot = episode % n_eps, unused slots zero."""
    T = int(meta["max_steps"])
    eps = sorted(set(d["episode"].tolist()) | set(d["reset_episode"].tolist()))
    assert len({e % n_eps for e in eps}) == len(eps)
    power = np.zeros((n_eps, T, R, U))
    for e in eps:
        a, f = quadriga_file_of_episode(meta["channel"], int(e))
        power[e % n_eps] = quadriga_power(a, f, T)[:, :, 0, 0, :]
    return power


def quadriga_episode_powers(meta, d):
    """The same files laid out the way the env surface wants them (QuadrigaChannel.register_power): one slot per env
    episode of the contiguous range the trace visits, slot = episode - first.  Returns (first_episode, power)."""
    T = int(meta["max_steps"])
    eps = sorted(set(d["episode"].tolist()) | set(d["reset_episode"].tolist()))
    first, last = int(eps[0]), int(eps[-1])
    power = np.zeros((last - first + 1, T, R, U))
    for e in range(first, last + 1):
        a, f = quadriga_file_of_episode(meta["channel"], e)
        power[e - first] = quadriga_power(a, f, T)[:, :, 0, 0, :]
    return first, power


# ---- BASELINE config 2 as a batch: 64 envs x 1000 steps recorded from the unmodified reference ------------------
def load_config2():
    f = np.load(os.path.join(GOLDEN, "config2_batch.npz"), allow_pickle=False)
    return {k: f[k] for k in f.files}


class Config2Frames:
    """SE / traffic frames of a batch fixture, regenerated env by env from the recipe (env e: seed seed_base + e;
    oracle/refharness/gen_config2.py, gen_batch_fixtures.py) and verified against the stored checksums."""

    def __init__(self, d, envs, seed_base=10, channel="mimic"):
        st, us = load_scenarios()
        self.d, self.envs = d, list(envs)
        self.src = [FrameSource(seed_base + e, channel, st, us) for e in self.envs]
        self.k = 0
        self.n_reset = 0

    def reset_frames(self):
        ri = self.n_reset
        se = np.stack([s.reset_frame(int(self.d["reset_episode"][e, ri]) % 200) for s, e in zip(self.src, self.envs)])
        if not np.array_equal(bits_checksum(se), self.d["reset_se_checksum"][self.envs, ri]):
            raise AssertionError("regenerated reset SE frames do not match the fixture checksum")
        self.n_reset += 1
        return se

    def step_frames(self):
        k = self.k
        out = [s.step_frame(int(self.d["episode"][e, k]) % 200) for s, e in zip(self.src, self.envs)]
        se, tr = np.stack([o[0] for o in out]), np.stack([o[1] for o in out])
        if not np.array_equal(bits_checksum(se), self.d["se_checksum"][self.envs, k]):
            raise AssertionError("regenerated SE frames do not match the fixture checksum (NumPy RNG stream differs)")
        if not np.array_equal(tr, self.d["traffic_mbit"][self.envs, k].astype(np.float64) * 1e6):
            raise AssertionError("regenerated traffic frames do not match the fixture")
        self.k += 1
        return se, tr


def replay_config2(env, d, envs, n_steps=None, rtol=1e-5):
    """Drive `env` (reset(initial_episodes, se) / step(a0, intra, se, traffic) / buffers(), batch = len(envs)) with the
    fixture's actions and frames as ONE batch and compare every output the fixture holds: integers bit-exact at every
    step, rewards at every step and observations / occupancy / delay at every 8th step against float32 (rtol 1e-5,
    the tolerance of the path), buffers at the checkpoints."""
    envs = list(envs)
    frames = Config2Frames(d, envs)
    n = int(d["meta_n_steps"]) if n_steps is None else n_steps
    every = int(d["meta_obs_every"])
    ie = d["reset_episode"][envs, 0].astype(np.int32)
    out = env.reset(initial_episodes=ie, se=frames.reset_frames())

    def close(tag, got, want, k):
        np.testing.assert_allclose(np.asarray(got, np.float64), np.asarray(want, np.float64), rtol=rtol, atol=1e-30,
                                   err_msg=f"config 2: {tag} at step {k}")

    def exact(tag, got, want, k):
        got, want = np.asarray(got), np.asarray(want)
        assert np.array_equal(got.astype(np.int64), want.astype(np.int64)), f"config 2: {tag} differs at step {k}"

    for key in ("obs0", "obs_i"):
        close("reset " + key, out[key], d["reset_" + key][envs, 0], -1)
    for key in ("mask0", "mask_i", "sorted_slices"):
        exact("reset " + key, out[key], d["reset_" + key][envs, 0], -1)
    ck = 0
    for k in range(n):
        se, tr = frames.step_frames()
        out = env.step(d["a0"][envs, k], d["intra"][envs, k], se, tr)
        for key, gold in (("rb_start", "rb_start"), ("rb_count", "rb_count"), ("pkt_incoming", "pkt_incoming"),
                          ("pkt_throughputs", "pkt_throughputs"), ("sent", "pkt_effective_thr"),
                          ("dropped", "dropped_pkts"), ("mask0", "mask0"), ("mask_i", "mask_i"),
                          ("sorted_slices", "sorted_slices"), ("done", "done")):
            exact(key, out[key], d[gold][envs, k], k)
        close("reward", out["reward"], d["reward"][envs, k], k)
        if k % every == every - 1:
            j = k // every
            close("obs0", out["obs0"], d["obs0"][envs, j], k)
            close("obs_i", out["obs_i"], d["obs_i"][envs, j], k)
            close("occ", out["occ"], d["buffer_occupancies"][envs, j], k)
            close("lat", out["lat"], d["buffer_latencies"][envs, j], k)
        if ck < d["ckpt_at_step"].shape[1] and d["ckpt_at_step"][envs[0], ck] == k:
            buf = np.asarray(env.buffers())
            for i, e in enumerate(envs):
                for u in range(U):
                    dep = int(d["ckpt_depth"][e, ck, u])
                    assert np.array_equal(buf[i, u, :dep], d["ckpt_buf"][e, ck, u, :dep]), f"buffer env {e} UE {u} step {k}"
            ck += 1
    return ck


def load_batch(name):
    f = np.load(os.path.join(GOLDEN, f"batch_{name}.npz"), allow_pickle=False)
    d = {k: f[k] for k in f.files}
    meta = {k[5:]: (v.item() if v.shape == () else v) for k, v in d.items() if k.startswith("meta_")}
    return meta, d


def replay_batch(env, meta, d, rtol=1e-5):
    """Replay one of the 16-env fixtures of oracle/refharness/gen_batch_fixtures.py as ONE batch: injected frames, the
    recorded actions, every episode end followed by a reset of the whole batch (the envs run in lockstep).  `env`:
    reset(initial_episodes | None, se) / step(a0, intra, se, traffic) / buffers()."""
    E, n, every = int(meta["n_envs"]), int(meta["n_steps"]), int(meta["obs_every"])
    envs = list(range(E))
    frames = Config2Frames(d, envs, seed_base=int(meta["seed_base"]), channel=str(meta["channel"]))
    flat = np.isnan(d["reward"][0, 0, 1])

    def close(tag, got, want, k):
        np.testing.assert_allclose(np.asarray(got, np.float64), np.asarray(want, np.float64), rtol=rtol, atol=1e-30,
                                   err_msg=f"batch {meta['agent']}: {tag} at step {k}")

    def exact(tag, got, want, k):
        assert np.array_equal(np.asarray(got).astype(np.int64), np.asarray(want).astype(np.int64)), \
            f"batch {meta['agent']}: {tag} differs at step {k}"

    def check_reset(out, ri, k):
        close("reset obs0", out["obs0"], d["reset_obs0"][:, ri], k)
        for key in ("mask0", "mask_i", "sorted_slices"):
            exact("reset " + key, out[key], d["reset_" + key][:, ri], k)

    ri = 0
    out = env.reset(initial_episodes=d["reset_episode"][:, 0].astype(np.int32), se=frames.reset_frames())
    check_reset(out, 0, -1)
    ck = n_done = 0
    for k in range(n):
        se, tr = frames.step_frames()
        out = env.step(d["a0"][:, k], np.maximum(d["intra"][:, k], 0), se, tr)
        for key, gold in (("rb_start", "rb_start"), ("rb_count", "rb_count"), ("pkt_incoming", "pkt_incoming"),
                          ("pkt_throughputs", "pkt_throughputs"), ("sent", "pkt_effective_thr"),
                          ("dropped", "dropped_pkts"), ("mask0", "mask0"), ("mask_i", "mask_i"),
                          ("sorted_slices", "sorted_slices"), ("done", "done")):
            exact(key, out[key], d[gold][:, k], k)
        close("reward", out["reward"][:, 0] if flat else out["reward"], d["reward"][:, k, 0] if flat else d["reward"][:, k], k)
        if k % every == every - 1:
            j = k // every
            close("obs0", out["obs0"], d["obs0"][:, j], k)
            if "obs_i" in d:
                close("obs_i", out["obs_i"], d["obs_i"][:, j], k)
            close("occ", out["occ"], d["buffer_occupancies"][:, j], k)
            close("lat", out["lat"], d["buffer_latencies"][:, j], k)
        if ck < d["ckpt_at_step"].shape[1] and d["ckpt_at_step"][0, ck] == k:
            buf = np.asarray(env.buffers())
            for e in range(E):
                for u in range(U):
                    dep = int(d["ckpt_depth"][e, ck, u])
                    assert np.array_equal(buf[e, u, :dep], d["ckpt_buf"][e, ck, u, :dep]), f"buffer env {e} UE {u} step {k}"
            ck += 1
        if d["done"][0, k]:
            assert d["done"][:, k].all()
            n_done += 1
            ri += 1
            out = env.reset(initial_episodes=None, se=frames.reset_frames())
            check_reset(out, ri, k)
    return n_done
