"""Record BASELINE config 2 from the UNMODIFIED reference as ONE batch fixture (runs only in the build container).

    python -m oracle.refharness.gen_config2 [n_envs] [n_procs]     # writes tests/golden/config2_batch.npz

SURVEY.md 8(d) config 2: env e plays association episode e mod 200 with env seed 10 + e and actions drawn by
default_rng(1234 + e) (player_0 ~ U(-1, 1)^5, player_1..5 ~ integers{0, 1, 2}); MimicQuadriga channel, IBSched dict
surface, 1000 steps (one whole episode, the rollover included).  Every integer output of every step is stored
(RB runs, packet counts, masks, slice order); rewards of every step and the observation vectors of every 8th step are
stored as float32 (the tolerance of the path is 1e-5 relative); SE and traffic frames are stored as a recipe
(env seed) + per-frame checksums, like the single-env traces (tests/golden_inputs.py).  TEST INFRASTRUCTURE ONLY.
"""
import os
import sys
from multiprocessing import Pool

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

GOLDEN = os.path.join(ROOT, "tests", "golden")
WORK = "/tmp/irs_work"
N_STEPS, OBS_EVERY = 1000, 8


def one_env(e):
    from oracle.refharness import gen_golden, refenv
    from golden_inputs import bits_checksum

    env, ag = refenv.make_env(WORK, agent="ib_sched", channel="mimic", assoc="mult_slice", seed=10 + e,
                              initial_episode=e % 200)
    steps, resets, ckpt = refenv.run_trace(env, ag, N_STEPS, gen_golden.rand_actions(1234 + e), 10 + e, e % 200,
                                           checkpoint_every=250)
    out = {k: steps[k] for k in ("a0", "intra", "rb_start", "rb_count", "mask0", "mask_i", "sorted_slices", "done", "episode")}
    for k in ("pkt_incoming", "pkt_throughputs", "pkt_effective_thr", "dropped_pkts"):
        v = np.asarray(steps[k])
        assert np.array_equal(v, np.floor(v)) and np.abs(v).max() < 2 ** 31
        out[k] = v.astype(np.int32)
    out["traffic_mbit"] = (steps["traffic"] / 1e6).astype(np.int16)
    assert np.array_equal(out["traffic_mbit"].astype(np.float64) * 1e6, steps["traffic"])
    out["reward"] = steps["reward"].astype(np.float32)
    for k in ("obs0", "obs_i", "buffer_occupancies", "buffer_latencies"):
        out[k] = steps[k][OBS_EVERY - 1::OBS_EVERY].astype(np.float32)
    out["se_checksum"] = bits_checksum(steps["se"])
    out["reset_se_checksum"] = bits_checksum(resets["se"])
    out["reset_episode"] = resets["episode"]
    for k in ("obs0", "obs_i", "mask0", "mask_i", "sorted_slices"):
        out["reset_" + k] = resets[k]
    out["ckpt_at_step"], out["ckpt_buf"], out["ckpt_depth"] = ckpt["at_step"], ckpt["buf"], ckpt["depth"]
    return out


def main():
    n_envs = int(sys.argv[1]) if len(sys.argv) > 1 else 64
    n_procs = int(sys.argv[2]) if len(sys.argv) > 2 else min(os.cpu_count() or 1, 16)
    from oracle.refharness import refenv

    refenv.make_workdir(WORK, n_scenarios=200)
    with Pool(n_procs) as pool:
        outs = pool.map(one_env, range(n_envs))
    data = {k: np.stack([o[k] for o in outs]) for k in outs[0]}
    data["meta_n_envs"], data["meta_n_steps"], data["meta_obs_every"] = np.array(n_envs), np.array(N_STEPS), np.array(OBS_EVERY)
    data["meta_numpy"] = np.array(np.__version__)
    path = os.path.join(GOLDEN, "config2_batch.npz")
    np.savez_compressed(path, **data)
    print(f"{path}: {n_envs} envs x {N_STEPS} steps, {os.path.getsize(path) / 1e6:.1f} MB")


if __name__ == "__main__":
    main()
