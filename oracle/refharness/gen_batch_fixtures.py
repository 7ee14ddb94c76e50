"""Record small BATCH fixtures of the other agents / channels from the UNMODIFIED reference (build container only).

    python -m oracle.refharness.gen_batch_fixtures            # writes tests/golden/batch_<name>.npz

Like gen_config2.py (env e: association episode (7 e) mod 200, env seed 100 + e, actions from the agent's own policy or
default_rng(4321 + e)), but 16 envs x 300 steps of 100-step episodes, so that every env crosses two rollovers, for:
MAPF and MARR (policy-driven, flat surface), SchedTWC and SchedColORAN (their own observation / reward surfaces, flat
random actions), IBSched over FixedSE (exact ties: the configuration where summation order and the argsort tie rule
show).  Integers of every step, rewards of every step and observations of every 4th step (float32).
TEST INFRASTRUCTURE ONLY.
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
N_ENVS, N_STEPS, MAX_STEPS, OBS_EVERY = 16, 300, 100, 4
FIXTURES = {  # name: agent, channel, action source
    "mapf": ("mapf", "mimic", "policy"), "marr": ("marr", "mimic", "policy"),
    "twc": ("sched_twc", "mimic", "flat"), "colran": ("sched_colran", "mimic", "flat"),
    "ibsched_fixed": ("ib_sched", "fixed", "dict"),
}


def one_env(job):
    name, e = job
    agent, channel, src = FIXTURES[name]
    from oracle.refharness import gen_golden, refenv
    from golden_inputs import bits_checksum

    wd = f"{WORK}_s{MAX_STEPS}"
    env, ag = refenv.make_env(wd, agent=agent, channel=channel, assoc="mult_slice", seed=100 + e,
                              initial_episode=(7 * e) % 200)
    fn = {"policy": gen_golden.policy, "flat": gen_golden.rand_flat(4321 + e), "dict": gen_golden.rand_actions(4321 + e)}[src]
    steps, resets, ckpt = refenv.run_trace(env, ag, N_STEPS, fn, 100 + e, (7 * e) % 200, checkpoint_every=150)
    out = {k: steps[k] for k in ("a0", "intra", "rb_start", "rb_count", "mask0", "mask_i", "sorted_slices", "done", "episode")}
    for k in ("pkt_incoming", "pkt_throughputs", "pkt_effective_thr", "dropped_pkts"):
        v = np.asarray(steps[k])
        assert np.array_equal(v, np.floor(v)) and np.abs(v).max() < 2 ** 31
        out[k] = v.astype(np.int32)
    out["traffic_mbit"] = (steps["traffic"] / 1e6).astype(np.int16)
    assert np.array_equal(out["traffic_mbit"].astype(np.float64) * 1e6, steps["traffic"])
    out["reward"] = steps["reward"].astype(np.float32)
    obs_key = "agent_obs" if "agent_obs" in steps else "obs0"
    out["obs0"] = steps[obs_key][OBS_EVERY - 1::OBS_EVERY].astype(np.float32)
    out["reset_obs0"] = resets[obs_key].astype(np.float32)
    if "agent_obs" not in steps:
        out["obs_i"] = steps["obs_i"][OBS_EVERY - 1::OBS_EVERY].astype(np.float32)
    for k in ("buffer_occupancies", "buffer_latencies"):
        out[k] = steps[k][OBS_EVERY - 1::OBS_EVERY].astype(np.float32)
    out["se_checksum"] = bits_checksum(steps["se"])
    out["reset_se_checksum"] = bits_checksum(resets["se"])
    out["reset_episode"] = resets["episode"]
    for k in ("mask0", "mask_i", "sorted_slices"):
        out["reset_" + k] = resets[k]
    out["ckpt_at_step"], out["ckpt_buf"], out["ckpt_depth"] = ckpt["at_step"], ckpt["buf"], ckpt["depth"]
    return out


def main():
    from oracle.refharness import refenv

    refenv.make_workdir(WORK, n_scenarios=200)
    wd = f"{WORK}_s{MAX_STEPS}"
    refenv.make_workdir(wd, n_scenarios=0, max_number_steps=MAX_STEPS)
    os.makedirs(f"{wd}/associations/data", exist_ok=True)
    if not os.path.exists(f"{wd}/associations/data/mult_slice"):
        os.symlink(f"{WORK}/associations/data/mult_slice", f"{wd}/associations/data/mult_slice")
    names = sys.argv[1:] or list(FIXTURES)
    with Pool(min(os.cpu_count() or 1, 8)) as pool:
        for name in names:
            outs = pool.map(one_env, [(name, e) for e in range(N_ENVS)])
            data = {k: np.stack([o[k] for o in outs]) for k in outs[0]}
            agent, channel, src = FIXTURES[name]
            data.update(meta_agent=np.array(agent), meta_channel=np.array(channel), meta_n_envs=np.array(N_ENVS),
                        meta_n_steps=np.array(N_STEPS), meta_max_steps=np.array(MAX_STEPS), meta_obs_every=np.array(OBS_EVERY),
                        meta_seed_base=np.array(100), meta_numpy=np.array(np.__version__))
            path = os.path.join(GOLDEN, f"batch_{name}.npz")
            np.savez_compressed(path, **data)
            print(f"{path}: {N_ENVS} envs x {N_STEPS} steps, {os.path.getsize(path) / 1e6:.2f} MB, "
                  f"{int(data['done'].sum())} episode ends")


if __name__ == "__main__":
    main()
