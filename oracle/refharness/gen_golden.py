"""Record golden traces from the UNMODIFIED reference (runs only in the build container).

    python -m oracle.refharness.gen_golden            # writes tests/golden/*.npz

Each trace is also replayed through the C oracle (oracle/irs_oracle.c) and must agree
bit-for-bit before it is written, and its SE frames must be reproducible from the recipe by
tests/golden_inputs.py.  TEST INFRASTRUCTURE ONLY.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from oracle.refharness import refenv, shims  # noqa: E402

GOLDEN = os.path.join(ROOT, "tests", "golden")
WORK = "/tmp/irs_work"


def rand_actions(seed):
    rng = np.random.default_rng(seed)

    def f(k, obs, ag):
        d = {"player_0": rng.uniform(-1, 1, 5)}
        for i in range(1, 6):
            d[f"player_{i}"] = int(rng.integers(0, 3))
        return d

    return f


def rand_flat(seed):
    rng = np.random.default_rng(seed)
    return lambda k, obs, ag: rng.uniform(-1, 1, 5)


def policy(k, obs, ag):
    return ag.step(obs)


TRACES = {
    # name: agent, channel, assoc, seed, initial_episode, n_steps, max_steps, max_episodes, action source
    "cfg1_marr": ("marr", "mimic", "mult_slice", 10, 0, 1050, 1000, None, None),
    "mapf": ("mapf", "mimic", "mult_slice", 11, 5, 300, 1000, None, None),
    "ibsched_mimic": ("ib_sched", "mimic", "mult_slice", 12, 2, 500, 1000, None, 1),
    "ibsched_fixed": ("ib_sched", "fixed", "mult_slice", 13, 3, 400, 1000, None, 2),
    "seq_mimic": ("ib_sched", "mimic", "mult_slice_seq", 14, 98, 240, 40, None, 3),
    "seq_fixed": ("ib_sched", "fixed", "mult_slice_seq", 15, 99, 160, 40, None, 4),
    "sweep_scen": ("ib_sched", "mimic", "mult_slice", 16, 20, 600, 25, 36, 5),
    # QuadrigaChannel / QuadrigaChannelSeq over synthetic power files (shims._install_h5py)
    "quadriga": ("ib_sched", "quadriga", "mult_slice", 17, 3, 100, 30, None, 6),
    "quadriga_seq": ("ib_sched", "quadriga_seq", "mult_slice_seq", 18, 99, 100, 30, None, 7),
    # the baseline agents' surfaces (negative action seed: flat 5-vector actions)
    "twc": ("sched_twc", "mimic", "mult_slice", 19, 4, 300, 100, None, -8),
    "colran": ("sched_colran", "mimic", "mult_slice", 20, 8, 300, 100, None, -9),
    # enable_random_episodes=True (every training configuration of simu.py:220-328): 25-step episodes shuffled over
    # [3, 12); 14 rollovers wrap the 9-episode cycle; reset 8 passes options={"initial_episode": 6} (new permutation)
    "random_eps": ("ib_sched", "mimic", "mult_slice", 21, 3, 375, 25, 12, 10),
}
EXTRA = {"random_eps": dict(random_episodes=True, reperm={8: 6})}


def record(name):
    agent, channel, assoc, seed, init_ep, n_steps, max_steps, max_eps, act_seed = TRACES[name]
    wd = WORK if max_steps == 1000 else f"{WORK}_s{max_steps}"
    if max_steps != 1000:
        refenv.make_workdir(wd, n_scenarios=0, max_number_steps=max_steps)
        src = f"{WORK}/associations/data/mult_slice"
        dst = f"{wd}/associations/data"
        os.makedirs(dst, exist_ok=True)
        if not os.path.exists(f"{dst}/mult_slice"):
            os.symlink(src, f"{dst}/mult_slice")
    shims.QUADRIGA_STEPS["n"] = max_steps
    extra = EXTRA.get(name, {})
    env, ag = refenv.make_env(wd, agent=agent, channel=channel, assoc=assoc, seed=seed,
                              initial_episode=init_ep, max_episode_number=max_eps,
                              random_episodes=extra.get("random_episodes", False))
    fn = policy if act_seed is None else (rand_actions(act_seed) if act_seed > 0 else rand_flat(-act_seed))
    steps, resets, ckpt = refenv.run_trace(env, ag, n_steps, fn, seed, init_ep, reperm=extra.get("reperm"))
    if extra.get("random_episodes"):  # the env played what the predicted permutations say
        pos, cur = 0, -1
        for ri, (init, ep) in enumerate(zip(resets["initial"], resets["episode"])):
            if init != -1:
                cur, pos = cur + 1, 0
                if ri > 0:
                    pos = 1  # that reset also ends an episode: next() right after the new permutation
            elif ri > 0:
                pos += 1
            order = resets[f"order_{cur}"]
            assert ep == order[pos % len(order)], (ri, ep, order, pos)
    from golden_inputs import bits_checksum

    out = {
        "meta_agent": np.array(agent), "meta_channel": np.array(channel), "meta_assoc": np.array(assoc),
        "meta_seed": np.array(seed), "meta_initial_episode": np.array(init_ep), "meta_n_steps": np.array(n_steps),
        "meta_max_steps": np.array(max_steps), "meta_max_episodes": np.array(10000 if max_eps is None else max_eps),
        "meta_numpy": np.array(np.__version__), "meta_random_episodes": np.array(int(bool(extra.get("random_episodes")))),
        "se_checksum": bits_checksum(steps["se"]), "reset_se_checksum": bits_checksum(resets["se"]),
    }
    if channel.startswith("quadriga"):
        out["se_sum"] = steps["se"].reshape(len(steps["se"]), -1).sum(1)
        out["reset_se_sum"] = resets["se"].reshape(len(resets["se"]), -1).sum(1)
    for k, v in steps.items():
        if k != "se":
            out[k] = v
    for k, v in resets.items():
        if k != "se":
            out["reset_" + k] = v
    for k, v in ckpt.items():
        out["ckpt_" + k] = v
    return out, steps, resets


def main():
    refenv.make_workdir(WORK, n_scenarios=200)
    os.makedirs(GOLDEN, exist_ok=True)
    st, us = refenv.load_scenarios(WORK, 200)
    np.savez_compressed(f"{GOLDEN}/scenarios_seed10.npz", slice_type=st, ue_slice=us)
    # the slice-type table restated in the product must equal the reference's
    shims.install()
    from associations.mult_slice import MultSliceAssociation
    from intent_radio_sched_multi_slice_b200 import scenario as scn

    ref = MultSliceAssociation(None, 25, 1, 5)  # type: ignore
    for i, nm in enumerate(ref.slice_types):
        a, b = ref.slice_type_model[nm], scn.slice_req_from_type(i)
        assert a["name"] == b["name"] and a["priority"] == b["priority"], nm
        for k in b["ues"]:
            assert a["ues"][k] == b["ues"][k], (nm, k)
        pa = {p["name"]: (p["value"], p["operator"].__name__) for p in a["parameters"].values()}
        pb = {p["name"]: (p["value"], {"at_least": "greater_equal", "at_most": "less_equal"}[p["operator"]])
              for p in b["parameters"].values()}
        assert pa == pb, (nm, pa, pb)
    print("slice-type table matches associations/mult_slice.py")
    names = sys.argv[1:] or list(TRACES)
    import golden_inputs

    for name in names:
        out, steps, resets = record(name)
        np.savez_compressed(f"{GOLDEN}/trace_{name}.npz", **out)
        meta, d = golden_inputs.load_trace(name)
        rse, sse, stf = golden_inputs.regenerate(meta, d)
        assert np.array_equal(rse, resets["se"]) and np.array_equal(sse, steps["se"])
        sz = os.path.getsize(f"{GOLDEN}/trace_{name}.npz")
        print(f"{name}: {meta['n_steps']} steps, {len(d['reset_episode'])} resets, episodes "
              f"{sorted(set(d['episode'].tolist()))[:8]}..., {sz/1e6:.2f} MB, SE regeneration exact")


if __name__ == "__main__":
    main()
