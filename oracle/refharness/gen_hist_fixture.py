"""Record one episode history file with the UNMODIFIED reference (save_hist=True; metrics.py:73-100) and commit a
description of its STRUCTURE plus the file itself (6-step episodes keep it small) - the format contract that
intent_radio_sched_multi_slice_b200/history.py must meet.  Runs only in the build container.

    python -m oracle.refharness.gen_hist_fixture        # writes tests/golden/hist_ref_ib_sched_ep_4.npz, hist_ref_marr_ep_4.npz
"""
import os
import shutil
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from oracle.refharness import gen_golden, refenv, shims  # noqa: E402

GOLDEN = os.path.join(ROOT, "tests", "golden")
WORK = "/tmp/irs_work"
T = 6


def record(agent, action_fn, episode=4, seed=31):
    wd = f"{WORK}_hist"
    refenv.make_workdir(wd, n_scenarios=0, max_number_steps=T)
    dst = f"{wd}/associations/data"
    os.makedirs(dst, exist_ok=True)
    if not os.path.exists(f"{dst}/mult_slice"):
        os.symlink(f"{WORK}/associations/data/mult_slice", f"{dst}/mult_slice")
    yml = open(f"{wd}/env_config/mult_slice.yml").read().replace("hist_root_path: ./", f"hist_root_path: {wd}/")
    open(f"{wd}/env_config/mult_slice.yml", "w").write(yml)
    shims.install()
    from associations.mult_slice import MultSliceAssociation
    from channels.mimic_quadriga import MimicQuadriga
    from mobilities.simple import SimpleMobility
    from sixg_radio_mgmt import MARLCommEnv
    from traffics.mult_slice import MultSliceTraffic

    env = MARLCommEnv(MimicQuadriga, MultSliceTraffic, SimpleMobility, MultSliceAssociation, "mult_slice", agent, seed,
                      root_path=wd, initial_episode_number=episode, simu_name="fixture", save_hist=True,
                      enable_random_episodes=False, debug=False)
    from agents.ib_sched import IBSched
    from agents.marr import MARR

    ag = {"ib_sched": IBSched, "marr": MARR}[agent](env, 25, 5, 1, np.array([135]))
    env.set_agent_functions(ag.obs_space_format, ag.action_format, ag.calculate_reward, ag.get_obs_space(),
                            ag.get_action_space())
    shims.pin_stable_argsort()
    try:
        obs, _ = env.reset(seed=seed, options={"initial_episode": episode})
        for k in range(T):
            obs, *_ = env.step(action_fn(k, obs, ag))
    finally:
        shims.unpin_argsort()
    src = f"{wd}/hist/fixture/{agent}/ep_{episode}.npz"
    out = os.path.join(GOLDEN, f"hist_ref_{agent}_ep_{episode}.npz")
    shutil.copy(src, out)
    f = np.load(out, allow_pickle=True)
    print(out, os.path.getsize(out), "bytes")
    for k in f.files:
        a = f[k]
        print(f"  {k}: dtype {a.dtype} shape {a.shape}" + (f" element {type(a.flat[0]).__name__}" if a.dtype == object else ""))


def main():
    refenv.make_workdir(WORK, n_scenarios=200)
    record("ib_sched", gen_golden.rand_actions(3))
    record("marr", gen_golden.policy)


if __name__ == "__main__":
    main()
