"""How often does the reference's UNPINNED np.argsort differ from the stable order the oracle pins?

SURVEY.md section 7 hard part 1: agents/common.py:496 (`np.argsort(x)[::-1]` in round_int_equal_sum: which entries get
the +1 RBs) and agents/ib_sched.py:370 (sort_slices) call np.argsort with the default kind, whose tie order depends on
the NumPy build and the host's SIMD level (AVX-512 hosts use a vectorised sort).  The golden traces are recorded with
np.argsort pinned to kind="stable" (shims.pin_stable_argsort).  This script runs the unmodified reference with a
COUNTING wrapper: every call computes both orders, returns the stable one (so the run stays the pinned run) and
counts calls whose default order differs, and steps in which at least one such call happened.

    python -m oracle.refharness.argsort_divergence [steps]      # build container only
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from oracle.refharness import gen_golden, refenv, shims  # noqa: E402

WORK = "/tmp/irs_work"


def measure(agent, channel, seed, episode, steps, action_fn):
    env, ag = refenv.make_env(WORK, agent=agent, channel=channel, assoc="mult_slice", seed=seed, initial_episode=episode)
    orig = shims._orig_argsort
    stat = {"calls": 0, "ties": 0, "differ": 0, "dirty_step": False}

    def counting(a, axis=-1, kind=None, order=None, **kw):
        st = orig(a, axis=axis, kind="stable", order=order, **kw)
        if kind is None and np.ndim(a) == 1:
            stat["calls"] += 1
            arr = np.asarray(a)
            if len(np.unique(arr)) < arr.size:
                stat["ties"] += 1
            if not np.array_equal(orig(a, axis=axis, order=order, **kw), st):
                stat["differ"] += 1
                stat["dirty_step"] = True
        return st

    np.argsort = counting
    dirty = 0
    try:
        obs, _ = env.reset(seed=seed, options={"initial_episode": episode})
        for k in range(steps):
            stat["dirty_step"] = False
            obs, _, term, _, _ = env.step(action_fn(k, obs, ag))
            dirty += stat["dirty_step"]
            if (term if not isinstance(term, dict) else term["__all__"]):
                obs, _ = env.reset()
    finally:
        np.argsort = orig
    return stat["calls"], stat["ties"], stat["differ"], dirty


def main():
    steps = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
    refenv.make_workdir(WORK, n_scenarios=200)
    print(f"numpy {np.__version__}; avx512f on this host: {'avx512f' in open('/proc/cpuinfo').read()}")
    rows = [("ib_sched", "mimic", 12, 2, gen_golden.rand_actions(1)), ("ib_sched", "fixed", 13, 3, gen_golden.rand_actions(2)),
            ("marr", "mimic", 10, 0, gen_golden.policy), ("mapf", "mimic", 11, 5, gen_golden.policy)]
    for agent, channel, seed, ep, fn in rows:
        calls, ties, differ, dirty = measure(agent, channel, seed, ep, steps, fn)
        print(f"{agent:9s} {channel:6s}: {calls} argsort calls in {steps} steps, {ties} on vectors with ties, "
              f"{differ} where the default order differs from the stable one ({100.0 * differ / max(calls, 1):.3f} %), "
              f"{dirty} steps affected ({100.0 * dirty / steps:.2f} %)")


if __name__ == "__main__":
    main()
