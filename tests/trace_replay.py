"""Replay a golden trace through any env implementing reset()/step() with injected frames
(the CPU oracle in the CPU suite, the CUDA path in the GPU suite) and compare every output."""
from __future__ import annotations

import numpy as np

import golden_inputs

EXACT = ["rb_start", "rb_count", "pkt_incoming", "pkt_throughputs", "sent", "dropped", "mask0", "mask_i",
         "sorted_slices", "done"]
CLOSE = ["occ", "lat", "obs0", "obs_i"]
RTOL = 1e-5  # BASELINE.json north_star tolerance for throughput/observations/reward
GOLD_KEY = {"sent": "pkt_effective_thr", "dropped": "dropped_pkts", "occ": "buffer_occupancies",
            "lat": "buffer_latencies"}


def agent_config(meta):
    """(sort_slices, fixed_intra) of the reference agent (ib_sched.py:57; marr.py:66; mapf.py:130)."""
    return {"ib_sched": (1, -1), "marr": (0, 0), "mapf": (0, 1), "sched_twc": (0, 0), "sched_colran": (0, 0)}[meta["agent"]]


def agent_surface(meta):
    """cfg.surface of the reference agent: 0 IBSched family, 1 SchedTWC, 2 SchedColORAN."""
    return {"sched_twc": 1, "sched_colran": 2}.get(meta["agent"], 0)


def replay(name, make_env, tight=False):
    """make_env(meta) -> object with reset(initial_episode, se[1,U,R]) -> dict,
    step(a0[1,5], intra[1,5], se[1,U,R], traffic[1,U]) -> dict, buffers() -> [1,U,D]."""
    meta, d = golden_inputs.load_trace(name)
    rse, sse, stf = golden_inputs.regenerate(meta, d)
    env = make_env(meta)
    rtol = 1e-12 if tight else RTOL

    def cmp(tag, got, want, k, exact):
        got, want = np.asarray(got), np.asarray(want)
        if exact:
            assert np.array_equal(got.astype(want.dtype), want), f"{name}: {tag} differs at step {k}\n{got}\n{want}"
        else:
            np.testing.assert_allclose(got, want, rtol=rtol, atol=1e-300, err_msg=f"{name}: {tag} at step {k}")

    alt = "agent_obs" in d  # SchedTWC / SchedColORAN: the surface is the agent's own 50-vector; the IBSched dict of
    #                         their internal fake agent (computed before the frame's second push) is not observable

    def cmp_reset(out, ri, k):
        if alt:
            cmp("reset agent obs", out["obs0"][0], d["reset_agent_obs"][ri], k, False)
        for key in (() if alt else ("obs0", "obs_i")):
            cmp("reset " + key, out[key][0], d["reset_" + key][ri], k, False)
        for key in ("mask0", "mask_i", "sorted_slices"):
            cmp("reset " + key, out[key][0], d["reset_" + key][ri], k, True)

    ri = 0
    random_eps = bool(int(meta.get("random_episodes", 0)))
    n_perm = 0
    if random_eps:  # replay the permutation the reference drew from its np_random (comm_env.py:435-442)
        env.set_episode_order(d["reset_order_0"][None])
        n_perm = 1
    out = env.reset(initial_episode=int(meta["initial_episode"]), se=rse[0][None])
    cmp_reset(out, 0, -1)
    if random_eps:
        assert int(env.state()["episode"][0]) == int(d["reset_episode"][0])
    ck = 0
    for k in range(int(meta["n_steps"])):
        out = env.step(d["a0"][k][None], np.maximum(d["intra"][k], 0)[None], sse[k][None], stf[k][None])
        for key in EXACT:
            cmp(key, out[key][0], d[GOLD_KEY.get(key, key)][k], k, True)
        for key in CLOSE:
            if alt and key in ("obs0", "obs_i"):
                continue
            cmp(key, out[key][0], d[GOLD_KEY.get(key, key)][k], k, False)
        if alt:
            cmp("agent obs", out["obs0"][0], d["agent_obs"][k], k, False)
        r = d["reward"][k]
        if np.isnan(r[1]):
            cmp("reward", out["reward"][0, 0], r[0], k, False)
        else:
            cmp("reward", out["reward"][0], r, k, False)
        if ck < len(d["ckpt_at_step"]) and d["ckpt_at_step"][ck] == k:
            buf = np.asarray(env.buffers())[0]
            want = d["ckpt_buf"][ck]
            depth = d["ckpt_depth"][ck]
            for u in range(25):
                cmp(f"buffer[{u}]", buf[u, : depth[u]], want[u, : depth[u]], k, True)
            ck += 1
        if d["done"][k]:
            ri += 1
            init = int(d["reset_initial"][ri]) if random_eps else -1
            if random_eps and init != -1:
                env.set_episode_order(d[f"reset_order_{n_perm}"][None])
                n_perm += 1
            out = env.reset(initial_episode=init, se=rse[ri][None])
            cmp_reset(out, ri, k)
            if random_eps:
                assert int(env.state()["episode"][0]) == int(d["reset_episode"][ri]), (ri, k)
    assert ck == len(d["ckpt_at_step"])
    return env
