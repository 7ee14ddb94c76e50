"""GPU suite: randomized scenario banks at the edges of the domain - slices without requirements, active slices
without UEs, an env with no slice at all, all 25 UEs associated, tiny buffers that overflow and expire every step,
huge buffers, degenerate actions (all -1: the equal-split branch of scores_to_rbs, common.py:449-455; exact ties:
the stable-argsort tie rule of round_int_equal_sum, common.py:481-505) - CUDA path vs the CPU oracle, every output
of every step, both per-env kernel families."""
import numpy as np
import pytest

from gpu_adapter import CudaEnv, compare_outputs

pytestmark = pytest.mark.gpu


def _random_bank(rng, n):
    from intent_radio_sched_multi_slice_b200 import scenario as scn

    bank = scn.ScenarioBank()
    for i in range(n):
        req, sua = {}, np.zeros((5, 25))
        kind = i % 6
        free = list(rng.permutation(25))
        for s in range(5):
            r = {}
            if kind != 0 and rng.random() < 0.8:                      # kind 0: no slice at all
                r = scn.slice_req_from_type(int(rng.integers(0, 10)))
                if kind == 1:                                          # tiny buffers: overflow + expiry every step
                    r["ues"]["buffer_size"] = int(rng.integers(3, 40))
                    r["ues"]["buffer_latency"] = int(rng.integers(1, 4))
                    for par in r["parameters"].values():
                        if par["name"] == "latency":
                            par["value"] = 0.5
                if kind == 2:                                          # deep and huge
                    r["ues"]["buffer_latency"] = 400
                    r["ues"]["buffer_size"] = 1024000
                if kind == 3 and rng.random() < 0.5:                   # slice declared but without requirements
                    r = {}
            n_ues = 5 if kind == 4 else int(rng.integers(0, 6))       # kind 4: all 25 UEs; 0 UEs in an active slice
            if r == {} and kind != 3:
                n_ues = 0
            for _ in range(n_ues):
                if free:
                    sua[s, free.pop()] = 1
            req[f"slice_{s}"] = r
        bsa = None
        if kind == 5:  # activity flag decoupled from the requirements (basestation_slice_assoc given explicitly)
            bsa = (rng.random(5) < 0.6).astype(float)[None]
        bank.add(req, sua, bsa)
    return bank


@pytest.mark.parametrize("family", ["one_launch", "warp_per_env", "lane8", "thread_per_env"])
@pytest.mark.parametrize("seed", [0, 1])
def test_fuzzed_scenarios_match_oracle(seed, family, monkeypatch):
    from oracle import oracle

    if family != "one_launch":  # k_env_step (default for small batches) / warp-per-env / thread-per-env kernels
        monkeypatch.setenv("IRS_FUSED_MAX_ENVS", "0")
        monkeypatch.setenv("IRS_PER_ENV", {"warp_per_env": "warp", "lane8": "lane8", "thread_per_env": "thread"}[family])
    rng = np.random.default_rng(100 + seed)
    n_scen, E, steps = 48, 192, 45
    bank = _random_bank(rng, n_scen)
    kw = dict(channel_mode=2, traffic_mode=1, policy=0, fixed_intra=-1, sort_slices=1, assoc_mode=0, scen_mod=n_scen,
              max_steps=11, auto_reset=1, env_id_base=9, seed=4321 + seed, initial_episode=0)
    orc = oracle.OracleEnv(bank.array(), bank.poisson_cdf(), E, depth_cap=404, **kw)
    gpu = CudaEnv(bank, E, depth_cap=401, **kw)
    ie = (np.arange(E) % n_scen).astype(np.int32)
    for e in range(E):
        m = np.zeros(E, np.uint8)
        m[e] = 1
        want = orc.reset(mask=m, initial_episode=int(ie[e]))
    got = gpu.reset(initial_episodes=ie, initial_episode=0)
    compare_outputs(got, want, "reset", keys=["obs0", "obs_i", "mask0", "mask_i", "sorted_slices"])
    for k in range(steps):
        a0 = rng.uniform(-1, 1, (E, 5))
        a0[rng.random(E) < 0.15] = -1.0                                   # sum(a+1) == 0: equal split
        tie = rng.random(E) < 0.15
        a0[tie] = np.round(a0[tie] * 2) / 2                               # exact ties between slices
        a0[rng.random(E) < 0.05] = 1.0
        intra = rng.integers(0, 3, (E, 5)).astype(np.int32)
        want, got = orc.step(a0, intra), gpu.step(a0, intra)
        compare_outputs(got, want, f"seed {seed} step {k}")
    assert np.array_equal(gpu.buffers()[:, :, :401], orc.buffers()[:, :, :401])
    assert int(np.asarray(want["dropped"]).sum()) >= 0
    per_g, tot_g = gpu.episode_metrics()
    per_o, tot_o = orc.episode_metrics()
    np.testing.assert_allclose(per_g, per_o, rtol=1e-12, atol=1e-300)


@pytest.mark.parametrize("depth,size", [(140, 90_000), (530, 1_024_000), (530, 250_000)])
def test_rows_age_to_the_expiry_slot_like_oracle(depth, size):
    """Rows that grow slot by slot to their full depth (across the 128-slot chunk and the 512-slot group boundaries of the
    row walk), hit the capacity clip on the way and then expire at the old end every step: the live-extent bookkeeping of
    the row kernel (irs_common.cuh ROW_LIVE) against the dense oracle - every output of every step, buffers at checkpoints."""
    from intent_radio_sched_multi_slice_b200 import scenario as scn
    from oracle import oracle

    rng = np.random.default_rng(depth + size)
    bank = scn.ScenarioBank()
    for i in range(4):
        req, sua = {}, np.zeros((5, 25))
        for s in range(5):
            r = scn.slice_req_from_type(5)  # 100 Mbit/s of 65,536-bit packets per UE against ~40 Mbit/s of service
            r["ues"]["buffer_latency"] = depth - 1 - (s if i else 0)
            r["ues"]["buffer_size"] = size
            req[f"slice_{s}"] = r
            sua[s, 5 * s:5 * s + (5 if i != 3 else 3)] = 1
        bank.add(req, sua)
    E, steps = 12, int(2.2 * depth)
    kw = dict(channel_mode=2, traffic_mode=1, policy=0, fixed_intra=-1, sort_slices=1, assoc_mode=0, scen_mod=4,
              max_steps=1_000_000, auto_reset=1, env_id_base=3, seed=77, initial_episode=0)
    stride = (depth + 3) // 4 * 4
    orc = oracle.OracleEnv(bank.array(), bank.poisson_cdf(), E, depth_cap=stride, **kw)
    gpu = CudaEnv(bank, E, depth_cap=depth, **kw)
    ie = (np.arange(E) % 4).astype(np.int32)
    for e in range(E):
        m = np.zeros(E, np.uint8)
        m[e] = 1
        orc.reset(mask=m, initial_episode=int(ie[e]))
    gpu.reset(initial_episodes=ie, initial_episode=0)
    expired = 0
    for k in range(steps):
        a0 = rng.uniform(-1, 1, (E, 5))
        intra = rng.integers(0, 3, (E, 5)).astype(np.int32)
        want, got = orc.step(a0, intra), gpu.step(a0, intra)
        compare_outputs(got, want, f"depth {depth} step {k}")
        expired += int(np.asarray(want["dropped"]).sum() > 0)
        if k % 97 == 0 or k == steps - 1:
            assert np.array_equal(gpu.buffers()[:, :, :depth], orc.buffers()[:, :, :depth]), k
    word = gpu.eng.v["path"].cpu().numpy()
    live = (word >> 2) & 0x3FFF
    buf = gpu.buffers()
    oldest = ((buf != 0) * (np.arange(buf.shape[2]) + 1)).max(2)
    assert (oldest <= live).all() and (live.max() >= depth - 5) and expired > depth // 2
    # nothing is ever written at or beyond a row's depth (the padding of the row stride stays zero; compute-sanitizer is
    # not available on the GPU pool, so the bound is checked through the data)
    dep = np.array([[depth - (s if i else 0) for s in range(5) for _ in range(5)] for i in range(4)])[ie]
    assert (live <= dep).all() and not buf[np.arange(buf.shape[2])[None, None, :] >= dep[:, :, None]].any()
