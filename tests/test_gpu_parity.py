"""GPU suite (-m gpu): the CUDA path, called through the C ABI, against
  * the golden traces recorded from the unmodified reference (tests/golden/),
  * the reference's own known-answer vector Buffer.main() (buffer.py:133-149),
  * the CPU oracle on the same seeded inputs (batched, native generators, resets),
  * size-independent properties at full batch size.
Tolerances: integers / masks / buffers / RB runs bit-exact; float64 surfaces rtol 1e-5
(BASELINE.json north_star), checked at 1e-12 where the oracle is in the loop.
"""
import numpy as np
import pytest

import golden_inputs
import trace_replay
from gpu_adapter import OUT_KEYS, CudaEnv, compare_outputs

pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True, params=["one_launch", "warp_per_env", "lane8", "thread_per_env"])
def per_env_kernels(request, monkeypatch):
    """Every test of this file runs with the four ways a step is launched: k_env_step (the whole step of an env in one
    warp and ONE launch; the default up to 2,368 envs per launch), and k_ue_env between the per-env stages as one warp
    per env (k_action / k_env_tail), eight lanes per env (k_action_8 / k_env_tail_8; the default for mid-size launches
    and the chunks of irs_step_host) or one thread per env (k_action_t / k_env_tail_t, the default from 49,152 envs per
    launch)."""
    if request.param != "one_launch":
        monkeypatch.setenv("IRS_FUSED_MAX_ENVS", "0")
        monkeypatch.setenv("IRS_PER_ENV", {"warp_per_env": "warp", "lane8": "lane8", "thread_per_env": "thread"}[request.param])

TRACES = ["cfg1_marr", "mapf", "ibsched_mimic", "ibsched_fixed", "seq_mimic", "seq_fixed", "sweep_scen", "twc", "colran",
          "random_eps"]


def _cuda_env_factory(bank):
    def make(meta):
        sort, fixed = trace_replay.agent_config(meta)
        return CudaEnv(bank, 1, channel_mode=0, traffic_mode=0, sort_slices=sort, fixed_intra=fixed,
                       surface=trace_replay.agent_surface(meta),
                       assoc_mode=0 if meta["assoc"] == "mult_slice" else 1, max_steps=int(meta["max_steps"]),
                       max_episodes=int(meta["max_episodes"]), initial_episode=int(meta["initial_episode"]),
                       random_episodes=int(meta.get("random_episodes", 0)))
    return make


@pytest.mark.parametrize("name", TRACES)
def test_cuda_matches_reference_trace(name, bank):
    """Every step of every golden trace: RB runs, packet counts, buffers, masks bit-exact;
    occupancy/latency/obs/reward within 1e-5 relative (observed: 0)."""
    trace_replay.replay(name, _cuda_env_factory(bank), tight=False)


def test_cuda_trace_is_bit_tight(bank):
    """Same replay at rtol 1e-12: the float64 evaluation order matches NumPy's."""
    trace_replay.replay("ibsched_fixed", _cuda_env_factory(bank), tight=True)


def test_buffer_main_known_answer():
    """buffer.py:133-149: Buffer(max_pkts=20, max_age=5), arrivals below, 2 packets sent per step."""
    from intent_radio_sched_multi_slice_b200 import scenario as scn
    from test_oracle_golden import BUFFER_MAIN

    req = {f"slice_{s}": {} for s in range(5)}
    req["slice_0"] = {"name": "kat", "priority": 0, "parameters": {},
                      "ues": {"buffer_size": 20, "buffer_latency": 5, "message_size": 1, "traffic": 1,
                              "min_number_ues": 1, "max_number_ues": 1}}
    sua = np.zeros((5, 25))
    sua[0, 3] = 1
    bank = scn.ScenarioBank()
    bank.add(req, sua)
    env = CudaEnv(bank, 1, channel_mode=0, traffic_mode=0, sort_slices=1, fixed_intra=0, bandwidth=135.0,
                  depth_cap=101, scen_mod=1)
    se = np.zeros((1, 25, 135))
    se[0, 3, 0] = 2.0  # 1 Hz per RB x SE 2 / 1-bit packets = 2 packets per step
    env.reset(initial_episode=0, se=se)
    arrivals = [0, 10, 0, 5, 2, 2, 10, 15, 20, 10]
    a0 = np.ones((1, 5))
    for k, arr in enumerate(arrivals):
        tr = np.zeros((1, 25))
        tr[0, 3] = arr
        out = env.step(a0, np.zeros((1, 5), np.int32), se, tr)
        after_rx, after_tx, drop = BUFFER_MAIN[k]
        assert env.buffers()[0, 3, :6].tolist() == after_tx, k
        assert out["dropped"][0, 3] == drop
        assert out["sent"][0, 3] == sum(after_rx) - sum(after_tx)
        assert out["rb_count"][0, 3] == 135
    assert out["lat"][0, 3] == 2.3333333333333335


@pytest.mark.parametrize("cfg", [
    dict(policy=3, fixed_intra=-1, sort_slices=1, channel_mode=2, assoc_mode=0),
    dict(policy=1, fixed_intra=0, sort_slices=0, channel_mode=2, assoc_mode=0),
    dict(policy=2, fixed_intra=1, sort_slices=0, channel_mode=2, assoc_mode=0),
    dict(policy=3, fixed_intra=-1, sort_slices=1, channel_mode=1, assoc_mode=1),
    dict(policy=3, fixed_intra=0, sort_slices=0, channel_mode=2, assoc_mode=0, surface=1),  # SchedTWC surface
    dict(policy=3, fixed_intra=0, sort_slices=0, channel_mode=2, assoc_mode=0, surface=2),  # SchedColORAN surface
])
def test_batch_native_matches_oracle(bank, cfg):
    """256 envs, native counter-based generators, on-device policy, auto-reset across several
    short episodes (scenario changes included): every output of every step equals the CPU
    oracle's; buffers and episode metrics at the end."""
    from oracle import oracle

    E, steps = 256, 130
    kw = dict(traffic_mode=1, max_steps=37, max_episodes=10000, initial_episode=7, auto_reset=1,
              env_id_base=1000, seed=1234, scen_div=2, **cfg)
    orc = oracle.OracleEnv(bank.array(), bank.poisson_cdf(), E, depth_cap=404, **kw)
    gpu = CudaEnv(bank, E, depth_cap=401, **kw)
    ie = (np.arange(E) % 5 + 3).astype(np.int32)
    # the oracle resets env by env; start every env at its own episode
    for e in range(E):
        m = np.zeros(E, np.uint8)
        m[e] = 1
        want = orc.reset(mask=m, initial_episode=int(ie[e]))
    got = gpu.reset(initial_episodes=ie, initial_episode=0)
    compare_outputs(got, want, "reset", keys=["obs0", "obs_i", "mask0", "mask_i", "sorted_slices"])
    for k in range(steps):
        want = orc.step()
        got = gpu.step()
        compare_outputs(got, want, f"step {k}")
        assert np.array_equal(got["a0_used"], want["a0_used"]) and np.array_equal(got["intra_used"], want["intra_used"])
        assert np.array_equal(got["traffic_used"], want["traffic_used"])
    assert np.array_equal(gpu.buffers()[:, :, :401], orc.buffers()[:, :, :401])
    st_g, st_o = gpu.state(), orc.state()
    for key in st_o:
        assert np.array_equal(st_g[key], st_o[key]), key
    per_g, tot_g = gpu.episode_metrics()
    per_o, tot_o = orc.episode_metrics()
    np.testing.assert_allclose(per_g, per_o, rtol=1e-12, atol=1e-300)
    np.testing.assert_allclose(tot_g, tot_o, rtol=1e-12, atol=1e-300)
    assert per_g[:, 15].min() >= 3  # every env went through >= 3 resets


def test_config2_1024_envs_full_episode(bank):
    """BASELINE config 2: 1,024 batched mult_slice envs, env e on association episode e mod 200, actions pre-drawn per env
    with default_rng(1234 + e) (player_0 ~ U(-1,1)^5, player_1..5 ~ integers{0,1,2}), one whole 1000-step episode plus the
    rollover into the next: bit-exact buffers / RB runs / packet counts / masks and <= 1e-12 observations and rewards
    against the oracle at every step (native generators: the reference's own RNG stream cannot be batched)."""
    from oracle import oracle

    E, steps = 1024, 1012
    kw = dict(channel_mode=2, traffic_mode=1, policy=0, fixed_intra=-1, sort_slices=1, assoc_mode=0, scen_mod=200,
              max_steps=1000, max_episodes=10000, initial_episode=0, auto_reset=1, env_id_base=0, seed=10)
    orc = oracle.OracleEnv(bank.array(), bank.poisson_cdf(), E, depth_cap=404, **kw)
    gpu = CudaEnv(bank, E, depth_cap=401, **kw)
    ie = (np.arange(E) % 200).astype(np.int32)
    for e0 in range(200):
        m = ((np.arange(E) % 200) == e0).astype(np.uint8)
        want = orc.reset(mask=m, initial_episode=e0)
    got = gpu.reset(initial_episodes=ie, initial_episode=0)
    compare_outputs(got, want, "reset", keys=["obs0", "obs_i", "mask0", "mask_i", "sorted_slices"])
    gens = [np.random.default_rng(1234 + e) for e in range(E)]
    a0_all = np.stack([g.uniform(-1, 1, (steps, 5)) for g in gens], 1)                       # [steps, E, 5]
    intra_all = np.stack([g.integers(0, 3, (steps, 5)) for g in gens], 1).astype(np.int32)
    dones = 0
    for k in range(steps):
        want, got = orc.step(a0_all[k], intra_all[k]), gpu.step(a0_all[k], intra_all[k])
        compare_outputs(got, want, f"config 2 step {k}")
        dones += int(np.asarray(want["done"]).sum())
        if k % 250 == 249 or k == steps - 1:
            assert np.array_equal(gpu.buffers()[:, :, :401], orc.buffers()[:, :, :401]), k
    assert dones == E
    st_g, st_o = gpu.state(), orc.state()
    for key in st_o:
        assert np.array_equal(st_g[key], st_o[key]), key
    per_g, tot_g = gpu.episode_metrics()
    per_o, tot_o = orc.episode_metrics()
    np.testing.assert_allclose(per_g, per_o, rtol=1e-12, atol=1e-300)
    np.testing.assert_allclose(tot_g, tot_o, rtol=1e-12, atol=1e-300)


@pytest.mark.parametrize("channel_mode", [1, 2])
def test_config4_seq_scenarios_across_resets(bank, channel_mode):
    """BASELINE config 4: MultSliceAssociationSeq (scenario = episode // 100) with the fixed_se and mimic_quadriga channels;
    envs start at episode 98 or 99 and run four episodes, so every env crosses a scenario boundary (new slice set, new
    per-UE depth / buffer size / packet size) with the observation window carried over the reset (quirk Q1); a third of the
    envs is reset once more in mid-episode, so resets are out of phase inside the batch."""
    from oracle import oracle

    E, T, steps = 384, 150, 640
    kw = dict(channel_mode=channel_mode, traffic_mode=1, policy=3, fixed_intra=-1, sort_slices=1, assoc_mode=1, scen_div=100,
              max_steps=T, max_episodes=10000, initial_episode=98, auto_reset=1, env_id_base=50, seed=4)
    orc = oracle.OracleEnv(bank.array(), bank.poisson_cdf(), E, depth_cap=404, **kw)
    gpu = CudaEnv(bank, E, depth_cap=401, **kw)
    ie = (98 + np.arange(E) % 2).astype(np.int32)
    for ep in (98, 99):
        want = orc.reset(mask=(ie == ep).astype(np.uint8), initial_episode=ep)
    got = gpu.reset(initial_episodes=ie, initial_episode=0)
    compare_outputs(got, want, "reset", keys=["obs0", "obs_i", "mask0", "mask_i", "sorted_slices"])
    third = (np.arange(E) % 3 == 0).astype(np.uint8)
    for k in range(steps):
        if k == 61:  # explicit reset of a third of the batch: those envs move on to their next episode now
            want, got = orc.reset(mask=third), gpu.reset(mask=third)
            compare_outputs(got, want, "masked reset", keys=["obs0", "obs_i", "mask0", "mask_i", "sorted_slices"])
        want, got = orc.step(), gpu.step()
        compare_outputs(got, want, f"config 4 step {k}")
    st_g, st_o = gpu.state(), orc.state()
    for key in st_o:
        assert np.array_equal(st_g[key], st_o[key]), key
    assert len(np.unique(st_o["step"])) >= 2 and st_o["scen"].min() >= 1  # out of phase; every env left scenario 0
    assert np.array_equal(gpu.buffers()[:, :, :401], orc.buffers()[:, :, :401])
    per_g, tot_g = gpu.episode_metrics()
    per_o, tot_o = orc.episode_metrics()
    np.testing.assert_allclose(per_g, per_o, rtol=1e-12, atol=1e-300)


def test_config3_results_do_not_depend_on_the_sharding(bank):
    """BASELINE config 3: 16,384 envs with the IBSched dict surface, native generators keyed by the GLOBAL env id.  The batch
    run whole and run as 8 shards of 2,048 envs (env_id_base = rank * 2048, as distributed.shard hands them out) gives
    bit-identical per-env results, identical integer episode counters and float sums equal to rounding; a 64-env sample of
    every shard is checked against the oracle."""
    from oracle import oracle

    E, G, steps = 16384, 8, 40
    kw = dict(channel_mode=2, traffic_mode=1, policy=3, fixed_intra=-1, sort_slices=1, assoc_mode=0, scen_mod=200,
              max_steps=25, max_episodes=10000, initial_episode=0, auto_reset=1, seed=10)
    whole = CudaEnv(bank, E, depth_cap=401, env_id_base=0, **kw)
    ie = (np.arange(E) % 200).astype(np.int32)
    whole.reset(initial_episodes=ie, initial_episode=0)
    for k in range(steps):
        out_w = whole.step()
    per_w, tot_w = whole.episode_metrics()
    buf_w = whole.buffers()
    tot_s = np.zeros_like(np.asarray(tot_w))
    n = E // G
    for r in range(G):
        sl = slice(r * n, (r + 1) * n)
        shard = CudaEnv(bank, n, depth_cap=401, env_id_base=r * n, **kw)
        shard.reset(initial_episodes=ie[sl], initial_episode=0)
        sample = np.arange(r * n, r * n + 64)
        orc = oracle.OracleEnv(bank.array(), bank.poisson_cdf(), 64, depth_cap=404, env_id_base=r * n, **kw)
        for e in range(64):
            m = np.zeros(64, np.uint8)
            m[e] = 1
            orc.reset(mask=m, initial_episode=int(ie[sample[e]]))
        for k in range(steps):
            out_s = shard.step()
            want = orc.step()
            compare_outputs({key: np.asarray(v)[:64] for key, v in out_s.items()}, want, f"shard {r} step {k}")
        for key, v in out_s.items():
            assert np.array_equal(np.asarray(v), np.asarray(out_w[key])[sl]), (r, key)
        assert np.array_equal(shard.buffers(), buf_w[sl])
        per_s, t = shard.episode_metrics()
        assert np.array_equal(per_s, per_w[sl])
        tot_s += np.asarray(t)
        del shard
    from intent_radio_sched_multi_slice_b200 import capi
    for i, name in enumerate(capi.METRIC_NAMES):
        np.testing.assert_allclose(tot_s[i], np.asarray(tot_w)[i], rtol=1e-12, err_msg=name)


def test_config5_full_size_properties():
    """BASELINE config 5 at its full size: 65,536 envs x 25 UEs x depth 1024 (every UE on the deep uav_app_case_1 profile,
    7 GB of state).  Size-independent properties over the whole batch - packet conservation, the per-UE packet count and
    the live extent against the rows themselves, sent <= throughput, 135 RBs per env - and the first 48 envs against the
    oracle at every step."""
    import torch
    from intent_radio_sched_multi_slice_b200 import scenario as scn
    from oracle import oracle

    D = 1024
    bank = scn.ScenarioBank()
    for i in range(8):
        req, sua = {}, np.zeros((5, 25))
        for s in range(5):
            r = scn.slice_req_from_type(5)
            r["ues"]["buffer_latency"] = D - 1
            r["parameters"]["par1"]["value"] = D // 2
            req[f"slice_{s}"] = r
            sua[s, 5 * s:5 * s + 5] = 1
        bank.add(req, sua)
    E, steps, n_orc = 65536, 24, 48
    kw = dict(channel_mode=2, traffic_mode=1, policy=3, fixed_intra=-1, sort_slices=1, assoc_mode=0, scen_mod=8,
              max_steps=1_000_000, max_episodes=10000, initial_episode=0, auto_reset=1, seed=10, env_id_base=0)
    gpu = CudaEnv(bank, E, depth_cap=D, **kw)
    orc = oracle.OracleEnv(bank.array(), bank.poisson_cdf(), n_orc, depth_cap=D, **kw)
    orc.reset(initial_episode=0)
    gpu.eng.reset(initial_episode=0)
    inc = torch.zeros(E, 25, dtype=torch.int64, device="cuda")
    out = torch.zeros_like(inc)
    for k in range(steps):
        gpu.eng.step()
        raw = gpu.raw
        want = orc.step()
        surf = gpu.eng.surface()
        got = {key: (surf[key] if key in surf else raw[key])[:n_orc].cpu().numpy() for key in OUT_KEYS}
        compare_outputs(got, want, f"config 5 step {k}")
        inc += raw["pkt_incoming"].long()
        out += raw["sent"].long() + raw["dropped"].long()
        assert bool((raw["sent"].double() <= raw["pkt_throughputs"]).all())
        assert bool((raw["rb_count"].long().sum(1) == 135).all())
    assert gpu.eng.check_errors() is None
    buf = gpu.eng.buf
    in_buf = buf.sum(2, dtype=torch.int64)
    assert bool((inc == out + in_buf).all()) and bool((in_buf == gpu.eng.v["tot"].long()).all())
    word = gpu.eng.v["path"].long()
    live = (word >> 2) & 0x3FFF
    idx = torch.arange(buf.shape[2], device="cuda", dtype=torch.int32) + 1
    oldest = torch.zeros(E, 25, dtype=torch.int32, device="cuda")
    for c0 in range(0, E, 8192):  # chunked: the mask of the whole state would be another 7 GB
        oldest[c0:c0 + 8192] = ((buf[c0:c0 + 8192] != 0).int() * idx).amax(2)
    assert bool((oldest.long() <= live).all())
    assert np.array_equal(buf[:n_orc].cpu().numpy()[:, :, :D], orc.buffers()[:, :, :D])


def test_generated_scenarios_match_oracle():
    """assoc_mode 2: a new scenario per (env, episode) generated on the device at every reset
    (associations/mult_slice.py:359-416 distribution): the rows and every step output equal the oracle's."""
    from intent_radio_sched_multi_slice_b200 import scenario as scn
    from oracle import oracle

    tb = scn.ScenarioBank.templates()
    E, steps = 200, 60
    kw = dict(channel_mode=2, traffic_mode=1, policy=3, fixed_intra=-1, sort_slices=1, assoc_mode=2, max_steps=17,
              auto_reset=1, env_id_base=40, seed=77, initial_episode=3)
    orc = oracle.OracleEnv(tb.array(), tb.poisson_cdf(), E, depth_cap=404, **kw)
    gpu = CudaEnv(tb, E, depth_cap=401, **kw)
    want, got = orc.reset(initial_episode=3), gpu.reset(initial_episode=3)
    compare_outputs(got, want, "reset", keys=["obs0", "obs_i", "mask0", "mask_i", "sorted_slices"])
    assert np.array_equal(gpu.eng.scenario_rows(), orc.scenario_rows())
    for k in range(steps):
        want, got = orc.step(), gpu.step()
        compare_outputs(got, want, f"step {k}")
        if k % 17 == 16:
            assert np.array_equal(gpu.eng.scenario_rows(), orc.scenario_rows()), k
    assert np.array_equal(gpu.buffers()[:, :, :401], orc.buffers()[:, :, :401])
    assert np.array_equal(gpu.state()["scen"], orc.state()["scen"])


def test_config2_batch_fixture_from_the_reference(bank):
    """BASELINE config 2 against the REFERENCE itself: 64 envs x 1000 steps recorded from the unmodified reference
    (env e: episode e mod 200, seed 10 + e, actions default_rng(1234 + e); oracle/refharness/gen_config2.py) replayed
    as ONE CUDA batch with the recorded SE / traffic frames injected (channel_mode 0, traffic_mode 0): RB runs, packet
    counts, masks, slice order bit-exact at every step, rewards at every step, observations every 8th step, buffers
    at the checkpoints, the episode rollover included."""
    d = golden_inputs.load_config2()
    envs = list(range(64))
    gpu = CudaEnv(bank, 64, channel_mode=0, traffic_mode=0, sort_slices=1, fixed_intra=-1)

    class Env:
        def reset(self, initial_episodes, se):
            return gpu.reset(initial_episodes=initial_episodes, initial_episode=0, se=se)

        def step(self, a0, intra, se, tr):
            return gpu.step(a0, intra, se, tr)

        def buffers(self):
            return gpu.buffers()

    assert golden_inputs.replay_config2(Env(), d, envs) == 4
    assert bool(gpu.eng.v["out_done"].all())


@pytest.mark.parametrize("name", ["mapf", "marr", "twc", "colran", "ibsched_fixed"])
def test_batch_fixtures_from_the_reference(name, bank):
    """16 envs x 300 steps (three 100-step episodes per env) of MAPF / MARR / SchedTWC / SchedColORAN / IBSched over
    fixed_se recorded from the unmodified reference (oracle/refharness/gen_batch_fixtures.py), each replayed as ONE CUDA
    batch with the recorded frames injected: integers bit-exact at every step, rewards every step, the agent's
    observation every 4th step, buffers at the checkpoints, every rollover's reset observation."""
    meta, d = golden_inputs.load_batch(name)
    m = {"agent": str(meta["agent"])}
    sort, fixed = trace_replay.agent_config(m)
    gpu = CudaEnv(bank, int(meta["n_envs"]), channel_mode=0, traffic_mode=0, sort_slices=sort, fixed_intra=fixed,
                  surface=trace_replay.agent_surface(m), max_steps=int(meta["max_steps"]))

    class Env:
        def reset(self, initial_episodes, se):
            if initial_episodes is None:
                return gpu.reset(initial_episode=-1, se=se)
            return gpu.reset(initial_episodes=initial_episodes, initial_episode=0, se=se)

        def step(self, a0, intra, se, tr):
            return gpu.step(a0, intra, se, tr)

        def buffers(self):
            return gpu.buffers()

    assert golden_inputs.replay_batch(Env(), meta, d) == 3


def test_random_episodes_native_matches_oracle(bank):
    """enable_random_episodes with the built-in (Feistel) permutation: CUDA and oracle play the same shuffled cycles,
    per env, across wrap-arounds and a mid-run reset that draws a new permutation (comm_env.py:219-226, 434-444)."""
    E = 40
    kw = dict(channel_mode=2, traffic_mode=1, policy=3, fixed_intra=-1, sort_slices=1, max_steps=3, max_episodes=17,
              initial_episode=6, auto_reset=1, random_episodes=1, seed=21, env_id_base=100)
    from oracle import oracle
    orc = oracle.OracleEnv(bank.array(), bank.poisson_cdf(), E, depth_cap=404, **kw)
    gpu = CudaEnv(bank, E, depth_cap=401, **kw)
    assert np.array_equal(gpu.state()["episode"], orc.state()["episode"])  # drawn at construction (comm_env.py:225)
    compare_outputs(gpu.reset(), orc.reset(), "reset", keys=["obs0", "obs_i", "mask0", "mask_i", "sorted_slices"])
    eps = []
    for k in range(3 * 25):
        compare_outputs(gpu.step(), orc.step(), f"step {k}")
        assert np.array_equal(gpu.state()["episode"], orc.state()["episode"]), k
        if k % 3 == 2:
            eps.append(gpu.state()["episode"].copy())
    eps = np.array(eps)
    for e in range(E):  # 11 episodes per cycle, every one exactly once, same order in the second cycle
        assert sorted(eps[:11, e]) == list(range(6, 17)) and np.array_equal(eps[:11, e], eps[11:22, e])
    m = (np.arange(E) % 2).astype(np.uint8)
    compare_outputs(gpu.reset(mask=m, initial_episode=12), orc.reset(mask=m, initial_episode=12), "re-permute",
                    keys=["obs0", "obs_i", "mask0", "mask_i", "sorted_slices"])
    for k in range(3 * 7):
        compare_outputs(gpu.step(), orc.step(), f"after re-permute step {k}")
        assert np.array_equal(gpu.state()["episode"], orc.state()["episode"]), k
    assert (gpu.state()["episode"][m == 1] >= 12).all()
    assert np.array_equal(gpu.buffers()[:, :, :401], orc.buffers()[:, :, :401])


def test_episode_numbers_played_again_draw_fresh_noise(bank):
    """The reference keeps one np_random stream running across episodes (comm_env.py:631-679), so an episode NUMBER
    that comes round again sees new channel and traffic noise; the native generators' key carries the env's reset
    count for the same effect."""
    gpu = CudaEnv(bank, 4, channel_mode=2, traffic_mode=1, policy=1, fixed_intra=0, sort_slices=0, seed=3)
    gpu.eng.enable_raw_outputs()
    gpu.raw = gpu.eng._raw
    gpu.reset(initial_episode=5)
    gpu.step()
    a = gpu.raw["se_used"].cpu().numpy().copy()
    gpu.reset(initial_episode=5)
    gpu.step()
    b = gpu.raw["se_used"].cpu().numpy()
    active = a.sum(axis=2) > 0
    assert active.any() and not np.array_equal(a, b) and (np.abs(a - b)[active].mean() > 0.5)


def test_native_se_statistics(bank):
    """Native mimic_quadriga / Poisson draws follow the reference's distributions
    (channels/mimic_quadriga.py:37-56: |N(mu_u, 1.5)|, mu_u = |N(10, 7.5)|;
    traffics/mult_slice.py:28-31: Poisson(lambda) Mbit)."""
    E = 512
    gpu = CudaEnv(bank, E, channel_mode=2, traffic_mode=1, policy=1, fixed_intra=0, sort_slices=0)
    gpu.reset(initial_episode=0)
    gpu.step()
    se = gpu.raw["se_used"].cpu().numpy()  # [E,25,135]
    scen = bank.array()[0]
    active = scen["ue_slice"][:25] >= 0
    se_a = se[:, active, :]
    mu = se_a.mean(axis=2)  # per (env, ue): close to the UE's mean (|N| folding is negligible for mu >> 1.5)
    resid = (se_a - mu[..., None])[mu > 6]
    assert abs(resid.std() - 1.5) < 0.02
    # Kolmogorov-Smirnov: deviations from the row mean are exactly N(0, 1.5^2 (1 - 1/135)) when the draws are normal
    # and nothing folds at zero (mu > 8 against |z| <= 4.32 sigma); critical value at alpha = 0.001 and n = 2e5: 4.4e-3
    from scipy import stats
    z = (se_a - mu[..., None])[mu > 8].ravel()[:200_000] / (1.5 * np.sqrt(1 - 1 / 135))
    assert z.size == 200_000 and stats.kstest(z, "norm").statistic < 4.4e-3
    # the same statistic from NumPy draws of the reference's distributions
    rng = np.random.default_rng(0)
    mu_ref = np.abs(rng.normal(10, 7.5, 20000))
    rows_ref = np.abs(rng.normal(mu_ref[:, None], 1.5, (20000, 135))).mean(axis=1)
    assert abs(mu.mean() - rows_ref.mean()) < 0.3 and abs(mu.std() - rows_ref.std()) < 0.3
    assert (se[:, ~active, :] == 0).all()
    tr = gpu.raw["traffic_used"].cpu().numpy() / 1e6
    for s in range(5):
        sl = scen["sl"][s]
        if sl["n_ues"] > 0:
            x = tr[:, scen["ue_slice"][:25] == s].ravel()
            lam = sl["traffic"]
            assert abs(x.mean() - lam) < 4 * np.sqrt(lam / x.size) + 1e-9
            assert abs(x.var() - lam) < 0.15 * lam


def test_full_size_properties(bank):
    """BASELINE config 2/5 sizes: conservation of packets and the reference's runtime asserts
    (ues.py:218-234, common.py:456-459, ib_sched.py:345-347) as properties over 16,384 envs."""
    import torch

    E = 16384
    gpu = CudaEnv(bank, E, channel_mode=2, traffic_mode=1, policy=3, fixed_intra=-1, sort_slices=1,
                  auto_reset=0, seed=99)
    ie = (np.arange(E) % 200).astype(np.int32)
    gpu.eng.reset(initial_episodes=ie, initial_episode=0)
    inc_total = torch.zeros(E, 25, dtype=torch.int64, device="cuda")
    sent_total = torch.zeros_like(inc_total)
    drop_total = torch.zeros_like(inc_total)
    for k in range(60):
        gpu.eng.step()
        raw = gpu.raw
        inc_total += raw["pkt_incoming"].long()
        sent_total += raw["sent"].long()
        drop_total += raw["dropped"].long()
        assert bool((raw["sent"].double() <= raw["pkt_throughputs"]).all())        # ues.py:225-229
        assert bool(((raw["occ"] >= 0) & (raw["occ"] <= 1)).all())                 # ues.py:218-224
        assert bool((raw["rb_count"].long().sum(1) == 135).all())                  # ib_sched.py:345-347
        end = raw["rb_start"].long() + raw["rb_count"].long()
        assert bool((end <= 135).all())
    assert gpu.eng.check_errors() is None
    in_buf = gpu.eng.buf.long().sum(2)
    assert bool((inc_total == sent_total + drop_total + in_buf).all())             # packet conservation
    assert bool((in_buf == gpu.eng.v["tot"].long()).all())
    # live extent of every row (irs_common.cuh ROW_LIVE): an upper bound of the oldest non-empty age + 1 - the
    # row code never looks past it
    word = gpu.eng.v["path"].long()
    live = (word >> 2) & 0x3FFF
    idx = torch.arange(gpu.eng.buf.shape[2], device="cuda")
    oldest = ((gpu.eng.buf != 0).long() * (idx + 1)).max(2).values
    assert bool((oldest <= live).all())
    # no RB is given to two UEs (comm_env.py:593-596): runs of one env do not overlap
    rs, rc = gpu.raw["rb_start"].long(), gpu.raw["rb_count"].long()
    cover = torch.zeros(E, 136, dtype=torch.int64, device="cuda")
    cover.scatter_add_(1, rs, (rc > 0).long())
    cover.scatter_add_(1, rs + rc, -(rc > 0).long())
    assert bool((cover.cumsum(1).max(1).values <= 1).all())
