"""GPU suite: the reference-shaped env surface (intent_radio_sched_multi_slice_b200.env) replays
golden traces through the PUBLIC API - plugin classes passed by type (host tensor hooks feeding
the recorded SE / traffic frames), dict actions in, dict observations / rewards / terminated out -
and must reproduce the reference's returns, dict key order included (ib_sched.py:91,186-200;
marl_comm_env.py:51-66)."""
import numpy as np
import pytest

import golden_inputs

pytestmark = pytest.mark.gpu


def _trace_plugins(rse, sse, stf):
    """Channel / Traffic classes (reference plugin interface) that replay recorded frames."""
    state = {"reset": 0, "step": 0}

    class TraceChannel:
        def __init__(self, max_number_ues, max_number_basestations, num_available_rbs, rng, root_path, scenario_name):
            self.fresh = True

        def step(self, step_number, episode_number, mobilities, sched_decision=None):
            if self.fresh:  # first call after construction = the reset frame (comm_env.py:449-451)
                self.fresh = False
                out = rse[state["reset"]]
                state["reset"] += 1
            else:
                out = sse[state["step"]]
            return out[None]

    class TraceTraffic:
        def __init__(self, max_number_ues, rng, root_path):
            pass

        def step(self, slice_ue_assoc, slice_req, step_number, episode_number):
            return stf[min(state["step"], len(stf) - 1)]

    return TraceChannel, TraceTraffic, state


@pytest.mark.parametrize("name,agent,n", [("ibsched_mimic", "IBSched", 150), ("seq_fixed", "IBSched", 100),
                                          ("cfg1_marr", "MARR", 120), ("mapf", "MAPF", 120),
                                          ("quadriga", "IBSched", 100), ("quadriga_seq", "IBSched", 100),
                                          ("twc", "SchedTWC", 220), ("colran", "SchedColORAN", 220),
                                          ("random_eps", "IBSched", 375)])
def test_env_api_replays_reference_trace(name, agent, n, bank):
    from intent_radio_sched_multi_slice_b200 import agents, plugins
    from intent_radio_sched_multi_slice_b200.env import MARLCommEnv

    meta, d = golden_inputs.load_trace(name)
    rse, sse, stf = golden_inputs.regenerate(meta, d)
    Ch, Tr, state = _trace_plugins(rse, sse, stf)
    if meta["channel"].startswith("quadriga"):  # the built-in device channel: SE bank in HBM (channel_mode 3)
        Ch = plugins.QuadrigaChannel if meta["channel"] == "quadriga" else plugins.QuadrigaChannelSeq
        first, power = golden_inputs.quadriga_episode_powers(meta, d)  # slot = episode - first (first = 3 / 99 here)
        Ch.register_power(power, first_episode=first)
    Assoc = plugins.MultSliceAssociation if meta["assoc"] == "mult_slice" else plugins.MultSliceAssociationSeq
    env = MARLCommEnv(Ch, Tr, plugins.SimpleMobility, Assoc, "mult_slice", agent.lower(), int(meta["seed"]),
                      root_path="/nonexistent", initial_episode_number=int(meta["initial_episode"]),
                      max_episode_number=int(meta["max_episodes"]), save_hist=False, debug=True, scenario_bank=bank,
                      enable_random_episodes=bool(int(meta.get("random_episodes", 0))))
    env.max_number_steps = int(meta["max_steps"])
    ag = getattr(agents, agent)(env, 25, 5, 1, np.array([135]))
    env.set_agent_functions(ag.obs_space_format, ag.action_format, ag.calculate_reward, ag.get_obs_space(),
                            ag.get_action_space())
    random_eps, n_perm = bool(int(meta.get("random_episodes", 0))), 0
    if random_eps:  # the permutation the reference drew from its np_random (comm_env.py:435-442)
        env._engine().set_episode_order(d["reset_order_0"][None])
        n_perm = 1
    obs, info = env.reset(seed=int(meta["seed"]), options={"initial_episode": int(meta["initial_episode"])})
    assert info == {}

    def check_obs(obs, o0, m0, oi, mi, order, tag):
        if agent != "IBSched":
            assert isinstance(obs, np.ndarray) and obs.shape == (50,)
            np.testing.assert_allclose(obs, o0, rtol=1e-5, atol=0, err_msg=tag)
            return
        assert list(obs.keys()) == [f"player_{int(k)}" for k in order], tag  # insertion order = sorted slices
        np.testing.assert_allclose(obs["player_0"]["observations"], o0, rtol=1e-5, atol=0, err_msg=tag)
        assert obs["player_0"]["action_mask"].dtype == np.int8 and np.array_equal(obs["player_0"]["action_mask"], m0)
        for s in range(5):
            np.testing.assert_allclose(obs[f"player_{s + 1}"]["observations"], oi[s], rtol=1e-5, atol=0, err_msg=tag)
            assert np.array_equal(obs[f"player_{s + 1}"]["action_mask"], mi[s])

    if "agent_obs" in d:  # SchedTWC / SchedColORAN return their own 50-vector
        d["obs0"], d["reset_obs0"] = d["agent_obs"], d["reset_agent_obs"]
    check_obs(obs, d["reset_obs0"][0], d["reset_mask0"][0], d["reset_obs_i"][0], d["reset_mask_i"][0],
              d["reset_order"][0], "reset")
    ri = 0
    for k in range(n):
        state["step"] = k
        if agent == "IBSched":
            action = {"player_0": d["a0"][k]}
            action.update({f"player_{i + 1}": int(d["intra"][k][i]) for i in range(5)})
        else:
            action = d["a0"][k] if agent.startswith("Sched") else ag.step(obs)  # DEVICE_POLICY: MARR / MAPF on the device
        obs, reward, term, trunc, info = env.step(action)
        check_obs(obs, d["obs0"][k], d["mask0"][k], d["obs_i"][k], d["mask_i"][k], d["order"][k], f"step {k}")
        done = bool(d["done"][k])
        if agent == "IBSched":
            assert list(reward.keys()) == list(obs.keys())  # common.py:396-439 iterates over the formatted observation
            np.testing.assert_allclose([reward[f"player_{i}"] for i in range(6)], d["reward"][k], rtol=1e-5, atol=1e-300)
            assert term["__all__"] == done and trunc["__all__"] == done and term["player_3"] == done
        else:
            np.testing.assert_allclose(reward, d["reward"][k][0], rtol=1e-5, atol=1e-300)
            assert term is done and trunc is done
        assert info == {}
        if done:
            ri += 1
            if random_eps and int(d["reset_initial"][ri]) != -1:
                env.eng.set_episode_order(d[f"reset_order_{n_perm}"][None])
                n_perm += 1
                obs, _ = env.reset(options={"initial_episode": int(d["reset_initial"][ri])})
            else:
                obs, _ = env.reset()
            if random_eps:
                assert env.episode_number == int(d["reset_episode"][ri]), ri
            check_obs(obs, d["reset_obs0"][ri], d["reset_mask0"][ri], d["reset_obs_i"][ri], d["reset_mask_i"][ri],
                      d["reset_order"][ri], f"reset {ri}")
    env.close()


def test_env_api_errors_like_the_reference(bank):
    from intent_radio_sched_multi_slice_b200 import agents, plugins
    from intent_radio_sched_multi_slice_b200.env import MARLCommEnv

    env = MARLCommEnv(plugins.FixedSE, plugins.MultSliceTraffic, plugins.SimpleMobility, plugins.MultSliceAssociation,
                      "mult_slice", "ib_sched", 10, save_hist=False, scenario_bank=bank, root_path="/nonexistent")
    env.reset(seed=10, options={"initial_episode": 0})
    bad = {"player_0": np.zeros(5)}
    bad.update({f"player_{i}": 7 for i in range(1, 6)})
    with pytest.raises(ValueError, match="Invalid intra-slice scheduling action"):  # ib_sched.py:341-344
        env.step(bad)
    env.close()


def test_trace_channel_episode_outside_the_bank_raises(bank):
    """An env that reaches an episode whose channel file is not in the SE bank raises (debug=True) instead of
    replaying another episode's file (the reference opens a new file per episode, channels/quadriga.py:46-55)."""
    from intent_radio_sched_multi_slice_b200 import agents, plugins
    from intent_radio_sched_multi_slice_b200.env import MARLCommEnv

    meta, d = golden_inputs.load_trace("quadriga")
    first, power = golden_inputs.quadriga_episode_powers(meta, d)
    plugins.QuadrigaChannel.register_power(power[:2], first_episode=first)
    env = MARLCommEnv(plugins.QuadrigaChannel, plugins.MultSliceTraffic, plugins.SimpleMobility,
                      plugins.MultSliceAssociation, "mult_slice", "ib_sched", 10, save_hist=False, scenario_bank=bank,
                      root_path="/nonexistent", initial_episode_number=first, debug=True)
    env.max_number_steps = 3
    env.reset(options={"initial_episode": first})
    act = {"player_0": np.zeros(5)}
    act.update({f"player_{i}": 0 for i in range(1, 6)})
    for ep in range(2):
        for k in range(3):
            env.step(act)
        if ep == 0:
            env.reset()
    with pytest.raises(Exception, match="no channel file"):
        env.reset()
    env.close()


def test_batched_env_matches_single_envs(bank):
    """num_envs=8 with the built-in device plugins: env e of the batch equals a single env created
    with env_id_base=e (results depend on the global env id only, not on the batch)."""
    import torch

    from intent_radio_sched_multi_slice_b200 import plugins
    from intent_radio_sched_multi_slice_b200.env import BatchedMARLCommEnv

    kw = dict(save_hist=False, scenario_bank=bank, root_path="/nonexistent", debug=True)
    args = (plugins.MimicQuadriga, plugins.MultSliceTraffic, plugins.SimpleMobility, plugins.MultSliceAssociation,
            "mult_slice", "ib_sched", 33)
    big = BatchedMARLCommEnv(*args, num_envs=8, auto_reset=False, **kw)
    ie = torch.arange(8, dtype=torch.int32) * 7
    ob, _ = big.reset(initial_episodes=ie)
    rng = np.random.default_rng(5)
    acts = [(rng.uniform(-1, 1, (8, 5)), rng.integers(0, 3, (8, 5))) for _ in range(25)]
    outs = []
    for a0, intra in acts:
        ob, rew, term, trunc, _ = big.step((a0, intra))
        outs.append((ob["player_0"]["observations"].cpu().numpy().copy(), rew.cpu().numpy().copy()))
    for e in (0, 3, 7):
        one = BatchedMARLCommEnv(*args, num_envs=1, env_id_base=e, **kw)
        one.reset(options={"initial_episode": int(ie[e])})
        for k, (a0, intra) in enumerate(acts):
            action = {"player_0": a0[e]}
            action.update({f"player_{i + 1}": int(intra[e, i]) for i in range(5)})
            ob1, rew1, *_ = one.step(action)
            assert np.array_equal(ob1["player_0"]["observations"], outs[k][0][e])
            assert np.array_equal([rew1[f"player_{i}"] for i in range(6)], outs[k][1][e])
        one.close()
    big.close()


def test_history_file_matches_reference_trace(bank, tmp_path):
    """save_hist=True: the episode file hist/<simu>/<agent>/ep_N.npz (metrics.py:73-100) written from
    the on-device recorder holds, for every step of the episode, the values the unmodified reference
    produced (golden trace seq_fixed, 40-step episodes) under the reference's 16 keys and shapes."""
    from intent_radio_sched_multi_slice_b200 import agents, history, plugins
    from intent_radio_sched_multi_slice_b200.env import MARLCommEnv

    meta, d = golden_inputs.load_trace("seq_fixed")
    rse, sse, stf = golden_inputs.regenerate(meta, d)
    Ch, Tr, state = _trace_plugins(rse, sse, stf)
    env = MARLCommEnv(Ch, Tr, plugins.SimpleMobility, plugins.MultSliceAssociationSeq, "mult_slice", "ib_sched",
                      int(meta["seed"]), root_path="/nonexistent", initial_episode_number=int(meta["initial_episode"]),
                      max_episode_number=int(meta["max_episodes"]), save_hist=True, debug=True, scenario_bank=bank,
                      hist_root_path=str(tmp_path), hist_with_se=True, simu_name="golden")
    env.max_number_steps = T = int(meta["max_steps"])
    ag = agents.IBSched(env, 25, 5, 1, np.array([135]))
    env.set_agent_functions(ag.obs_space_format, ag.action_format, ag.calculate_reward, ag.get_obs_space(),
                            ag.get_action_space())
    env.reset(seed=int(meta["seed"]), options={"initial_episode": int(meta["initial_episode"])})
    n_eps = 2
    for k in range(n_eps * T):
        state["step"] = k
        action = {"player_0": d["a0"][k]}
        action.update({f"player_{i + 1}": int(d["intra"][k][i]) for i in range(5)})
        *_, term, _, _ = env.step(action)
        if term["__all__"]:
            env.reset()
    env.close()
    for e in range(n_eps):
        ep = int(meta["initial_episode"]) + e
        h = history.read(str(tmp_path), "golden", "ib_sched", ep)
        sl = slice(e * T, (e + 1) * T)
        assert set(history.KEYS) - {"obs"} == set(h.keys())
        assert h["pkt_effective_thr"].dtype == np.int64 and h["pkt_effective_thr"].shape == (T, 25)
        for key, gold in (("pkt_incoming", "pkt_incoming"), ("pkt_throughputs", "pkt_throughputs"),
                          ("pkt_effective_thr", "pkt_effective_thr"), ("dropped_pkts", "dropped_pkts")):
            assert np.array_equal(h[key], d[gold][sl]), key
        np.testing.assert_allclose(h["buffer_occupancies"], d["buffer_occupancies"][sl], rtol=1e-12)
        np.testing.assert_allclose(h["buffer_latencies"], d["buffer_latencies"][sl], rtol=1e-12)
        assert h["sched_decision"].shape == (T, 1, 25, 135) and h["spectral_efficiencies"].shape == (T, 1, 25, 135)
        assert np.array_equal(h["sched_decision"][:, 0].sum(-1), d["rb_count"][sl])
        first = np.argmax(h["sched_decision"][:, 0], -1)
        assert np.array_equal(np.where(d["rb_count"][sl] > 0, first, 0), d["rb_start"][sl])
        assert np.array_equal(h["spectral_efficiencies"][:, 0], sse[sl])
        assert h["slice_ue_assoc"].shape == (T, 5, 25) and h["basestation_slice_assoc"].shape == (T, 1, 5)
        assert h["mobility"].shape == (T, 25, 2) and h["basestation_ue_assoc"].shape == (T, 1, 25)
        scen = golden_inputs.scen_of_episode(meta, ep)
        st, us = golden_inputs.load_scenarios()
        for s in range(5):
            assert np.array_equal(np.flatnonzero(h["slice_ue_assoc"][0, s]), np.flatnonzero(us[scen] == s))
            assert (h["slice_req"][0][f"slice_{s}"] != {}) == (st[scen, s] >= 0) == bool(h["basestation_slice_assoc"][0, 0, s])
        for k in (0, T // 2, T - 1):
            np.testing.assert_allclose([h["reward"][k][f"player_{i}"] for i in range(6)], d["reward"][sl][k], rtol=1e-12)
            assert np.array_equal(h["agent_action"][k]["player_0"], d["a0"][sl][k])
            assert [h["agent_action"][k][f"player_{i + 1}"] for i in range(5)] == d["intra"][sl][k].tolist()


@pytest.mark.parametrize("agent,fixture", [("IBSched", "ib_sched"), ("MARR", "marr")])
def test_history_file_has_the_structure_of_the_reference_file(agent, fixture, bank, tmp_path):
    """The file history.py writes against the file the UNMODIFIED reference wrote with save_hist=True
    (tests/golden/hist_ref_*_ep_4.npz, oracle/refharness/gen_hist_fixture.py): same keys in the same order, same
    dtypes and shapes, same nesting of the pickled dicts (reward / obs / agent_action / slice_req with callable
    operators) - compared through NumPy, not through this package's own reader."""
    import os

    from intent_radio_sched_multi_slice_b200 import agents, plugins
    from intent_radio_sched_multi_slice_b200.env import MARLCommEnv
    from test_reference_readers_cpu import GOLDEN, history_structure

    env = MARLCommEnv(plugins.MimicQuadriga, plugins.MultSliceTraffic, plugins.SimpleMobility,
                      plugins.MultSliceAssociation, "mult_slice", fixture, 31, root_path="/nonexistent",
                      initial_episode_number=4, save_hist=True, debug=True, scenario_bank=bank,
                      hist_root_path=str(tmp_path), simu_name="fixture")
    env.max_number_steps = 6
    ag = getattr(agents, agent)(env, 25, 5, 1, np.array([135]))
    env.set_agent_functions(ag.obs_space_format, ag.action_format, ag.calculate_reward, ag.get_obs_space(),
                            ag.get_action_space())
    obs, _ = env.reset(seed=31, options={"initial_episode": 4})
    rng = np.random.default_rng(3)
    for k in range(6):
        if agent == "IBSched":
            a = {"player_0": rng.uniform(-1, 1, 5)}
            a.update({f"player_{i}": int(rng.integers(0, 3)) for i in range(1, 6)})
        else:
            a = ag.step(obs)
        obs, *_ = env.step(a)
    env.close()
    mine = os.path.join(str(tmp_path), "hist", "fixture", fixture, "ep_4.npz")
    keys, st = history_structure(mine)
    ref_keys, ref_st = history_structure(os.path.join(GOLDEN, f"hist_ref_{fixture}_ep_4.npz"))
    assert keys == ref_keys
    for k in ref_keys:
        assert st[k] == ref_st[k], (k, st[k], ref_st[k])
    f, g = np.load(mine, allow_pickle=True), np.load(os.path.join(GOLDEN, f"hist_ref_{fixture}_ep_4.npz"), allow_pickle=True)
    assert f["spectral_efficiencies"].any()              # recorded by default (results/gen_results.py:260-322 reads it)
    for k in ("slice_ue_assoc", "basestation_slice_assoc", "basestation_ue_assoc", "mobility"):
        assert np.array_equal(f[k], g[k]), k             # same scenario (episode 4 of the seed-10 associations)
