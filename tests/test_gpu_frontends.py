"""GPU suite: the front ends a trainer holds - the reference's own training arguments (enable_random_episodes),
the SB3 / RLlib vector adapters over the real batch, and ARBITRARY agent callbacks on the host path (the scheduling
matrix goes to the device as it is, irs_actions.sched_bits)."""
import numpy as np
import pytest

import golden_inputs
from gpu_adapter import CudaEnv, compare_outputs

pytestmark = pytest.mark.gpu


def _pack(sched):
    """dense 0/1 [E, 25, 135] -> uint32 [E, 25, 5] (bit r of the row = RB r)."""
    m = np.zeros(sched.shape[:2] + (160,), np.uint64)
    m[:, :, :135] = sched != 0
    return (m.reshape(sched.shape[0], 25, 5, 32) * (1 << np.arange(32, dtype=np.uint64))).sum(axis=3).astype(np.uint32)


@pytest.mark.parametrize("native", [True, False])
def test_given_scheduling_matrix_matches_oracle(bank, native):
    """Non-contiguous scheduling matrices (what an arbitrary action_format may produce): every RB to a random
    associated UE.  Native generators, and injected SE / traffic frames."""
    from oracle import oracle

    E, steps = 48, 40
    rng = np.random.default_rng(11)
    kw = dict(channel_mode=2 if native else 0, traffic_mode=1 if native else 0, policy=0, fixed_intra=-1, sort_slices=1,
              max_steps=15, auto_reset=1, seed=4, env_id_base=9, initial_episode=3)
    orc = oracle.OracleEnv(bank.array(), bank.poisson_cdf(), E, depth_cap=404, **kw)
    gpu = CudaEnv(bank, E, depth_cap=401, **kw)
    se0 = None if native else np.abs(rng.normal(9, 4, (E, 25, 135)))
    compare_outputs(gpu.reset(se=se0), orc.reset(se=se0), "reset", keys=["obs0", "obs_i", "mask0", "mask_i", "sorted_slices"])
    keys = [k for k in ("pkt_incoming", "pkt_throughputs", "sent", "dropped", "occ", "lat", "rb_count", "obs0", "mask0",
                        "obs_i", "mask_i", "sorted_slices", "reward", "done")]
    for k in range(steps):
        rows = orc.scenario_rows()
        sched = np.zeros((E, 25, 135))
        for e in range(E):
            ues = np.flatnonzero(rows[e]["ue_slice"][:25] >= 0)
            sched[e, rng.choice(ues, 135), np.arange(135)] = 1
            sched[e, :, rng.integers(0, 135, 10)] = 0           # some RBs stay unused
        bits = _pack(sched)
        se = None if native else np.abs(rng.normal(9, 4, (E, 25, 135)))
        tr = None if native else rng.poisson(20, (E, 25)) * 1e6
        want = orc.step(se=se, traffic=tr, sched_bits=bits)
        gpu.eng.step(se=se, traffic=tr, sched_bits=bits)
        got = gpu._out()
        compare_outputs(got, want, f"step {k}", keys=keys)
        assert np.array_equal(got["rb_count"], sched.sum(2))
    assert np.array_equal(gpu.buffers()[:, :, :401], orc.buffers()[:, :, :401])


def test_shared_rb_is_flagged(bank):
    gpu = CudaEnv(bank, 2, channel_mode=1, traffic_mode=1)
    gpu.reset(initial_episode=0)
    sched = np.zeros((2, 25, 135))
    sched[:, 0, :60] = 1
    sched[1, 1, 59] = 1                                         # env 1: RB 59 given to UEs 0 and 1
    gpu.eng.step(sched_bits=_pack(sched))
    assert gpu.eng.check_errors() == (1, 16)


def test_arbitrary_agent_callbacks_on_the_host_path(bank):
    """set_agent_functions with callables that are not a built-in surface (comm_env.py:484-512): action_format's
    matrix is what the device schedules, obs_space_format / calculate_reward see the reference's raw frame.  The
    recorded RB runs of a reference trace are fed back as the scheduling matrix, so every per-UE value of the
    frame must equal the trace."""
    from intent_radio_sched_multi_slice_b200 import plugins
    from intent_radio_sched_multi_slice_b200.env import MARLCommEnv
    from test_gpu_env_surface import _trace_plugins

    meta, d = golden_inputs.load_trace("ibsched_mimic")
    rse, sse, stf = golden_inputs.regenerate(meta, d)
    Ch, Tr, state = _trace_plugins(rse, sse, stf)
    env = MARLCommEnv(Ch, Tr, plugins.SimpleMobility, plugins.MultSliceAssociation, "mult_slice", "custom",
                      int(meta["seed"]), root_path="/nonexistent", initial_episode_number=int(meta["initial_episode"]),
                      save_hist=False, debug=True, scenario_bank=bank)
    seen = {}

    def action_format(action):                                 # any callable: here, "the action IS the run table"
        start, count = action
        r = np.arange(135)[None, :]
        return ((r >= start[:, None]) & (r < (start + count)[:, None])).astype(float)[None]

    def obs_space_format(frame):
        seen["frame"] = frame
        return np.concatenate([frame["buffer_occupancies"], frame["buffer_latencies"]])

    def calculate_reward(frame):
        return -float(np.sum(frame["dropped_pkts"]))

    env.set_agent_functions(obs_space_format, action_format, calculate_reward, None, None)
    obs, info = env.reset(seed=int(meta["seed"]), options={"initial_episode": int(meta["initial_episode"])})
    assert obs.shape == (50,) and not obs.any() and info == {}
    f = seen["frame"]
    assert np.array_equal(f["spectral_efficiencies"][0], rse[0]) and f["sched_decision"].shape == (1, 25, 135)
    par = next(p for s in f["slice_req"].values() if s for p in s["parameters"].values())
    assert callable(par["operator"])                          # agents/common.py:124-135 calls it
    for k in range(120):
        state["step"] = k
        obs, rew, term, trunc, info = env.step((d["rb_start"][k].astype(int), d["rb_count"][k].astype(int)))
        f = seen["frame"]
        assert set(f) == {"pkt_incoming", "pkt_throughputs", "pkt_effective_thr", "buffer_occupancies", "buffer_latencies",
                          "dropped_pkts", "mobility", "spectral_efficiencies", "basestation_ue_assoc",
                          "basestation_slice_assoc", "slice_ue_assoc", "sched_decision", "slice_req", "agent_action"}
        for key in ("pkt_incoming", "pkt_throughputs", "pkt_effective_thr", "dropped_pkts"):
            assert np.array_equal(f[key], d[key][k]), (key, k)
        np.testing.assert_allclose(f["buffer_occupancies"], d["buffer_occupancies"][k], rtol=1e-12)
        np.testing.assert_allclose(f["buffer_latencies"], d["buffer_latencies"][k], rtol=1e-12)
        assert np.array_equal(f["spectral_efficiencies"][0], sse[k])
        assert rew == -float(d["dropped_pkts"][k].sum()) and term is False and trunc is False
        assert np.array_equal(obs[:25], f["buffer_occupancies"])
    # an RB given to two UEs raises like comm_env.py:593-596
    with pytest.raises(Exception, match="same RB for more than one UE"):
        env.step((np.zeros(25, int), np.full(25, 3)))
    env.close()


def test_env_with_the_reference_training_arguments(bank):
    """simu.py:201-338 constructs every training env with enable_random_episodes=True: the batch plays per-env
    shuffled cycles of [initial_episode_number, max_episode_number)."""
    from intent_radio_sched_multi_slice_b200 import agents, plugins
    from intent_radio_sched_multi_slice_b200.env import BatchedMARLCommEnv

    env = BatchedMARLCommEnv(plugins.MimicQuadriga, plugins.MultSliceTraffic, plugins.SimpleMobility,
                             plugins.MultSliceAssociation, "mult_slice", "ib_sched", 10, root_path="/nonexistent",
                             initial_episode_number=10, max_episode_number=18, save_hist=False,
                             enable_random_episodes=True, scenario_bank=bank, num_envs=16, debug=True)
    env.max_number_steps = 2
    ag = agents.IBSchedSB3(env, 25, 5, 1, np.array([135]))
    env.set_agent_functions(ag.obs_space_format, ag.action_format, ag.calculate_reward, ag.get_obs_space(), ag.get_action_space())
    env.reset()
    eps = [env.episode_number.cpu().numpy().copy()]
    for k in range(2 * 16):
        env.step(np.zeros((16, 5)))
        if k % 2 == 1:
            eps.append(env.episode_number.cpu().numpy().copy())
    eps = np.array(eps)
    for e in range(16):
        assert sorted(eps[:8, e]) == list(range(10, 18)) and np.array_equal(eps[:8, e], eps[8:16, e])
    assert len({tuple(eps[:8, e]) for e in range(16)}) > 8
    env.close()


def test_sb3_vec_env_over_the_batch(bank):
    """SB3VecEnv against the same batch stepped with the device's own auto-reset: same observations and rewards,
    and the terminal observation of a finished env is what a non-resetting step returns."""
    import torch

    from intent_radio_sched_multi_slice_b200 import agents, plugins, vector
    from intent_radio_sched_multi_slice_b200.env import BatchedMARLCommEnv

    args = (plugins.MimicQuadriga, plugins.MultSliceTraffic, plugins.SimpleMobility, plugins.MultSliceAssociation,
            "mult_slice", "sb3", 5)
    kw = dict(root_path="/nonexistent", save_hist=False, scenario_bank=bank, num_envs=12, initial_episode_number=4,
              debug=True)
    venv = vector.SB3VecEnv(args, obs_dtype=None, **kw)
    ref = BatchedMARLCommEnv(*args, auto_reset=True, **kw)
    ag = agents.IBSchedSB3(ref, 25, 5, 1, np.array([135]))
    ref.set_agent_functions(ag.obs_space_format, ag.action_format, ag.calculate_reward, ag.get_obs_space(), ag.get_action_space())
    venv.env.max_number_steps = ref.max_number_steps = 4
    obs = venv.reset()
    robs, _ = ref.reset()
    assert obs.shape == (12, 50) and np.array_equal(obs, robs.cpu().numpy())
    rng = np.random.default_rng(2)
    n_done = 0
    for k in range(11):
        a = rng.uniform(-1, 1, (12, 5))
        obs, rew, dones, infos = venv.step(a)
        robs, rrew, rdone, _, _ = ref.step(a)
        assert np.array_equal(obs, robs.cpu().numpy()), k       # after a rollover both hold the reset observation
        np.testing.assert_allclose(rew, rrew.cpu().numpy().astype(np.float32))
        assert np.array_equal(dones, rdone.cpu().numpy())
        if dones.any():
            n_done += 1
            assert all(i["terminal_observation"].shape == (50,) and i["TimeLimit.truncated"] for i in infos)
            assert not np.array_equal(infos[0]["terminal_observation"], obs[0])
    assert n_done == 2
    venv.close(); ref.close()


def test_rllib_vector_env_over_the_batch(bank):
    from intent_radio_sched_multi_slice_b200 import plugins, vector
    from intent_radio_sched_multi_slice_b200.env import MARLCommEnv

    args = (plugins.FixedSE, plugins.MultSliceTraffic, plugins.SimpleMobility, plugins.MultSliceAssociation,
            "mult_slice", "ib_sched", 5)
    kw = dict(root_path="/nonexistent", save_hist=False, scenario_bank=bank, initial_episode_number=7, debug=True)
    venv = vector.RLlibVectorEnv(args, num_envs=3, **kw)
    venv.env.max_number_steps = 3
    obs, _ = venv.vector_reset()
    rng = np.random.default_rng(0)
    acts = []
    for k in range(3):
        a = [{"player_0": rng.uniform(-1, 1, 5), **{f"player_{i}": int(rng.integers(0, 3)) for i in range(1, 6)}} for _ in range(3)]
        acts.append(a)
        obs, rew, term, trunc, infos = venv.vector_step(a)
    assert all(t["__all__"] for t in term) and set(rew[0]) == {f"player_{i}" for i in range(6)}
    # env 1 of the batch = a single reference-shaped env with the same global id
    one = MARLCommEnv(*args, env_id_base=1, **kw)
    one.max_number_steps = 3
    from intent_radio_sched_multi_slice_b200 import agents
    ag = agents.IBSched(one, 25, 5, 1, np.array([135]))
    one.set_agent_functions(ag.obs_space_format, ag.action_format, ag.calculate_reward, ag.get_obs_space(), ag.get_action_space())
    one.reset()
    for k in range(3):
        o1, r1, t1, _, _ = one.step(acts[k][1])
    assert list(o1.keys()) == list(obs[1].keys())
    for p in o1:
        assert np.array_equal(o1[p]["observations"], obs[1][p]["observations"])
    assert r1 == rew[1] and t1 == term[1]
    o, info = venv.reset_at(1)
    assert o["player_0"]["observations"].shape == (50,) and int(venv.env.episode_number[1]) == 8
    venv.close(); one.close()
