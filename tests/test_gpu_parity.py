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
from gpu_adapter import CudaEnv, compare_outputs

pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True, params=["warp_per_env", "thread_per_env"])
def per_env_kernels(request, monkeypatch):
    """Every test of this file runs with both families of per-env kernels: k_action / k_env_tail (one warp
    per env, small batches) and k_action_t / k_env_tail_t + k_rollover (one thread per env, the default from
    8,192 envs per launch; irs_serial.cuh)."""
    monkeypatch.setenv("IRS_SERIAL_MIN_ENVS", "1" if request.param == "thread_per_env" else str(1 << 30))

TRACES = ["cfg1_marr", "mapf", "ibsched_mimic", "ibsched_fixed", "seq_mimic", "seq_fixed", "sweep_scen", "twc", "colran"]


def _cuda_env_factory(bank):
    def make(meta):
        sort, fixed = trace_replay.agent_config(meta)
        return CudaEnv(bank, 1, channel_mode=0, traffic_mode=0, sort_slices=sort, fixed_intra=fixed,
                       surface=trace_replay.agent_surface(meta),
                       assoc_mode=0 if meta["assoc"] == "mult_slice" else 1, max_steps=int(meta["max_steps"]),
                       max_episodes=int(meta["max_episodes"]), initial_episode=int(meta["initial_episode"]))
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
    # live extent of every row (irs_common.cuh ROW_LIVE): exactly the oldest non-empty slot + 1 for a
    # streamed row, an upper bound otherwise - the row kernels never look past it
    word = gpu.eng.v["path"].long()
    live = (word >> 2) & 0x3FFF
    idx = torch.arange(gpu.eng.buf.shape[2], device="cuda")
    oldest = ((gpu.eng.buf != 0).long() * (idx + 1)).max(2).values
    assert bool((oldest <= live).all())
    assert bool((oldest == live)[(word & 3) == 2].all())
    # no RB is given to two UEs (comm_env.py:593-596): runs of one env do not overlap
    rs, rc = gpu.raw["rb_start"].long(), gpu.raw["rb_count"].long()
    cover = torch.zeros(E, 136, dtype=torch.int64, device="cuda")
    cover.scatter_add_(1, rs, (rc > 0).long())
    cover.scatter_add_(1, rs + rc, -(rc > 0).long())
    assert bool((cover.cumsum(1).max(1).values <= 1).all())
