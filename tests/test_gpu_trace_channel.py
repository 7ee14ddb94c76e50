"""GPU suite: the device-resident SE trace channel (channel_mode 3) - QuadrigaChannel /
QuadrigaChannelSeq of the reference (channels/quadriga.py:38-87, channels/quadriga_seq.py:28-39)
as a bank in HBM - against the traces recorded from the unmodified reference, and against the CPU
oracle on a batch with native traffic, random actions and auto-resets."""
import numpy as np
import pytest

import golden_inputs
import trace_replay
from gpu_adapter import CudaEnv, compare_outputs
from test_oracle_golden import BankEnv

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", ["quadriga", "quadriga_seq"])
def test_cuda_inject_matches_quadriga_trace(name, bank):
    from test_gpu_parity import _cuda_env_factory

    trace_replay.replay(name, _cuda_env_factory(bank), tight=False)


@pytest.mark.parametrize("name", ["quadriga", "quadriga_seq"])
def test_cuda_power_to_se_and_bank_mode_match_reference(name, bank):
    """irs_power_to_se reproduces the frames the reference computed (<= 2 ulp: CUDA log2 vs NumPy's),
    and a channel_mode 3 env over that bank reproduces every output of the reference trace."""
    import torch

    meta, d = golden_inputs.load_trace(name)
    rse, sse, _ = golden_inputs.regenerate(meta, d)
    power = golden_inputs.quadriga_bank_powers(meta, d)
    made = {}

    def make(meta_):
        sort, fixed = trace_replay.agent_config(meta_)
        env = CudaEnv(bank, 1, channel_mode=3, traffic_mode=0, sort_slices=sort, fixed_intra=fixed,
                      assoc_mode=0 if meta_["assoc"] == "mult_slice" else 1, max_steps=int(meta_["max_steps"]),
                      max_episodes=int(meta_["max_episodes"]), initial_episode=int(meta_["initial_episode"]))
        se_bank = env.eng.power_to_se(torch.from_numpy(power).to(env.eng.device), layout=0)
        env.eng.set_se_trace(se_bank, wrap=True)
        made["bank"] = se_bank.cpu().numpy()
        return BankEnv(env)

    trace_replay.replay(name, make, tight=False)
    k_in_ep = 0
    for k in range(int(meta["n_steps"])):
        np.testing.assert_allclose(made["bank"][int(d["episode"][k]) % 8, k_in_ep], sse[k], rtol=5e-16, atol=0)
        k_in_ep = 0 if d["done"][k] else k_in_ep + 1
    # layout 1 ([25][135] input) is the same transform without the transposition
    env = CudaEnv(bank, 1, channel_mode=1, traffic_mode=1)
    p1 = torch.from_numpy(np.ascontiguousarray(power.transpose(0, 1, 3, 2))).cuda()
    assert np.array_equal(env.eng.power_to_se(p1, layout=1).cpu().numpy(), made["bank"])


@pytest.mark.parametrize("raw_se", [False, True])
def test_batch_trace_bank_matches_oracle(bank, raw_se):
    """300 envs over a 16-episode x 12-step bank, native Poisson traffic, random on-device
    policy, auto-reset: every output of every step equals the oracle's (also with the per-RB SE
    output enabled, which switches to the kernel variant that reads every row)."""
    import torch

    from oracle import oracle

    E, steps, n_eps, T = 300, 50, 16, 12
    rng = np.random.default_rng(3)
    se_bank = np.abs(rng.normal(8, 4, (n_eps, T, 25, 135)))
    kw = dict(channel_mode=3, traffic_mode=1, policy=3, fixed_intra=-1, sort_slices=1, assoc_mode=0, max_steps=T,
              auto_reset=1, env_id_base=77, seed=99, initial_episode=0)
    orc = oracle.OracleEnv(bank.array(), bank.poisson_cdf(), E, depth_cap=404, **kw)
    orc.set_se_trace(se_bank, wrap=True)
    gpu = CudaEnv(bank, E, depth_cap=401, **kw)
    if not raw_se:  # drop the per-RB output: the production kernel skips rows of unassociated UEs
        gpu.raw = gpu.eng.enable_raw_outputs([n for n in gpu.raw if n != "se_used"])
        gpu.raw["se_used"] = None
    gpu.eng.set_se_trace(torch.from_numpy(se_bank).to(gpu.eng.device), wrap=True)
    ie = (np.arange(E) % 7 + 2).astype(np.int32)
    for e in range(E):
        m = np.zeros(E, np.uint8)
        m[e] = 1
        want = orc.reset(mask=m, initial_episode=int(ie[e]))
    got = gpu.reset(initial_episodes=ie, initial_episode=0)
    compare_outputs(got, want, "reset", keys=["obs0", "obs_i", "mask0", "mask_i", "sorted_slices"])
    if raw_se:
        orc.want_se()
    for k in range(steps):
        want, got = orc.step(), gpu.step()
        compare_outputs(got, want, f"step {k}")
        if raw_se:
            assert np.array_equal(gpu.raw["se_used"].cpu().numpy(), want["se_used"]), k
    assert np.array_equal(gpu.buffers()[:, :, :401], orc.buffers()[:, :, :401])
