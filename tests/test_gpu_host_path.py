"""GPU suite: irs_step_host (host actions in, host surface out) must return exactly what the
device-resident irs_step produces - also when the batch is pipelined in chunks (kernels of chunk
c+1 overlap the device->host copy of chunk c) and the chunk size does not divide the batch."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("chunks,extra", [(1, {}), (3, {}), (8, {}), (3, dict(surface=1, fixed_intra=0, sort_slices=0)),
                                          (5, dict(surface=2, fixed_intra=0, sort_slices=0, assoc_mode=2)),
                                          (0, dict(E=40_000)), (0, dict(E=70_001))])
def test_step_host_equals_device_step(bank, chunks, extra, monkeypatch):
    """chunks = 0: the DEFAULT chunk plan of irs_step_host (whole waves of the row kernel, batches >= 16,384 envs) - the
    code the e2e figure of bench.py runs - at 40,000 envs (below the thread-per-env threshold: every chunk runs the
    warp-per-env kernels) and at 70,001 (the device-resident step runs the thread-per-env kernels, the chunks do not)."""
    import torch

    from intent_radio_sched_multi_slice_b200 import capi
    from intent_radio_sched_multi_slice_b200.engine import BatchEngine

    extra = dict(extra)
    E = extra.pop("E", 1000)  # 1000: not a multiple of 3 or 8
    kw = dict(channel_mode=2, traffic_mode=1, sort_slices=1, fixed_intra=-1, policy=0, scen_mod=200, max_steps=9,
              auto_reset=1, seed=21, env_id_base=3, depth_cap=401)
    kw.update(extra)  # the SchedTWC / SchedColORAN surfaces (extra kernels around the rollover) and generated scenarios
    if kw.get("assoc_mode") == 2:  # scenarios drawn on the device from the 10 slice-type templates
        from intent_radio_sched_multi_slice_b200 import scenario as scn
        bank = scn.ScenarioBank.templates()
        kw.pop("scen_mod")
    if chunks:
        monkeypatch.setenv("IRS_HOST_CHUNKS", str(chunks))
    else:
        monkeypatch.delenv("IRS_HOST_CHUNKS", raising=False)
    host = BatchEngine(bank, E, **kw)
    monkeypatch.delenv("IRS_HOST_CHUNKS", raising=False)
    dev = BatchEngine(bank, E, **kw)
    ep = (np.arange(E) % 200).astype(np.int32)
    host.reset(initial_episodes=ep)
    dev.reset(initial_episodes=ep)
    pin = lambda *s, dt: torch.empty(*s, dtype=dt).pin_memory()
    hs = {"obs0": pin(E, 50, dt=torch.float64), "obsi": pin(E, 5, 19, dt=torch.float64),
          "mask0": pin(E, 5, dt=torch.int8), "maski": pin(E, 5, 5, dt=torch.int8),
          "reward": pin(E, 6, dt=torch.float64), "done": pin(E, dt=torch.uint8)}
    h32 = {"obs0_f32": pin(E, 50, dt=torch.float32), "obsi_f32": pin(E, 5, 19, dt=torch.float32),
           "reward_f32": pin(E, 6, dt=torch.float32)}
    surf = capi.HostSurface(*[hs[n].data_ptr() for n in capi.SURFACE_FIELDS],
                            *[h32[n].data_ptr() for n in capi.SURFACE_F32_FIELDS])
    rng = np.random.default_rng(5)
    names = {"obs0": "obs0", "obsi": "obs_i", "mask0": "mask0", "maski": "mask_i", "reward": "reward", "done": "done"}
    for k in range(25 if E <= 1000 else 11):  # crosses two auto-resets (max_steps = 9) / one
        a0 = torch.from_numpy(rng.uniform(-1, 1, (E, 5))).pin_memory()
        intra = torch.from_numpy(rng.integers(0, 3, (E, 5)).astype(np.int32)).pin_memory()
        host.step_host(a0, intra, surf)
        dev.step(a0, intra)
        torch.cuda.synchronize()
        s = dev.surface()
        for hn, dn in names.items():
            assert np.array_equal(hs[hn].numpy(), s[dn].cpu().numpy()), f"step {k}: {hn} differs (chunks={chunks})"
        for hn, dn in (("obs0_f32", "obs0"), ("obsi_f32", "obs_i"), ("reward_f32", "reward")):  # binary32 copies
            assert np.array_equal(h32[hn].numpy(), s[dn].cpu().numpy().astype(np.float32)), f"step {k}: {hn}"
    assert host.check_errors() is None
    assert torch.equal(host.buf, dev.buf)
    host.close()
    dev.close()


def test_step_is_cuda_graph_capturable(bank):
    """irs_step allocates nothing, copies nothing and never synchronises (include/irs_b200.h): a whole step is
    captured into a CUDA graph and the replays produce exactly what eager launches produce."""
    import torch

    from intent_radio_sched_multi_slice_b200.engine import BatchEngine

    E = 512
    kw = dict(channel_mode=2, traffic_mode=1, sort_slices=1, fixed_intra=-1, policy=3, scen_mod=200, max_steps=7,
              auto_reset=1, seed=5, depth_cap=401)
    a, b = BatchEngine(bank, E, **kw), BatchEngine(bank, E, **kw)
    ep = (np.arange(E) % 200).astype(np.int32)
    a.reset(initial_episodes=ep)
    b.reset(initial_episodes=ep)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    side = torch.cuda.Stream()
    with torch.cuda.stream(side):
        a.step()              # warm-up on the capture stream
        b_steps = 1
        side.synchronize()
        with torch.cuda.graph(g, stream=side):
            a.step()
    for _ in range(20):       # crosses two auto-resets
        g.replay()
    torch.cuda.synchronize()
    for _ in range(b_steps + 20):
        b.step()
    torch.cuda.synchronize()
    # the capture itself does not execute: a ran 1 eager + 20 replayed steps
    sa, sb = a.surface(), b.surface()
    for k in ("obs0", "obs_i", "mask0", "mask_i", "reward", "done"):
        assert torch.equal(sa[k], sb[k]), k
    assert torch.equal(a.buf, b.buf) and torch.equal(a.v["episode"], b.v["episode"])
    a.close()
    b.close()


def test_engine_graph_mode_is_bit_identical(bank):
    """BatchEngine(graph=True): every step is a replay of one captured graph of irs_step, actions staged into the
    state blob.  Same results as eager launches, with supplied actions and across auto-resets (BASELINE config 2 size)."""
    import torch

    from intent_radio_sched_multi_slice_b200.engine import BatchEngine

    E = 1024
    kw = dict(channel_mode=2, traffic_mode=1, sort_slices=1, fixed_intra=-1, policy=0, scen_mod=200, max_steps=6,
              auto_reset=1, seed=5, depth_cap=401)
    a, b = BatchEngine(bank, E, graph=True, **kw), BatchEngine(bank, E, **kw)
    ep = (np.arange(E) % 200).astype(np.int32)
    a.reset(initial_episodes=ep)
    b.reset(initial_episodes=ep)
    g = torch.Generator(device="cuda")
    g.manual_seed(1)
    for k in range(15):
        a0 = torch.rand(E, 5, generator=g, device="cuda", dtype=torch.float64) * 2 - 1
        intra = torch.randint(0, 3, (E, 5), generator=g, device="cuda", dtype=torch.int32)
        a.step(a0, intra)
        b.step(a0, intra)
    torch.cuda.synchronize()
    sa, sb = a.surface(), b.surface()
    for k in ("obs0", "obs_i", "mask0", "mask_i", "reward", "done"):
        assert torch.equal(sa[k], sb[k]), k
    assert torch.equal(a.buf, b.buf) and torch.equal(a.v["episode"], b.v["episode"]) and a._graph is not None
    a.close()
    b.close()
