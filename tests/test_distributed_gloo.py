"""CPU suite: the N>1 host logic with world_size=2 over gloo - sharding of global env ids, the
episode-metric all-reduce and the max-over-ranks timing helper.  The per-rank "engine" here is
the CPU oracle (tests only): what is checked is that sharded runs reduce to the unsharded one."""
import os
import sys

import numpy as np
import pytest
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_shard_partitions_contiguously():
    from intent_radio_sched_multi_slice_b200 import distributed as D

    for total, world in ((16384, 8), (65536, 4), (10, 3), (7, 7)):
        spans = [D.shard(total, r, world) for r in range(world)]
        assert spans[0][0] == 0 and sum(n for _, n in spans) == total
        for (b0, n0), (b1, _) in zip(spans, spans[1:]):
            assert b0 + n0 == b1
        assert max(n for _, n in spans) - min(n for _, n in spans) <= 1
    with pytest.raises(ValueError):
        D.shard(2, 0, 4)


def _worker(rank, world, port, total_envs, steps, out_path):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    import torch.distributed as dist

    from intent_radio_sched_multi_slice_b200 import distributed as D
    from intent_radio_sched_multi_slice_b200 import scenario as scn
    from oracle import oracle

    D.init("gloo")
    st, us = scn.generate_scenarios(12, seed=4)
    bank = scn.ScenarioBank()
    for i in range(12):
        bank.add_typed(st[i], us[i])
    base, local = D.shard(total_envs)
    env = oracle.OracleEnv(bank.array(), bank.poisson_cdf(), local, depth_cap=404, channel_mode=2, traffic_mode=1,
                           policy=3, max_steps=9, auto_reset=1, seed=21, scen_mod=12, env_id_base=base)
    env.reset(initial_episode=1)
    for _ in range(steps):
        env.step()
    _, tot = env.episode_metrics()
    red = D.all_reduce_metrics(tot)
    slow = D.max_over_ranks(1.0 + rank)
    if rank == 0:
        np.save(out_path, np.array([red[k] for k in red] + [slow]))
    dist.barrier()
    dist.destroy_process_group()


def test_sharded_metrics_equal_unsharded(tmp_path):
    """2 ranks x 24 envs == 1 process x 48 envs: integer counters bit-equal, float sums to 1e-12
    (the generators are keyed by global env id, so sharding changes nothing but the sum order)."""
    from intent_radio_sched_multi_slice_b200 import capi
    from intent_radio_sched_multi_slice_b200 import scenario as scn
    from oracle import oracle

    total, steps = 48, 30
    out = str(tmp_path / "red.npy")
    port = 29500 + (os.getpid() % 2000)
    mp.spawn(_worker, args=(2, port, total, steps, out), nprocs=2, join=True)
    red = np.load(out)
    assert red[-1] == 2.0  # max over ranks of (1 + rank)
    st, us = scn.generate_scenarios(12, seed=4)
    bank = scn.ScenarioBank()
    for i in range(12):
        bank.add_typed(st[i], us[i])
    env = oracle.OracleEnv(bank.array(), bank.poisson_cdf(), total, depth_cap=404, channel_mode=2, traffic_mode=1,
                           policy=3, max_steps=9, auto_reset=1, seed=21, scen_mod=12, env_id_base=0)
    env.reset(initial_episode=1)
    for _ in range(steps):
        env.step()
    _, tot = env.episode_metrics()
    names = capi.METRIC_NAMES
    for k in ("sent_pkts", "dropped_pkts", "incoming_pkts", "violations", "violations_priority", "env_steps", "episodes"):
        assert red[names.index(k)] == tot[names.index(k)], k
    np.testing.assert_allclose(red[:16], tot, rtol=1e-12)
    assert tot[names.index("episodes")] == total * (steps // 9)
