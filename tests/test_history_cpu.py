"""CPU suite: host logic of the episode-history writer (no device needed)."""
import numpy as np


def test_history_keys_are_the_reference_metrics_keys():
    from intent_radio_sched_multi_slice_b200 import history

    # sixg_radio_mgmt/sixg_radio_mgmt/metrics.py:43-60 (insertion order of Metrics.metrics_hist)
    assert history.KEYS == ["pkt_incoming", "pkt_throughputs", "pkt_effective_thr", "buffer_occupancies",
                            "buffer_latencies", "dropped_pkts", "mobility", "spectral_efficiencies",
                            "basestation_ue_assoc", "basestation_slice_assoc", "slice_ue_assoc", "sched_decision",
                            "reward", "slice_req", "obs", "agent_action"]


def test_dense_sched_expands_rb_runs():
    from intent_radio_sched_multi_slice_b200 import history

    start = np.zeros((2, 25), np.int16)
    count = np.zeros((2, 25), np.int16)
    start[0, 3], count[0, 3] = 10, 125
    start[0, 7], count[0, 7] = 0, 10
    start[1, 24], count[1, 24] = 134, 1
    d = history.dense_sched(start, count)
    assert d.shape == (2, 1, 25, 135) and d.dtype == np.float64
    assert d[0, 0, 3, 10:].all() and not d[0, 0, 3, :10].any() and d[0, 0, 7, :10].all() and d[0, 0, 7].sum() == 10
    assert d[0, 0].sum(0).max() == 1  # comm_env.py:593-596: an RB goes to at most one UE
    assert d[1].sum() == 1 and d[1, 0, 24, 134] == 1
