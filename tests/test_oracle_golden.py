"""CPU suite: the oracle (oracle/irs_oracle.c) against every golden vector we hold for the
path - the reference's Buffer.main() known answer (buffer.py:133-149, values recorded in
SURVEY.md 8c) and the traces recorded from the unmodified reference (tests/golden/)."""
import numpy as np
import pytest

import golden_inputs
import trace_replay
from oracle import oracle

TRACES = ["cfg1_marr", "mapf", "ibsched_mimic", "ibsched_fixed", "seq_mimic", "seq_fixed", "sweep_scen"]

# python sixg_radio_mgmt/sixg_radio_mgmt/buffer.py  (rows after receive -> after send)
BUFFER_MAIN = [
    ([0, 0, 0, 0, 0, 0], [0, 0, 0, 0, 0, 0], 0), ([10, 0, 0, 0, 0, 0], [8, 0, 0, 0, 0, 0], 0),
    ([0, 8, 0, 0, 0, 0], [0, 6, 0, 0, 0, 0], 0), ([5, 0, 6, 0, 0, 0], [5, 0, 4, 0, 0, 0], 0),
    ([2, 5, 0, 4, 0, 0], [2, 5, 0, 2, 0, 0], 0), ([2, 2, 5, 0, 2, 0], [2, 2, 5, 0, 0, 0], 0),
    ([10, 2, 2, 5, 0, 0], [10, 2, 2, 3, 0, 0], 0), ([3, 10, 2, 2, 3, 0], [3, 10, 2, 2, 1, 0], 12),
    ([2, 3, 10, 2, 2, 1], [2, 3, 10, 2, 1, 0], 18), ([2, 2, 3, 10, 2, 1], [2, 2, 3, 10, 1, 0], 8),
]


def test_buffer_main_known_answer():
    rows, dropped, sent, delay = oracle.buffer_main(20, 5, [0, 10, 0, 5, 2, 2, 10, 15, 20, 10], 2)
    for k, (after_rx, after_tx, drop) in enumerate(BUFFER_MAIN):
        assert rows[k, 0].tolist() == after_rx and rows[k, 1].tolist() == after_tx and dropped[k] == drop
    assert delay == 2.3333333333333335


def test_numpy_pairwise_sum_model():
    rng = np.random.default_rng(0)
    for n in (1, 5, 7, 8, 9, 64, 71, 127, 128, 129, 135, 136, 300):
        for _ in range(200):
            a = np.abs(rng.normal(10, 5, n)) * rng.choice([1.0, 740740.7407407408])
            assert oracle.np_sum(a) == np.sum(a), n
    a = np.full(135, 2.0 * (1e8 / 135))
    a[27:] = 0
    assert oracle.np_sum(a) == np.sum(a) == 4.0e7  # exact only in NumPy's order


def test_round_int_equal_sum_tie_rule():
    # stable argsort reversed: ties -> higher index first
    assert oracle.round_int_equal_sum([5.4, 5.4, 5.4, 5.4, 5.4], 27).tolist() == [5, 5, 5, 6, 6]
    assert oracle.round_int_equal_sum([0, 13.5, 0, 13.5, 0], 27).tolist() == [0, 13, 0, 14, 0]
    assert oracle.round_int_equal_sum([1.0, 26.0, 0, 0, 0], 27).tolist() == [1, 26, 0, 0, 0]


def make_oracle_env(bank):
    def make(meta):
        sort, fixed = trace_replay.agent_config(meta)
        env = oracle.OracleEnv(bank.array(), bank.poisson_cdf(), 1, channel_mode=0, traffic_mode=0,
                               sort_slices=sort, fixed_intra=fixed,
                               assoc_mode=0 if meta["assoc"] == "mult_slice" else 1,
                               max_steps=int(meta["max_steps"]), max_episodes=int(meta["max_episodes"]),
                               initial_episode=int(meta["initial_episode"]))
        return env
    return make


@pytest.mark.parametrize("name", TRACES)
def test_oracle_matches_reference_trace(name, bank):
    trace_replay.replay(name, make_oracle_env(bank), tight=True)


def test_golden_traces_cover_the_edge_cases():
    """The fixtures must exercise what the domain offers: drops, overflow clips, empty buffers,
    every intra scheduler, every slice type, resets and a scenario change."""
    st, _ = golden_inputs.load_scenarios()
    seen_types, intra = set(), set()
    drops = 0
    for name in TRACES:
        meta, d = golden_inputs.load_trace(name)
        drops += int(d["dropped_pkts"].sum() > 0)
        intra |= set(np.unique(d["intra"]).tolist())
        for ep in np.unique(d["episode"]):
            seen_types |= set(st[golden_inputs.scen_of_episode(meta, int(ep))].tolist())
    assert drops >= 3 and {0, 1, 2} <= intra and set(range(10)) <= seen_types
