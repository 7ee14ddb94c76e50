"""CPU suite: the two on-disk formats against the REFERENCE's own readers and writers.

  * scenario.export_reference_episode -> associations/mult_slice.py:490-508 load_episode_data + :416-432 step (replay
    mode), run from /root/reference when it is present (build container; skipped on the GPU box);
  * the episode-history fixture the unmodified reference wrote with save_hist=True (metrics.py:73-100; recorded by
    oracle/refharness/gen_hist_fixture.py) holds the structure history.py documents - the GPU suite compares the files
    history.py writes against the same fixture.
"""
import os

import numpy as np
import pytest

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
HAVE_REF = os.path.isdir("/root/reference/associations")


def structure(a, depth=0):
    """dtype / shape / nested key structure of a history entry (values dropped)."""
    if isinstance(a, np.ndarray) and a.dtype == object:
        return ("objarray", a.shape, structure(a.flat[0], depth + 1))
    if isinstance(a, np.ndarray):
        return ("array", str(a.dtype), a.shape[1:] if depth == 0 else a.shape)
    if isinstance(a, dict):
        return ("dict", tuple((k, structure(v, depth + 1)) for k, v in a.items()))
    if callable(a):
        return ("callable", getattr(a, "__name__", ""))
    return (type(a).__name__,)


def history_structure(path):
    f = np.load(path, allow_pickle=True)
    out = {}
    for k in f.files:
        a = f[k]
        if k == "slice_req":  # per-slot dicts differ between scenarios: compare one non-empty slot's shape
            req = a.flat[0]
            slot = next(v for v in req.values() if v)
            par = next(iter(slot["parameters"].values()))
            out[k] = ("objarray", a.shape[1:], sorted(req), sorted(slot), sorted(slot["ues"]) , sorted(par),
                      callable(par["operator"]))
        elif k == "obs" and a.dtype == object:  # key ORDER follows the sorted slices: compare as a set + player_0 first
            o = a.flat[0]
            out[k] = ("objarray", a.shape[1:], list(o)[0], sorted(o), structure(o["player_0"], 1), structure(o["player_3"], 1))
        else:
            out[k] = structure(a)
    return list(f.files), out


@pytest.mark.parametrize("agent", ["ib_sched", "marr"])
def test_reference_history_fixture_has_the_documented_structure(agent):
    from intent_radio_sched_multi_slice_b200 import history

    keys, st = history_structure(os.path.join(GOLDEN, f"hist_ref_{agent}_ep_4.npz"))
    assert keys == history.KEYS                       # insertion order of Metrics.metrics_hist (metrics.py:43-60)
    assert st["pkt_effective_thr"] == ("array", "int64", (25,)) and st["dropped_pkts"] == ("array", "int64", (25,))
    assert st["sched_decision"] == ("array", "float64", (1, 25, 135)) == st["spectral_efficiencies"]
    assert st["slice_req"][-1] is True                # operators are pickled NumPy ufuncs
    if agent == "ib_sched":
        assert st["reward"][0] == "objarray" and st["obs"][2] == "player_0"
    else:
        assert st["reward"] == ("array", "float64", ()) and st["obs"] == ("array", "float64", (50,))


@pytest.mark.skipif(not HAVE_REF, reason="/root/reference is only present in the build container")
def test_exported_episode_is_replayed_by_the_reference_association(tmp_path):
    """export_reference_episode -> the unmodified MultSliceAssociation (replay mode) returns the associations and
    slice requirements the scenario bank compiles from the same (slice types, UE slices) pair."""
    from intent_radio_sched_multi_slice_b200 import scenario as scn
    from oracle.refharness import shims

    shims.install()
    from associations.mult_slice import MultSliceAssociation
    from sixg_radio_mgmt import UEs

    st, us = scn.generate_scenarios(5, seed=8)
    d = tmp_path / "associations" / "data" / "mult_slice"
    for i in range(5):
        scn.export_reference_episode(str(d), i, st[i], us[i], number_steps=4)
    ues = UEs(25, np.full(25, 100), np.full(25, 10240), np.full(25, 65536))
    assoc = MultSliceAssociation(ues, 25, 1, 5, np.random.default_rng(0), str(tmp_path), scenario_name="mult_slice")
    for ep in (3, 0, 4):
        bua, bsa, sua, req = assoc.step(np.zeros((1, 25)), np.zeros((1, 5)), np.zeros((5, 25)), {}, 0, ep)
        want = scn.ScenarioBank()
        want.add_typed(st[ep], us[ep])
        got = scn.ScenarioBank()
        for v in req.values():  # the reference holds operators as ufuncs: the bank takes their names
            for par in (v.get("parameters", {}) if v else {}).values():
                par["operator"] = scn._operator_name(par["operator"])
        got.add(req, sua, bsa)
        assert np.array_equal(got.array(), want.array()), ep
        assert np.array_equal(bua[0], (np.asarray(us[ep][:25]) >= 0).astype(float))
        # update_ues (associations/mult_slice.py:468-488) pushed the slices' buffer parameters into the UEs
        for u in range(25):
            s = int(us[ep][u])
            if s >= 0:
                t = scn.SLICE_TYPES[int(st[ep][s])]
                assert ues.buffers[u].max_packets_buffer == t[5] and ues.buffers[u].max_packets_age == t[6]
