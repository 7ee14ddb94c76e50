"""CPU suite: scenario generator (associations/mult_slice.py generator mode, counter-based) in the oracle, and the
ep_N.npz exporter / importer (gen_assoc_mult_slice.py:229-237, associations/mult_slice.py:490-508)."""
import numpy as np


def test_generated_scenarios_follow_the_reference_distribution():
    from intent_radio_sched_multi_slice_b200 import scenario as scn
    from oracle import oracle

    bank = scn.ScenarioBank.templates()
    E = 3000
    env = oracle.OracleEnv(bank.array(), bank.poisson_cdf(), E, depth_cap=404, channel_mode=1, traffic_mode=1,
                           policy=3, assoc_mode=2, max_steps=4, auto_reset=1, seed=5)
    env.reset(initial_episode=0)
    first = env.scenario_rows()
    n_slices, types = [], []
    for r in first:
        st, us = scn.decode_row(r)
        used = st >= 0
        n_slices.append(int(used.sum()))
        assert len(set(st[used].tolist())) == used.sum()              # types without replacement (:454-466)
        for s in range(5):
            n = int((us == s).sum())
            assert r["sl"][s]["n_ues"] == n
            if used[s]:
                assert scn.SLICE_TYPES[st[s]][9] <= n <= scn.SLICE_TYPES[st[s]][10]   # :383-396
            else:
                assert n == 0
        types += st[used].tolist()
    cnt = np.bincount(n_slices, minlength=6)
    assert cnt[:3].sum() == 0 and (np.abs(cnt[3:] / E - 1 / 3) < 0.04).all()             # uniform 3..5 (:361-365)
    assert (np.abs(np.bincount(types, minlength=10) / len(types) - 0.1) < 0.02).all()
    for _ in range(5):  # crosses a reset: every env gets a NEW scenario for its next episode
        env.step()
    second = env.scenario_rows()
    assert (env.state()["episode"] == 1).all()
    changed = sum(not np.array_equal(a["ue_slice"], b["ue_slice"]) for a, b in zip(first, second))
    assert changed > 0.99 * E


def test_export_import_round_trip(tmp_path):
    from intent_radio_sched_multi_slice_b200 import plugins
    from intent_radio_sched_multi_slice_b200 import scenario as scn

    st, us = scn.generate_scenarios(6, seed=3)
    d = tmp_path / "associations" / "data" / "mult_slice"
    for i in range(6):
        scn.export_reference_episode(str(d), i, st[i], us[i], number_steps=3)
    f = np.load(d / "ep_2.npz", allow_pickle=True)
    assert set(f.files) == {"hist_basestation_ue_assoc", "hist_basestation_slice_assoc", "hist_slice_ue_assoc",
                            "hist_slice_req", "hist_slices_lifetime", "hist_slices_to_use"}
    assert f["hist_slice_ue_assoc"].shape == (3, 5, 25) and f["hist_basestation_slice_assoc"].shape == (3, 1, 5)
    back = plugins.load_reference_episodes(str(d), 6)
    want = scn.ScenarioBank()
    for i in range(6):
        want.add_typed(st[i], us[i])
    assert np.array_equal(back.array(), want.array())
