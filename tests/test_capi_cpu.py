"""CPU suite: the C-ABI library builds for sm_100a, loads, and exports every symbol that
include/irs_b200.h declares; handle creation and the state layout work without a GPU (no
compute call is made here)."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from intent_radio_sched_multi_slice_b200 import build, capi, scenario

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    hdr = open(os.path.join(ROOT, "include", "irs_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    return sorted(set(re.findall(r"\b(irs_[a-z_]+)\s*\(", hdr)))


def test_library_builds_and_exports_every_declared_symbol():
    path = build.build()
    assert os.path.exists(path)
    lib = capi.load()
    declared = _declared_symbols()
    assert len(declared) >= 14
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in irs_b200.h but not exported"
    assert set(declared) == set(capi.SYMBOLS), set(declared) ^ set(capi.SYMBOLS)
    assert lib.irs_version() == 1


def test_struct_sizes_match_the_header():
    assert C.sizeof(capi.Config) == 18 * 4 + 8 + 8
    assert scenario.SCENARIO_DTYPE.itemsize == 432 and scenario.SLICE_DTYPE.itemsize == 80
    assert C.sizeof(capi.RawOutputs) == 13 * 8 and C.sizeof(capi.HostSurface) == 9 * 8


def _cfg(**kw):
    d = dict(num_envs=64, depth_cap=401, channel_mode=2, traffic_mode=1, sort_slices=1, fixed_intra=-1, policy=0,
             assoc_mode=0, scen_mod=200, scen_div=100, n_scen=200, max_steps=1000, max_episodes=10000,
             initial_episode=0, auto_reset=0, env_id_base=0, surface=0, random_episodes=0, seed=10, bandwidth=1e8)
    d.update(kw)
    return capi.Config(*[d[n] for n, _ in capi.Config._fields_])


def test_create_layout_destroy_without_gpu():
    lib = capi.load()
    h = C.c_void_p()
    cfg = _cfg()
    assert lib.irs_create(C.byref(cfg), 7, C.byref(h)) == 0
    lay = capi.Layout()
    assert lib.irs_state_layout(h, C.byref(lay)) == 0
    assert lay.row_stride == 404
    offs = [getattr(lay, n) for n in capi._LAYOUT_ARRAYS]
    assert offs == sorted(offs) and all(o % 256 == 0 for o in offs)
    assert lay.buf == 0 and lay.ue_i == 64 * 25 * 404 * 4
    assert lay.total_bytes > lay.metrics_total
    # stepping an unbound handle is an error code, not a crash
    assert lib.irs_step(h, None, None, None, None) == capi.IRS_ERR_UNBOUND
    assert b"irs_bind" in lib.irs_last_error(h)
    assert lib.irs_destroy(h) == 0


@pytest.mark.parametrize("bad", [dict(num_envs=0), dict(channel_mode=4), dict(surface=3), dict(policy=9), dict(fixed_intra=3),
                                 dict(bandwidth=0.0), dict(assoc_mode=1, scen_div=0), dict(depth_cap=16384),
                                 dict(random_episodes=2)])
def test_create_rejects_bad_config(bad):
    lib = capi.load()
    h = C.c_void_p()
    cfg = _cfg(**bad)
    assert lib.irs_create(C.byref(cfg), 7, C.byref(h)) == capi.IRS_ERR_ARG
    assert not h.value


def test_engine_refuses_to_run_without_cuda(bank):
    import torch

    if torch.cuda.is_available():
        pytest.skip("CUDA present")
    from intent_radio_sched_multi_slice_b200.engine import BatchEngine

    with pytest.raises(RuntimeError, match="no CPU path"):
        BatchEngine(bank, 4)


def test_scenario_bank_compiles_the_reference_scenarios(bank):
    arr = bank.array()
    assert arr.shape == (200,) and bank.max_depth() == 401
    n_act = (arr["sl"]["flags"] & scenario.F_ACTIVE != 0).sum(axis=1)
    assert n_act.min() >= 3 and n_act.max() <= 5
    assert ((arr["ue_slice"][:, :25] >= 0).sum(axis=1) == arr["sl"]["n_ues"].sum(axis=1)).all()
    cdf = bank.poisson_cdf()
    assert cdf.shape[1] == 512 and np.all(np.diff(cdf, axis=1) >= 0) and np.allclose(cdf[:, -1], 1.0)


def test_noalloc_entry_point_rejects_bad_sizes_before_touching_the_device():
    """irs_check_errors_noalloc validates its arguments first (no CUDA call is made for a rejected request): null
    arrays, num_envs < 1 and num_envs beyond the static scratch's IRS_NOALLOC_MAX_ENVS, with a message in errbuf."""
    lib = capi.load()
    hdr = open(os.path.join(ROOT, "include", "irs_b200.h")).read()
    assert int(re.search(r"#define IRS_NOALLOC_MAX_ENVS (\d+)", hdr).group(1)) == capi.IRS_NOALLOC_MAX_ENVS
    msg = C.create_string_buffer(256)
    fake = C.c_void_p(256)  # never dereferenced: every call below is rejected before a launch
    assert lib.irs_check_errors_noalloc(None, 16, fake, None, msg, 256) == capi.IRS_ERR_ARG
    assert lib.irs_check_errors_noalloc(fake, 16, None, None, msg, 256) == capi.IRS_ERR_ARG
    assert lib.irs_check_errors_noalloc(fake, 0, fake, None, msg, 256) == capi.IRS_ERR_ARG
    assert lib.irs_check_errors_noalloc(fake, capi.IRS_NOALLOC_MAX_ENVS + 1, fake, None, msg, 256) == capi.IRS_ERR_ARG
    assert b"IRS_NOALLOC_MAX_ENVS" in msg.value and str(capi.IRS_NOALLOC_MAX_ENVS + 1).encode() in msg.value
    assert lib.irs_check_errors_noalloc(fake, capi.IRS_NOALLOC_MAX_ENVS + 1, fake, None, None, 0) == capi.IRS_ERR_ARG
