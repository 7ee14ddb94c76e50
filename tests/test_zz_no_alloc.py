"""irs_check_errors_noalloc (include/irs_b200.h): the error scan of irs_check_errors with its scratch in a static device
array of the library, for callers that cannot allocate.  Called through the C-ABI and compared with irs_check_errors on
the same error words, two calls in a row (the static scratch keeps the first call's contents).  Collected last on
purpose."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _noalloc(lib, err, n, out):
    import torch

    msg = C.create_string_buffer(256)
    rc = lib.irs_check_errors_noalloc(err.data_ptr(), n, out.data_ptr(), torch.cuda.current_stream().cuda_stream, msg, 256)
    return rc, msg.value.decode()


def _engine(bank, E):
    import torch

    from intent_radio_sched_multi_slice_b200.engine import BatchEngine

    eng = BatchEngine(bank, E, device=torch.device("cuda", 0), depth_cap=max(bank.max_depth(), 101), channel_mode=2,
                      traffic_mode=1, sort_slices=1, fixed_intra=-1, policy=0, assoc_mode=0, scen_mod=200, max_steps=1000,
                      max_episodes=10000, auto_reset=1, seed=3, env_id_base=0)
    eng.reset(initial_episode=0)
    return eng


def test_matches_irs_check_errors_two_calls_in_a_row(bank):
    import torch

    from intent_radio_sched_multi_slice_b200 import capi

    E = 3000
    eng = _engine(bank, E)
    lib = capi.load()
    dev = eng.device
    out = torch.full((2,), 12345, dtype=torch.int32, device=dev)
    g = torch.Generator(device=dev)
    g.manual_seed(5)
    a0 = torch.rand(E, 5, generator=g, device=dev, dtype=torch.float64) * 2 - 1
    intra = torch.randint(0, 3, (E, 5), generator=g, device=dev, dtype=torch.int32)

    # call 1: a clean step, then a step in which envs 1777 and 2500 pass an intra-slice algorithm that does not exist
    # (ValueError of agents/ib_sched.py:341-344 -> error bit 1)
    eng.step(a0, intra)
    assert eng.check_errors() is None
    rc, msg = _noalloc(lib, eng.v["err"], E, out)
    assert rc == 0, msg
    assert out.cpu().tolist() == [-1, 0]
    bad = intra.clone()
    bad[1777] = 3
    bad[2500] = 3
    eng.step(a0, bad)
    want = eng.check_errors()
    assert want is not None and want[0] == 1777 and want[1] != 0
    rc, msg = _noalloc(lib, eng.v["err"], E, out)
    assert rc == 0, msg
    assert tuple(out.cpu().tolist()) == want

    # call 2, right after: other error words in other blocks of the scan (the scratch still holds call 1's slots)
    eng.v["err"].zero_()
    eng.v["err"][2999] = 4
    want = eng.check_errors()
    assert want == (2999, 4)
    rc, msg = _noalloc(lib, eng.v["err"], E, out)
    assert rc == 0, msg
    assert tuple(out.cpu().tolist()) == want

    # and none at all again
    eng.v["err"].zero_()
    assert eng.check_errors() is None
    rc, msg = _noalloc(lib, eng.v["err"], E, out)
    assert rc == 0, msg
    assert out.cpu().tolist() == [-1, 0]


def test_any_error_array_and_the_size_limit():
    """err_dev is just an i32 array: sizes that are not a multiple of the 256-env block, a large scan followed by a
    small one (stale slots beyond the small scan's range must not be read), and the stated maximum."""
    import torch

    from intent_radio_sched_multi_slice_b200 import capi

    lib = capi.load()
    dev = torch.device("cuda", 0)
    out = torch.zeros(2, dtype=torch.int32, device=dev)
    rng = np.random.default_rng(11)
    with torch.cuda.device(dev):
        for n, n_bad in [(capi.IRS_NOALLOC_MAX_ENVS, 5), (70001, 3), (257, 1), (1, 1), (255, 0), (70001, 0)]:
            err = np.zeros(n, np.int32)
            pos = np.sort(rng.choice(n, n_bad, replace=False)) if n_bad else np.zeros(0, np.int64)
            if n == 257:
                pos = np.array([256])
            err[pos] = rng.integers(1, 32, len(pos))
            err_t = torch.from_numpy(err).to(dev)
            for _ in range(2):
                rc, msg = _noalloc(lib, err_t, n, out)
                assert rc == 0, msg
                want = [int(pos[0]), int(err[pos[0]])] if n_bad else [-1, 0]
                assert out.cpu().tolist() == want, (n, n_bad)
        big = torch.zeros(8, dtype=torch.int32, device=dev)
        rc, msg = _noalloc(lib, big, capi.IRS_NOALLOC_MAX_ENVS + 1, out)
        assert rc == capi.IRS_ERR_ARG and "IRS_NOALLOC_MAX_ENVS" in msg
        rc, msg = _noalloc(lib, big, 0, out)
        assert rc == capi.IRS_ERR_ARG
