"""CPU suite: the native (counter-based) generators as restated in the oracle - the 16-bit inverse-CDF normal
(csrc/irs_rng.cuh normal_from, table from scripts/gen_normal_table.py), the Philox words feeding it, and the
table-free episode permutation of enable_random_episodes (comm_env.py:219-226, 434-444).  The CUDA side is
compared with the oracle bit for bit in the GPU suite; here the oracle's generators are checked against the
distributions of the reference (channels/mimic_quadriga.py:37-56: N(mu, 1.5); np_random.permutation)."""
import numpy as np
import pytest
from scipy import stats
from scipy.special import ndtri

from oracle import oracle


def test_normal16_is_the_quantile_of_its_cell():
    z = np.array([oracle.normal16(h) for h in range(65536)], np.float64)
    assert np.array_equal(z[:32768], -z[32768:])                       # sign bit
    assert (np.diff(z[:32768]) > 0).all()                              # monotone in the magnitude word
    t = 32768 - np.arange(32768)
    want = ndtri(1.0 - (t - 0.5) / 65536.0)                            # P(|Z| > z) = (t - 0.5) / 32768
    assert np.abs(z[:32768] - want).max() < 1e-4
    assert abs(z.mean()) < 1e-12 and abs(z.var() - 1.0) < 1e-4 and 4.2 < z.max() < 4.4


def test_normal16_over_philox_words_is_standard_normal():
    """Kolmogorov-Smirnov against N(0, 1) over 2^20 draws (critical value at alpha = 0.001: 1.9e-3)."""
    words = np.concatenate([oracle.philox([i, 7, 3, 1], 0x1234, 0x9abc) for i in range(1 << 17)])
    halves = np.concatenate([words & 0xFFFF, words >> 16])
    lut = np.array([oracle.normal16(h) for h in range(65536)])
    d = stats.kstest(lut[halves], "norm").statistic
    assert d < 1.9e-3, d
    # the half-words themselves are uniform
    assert stats.kstest((halves + 0.5) / 65536.0, "uniform").statistic < 1.9e-3


@pytest.mark.parametrize("n", [1, 2, 3, 9, 16, 17, 187, 200, 1000, 9997])
def test_feistel_permutation_is_a_bijection(n):
    rng = np.random.default_rng(n)
    for _ in range(3):
        keys = rng.integers(0, 2 ** 32, 4, dtype=np.uint64).astype(np.uint32)
        p = [oracle.feistel_perm(i, n, keys) for i in range(n)]
        assert sorted(p) == list(range(n))


def test_feistel_permutations_look_uniform():
    """Over many keys every element lands on every position about equally often (chi-square, n = 12)."""
    n, trials = 12, 6000
    rng = np.random.default_rng(5)
    counts = np.zeros((n, n))
    for _ in range(trials):
        keys = rng.integers(0, 2 ** 32, 4, dtype=np.uint64).astype(np.uint32)
        for i in range(n):
            counts[i, oracle.feistel_perm(i, n, keys)] += 1
    chi2 = ((counts - trials / n) ** 2 / (trials / n)).sum()
    assert chi2 < stats.chi2.ppf(0.9999, (n - 1) * (n - 1)), chi2


def test_random_episode_cycle_in_the_oracle(bank):
    """enable_random_episodes with the built-in permutation: every episode of [initial, max) once per cycle, the same
    order in the next cycle (itertools.cycle), a new permutation when a reset passes an initial episode."""
    E = 6
    env = oracle.OracleEnv(bank.array(), bank.poisson_cdf(), E, channel_mode=2, traffic_mode=1, policy=1, fixed_intra=0,
                           sort_slices=0, max_steps=2, max_episodes=11, initial_episode=4, auto_reset=1, random_episodes=1,
                           seed=5)
    env.reset()
    seen = [env.state()["episode"].copy()]
    for k in range(2 * 7 * 2):
        env.step()
        if k % 2 == 1:
            seen.append(env.state()["episode"].copy())
    seen = np.array(seen)                          # [1 + 14, E]
    for e in range(E):
        assert sorted(seen[:7, e]) == list(range(4, 11))
        assert np.array_equal(seen[:7, e], seen[7:14, e])
    assert len({tuple(seen[:7, e]) for e in range(E)}) > 1          # envs do not share an order
    env.reset(initial_episode=8)                   # comm_env.py:435-442
    again = [env.state()["episode"].copy()]
    for k in range(3 * 2):
        env.step()
        if k % 2 == 1:
            again.append(env.state()["episode"].copy())
    again = np.array(again)
    for e in range(E):
        assert sorted(again[:3, e]) == [8, 9, 10] and again[3, e] == again[0, e]


def test_host_mirror_of_the_episode_permutation(bank):
    """scenario.shuffled_episode (used by the env when a HOST channel plugin must be asked for the next episode's reset
    frame before the device has moved on) equals the permutation the oracle / the kernels play."""
    from intent_radio_sched_multi_slice_b200 import scenario as scn

    for c, k0, k1 in [([1, 2, 3, 4], 5, 6), ([0xFFFFFFFF, 0, 7, 99], 0x12345678, 0x9ABCDEF0)]:
        assert scn._philox4x32_7(c, k0, k1) == [int(x) for x in oracle.philox(c, k0, k1)]
    E, seed = 5, (7 << 32) + 11
    env = oracle.OracleEnv(bank.array(), bank.poisson_cdf(), E, channel_mode=1, traffic_mode=1, policy=1, fixed_intra=0,
                           sort_slices=0, max_steps=1, max_episodes=23, initial_episode=4, auto_reset=1, random_episodes=1,
                           seed=seed, env_id_base=40)
    env.reset()
    for k in range(25):
        ep = env.state()["episode"]
        for e in range(E):
            assert ep[e] == scn.shuffled_episode(seed, 40 + e, 0, 4, 19, k % 19), (k, e)
        env.step()
