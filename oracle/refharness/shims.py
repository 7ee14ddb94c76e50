"""Import shims that let the UNMODIFIED reference run in this container.

TEST INFRASTRUCTURE ONLY.  Used by `oracle/refharness/gen_golden.py` (which runs
only in the build container where `/root/reference` is mounted) to record golden
traces.  Nothing in the product package imports this file.

The reference imports `gymnasium`, `pettingzoo`, `ray`, `matplotlib`, none of which
are installed here; none of them takes part in the arithmetic of the hot path
(SURVEY.md 8c).  The shims below only provide the names the reference touches at
import / construction time:

* `gymnasium.Env.reset(seed)` seeds `np_random = Generator(PCG64(SeedSequence(seed)))`
  exactly like gymnasium 0.28.1 `seeding.np_random` (used at
  `sixg_radio_mgmt/sixg_radio_mgmt/comm_env.py:413,631`).
* `gymnasium.spaces.{Space,Box,Dict,Discrete}` are inert containers
  (`agents/ib_sched.py:394-470`).
* `ray.rllib.env.multi_agent_env.MultiAgentEnv` is an empty base
  (`sixg_radio_mgmt/sixg_radio_mgmt/marl_comm_env.py:7,29`).

The reference's `agents/`, `channels/`, ... directories have no `__init__.py` and
this image ships an unrelated `site-packages/agents`; they are pre-registered as
path-only packages.
"""
import sys
import types

import numpy as np

REF = "/root/reference"


def _mod(name, **attrs):
    m = types.ModuleType(name)
    m.__dict__.update(attrs)
    sys.modules[name] = m
    return m


class _Space:
    def __init__(self, *a, **k):
        self.args, self.kwargs = a, k
        self.shape = k.get("shape")
        self.dtype = k.get("dtype")


class _Box(_Space):
    def __init__(self, low=None, high=None, shape=None, dtype=None, **k):
        super().__init__(low=low, high=high, shape=shape, dtype=dtype)
        self.low, self.high = low, high


class _Dict(_Space):
    def __init__(self, spaces=None, **k):
        super().__init__()
        self.spaces = dict(spaces or {})

    def __getitem__(self, k):
        return self.spaces[k]


class _Discrete(_Space):
    def __init__(self, n, **k):
        super().__init__()
        self.n = n


class _Env:
    """gymnasium.Env stand-in: only the seeding contract matters."""

    _np_random = None

    @property
    def np_random(self):
        if self._np_random is None:
            self._np_random = np.random.Generator(np.random.PCG64(np.random.SeedSequence()))
        return self._np_random

    @np_random.setter
    def np_random(self, v):
        self._np_random = v

    def reset(self, *, seed=None, options=None):
        if seed is not None:
            self._np_random = np.random.Generator(np.random.PCG64(np.random.SeedSequence(seed)))


def install():
    if "gymnasium" in sys.modules and getattr(sys.modules["gymnasium"], "_irs_shim", False):
        return
    spaces = _mod("gymnasium.spaces", Space=_Space, Box=_Box, Dict=_Dict, Discrete=_Discrete)
    g = _mod("gymnasium", Env=_Env, spaces=spaces, _irs_shim=True)
    g.Space = _Space

    class _AEC:
        pass

    pz_utils = _mod("pettingzoo.utils", agent_selector=object, wrappers=types.SimpleNamespace())
    _mod("pettingzoo", AECEnv=_AEC, utils=pz_utils)

    class _MAE:
        def __init__(self, *a, **k):
            pass

    _mod("ray")
    _mod("ray.rllib")
    _mod("ray.rllib.env")
    _mod("ray.rllib.env.multi_agent_env", MultiAgentEnv=_MAE)

    class _Any:
        def __getattr__(self, k):
            return _Any()

        def __call__(self, *a, **k):
            return _Any()

    for name in ("matplotlib", "matplotlib.figure", "matplotlib.pyplot", "matplotlib.pylab"):
        m = _mod(name)
        m.__getattr__ = lambda k: _Any()  # type: ignore
    _install_h5py()
    _install_sb3()
    for pkg in ("agents", "channels", "traffics", "associations", "mobilities"):
        m = types.ModuleType(pkg)
        m.__path__ = [f"{REF}/{pkg}"]
        sys.modules[pkg] = m
    for p in (REF, f"{REF}/sixg_radio_mgmt"):
        if p not in sys.path:
            sys.path.insert(0, p)


def _install_sb3():
    """stable_baselines3 is absent; agents/sched_twc.py and agents/sched_colran.py import PPO / SAC and callback
    classes at module level but their env adapters (obs_space_format, action_format, calculate_reward) never touch
    them: inert stand-ins."""
    class _Inert:
        def __init__(self, *a, **k):
            pass

    class _Event(_Inert):
        pass

    cb = _mod("stable_baselines3.common.callbacks", CheckpointCallback=_Inert, BaseCallback=_Inert, EvalCallback=_Inert,
              EventCallback=_Event)
    ev = _mod("stable_baselines3.common.evaluation", evaluate_policy=lambda *a, **k: None)
    ve = _mod("stable_baselines3.common.vec_env", DummyVecEnv=_Inert, VecEnv=_Inert, sync_envs_normalization=lambda *a, **k: None)
    common = _mod("stable_baselines3.common", callbacks=cb, evaluation=ev, vec_env=ve)
    ppo = _mod("stable_baselines3.ppo.ppo", PPO=_Inert)
    sac = _mod("stable_baselines3.sac.sac", SAC=_Inert)
    _mod("stable_baselines3.ppo", ppo=ppo, PPO=_Inert)
    _mod("stable_baselines3.sac", sac=sac, SAC=_Inert)
    _mod("stable_baselines3", common=common, PPO=_Inert, SAC=_Inert)


QUADRIGA_STEPS = {"n": 1000}  # frames per synthetic target_cell_power.mat


def _install_h5py():
    """h5py is absent and the QuaDRiGa channel files are an external MATLAB product.  The stand-in
    serves the synthetic recipe of tests/golden_inputs.quadriga_power for the path the reference
    opens (channels/quadriga.py:52-55: .../assoc_A/ep_E/target_cell_power.mat); the reference's
    own indexing, transform and transposition (quadriga.py:56-77) run unmodified on it."""
    import os
    import re

    sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))), "tests"))

    class File:
        def __init__(self, path, mode="r"):
            import golden_inputs

            m = re.search(r"assoc_(\d+)/ep_(\d+)/target_cell_power.mat$", path)
            assert m, path
            self.data = golden_inputs.quadriga_power(int(m.group(1)), int(m.group(2)), QUADRIGA_STEPS["n"])

        def get(self, name):
            assert name == "target_cell_power"
            return self.data

        def close(self):
            self.data = None

    _mod("h5py", File=File)


_orig_argsort = np.argsort


def pin_stable_argsort():
    """SURVEY.md section 7 hard part 1: the reference's np.argsort tie order is build
    dependent (AVX-512 quicksort); the oracle pins kind='stable'."""

    def stable(a, axis=-1, kind=None, order=None, **kw):
        return _orig_argsort(a, axis=axis, kind="stable", order=order, **kw)

    np.argsort = stable


def unpin_argsort():
    np.argsort = _orig_argsort
