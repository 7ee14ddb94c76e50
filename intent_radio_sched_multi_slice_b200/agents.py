"""Agent surfaces with the reference's names.

In the reference an agent object hands three callbacks and two spaces to the env
(`simu.py:405-418` -> `CommunicationEnv.set_agent_functions`, comm_env.py:484-512):
`obs_space_format`, `action_format`, `calculate_reward`.  For the agents of the mult_slice
scenarios all three are the IBSched adapters (agents/ib_sched.py:63-349), differing only in
  * whether slices are sorted (IBSched.enable_sort_slices, ib_sched.py:57; MARR/MAPF switch it
    off on their `fake_agent`, marr.py:29-32, mapf.py:30-33),
  * the intra-slice scheduler: per-slice action (IBSched), fixed "rr" (MARR, SB3) or "pf" (MAPF)
    (marr.py:62-70, mapf.py:126-134, sb3_sched.py:164-177),
  * the shape of what is returned: the full dict (IBSched) or only player_0's 50-vector and a
    scalar reward (marr.py:52-60, mapf.py:116-124, sb3_sched.py:159-187).
On the device these are kernel modes, so the classes below carry the mode and the spaces; their
bound methods are accepted by `set_agent_functions` as selectors of the built-in surface.
"""
from __future__ import annotations

from dataclasses import dataclass, field

import numpy as np

try:  # gymnasium is optional (absent in the build image)
    from gymnasium import spaces as _spaces
except Exception:  # pragma: no cover - depends on the environment
    _spaces = None


@dataclass
class Box:
    low: float
    high: float
    shape: tuple
    dtype: type = np.float64


@dataclass
class Discrete:
    n: int


@dataclass
class Dict:
    spaces: dict = field(default_factory=dict)

    def __getitem__(self, k):
        return self.spaces[k]


def _box(low, high, shape, dtype):
    return _spaces.Box(low=low, high=high, shape=shape, dtype=dtype) if _spaces else Box(low, high, shape, dtype)


def _discrete(n):
    return _spaces.Discrete(n) if _spaces else Discrete(n)


def _dict(d):
    return _spaces.Dict(d) if _spaces else Dict(d)


class DevicePolicy:
    """Returned by MARR.step / MAPF.step: "compute the action on the device" (the policy is a
    few arithmetic operations on state that already lives in HBM)."""

    def __repr__(self):
        return "<device policy>"


DEVICE_POLICY = DevicePolicy()


class _AgentBase:
    sort_slices, fixed_intra, policy, flat = 1, -1, 0, False
    surface = 0  # irs_config.surface: which observation / reward kernel fills obs0 and reward[0]
    # ib_sched.py:49-58
    max_obs_memory, max_number_ues_slice, intent_overfulfillment_rate, rbs_per_rbg = 10, 5, 0.2, 5

    def __init__(self, env=None, max_number_ues=25, max_number_slices=5, max_number_basestations=1,
                 num_available_rbs=np.array([135]), **kw):
        self.env = env
        self.max_number_ues, self.max_number_slices = max_number_ues, max_number_slices
        self.max_number_basestations, self.num_available_rbs = max_number_basestations, num_available_rbs
        if (max_number_ues, max_number_slices, max_number_basestations, int(np.asarray(num_available_rbs)[0])) != (25, 5, 1, 135):
            raise ValueError("the device kernels are built for the mult_slice shape: 25 UEs, 5 slices, 1 BS, 135 RBs")

    # the three callbacks: selectors, never executed
    def obs_space_format(self, obs_space):
        raise RuntimeError("built-in surface: evaluated by the obs/reward kernel, not on the host")

    def action_format(self, action):
        raise RuntimeError("built-in surface: evaluated by the action kernel, not on the host")

    def calculate_reward(self, obs_space):
        raise RuntimeError("built-in surface: evaluated by the obs/reward kernel, not on the host")

    def get_action_space(self):
        if self.flat:  # marr.py:72-76
            return _box(-1, 1, (self.max_number_slices,), np.float64)
        d = {"player_0": _box(-1, 1, (self.max_number_slices,), np.float64)}  # ib_sched.py:394-411
        for s in range(1, self.max_number_slices + 1):
            d[f"player_{s}"] = _discrete(3)
        return _dict(d)

    def get_obs_space(self):
        inter = _box(-1, np.inf, (self.max_number_slices * 10,), np.float64)
        if self.flat:
            return inter
        mask = lambda n: _box(0, 1, (n,), np.int8)
        d = {"player_0": _dict({"observations": inter, "action_mask": mask(self.max_number_slices)})}
        for s in range(1, self.max_number_slices + 1):  # ib_sched.py:413-470
            d[f"player_{s}"] = _dict({"observations": _box(-1, np.inf, (19,), np.float64),
                                      "action_mask": mask(self.max_number_ues_slice)})
        return _dict(d)


class IBSched(_AgentBase):
    """agents/ib_sched.py:21-473: dict observation with action masks, six rewards, per-slice
    intra-slice scheduler chosen by player_1..5."""

    def step(self, obs):
        raise NotImplementedError("IBSched actions come from the RL policy (agents/ray_agent.py)")


class MARR(_AgentBase):
    """agents/marr.py:12-82: equal inter-slice scores, round-robin inside slices."""
    sort_slices, fixed_intra, policy, flat = 0, 0, 1, True

    def step(self, obs):
        return DEVICE_POLICY


class MAPF(_AgentBase):
    """agents/mapf.py:13-146: proportional-fair inter-slice scores and intra-slice scheduler."""
    sort_slices, fixed_intra, policy, flat = 0, 1, 2, True

    def step(self, obs):
        return DEVICE_POLICY


class IBSchedSB3(_AgentBase):
    """agents/sb3_sched.py:15-200 env adapters: flat player_0 observation, scalar reward,
    round-robin intra-slice scheduler, slices sorted."""
    sort_slices, fixed_intra, policy, flat = 1, 0, 0, True


class IBSchedSB3PF(IBSchedSB3):
    """agents/sb3_pf_sched.py: same with proportional-fair intra-slice scheduling."""
    fixed_intra = 1


class SchedTWC(_AgentBase):
    """agents/sched_twc.py:20-449 env adapters: 50-value per-slice metric vector (requirements, mean SE, served and
    effective throughput, buffer occupancy and latency, packet-loss rate, requested throughput), reward = weighted
    mean of the negative slice intent drifts, round-robin inside slices, slices not sorted, every frame pushed twice
    into the observation window (:176-179)."""
    sort_slices, fixed_intra, policy, flat, surface = 0, 0, 0, True, 1

    def get_obs_space(self):  # sched_twc.py:425-428: declared with the fake agent's player_0 shape (50,)
        return _box(-1, np.inf, (self.max_number_slices * 10,), np.float64)


class SchedColORAN(SchedTWC):
    """agents/sched_colran.py:20-455: the same observation; reward = sum over slices of +throughput / 200 Mbps for eMBB
    slice types and -buffered Mbps / 2000 for URLLC ones (:347-417)."""
    surface = 2


_BY_NAME = {c.__name__: c for c in (IBSched, MARR, MAPF, IBSchedSB3, IBSchedSB3PF, SchedTWC, SchedColORAN)}


def surface_of(callback):
    """Map a callback given to set_agent_functions to a built-in surface class: a bound method
    of one of our agents, or of a reference agent of the same class name."""
    owner = getattr(callback, "__self__", None)
    if owner is None:
        return None
    for klass in type(owner).__mro__:
        if klass.__name__ in _BY_NAME:
            return _BY_NAME[klass.__name__]
    return None
