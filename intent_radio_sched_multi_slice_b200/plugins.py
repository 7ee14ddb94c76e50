"""The four plugin hooks of the reference env, with the reference's class names.

The reference passes plugin CLASSES to the env, which re-instantiates them at every reset
(`sixg_radio_mgmt/sixg_radio_mgmt/comm_env.py:626-681`) and calls their `step` once per TTI:

    Association(ues, max_ues, max_bs, max_slices, rng, root_path, scenario_name)
        .step(bs_ue, bs_slice, slice_ue, slice_req, step, episode) -> 4-tuple   association.py:30-109
    Channel(max_ues, max_bs, num_rbs, rng, root_path, scenario_name)
        .step(step, episode, mobilities, sched=None) -> B x U x R               channel.py:31-92
    Traffic(max_ues, rng, root_path).step(slice_ue_assoc, slice_req, step, episode) -> U bits
                                                                                 traffic.py:23-62
    Mobility(max_ues, rng, root_path).step(step, episode) -> U x 2              mobility.py:23-58

The classes below are the built-ins of the mult_slice scenarios.  They run ON THE DEVICE: the
env recognises them by their `device_mode` attribute and selects the matching kernel mode, so
their Python `step` is never called on the hot path (it exists, with the reference's semantics,
for users who want the values on the host).  Any other class implementing one of the four
interfaces is honoured through the host "tensor hook" path of `BatchedMARLCommEnv` (its output
is uploaded every step - slow, and exactly the path the parity tests use to inject traces).
"""
from __future__ import annotations

import os

import numpy as np

from . import scenario as scn


# ---- channels -----------------------------------------------------------------------------
class MimicQuadriga:
    """channels/mimic_quadriga.py:8-58: per episode mu_u = |N(10, 7.5)|, per step
    SE[u, r] = |N(mu_u, 1.5)|.  Device mode: native counter-based generators (irs_rng.cuh)."""
    device_mode = {"channel_mode": 2}

    def __init__(self, max_number_ues, max_number_basestations, num_available_rbs, rng=np.random.default_rng(),
                 root_path="", scenario_name=""):
        self.max_number_ues, self.max_number_basestations = max_number_ues, max_number_basestations
        self.num_available_rbs, self.rng = num_available_rbs, rng
        self.current_episode, self.ues_mean_se = -1, None

    def step(self, step_number, episode_number, mobilities, sched_decision=None):
        if self.current_episode != episode_number:
            self.current_episode = episode_number
            self.ues_mean_se = np.abs(self.rng.normal(10, 7.5, size=(self.max_number_ues,)))
        se = np.zeros((self.max_number_basestations, self.max_number_ues, int(self.num_available_rbs[0])))
        for u in range(self.max_number_ues):
            se[0, u] = np.abs(self.rng.normal(self.ues_mean_se[u], 1.5, size=se.shape[2]))
        return se


class FixedSE:
    """channels/fixed_se.py:8-41: SE = 2.0 everywhere."""
    device_mode = {"channel_mode": 1}

    def __init__(self, max_number_ues, max_number_basestations, num_available_rbs, rng=np.random.default_rng(),
                 root_path="", scenario_name=""):
        self.shape = (max_number_basestations, max_number_ues, int(num_available_rbs[0]))

    def step(self, step_number, episode_number, mobilities, sched_decision=None):
        return 2.0 * np.ones(self.shape)


class QuadrigaChannel:
    """channels/quadriga.py:8-87: per step, SE = log2(1 + ((100 W / 135) * target_cell_power) / 1e-13)
    from `<root>/mult_slice_channel_generation/results/mult_slice/freq_channel/assoc_A/ep_E/
    target_cell_power.mat`, with (A, E) = (episode, 0).

    Device mode: the files of `episodes` are transformed once (irs_power_to_se) into a float64
    [n_episodes, n_steps, 25, 135] bank that stays in HBM (irs_set_se_trace, channel_mode 3); an env in
    episode `ep` reads bank slot `ep - first_episode` (an episode outside the bank is an error, never another
    episode's file).  Supply the powers either as .mat files under root_path (needs h5py, like the reference) or
    through `register_power(power, first_episode)` with an array [n_episodes, n_steps, 135, 25] (the file order) -
    e.g. synthetic or pre-converted data."""
    device_mode = {"channel_mode": 3}
    transmission_power, thermal_noise_power = 100.0, 10e-14  # quadriga.py:33-34
    channel_eps_per_scenario = 100
    _registered = None
    _registered_first = 0

    def __init__(self, max_number_ues, max_number_basestations, num_available_rbs, rng=np.random.default_rng(),
                 root_path="", scenario_name="mult_slice"):
        self.root_path = root_path
        self.num_available_rbs = num_available_rbs
        self.current_episode_number, self._power = -1, None

    @classmethod
    def register_power(cls, power, first_episode=0):
        """Use an in-memory array [n_episodes, n_steps, 135, 25] instead of .mat files; power[i] is the file of env
        episode first_episode + i."""
        power = np.ascontiguousarray(power, np.float64)
        if power.ndim != 4 or power.shape[2:] != (135, 25):
            raise ValueError("power must be [n_episodes, n_steps, 135, 25] (target_cell_power.mat order)")
        cls._registered = power
        cls._registered_first = int(first_episode)

    @classmethod
    def choose_episode(cls, episode_number):
        """(association, episode) of the file an env episode reads (quadriga.py:79-87)."""
        return episode_number, 0

    @classmethod
    def channels_path(cls, root_path):
        return os.path.join(root_path, "mult_slice_channel_generation", "results", "mult_slice", "freq_channel")

    @classmethod
    def read_file(cls, root_path, episode_number):
        """One file as [n_steps, 135, 25] float64 (h5py order of simu.m:103's array, singletons dropped)."""
        try:
            import h5py
        except ImportError as exc:
            raise RuntimeError("reading target_cell_power.mat needs h5py; use register_power(array) instead") from exc
        a, e = cls.choose_episode(episode_number)
        with h5py.File(os.path.join(cls.channels_path(root_path), f"assoc_{a}", f"ep_{e}", "target_cell_power.mat"), "r") as f:
            return np.asarray(f["target_cell_power"], np.float64)[:, :, 0, 0, :]

    @classmethod
    def power_bank(cls, root_path, episodes):
        """Powers of `episodes` (env episode numbers) as [len(episodes), n_steps, 135, 25]."""
        if cls._registered is not None:
            idx = [int(ep) - cls._registered_first for ep in episodes]
            if min(idx) < 0 or max(idx) >= len(cls._registered):
                raise ValueError("an episode lies outside the registered power array")
            return cls._registered[idx]
        return np.stack([cls.read_file(root_path, ep) for ep in episodes])

    def step(self, step_number, episode_number, mobilities, sched_decision=None):
        """Host-side value with the reference's semantics (not used on the device path)."""
        if episode_number != self.current_episode_number:
            self.current_episode_number = episode_number
            bank = type(self)._registered
            self._power = (bank[episode_number - type(self)._registered_first] if bank is not None
                           else self.read_file(self.root_path, episode_number))
        p = self._power[step_number]
        se = np.log2(1 + np.divide((self.transmission_power / self.num_available_rbs[0]) * p,
                                   (np.zeros_like(p) + self.thermal_noise_power)))
        return np.array([se.transpose()])


class QuadrigaChannelSeq(QuadrigaChannel):
    """channels/quadriga_seq.py:8-39: (association, episode) = (episode // 100, episode % 100)."""
    _registered = None
    _registered_first = 0

    @classmethod
    def choose_episode(cls, episode_number):
        return episode_number // cls.channel_eps_per_scenario, episode_number % cls.channel_eps_per_scenario


# ---- traffic --------------------------------------------------------------------------------
class MultSliceTraffic:
    """traffics/mult_slice.py:8-34: Poisson(slice traffic in Mbps) * 1e6 bits per UE of every
    slice with requirements.  Device mode: native Poisson by table inversion."""
    device_mode = {"traffic_mode": 1}

    def __init__(self, max_number_ues, rng=np.random.default_rng(), root_path=""):
        self.max_number_ues, self.rng = max_number_ues, rng

    def step(self, slice_ue_assoc, slice_req, step_number, episode_number):
        out = np.zeros(self.max_number_ues)
        for s, name in enumerate(slice_req):
            if slice_req[name] == {}:
                continue
            idx = slice_ue_assoc[s].nonzero()[0]
            out[idx] = self.rng.poisson(slice_req[name]["ues"]["traffic"], len(idx)) * 1e6
        return out


# ---- mobility -------------------------------------------------------------------------------
class SimpleMobility:
    """mobilities/simple.py:8-16: constant positions; unused by the synthetic channels."""
    device_mode = {}

    def __init__(self, max_number_ues, rng=np.random.default_rng(), root_path=""):
        self.max_number_ues = max_number_ues

    def step(self, step_number, episode_number):
        return np.ones((self.max_number_ues, 2))


# ---- associations -----------------------------------------------------------------------------
def load_reference_episodes(data_dir: str, limit: int = 200) -> scn.ScenarioBank | None:
    """Compile the reference's ep_N.npz files (gen_assoc_mult_slice.py:229-237) into a bank.
    Returns None when the directory holds none (they are an external download, README.md:35)."""
    bank = scn.ScenarioBank()
    for n in range(limit):
        path = os.path.join(data_dir, f"ep_{n}.npz")
        if not os.path.exists(path):
            break
        f = np.load(path, allow_pickle=True)
        req = dict(f["hist_slice_req"][0])
        for k, v in req.items():  # operators are stored as numpy ufuncs; the bank wants names
            for par in (v.get("parameters", {}) if v else {}).values():
                par["operator"] = scn._operator_name(par["operator"])
        bank.add(req, f["hist_slice_ue_assoc"][0], f["hist_basestation_slice_assoc"][0])
    return bank if bank.rows else None


class MultSliceAssociation:
    """associations/mult_slice.py:8-508 in replay mode: network scenario = episode % 200
    (`choose_episode` :444-452), constant within an episode.  Device mode: the scenario bank is
    uploaded once and the rollover kernel indexes it."""
    device_mode = {"assoc_mode": 0, "scen_mod": 200}
    data_subdir = os.path.join("associations", "data", "mult_slice")

    @classmethod
    def scenario_bank(cls, root_path: str, n: int = 200, seed: int = 10) -> scn.ScenarioBank:
        bank = load_reference_episodes(os.path.join(root_path, cls.data_subdir), n)
        if bank is None:  # same distribution, generated natively (scenario.generate_scenarios)
            st, us = scn.generate_scenarios(n, seed)
            bank = scn.ScenarioBank()
            for i in range(n):
                bank.add_typed(st[i], us[i])
        return bank


class MultSliceAssociationSeq(MultSliceAssociation):
    """associations/mult_slice_seq.py:8-46: scenario = episode // 100."""
    device_mode = {"assoc_mode": 1, "scen_div": 100}


class MultSliceAssociationGen(MultSliceAssociation):
    """associations/mult_slice.py in GENERATOR mode (:359-416): a new network scenario for every env and episode -
    3..5 slices, slots and slice types without replacement, UEs per slice uniform in the type's [min, max], random
    UE ids - drawn on the device by the counter-based generator at every reset (assoc_mode 2).  No ep_N.npz files."""
    device_mode = {"assoc_mode": 2}

    @classmethod
    def scenario_bank(cls, root_path: str = "", n: int = 0, seed: int = 10) -> scn.ScenarioBank:
        return scn.ScenarioBank.templates()


BUILTIN = {c.__name__: c for c in (MultSliceAssociationGen, MimicQuadriga, FixedSE, QuadrigaChannel, QuadrigaChannelSeq, MultSliceTraffic, SimpleMobility, MultSliceAssociation,
                                   MultSliceAssociationSeq)}


def device_mode_of(cls) -> dict | None:
    """Kernel mode of a plugin class: ours, or a reference class of the same name (so the
    reference's own `channels.mimic_quadriga.MimicQuadriga` etc. can be passed unchanged)."""
    if hasattr(cls, "device_mode"):
        return dict(cls.device_mode)
    twin = BUILTIN.get(getattr(cls, "__name__", ""))
    return dict(twin.device_mode) if twin is not None else None
