"""BatchedMARLCommEnv: the reference's env surface over the CUDA engine.

Mirrors `MARLCommEnv` / `CommunicationEnv` (sixg_radio_mgmt/sixg_radio_mgmt/marl_comm_env.py:12-90,
comm_env.py:79-104, 271-512): same constructor arguments and meaning, `reset(seed, options)`,
`step(action)`, `set_agent_functions(...)`, the plugin classes passed by type.  With
`num_envs=1` the returns have exactly the reference's structure (dict keys, dict order, numpy
dtypes, the `__all__` entry); with `num_envs>1` the same fields come back as device tensors
with a leading env axis and finished envs are reset in the same call (VecEnv convention).
"""
from __future__ import annotations

import os
import warnings

import numpy as np
import torch
import yaml

from . import agents as agents_mod
from . import plugins
from .engine import BatchEngine

HERE = os.path.dirname(os.path.abspath(__file__))


class BatchedMARLCommEnv:
    metadata = {"render_modes": []}

    def __init__(self, ChannelClass=plugins.MimicQuadriga, TrafficClass=plugins.MultSliceTraffic,
                 MobilityClass=plugins.SimpleMobility, AssociationClass=plugins.MultSliceAssociation,
                 config_file="mult_slice", agent_name="agent", seed=10, action_format=None, obs_space_format=None,
                 calculate_reward=None, obs_space=None, action_space=None, debug=True, root_path=".",
                 initial_episode_number=0, max_episode_number=None, simu_name=None, save_hist=True,
                 enable_random_episodes=False, *, num_envs=1, device="cuda:0", scenario_bank=None,
                 auto_reset=None, env_id_base=0, channel_episodes=None, hist_envs=None, hist_with_se=False,
                 hist_root_path=None):
        path = os.path.join(root_path, "env_config", f"{config_file}.yml")
        if not os.path.exists(path):
            path = os.path.join(HERE, "env_config", f"{config_file}.yml")
        with open(path) as fh:  # comm_env.py:139-216
            data = yaml.safe_load(fh)
        bs = data["basestations"]
        self.max_number_basestations = bs["max_number_basestations"]
        self.max_number_slices = data["slices"]["max_number_slices"]
        self.max_number_ues = data["ues"]["max_number_ues"]
        self.bandwidths = np.array(bs["bandwidths"])
        self.num_available_rbs = np.array(bs["num_available_rbs"])
        if (self.max_number_basestations, self.max_number_slices, self.max_number_ues,
                int(self.num_available_rbs[0])) != (1, 5, 25, 135):
            raise ValueError("the device kernels are built for 1 base station, 5 slices, 25 UEs, 135 RBs")
        self.max_number_steps = data["simulation"]["max_number_steps"]
        self.max_number_episodes = (data["simulation"]["max_number_episodes"] if max_episode_number is None
                                    else max_episode_number)
        self.simu_name = simu_name if simu_name is not None else data["simulation"]["simu_name"]
        self.root_path, self.agent_name, self.seed, self.debug = root_path, agent_name, seed, debug
        self.initial_episode_number = initial_episode_number
        if enable_random_episodes:
            raise NotImplementedError("enable_random_episodes (comm_env.py:219-226) is not built; pass explicit "
                                      "initial episodes per env instead")
        # save_hist (comm_env.py:345-350): hist/<simu>/<agent>/ep_N.npz per finished episode, for the envs in
        # `hist_envs` (default: the single env of a MARLCommEnv; nothing for a batch unless asked)
        self.save_hist = bool(save_hist)
        self.hist_envs = hist_envs
        self.hist_with_se = bool(hist_with_se)
        self.hist_root_path = hist_root_path if hist_root_path is not None else \
            data["simulation"].get("hist_root_path", "./")  # comm_env.py:211,682
        self._recorder = None
        self.num_envs, self.device = int(num_envs), torch.device(device)
        self.ChannelClass, self.TrafficClass = ChannelClass, TrafficClass
        self.MobilityClass, self.AssociationClass = MobilityClass, AssociationClass
        self._modes = {}
        self._host_channel = self._host_traffic = None
        for cls, kind in ((ChannelClass, "channel"), (TrafficClass, "traffic"), (AssociationClass, "assoc")):
            mode = plugins.device_mode_of(cls)
            if mode is None:
                if kind == "assoc":
                    raise NotImplementedError(f"{cls.__name__}: pass scenario_bank= for a custom association")
                mode = {f"{kind}_mode": 0}  # host tensor hook: the class is called every step
                setattr(self, f"_host_{kind}", cls)
            self._modes.update(mode)
        if scenario_bank is None:
            acls = AssociationClass if hasattr(AssociationClass, "scenario_bank") else plugins.MultSliceAssociation
            scenario_bank = acls.scenario_bank(root_path)
        self.bank = scenario_bank
        self.auto_reset = (self.num_envs > 1) if auto_reset is None else bool(auto_reset)
        self.env_id_base = env_id_base
        self.channel_episodes = channel_episodes  # channel_mode 3: env episodes whose files form the SE bank
        self.agents = {f"player_{i}" for i in range(self.max_number_slices + 1)}  # marl_comm_env.py:20-24
        self._agent_ids = set(self.agents)
        self.comm_env = self  # the reference exposes the wrapped CommunicationEnv here
        self._surface = agents_mod.IBSched
        self.observation_space, self.action_space = obs_space, action_space
        self.eng = None
        self._first_reset = True
        self._np_rngs = None
        if obs_space_format is not None:
            self.set_agent_functions(obs_space_format, action_format, calculate_reward, obs_space, action_space)

    # ---- comm_env.py:484-512 ------------------------------------------------------------------
    def set_agent_functions(self, obs_space_format, action_format, calculate_reward, obs_space=None,
                            action_space=None):
        surf = {agents_mod.surface_of(cb) for cb in (obs_space_format, action_format, calculate_reward)}
        if len(surf) != 1 or None in surf:
            raise NotImplementedError(
                "set_agent_functions: only the built-in agent surfaces (IBSched, MARR, MAPF, IBSchedSB3[PF]) run "
                "on the device; pass the bound methods of one of those agent objects")
        self._surface = surf.pop()
        self.observation_space = obs_space if obs_space is not None else self._surface().get_obs_space()
        self.action_space = action_space if action_space is not None else self._surface().get_action_space()
        if self.eng is not None:
            self.eng.close()
            self.eng = None
            self._recorder = None

    def _engine(self):
        if self.eng is None:
            s = self._surface
            kw = dict(channel_mode=2, traffic_mode=1, assoc_mode=0, scen_mod=200, scen_div=100)
            kw.update(self._modes)
            self.eng = BatchEngine(self.bank, self.num_envs, device=self.device, sort_slices=s.sort_slices,
                                   fixed_intra=s.fixed_intra, policy=s.policy, surface=s.surface, max_steps=self.max_number_steps,
                                   max_episodes=self.max_number_episodes, initial_episode=self.initial_episode_number,
                                   auto_reset=int(self.auto_reset), env_id_base=self.env_id_base, seed=self.seed,
                                   bandwidth=float(self.bandwidths[0]), **kw)
            if kw["channel_mode"] == 3:  # QuadrigaChannel[Seq]: transform the power files once, keep SE in HBM
                ch = self.ChannelClass if hasattr(self.ChannelClass, "power_bank") else \
                    plugins.BUILTIN[self.ChannelClass.__name__]
                eps = self.channel_episodes if self.channel_episodes is not None else \
                    range(self.initial_episode_number, self.initial_episode_number + 1)
                power = torch.from_numpy(ch.power_bank(self.root_path, eps)).to(self.device)
                self.eng.set_se_trace(self.eng.power_to_se(power, layout=0, tx_power=ch.transmission_power,
                                                           noise_power=ch.thermal_noise_power))
        return self.eng

    # ---- host tensor hooks (slow path; any Channel/Traffic class) -------------------------------
    def _hook_frames(self, reset_mask=None):
        """Call user plugin classes on the host for every env (comm_env.py:302-311, 449-473)."""
        eng = self._engine()
        se = tr = None
        if self._host_channel is None and self._host_traffic is None:
            return None, None
        step, episode, scen = (eng.v[k].cpu().numpy() for k in ("step", "episode", "scen"))
        if self._np_rngs is None:
            self._np_rngs = [np.random.default_rng([self.seed, self.env_id_base + e]) for e in range(self.num_envs)]
            self._ch = [None] * self.num_envs
            self._tf = [None] * self.num_envs
        bank = self.bank.array()
        if self._host_channel is not None:
            se = np.zeros((self.num_envs, 25, 135))
        if self._host_traffic is not None:
            tr = np.zeros((self.num_envs, 25))
        for e in range(self.num_envs):
            fresh = self._ch[e] is None or (reset_mask is not None and reset_mask[e])
            if self._host_channel is not None:
                if fresh:  # plugins are re-created at every reset (comm_env.py:626-681)
                    self._ch[e] = self._host_channel(25, 1, self.num_available_rbs, self._np_rngs[e], self.root_path,
                                                     self.simu_name)
                se[e] = self._ch[e].step(int(step[e]), int(episode[e]), np.ones((25, 2)), None)[0]
            if self._host_traffic is not None:
                if fresh:
                    self._tf[e] = self._host_traffic(25, self._np_rngs[e], self.root_path)
                row = bank[scen[e]]
                sua = np.zeros((5, 25))
                for u in range(25):
                    if row["ue_slice"][u] >= 0:
                        sua[row["ue_slice"][u], u] = 1
                req = {f"slice_{s}": ({"ues": {"traffic": float(row["sl"][s]["traffic"])}}
                                      if row["sl"][s]["flags"] & 2 else {}) for s in range(5)}
                tr[e] = self._tf[e].step(sua, req, int(step[e]), int(episode[e]))
        return se, tr

    def _predict_episode(self, step, ep, initial_episode):
        """comm_env.py:415-433 for one env (host mirror, used only by the host hook path)."""
        init = initial_episode if initial_episode != -1 else self.initial_episode_number
        last = self.max_number_episodes - 1
        if (step == 0 and ep == 0) or (step == self.max_number_steps and ep == last) or initial_episode != -1:
            return init
        if step == 0 and ep <= last:
            return ep
        if ep < last:
            return ep + 1
        raise Exception("Episode number received is greater than the maximum number of episodes")

    # ---- comm_env.py:375-482 / marl_comm_env.py:31-40 ------------------------------------------
    def reset(self, seed=None, options=None, *, mask=None, initial_episodes=None):
        """reset(seed=None, options={"initial_episode": -1}) -> (obs, {}).  `mask` (bool[E]) and
        `initial_episodes` (int[E]) are the batched extensions."""
        eng = self._engine()
        if seed is not None:  # gym semantics: an explicit seed re-seeds the generators
            self.seed = int(seed)
            from . import capi
            capi.check(eng.h, eng.lib.irs_set_seed(eng.h, self.seed))
            self._np_rngs = None
        initial_episode = -1 if not options else int(options.get("initial_episode", -1))
        se = None
        if self._host_channel is not None:
            st, ep = eng.v["step"].cpu().numpy(), eng.v["episode"].cpu().numpy()
            m = np.ones(self.num_envs, bool) if mask is None else np.asarray(mask, bool)
            saved = (st.copy(), ep.copy())
            for e in np.flatnonzero(m):  # the channel's reset frame belongs to the NEW episode, step 0
                ie = int(initial_episodes[e]) if initial_episodes is not None else initial_episode
                ep[e], st[e] = self._predict_episode(int(st[e]), int(ep[e]), ie), 0
            eng.v["step"].copy_(torch.from_numpy(st)); eng.v["episode"].copy_(torch.from_numpy(ep))
            se, _ = self._hook_frames(reset_mask=m)
            eng.v["step"].copy_(torch.from_numpy(saved[0])); eng.v["episode"].copy_(torch.from_numpy(saved[1]))
        eng.reset(mask=mask, initial_episode=initial_episode, initial_episodes=initial_episodes, se=se)
        self._check()
        rec = self._history()
        if rec is not None:
            rec.mark_reset()
        return self._format_obs(), {}

    def _history(self):
        if not self.save_hist:
            return None
        if self._recorder is None:
            ids = self.hist_envs if self.hist_envs is not None else ([0] if self.num_envs == 1 else [])
            if len(ids) == 0:
                self.save_hist = False
                return None
            from .history import HistoryRecorder
            self._recorder = HistoryRecorder(self._engine(), self.bank, ids, self.max_number_steps, self.hist_root_path,
                                             self.simu_name, self.agent_name, flat_surface=self._surface.flat,
                                             with_se=self.hist_with_se)
        return self._recorder

    # ---- comm_env.py:271-373 / marl_comm_env.py:42-66 ------------------------------------------
    def step(self, action):
        eng = self._engine()
        a0, intra = self._parse_action(action)
        se, tr = self._hook_frames()
        rec = self._history()
        eng.step(a0, intra, se, tr)
        self._check()
        if rec is not None:
            rec.record()
            rec.flush()
        obs = self._format_obs()
        s = eng.surface()
        if self.num_envs > 1:
            rew = s["reward"][:, 0] if self._surface.flat else s["reward"]
            done = s["done"].bool()
            return obs, rew, done, done.clone(), {}
        r = s["reward"][0].cpu().numpy()
        done = bool(s["done"][0].item())
        if self._surface.flat or not isinstance(action, dict):
            # array actions -> scalar reward and plain bools (marl_comm_env.py:60-66)
            return obs, float(r[0]), done, done, {}
        reward = {f"player_{i}": float(r[i]) for i in range(6)}
        term = {f"player_{i}": done for i in range(6)}
        term["__all__"] = done
        return obs, reward, term, dict(term), {}

    def _parse_action(self, action):
        E = self.num_envs
        if action is None or isinstance(action, agents_mod.DevicePolicy):
            if self._surface.policy == 0:
                raise ValueError("this agent surface has no on-device policy: pass an action")
            return None, None
        if isinstance(action, dict):
            a0 = action["player_0"]
            if "player_1" in action:
                cols = [torch.as_tensor(np.asarray(action[f"player_{i}"]) if not isinstance(action[f"player_{i}"], torch.Tensor)
                                        else action[f"player_{i}"]).reshape(E) for i in range(1, 6)]
                intra = torch.stack(cols, dim=1).to(torch.int32)
            else:
                intra = None
        elif isinstance(action, (tuple, list)) and len(action) == 2 and not np.isscalar(action[0]):
            a0, intra = action
        else:
            a0, intra = action, None
        a0 = torch.as_tensor(np.asarray(a0) if not isinstance(a0, torch.Tensor) else a0).reshape(E, 5)
        if intra is None and self._surface.fixed_intra < 0:
            raise ValueError("IBSched surface: the action needs player_1..player_5 (ib_sched.py:394-411)")
        if intra is not None:
            intra = torch.as_tensor(intra).reshape(E, 5)
            if self.debug and bool(((intra < 0) | (intra > 2)).any()) and self._surface.fixed_intra < 0:
                raise ValueError("Invalid intra-slice scheduling action")  # ib_sched.py:341-344
        return a0, intra

    def _check(self):
        """debug=True of the reference (comm_env.py:313-319): surface device-side invariant flags."""
        if not self.debug:
            return
        bad = self.eng.check_errors()
        if bad is not None:
            env, code = bad
            from . import capi
            if code & capi.ERR_INTRA_ACTION:
                raise ValueError("Invalid intra-slice scheduling action")
            if code & capi.ERR_EPISODE_RANGE:
                raise Exception("Episode number received is greater than the maximum number of episodes")
            raise Exception(f"env {env}: scheduling decision violates the RB invariants (code {code})")

    def _format_obs(self):
        s = self.eng.surface()
        if self.num_envs > 1:
            if self._surface.flat:
                return s["obs0"]
            return {"player_0": {"observations": s["obs0"], "action_mask": s["mask0"]},
                    "players": {"observations": s["obs_i"], "action_mask": s["mask_i"]},
                    "sorted_slices": s["sorted_slices"]}
        obs0 = s["obs0"][0].cpu().numpy()
        if self._surface.flat:
            return obs0  # marr.py:52-55: only player_0's vector
        mask0 = s["mask0"][0].cpu().numpy()
        obs_i, mask_i = s["obs_i"][0].cpu().numpy(), s["mask_i"][0].cpu().numpy()
        order = s["sorted_slices"][0].cpu().numpy()
        out = {"player_0": {"observations": obs0, "action_mask": mask0}}
        for sl in order:  # dict insertion order = sorted slice order (ib_sched.py:91,186-200)
            out[f"player_{int(sl) + 1}"] = {"observations": obs_i[sl], "action_mask": mask_i[sl]}
        return out

    # ---- reference attributes -------------------------------------------------------------------
    @property
    def step_number(self):
        v = self._engine().v["step"]
        return int(v[0].item()) if self.num_envs == 1 else v

    @property
    def episode_number(self):
        v = self._engine().v["episode"]
        return int(v[0].item()) if self.num_envs == 1 else v

    def episode_metrics(self):
        """Named episode statistics summed over this process's envs (replaces the offline pass
        over hist/*.npz in results/gen_results.py:791-809, 874-1022)."""
        from . import capi
        return dict(zip(capi.METRIC_NAMES, self._engine().episode_metrics().tolist()))

    def close(self):
        if self.eng is not None:
            self.eng.close()
            self.eng = None


class MARLCommEnv(BatchedMARLCommEnv):
    """Drop-in name: one env instance with exactly the reference's return structures."""

    def __init__(self, *args, **kwargs):
        kwargs.setdefault("num_envs", 1)
        super().__init__(*args, **kwargs)
