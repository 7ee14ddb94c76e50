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
                 auto_reset=None, env_id_base=0, channel_episodes=None, hist_envs=None, hist_with_se=True,
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
        # comm_env.py:219-226, 434-444: every env plays its own shuffled cycle of [initial_episode, max_episodes)
        # (a counter-based permutation keyed by (seed, global env id); see irs_b200.h cfg.random_episodes)
        self.enable_random_episodes = bool(enable_random_episodes)
        # save_hist (comm_env.py:345-350): hist/<simu>/<agent>/ep_N.npz per finished episode, for the envs in
        # `hist_envs` (default: the single env of a MARLCommEnv; nothing for a batch unless asked)
        self.save_hist = bool(save_hist)
        self.hist_envs = hist_envs
        # spectral_efficiencies in the history files: results/gen_results.py:260-322 computes throughputs from this
        # key, so it is recorded by default (27 MB of HBM per recorded env and 1000-step episode); False writes zeros
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
        hooked = self._host_channel is not None or self._host_traffic is not None
        if hooked and auto_reset:
            # the device rollover cannot call a host plugin for the new episode's step-0 frame (comm_env.py:449-451)
            raise ValueError("auto_reset=True needs device-side channel and traffic plugins; with host plugin hooks "
                             "reset finished envs with reset(mask=done)")
        self.auto_reset = (self.num_envs > 1 and not hooked) if auto_reset is None else bool(auto_reset)
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
        self._callbacks = None
        if obs_space_format is not None:
            self.set_agent_functions(obs_space_format, action_format, calculate_reward, obs_space, action_space)

    # ---- comm_env.py:484-512 ------------------------------------------------------------------
    def set_agent_functions(self, obs_space_format, action_format, calculate_reward, obs_space=None,
                            action_space=None):
        surf = {agents_mod.surface_of(cb) for cb in (obs_space_format, action_format, calculate_reward)}
        if len(surf) == 1 and None not in surf:  # a built-in surface: evaluated by the kernels
            self._callbacks = None
            self._surface = surf.pop()
            self.observation_space = obs_space if obs_space is not None else self._surface().get_obs_space()
            self.action_space = action_space if action_space is not None else self._surface().get_action_space()
        else:
            # Arbitrary callables (comm_env.py:484-512): honoured on the HOST path.  action_format runs here and its
            # scheduling matrix goes to the device as it is (irs_actions.sched_bits); UEs.step runs on the device; the
            # raw frame comes back and obs_space_format / calculate_reward run here on the reference's dict
            # (comm_env.py:320-341).  One Python call per env and step: the slow path, for agents that are not built in.
            if not all(callable(cb) for cb in (obs_space_format, action_format, calculate_reward)):
                raise ValueError("set_agent_functions needs three callables (comm_env.py:484-512)")
            self._callbacks = (obs_space_format, action_format, calculate_reward)
            self._surface = agents_mod.IBSched  # still written by the tail kernel; not returned on this path
            # the callbacks see the raw frame of the step that just ran, so a finished env must not be rolled over
            # inside that step (its SE frame would already be the next episode's): reset with reset(mask=done)
            self.auto_reset = False
            self.observation_space, self.action_space = obs_space, action_space
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
                                   auto_reset=int(self.auto_reset and not self._records()), env_id_base=self.env_id_base,
                                   seed=self.seed,
                                   bandwidth=float(self.bandwidths[0]), random_episodes=int(self.enable_random_episodes),
                                   **kw)
            if kw["channel_mode"] == 3:  # QuadrigaChannel[Seq]: transform the power files once, keep SE in HBM
                ch = self.ChannelClass if hasattr(self.ChannelClass, "power_bank") else \
                    plugins.BUILTIN[self.ChannelClass.__name__]
                eps = self.channel_episodes
                if eps is None:  # every episode the run can visit (quadriga.py:46-55 opens one file per episode)
                    if getattr(ch, "_registered", None) is not None:
                        eps = range(ch._registered_first, ch._registered_first + len(ch._registered))
                    elif self.max_number_episodes - self.initial_episode_number <= 64:
                        eps = range(self.initial_episode_number, self.max_number_episodes)
                    else:
                        raise ValueError(
                            f"{ch.__name__}: the run can visit episodes {self.initial_episode_number}.."
                            f"{self.max_number_episodes - 1}; pass channel_episodes=range(first, last + 1) (the files "
                            "kept in HBM, 27 MB per 1000-step episode) or a smaller max_episode_number")
                eps = [int(x) for x in eps]
                if eps != list(range(eps[0], eps[0] + len(eps))):
                    raise ValueError("channel_episodes must be a contiguous ascending range of env episode numbers")
                power = torch.from_numpy(ch.power_bank(self.root_path, eps)).to(self.device)
                # slot i = episode eps[0] + i; an env that reaches an episode outside the bank raises (debug) instead
                # of replaying another episode's file
                self.eng.set_se_trace(self.eng.power_to_se(power, layout=0, tx_power=ch.transmission_power,
                                                           noise_power=ch.thermal_noise_power), first_episode=eps[0])
            if self._callbacks is not None:
                self.eng.enable_raw_outputs()
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

    def _predict_episode(self, step, ep, initial_episode, e=0, cycle=None):
        """comm_env.py:415-444 for one env (host mirror, used only by the host hook path: a host Channel plugin is
        asked for the reset frame of the NEXT episode before the device has moved to it).  `cycle` = this env's
        (position, permutation number, first episode) when enable_random_episodes."""
        if self.enable_random_episodes:
            from . import scenario as scn
            pos, gen, base = cycle
            moved = False
            if initial_episode != -1:
                base, gen, pos, moved = initial_episode, gen + 1, 0, True
            if step == self.max_number_steps:
                pos, moved = pos + 1, True
            n = self.max_number_episodes - base
            if n <= 0:
                raise Exception("Episode number received is greater than the maximum number of episodes")
            if self.eng is not None and getattr(self.eng, "_order", None) is not None:
                order = self.eng._order[e].cpu().numpy()
                return int(order[pos % len(order)]) if moved else ep
            return scn.shuffled_episode(int(self.seed), self.env_id_base + e, gen, base, n, pos % n) if moved else ep
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
            cyc = [eng.v[k].cpu().numpy() for k in ("ep_pos", "ep_gen", "ep_base")] if self.enable_random_episodes else None
            for e in np.flatnonzero(m):  # the channel's reset frame belongs to the NEW episode, step 0
                ie = int(initial_episodes[e]) if initial_episodes is not None else initial_episode
                cy = (int(cyc[0][e]), int(cyc[1][e]), int(cyc[2][e])) if cyc is not None else None
                ep[e], st[e] = self._predict_episode(int(st[e]), int(ep[e]), ie, int(e), cy), 0
            eng.v["step"].copy_(torch.from_numpy(st)); eng.v["episode"].copy_(torch.from_numpy(ep))
            se, _ = self._hook_frames(reset_mask=m)
            eng.v["step"].copy_(torch.from_numpy(saved[0])); eng.v["episode"].copy_(torch.from_numpy(saved[1]))
        eng.reset(mask=mask, initial_episode=initial_episode, initial_episodes=initial_episodes, se=se)
        self._check()
        rec = self._history()
        if rec is not None:
            rec.mark_reset()
        if self._callbacks is not None:
            frames = self._raw_frames(None, None, reset=True)
            obs = [self._callbacks[0](f) for f in frames]
            return (obs[0] if self.num_envs == 1 else obs), {}
        return self._format_obs(), {}

    def _records(self):
        """True when episode history files are written for some env of the batch (save_hist)."""
        if not self.save_hist:
            return False
        ids = self.hist_envs if self.hist_envs is not None else ([0] if self.num_envs == 1 else [])
        return len(ids) > 0

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
    # ---- arbitrary agent callbacks: the host path (comm_env.py:300, 320-341) ------------------------
    def _raw_frames(self, scheds, actions, reset=False):
        """The reference's raw observation dict (ues.py:236-243 + comm_env.py:327-338) of every env, from the
        device's per-UE outputs.  At reset the metrics are zero (comm_env.py:455-478; `pkt_incoming` of the reset
        frame - a draw the reference shows but never enqueues - is zero here)."""
        from . import scenario as scn
        eng = self.eng
        raw = {k: v.cpu().numpy() for k, v in eng._raw.items()}
        rows = eng.scenario_rows()
        scen = eng.v["scen"].cpu().numpy()
        frames = []
        for e in range(self.num_envs):
            if scen[e] < len(self.bank.reqs):
                req, sua = self.bank.reqs[scen[e]], self.bank.assocs[scen[e]]
            else:  # generated on the device (assoc_mode 2)
                req, sua = scn.req_from_row(rows[e])
            f = {"mobility": np.ones((25, 2)), "spectral_efficiencies": raw["se_used"][e][None],
                 "basestation_ue_assoc": (sua.sum(axis=0) > 0).astype(np.float64)[None],
                 "basestation_slice_assoc": np.array([[1.0 if req.get(f"slice_{s}", {}) != {} else 0.0 for s in range(5)]]),
                 "slice_ue_assoc": sua.copy(), "slice_req": scn.callable_req(req)}
            if reset:
                z = np.zeros(25)
                f.update({"sched_decision": np.zeros((1, 25, 135)), "pkt_incoming": z.copy(), "pkt_throughputs": z.copy(),
                          "pkt_effective_thr": z.copy(), "buffer_occupancies": z.copy(), "buffer_latencies": z.copy(),
                          "dropped_pkts": z.copy()})
            else:
                f.update({"sched_decision": scheds[e], "pkt_incoming": raw["pkt_incoming"][e],
                          "pkt_throughputs": raw["pkt_throughputs"][e],
                          "pkt_effective_thr": raw["sent"][e].astype(np.int64), "buffer_occupancies": raw["occ"][e],
                          "buffer_latencies": raw["lat"][e], "dropped_pkts": raw["dropped"][e].astype(np.int64),
                          "agent_action": actions[e]})
            frames.append(f)
        return frames

    def _step_callbacks(self, action):
        eng = self._engine()
        obs_fmt, act_fmt, rew_fn = self._callbacks
        E = self.num_envs
        acts = [action] if E == 1 else list(action)
        if len(acts) != E:
            raise ValueError(f"expected one action per env ({E})")
        scheds, bits = [], np.zeros((E, 25, 5), np.uint32)
        weights = (1 << np.arange(32, dtype=np.uint64))
        for e in range(E):
            sd = np.asarray(act_fmt(acts[e]), np.float64)  # comm_env.py:300
            if sd.shape != (1, 25, 135):
                raise ValueError("action_format must return the scheduling matrix [1, 25, 135] (comm_env.py:271-296)")
            scheds.append(sd)
            m = np.zeros((25, 160), np.uint64)
            m[:, :135] = sd[0] != 0
            bits[e] = (m.reshape(25, 5, 32) * weights).sum(axis=2).astype(np.uint32)
        se, tr = self._hook_frames()
        eng.step(se=se, traffic=tr, sched_bits=bits)
        self._check()
        frames = self._raw_frames(scheds, acts)
        obs = [obs_fmt(f) for f in frames]      # comm_env.py:340
        rew = [rew_fn(f) for f in frames]       # comm_env.py:341
        done = eng.v["out_done"].cpu().numpy().astype(bool)
        if E > 1:
            return obs, rew, done, done.copy(), [{} for _ in range(E)]
        d = bool(done[0])
        if isinstance(action, dict):            # marl_comm_env.py:51-66
            term = {a: d for a in sorted(self.agents)}
            term["__all__"] = d
            return obs[0], rew[0], term, dict(term), {}
        return obs[0], rew[0], d, d, {}

    def step(self, action):
        if self._callbacks is not None:
            return self._step_callbacks(action)
        eng = self._engine()
        a0, intra = self._parse_action(action)
        se, tr = self._hook_frames()
        rec = self._history()
        eng.step(a0, intra, se, tr)
        self._check()
        if rec is not None:
            rec.record()
            rec.flush()
            if self.auto_reset:
                # With a recorder the device does not roll finished envs over inside the step (the last frame's
                # observation and SE would be overwritten before they are recorded): one masked reset here instead.
                done = eng.v["out_done"].bool()
                if bool(done.any()):
                    eng.reset(mask=done)
                    self._check()
                    rec.mark_reset()
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
        # calculate_reward_no_mask iterates over the formatted observation (common.py:396-439): the reward dict has the
        # observation's key order - player_0, then the slices in sorted order
        reward = {k: float(r[int(k.split("_")[1])]) for k in obs}
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
            if code & capi.ERR_RB_SHARED:  # comm_env.py:593-596
                raise Exception("Scheduling decision allocated the same RB for more than one UE")
            if code & capi.ERR_TRACE_EPISODE:
                raise Exception(f"env {env}: its episode has no channel file in the SE bank (channel_episodes)")
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
