"""Vector front ends a trainer can hold: the batch of `BatchedMARLCommEnv` behind the vector-env interfaces of
Stable-Baselines3 and RLlib, the two consumers of the reference env.

The reference builds ONE env per worker: SB3 wraps it in a `DummyVecEnv` behind `PPO("MlpPolicy", env)`
(agents/sb3_sched.py:105-111, flat `player_0` surface of :159-187), RLlib registers a `MultiAgentEnv`
(sixg_radio_mgmt/sixg_radio_mgmt/marl_comm_env.py:12-66; agents/ray_agent.py:42, 10 env runners x 1 env :296-300).
Here thousands of env instances live in one batch, so the vector interface is the natural seam: the classes below are
duck-typed to `stable_baselines3.common.vec_env.VecEnv` and to RLlib's `VectorEnv` (`vector_reset` / `reset_at` /
`vector_step`) without importing either library (neither is a dependency of this package).

Episode ends follow each library's convention: the env that finished is reset in the same call, its last observation
is handed out as `infos[i]["terminal_observation"]` (SB3) or returned by `vector_step` before `reset_at` (RLlib).
The device does not auto-reset under these adapters (the terminal observation would be overwritten by the reset
observation in the same launch); finished envs are reset with one masked `irs_reset`.
"""
from __future__ import annotations

import numpy as np
import torch

from . import agents as agents_mod
from .env import BatchedMARLCommEnv


def _make_env(env_or_args, surface, kwargs):
    if isinstance(env_or_args, BatchedMARLCommEnv) or hasattr(env_or_args, "num_envs"):
        env = env_or_args
        if env.auto_reset:
            raise ValueError("pass an env created with auto_reset=False: the adapter resets finished envs itself")
    else:
        kwargs = dict(kwargs)
        kwargs["auto_reset"] = False
        env = BatchedMARLCommEnv(*env_or_args, **kwargs)
    if surface is not None:
        ag = surface(env, 25, 5, 1, np.array([135]))
        env.set_agent_functions(ag.obs_space_format, ag.action_format, ag.calculate_reward, ag.get_obs_space(),
                                ag.get_action_space())
    return env


class SB3VecEnv:
    """`stable_baselines3.common.vec_env.VecEnv`-shaped view of the batch for the flat agents
    (agents/sb3_sched.py:159-187: observation = player_0's 50-vector, action = 5 inter-slice scores, scalar reward;
    intra-slice scheduler fixed to round-robin, `IBSchedSB3PF` = proportional fair).

        venv = SB3VecEnv((Channel, Traffic, Mobility, Association, "mult_slice", "sb3_ib_sched", seed),
                         num_envs=4096, enable_random_episodes=True, ...)
        model = PPO("MlpPolicy", venv)          # what agents/sb3_sched.py:105-111 does with one env

    `reset()` returns float32 observations [num_envs, 50]; `step(actions)` returns (obs, rewards, dones, infos) with
    numpy arrays as SB3 expects; `obs_dtype=None` keeps float64.  `device_tensors=True` returns torch tensors on the
    GPU instead (for a trainer that keeps rollouts in HBM)."""

    def __init__(self, env_or_args, *, surface=agents_mod.IBSchedSB3, obs_dtype=np.float32, device_tensors=False,
                 **kwargs):
        self.env = _make_env(env_or_args, surface, kwargs)
        if not self.env._surface.flat:
            raise ValueError("SB3VecEnv needs a flat agent surface (IBSchedSB3, IBSchedSB3PF, MARR, MAPF, SchedTWC, ...)")
        self.num_envs = self.env.num_envs
        self.observation_space, self.action_space = self.env.observation_space, self.env.action_space
        self.obs_dtype, self.device_tensors = obs_dtype, device_tensors
        self.render_mode = None
        self.reset_infos = [{} for _ in range(self.num_envs)]
        self._actions = None
        self._seeds = None
        self._options = [{} for _ in range(self.num_envs)]

    # -- VecEnv API ------------------------------------------------------------------------------
    def _out(self, t):
        if self.device_tensors:
            return t if self.obs_dtype is None else t.to(getattr(torch, np.dtype(self.obs_dtype).name))
        a = t.cpu().numpy()
        return a if self.obs_dtype is None else a.astype(self.obs_dtype)

    def seed(self, seed=None):
        self._seeds = seed
        return [None if seed is None else seed + i for i in range(self.num_envs)]

    def set_options(self, options=None):
        self._options = [dict(options or {}) for _ in range(self.num_envs)] if not isinstance(options, list) else options

    def reset(self):
        opt = self._options[0] if self._options else {}
        obs, _ = self.env.reset(seed=self._seeds, options=opt or None)
        self._seeds = None
        self._options = [{} for _ in range(self.num_envs)]
        self.reset_infos = [{} for _ in range(self.num_envs)]
        return self._out(obs if torch.is_tensor(obs) else torch.as_tensor(obs)[None])

    def step_async(self, actions):
        self._actions = actions

    def step_wait(self):
        env = self.env
        obs, rew, done, _, _ = env.step(self._actions)
        if env.num_envs == 1:
            obs, rew, done = torch.as_tensor(obs)[None], torch.as_tensor([rew]), torch.as_tensor([done])
        infos = [{} for _ in range(self.num_envs)]
        done_h = done.cpu().numpy().astype(bool) if torch.is_tensor(done) else np.asarray(done, bool)
        if done_h.any():
            term = self._out(obs[torch.as_tensor(done_h, device=obs.device)] if torch.is_tensor(obs) else obs)
            for j, i in enumerate(np.flatnonzero(done_h)):
                # the episode ended on the step limit (comm_env.py:370): a time-limit truncation in SB3's terms
                infos[i] = {"terminal_observation": term[j], "TimeLimit.truncated": True}
            new_obs, _ = env.reset(mask=done_h)
            obs = new_obs if torch.is_tensor(new_obs) else torch.as_tensor(new_obs)[None]
        rew_o = rew if self.device_tensors else rew.cpu().numpy().astype(np.float32)
        return self._out(obs), rew_o, (done if self.device_tensors else done_h), infos

    def step(self, actions):
        self.step_async(actions)
        return self.step_wait()

    def close(self):
        self.env.close()

    def get_attr(self, attr_name, indices=None):
        n = self.num_envs if indices is None else len(np.atleast_1d(indices))
        return [getattr(self.env, attr_name)] * n

    def set_attr(self, attr_name, value, indices=None):
        setattr(self.env, attr_name, value)

    def env_method(self, method_name, *args, indices=None, **kwargs):
        n = self.num_envs if indices is None else len(np.atleast_1d(indices))
        return [getattr(self.env, method_name)(*args, **kwargs)] * n

    def env_is_wrapped(self, wrapper_class, indices=None):
        n = self.num_envs if indices is None else len(np.atleast_1d(indices))
        return [False] * n

    def get_images(self):
        return [None] * self.num_envs

    def render(self, mode=None):
        return None

    @property
    def unwrapped(self):
        return self


class RLlibVectorEnv:
    """RLlib `VectorEnv`-shaped view of the batch with the reference's multi-agent dicts
    (marl_comm_env.py:31-66): every sub-env speaks `{"player_0": ..., "player_1": ...}` as `MARLCommEnv` does.

        vector_reset()            -> ([obs_dict] * E, [info] * E)
        reset_at(i)               -> (obs_dict, info)
        vector_step([action_dict] * E) -> (obs, rewards, terminateds, truncateds, infos), lists of per-env dicts with
                                     the "__all__" entries of marl_comm_env.py:51-66

    `vector_step` returns the observation the step produced, also for envs that finished; RLlib then calls
    `reset_at(i)` for those (its VectorEnv contract), which runs a masked reset on the device."""

    def __init__(self, env_or_args, *, surface=agents_mod.IBSched, **kwargs):
        self.env = _make_env(env_or_args, surface, kwargs)
        self.num_envs = self.env.num_envs
        self.observation_space, self.action_space = self.env.observation_space, self.env.action_space
        self._agent_ids = set(self.env.agents)

    # -- conversions -------------------------------------------------------------------------------
    def _obs_list(self, idx=None):
        s = self.env.eng.surface()
        ids = range(self.num_envs) if idx is None else idx
        obs0, m0 = s["obs0"].cpu().numpy(), s["mask0"].cpu().numpy()
        if self.env._surface.flat:
            return [obs0[e] for e in ids]
        oi, mi, order = s["obs_i"].cpu().numpy(), s["mask_i"].cpu().numpy(), s["sorted_slices"].cpu().numpy()
        out = []
        for e in ids:
            o = {"player_0": {"observations": obs0[e], "action_mask": m0[e]}}
            for sl in order[e]:  # insertion order = sorted slice order (ib_sched.py:91,186-200)
                o[f"player_{int(sl) + 1}"] = {"observations": oi[e, sl], "action_mask": mi[e, sl]}
            out.append(o)
        return out

    def _batch_actions(self, actions):
        E = self.num_envs
        if self.env._surface.flat:
            return np.stack([np.asarray(a, np.float64) for a in actions]).reshape(E, 5)
        a0 = np.stack([np.asarray(a["player_0"], np.float64) for a in actions])
        intra = np.array([[int(a[f"player_{i}"]) for i in range(1, 6)] for a in actions], np.int32)
        return a0, intra

    # -- VectorEnv API -----------------------------------------------------------------------------
    def vector_reset(self, *, seeds=None, options=None):
        seed = None if seeds is None else seeds[0]
        opt = None if options is None else options[0]
        self.env.reset(seed=seed, options=opt)
        return self._obs_list(), [{} for _ in range(self.num_envs)]

    def reset_at(self, index=None, *, seed=None, options=None):
        index = 0 if index is None else int(index)
        m = np.zeros(self.num_envs, bool)
        m[index] = True
        self.env.reset(seed=seed, options=options, mask=m)
        return self._obs_list([index])[0], {}

    def vector_step(self, actions):
        env = self.env
        batch = self._batch_actions(actions)
        a0, intra = (batch, None) if env._surface.flat else batch
        se, tr = env._hook_frames()
        env.eng.step(a0, intra, se, tr)
        env._check()
        s = env.eng.surface()
        rew = s["reward"].cpu().numpy()
        done = s["done"].cpu().numpy().astype(bool)
        obs = self._obs_list()
        if env._surface.flat:
            return obs, [float(r[0]) for r in rew], [bool(d) for d in done], [bool(d) for d in done], \
                [{} for _ in range(self.num_envs)]
        rewards = [{f"player_{i}": float(r[i]) for i in range(6)} for r in rew]
        terms = []
        for d in done:
            t = {f"player_{i}": bool(d) for i in range(6)}
            t["__all__"] = bool(d)
            terms.append(t)
        return obs, rewards, terms, [dict(t) for t in terms], [{} for _ in range(self.num_envs)]

    def get_sub_environments(self):
        return []

    def try_render_at(self, index=None):
        return None

    def close(self):
        self.env.close()
