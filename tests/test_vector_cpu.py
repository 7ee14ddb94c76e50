"""CPU suite: the SB3 / RLlib vector adapters (intent_radio_sched_multi_slice_b200.vector) against a stand-in batch
env with the surface of BatchedMARLCommEnv - the contracts a trainer relies on (stable_baselines3 VecEnv: stacked
observations, `terminal_observation` + same-call reset of finished envs; RLlib VectorEnv: per-env multi-agent dicts
with "__all__", reset_at).  The real batch under the adapters is exercised in the GPU suite."""
import numpy as np
import torch

from intent_radio_sched_multi_slice_b200 import agents, vector


class _FakeEngine:
    def __init__(self, env):
        self.env = env

    def surface(self):
        e = self.env
        return {"obs0": e.obs0, "mask0": e.mask0, "obs_i": e.obs_i, "mask_i": e.mask_i, "reward": e.reward,
                "done": e.done, "sorted_slices": e.order}

    def step(self, a0, intra, se, tr):
        self.env._advance(torch.as_tensor(np.asarray(a0)))


class FakeBatch:
    """obs0[e] = (episode, step, last action sum, ...); episodes are `T` steps long."""
    auto_reset = False

    def __init__(self, E, T, surface):
        self.num_envs, self.T, self._surface = E, T, surface
        self.agents = {f"player_{i}" for i in range(6)}
        ag = surface()
        self.observation_space, self.action_space = ag.get_obs_space(), ag.get_action_space()
        self.step_n = torch.zeros(E, dtype=torch.int64)
        self.episode = torch.zeros(E, dtype=torch.int64)
        self.obs0 = torch.zeros(E, 50, dtype=torch.float64)
        self.mask0 = torch.ones(E, 5, dtype=torch.int8)
        self.obs_i = torch.zeros(E, 5, 19, dtype=torch.float64)
        self.mask_i = torch.ones(E, 5, 5, dtype=torch.int8)
        self.reward = torch.zeros(E, 6, dtype=torch.float64)
        self.done = torch.zeros(E, dtype=torch.uint8)
        self.order = torch.tensor([[2, 0, 1, 4, 3]] * E, dtype=torch.int8)
        self.eng = _FakeEngine(self)
        self.resets = []

    def set_agent_functions(self, *a):
        pass

    def _hook_frames(self):
        return None, None

    def _check(self):
        pass

    def _advance(self, a0):
        self.step_n += 1
        self.obs0[:, 0], self.obs0[:, 1], self.obs0[:, 2] = self.episode.double(), self.step_n.double(), a0.double().sum(1)
        self.obs_i[:, :, 0] = self.step_n.double()[:, None]
        self.reward[:, 0] = self.step_n.double() + 0.5
        self.done[:] = (self.step_n == self.T).to(torch.uint8)

    def reset(self, seed=None, options=None, mask=None):
        m = torch.ones(self.num_envs, dtype=torch.bool) if mask is None else torch.as_tensor(np.asarray(mask, bool))
        self.resets.append(m.clone())
        self.episode[m] += (self.step_n[m] > 0).long()
        self.step_n[m] = 0
        self.obs0[m] = 0
        self.obs0[m, 0] = self.episode[m].double()
        self.done[m] = 0
        return (self.obs0 if self._surface.flat else None), {}

    def step(self, action):
        self._advance(torch.as_tensor(np.asarray(action)))
        return self.obs0, self.reward[:, 0], self.done.bool(), self.done.bool().clone(), {}

    def close(self):
        pass


def test_sb3_vec_env_contract():
    E, T = 5, 3
    fake = FakeBatch(E, T, agents.IBSchedSB3)
    fake.step_n[:] = torch.tensor([0, 1, 2, 0, 2])  # envs 2 and 4 finish on the first step
    venv = vector.SB3VecEnv(fake, surface=None)
    assert venv.num_envs == E and venv.action_space.shape == (5,) and venv.observation_space.shape == (50,)
    fake.step_n[:] = 0
    obs = venv.reset()
    assert isinstance(obs, np.ndarray) and obs.shape == (E, 50) and obs.dtype == np.float32
    fake.step_n[:] = torch.tensor([0, 1, 2, 0, 2])
    acts = np.full((E, 5), 0.25)
    venv.step_async(acts)
    obs, rew, dones, infos = venv.step_wait()
    assert obs.shape == (E, 50) and rew.shape == (E,) and rew.dtype == np.float32 and dones.dtype == bool
    assert dones.tolist() == [False, False, True, False, True] and len(infos) == E
    for i in range(E):
        if dones[i]:
            t = infos[i]["terminal_observation"]
            assert t.shape == (50,) and t[1] == T and t[2] == 1.25 and infos[i]["TimeLimit.truncated"] is True
            assert obs[i, 1] == 0 and obs[i, 0] == 1          # already the next episode's reset observation
        else:
            assert infos[i] == {} and obs[i, 2] == 1.25
    assert fake.resets[-1].tolist() == [False, False, True, False, True]  # ONE masked reset for the finished envs
    assert venv.env_is_wrapped(object) == [False] * E and venv.get_attr("T") == [T] * E
    assert venv.seed(3) == [3, 4, 5, 6, 7]


def test_rllib_vector_env_contract():
    E, T = 3, 2
    fake = FakeBatch(E, T, agents.IBSched)
    venv = vector.RLlibVectorEnv(fake, surface=None)
    obs, infos = venv.vector_reset()
    assert len(obs) == E and infos == [{}] * E
    assert list(obs[0].keys()) == ["player_0", "player_3", "player_1", "player_2", "player_5", "player_4"]  # sorted order
    assert obs[0]["player_0"]["observations"].shape == (50,) and obs[0]["player_2"]["action_mask"].dtype == np.int8
    act = {"player_0": np.zeros(5)}
    act.update({f"player_{i}": 1 for i in range(1, 6)})
    for k in range(T):
        obs, rew, term, trunc, infos = venv.vector_step([act] * E)
    assert set(rew[0]) == {f"player_{i}" for i in range(6)} and rew[1]["player_0"] == T + 0.5
    assert term[2]["__all__"] is True and trunc[2]["player_4"] is True and set(term[0]) == set(rew[0]) | {"__all__"}
    assert obs[1]["player_1"]["observations"][0] == T      # the step's own observation, not a reset one
    o, info = venv.reset_at(1)
    assert info == {} and fake.resets[-1].tolist() == [False, True, False] and o["player_0"]["observations"][0] == 1
    assert venv.get_sub_environments() == []
