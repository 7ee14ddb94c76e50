"""Episode history files compatible with the reference's `Metrics.read`
(`sixg_radio_mgmt/sixg_radio_mgmt/metrics.py:43-60,73-145`; consumers
`results/gen_results.py:845-1022`): `<root>/hist/<simu>/<agent>/ep_<N>.npz` with the 16 keys,
one entry per step of the episode.

The reference appends the whole raw observation of its single env to Python lists every step
(`comm_env.py:343-344`).  Here the per-step values of a chosen subset of the batch are scattered
into device-resident [n_recorded, max_steps, ...] tensors right after each step (asynchronously,
on the engine's stream) and only a finished episode crosses PCIe.  The RB allocation is kept as
(start, count) runs and expanded to the reference's dense 0/1 matrix when the file is written.
"""
from __future__ import annotations

import os

import numpy as np
import torch

from . import capi

U, S, R = capi.MAX_UES, capi.MAX_SLICES, capi.NUM_RBS
_PER_UE = {"pkt_incoming": torch.float64, "pkt_throughputs": torch.float64, "sent": torch.int32,
           "dropped": torch.int32, "occ": torch.float64, "lat": torch.float64, "rb_start": torch.int16,
           "rb_count": torch.int16}
KEYS = ["pkt_incoming", "pkt_throughputs", "pkt_effective_thr", "buffer_occupancies", "buffer_latencies",
        "dropped_pkts", "mobility", "spectral_efficiencies", "basestation_ue_assoc", "basestation_slice_assoc",
        "slice_ue_assoc", "sched_decision", "reward", "slice_req", "obs", "agent_action"]


def dense_sched(rb_start, rb_count):
    """(start, count) runs [T, 25] -> the reference's sched_decision [T, 1, 25, 135] of 0/1 float64."""
    r = np.arange(R)[None, None, :]
    s, c = np.asarray(rb_start, np.int64)[:, :, None], np.asarray(rb_count, np.int64)[:, :, None]
    return ((r >= s) & (r < s + c)).astype(np.float64)[:, None]


class HistoryRecorder:
    def __init__(self, eng, bank, env_ids, max_steps: int, root_path=".", simu_name="mult_slice", agent_name="agent",
                 flat_surface=False, with_se=False):
        self.eng, self.bank, self.T = eng, bank, int(max_steps)
        self.ids = torch.as_tensor(list(env_ids), dtype=torch.long, device=eng.device)
        self.n = len(self.ids)
        self.root_path, self.simu_name, self.agent_name = root_path, simu_name, agent_name
        self.flat, self.with_se = flat_surface, with_se
        names = list(_PER_UE) + ["a0_used", "intra_used", "sorted_slices"] + (["se_used"] if with_se else [])
        have = eng._raw or {}
        self.raw = eng.enable_raw_outputs(sorted(set(names) | set(have)))
        n, T, dev = self.n, self.T, eng.device
        self.h = {k: torch.zeros((n, T, U), dtype=dt, device=dev) for k, dt in _PER_UE.items()}
        self.h["a0"] = torch.zeros((n, T, S), dtype=torch.float64, device=dev)
        self.h["intra"] = torch.zeros((n, T, S), dtype=torch.int32, device=dev)
        self.h["reward"] = torch.zeros((n, T, 6), dtype=torch.float64, device=dev)
        self.h["obs0"] = torch.zeros((n, T, capi.OBS0), dtype=torch.float64, device=dev)
        self.h["obsi"] = torch.zeros((n, T, S, capi.OBSI), dtype=torch.float64, device=dev)
        self.h["mask0"] = torch.zeros((n, T, S), dtype=torch.int8, device=dev)
        self.h["maski"] = torch.zeros((n, T, S, 5), dtype=torch.int8, device=dev)
        self.h["order"] = torch.zeros((n, T, S), dtype=torch.int8, device=dev)
        if with_se:
            self.h["se"] = torch.zeros((n, T, U, R), dtype=torch.float64, device=dev)
        self.episode = torch.zeros(n, dtype=torch.int32, device=dev)   # episode / scenario the open block belongs to
        self.scen = torch.zeros(n, dtype=torch.int32, device=dev)
        self.scen_rows = None  # the scenario rows themselves (they may have been generated on the device: assoc_mode 2)
        self.ar = torch.arange(n, device=dev)
        self.written: list[str] = []

    def mark_reset(self):
        """Call after a reset: the recorded envs' current episode / scenario open a new block."""
        v = self.eng.v
        self.episode.copy_(v["episode"][self.ids])
        self.scen.copy_(v["scen"][self.ids])
        self.scen_rows = self.eng.scenario_rows()[self.ids.cpu().numpy()]

    def record(self):
        """Call right after a step (asynchronous; no host synchronisation)."""
        v, ids, ar = self.eng.v, self.ids, self.ar
        done = v["out_done"][ids].bool()
        # the step just taken is number step-1 of its episode; a finished env may already have been
        # rolled over by auto-reset (step counter back at 0): that step was the last one
        slot = torch.where(done, torch.full_like(ids, self.T - 1), v["step"][ids].long() - 1).clamp_(0, self.T - 1)
        for k in _PER_UE:
            self.h[k][ar, slot] = self.raw[k][ids]
        self.h["a0"][ar, slot] = self.raw["a0_used"][ids]
        self.h["intra"][ar, slot] = self.raw["intra_used"][ids]
        self.h["reward"][ar, slot] = v["out_reward"][ids]
        self.h["obs0"][ar, slot] = v["out_obs0"][ids]
        self.h["obsi"][ar, slot] = v["out_obsi"][ids]
        self.h["mask0"][ar, slot] = v["out_mask0"][ids]
        self.h["maski"][ar, slot] = v["out_maski"][ids]
        self.h["order"][ar, slot] = self.raw["sorted_slices"][ids]
        if self.with_se:
            self.h["se"][ar, slot] = self.raw["se_used"][ids]
        self._done = done

    def flush(self):
        """Write the block of every recorded env whose episode just ended (one small D2H sync)."""
        done = self._done.cpu().numpy()
        out = []
        for i in np.flatnonzero(done):
            out.append(self._write(int(i)))
        return out

    def _write(self, i: int) -> str:
        T = self.T
        h = {k: t[i].cpu().numpy() for k, t in self.h.items()}
        scen, ep = int(self.scen[i].item()), int(self.episode[i].item())
        if scen < len(self.bank.reqs):
            sua, req = self.bank.assocs[scen], self.bank.reqs[scen]
        else:  # a scenario generated on the device (assoc_mode 2): the dicts are rebuilt from its dense row
            from . import scenario as scn
            req, sua = scn.req_from_row(self.scen_rows[i])
        bsa = np.array([[1.0 if req.get(f"slice_{s}", {}) != {} else 0.0 for s in range(S)]])
        rep = lambda a: np.repeat(np.asarray(a)[None], T, axis=0)
        reward = np.empty(T, object)
        obs = np.empty(T, object)
        action = np.empty(T, object)
        for k in range(T):
            if self.flat:
                reward[k], obs[k], action[k] = float(h["reward"][k, 0]), h["obs0"][k], h["a0"][k]
                continue
            order = [0] + [int(sl) + 1 for sl in h["order"][k]]  # common.py:396-439: the observation's key order
            reward[k] = {f"player_{p}": h["reward"][k, p] for p in order}  # np.float64 values, like the reference
            o = {"player_0": {"observations": h["obs0"][k], "action_mask": h["mask0"][k]}}
            for sl in h["order"][k]:  # insertion order = sorted slices (ib_sched.py:91,186-200)
                o[f"player_{int(sl) + 1}"] = {"observations": h["obsi"][k, sl], "action_mask": h["maski"][k, sl]}
            obs[k] = o
            a = {"player_0": h["a0"][k]}
            a.update({f"player_{p + 1}": int(h["intra"][k, p]) for p in range(S)})
            action[k] = a
        if self.flat:
            reward, obs, action = reward.astype(np.float64), np.stack(list(obs)), np.stack(list(action))
            if self.eng.cfg.policy == 1:  # MARR.step returns integer scores +-1 (agents/marr.py:40-47)
                action = action.astype(np.int64)
        from . import scenario as scn
        req = scn.callable_req(req)  # operators as NumPy ufuncs, like the dicts the reference pickles
        slice_req = np.empty(T, object)
        for k in range(T):
            slice_req[k] = req
        se = h["se"][:, None] if self.with_se else np.zeros((T, 1, U, R))
        hist = {
            "pkt_incoming": h["pkt_incoming"], "pkt_throughputs": h["pkt_throughputs"],
            "pkt_effective_thr": h["sent"].astype(np.int64), "buffer_occupancies": h["occ"],
            "buffer_latencies": h["lat"], "dropped_pkts": h["dropped"].astype(np.int64),
            "mobility": np.ones((T, U, 2)), "spectral_efficiencies": se,
            "basestation_ue_assoc": rep((sua.sum(axis=0) > 0).astype(np.float64)[None]),
            "basestation_slice_assoc": rep(bsa), "slice_ue_assoc": rep(sua),
            "sched_decision": dense_sched(h["rb_start"], h["rb_count"]), "reward": reward, "slice_req": slice_req,
            "obs": obs, "agent_action": action,
        }
        path = os.path.join(self.root_path, "hist", self.simu_name, self.agent_name)
        os.makedirs(path, exist_ok=True)
        fn = os.path.join(path, f"ep_{ep}.npz")
        np.savez_compressed(fn, **hist)  # metrics.py:96-98
        self.written.append(fn)
        return fn


def read(root_path: str, simu_name: str, agent_name: str, episode_number: int) -> dict:
    """Metrics.read (metrics.py:102-145): the same 15 keys it returns."""
    data = np.load(os.path.join(root_path, "hist", simu_name, agent_name, f"ep_{episode_number}.npz"), allow_pickle=True)
    return {k: data[k] for k in KEYS if k != "obs"}
