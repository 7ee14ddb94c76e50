"""Adapters that give the CUDA engine and the CPU oracle the same reset()/step() face, so the
parity tests drive both with identical inputs (tests only)."""
from __future__ import annotations

import numpy as np

OUT_KEYS = ["pkt_incoming", "pkt_throughputs", "sent", "dropped", "occ", "lat", "rb_start", "rb_count",
            "obs0", "mask0", "obs_i", "mask_i", "sorted_slices", "reward", "done"]


class CudaEnv:
    """BatchEngine with numpy in / numpy out (synchronises every call: tests only)."""

    def __init__(self, bank, num_envs=1, **kw):
        from intent_radio_sched_multi_slice_b200.engine import BatchEngine

        self.eng = BatchEngine(bank, num_envs, **kw)
        self.raw = self.eng.enable_raw_outputs()
        self.E = num_envs

    def _out(self):
        import torch

        torch.cuda.synchronize()
        pos = self.eng.check_errors()
        assert pos is None, f"device error flag: env {pos[0]} code {pos[1]}"
        s = self.eng.surface()
        out = {k: s[k].cpu().numpy() for k in ("obs0", "mask0", "obs_i", "mask_i", "reward", "done")}
        for k in ("pkt_incoming", "pkt_throughputs", "sent", "dropped", "occ", "lat", "rb_start", "rb_count",
                  "sorted_slices", "a0_used", "intra_used", "traffic_used"):
            out[k] = self.raw[k].cpu().numpy()
        return out

    def reset(self, mask=None, initial_episode=-1, se=None, initial_episodes=None):
        if se is not None and self.E > 1 and np.asarray(se).shape[0] == 1:
            se = np.repeat(se, self.E, axis=0)
        self.eng.reset(mask=mask, initial_episode=initial_episode, initial_episodes=initial_episodes, se=se)
        return self._out()

    def step(self, a0=None, intra=None, se=None, traffic=None):
        self.eng.step(a0, intra, se, traffic)
        return self._out()

    def buffers(self):
        return self.eng.buffers().cpu().numpy()

    def set_episode_order(self, order):
        self.eng.set_episode_order(order)

    def state(self):
        v = self.eng.v
        return {k: v[k].cpu().numpy() for k in ("step", "episode", "scen")}

    def episode_metrics(self):
        per = self.eng.v["acc"].cpu().numpy()
        return per, self.eng.episode_metrics()


def compare_outputs(got, want, tag, rtol=1e-12, keys=OUT_KEYS):
    for k in keys:
        g, w = np.asarray(got[k]), np.asarray(want[k])
        if g.dtype.kind in "iub" and w.dtype.kind in "iub":
            assert np.array_equal(g.astype(np.int64), w.astype(np.int64)), f"{tag}: {k} differs\n{g}\n{w}"
        elif k in ("pkt_incoming", "pkt_throughputs"):
            assert np.array_equal(g, w), f"{tag}: {k} differs"
        else:
            np.testing.assert_allclose(g, w, rtol=rtol, atol=1e-300, err_msg=f"{tag}: {k}")
