"""BatchEngine: E env instances as structure-of-arrays state in HBM, stepped through the C ABI.

PyTorch owns the memory (one uint8 state blob, viewed as typed tensors) and the stream; the
kernels live in libirs_b200.so.  This is the layer `BatchedMARLCommEnv` (env.py) and bench.py
sit on.  There is no CPU path: constructing an engine without CUDA raises.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import capi
from .scenario import ScenarioBank

U, S, R = capi.MAX_UES, capi.MAX_SLICES, capi.NUM_RBS

_VIEWS = {  # name -> (torch dtype, trailing shape)
    "ue_i": (torch.int32, (U, len(capi.UI_FIELDS))), "ue_d": (torch.float64, (U, len(capi.UD_FIELDS))),
    "ring": (torch.int32, (U, capi.WINDOW, 2)), "rb": (torch.int16, (U, 2)),
    "step": (torch.int32, ()), "episode": (torch.int32, ()), "scen": (torch.int32, ()),
    "win_len": (torch.int32, ()), "win_head": (torch.int32, ()), "err": (torch.int32, ()),
    "visit": (torch.int32, ()), "ep_pos": (torch.int32, ()), "ep_gen": (torch.int32, ()), "ep_base": (torch.int32, ()),
    "wsum": (torch.int64, (U,)),
    "sorted": (torch.int8, (S,)), "acc": (torch.float64, (capi.NUM_METRICS,)),
    "in_a0": (torch.float64, (S,)), "in_intra": (torch.int32, (S,)),
    "out_obs0": (torch.float64, (capi.OBS0,)), "out_obsi": (torch.float64, (S, capi.OBSI)),
    "out_mask0": (torch.int8, (S,)), "out_maski": (torch.int8, (S, 5)),
    "out_reward": (torch.float64, (6,)), "out_done": (torch.uint8, ()),
}

_RAW_SPECS = {
    "pkt_incoming": (torch.float64, (U,)), "pkt_throughputs": (torch.float64, (U,)),
    "sent": (torch.int32, (U,)), "dropped": (torch.int32, (U,)), "occ": (torch.float64, (U,)),
    "lat": (torch.float64, (U,)), "rb_start": (torch.int16, (U,)), "rb_count": (torch.int16, (U,)),
    "sorted_slices": (torch.int8, (S,)), "a0_used": (torch.float64, (S,)), "intra_used": (torch.int32, (S,)),
    "se_used": (torch.float64, (U, R)), "traffic_used": (torch.float64, (U,)),
}


def _nbytes(dtype):
    return torch.empty((), dtype=dtype).element_size()


class BatchEngine:
    def __init__(self, bank: ScenarioBank, num_envs: int, *, device="cuda:0", depth_cap=None,
                 channel_mode=2, traffic_mode=1, sort_slices=1, fixed_intra=-1, policy=0, assoc_mode=0,
                 scen_mod=200, scen_div=100, max_steps=1000, max_episodes=10000, initial_episode=0,
                 auto_reset=0, env_id_base=0, seed=10, bandwidth=1e8, surface=0, random_episodes=0,
                 graph=False):
        if not torch.cuda.is_available():
            raise RuntimeError("intent_radio_sched_multi_slice_b200 needs a CUDA device (there is no CPU path)")
        self.lib = capi.load()
        self.device = torch.device(device)
        self.E = int(num_envs)
        self.bank_array = np.ascontiguousarray(bank.array())
        self.cdf = np.ascontiguousarray(bank.poisson_cdf(), np.float64)
        if depth_cap is None:
            depth_cap = max(bank.max_depth(), 101)  # 101: default depth of unassociated UEs (comm_env.py:184-198)
        self.cfg = capi.Config(self.E, int(depth_cap), channel_mode, traffic_mode, sort_slices, fixed_intra, policy,
                               assoc_mode, scen_mod, scen_div, len(self.bank_array), max_steps, max_episodes,
                               initial_episode, auto_reset, env_id_base, surface, int(random_episodes), seed, bandwidth)
        self.h = C.c_void_p()
        rc = self.lib.irs_create(C.byref(self.cfg), self.cdf.shape[0], C.byref(self.h))
        if rc:
            raise capi.IrsError(rc, "irs_create rejected the configuration")
        self.layout = capi.Layout()
        capi.check(self.h, self.lib.irs_state_layout(self.h, C.byref(self.layout)))
        self.row_stride = self.layout.row_stride
        with torch.cuda.device(self.device):
            self.blob = torch.empty(self.layout.total_bytes + 256, dtype=torch.uint8, device=self.device)
            off = (-self.blob.data_ptr()) % 256
            self.blob = self.blob[off:off + self.layout.total_bytes]
            capi.check(self.h, self.lib.irs_bind(self.h, self.blob.data_ptr(), self.layout.total_bytes, self._stream()))
            capi.check(self.h, self.lib.irs_load_scenarios(
                self.h, self.bank_array.ctypes.data, len(self.bank_array), self.cdf.ctypes.data, self.cdf.shape[0],
                self._stream()))
        self.v = {name: self._view(name, dt, shp) for name, (dt, shp) in _VIEWS.items()}
        for k, name in enumerate(capi.UI_FIELDS):  # strided views of the per-UE records
            self.v[name] = self.v["ue_i"][:, :, k]
        for k, name in enumerate(capi.UD_FIELDS):
            self.v[name] = self.v["ue_d"][:, :, k]
        self.v["rb_start"], self.v["rb_count"] = self.v["rb"][:, :, 0], self.v["rb"][:, :, 1]
        self.ring_rows = self._view("buf", torch.int32, (U, self.row_stride))  # physical ring rows (slot = (step - age) mod depth)
        self.metrics_total = self._view1("metrics_total", torch.float64, capi.NUM_METRICS)
        self._raw = None
        self._raw_struct = None
        self._keep = None
        # graph=True: step() replays ONE captured CUDA graph of irs_step (its launches are fixed-pointer and allocate
        # nothing).  Actions are staged into the blob's in_a0 / in_intra arrays, which the captured launch reads.
        # For small batches (BASELINE config 2: 1,024 envs) the step is launch-bound, and a replay costs one launch.
        self.graph = bool(graph)
        self._graph = None
        self._graph_key = None

    # -- plumbing ---------------------------------------------------------------------------
    def _stream(self):
        return torch.cuda.current_stream(self.device).cuda_stream

    def _view1(self, name, dtype, n):
        off = getattr(self.layout, name)
        return self.blob[off:off + n * _nbytes(dtype)].view(dtype)

    def _view(self, name, dtype, shape):
        n = self.E * int(np.prod(shape, dtype=np.int64)) if shape else self.E
        return self._view1(name, dtype, n).view((self.E,) + tuple(shape))

    def enable_raw_outputs(self, names=None):
        """Allocate the optional per-UE outputs (ues.py:236-243 raw metrics + inputs used)."""
        names = list(_RAW_SPECS) if names is None else list(names)
        self._raw = {n: torch.zeros((self.E,) + _RAW_SPECS[n][1], dtype=_RAW_SPECS[n][0], device=self.device)
                     for n in names}
        self._raw_struct = capi.RawOutputs(*[self._raw[n].data_ptr() if n in self._raw else None
                                             for n in capi.RAW_FIELDS])
        return self._raw

    def _raw_ref(self):
        return C.byref(self._raw_struct) if self._raw_struct is not None else None

    def _dev(self, x, dtype):
        if x is None:
            return None
        if isinstance(x, torch.Tensor):
            t = x.to(device=self.device, dtype=dtype).contiguous()
        else:
            t = torch.from_numpy(np.ascontiguousarray(x)).to(device=self.device, dtype=dtype)
        return t

    def buffers(self, out=None):
        """Buffer.buffer of every UE as the reference indexes it (buffer.py:47, index = age): int32
        [E, 25, row_stride], de-rotated from the ring rows by irs_read_buffers."""
        if out is None:
            out = torch.empty((self.E, U, self.row_stride), dtype=torch.int32, device=self.device)
        with torch.cuda.device(self.device):
            capi.check(self.h, self.lib.irs_read_buffers(self.h, out.data_ptr(), self._stream()))
        return out

    @property
    def buf(self):
        return self.buffers()

    def set_episode_order(self, order):
        """enable_random_episodes with a given order: int32 [E, n] (irs_set_episode_order); None = built-in."""
        if order is None:
            self._order = None
            with torch.cuda.device(self.device):
                capi.check(self.h, self.lib.irs_set_episode_order(self.h, None, 0, self._stream()))
            return
        t = self._dev(order, torch.int32)
        if t.dim() != 2 or t.shape[0] != self.E:
            raise ValueError("order must be int32 [num_envs, n]")
        self._order = t
        with torch.cuda.device(self.device):
            capi.check(self.h, self.lib.irs_set_episode_order(self.h, t.data_ptr(), t.shape[1], self._stream()))

    # -- channel_mode 3: device-resident SE trace (channels/quadriga.py) ------------------------
    def set_se_trace(self, se_bank: torch.Tensor, first_episode: int = 0, wrap: bool = False):
        """Bind a float64 [n_episodes, n_steps, 25, 135] SE bank living on this engine's device
        (irs_set_se_trace): slot i = the file of env episode first_episode + i; `wrap` maps episodes onto slots
        modulo n_episodes instead of flagging an episode outside the bank.  The engine keeps a reference so the
        memory outlives its use."""
        if se_bank.dtype != torch.float64 or se_bank.dim() != 4 or tuple(se_bank.shape[2:]) != (U, R):
            raise ValueError("se_bank must be float64 [n_episodes, n_steps, 25, 135]")
        if se_bank.device != self.device or not se_bank.is_contiguous():
            raise ValueError("se_bank must be a contiguous tensor on the engine's device")
        self._se_bank = se_bank
        capi.check(self.h, self.lib.irs_set_se_trace(self.h, se_bank.data_ptr(), se_bank.shape[0], se_bank.shape[1],
                                                     int(first_episode), int(bool(wrap))))

    def power_to_se(self, power: torch.Tensor, layout: int = 0, tx_power: float = 100.0, noise_power: float = 10e-14,
                    out: torch.Tensor | None = None):
        """QuadrigaChannel.step's transform (channels/quadriga.py:56-74) over a device tensor of
        target-cell powers [..., 135, 25] (layout 0, the .mat file order) or [..., 25, 135]
        (layout 1); returns float64 [..., 25, 135] spectral efficiencies."""
        want = (R, U) if layout == 0 else (U, R)
        if power.dtype != torch.float64 or tuple(power.shape[-2:]) != want or power.device != self.device:
            raise ValueError(f"power must be float64 [..., {want[0]}, {want[1]}] on the engine's device")
        power = power.contiguous()
        shape = tuple(power.shape[:-2]) + (U, R)
        if out is None:
            out = torch.empty(shape, dtype=torch.float64, device=self.device)
        elif tuple(out.shape) != shape or out.dtype != torch.float64 or not out.is_contiguous() or out.device != self.device:
            raise ValueError("out must be a contiguous float64 tensor of shape [..., 25, 135] on the engine's device")
        n = power.numel() // (U * R)
        with torch.cuda.device(self.device):
            rc = self.lib.irs_power_to_se(power.data_ptr(), out.data_ptr(), n, int(layout), float(tx_power),
                                          float(noise_power), self._stream())
        capi.check(self.h, rc)
        self._keep_power = power
        return out

    # -- the path ---------------------------------------------------------------------------
    def reset(self, mask=None, initial_episode=-1, initial_episodes=None, se=None):
        """CommunicationEnv.reset for the masked envs (comm_env.py:375-482)."""
        m = self._dev(mask, torch.uint8)
        ie = self._dev(initial_episodes, torch.int32)
        se_t = self._dev(se, torch.float64)
        inj = capi.Inject(se_t.data_ptr() if se_t is not None else None, None)
        self._keep = (m, ie, se_t)
        with torch.cuda.device(self.device):
            capi.check(self.h, self.lib.irs_reset(
                self.h, m.data_ptr() if m is not None else None, int(initial_episode),
                ie.data_ptr() if ie is not None else None, C.byref(inj), self._raw_ref(), self._stream()))

    def _step_graph(self, a0, intra):
        """One step as a replay of the captured graph (native generators / device channel only: nothing is injected)."""
        key = (a0 is not None, intra is not None, id(self._raw_struct))  # the capture holds the raw-output pointers
        if a0 is not None:
            self.v["in_a0"].copy_(a0 if isinstance(a0, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(a0, np.float64)),
                                  non_blocking=True)
        if intra is not None:
            self.v["in_intra"].copy_(intra if isinstance(intra, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(intra, np.int32)),
                                     non_blocking=True)
        if self._graph is None or self._graph_key != key:
            act = capi.Actions(self.v["in_a0"].data_ptr() if a0 is not None else None,
                               self.v["in_intra"].data_ptr() if intra is not None else None)
            inj = capi.Inject(None, None)
            cur = torch.cuda.current_stream(self.device)
            side = torch.cuda.Stream(self.device)
            side.wait_stream(cur)
            g = torch.cuda.CUDAGraph()
            with torch.cuda.device(self.device), torch.cuda.graph(g, stream=side):
                capi.check(self.h, self.lib.irs_step(self.h, C.byref(act), C.byref(inj), self._raw_ref(),
                                                     torch.cuda.current_stream(self.device).cuda_stream))
            cur.wait_stream(side)
            self._graph, self._graph_key, self._graph_keep = g, key, (act, inj)
        self._graph.replay()

    def step(self, a0=None, intra=None, se=None, traffic=None, sched_bits=None):
        """CommunicationEnv.step for every env (comm_env.py:271-373).  Asynchronous.  `sched_bits` (int32/uint32
        [E, 25, 5], bit r of a UE's 135-bit row = RB r) gives the scheduling matrix directly instead of an action."""
        if self.graph and se is None and traffic is None and sched_bits is None:
            return self._step_graph(a0, intra)
        a0_t, in_t = self._dev(a0, torch.float64), self._dev(intra, torch.int32)
        se_t, tr_t = self._dev(se, torch.float64), self._dev(traffic, torch.float64)
        sb_t = None
        if sched_bits is not None:
            sb = sched_bits if isinstance(sched_bits, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(sched_bits).view(np.int32))
            sb_t = sb.to(device=self.device).contiguous()
            if tuple(sb_t.shape) != (self.E, U, 5) or sb_t.element_size() != 4:
                raise ValueError("sched_bits must be 32-bit words [num_envs, 25, 5]")
        act = capi.Actions(a0_t.data_ptr() if a0_t is not None else None, in_t.data_ptr() if in_t is not None else None,
                           sb_t.data_ptr() if sb_t is not None else None)
        inj = capi.Inject(se_t.data_ptr() if se_t is not None else None, tr_t.data_ptr() if tr_t is not None else None)
        self._keep = (a0_t, in_t, se_t, tr_t, sb_t)
        with torch.cuda.device(self.device):
            capi.check(self.h, self.lib.irs_step(self.h, C.byref(act), C.byref(inj), self._raw_ref(), self._stream()))

    def step_profiled(self, a0=None, intra=None, se=None, traffic=None):
        """One step with CUDA events around each kernel; returns their durations in ms
        (action->RB, per-UE rows, obs/reward/rollover).  Synchronous; for measurement."""
        a0_t, in_t = self._dev(a0, torch.float64), self._dev(intra, torch.int32)
        se_t, tr_t = self._dev(se, torch.float64), self._dev(traffic, torch.float64)
        act = capi.Actions(a0_t.data_ptr() if a0_t is not None else None, in_t.data_ptr() if in_t is not None else None)
        inj = capi.Inject(se_t.data_ptr() if se_t is not None else None, tr_t.data_ptr() if tr_t is not None else None)
        ms = (C.c_float * 3)()
        with torch.cuda.device(self.device):
            capi.check(self.h, self.lib.irs_step_profiled(self.h, C.byref(act), C.byref(inj), self._raw_ref(),
                                                          self._stream(), ms))
        return [float(x) for x in ms]

    def run(self, n_steps: int):
        with torch.cuda.device(self.device):
            capi.check(self.h, self.lib.irs_run(self.h, int(n_steps), self._stream()))

    def step_host(self, a0_host, intra_host, surface):
        """irs_step_host: host actions in, host surface out (synchronous)."""
        act = capi.Actions(a0_host.data_ptr() if a0_host is not None else None,
                           intra_host.data_ptr() if intra_host is not None else None)
        with torch.cuda.device(self.device):
            capi.check(self.h, self.lib.irs_step_host(self.h, C.byref(act), C.byref(surface), self._stream()))

    def reset_host(self, initial_episode, surface):
        with torch.cuda.device(self.device):
            capi.check(self.h, self.lib.irs_reset_host(self.h, int(initial_episode), C.byref(surface), self._stream()))

    def episode_metrics(self):
        out = np.zeros(capi.NUM_METRICS)
        with torch.cuda.device(self.device):
            capi.check(self.h, self.lib.irs_episode_metrics(self.h, out.ctypes.data, self._stream()))
        return out

    def episode_metrics_device(self):
        """The same sums left on the device (a float64 [16] view of the state blob, valid until the next call):
        no host round trip - distributed.all_reduce_metrics reduces it in place over NCCL."""
        with torch.cuda.device(self.device):
            capi.check(self.h, self.lib.irs_episode_metrics(self.h, None, self._stream()))
        return self.metrics_total

    def check_errors(self):
        first, code = C.c_int32(-1), C.c_int32(0)
        with torch.cuda.device(self.device):
            rc = self.lib.irs_check_errors(self.h, C.byref(first), C.byref(code), self._stream())
        if rc == capi.IRS_ERR_ENV:
            return first.value, code.value
        capi.check(self.h, rc)
        return None

    def scenario_rows(self):
        """The scenario row every env is playing, as a NumPy record array (scenario.SCENARIO_DTYPE).  With
        assoc_mode 2 these are the rows generated on the device."""
        from .scenario import SCENARIO_DTYPE
        n = len(self.bank_array) + (2 * self.E if self.cfg.assoc_mode == 2 else 0)
        raw = self.blob[self.layout.bank:self.layout.bank + n * SCENARIO_DTYPE.itemsize].cpu().numpy()
        return raw.view(SCENARIO_DTYPE)[self.v["scen"].cpu().numpy()]

    def surface(self):
        """The env surface tensors (device views into the state blob)."""
        v = self.v
        return {"obs0": v["out_obs0"], "obs_i": v["out_obsi"], "mask0": v["out_mask0"], "mask_i": v["out_maski"],
                "reward": v["out_reward"], "done": v["out_done"], "sorted_slices": v["sorted"]}

    def close(self):
        if getattr(self, "h", None) is not None and self.h:
            torch.cuda.synchronize(self.device)
            self.lib.irs_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
