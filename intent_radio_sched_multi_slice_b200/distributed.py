"""Multi-GPU: env instances shard across ranks; the only collective is the metric reduction.

Env instances share no state (`create_scenario` builds private UEs/plugins per env,
sixg_radio_mgmt/sixg_radio_mgmt/comm_env.py:616-682), so rank r simply owns the contiguous
global env ids [r*E_local, (r+1)*E_local) - `env_id_base` keys the counter-based generators
by GLOBAL id, so results do not depend on the number of GPUs.  No data-path collective exists.
The reference's per-episode history files (metrics.py:73-100) are replaced by per-env
accumulators; their global sum is ONE small all-reduce (16 float64) per episode: NCCL over
NVLink on GPUs, gloo in the CPU tests.
"""
from __future__ import annotations

import os

import numpy as np
import torch
import torch.distributed as dist

from . import capi


def rank_world():
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))


def shard(total_envs: int, rank: int | None = None, world: int | None = None):
    """(env_id_base, local_envs) of this rank: contiguous, remainder spread over the first ranks."""
    if rank is None or world is None:
        rank, world = rank_world()
    if total_envs < world:
        raise ValueError("fewer envs than ranks")
    q, r = divmod(total_envs, world)
    local = q + (1 if rank < r else 0)
    base = rank * q + min(rank, r)
    return base, local


def init(backend: str | None = None, device: torch.device | None = None):
    """Initialise torch.distributed from the torchrun environment (127.0.0.1 rendezvous)."""
    if dist.is_initialized() or int(os.environ.get("WORLD_SIZE", "1")) <= 1:
        return
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    backend = backend or ("nccl" if torch.cuda.is_available() else "gloo")
    kw = {"device_id": device} if (backend == "nccl" and device is not None) else {}
    dist.init_process_group(backend, **kw)


def all_reduce_metrics(local, device=None) -> dict:
    """Sum the 16 episode-metric accumulators over all ranks.  `local`: a tensor already on the device - e.g.
    BatchEngine.episode_metrics_device(), the [16] result of the fixed-order reduction kernel inside the state blob,
    which NCCL then reduces IN PLACE on the same stream (no host round trip before the collective) - or an
    array-like [16].  Counters (packets, steps, violations, episodes) are integer-valued float64 < 2^53 and therefore
    exact in any reduction order; the float sums (rewards, Mbit, distances) agree to rounding."""
    t = local if isinstance(local, torch.Tensor) else torch.as_tensor(np.asarray(local), dtype=torch.float64)
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        if dist.get_backend() == "nccl" and not t.is_cuda:
            t = t.to(device if device is not None else torch.device("cuda", torch.cuda.current_device()))
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return dict(zip(capi.METRIC_NAMES, t.cpu().tolist()))


def max_over_ranks(x: float, device=None) -> float:
    """Device-timed durations are reported as the max over ranks."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return float(x)
    t = torch.tensor([x], dtype=torch.float64)
    if dist.get_backend() == "nccl":
        t = t.to(device if device is not None else torch.device("cuda", torch.cuda.current_device()))
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def pin_to_device_numa(device_index: int) -> bool:
    """Restrict this process to the CPUs NVML reports as closest to the GPU (its NUMA node), so that the pinned host
    buffers of `irs_step_host` are allocated on that node.  Matters when all 8 GPUs of a box stream their surface to the
    host at once (8 x 81 MB per step at 65,536 envs per GPU): without it the copies cross the socket interconnect.
    Returns False (and changes nothing) when NVML is not available."""
    try:
        import pynvml

        pynvml.nvmlInit()
        try:
            p = torch.cuda.get_device_properties(device_index)
            h = pynvml.nvmlDeviceGetHandleByPciBusId(f"{p.pci_domain_id:08x}:{p.pci_bus_id:02x}:{p.pci_device_id:02x}.0")
        except Exception:
            h = pynvml.nvmlDeviceGetHandleByIndex(device_index)
        pynvml.nvmlDeviceSetCpuAffinity(h)
        return True
    except Exception:
        return False
