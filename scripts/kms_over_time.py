"""Per-kernel ms as the simulation progresses (diagnostic)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, numpy as np
import bench
from intent_radio_sched_multi_slice_b200.engine import BatchEngine
a = bench.parse()
E = a.envs
bank, wl = bench.make_bank(a.depth)
eng = BatchEngine(bank, E, depth_cap=max(bank.max_depth(), 101), **bench.engine_kwargs(a, 0))
dev = eng.device
gid = torch.arange(E, device=dev, dtype=torch.int32)
eng.reset(initial_episodes=(gid % 200).to(torch.int32), initial_episode=0)
g = torch.Generator(device=dev); g.manual_seed(1234)
pool = [(torch.rand(E, 5, generator=g, device=dev, dtype=torch.float64) * 2 - 1,
         torch.randint(0, 3, (E, 5), generator=g, device=dev, dtype=torch.int32)) for _ in range(4)]
for k in range(a.steps):
    if k in (2, 5, 10, 20, 50, 100, 200, 300, 500, 800) or k == a.steps - 1:
        ms = eng.step_profiled(*pool[k % 4])
        nz = float((eng.buf != 0).float().mean().item()) if E <= 65536 else -1
        print(k, ["%.3f" % x for x in ms], "nonzero slot frac %.4f" % nz, "mean tot %.1f" % eng.v["tot"].float().mean().item(), flush=True)
    else:
        eng.step(*pool[k % 4])
