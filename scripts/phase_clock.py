"""Where a k_env_step block spends its time (debug build only).

    IRS_LIB_PATH=$PWD/build_variants/lib_phaseclock.so python scripts/phase_clock.py [envs]

build_variants/lib_phaseclock.so is the library compiled with -DIRS_PHASE_CLOCK (IRS_NVCC_EXTRA): lane 0 of each
warp stamps the SM clock at the phase boundaries of ue_env_body<FULL_STEP>; this script runs BASELINE config 2 the
way bench.py does (L2 flushed between steps), reads the stamps of the last launch and prints the median cycles per
phase over the blocks."""
import ctypes as C
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from intent_radio_sched_multi_slice_b200 import capi  # noqa: E402
from intent_radio_sched_multi_slice_b200.engine import BatchEngine  # noqa: E402

E = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
warm = os.environ.get("PHASE_WARM", "0") == "1"
dev = torch.device("cuda", 0)
bank, _ = bench.make_bank(0)  # the mult_slice scenario mix, as bench.py --workload cfg2
eng = BatchEngine(bank, E, device=dev, depth_cap=max(bank.max_depth(), 101), channel_mode=2, traffic_mode=1, sort_slices=1,
                  fixed_intra=-1, policy=0, assoc_mode=0, scen_mod=200, max_steps=1000, max_episodes=10000, auto_reset=1,
                  seed=10, env_id_base=0)
gid = torch.arange(E, device=dev, dtype=torch.int32)
eng.reset(initial_episodes=(gid % 200).to(torch.int32), initial_episode=0)
g = torch.Generator(device=dev)
g.manual_seed(1234)
pool = [(torch.rand(E, 5, generator=g, device=dev, dtype=torch.float64) * 2 - 1,
         torch.randint(0, 3, (E, 5), generator=g, device=dev, dtype=torch.int32)) for _ in range(4)]
flush = torch.empty(int(256e6), dtype=torch.uint8, device=dev)
for k in range(300):
    eng.step(*pool[k % 4])
lib = capi.load()
fn = lib.irs_debug_phase_clock
fn.argtypes = [C.c_void_p, C.c_int]
names = ["phase 0 + action (warp 0) / wait (warp 1)", "normal-table wait", "pass A (mu, tail quad, Poisson)",
         "pair passes", "buffer arithmetic + ring + stores + barrier", "tail (obs, reward, rollover)"]
acc = []
for k in range(20):
    if not warm:
        flush.fill_(k & 255)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    eng.step(*pool[k % 4])
    ev1.record()
    torch.cuda.synchronize()
    out = np.zeros((E, 4, 8), np.int64)
    rc = fn(out.ctypes.data, E)
    assert rc == 0, rc
    acc.append((ev0.elapsed_time(ev1), out))
ms = np.median([a for a, _ in acc])
print(f"envs {E}  L2 {'warm' if warm else 'flushed'}  step {ms * 1e3:.1f} us (CUDA events, eager launch)")
o = np.stack([b for _, b in acc[5:]])  # [runs, E, warp, stamp]
# SM clocks are per SM: only differences inside a block are meaningful
for w in (0, 1):
    d = np.diff(o[:, :, w, :7], axis=-1).astype(np.float64)
    n = 6 if w == 0 else 5
    print(f"warp {w}: median cycles per phase (p10 / p50 / p90 over blocks and runs)")
    for i in range(n):
        x = d[..., i].ravel()
        print(f"  {names[i]:52s} {np.percentile(x, 10):8.0f} {np.percentile(x, 50):8.0f} {np.percentile(x, 90):8.0f}")
    tot = (o[:, :, w, n] - o[:, :, w, 0]).ravel()
    print(f"  {'block total':52s} {np.percentile(tot, 10):8.0f} {np.percentile(tot, 50):8.0f} {np.percentile(tot, 90):8.0f}")
sub = {2: ("action core", ["ue_slice / masks / actions loaded", "scores_to_rbs (inter-slice RBs)", "round-robin shares, ta, group_max",
                            "snt, w, group_sum, proportional-fair rounding", "group_sum, max-throughput rounding",
                            "select, prefix, RB-sum check"]),
       3: ("tail core", ["window scalars, ue_slice masks, intent_drift_ue", "4 slice means + 2 int sums", "slice view, sort_slices",
                         "obs / mask stores (stamp 4 - 3)", "reward", "episode metrics"])}
for row, (title, nm) in sub.items():
    d = np.diff(o[:, :, row, :7], axis=-1).astype(np.float64)
    print(f"{title}: cycles per section (p10 / p50 / p90)")
    for i in range(6):
        x = d[..., i].ravel()
        print(f"  {nm[i]:52s} {np.percentile(x, 10):8.0f} {np.percentile(x, 50):8.0f} {np.percentile(x, 90):8.0f}")
    tot = (o[:, :, row, 6] - o[:, :, row, 0]).ravel()
    print(f"  {'core total':52s} {np.percentile(tot, 10):8.0f} {np.percentile(tot, 50):8.0f} {np.percentile(tot, 90):8.0f}")
x = (o[:, :, 2, 0] - o[:, :, 0, 0]).ravel()
print(f"entry -> action core (phase 0 loads): p50 {np.percentile(x, 50):.0f}")
x = (o[:, :, 0, 1] - o[:, :, 2, 6]).ravel()
print(f"action core end -> barrier passed: p50 {np.percentile(x, 50):.0f}")
