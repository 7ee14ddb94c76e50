#!/usr/bin/env python
"""bench.py - UE-steps/s of the batched CommunicationEnv.step path (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]
                    [--envs E_per_gpu] [--depth D] [--workload mix|cfg2]

One "step" = one env step (action->RB, arrivals, SE, throughput, buffers, obs, masks, reward,
rollover) of every env of the batch.  Prints ONE JSON line (rank 0).  See DESIGN.md
"Measurement" for the definitions of value / e2e / roofline / cpu_baseline.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

U = 25
FIXED_ROW_BYTES = 140  # per UE row besides the buffer: int/f64 records R+W, window-ring slot R+W, RB run, sum(age*n) R+W


def parse():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=200)
    p.add_argument("--warmup", type=int, default=20)
    p.add_argument("--impl", default="b200", choices=["b200", "reference"])
    p.add_argument("--envs", type=int, default=65536, help="env instances per GPU (weak scaling)")
    p.add_argument("--fill", type=int, default=-1,
                   help="untimed steps before the warm-up that fill the packet buffers to their steady state "
                        "(default: the deepest row of the workload)")
    p.add_argument("--depth", type=int, default=0,
                   help="buffer depth of the sweep workload (BASELINE config 5: 401, 1024, 4096); default 1024")
    p.add_argument("--workload", default="sweep", choices=["sweep", "mix", "cfg2", "trace"],
                   help="sweep (default): BASELINE config 5, 65,536 envs x 25 UEs x deep buffers, the HBM-roofline "
                        "configuration the metric is quoted on at 1/2/4/8 GPUs; mix: the 200-scenario mult_slice mix "
                        "(D 21..401); cfg2: BASELINE config 2 (1,024 envs); trace: QuadrigaChannel-shaped SE bank in HBM")
    p.add_argument("--trace-steps", type=int, default=16, help="frames per episode of the SE bank (wraps around)")
    p.add_argument("--cpu-envs", type=int, default=0,
                   help="env instances of the CPU baseline sample (default 512 for the sweep, whose buffers take "
                        "thousands of steps to fill, else 2048)")
    p.add_argument("--d2h-probe", action="store_true",
                   help="no kernels: every rank copies one step's float64 surface (81.2 MB at 65,536 envs) device->pinned "
                        "host in a loop; prints the aggregate GB/s, the ceiling of the e2e figure at N GPUs")
    p.add_argument("--no-graph", action="store_true", help="small batches: eager launches instead of the graph replay (profiling)")
    p.add_argument("--no-cpu-baseline", action="store_true")
    p.add_argument("--no-e2e", action="store_true")
    a = p.parse_args()
    if a.workload == "sweep" and a.depth <= 0:
        a.depth = 1024
    if a.workload != "sweep" and a.depth > 0:
        a.workload = "sweep"  # --depth D alone selects the sweep
    return a


def make_bank(depth: int):
    from intent_radio_sched_multi_slice_b200 import scenario as scn

    bank = scn.ScenarioBank()
    st, us = scn.generate_scenarios(200, seed=10)
    if depth <= 0:
        for i in range(len(st)):
            bank.add_typed(st[i], us[i])
        return bank, "mult_slice scenario mix (200 scenarios, 10 slice types, D 21..401)"
    # BASELINE config 5: every UE gets a deep profile derived from uav_app_case_1
    for i in range(len(st)):
        req = {}
        sua = np.zeros((5, 25))
        for s in range(5):
            req[f"slice_{s}"] = scn.slice_req_from_type(5)
            req[f"slice_{s}"]["ues"]["buffer_latency"] = depth - 1
            req[f"slice_{s}"]["parameters"]["par1"]["value"] = depth // 2
            sua[s, 5 * s:5 * s + 5] = 1
        bank.add(req, sua)
    return bank, (f"BASELINE config 5 (scale sweep): envs x 25 UEs x packet-buffer depth {depth} "
                  "(every UE on the uav_app_case_1 profile, native mimic_quadriga SE + Poisson traffic, random actions)")


def engine_kwargs(args, rank):
    return dict(channel_mode=3 if args.workload == "trace" else 2, traffic_mode=1, sort_slices=1, fixed_intra=-1, policy=0, assoc_mode=0,
                scen_mod=200, max_steps=episode_steps(args), max_episodes=10000, auto_reset=1, seed=10,
                env_id_base=rank * args.envs)


def episode_steps(args):
    """mult_slice.yml's 1000-step episodes; the depth sweep never rolls over, so that its rows stay
    fully live (aged across the whole row, expiring at the old end) for the whole measurement."""
    return 1_000_000 if args.workload == "sweep" else 1000


def fill_to_steady_state(args, bank, eng, pool):
    """Untimed steps that bring the packet buffers to their steady state before anything is measured.
    The row kernel only touches a row's live extent (slots up to the oldest packet), which keeps growing
    for about 1.7 x depth steps in the sweep profile (backlog grows, the capacity clip sets in, then the
    oldest packets reach the expiry age).  Steps in blocks of 128 until the mean live extent moves by
    less than 0.5 % per block (or --fill steps exactly, when given)."""
    if args.fill >= 0:
        for k in range(args.fill):
            eng.step(*pool[k % 4])
        return args.fill
    cap = min(4 * bank.max_depth(), episode_steps(args) // 2)  # mult_slice episodes: stay clear of the rollover
    n, prev = 0, -1.0
    while n < cap:
        for k in range(128):
            eng.step(*pool[k % 4])
        n += 128
        live = float(((eng.v["path"] >> 2) & 0x3FFF).float().mean().item())
        if prev >= 0 and abs(live - prev) <= 0.005 * max(live, 1.0):
            break
        prev = live
    return n


class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), line.strip()))

    def stop(self, t0, t1):
        if not self.proc:
            return None
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], 0, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for t, line in self.rows:
            f = [x.strip() for x in line.split(",")]
            if len(f) < 6:
                continue
            try:
                mx = max(mx, int(f[1]))
                if t0 - 0.05 <= t <= t1 + 0.15:
                    sm.append(int(f[0]))
                    reasons |= {n for n, v in zip(names, f[2:6]) if v.lower().startswith("active")}
            except ValueError:
                continue
        if not sm:
            sm = [int(x.split(",")[0]) for _, x in self.rows[-3:] if x.split(",")[0].strip().isdigit()]
        return {"sm_mhz": int(np.median(sm)) if sm else None, "sm_max_mhz": mx or None, "reasons": sorted(reasons),
                "samples": len(sm)}


def cpu_fill_steps(args, bank):
    """Steps the GPU arm's plateau detection settles on (config.fill_steps of its bench lines): the CPU
    sample is brought to the same steady state before it is timed (its drain loop is FASTER on full rows)."""
    if args.fill >= 0:
        return args.fill
    if args.workload != "sweep":
        return 512
    return {401: 1664, 1024: 2944, 4096: 4736}.get(bank.max_depth(), 3 * bank.max_depth())


def cpu_oracle_run(args, bank, n_envs, warmup, steps=None, budget_s=12.0, fill=0, min_wall_s=0.0):
    """The oracle port (oracle/irs_oracle.c, OpenMP over envs) on the host cores: same scenario
    bank, generators, agent surface and action distribution as the GPU arm."""
    n_envs = n_envs or (512 if args.workload == "sweep" else 2048)
    from oracle import oracle

    kw = engine_kwargs(args, 0)
    kw["env_id_base"] = 0
    cores = os.cpu_count() or 1
    cores = oracle.set_threads(cores)  # explicit: torchrun exports OMP_NUM_THREADS=1
    env = oracle.OracleEnv(bank.array(), bank.poisson_cdf(), n_envs,
                           depth_cap=max(bank.max_depth(), 101), **kw)
    rng = np.random.default_rng(1234)
    pool = [(rng.uniform(-1, 1, (n_envs, 5)), rng.integers(0, 3, (n_envs, 5)).astype(np.int32)) for _ in range(4)]
    m = np.ones(n_envs, np.uint8)
    if args.workload == "trace":  # one bank slot per env, like the GPU arm
        power = 10.0 ** rng.uniform(-13.5, -9.5, (n_envs, args.trace_steps, 135, 25))
        env.set_se_trace(oracle.power_to_se(power, layout=0), wrap=True)
        del power
    for e0 in range(200):  # stagger the scenarios like the GPU arm: env e starts at episode e % 200
        m[:] = (np.arange(n_envs) % 200) == e0
        if m.any():
            env.reset(mask=m, initial_episode=e0)
    for k in range(fill + warmup):
        env.step(*pool[k % 4])
    if steps is None:
        t = time.perf_counter()
        for k in range(3):
            env.step(*pool[k % 4])
        per = (time.perf_counter() - t) / 3
        steps = int(max(5, min(2000, budget_s / max(per, 1e-6))))
    blocks, wall = [], 0.0
    while not blocks or (wall < min_wall_s and len(blocks) < 200):
        t = time.perf_counter()
        for k in range(steps):
            env.step(*pool[k % 4])
        blocks.append(time.perf_counter() - t)
        wall += blocks[-1]
    dt = float(np.median(blocks))
    env.close()
    rep = f" (median of {len(blocks)} blocks of {steps} steps, {wall:.1f} s measured)" if len(blocks) > 1 else ""
    return {"value": n_envs * U * steps / dt, "unit": "UE-steps/s", "cores": cores, "kind": "port",
            "sample": f"{n_envs} envs x {steps} steps of the same workload after {fill} untimed fill steps, "
                      f"oracle/irs_oracle.c with OpenMP on {cores} threads" + rep, "envs": n_envs,
            "seconds": dt, "ms_per_step": dt / steps * 1e3, "steps": steps, "blocks": len(blocks)}


def run_reference(args):
    """--impl reference: the reference is pure Python and is not present on the GPU box
    (/root/reference does not travel), so its CPU implementation of the path is timed as the C
    restatement that tests/ pin against it (kind "port"), on all host threads."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    bank, wl = make_bank(args.depth)
    n_envs = args.cpu_envs
    # the --steps block is repeated until at least 4 s of wall time have been measured and the MEDIAN block is
    # reported: a 20-step block is 85 ms on this sample, and one such window moved the driver's ratio by 30 %
    r = cpu_oracle_run(args, bank, n_envs, args.warmup, steps=args.steps, fill=cpu_fill_steps(args, bank), min_wall_s=4.0)
    n_envs = r["envs"]
    line = {
        "impl": "reference", "metric": "UE-steps/sec (env-steps/sec x UEs)", "value": r["value"], "unit": "UE-steps/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": r["ms_per_step"],
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64+i32", "data": "synthetic",
        "config": {"workload": wl, "envs_per_step": n_envs, "ues": U, "surface": "IBSched dict obs + masks + 6 rewards",
                   "note": "bounded CPU sample of the GPU arm's workload; per-env cost is independent of batch size"},
        "cpu_baseline": {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")},
        "e2e": {"value": r["value"], "unit": "UE-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


def d2h_copy_rate(torch, dev, nbytes, barrier, world, dist, reps=20, pieces=1):
    """Device -> pinned host copies of `nbytes` (as `pieces` back-to-back copies of nbytes / pieces) in a loop, no
    kernels: GB/s of this rank and of all ranks together (time = max over ranks).  What the host side can absorb is the
    ceiling of the e2e figure."""
    src = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    dst = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
    per = (nbytes + pieces - 1) // pieces
    spans = [(i * per, min(nbytes, (i + 1) * per)) for i in range(pieces)]

    def once():
        for a, b in spans:
            dst[a:b].copy_(src[a:b], non_blocking=True)

    for _ in range(3):
        once()
    barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for _ in range(reps):
        once()
    ev1.record()
    barrier()
    ms = ev0.elapsed_time(ev1)
    if world > 1:
        t = torch.tensor([ms], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    rate = nbytes * reps / (ms * 1e-3) / 1e9
    return {"gbs_per_gpu": rate, "gbs_total": rate * world, "bytes_per_copy": nbytes, "pieces": pieces, "copies": reps,
            "ms_per_copy": ms / reps}


def ncu_facts(workload, envs, depth):
    """What one `ncu --set full` capture of k_ue_env on this workload measured (profiles/ue_env_ncu.json, written by
    scripts/ncu_to_json.py from the committed captures): warp instructions and DRAM bytes per launch are properties of
    the launch, not of the run, so they are combined with the LIVE kernel time below."""
    try:
        for tj in json.load(open(os.path.join(ROOT, "profiles", "ue_env_ncu.json"))):
            if tj.get("envs") == envs and tj.get("depth", 0) == depth and tj.get("workload") == workload:
                return tj
    except Exception:
        pass
    return None


def run_b200(args):
    import torch
    import torch.distributed as dist

    from intent_radio_sched_multi_slice_b200 import capi
    from intent_radio_sched_multi_slice_b200.engine import BatchEngine

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    numa_pinned = False
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
        if os.environ.get("IRS_BENCH_NUMA", "1") != "0":  # host buffers next to the GPU (see distributed.pin_to_device_numa)
            from intent_radio_sched_multi_slice_b200 import distributed as D0
            numa_pinned = D0.pin_to_device_numa(local)
    if args.workload == "cfg2":
        args.envs = 1024
    E, K, Wm = args.envs, args.steps, max(args.warmup, 3)
    bank, wl = make_bank(args.depth)
    # small batches are launch-bound: the engine replays one captured graph per step (BatchEngine(graph=True))
    eng = BatchEngine(bank, E, device=dev, depth_cap=max(bank.max_depth(), 101), graph=E <= 4096 and not args.no_graph, **engine_kwargs(args, rank))
    gid = torch.arange(E, device=dev, dtype=torch.int32) + rank * E
    trace_bytes = 0
    if args.workload == "trace":
        # QuadrigaChannel-shaped input: every env reads its own episode file, so each step streams
        # E distinct 25 x 135 float64 frames from HBM.  Synthetic target-cell powers (10^U(-13.5,-9.5) W)
        # go through the library's own transform (channels/quadriga.py:56-74) chunk by chunk.
        T = args.trace_steps
        se_bank = torch.empty((E, T, U, 135), dtype=torch.float64, device=dev)
        g0 = torch.Generator(device=dev)
        g0.manual_seed(77 + rank)
        chunk = max(1, (1 << 28) // (T * U * 135))
        for c0 in range(0, E, chunk):
            n = min(chunk, E - c0)
            power = torch.pow(10.0, torch.rand((n, T, 135, U), generator=g0, device=dev, dtype=torch.float64) * 4.0 - 13.5)
            eng.power_to_se(power, layout=0, out=se_bank[c0:c0 + n])
        torch.cuda.synchronize(dev)
        del power
        eng.set_se_trace(se_bank, wrap=True)
        trace_bytes = se_bank.numel() * 8
        wl = (f"SE trace channel (QuadrigaChannel-shaped): one {T}-frame episode file per env resident in HBM, "
              "mult_slice scenario mix (200 scenarios, D 21..401)")
    # env e starts at episode gid: scenario gid % 200 (and SE bank slot gid % E in the trace workload)
    eng.reset(initial_episodes=gid.to(torch.int32) if args.workload == "trace" else (gid % 200).to(torch.int32),
              initial_episode=0)
    g = torch.Generator(device=dev)
    g.manual_seed(1234 + rank)
    pool = [(torch.rand(E, 5, generator=g, device=dev, dtype=torch.float64) * 2 - 1,
             torch.randint(0, 3, (E, 5), generator=g, device=dev, dtype=torch.int32)) for _ in range(4)]
    state_bytes = eng.layout.total_bytes
    flush = None
    if state_bytes < 2 * 126e6:  # state fits L2: flush it between timed steps (config 2)
        flush = torch.empty(int(256e6), dtype=torch.uint8, device=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    surface_bytes = E * (50 * 8 + 95 * 8 + 5 + 25 + 6 * 8 + 1)  # float64 obs0 / obs_i, int8 masks, 6 rewards, done
    if args.d2h_probe:
        r = d2h_copy_rate(torch, dev, surface_bytes, barrier, world, dist, reps=50)
        r["by_pieces"] = {str(p_): d2h_copy_rate(torch, dev, surface_bytes, barrier, world, dist, reps=30, pieces=p_)["gbs_per_gpu"]
                          for p_ in (4, 16, 64)}
        r["by_size_mb"] = {str(mb): d2h_copy_rate(torch, dev, mb << 20, barrier, world, dist, reps=30)["gbs_per_gpu"]
                           for mb in (1, 4, 16, 64, 256)}
        if rank == 0:
            print(json.dumps({"probe": "d2h", "n_gpus": world, "envs_per_gpu": E, **r,
                              "note": "device -> pinned host, no kernels running; aggregate = per-GPU rate x ranks, timed as "
                                      "the max over ranks", "host_numa_pinned": numa_pinned}))
        if world > 1:
            dist.destroy_process_group()
        return

    n_fill = fill_to_steady_state(args, bank, eng, pool)
    for k in range(Wm):
        eng.step(*pool[k % 4])
    barrier()
    sampler = ClockSampler(local)
    sampler.start()
    time.sleep(0.25)

    # ---- value: K steps, inputs resident in HBM, CUDA events on the launching stream ----------
    t0 = time.time()
    if flush is None:
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ev0.record()
        for k in range(K):
            eng.step(*pool[k % 4])
        ev1.record()
        barrier()
        ms_total = ev0.elapsed_time(ev1)
    else:
        evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
        for k in range(K):
            flush.fill_(k & 255)
            evs[k][0].record()
            eng.step(*pool[k % 4])
            evs[k][1].record()
        barrier()
        ms_total = sum(a.elapsed_time(b) for a, b in evs)
    t1 = time.time()
    clocks = sampler.stop(t0, t1)

    # ---- roofline of the dominant kernel (k_ue_step): per-kernel CUDA events, live ----------
    arr = bank.array()
    depth_tab = np.full((len(arr), U), 101, np.int64)  # buffer depth of every (scenario, UE) row
    for i in range(len(arr)):
        for u in range(U):
            s_ = int(arr["ue_slice"][i, u])
            if s_ >= 0:
                depth_tab[i, u] = int(arr["sl"]["buffer_latency"][i, s_]) + 1
    depth_tab_t = torch.from_numpy(depth_tab).to(dev)
    assoc_tab_t = torch.from_numpy((arr["ue_slice"][:, :U] >= 0).astype(np.int64)).to(dev)
    P = max(3, min(K, 20))
    kms = np.zeros(3)
    ue_bytes = 0.0
    path_frac = np.zeros(4)
    live_frac = 0.0
    for k in range(P):
        if flush is not None:
            flush.fill_(k & 255)
        kms += np.array(eng.step_profiled(*pool[k % 4]))
        # bytes the row kernel had to move in THIS launch: per-UE records for every row, plus the
        # buffer row by the path the kernel took (DESIGN.md "Algorithmic bytes")
        word = eng.v["path"].long()  # path | live extent << 2 | slots touched << 16 (irs_common.cuh)
        path, dep = word & 3, word >> 16
        live_frac += float(((word >> 2) & 0x3fff).float().sum().item() / depth_tab_t[eng.v["scen"].long()].sum().item())
        row_b = 4 * dep  # every 4-byte ring-slot access the step made (expiring slot, look-ahead, drain, new arrivals)
        if args.workload == "trace":  # + the 135 float64 SE values of every associated UE's row
            row_b = row_b + 1080 * assoc_tab_t[eng.v["scen"].long()]
        ue_bytes += float((row_b + FIXED_ROW_BYTES).sum().item())
        path_frac += np.array([(path == i).float().mean().item() for i in range(4)])
    kms /= P
    ue_bytes /= P
    path_frac /= P
    live_frac /= P
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    achieved = ue_bytes / (kms[1] * 1e-3) / 1e9
    facts = ncu_facts(args.workload, E, args.depth)
    traffic = facts["dram_bytes_per_launch"] if facts else None
    rows = E * U
    hbm = {"achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic,
           "peak_source": "MEASURED_PEAKS.json hbm_gbs (measured copy)" if peaks else "fallback 6650 GB/s",
           "algorithmic_bytes_per_launch": ue_bytes,
           "bytes_per_ue_row": ("%d (per-UE records) + ring row: 4 per slot access (expiring slot, the 4-slot look-ahead at "
                                "the old end, slots the drain changes, the arrivals slot)" % FIXED_ROW_BYTES)
           + (" + 1080 (SE row) per associated UE" if args.workload == "trace" else ""),
           # what SURVEY 8(d)'s dense age-shift formulation (np.roll of every row, buffer.py:61) would move: stated for
           # reference, never divided by a time (the ring rows do not move these bytes)
           "survey_8d_dense_shift_bytes_per_launch": float(sum(8 * int(d) + 136 for d in depth_tab[(np.arange(E) + rank * E) % len(arr)].ravel()))
           if args.workload != "trace" else None}
    sm_mhz = (clocks or {}).get("sm_mhz") or float(peaks.get("sm_max_mhz", 1965.0))
    n_sm = torch.cuda.get_device_properties(dev).multi_processor_count
    if args.workload != "trace" and facts and facts.get("inst_per_launch"):
        # With ring rows the kernel moves ~0.45 KB per UE row and draws 135 normals for it: it is bound by instruction
        # issue (one warp instruction per SM sub-partition and clock), not by HBM.  Achieved = warp instructions of
        # one launch (ncu smsp__inst_executed.sum of the same launch shape) / the LIVE kernel time.
        ginst = facts["inst_per_launch"] / (kms[1] * 1e-3) / 1e9
        peak_i = n_sm * 4 * sm_mhz * 1e6 / 1e9
        roofline = {"bound": "issue", "kernel": "k_ue_env", "achieved": ginst, "peak": peak_i, "unit": "G warp-inst/s",
                    "frac": ginst / peak_i, "traffic": traffic,
                    "peak_source": f"{n_sm} SMs x 4 schedulers x {sm_mhz:.0f} MHz (SM clock sampled during the timed region)",
                    "inst_per_launch": facts["inst_per_launch"], "ncu_issue_active_pct": facts.get("issue_active_pct"),
                    "ncu_source": facts.get("source"), "hbm": hbm}
    else:
        roofline = dict(hbm, bound="hbm", kernel="k_ue_env_mem" if args.workload == "trace" else "k_ue_env")
    roofline.update({"kernel_ms": kms[1], "kernel_ms_all": {"k_action": kms[0], "k_ue_env": kms[1], "k_env_tail": kms[2]},
                     "row_paths": {"no_old_packets": path_frac[0], "window_cleared": path_frac[1], "lane_drain": path_frac[2],
                                   "warp_drain": path_frac[3]},
                     "live_extent_fraction": live_frac})

    # ---- e2e: host actions in, host surface out, through irs_step_host ------------------------
    e2e = e2e_f32 = None
    if not args.no_e2e:
        Ke = max(3, min(K, 30))
        pin = lambda *s, dt: torch.empty(*s, dtype=dt).pin_memory()
        h_pool = [(a.cpu().pin_memory(), b.cpu().pin_memory()) for a, b in pool]
        hs = {"obs0": pin(E, 50, dt=torch.float64), "obsi": pin(E, 5, 19, dt=torch.float64),
              "mask0": pin(E, 5, dt=torch.int8), "maski": pin(E, 5, 5, dt=torch.int8),
              "reward": pin(E, 6, dt=torch.float64), "done": pin(E, dt=torch.uint8)}
        surf = capi.HostSurface(*[hs[n].data_ptr() for n in capi.SURFACE_FIELDS])
        if os.environ.get("IRS_BENCH_SURF"):  # experiment: only the first k surface arrays (per-copy overhead of the chunked D2H)
            keep = int(os.environ["IRS_BENCH_SURF"])
            surf = capi.HostSurface(*[hs[n].data_ptr() if i < keep else None for i, n in enumerate(capi.SURFACE_FIELDS)])
        h2d = sum(a.numel() * a.element_size() + b.numel() * b.element_size() for a, b in h_pool[:1])
        d2h = sum(t.numel() * t.element_size() for t in hs.values())
        def time_host(surface, bytes_out):
            for k in range(3):
                eng.step_host(h_pool[k % 4][0], h_pool[k % 4][1], surface)
            barrier()
            ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            ev0.record()
            tw = time.perf_counter()
            for k in range(Ke):
                eng.step_host(h_pool[k % 4][0], h_pool[k % 4][1], surface)
            ev1.record()
            barrier()
            ms = max(ev0.elapsed_time(ev1), (time.perf_counter() - tw) * 1e3)
            if os.environ.get("IRS_BENCH_DEBUG"):
                per = []
                for k in range(10):
                    tq = time.perf_counter()
                    eng.step_host(h_pool[k % 4][0], h_pool[k % 4][1], surface)
                    per.append((time.perf_counter() - tq) * 1e3)
                print(f"e2e debug: events {ev0.elapsed_time(ev1) / Ke:.3f} ms/step, wall {ms / Ke:.3f}, single calls {sorted(per)}", file=sys.stderr)
            if world > 1:
                t = torch.tensor([ms], device=dev, dtype=torch.float64)
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                ms = float(t.item())
            return {"value": world * E * U * Ke / (ms * 1e-3), "unit": "UE-steps/s", "h2d_bytes_per_step": h2d,
                    "d2h_bytes_per_step": bytes_out, "steps": Ke, "ms_per_step": ms / Ke}

        probe = d2h_copy_rate(torch, dev, d2h, barrier, world, dist)
        e2e = time_host(surf, d2h)
        e2e["api"] = "irs_step_host (pinned host actions in, float64 obs/masks/rewards/done out, stream drained)"
        # the ceiling of this figure: the same bytes copied device -> pinned host with no kernels running, all ranks at once
        e2e["d2h_copy_only"] = {"gbs_per_gpu": probe["gbs_per_gpu"], "gbs_total": probe["gbs_total"],
                                "ms_per_step": probe["ms_per_copy"], "frac_of_ceiling": probe["ms_per_copy"] / e2e["ms_per_step"]}
        # the same call with the optional binary32 surface (converted on the device): half the bytes over PCIe
        h32 = {"obs0_f32": pin(E, 50, dt=torch.float32), "obsi_f32": pin(E, 5, 19, dt=torch.float32),
               "reward_f32": pin(E, 6, dt=torch.float32)}
        surf32 = capi.HostSurface(None, None, hs["mask0"].data_ptr(), hs["maski"].data_ptr(), None, hs["done"].data_ptr(),
                                  *[h32[n].data_ptr() for n in capi.SURFACE_F32_FIELDS])
        d2h32 = sum(t.numel() * t.element_size() for t in list(h32.values()) + [hs["mask0"], hs["maski"], hs["done"]])
        e2e_f32 = time_host(surf32, d2h32)
        e2e_f32["api"] = "irs_step_host with the float32 surface option (obs/rewards converted on the device)"

    if world > 1:
        t = torch.tensor([ms_total], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_total = float(t.item())
    err = eng.check_errors()
    active = float((torch.from_numpy((arr["ue_slice"][:, :25] >= 0).sum(1)).to(dev)[eng.v["scen"].long()]).double().mean().item())
    from intent_radio_sched_multi_slice_b200 import distributed as D
    metrics_named = D.all_reduce_metrics(eng.episode_metrics_device(), dev)  # the path's one collective (NCCL when N > 1)

    # kernels per irs_step (irs_capi.cu launch_step): ONE (k_env_step) up to IRS_FUSED_MAX_ENVS envs (default 2,368), else
    # k_action_8 | k_action_t, k_ue_env, k_env_tail_8 | k_env_tail_t and k_rollover (auto-reset)
    launches_per_step = 1 if E <= int(os.environ.get("IRS_FUSED_MAX_ENVS", "2368")) else 4
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cpu = cpu_oracle_run(args, bank, args.cpu_envs, warmup=5, fill=n_fill)
        cpu = {k: cpu[k] for k in ("value", "unit", "cores", "kind", "sample")}
    if rank == 0:
        value = world * E * U * K / (ms_total * 1e-3)
        line = {
            "metric": "UE-steps/sec (env-steps/sec x UEs)", "value": value, "unit": "UE-steps/s", "n_gpus": world,
            "steps": K, "warmup": Wm, "ms_per_step": ms_total / K, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "i32 buffers + f64", "data": "synthetic",
            "config": {"workload": wl, "envs_per_gpu": E, "ues_per_env": U, "active_ues_per_env": active,
                       "surface": "IBSched dict obs + action masks + 6 rewards, random actions supplied on device",
                       "channel": "SE trace bank in HBM (channel_mode 3)" if args.workload == "trace" else "native mimic_quadriga",
                       "trace_bank_bytes": trace_bytes, "traffic": "native Poisson",
                       "state_bytes_per_gpu": state_bytes, "fill_steps": n_fill, "episode_steps": episode_steps(args),
                       "host_numa_pinned": numa_pinned, "cuda_graph": bool(eng.graph),
                       "l2": "state larger than L2" if flush is None else "L2 flushed (256 MB fill) between timed steps"},
            "e2e": e2e, "e2e_f32_surface": e2e_f32, "roofline": roofline, "cpu_baseline": cpu, "gpu_launches": launches_per_step * K, "clocks": clocks,
            "active_ue_steps_per_s": value * active / U,  # SURVEY 8(d) secondary figure: only UEs associated to a slice
            "env_errors": err, "episode_metrics": metrics_named,
        }
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_b200(a)
