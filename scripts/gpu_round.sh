#!/bin/bash
# One GPU-box pass: parity tests, smoke, both bench arms, the other workloads, ncu launch list + full captures.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err
python bench.py --impl reference --steps 20 --warmup 3 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
python bench.py --workload mix > gpurun_out/bench_mix.json 2> gpurun_out/bench_mix.err
python bench.py --workload trace > gpurun_out/bench_trace.json 2> gpurun_out/bench_trace.err
python bench.py --workload cfg2 --no-cpu-baseline > gpurun_out/bench_cfg2.json 2> gpurun_out/bench_cfg2.err
python bench.py --depth 401 --steps 100 --no-cpu-baseline > gpurun_out/bench_d401.json 2> gpurun_out/bench_d401.err
python bench.py --depth 4096 --envs 16384 --steps 50 --no-cpu-baseline > gpurun_out/bench_d4096.json 2> gpurun_out/bench_d4096.err
# ncu passes: the buffers are first filled to their steady state (the fill lengths are what bench.py's plateau detection
# settles on, see config.fill_steps of the bench lines); unprofiled launches are skipped with -s (counts matching kernels)
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:^k_ -s 8843 -c 15 --csv --log-file gpurun_out/launches.csv \
    python bench.py --fill 2944 --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/ncu_launches.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_ue_env -s 2950 -c 1 -f -o gpurun_out/k_ue_env_sweep \
    python bench.py --fill 2944 --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/ncu_full_sweep.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_ue_env -s 518 -c 1 -f -o gpurun_out/k_ue_env_mix \
    python bench.py --workload mix --fill 512 --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/ncu_full_mix.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_ue_env -s 518 -c 1 -f -o gpurun_out/k_ue_env_trace \
    python bench.py --workload trace --fill 512 --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/ncu_full_trace.log 2>&1
# a history file written by the recorder, for a read-back with the reference's own Metrics.read in the build container
python - > gpurun_out/hist.log 2>&1 <<'PY'
import sys, numpy as np
sys.path.insert(0, "tests")
from intent_radio_sched_multi_slice_b200 import agents, plugins
from intent_radio_sched_multi_slice_b200.env import MARLCommEnv
env = MARLCommEnv(plugins.MimicQuadriga, plugins.MultSliceTraffic, plugins.SimpleMobility, plugins.MultSliceAssociation,
                  "mult_slice", "ib_sched", 10, root_path="/nonexistent", save_hist=True, hist_root_path="gpurun_out/",
                  simu_name="demo", debug=True)
env.max_number_steps = 50
ag = agents.IBSched(env, 25, 5, 1, np.array([135]))
env.set_agent_functions(ag.obs_space_format, ag.action_format, ag.calculate_reward, ag.get_obs_space(), ag.get_action_space())
env.reset(seed=10, options={"initial_episode": 0})
rng = np.random.default_rng(0)
for k in range(50):
    a = {"player_0": rng.uniform(-1, 1, 5)}
    a.update({f"player_{i}": int(rng.integers(0, 3)) for i in range(1, 6)})
    env.step(a)
print("written", env._recorder.written)
PY
ls -la gpurun_out
