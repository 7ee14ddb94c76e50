"""Print the headline fields of bench JSON lines: python scripts/show_bench.py gpurun_out/bench*.json"""
import json, sys
for p in sys.argv[1:]:
    try:
        lines = [l for l in open(p).read().strip().splitlines() if l.startswith("{")]
        d = json.loads(lines[-1])
    except Exception as exc:
        print(p, "unreadable:", exc)
        continue
    r = d.get("roofline") or {}
    e = d.get("e2e") or {}
    print(f"{p}: value {d.get('value', 0):.3e} ms/step {d.get('ms_per_step', 0):.3f} e2e {e.get('value', 0):.3e} "
          f"kernels {({k: round(v, 3) for k, v in (r.get('kernel_ms_all') or {}).items()})} frac {r.get('frac')} "
          f"paths {({k: round(v, 3) for k, v in (r.get('row_paths') or {}).items()})} cpu {(d.get('cpu_baseline') or {}).get('value')}")
