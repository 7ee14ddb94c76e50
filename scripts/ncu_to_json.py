"""Record what an `ncu --set full` capture of k_ue_env measured into profiles/ue_env_ncu.json (read by bench.py).
usage: python scripts/ncu_to_json.py <capture.ncu-rep> <workload> <envs> <depth> [source note]
Instruction and DRAM-byte counts are properties of the launch; bench.py divides them by the live kernel time."""
import csv, io, json, os, subprocess, sys

rep, workload, envs, depth = sys.argv[1], sys.argv[2], int(sys.argv[3]), int(sys.argv[4])
note = sys.argv[5] if len(sys.argv) > 5 else os.path.basename(rep)
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, r = rows[0], rows[1], rows[2]
col = {h: i for i, h in enumerate(hdr)}


def val(name, scale_by_unit=True):
    v = float(r[col[name]].replace(",", ""))
    u = units[col[name]]
    mul = {"Mbyte": 1e6, "Gbyte": 1e9, "Kbyte": 1e3, "byte": 1, "us": 1e-6, "ms": 1e-3, "ns": 1e-9}.get(u, 1)
    return v * mul if scale_by_unit else v


entry = {"workload": workload, "envs": envs, "depth": depth, "kernel": r[col["Kernel Name"]],
         "inst_per_launch": val("smsp__inst_executed.sum"),
         "dram_bytes_per_launch": val("dram__bytes_read.sum") + val("dram__bytes_write.sum"),
         "issue_active_pct": val("smsp__issue_active.avg.pct_of_peak_sustained_active", False),
         "ncu_kernel_seconds": val("gpu__time_duration.sum"), "source": note}
path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "profiles", "ue_env_ncu.json")
data = json.load(open(path)) if os.path.exists(path) else []
data = [d for d in data if (d["workload"], d["envs"], d["depth"]) != (workload, envs, depth)] + [entry]
json.dump(data, open(path, "w"), indent=1)
print(entry)
