"""Turns one GPU session's raw captures under gpurun_out/ into the tracked summaries under profiles/ (prefix given):
launch list shares, ncu --set full summaries and hot source lines of the two kernels, the SASS evidence for the
bulk copies, the bench lines, and profiles/ncu_traffic.json (DRAM bytes per launch, read by bench.py).
usage: python tools/profile_summaries.py r02a"""
import csv, json, os, subprocess, sys
from collections import OrderedDict

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
G = os.path.join(ROOT, "gpurun_out")
P = os.path.join(ROOT, "profiles")
prefix = sys.argv[1]
commit = subprocess.run(["git", "rev-parse", "--short", "HEAD"], capture_output=True, text=True, cwd=ROOT).stdout.strip()


def sh(cmd):
    return subprocess.run(cmd, capture_output=True, text=True, cwd=ROOT).stdout


# ---- launch list
rows = [r for r in csv.reader(open(os.path.join(G, "r02_launches.csv"))) if len(r) > 10 and r[0].isdigit()]
agg = OrderedDict()
for r in rows:
    name = r[4].split("(")[0].replace("void ", "")
    key = (name[:90], r[8])
    a = agg.setdefault(key, [0, 0.0])
    a[0] += 1
    a[1] += float(r[14]) / 1000.0
total = sum(v[1] for v in agg.values())
with open(os.path.join(P, f"{prefix}_launches_summary.txt"), "w") as f:
    f.write(f"kernel, grid, launches, mean_us, total_us, share  (commit {commit}; ncu --profile-from-start off --metrics gpu__time_duration.sum "
            "--clock-control none, python tools/profile_steady.py: the bench's steady state, one step as two whole-batch kernels and one step "
            "as the sliced graph; cold-cache, serialised: shares matter, not absolutes)\n")
    for (name, grid), (n, us) in agg.items():
        f.write(f"{name}, {grid}, {n}, {us / n:.1f}, {us:.1f}, {100 * us / total:.1f}%\n")
os.replace(os.path.join(G, "r02_launches.csv"), os.path.join(P, f"{prefix}_launches.csv")) if os.path.exists(os.path.join(G, "r02_launches.csv")) else None

# ---- ncu summaries + hot lines
traffic = {"source": f"profiles/{prefix}_ncu_full_summary.txt (ncu --set full --clock-control none, one launch over 4096 envs x 24 agents in the bench's steady state, commit {commit})"}
with open(os.path.join(P, f"{prefix}_ncu_full_summary.txt"), "w") as f:
    for tag, warps in (("step", 4096), ("obs", 98304)):
        rep = os.path.join(G, f"r02_{tag}.ncu-rep")
        out = sh([sys.executable, "tools/ncu_summary.py", rep])
        f.write(f"==== {tag} kernel (commit {commit}, steady state, L2 flushed before the launch)\n{out}\n")
        rd = wr = 0.0
        for line in out.splitlines():
            parts = line.split()
            if parts and parts[0] == "dram__bytes_read.sum": rd = float(parts[1]) * {"Mbyte": 1e6, "Kbyte": 1e3, "Gbyte": 1e9, "byte": 1}[parts[2]]
            if parts and parts[0] == "dram__bytes_write.sum": wr = float(parts[1]) * {"Mbyte": 1e6, "Kbyte": 1e3, "Gbyte": 1e9, "byte": 1}[parts[2]]
        traffic[f"mg_{tag}_kernel_bytes"] = int(rd + wr)
        hot = sh([sys.executable, "tools/hot_lines.py", rep, str(warps)])
        open(os.path.join(P, f"{prefix}_{tag}_kernel_hot_lines.txt"), "w").write(f"commit {commit}; per warp = per {'env' if tag == 'step' else 'row'}\n" + "\n".join(hot.splitlines()[:64]) + "\n")
traffic["whole_step_bytes"] = traffic["mg_step_kernel_bytes"] + traffic["mg_obs_kernel_bytes"]
json.dump(traffic, open(os.path.join(P, "ncu_traffic.json"), "w"), indent=1)

# ---- SASS evidence
sass = sh(["cuobjdump", "-sass", "metta_b200/csrc/libmettagrid_b200.so"])
cur, counts = None, OrderedDict()
import re
for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = sh(["c++filt", m.group(1)]).strip().split("(")[0]
        continue
    m = re.search(r"\b(UBLKCP\.\S+|SYNCS\.\S+|UTMALDG\S*|UTMASTG\S*|REDUX\S*|RED\.E\.\S+)", line)
    if m and cur:
        counts.setdefault(cur, {}).setdefault(m.group(1).rstrip(";"), 0)
        counts[cur][m.group(1).rstrip(";")] += 1
with open(os.path.join(P, f"{prefix}_sass_sm100.txt"), "w") as f:
    f.write(f"cuobjdump -sass libmettagrid_b200.so (commit {commit}, sm_100a): bulk-copy (TMA), mbarrier, warp-reduction and reduction-atomic instructions per kernel\n")
    for k, v in counts.items():
        f.write(f"{k}: " + ", ".join(f"{n} x {i}" for i, n in v.items()) + "\n")

# ---- bench lines
for name in ("bench_r02.json", "bench_r02_reference.json"):
    src = os.path.join(G, name)
    if os.path.exists(src):
        line = open(src).read().strip().splitlines()[-1]
        json.loads(line)
        open(os.path.join(P, f"{prefix}_{name.replace('_r02', '')}"), "w").write(line + "\n")
print(open(os.path.join(P, f"{prefix}_launches_summary.txt")).read())
print(json.dumps(traffic, indent=1))
