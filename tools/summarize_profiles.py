"""Turn gpurun_out/ ncu captures into the small, committed summaries under profiles/."""
import csv, json, os, subprocess, sys, collections
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tag = sys.argv[1] if len(sys.argv) > 1 else "r1"
launches = os.path.join(ROOT, "gpurun_out", f"launches_{tag}.csv")
rep = os.path.join(ROOT, "gpurun_out", f"prof_{tag}_final.ncu-rep")
out_dir = os.path.join(ROOT, "profiles")

# ---- launch list: per kernel count / mean / share
rows = [r for r in csv.reader(open(launches)) if len(r) > 5]
h = rows[[i for i, r in enumerate(rows) if r[0] == "ID"][0]]
data = rows[rows.index(h) + 1:]
ki, vi = h.index("Kernel Name"), h.index("Metric Value")
agg = collections.OrderedDict()
for r in data:
    agg.setdefault(r[ki].split("(")[0], []).append(float(r[vi].replace(",", "")) / 1000.0)
ours = {k: v for k, v in agg.items() if k in ("k_prep",) or "k_logmel" in k}
step_us = sum(sum(v) / len(v) for v in ours.values())
with open(os.path.join(out_dir, f"{tag}_launch_list_summary.md"), "w") as f:
    f.write(f"# ncu launch list ({tag}): `ncu --metrics gpu__time_duration.sum --clock-control none -c 400 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-sweep`\n\n")
    f.write("Per-launch times under ncu are cold-cache and serialised: compare SHARES, not absolutes.\n\n")
    f.write("| kernel | launches | mean us | min us | share of our kernels per step |\n|---|---|---|---|---|\n")
    for k, v in agg.items():
        share = f"{(sum(v)/len(v))/step_us*100:.1f} %" if k in ours else "(torch: input generation / e2e copies)"
        f.write(f"| `{k[:60]}` | {len(v)} | {sum(v)/len(v):.1f} | {min(v):.1f} | {share} |\n")

# ---- full capture: key metrics of k_logmel
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rr = list(csv.reader(raw.splitlines()))
hh = rr[0]
want = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "sm__inst_executed_pipe_tensor_subpipe_hmma.avg.pct_of_peak_sustained_active", "lts__t_sectors_srcunit_tex_op_read.sum",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__cycles_elapsed.max", "smsp__inst_executed.sum", "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic",
        "launch__grid_size", "launch__block_size", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed"]
met = {}
for i, c in enumerate(hh):
    if c in want:
        vals = [r[i] for r in rr[2:] if len(r) > i]
        met[c] = {"unit": rr[1][i], "per_launch": vals}
def f(x): return float(x.replace(",", ""))
rd = [f(v) for v in met["dram__bytes_read.sum"]["per_launch"]]
wr = [f(v) for v in met["dram__bytes_write.sum"]["per_launch"]]
scale = {"Mbyte": 1e6, "Gbyte": 1e9, "Kbyte": 1e3, "byte": 1}
rb = sum(rd) / len(rd) * scale[met["dram__bytes_read.sum"]["unit"]]
wb = sum(wr) / len(wr) * scale[met["dram__bytes_write.sum"]["unit"]]
kname = next((r[hh.index("Kernel Name")] for r in rr[2:] if len(r) > 5), "k_logmel2") if "Kernel Name" in hh else "k_logmel2"
commit = subprocess.run(["git", "rev-parse", "--short", "HEAD"], capture_output=True, text=True, cwd=ROOT).stdout.strip()
summary = {"kernel": kname, "workload": "64x10s (55168 frames) per launch", "source": os.path.basename(rep), "commit": commit,
           "dram_bytes_per_launch": rb + wb, "dram_read_bytes": rb, "dram_write_bytes": wb,
           "algorithmic_bytes_per_launch": 55168 * 1344, "metrics": met}
json.dump(summary, open(os.path.join(out_dir, "k_logmel_ncu_metrics.json"), "w"), indent=1)
print(json.dumps({k: v["per_launch"] for k, v in met.items()}, indent=0)[:1500])
print("dram bytes/launch", rb + wb)
