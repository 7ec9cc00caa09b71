"""Turn the .ncu-rep files of tools/profile_round.sh (under gpurun_out/) into the tracked summaries
under profiles/: raw-page CSV per kernel, and profiles/ncu_summary.json with the per-unit figures
bench.py quotes (DRAM bytes and executed MUFU.EX2 per image).

    python tools/ncu_summarize.py r02
"""
import csv, json, os, subprocess, sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
R = sys.argv[1] if len(sys.argv) > 1 else "r02"
G, P = os.path.join(ROOT, "gpurun_out"), os.path.join(ROOT, "profiles")


def raw(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    return out, dict(zip(rows[0], rows[2])), dict(zip(rows[0], rows[1]))


def sass_counts(rep, mnemonic):
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))[2:]
    return sum(int(r[6]) for r in rows if mnemonic in r[1])   # thread instructions executed


summary = {}
KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "launch__registers_per_thread",
        "launch__grid_size", "launch__block_size", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fmaheavy.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum"]
for name, units, unit_name in [(f"{R}_jet2_kernel", None, "jet2"), (f"{R}_survey_tile_012", 16, "survey_tile_012"),
                               (f"{R}_survey_tile_212", 16, "survey_tile")]:
    rep = os.path.join(G, name + ".ncu-rep")
    if not os.path.exists(rep):
        print("missing", rep)
        continue
    text, d, u = raw(rep)
    open(os.path.join(P, name + "_ncu_full_raw.csv"), "w").write(text)
    ex2 = sass_counts(rep, "MUFU.EX2")
    if units is None:   # jet2: images per launch = grid / 2 clusters ... read from the bench's chunk (16384)
        units = 16384
    def f(k):
        v = float(d[k].replace(",", ""))
        return v
    mult = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    dram = f("dram__bytes_read.sum") * mult[u["dram__bytes_read.sum"]] + f("dram__bytes_write.sum") * mult[u["dram__bytes_write.sum"]]
    tmult = {"ns": 1e-9, "us": 1e-6, "ms": 1e-3, "s": 1.0}
    entry = {"source": f"ncu --set full, profiles/{name}_ncu_full_raw.csv ({units} images in the launch)",
             "launch_seconds": f("gpu__time_duration.sum") * tmult[u["gpu__time_duration.sum"]],
             "dram_bytes_per_image": dram / units, "executed_ex2_per_image": ex2 / units,
             "metrics": {k: d[k] + " " + u[k] for k in KEYS if k in d},
             "stalls_per_issue": {k.split("stalled_")[1].split("_per_issue")[0]: float(d[k]) for k in d
                                  if "issue_stalled" in k and k.endswith("per_issue_active.ratio") and float(d[k]) >= 0.05}}
    summary[unit_name] = entry
if "survey_tile" in summary:
    summary["survey_tile"]["executed_ex2_per_image_2000"] = summary["survey_tile"]["executed_ex2_per_image"]
json.dump(summary, open(os.path.join(P, "ncu_summary.json"), "w"), indent=1)
print(json.dumps({k: {kk: vv for kk, vv in v.items() if kk != "metrics"} for k, v in summary.items()}, indent=1))
for f in os.listdir(G):
    if f.startswith(R + "_launches_") and f.endswith(".csv"):
        open(os.path.join(P, f), "w").write(open(os.path.join(G, f)).read())
