"""Turn the scratch outputs of tools/gpu_round.sh (gpurun_out/) into the tracked summaries under profiles/.

    python tools/summarize_profiles.py r01

* profiles/<round>_launches.csv            kernel, launches, total us, us/launch, share   (ncu gpu__time_duration.sum pass)
* profiles/<round>_<kernel>_ncu_full.txt   selected metrics of the `ncu --set full` capture, per launch
* profiles/decode_attn_traffic.json        dram bytes per launch of the dominant kernel (read by bench.py -> roofline.traffic)
"""
import collections
import csv
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "gpurun_out")
PROF = os.path.join(ROOT, "profiles")

WANT = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__bytes_read.sum.per_second",
    "dram__throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "launch__grid_size", "launch__block_size", "launch__cluster_size", "launch__registers_per_thread",
    "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_registers",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "lts__t_sector_hit_rate.pct", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
    "sm__inst_executed_pipe_tensor.sum", "sm__pipe_tensor_subpipe_imma_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
]


def launches(tag, src="launches.csv", skip_setup=True):
    path = os.path.join(OUT, src)
    if not os.path.exists(path):
        return
    lines = [l for l in open(path) if not l.startswith("==")]
    agg = collections.OrderedDict()
    for row in csv.DictReader(lines):
        if row.get("Metric Name") != "gpu__time_duration.sum":
            continue
        name = row["Kernel Name"]
        if skip_setup and ("distribution_elementwise" in name or "FillFunctor" in name or "copy_kernel" in name):
            continue  # weight / KV initialisation of the harness, not part of a step
        v = float(row["Metric Value"].replace(",", ""))
        v = {"ns": v / 1e3, "us": v, "ms": v * 1e3}.get(row["Metric Unit"], v)
        a = agg.setdefault(name[:110], [0, 0.0])
        a[0] += 1
        a[1] += v
    tot = sum(v[1] for v in agg.values())
    with open(os.path.join(PROF, f"{tag}_launches.csv"), "w") as fh:
        fh.write("# ncu --metrics gpu__time_duration.sum --clock-control none python tools/profile_step.py --layers 2 --steps 3\n")
        fh.write("# (cold-cache, serialised per-launch times; harness initialisation kernels dropped)\n")
        fh.write("kernel,launches,total_us,us_per_launch,share_pct\n")
        for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            fh.write(f"\"{k}\",{v[0]},{v[1]:.1f},{v[1] / v[0]:.2f},{100 * v[1] / tot:.1f}\n")


def full(tag, rep, kernel_label, traffic_json=None):
    path = os.path.join(OUT, rep)
    if not os.path.exists(path):
        return
    raw = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    with open(os.path.join(PROF, f"{tag}_{kernel_label}_ncu_full.txt"), "w") as fh:
        fh.write(f"# ncu --set full --clock-control none --import-source on  ({rep}); one block per captured launch\n")
        for r in rows[2:]:
            fh.write(f"--- {r[hdr.index('Kernel Name')]}\n")
            for w in WANT:
                if w in hdr:
                    i = hdr.index(w)
                    fh.write(f"{w:86s} {r[i]:>16s} {units[i]}\n")
    if traffic_json:
        def val(r, name):
            i = hdr.index(name)
            m = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[units[i]]
            return float(r[i].replace(",", "")) * m
        tr = [val(r, "dram__bytes_read.sum") + val(r, "dram__bytes_write.sum") for r in rows[2:]]
        json.dump({"kernel": kernel_label, "dram_bytes_per_launch": sum(tr) / len(tr), "launches_captured": len(tr),
                   "source": f"profiles/{tag}_{kernel_label}_ncu_full.txt"},
                  open(os.path.join(PROF, traffic_json), "w"))


if __name__ == "__main__":
    tag = sys.argv[1] if len(sys.argv) > 1 else "r01"
    os.makedirs(PROF, exist_ok=True)
    launches(tag)
    full(tag, "decode_attn_full.ncu-rep", "decode_attn", "decode_attn_traffic.json")
    for extra in sys.argv[2:]:
        rep, label = extra.split(":")
        full(tag, rep, label)
