#!/usr/bin/env python
"""Turn what a GPU visit left in gpurun_out/ into the tracked summaries under profiles/:
  <tag>_launches.md      per-kernel totals of the ncu launch list (gpu__time_duration.sum) of the bench command
  <tag>_<kernel>_full.md key counters of the `ncu --set full` capture of one launch of a kernel
  <tag>_bench.json       the bench lines of the same visit
usage: tools/summarize_profiles.py r01 [gpurun_out [output dir]]"""
import collections
import csv
import json
import os
import re
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

KEYS = [
    "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
    "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "launch__waves_per_multiprocessor",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__inst_executed.avg.per_cycle_active",
    "smsp__warps_eligible.avg.per_cycle_active", "smsp__thread_inst_executed_per_inst_executed.ratio",
    "smsp__thread_inst_executed_per_inst_executed.pct",
    "gpu__compute_memory_throughput.avg.pct_of_peak_sustained_elapsed",
    "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__t_bytes.sum", "lts__t_sector_hit_rate.pct", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "l1tex__t_bytes.sum", "l1tex__t_sector_hit_rate.pct", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
    "smsp__inst_executed_op_local_ld.sum", "smsp__inst_executed_op_local_st.sum",
    "smsp__inst_executed_op_global_ld.sum", "smsp__inst_executed_op_global_st.sum",
    "smsp__inst_executed.sum", "sm__inst_executed_pipe_fp64.sum", "sm__inst_executed_pipe_fma.sum",
    "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
    "smsp__sass_average_branch_targets_threads_uniform.pct",
]


TRAFFIC = {}
LIST_AVG_MS = {}     # kernel -> average launch duration in the launch list (same serialised, cold-cache conditions as a full capture)


def short(name):
    name = re.sub(r"<unnamed>::", "", name)
    name = re.sub(r"^void ", "", name)
    return re.sub(r"\(.*$", "", name)


def launches(tag, src, out):
    path = os.path.join(src, "launches.csv")
    if not os.path.exists(path):
        return
    rows = [r for r in csv.reader(l for l in open(path) if l.startswith('"'))]
    hdr = rows[0]
    k, v = hdr.index("Kernel Name"), hdr.index("Metric Value")
    tot = collections.defaultdict(float); cnt = collections.Counter()
    for r in rows[1:]:
        n = short(r[k]); tot[n] += float(r[v].replace(",", "")) * 1e-6; cnt[n] += 1
    total = sum(tot.values())
    for n in tot:
        LIST_AVG_MS[n] = tot[n] / cnt[n]
    with open(os.path.join(out, f"{tag}_launches.md"), "w") as f:
        f.write(f"# {tag}: ncu launch list of `python bench.py --steps 1 --warmup 1 --no-cpu-baseline`\n\n")
        f.write("`ncu --metrics gpu__time_duration.sum --clock-control none` (per-launch times are serialised and "
                "cold-cache: compare SHARES with bench.py's live `roofline.share_of_step`, not absolutes).\n"
                "The list covers the warm-up step, the timed step, the roofline counting pass and the e2e passes "
                "(scene upload + BVH build kernels).\n\n")
        f.write(f"{sum(cnt.values())} launches, {total:.1f} ms of kernel time\n\n| kernel | launches | total ms | share | avg us |\n|---|---|---|---|---|\n")
        for n in sorted(tot, key=tot.get, reverse=True):
            f.write(f"| `{n}` | {cnt[n]} | {tot[n]:.2f} | {100 * tot[n] / total:.1f}% | {1e3 * tot[n] / cnt[n]:.1f} |\n")
        wf = {n: t for n, t in tot.items() if n.startswith("k_wf_")}
        wt = sum(wf.values())
        f.write("\nWavefront stages only (what one render step consists of):\n\n| stage kernel | share of wavefront kernel time |\n|---|---|\n")
        for n in sorted(wf, key=wf.get, reverse=True):
            f.write(f"| `{n}` | {100 * wf[n] / wt:.1f}% |\n")


def full(tag, src, out):
    for rep in sorted(os.listdir(src)):
        if not rep.endswith(".ncu-rep"):
            continue
        raw = subprocess.run(["ncu", "-i", os.path.join(src, rep), "--page", "raw", "--csv"], stdout=subprocess.PIPE,
                             stderr=subprocess.DEVNULL, text=True).stdout
        rows = list(csv.reader(raw.splitlines()))
        if len(rows) < 3:
            continue
        hdr, units, vals = rows[0], rows[1], rows[2]
        d = {h: (v, u) for h, u, v in zip(hdr, units, vals)}
        name = rep[:-8]
        with open(os.path.join(out, f"{tag}_{name}.md"), "w") as f:
            f.write(f"# {tag}: `ncu --set full --clock-control none` of one launch of `{short(d['Kernel Name'][0])}`\n\n")
            what = ("`python tools/sweep_pool.py hetvol 8 16777216` (hetvol 768x576, 8 spp, volumetric wavefront engine); launch `-s 2 -c 1`"
                    if "track" in name else
                    "`python bench.py --steps 1 --warmup 1` (sponza 768x575, 64 spp, wavefront engine); a mid-render launch (`-s 6 -c 1`)")
            f.write(f"Captured inside {what}.\n\n| metric | value | unit |\n|---|---|---|\n")
            for k in KEYS:
                if k in d:
                    f.write(f"| {k} | {d[k][0]} | {d[k][1]} |\n")
            try:
                tr = (float(d["dram__bytes_read.sum"][0].replace(",", "")) * {"Mbyte": 1e6, "Gbyte": 1e9, "Kbyte": 1e3, "byte": 1}[d["dram__bytes_read.sum"][1]]
                      + float(d["dram__bytes_write.sum"][0].replace(",", "")) * {"Mbyte": 1e6, "Gbyte": 1e9, "Kbyte": 1e3, "byte": 1}[d["dram__bytes_write.sum"][1]])
                TRAFFIC[short(d["Kernel Name"][0])] = {"dram_bytes_per_launch": tr, "capture": f"{tag}_{name}.md",
                                                       "launch_ms": float(d["gpu__time_duration.sum"][0].replace(",", "")) * {"ms": 1, "us": 1e-3, "s": 1e3, "ns": 1e-6}[d["gpu__time_duration.sum"][1]]}
                def num(k):
                    return float(d[k][0].replace(",", "")) if k in d else None
                ent = TRAFFIC[short(d["Kernel Name"][0])]
                # the issue-slot view of the same launch (bench.py `roofline.issue`): warp instructions, lanes per instruction
                ent.update({"warp_inst_per_launch": num("smsp__inst_executed.sum"),
                            "lanes_per_inst": num("smsp__thread_inst_executed_per_inst_executed.ratio"),
                            "issue_active_pct": num("smsp__issue_active.avg.pct_of_peak_sustained_active"),
                            "warps_active_pct": num("sm__warps_active.avg.pct_of_peak_sustained_active"),
                            "l2_hit_pct": num("lts__t_sector_hit_rate.pct"), "l1_hit_pct": num("l1tex__t_sector_hit_rate.pct"),
                            "registers": num("launch__registers_per_thread")})
            except Exception as e:
                print("traffic:", e)
            f.write("\nWarp stall reasons (% of warp-active cycles per issue slot):\n\n| reason | pct |\n|---|---|\n")
            st = [(h, float(v.replace(",", ""))) for h, (v, u) in d.items()
                  if h.startswith("smsp__average_warps_issue_stalled_") and h.endswith("_per_issue_active.ratio")]
            for h, v in sorted(st, key=lambda x: -x[1])[:10]:
                f.write(f"| {h[len('smsp__average_warps_issue_stalled_'):-len('_per_issue_active.ratio')]} | {v:.3f} |\n")


def main():
    tag = sys.argv[1]
    src = sys.argv[2] if len(sys.argv) > 2 else os.path.join(ROOT, "gpurun_out")
    out = sys.argv[3] if len(sys.argv) > 3 else os.path.join(ROOT, "profiles")
    os.makedirs(out, exist_ok=True)
    launches(tag, src, out)
    full(tag, src, out)
    for k, t in TRAFFIC.items():
        # the captured launch is one of the early, larger ones; the launch list gives the duration of the average
        # launch under the same conditions, so bench.py can scale the capture's DRAM bytes to an average launch
        if k in LIST_AVG_MS:
            t["launch_list_avg_ms"] = LIST_AVG_MS[k]
            t["dram_bytes_per_avg_launch"] = t["dram_bytes_per_launch"] * LIST_AVG_MS[k] / t["launch_ms"]
    if TRAFFIC:
        json.dump(TRAFFIC, open(os.path.join(out, "traffic.json"), "w"), indent=1)
    lines = {}
    for n in ("bench.json", "bench_ref.json", "stages_sponza.json"):
        p = os.path.join(src, n)
        if os.path.exists(p) and os.path.getsize(p):
            try:
                lines[n] = json.loads(open(p).read())
            except Exception:
                lines[n] = open(p).read()
    json.dump(lines, open(os.path.join(out, f"{tag}_bench.json"), "w"), indent=1)
    for n in ("pytest_gpu.log", "smoke.log"):
        p = os.path.join(src, n)
        if os.path.exists(p):
            shutil.copy(p, os.path.join(out, f"{tag}_{n}"))


if __name__ == "__main__":
    main()
