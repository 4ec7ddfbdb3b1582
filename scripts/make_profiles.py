"""Turn the raw files a GPU run left in gpurun_out/ into the tracked summaries under profiles/.

Inputs (written by the commands quoted in profiles/README.md):
  gpurun_out/<tag>_bench.json       one line of `python bench.py`
  gpurun_out/<tag>_bench_ref.json   one line of `python bench.py --impl reference`
  gpurun_out/<tag>_bench_n<N>.json  one line of the N-GPU run under torchrun (optional)
  gpurun_out/<tag>_launches.csv     ncu --metrics gpu__time_duration.sum launch list of the bench command
  gpurun_out/<tag>_series_flow_full.ncu-rep   one ncu --set full capture of series_flow on the bench command (optional)
Usage: python scripts/make_profiles.py r02
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
TAG = sys.argv[1] if len(sys.argv) > 1 else "r02"
KERNEL = "series_flow"


def last_json_line(path):
    with open(path) as f:
        lines = [ln for ln in f.read().splitlines() if ln.startswith("{")]
    return json.loads(lines[-1])


def launches():
    rows = [r for r in csv.reader(open(os.path.join(OUT, "%s_launches.csv" % TAG))) if len(r) > 5]
    hdr = rows[0]
    ix = {h: i for i, h in enumerate(hdr)}
    seq = []
    for r in rows[1:]:
        try:
            us = float(r[ix["Metric Value"]]) * (1e-3 if r[ix["Metric Unit"]] == "ns" else 1.0)
        except ValueError:
            continue
        name = r[ix["Kernel Name"]].split("(")[0].split("::")[-1]
        seq.append((name, us))
    # one device-resident step = classify_series ... series_long with the big series_flow launch inside
    steps, cur = [], []
    for name, us in seq:
        if name == "classify_series" and cur:
            steps.append(cur)
            cur = []
        cur.append((name, us))
    if cur:
        steps.append(cur)
    dev_steps = [s for s in steps if max(us for _, us in s) > 10000.0]   # whole-workload steps, not e2e chunks
    agg = collections.OrderedDict()
    for s in dev_steps:
        seen = collections.Counter()
        for name, us in s:
            seen[name] += 1
            key = name if seen[name] == 1 else "%s (launch %d: second attempt)" % (name, seen[name])
            agg.setdefault(key, []).append(us)
    total = sum(sum(v) / len(v) for v in agg.values())
    table = [{"kernel": k, "launches_per_step": 1, "avg_us": round(sum(v) / len(v), 1),
              "share_of_step": round(sum(v) / len(v) / total, 4)} for k, v in agg.items()]
    return {"command": "ncu --metrics gpu__time_duration.sum --clock-control none -k regex:<library kernels> "
                       "-c 160 --csv python bench.py --quick --no-cpu-baseline",
            "note": "per-launch times under ncu are cold-cache and serialised: shares, not absolutes",
            "device_resident_steps_seen": len(dev_steps), "sum_us": round(total, 1), "kernels": table}


def main():
    os.makedirs(PROF, exist_ok=True)
    bench = last_json_line(os.path.join(OUT, "%s_bench.json" % TAG))
    json.dump(bench, open(os.path.join(PROF, "%s_bench_line.json" % TAG), "w"), indent=1)
    ref_path = os.path.join(OUT, "%s_bench_ref.json" % TAG)
    if os.path.exists(ref_path):
        json.dump(last_json_line(ref_path), open(os.path.join(PROF, "%s_bench_reference_line.json" % TAG), "w"),
                  indent=1)
    ll = launches()
    json.dump(ll, open(os.path.join(PROF, "%s_launch_list.json" % TAG), "w"), indent=1)
    for n in (2, 4, 8):
        pn = os.path.join(OUT, "%s_bench_n%d.json" % (TAG, n))
        if os.path.exists(pn):
            json.dump(last_json_line(pn), open(os.path.join(PROF, "%s_bench_line_n%d.json" % (TAG, n)), "w"), indent=1)
    rep = os.path.join(OUT, "%s_%s_full.ncu-rep" % (TAG, KERNEL))
    if os.path.exists(rep):
        raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
        rows = list(csv.reader(raw.splitlines()))
        hdr, units, vals = rows[0], rows[1], rows[-1]
        d = {h: (v, u) for h, v, u in zip(hdr, vals, units)}
        keep = ["Kernel Name", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
                "launch__shared_mem_per_block_dynamic", "gpu__time_duration.sum", "dram__bytes_read.sum",
                "dram__bytes_write.sum", "smsp__inst_executed.sum",
                "smsp__issue_active.avg.pct_of_peak_sustained_active",
                "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__warps_eligible.avg.per_cycle_active",
                "smsp__average_warp_latency_per_inst_issued.ratio",
                "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
                "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
                "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
                "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
                "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
                "sm__icc_request_hit_rate.pct", "lts__t_sector_hit_rate.pct",
                "dram__throughput.avg.pct_of_peak_sustained_elapsed"]
        out = {k: {"value": d[k][0], "unit": d[k][1]} for k in keep if k in d}
        out["pc_sampling_stall_reasons"] = {h.replace("smsp__pcsamp_warps_issue_stalled_", ""): d[h][0]
                                            for h in hdr if h.startswith("smsp__pcsamp_warps_issue_stalled")}
        scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
        traffic = sum(float(d[k][0]) * scale[d[k][1]] for k in ("dram__bytes_read.sum", "dram__bytes_write.sum"))
        out["dram_bytes_per_launch"] = traffic
        algo = float(bench["roofline"]["algorithmic_bytes_per_launch"])
        out["algorithmic_bytes_per_launch"] = algo
        out["dram_bytes_per_algorithmic_byte"] = traffic / algo
        out["command"] = ("ncu --set full --clock-control none --import-source on -k regex:%s -s 8 -c 1 "
                          "python bench.py --quick --no-cpu-baseline" % KERNEL)
        json.dump(out, open(os.path.join(PROF, "%s_%s_ncu_full.json" % (TAG, KERNEL)), "w"), indent=1)
        json.dump({"kernel": KERNEL, "dram_bytes_per_algorithmic_byte": traffic / algo,
                   "dram_bytes_per_launch_of_capture": traffic, "algorithmic_bytes_of_capture": algo,
                   "source": "profiles/%s_%s_ncu_full.json (ncu --set full on the bench command, not this run)"
                             % (TAG, KERNEL)},
                  open(os.path.join(PROF, "roofline_traffic.json"), "w"), indent=1)
    print(json.dumps(ll, indent=1))
    print("bench:", bench["ms_per_step"], bench["roofline"]["frac"], bench["e2e"]["value"])


if __name__ == "__main__":
    main()
