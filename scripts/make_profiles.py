"""Turn the raw files a GPU run left in gpurun_out/ into the tracked summaries under profiles/.

Inputs (written by the commands quoted in profiles/README.md):
  gpurun_out/r01_bench.json       one line of `python bench.py`
  gpurun_out/r01_bench_ref.json   one line of `python bench.py --impl reference`
  gpurun_out/r01_launches.csv     ncu --metrics gpu__time_duration.sum launch list of the bench command
  gpurun_out/r01_series_pipe_full.ncu-rep   one ncu --set full capture of series_pipe (optional)
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
TAG = sys.argv[1] if len(sys.argv) > 1 else "r01"


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
    # one device-resident step = classify_series ... series_long with the big series_pipe launch inside
    steps, cur = [], []
    for name, us in seq:
        if name == "classify_series" and cur:
            steps.append(cur)
            cur = []
        cur.append((name, us))
    if cur:
        steps.append(cur)
    dev_steps = [s for s in steps if max(us for _, us in s) > 800.0]
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
                       "--csv python bench.py --steps 3 --warmup 3 --no-cpu-baseline",
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
    rep = os.path.join(OUT, "%s_series_pipe_full.ncu-rep" % TAG)
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
                "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum"]
        out = {k: {"value": d[k][0], "unit": d[k][1]} for k in keep if k in d}
        out["pc_sampling_stall_reasons"] = {h.replace("smsp__pcsamp_warps_issue_stalled_", ""): d[h][0]
                                            for h in hdr if h.startswith("smsp__pcsamp_warps_issue_stalled")}
        scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
        traffic = sum(float(d[k][0]) * scale[d[k][1]] for k in ("dram__bytes_read.sum", "dram__bytes_write.sum"))
        out["dram_bytes_per_launch"] = traffic
        out["command"] = ("ncu --set full --clock-control none --import-source on -k regex:series_pipe -s 8 -c 1 "
                          "python bench.py --steps 3 --warmup 3 --no-cpu-baseline")
        json.dump(out, open(os.path.join(PROF, "%s_series_pipe_ncu_full.json" % TAG), "w"), indent=1)
        json.dump({"kernel": "series_pipe", "dram_bytes_per_launch": traffic,
                   "source": "profiles/%s_series_pipe_ncu_full.json" % TAG},
                  open(os.path.join(PROF, "roofline_traffic.json"), "w"), indent=1)
    print(json.dumps(ll, indent=1))
    print("bench:", bench["ms_per_step"], bench["roofline"]["frac"], bench["e2e"]["value"])


if __name__ == "__main__":
    main()
