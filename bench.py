#!/usr/bin/env python
"""bench.py — samples/s rolled up by the monitoring-rollup hot path (BASELINE.json metric) on N B200.

One "step" = one pass of the hot path over one batch of synthetic input: the window statistics + p50/p95/p99
of every series of BASELINE.json configs[1] (100-node cluster incl. per-GPU DCGM series, 7 days at 15 s:
10 400 series x 40 320 samples = 419 M samples, 3.35 GB) — per rank (weak scaling: series are sharded by ID,
every rank owns one such shard) — followed, for N > 1, by the fleet-wide quantile merge (all-gather of key
histograms over NCCL).
  value : device-resident (inputs in HBM), CUDA events on the launching stream, max over ranks.
  e2e   : the same step through the host-pointer C-ABI call (ko_series_rollup: pinned host -> chunked H2D ->
          kernels -> D2H results) plus the node / pod rollup of the config through the C ABI.
  --impl reference : the oracle port (plain C, OpenMP, all host threads) on a bounded sample — the reference
          itself is Python and has no range-window code, so the port is the CPU arm.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

QS = [0.5, 0.95, 0.99]
METRIC = "samples_per_sec_rolled_up"
UNIT = "samples/s"
BYTES_PER_SAMPLE = 8.0          # algorithmic bytes per f64 sample


def measured_peak_gbs():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 50 ms during the timed region."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "50"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx.append(float(r[1]))
                for nm, v in zip(names, r[2:6]):
                    if v.lower().startswith("active"):
                        reasons.add(nm)
            except Exception:
                pass
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def cpu_port_baseline(values, offsets, target_seconds=12.0):
    """The C oracle (all host threads) on a bounded sample of the workload; grows the sample until the run
    takes about target_seconds."""
    from oracle import c_oracle
    cores = os.cpu_count() or 1
    n_total = len(offsets) - 1
    n = min(n_total, 64)
    t0 = time.perf_counter()
    c_oracle.series_rollup(values[:offsets[n]], offsets[:n + 1], QS, nthreads=cores)
    dt = max(time.perf_counter() - t0, 1e-4)
    n = int(min(n_total, max(64, n * target_seconds / dt)))
    c_oracle.series_rollup(values[:offsets[n]], offsets[:n + 1], QS, nthreads=cores)   # warm (page faults, threads)
    t0 = time.perf_counter()
    c_oracle.series_rollup(values[:offsets[n]], offsets[:n + 1], QS, nthreads=cores)
    dt1 = max(time.perf_counter() - t0, 1e-4)
    reps = int(max(1, min(200, round(target_seconds / dt1))))   # the whole workload is short on many cores: repeat it
    t0 = time.perf_counter()
    for _ in range(reps):
        c_oracle.series_rollup(values[:offsets[n]], offsets[:n + 1], QS, nthreads=cores)
    dt = time.perf_counter() - t0
    samples = int(offsets[n] - offsets[0])
    return {"value": samples * reps / dt, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": "%d series x %d samples of the same workload, %d passes (%.1f s of CPU work, C oracle "
                      "oracle/korollup_oracle.c with OpenMP on %d threads; the reference has no range-window "
                      "code, SURVEY.md 0.3)" % (n, samples // max(n, 1), reps, dt, cores)}


def run_reference(args, rank, world):
    """--impl reference: the oracle port on the host cores (rank 0 only)."""
    if rank != 0:
        return
    from kubeoperator_b200 import synth
    from oracle import c_oracle
    c_oracle.build()
    cores = os.cpu_count() or 1
    n = 256 if args.sample_series <= 0 else args.sample_series
    d = synth.make_series("c2", seed=1, max_series=n)
    values, offsets = d["values"], d["offsets"]
    samples = int(offsets[-1])
    for _ in range(args.warmup):
        c_oracle.series_rollup(values, offsets, QS, nthreads=cores)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        c_oracle.series_rollup(values, offsets, QS, nthreads=cores)
    dt = time.perf_counter() - t0
    v = samples * args.steps / dt
    sample = ("%d series x 40320 samples of configs[1] per step, C oracle port (oracle/korollup_oracle.c, OpenMP, "
              "%d threads); the Python reference has no range-window implementation" % (n, cores))
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": "configs[1]: 100-node cluster incl. per-GPU DCGM series, 7 d at 15 s; bounded sample",
                   "series_per_step": n, "samples_per_series": 40320, "quantiles": QS},
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))


def run_ours(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist
    from kubeoperator_b200 import Rollup, fleet, synth

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — libkorollup has no CPU fallback (use --impl reference "
                         "for the CPU arm)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    rollup = Rollup(local_rank)
    cfg = synth.CONFIGS["c2"]
    n_series = args.series if args.series > 0 else synth.n_series(cfg)
    L = cfg["L"]
    d_values, d_offsets, kinds = synth.make_series_torch("c2", 1 + rank, dev,
                                                         n_series_override=n_series if args.series > 0 else None)
    n_series = len(kinds)
    samples = n_series * L
    outs = {k: torch.zeros(n_series, dtype=torch.float64, device=dev) for k in ("sum", "mean", "min", "max", "last")}
    d_count = torch.zeros(n_series, dtype=torch.int64, device=dev)
    d_quant = torch.zeros((n_series, len(QS)), dtype=torch.float64, device=dev)
    d_hist = torch.zeros(fleet.HIST_BINS, dtype=torch.int64, device=dev)
    n_nodes = min(cfg["nodes"], n_series // 4)
    node_idx = torch.arange(0, 4 * n_nodes, 4, device=dev)
    stream = torch.cuda.current_stream().cuda_stream
    fleet_q = [None]
    fleet_ws = {}
    fleet_rollup = Rollup(local_rank) if world > 1 else None   # own context: keeps the kernel timers of `rollup`

    def step_dev():
        rollup.series_rollup_dev(d_values.data_ptr(), d_offsets.data_ptr(), n_series, QS, outs["sum"].data_ptr(),
                                 outs["mean"].data_ptr(), outs["min"].data_ptr(), outs["max"].data_ptr(),
                                 outs["last"].data_ptr(), d_count.data_ptr(), d_quant.data_ptr(), stream)
        if world > 1:   # fleet-wide p95 of the per-node mean CPU usage
            node_vals = outs["mean"].index_select(0, node_idx)
            if n_nodes <= fleet.SMALL_FLEET_CAP:     # few values: one all-gather, select on the GPU, no read-back
                fleet_q[0] = fleet.fleet_quantile_device(node_vals, 0.95, n_nodes, fleet_rollup, dist, fleet_ws)
            else:                                    # histogram all-gather, refined 16 bits per round
                nv = int((node_vals == node_vals).sum().item())

                def hist_fn(shift, mask, prefix):
                    rollup.key_histogram_dev(node_vals.data_ptr(), node_vals.numel(), shift, mask, prefix,
                                             d_hist.data_ptr(), stream)
                    return d_hist
                fleet_q[0] = fleet.fleet_quantile(hist_fn, nv, 0.95, dist, dev)

    def sync_all():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---------------- device-resident timing (value)
    for _ in range(max(args.warmup, 3)):
        step_dev()
    sync_all()
    c0 = rollup.counters()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sync_all()
    e0.record()
    for _ in range(args.steps):
        step_dev()
    e1.record()
    sync_all()
    ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms_total = float(ms.item())
    c1 = rollup.counters()
    launches = c1["kernel_launches"] - c0["kernel_launches"]
    value = world * samples * args.steps / (ms_total * 1e-3)

    # ---------------- dominant kernel (series_pipe, plus series_long over the deferred series), CUDA events on its stream inside the library
    kms, pms = [], []
    for _ in range(5):
        step_dev()
        kms.append(rollup.last_series_kernel_ms())
        pms.append(rollup.last_series_pipe_ms())
    k_ms = float(np.mean(kms))      # series_plan + series_pipe + second attempt + series_long
    p_ms = float(np.mean(pms))      # series_pipe alone: the dominant kernel
    peak, peak_src = measured_peak_gbs()
    achieved = BYTES_PER_SAMPLE * samples / (p_ms * 1e-3) / 1e9
    achieved_all = BYTES_PER_SAMPLE * samples / (k_ms * 1e-3) / 1e9
    traffic = None
    try:
        with open(os.path.join(ROOT, "profiles", "roofline_traffic.json")) as f:
            traffic = json.load(f).get("dram_bytes_per_launch")
    except Exception:
        pass

    # ---------------- end to end through the host-pointer C ABI
    host = torch.empty(samples, dtype=torch.float64, pin_memory=True)
    host.copy_(d_values)
    torch.cuda.synchronize()
    h_values = host.numpy()
    h_offsets = d_offsets.cpu().numpy()
    pods = synth.make_pods(7 + rank, cfg["pods"])
    cluster_of_node = np.zeros(n_nodes, np.int32)

    def step_e2e():
        r = rollup.series_rollup(h_values, h_offsets, QS)
        last = np.nan_to_num(r["last"][:4 * n_nodes])
        c = rollup.cluster_rollup(last[1::4], last[2::4], last[0::4], last[3::4], cluster_of_node, 1, 80.0)
        p = rollup.pod_rollup(pods["container_restart"], pods["container_ready"], pods["pod_offsets"],
                              pods["pod_phase_ok"])
        return r, c, p

    for _ in range(2):
        step_e2e()
    e2e_steps = max(3, min(args.steps, 10))
    sync_all()
    c0 = rollup.counters()
    times = []
    t_begin = time.perf_counter()
    for _ in range(e2e_steps):
        t0 = time.perf_counter()
        step_e2e()
        times.append(time.perf_counter() - t0)
    sync_all()
    t_total = time.perf_counter() - t_begin
    c1e = rollup.counters()
    clocks = sampler.stop() if rank == 0 else None     # sampled across both timed regions (device-resident, e2e)
    tt = torch.tensor([t_total], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    e2e_value = world * samples * e2e_steps / float(tt.item())

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms_total / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": "configs[1]: 100-node cluster incl. per-GPU DCGM series, 7 d at 15 s "
                                   "(per rank; series sharded by ID)",
                       "series_per_rank": n_series, "samples_per_series": L, "samples_per_rank": samples,
                       "quantiles": QS, "stats": ["sum", "mean", "min", "max", "last", "count"],
                       "l2": "inputs (%.2f GB per rank) are larger than the 126 MB L2; no flush needed" % (samples * 8 / 1e9),
                       "multi_gpu": "fleet p95 of per-node mean CPU usage: one NCCL all-gather of the per-node means, selected on the GPU (histogram rounds above 65536 nodes per rank)"
                                    if world > 1 else "none (N=1: no NCCL)",
                       "data_gen": "torch on device, distributions of kubeoperator_b200/synth.py"},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "traffic": traffic, "kernel": "series_pipe",
                         "kernel_ms": p_ms, "algorithmic_bytes_per_launch": BYTES_PER_SAMPLE * samples,
                         "peak_source": peak_src,
                         "all_long_series_kernels": {
                             "what": "series_plan + series_pipe + second attempt + series_long (the whole step "
                                     "but classify_series / series_small)",
                             "ms": k_ms, "achieved": achieved_all, "frac": achieved_all / peak}},
            "e2e": {"value": e2e_value, "unit": UNIT,
                    "h2d_bytes_per_step": (c1e["h2d_bytes"] - c0["h2d_bytes"]) // e2e_steps,
                    "d2h_bytes_per_step": (c1e["d2h_bytes"] - c0["d2h_bytes"]) // e2e_steps,
                    "p50_ms": float(np.median(times)) * 1e3, "steps": e2e_steps,
                    "includes": "ko_series_rollup (pinned host -> H2D -> kernels -> D2H) + ko_cluster_rollup + "
                                "ko_pod_rollup through the C ABI"},
            "gpu_launches": int(launches), "clocks": clocks,
            "series_needing_extra_pass": int(c1e["series_fallback"]),
        }
        if world > 1:
            line["fleet_p95_node_cpu_mean"] = float(fleet_q[0].item()) if hasattr(fleet_q[0], "item") else fleet_q[0]
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_port_baseline(h_values, h_offsets)
        else:
            line["cpu_baseline"] = None
        print(json.dumps(line))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if fleet_rollup is not None:
        fleet_rollup.close()
    rollup.close()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--series", type=int, default=0, help="override the number of series per rank (debug)")
    ap.add_argument("--sample-series", type=int, default=0, help="series per step of --impl reference")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    if args.impl == "reference":
        run_reference(args, rank, world)
    else:
        run_ours(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
