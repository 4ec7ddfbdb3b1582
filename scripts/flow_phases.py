"""Where CTA 0 of series_flow spends its cycles, per role (needs a build with KO_NVCC_EXTRA=-DKO_FLOW_TIMING).
python scripts/flow_phases.py L n_series [nq]"""
import ctypes
import sys

sys.path.insert(0, ".")
import torch
from kubeoperator_b200 import Rollup, synth, korollup

L = int(sys.argv[1]) if len(sys.argv) > 1 else 40320
n = int(sys.argv[2]) if len(sys.argv) > 2 else 2960
nq = int(sys.argv[3]) if len(sys.argv) > 3 else 3
QS = [0.5, 0.95, 0.99][:nq]
r = Rollup(0)
lib = korollup.load_library()
cfg = "c2" if L <= 40320 else "c3"
import os, numpy as np
kk = os.environ.get("KO_KINDS")
kinds = synth.mix_kinds(cfg, n) if not kk else np.resize(np.array([int(x) for x in kk.split(",")], np.int8), n)
vals, offs, kinds = synth.make_series_torch(cfg, 1, "cuda", L=L, kinds=kinds)
q = torch.zeros((n, 3), dtype=torch.float64, device="cuda")
s = torch.zeros(n, dtype=torch.float64, device="cuda")
out = (ctypes.c_ulonglong * 64)()
for it in range(3):
    if it == 2:
        torch.cuda.synchronize()
        lib.ko_debug_flow_phases(out, 1)
    r.series_rollup_dev(vals.data_ptr(), offs.data_ptr(), n, QS, s.data_ptr(), 0, 0, 0, 0, 0, q.data_ptr(),
                        torch.cuda.current_stream().cuda_stream)
torch.cuda.synchronize()
lib.ko_debug_flow_phases(out, 0)
pms = r.last_series_pipe_ms()
names = {0: "producer wait desc", 1: "producer wait empty stage", 2: "producer issue", 5: "planner wait finish",
         6: "planner plan (total)", 10: "  plan: gather", 11: "  plan: round 1 + targets", 12: "  plan: round 2",
         13: "  plan: heavy levels", 14: "  plan: thread 0", 20: "finish wait plan", 21: "finish wait stream",
         22: "finish work", 23: "  finish A: partials, last sample", 24: "  finish B: start state", 25: "  finish C: rounds (settling part)", 26: "  finish C: number of rounds", 27: "  finish C: spilled candidates + sync", 28: "  finish C: clearing the bins", 29: "  finish C: prefilter + candidates in shared memory", 30: "stream0 wait plan", 31: "stream0 series work (all)", 32: "  stream0 wait full stage",
         33: "  stream0 scan", 34: "  stream0 store candidates"}
per = max(1, n // 148)
print("L=%d n=%d nq=%d series_flow %.3f ms (%.0f GB/s), ~%d series per CTA; kcycles per series in CTA 0:" %
      (L, n, nq, pms, vals.numel() * 8 / pms / 1e6, per))
for k in sorted(names):
    print("  %-30s %9.1f" % (names[k], out[k] / 1e3 / per))
