"""Diagnostic (debug build only): why series leave the fast path, on the bench workload."""
import ctypes, os, sys
sys.path.insert(0, ".")
os.environ["KOROLLUP_LIB"] = os.path.abspath(os.environ.get("KO_DBG_LIB", "scripts/libkorollup_dbg.so"))
import torch
from kubeoperator_b200 import Rollup, synth, korollup
QS = [0.5, 0.95, 0.99]
r = Rollup(0)
lib = korollup.load_library()
n = int(sys.argv[1]) if len(sys.argv) > 1 else 10400
seed = int(sys.argv[2]) if len(sys.argv) > 2 else 1
vals, offs, kinds = synth.make_series_torch("c2", seed, "cuda", n_series_override=n)
n = len(kinds)
q = torch.zeros((n, 3), dtype=torch.float64, device="cuda")
s = torch.zeros(n, dtype=torch.float64, device="cuda")
buf = (ctypes.c_ulonglong * 48)()
lib.ko_debug_phases(buf, 1)
r.series_rollup_dev(vals.data_ptr(), offs.data_ptr(), n, QS, s.data_ptr(), 0, 0, 0, 0, 0, q.data_ptr(),
                    torch.cuda.current_stream().cuda_stream)
torch.cuda.synchronize()
lib.ko_debug_phases(buf, 1)
names = ["few valid samples", "too many intervals", "no levels plan", "bracket missed rank", "overflow queue full",
         "sub-bracket list full", "no candidate at rank", "rank on unlisted level"]
for w in (0, 1):
    print("attempt %d: " % (w + 1) + ", ".join("%s=%d" % (nm, buf[32 + 8 * w + i]) for i, nm in enumerate(names)))
print("kernel_ms", r.last_series_kernel_ms())
