"""Debug build only: diag_long's batch, with the reasons series leave the fast path and the per-kernel times."""
import ctypes, os, sys
import numpy as np
sys.path.insert(0, ".")
os.environ["KOROLLUP_LIB"] = os.path.abspath("scripts/libkorollup_dbg.so")
import torch
from kubeoperator_b200 import Rollup, synth, korollup
L = int(sys.argv[1]) if len(sys.argv) > 1 else 40320
n = int(sys.argv[2]) if len(sys.argv) > 2 else 3100
QS = [0.5, 0.95, 0.99]
r = Rollup(0)
lib = korollup.load_library()
rng = np.random.default_rng(0)
vals = np.concatenate([synth._one_series(rng, i % 4 if i % 4 != 2 else 0, L) for i in range(n)])
offs = np.arange(n + 1, dtype=np.int64) * L
dv, do = torch.from_numpy(vals).cuda(), torch.from_numpy(offs).cuda()
q = torch.zeros((n, 3), dtype=torch.float64, device="cuda")
s = torch.zeros(n, dtype=torch.float64, device="cuda")
buf = (ctypes.c_ulonglong * 48)()
for _ in range(2):
    lib.ko_debug_phases(buf, 1)
    r.series_rollup_dev(dv.data_ptr(), do.data_ptr(), n, QS, s.data_ptr(), 0, 0, 0, 0, 0, q.data_ptr(),
                        torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
lib.ko_debug_phases(buf, 1)
names = ["few valid samples", "too many intervals", "no levels plan", "bracket missed rank", "overflow queue full",
         "sub-bracket list full", "no candidate at rank", "rank on unlisted level"]
for w in (0, 1):
    print("attempt %d: " % (w + 1) + ", ".join("%s=%d" % (nm, buf[32 + 8 * w + i]) for i, nm in enumerate(names)))
print("L=%d n=%d all_ms=%.3f pipe_ms=%.3f" % (L, n, r.last_series_kernel_ms(), r.last_series_pipe_ms()))
