import ctypes, os, sys
import numpy as np
sys.path.insert(0, ".")
os.environ["KOROLLUP_LIB"] = os.path.abspath("scripts/libkorollup_dbg.so")
import torch
from kubeoperator_b200 import Rollup, synth, korollup
L, n = 40320, 2960
r = Rollup(0)
lib = korollup.load_library()
rng = np.random.default_rng(0)
vals = np.concatenate([synth._one_series(rng, 0, L) for _ in range(n)])
offs = np.arange(n + 1, dtype=np.int64) * L
dv, do = torch.from_numpy(vals).cuda(), torch.from_numpy(offs).cuda()
q = torch.zeros((n, 4), dtype=torch.float64, device="cuda")
s = torch.zeros(n, dtype=torch.float64, device="cuda")
names = ["fetch", "sample+sort", "plan", "stream", "reduce+decide", "select", "finish"]
for qs in ([], [0.5], [0.5, 0.95, 0.99]):
    buf = (ctypes.c_ulonglong * 48)()
    r.series_rollup_dev(dv.data_ptr(), do.data_ptr(), n, qs, s.data_ptr(), 0, 0, 0, 0, 0, q.data_ptr() if qs else 0,
                        torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    lib.ko_debug_phases(buf, 1)
    r.series_rollup_dev(dv.data_ptr(), do.data_ptr(), n, qs, s.data_ptr(), 0, 0, 0, 0, 0, q.data_ptr() if qs else 0,
                        torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    ms = r.last_series_kernel_ms()
    lib.ko_debug_phases(buf, 1)
    tot = sum(buf[:7])
    print("qs=%s kernel_ms=%.3f CTA0 cycles total=%d (%.1f us at 1.9GHz)" % (qs, ms, tot, tot / 1900.0))
    print("   " + "  ".join("%s=%.1f%%(%.1fus)" % (nm, 100.0 * buf[i] / max(tot, 1), buf[i] / 1900.0) for i, nm in enumerate(names)))
    print("   finish parts (us): prologue=%.1f groups=%.1f final_sync=%.1f | scan1=%.1f totals=%.1f walk+zero=%.1f scan2=%.1f rank=%.1f" % tuple(buf[i] / 1900.0 for i in (13, 14, 7, 24, 25, 26, 27, 28)))
    print("   for_each calls=%d max nmax=%d" % (buf[23], buf[22]))
    print("   for_each loops: %.1f us total, %d iterations -> %.0f cycles/iteration" % (buf[20] / 1900.0, buf[21], buf[20] / max(buf[21], 1)))
    print("   pipe CTA0 (us): planner=%.1f stream=%.1f finish=%.1f stream_wait_plan=%.1f finish_wait_pool=%.1f stream_wait_data=%.1f" % tuple(buf[i] / 1900.0 for i in (8, 9, 10, 11, 12, 16)))
