import sys, time, numpy as np
sys.path.insert(0, ".")
import torch
from kubeoperator_b200 import Rollup
QS=[0.5,0.95,0.99]
r=Rollup(0); rng=np.random.default_rng(5)
for L in (20,240):
    n=1000000
    values=rng.beta(2.0,5.0,n*L); offsets=np.arange(n+1,dtype=np.int64)*L
    dv,do=torch.from_numpy(values).cuda(),torch.from_numpy(offsets).cuda()
    outs=[torch.zeros(n,dtype=torch.float64,device="cuda") for _ in range(5)]
    cnt=torch.zeros(n,dtype=torch.int64,device="cuda"); qt=torch.zeros((n,3),dtype=torch.float64,device="cuda")
    st=torch.cuda.current_stream().cuda_stream
    def dev():
        r.series_rollup_dev(dv.data_ptr(),do.data_ptr(),n,QS,*[o.data_ptr() for o in outs],cnt.data_ptr(),qt.data_ptr(),st); torch.cuda.synchronize()
    dev(); best=1e9
    for _ in range(5):
        t=time.perf_counter(); dev(); best=min(best,time.perf_counter()-t)
    print(L, "ms %.3f"%(best*1e3), "GB/s %.0f"%(n*L*8/best/1e9))
