"""Two ranks: where the fleet step's time goes.  torchrun --nproc-per-node 2 scripts/diag_fleet2.py"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402
from kubeoperator_b200 import Rollup, fleet, synth  # noqa: E402

QS = [0.5, 0.95, 0.99]


def main():
    rank, lr = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(lr)
    dev = torch.device("cuda", lr)
    dist.init_process_group("nccl", device_id=dev)
    world = dist.get_world_size()
    r, fr = Rollup(lr), Rollup(lr)
    d_values, d_offsets, kinds = synth.make_series_torch("c2", 1 + rank, dev)
    n = len(kinds)
    outs = [torch.zeros(n, dtype=torch.float64, device=dev) for _ in range(5)]
    cnt = torch.zeros(n, dtype=torch.int64, device=dev)
    qt = torch.zeros((n, 3), dtype=torch.float64, device=dev)
    node_idx = torch.arange(0, 400, 4, device=dev)
    stream = torch.cuda.current_stream().cuda_stream
    ws = {}
    send = torch.zeros(100, dtype=torch.float64, device=dev)
    recv = torch.zeros(100 * world, dtype=torch.float64, device=dev)

    def local():
        r.series_rollup_dev(d_values.data_ptr(), d_offsets.data_ptr(), n, QS, *[o.data_ptr() for o in outs],
                            cnt.data_ptr(), qt.data_ptr(), stream)

    def local_gather():
        local()
        dist.all_gather_into_tensor(recv, send)

    def gather_only():
        dist.all_gather_into_tensor(recv, send)

    def full():
        local()
        nv = outs[1].index_select(0, node_idx)
        fleet.fleet_quantile_device(nv, 0.95, 100, fr, dist, ws)

    def timed(fn, reps=30):
        for _ in range(5):
            fn()
        dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / reps

    for name, fn in (("local", local), ("all_gather only", gather_only), ("local + all_gather", local_gather),
                     ("full step", full)):
        t = timed(fn)
        print("rank %d  %-20s %.3f ms" % (rank, name, t), flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
