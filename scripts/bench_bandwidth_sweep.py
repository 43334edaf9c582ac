#!/usr/bin/env python
"""BASELINE.json config 5: peer-aggregation bandwidth sweep for a widened SAE (1-512 MB of parameters per
client): the in-kernel P2P weighted reduce (reduce-scatter + all-gather by pulling, and the naive pull-all)
against NCCL (all_gather of the models + local weighted sum, and local partial + all_reduce).

  torchrun --nproc-per-node 2 scripts/bench_bandwidth_sweep.py [--sizes-mb 1 4 16 64 256 512] [--clients-per-gpu 1]

Device-timed with CUDA events, max over ranks.  Reports the achieved algorithm bandwidth (model bytes /
time) and, for the P2P reduce, the fraction of the NVLink roof: 2 (N-1)/N * P bytes per GPU at 770 GB/s
(measured peer-copy bandwidth per direction on this pool, B200_PROFILING.md).
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from fedmse_decentralized_b200.parallel import init_distributed, shutdown
from fedmse_decentralized_b200.parallel.p2p_reduce import P2PReducer

NVLINK_GBS = 770.0


def timeit(fn, iters, comm):
    for _ in range(3):
        fn()
    torch.cuda.synchronize(); comm.barrier(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / iters], device=comm.device)
    comm.max_(t)
    return float(t[0])


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sizes-mb", type=float, nargs="+", default=[1, 4, 16, 64, 256, 512])
    ap.add_argument("--clients-per-gpu", type=int, default=1)
    ap.add_argument("--iters", type=int, default=10)
    a = ap.parse_args()
    comm = init_distributed()
    W, m, dev = comm.world, a.clients_per_gpu, comm.device
    for mb in a.sizes_mb:
        P = int(mb * 2 ** 20 / 4) // (4 * W) * (4 * W)
        red = P2PReducer(comm, P, m)
        g = torch.Generator(device="cpu").manual_seed(comm.rank)
        red.rows.tensor.copy_(torch.randn(m, P, generator=g).to(dev))
        w_all = torch.rand(W * m, generator=torch.Generator().manual_seed(7)).to(dev)
        w_all /= w_all.sum()
        w_loc = w_all[comm.rank * m:(comm.rank + 1) * m].contiguous()
        torch.cuda.synchronize(); comm.barrier()
        # reference result through NCCL
        part = (red.rows.tensor * w_loc[:, None]).sum(0)
        want = part.clone()
        if comm.is_dist:
            dist.all_reduce(want)
        got = red.allreduce(w_loc).clone()
        torch.cuda.synchronize(); comm.barrier()
        got2 = red.pull_all(w_all).clone()
        torch.cuda.synchronize()
        has_nvls = comm.is_dist and red.enable_nvls()
        err3 = float("nan")
        if has_nvls:
            got3 = red.allreduce_nvls(w_loc).clone()
            torch.cuda.synchronize(); comm.barrier()
            err3 = float((got3 - want).abs().max() / want.abs().max())
        err = float((got - want).abs().max() / want.abs().max())
        err2 = float((got2 - want).abs().max() / want.abs().max())

        def nccl_allgather():
            buf = torch.empty(W * m, P, device=dev)
            if comm.is_dist:
                dist.all_gather_into_tensor(buf, red.rows.tensor)
            else:
                buf.copy_(red.rows.tensor)
            return (buf * w_all[:, None]).sum(0)

        def nccl_allreduce():
            p = (red.rows.tensor * w_loc[:, None]).sum(0)
            if comm.is_dist:
                dist.all_reduce(p)
            return p

        def pull_all():
            comm.barrier() if False else None
            return red.pull_all(w_all)

        t_p2p = timeit(lambda: red.allreduce(w_loc), a.iters, comm)
        t_nvls = timeit(lambda: red.allreduce_nvls(w_loc), a.iters, comm) if has_nvls else float("nan")
        t_pull = timeit(pull_all, a.iters, comm)
        t_ag = timeit(nccl_allgather, a.iters, comm)
        t_ar = timeit(nccl_allreduce, a.iters, comm)
        roof_ms = (2 * (W - 1) / W * P * 4) / (NVLINK_GBS * 1e9) * 1e3 if W > 1 else float("nan")
        if comm.rank == 0:
            print(json.dumps({"model_mb": P * 4 / 2 ** 20, "gpus": W, "clients_per_gpu": m,
                              "p2p_rs_ag_ms": round(t_p2p, 4), "p2p_pull_all_ms": round(t_pull, 4),
                              "nccl_allgather_sum_ms": round(t_ag, 4), "nccl_partial_allreduce_ms": round(t_ar, 4),
                              "speedup_vs_nccl_allgather": round(t_ag / t_p2p, 2),
                              "speedup_vs_nccl_allreduce": round(t_ar / t_p2p, 2),
                              "nvlink_roof_ms": round(roof_ms, 4), "frac_of_nvlink_roof": round(roof_ms / t_p2p, 3),
                              "nvls_ms": round(t_nvls, 4) if has_nvls else None,
                              "nvls_speedup_vs_nccl_allreduce": round(t_ar / t_nvls, 2) if has_nvls else None,
                              "nvls_frac_of_nvlink_roof": round(roof_ms / t_nvls, 3) if has_nvls else None,
                              "max_rel_err_nvls": err3 if has_nvls else None,
                              "max_rel_err": err, "max_rel_err_pull_all": err2}))
        red.close()
        del red
        torch.cuda.empty_cache()
    shutdown()


if __name__ == "__main__":
    main()
