"""W16: NVLink micro-benchmarks for the ps data plane (run with torchrun, >= 2 ranks).

Rank 1 acts as the worker, rank 0 owns the ps shard.  Reports GB/s of
  pull      peer loads fp32 -> local fp32+bf16 (ops.pull_convert)
  push      coalesced red.global.add.v4.f32 (ops.push_axpy)
  adagrad   atom.add + red.add per element (ops.push_adagrad)
  gemm-red  the wgrad GEMM whose epilogue pushes -lr*dW into the remote shard,
            against the same GEMM pushing into local memory
  HOTSPOT   (> 2 ranks) every other rank pulls from / pushes into rank 0 at once: what GPU 0's one
            NVLink port carries (plain red.v4 push, and the wgrad GEMM's TMA reduce-add epilogue)
  SYNCBW    the fused reduce-scatter + update + all-gather kernel on 256 MB of gradients per rank,
            peer loads / stores vs NVLS (multimem.ld_reduce / multimem.st)
All timings are CUDA events on the worker's stream (max over ranks where several take part).
"""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def timeit(fn, iters=10, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(iters):
        fn()
    e.record()
    torch.cuda.synchronize()
    return s.elapsed_time(e) / iters


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    import dtg_b200  # noqa: F401
    from dtg_b200 import ops
    from dtg_b200.parallel.fabric import Fabric
    fabric = Fabric.from_torch_distributed(dev)
    n = 64 * 1024 * 1024                      # 256 MB fp32
    shard = fabric.alloc("shard", n * 4)
    acc = fabric.alloc("acc", n * 4)
    remote = shard.tensor(0, torch.float32, 0, n)      # rank 0's memory
    remote_acc = acc.tensor(0, torch.float32, 0, n)
    res = {}
    dist.barrier()
    if rank == 1:
        loc32 = torch.empty(n, device=dev); loc16 = torch.empty(n, device=dev, dtype=torch.bfloat16)
        g = torch.randn(n, device=dev)
        ms = timeit(lambda: ops.pull_convert(remote, loc16, loc32))
        res["pull_GBps"] = n * 4 / ms / 1e6
        ms = timeit(lambda: ops.push_axpy(remote, g, -1e-3))
        res["push_axpy_GBps"] = n * 4 / ms / 1e6
        ms = timeit(lambda: ops.push_adagrad(remote, remote_acc, g, 1e-3, 1.0), iters=3, warm=1)
        res["push_adagrad_GBps"] = n * 8 / ms / 1e6
        M = N = 4096; K = 8192
        dy = torch.randn(K, M, device=dev).bfloat16(); x = torch.randn(K, N, device=dev).bfloat16()
        rview = remote[:M * N].view(M, N)
        lview = torch.zeros(M, N, device=dev)
        ms_l = timeit(lambda: ops.gemm(dy, x, lview, a_mn=True, b_mn=True, alpha=-1e-3,
                                       mode=ops.EPI_F32_RED))
        ms_r = timeit(lambda: ops.gemm(dy, x, rview, a_mn=True, b_mn=True, alpha=-1e-3,
                                       mode=ops.EPI_F32_RED))
        res["gemm_red_local_ms"] = ms_l
        res["gemm_red_remote_ms"] = ms_r
        res["gemm_red_remote_GBps"] = M * N * 4 / ms_r / 1e6
        ms_c = timeit(lambda: ops.gemm(dy, x, rview, a_mn=True, b_mn=True, alpha=-1e-3,
                                       mode=ops.EPI_F32_RED_COAL))
        res["gemm_red_coalesced_remote_ms"] = ms_c
        print("NVLINK " + json.dumps(res), flush=True)
    torch.cuda.synchronize()
    dist.barrier()

    # ---- hot spot: every other rank pulls from / pushes into rank 0 at the same time --------
    # (BASELINE config 2 as written: one ps GPU; the ceiling is GPU 0's one NVLink port)
    if world > 2:
        m = 32 * 1024 * 1024                      # 128 MB per worker
        src = remote[:m]
        l32 = torch.empty(m, device=dev); g2 = torch.randn(m, device=dev)

        def all_time(fn, iters=5):
            for _ in range(2):
                if rank != 0:
                    fn()
            torch.cuda.synchronize(); dist.barrier()
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record()
            if rank != 0:
                for _ in range(iters):
                    fn()
            e.record(); torch.cuda.synchronize()
            t = torch.tensor([s.elapsed_time(e) / iters], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return float(t.item())
        hs = {"workers": world - 1}
        ms = all_time(lambda: ops.pull_convert(src, None, l32))
        hs["pull_through_gpu0_port_GBps"] = (world - 1) * m * 4 / ms / 1e6
        ms = all_time(lambda: ops.push_axpy(src, g2, -1e-3))
        hs["push_red_v4_into_gpu0_GBps"] = (world - 1) * m * 4 / ms / 1e6
        Mh = Nh = 4096
        dyh = torch.randn(8192, Mh, device=dev).bfloat16(); xh = torch.randn(8192, Nh, device=dev).bfloat16()
        rv = remote[:Mh * Nh].view(Mh, Nh)
        ms = all_time(lambda: ops.gemm(dyh, xh, rv, a_mn=True, b_mn=True, alpha=-1e-3,
                                       mode=ops.EPI_F32_RED_COAL))
        hs["wgrad_gemm_tma_reduce_into_gpu0_ms"] = ms
        hs["wgrad_gemm_tma_reduce_into_gpu0_GBps"] = (world - 1) * Mh * Nh * 4 / ms / 1e6
        if rank == 0:
            print("HOTSPOT " + json.dumps(hs), flush=True)
    dist.barrier()

    # ---- fused reduce-scatter + update + all-gather: peer loads/stores vs NVLS multimem -----
    from dtg_b200.parallel.peer_store import PeerSync
    nn = 64 * 1024 * 1024                          # 256 MB of fp32 gradients per rank
    out = {}
    first = 0
    for nv in ("0", "1"):
        os.environ["DTG_NVLS"] = nv
        try:
            sy = PeerSync(fabric, nn, lr=0.01, first_step=first)
        except Exception as ex:                      # no multicast on this host
            out["nvls%s_error" % nv] = str(ex)[:120]
            continue
        first += 1000
        sy.load_master(torch.zeros(nn, device=dev))
        sy.g32.normal_()
        for _ in range(2):
            sy.step()
        torch.cuda.synchronize(); dist.barrier()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        for _ in range(5):
            sy.step()
        e.record(); torch.cuda.synchronize()
        t = torch.tensor([s.elapsed_time(e) / 5], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
        f = (world - 1) / world
        out["nvls_%s" % nv] = {"used_multimem": bool(sy.nvls), "ms": ms,
                               "bus_GBps_per_gpu": (nn * 4 * f + nn * 6 * f) / ms / 1e6,
                               "status": int(sy.status.item())}
        del sy
    if rank == 0:
        print("SYNCBW " + json.dumps({"world": world, "elements": nn, **out}), flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
