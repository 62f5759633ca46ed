"""What would a pull fused into the forward GEMM cost?  Measured, not modelled: the forward GEMM
of one wide-MLP layer (8192 x 4096 x 4096) with its WEIGHT OPERAND READ STRAIGHT FROM THE PEER
GPU's HBM by the GEMM's own TMA loads (plain, non-multicast TMA from peer memory), against the
same GEMM on a local bf16 mirror plus the one-pass pull_convert that produces the mirror.
The bf16 peer read is the friendly case: the real thing would read fp32 (twice the bytes) and
convert in a producer warpgroup.

    torchrun --nproc-per-node 2 --master-addr 127.0.0.1 tools/fused_pull_probe.py
"""
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


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
    M, N, K = 8192, 4096, 4096
    wbuf = fabric.alloc("w_bf16", N * K * 2)
    w32buf = fabric.alloc("w_f32", N * K * 4)
    wbuf.local(torch.bfloat16, 0, N * K).copy_((torch.randn(N * K, device=dev) * 0.02).bfloat16())
    w32buf.local(torch.float32, 0, N * K).copy_(torch.randn(N * K, device=dev) * 0.02)
    torch.cuda.synchronize()
    dist.barrier()
    peer = (rank + 1) % world
    W_local = wbuf.local(torch.bfloat16, 0, N * K).view(N, K)
    W_peer = wbuf.tensor(peer, torch.bfloat16, 0, N * K).view(N, K)
    W32_peer = w32buf.tensor(peer, torch.float32, 0, N * K)
    x = torch.randn(M, K, device=dev).bfloat16()
    y = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
    mirror = torch.empty(N * K, device=dev, dtype=torch.bfloat16)

    def timed(fn, reps=10):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / reps], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    out = {"shape": [M, N, K], "world": world}
    out["gemm_local_weights_ms"] = timed(lambda: ops.gemm(x, W_local, y))
    out["gemm_weights_read_from_peer_by_TMA_ms"] = timed(lambda: ops.gemm(x, W_peer, y))
    out["pull_convert_fp32_peer_to_bf16_ms"] = timed(lambda: ops.pull_convert(W32_peer, mirror, None))
    ref = x.float() @ W_peer.float().t()
    ops.gemm(x, W_peer, y)
    torch.cuda.synchronize()
    out["peer_gemm_rel_err"] = ((y.float() - ref).abs().max() / ref.abs().max()).item()
    out["separate_total_ms"] = out["gemm_local_weights_ms"] + out["pull_convert_fp32_peer_to_bf16_ms"]
    tiles_m = M // 256
    out["peer_bytes_read_by_fused_gemm_MB"] = tiles_m * N * K * 2 / 1e6
    out["peer_bytes_read_by_pull_MB"] = N * K * 4 / 1e6
    if rank == 0:
        print("FUSEDPULL " + json.dumps(out))
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
