"""Where a Hogwild step's time goes at N GPUs: pull, forward, backward with the fused remote
push, backward with local gradient stores, standalone push -- each phase timed with CUDA events
(max over ranks), all ranks in the same phase at the same time (worst-case NVLink contention).
The difference between the two backward rows is what the push costs inside the wgrad GEMMs.

    torchrun --nproc-per-node N --master-addr 127.0.0.1 tools/step_breakdown.py [--ps_layout striped|gpu0|sharded]
"""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--ps_layout", default="striped")
    ap.add_argument("--reps", type=int, default=10)
    args = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    import dtg_b200  # noqa: F401
    from dtg_b200.models.mlp_native import NativeMLP
    from dtg_b200.parallel.engines import SchemeConfig
    from dtg_b200.parallel.fabric import Fabric
    from dtg_b200.parallel.gpu_engines import GpuWorker
    fabric = Fabric.from_torch_distributed(dev) if world > 1 else Fabric.single(dev)
    striped = args.ps_layout == "striped" and world > 1 and fabric.striped_supported()
    sharded = args.ps_layout == "sharded"
    nsh = world if sharded else 1
    model = NativeMLP([784, 4096, 4096, 4096, 4096, 10], 8192, dev, num_shards=nsh,
                      sharding="balanced" if nsh > 1 else "round_robin")
    w = GpuWorker(model, SchemeConfig(scheme="hogwild", lr=1e-3), fabric,
                  ps_ranks=list(range(nsh)) if sharded else [0] * nsh, prefetch_pull=False,
                  striped_ps=striped)
    x, y = model.synthetic_batch()
    m, st = w.model, w.store

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    def timed(fn):
        for _ in range(2):
            fn()
        barrier()
        ts = []
        for _ in range(args.reps):
            barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); fn(); e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        ts.sort()
        t = torch.tensor([ts[len(ts) // 2]], device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    m.set_input(x)
    out = {"world": world, "ps_layout": "striped" if striped else args.ps_layout,
           "params_MB_fp32": model.layout.total * 4 / 1e6}
    out["pull_bf16_plus_bias_ms"] = timed(lambda: st.pull(m.p32, m.p16, w._pull_f32))
    out["pull_fp32_and_bf16_ms"] = timed(lambda: st.pull(m.p32, m.p16))
    out["forward_ms"] = timed(lambda: m.forward(y))
    out["backward_fused_push_ms"] = timed(lambda: m.backward(w._push_sink))
    out["backward_local_store_ms"] = timed(lambda: m.backward(m.local_sink))
    out["standalone_push_ms"] = timed(lambda: st.push_sgd(m.g32, -1e-3))
    out["full_step_serial_ms"] = timed(lambda: w.step(x, y))
    remote = (world - 1) / world if (striped or sharded) else (0.0 if rank == 0 else 1.0)
    gb = model.layout.total * 4 / 1e9
    out["pull_GBps_per_gpu_remote"] = remote * gb / (out["pull_bf16_plus_bias_ms"] / 1e3)
    out["push_GBps_per_gpu_remote_standalone"] = remote * gb / (out["standalone_push_ms"] / 1e3)
    d = out["backward_fused_push_ms"] - out["backward_local_store_ms"]
    out["push_cost_inside_backward_ms"] = d
    if rank == 0:
        print("BREAKDOWN " + json.dumps(out))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
