"""torchrun worker used by tests/test_gpu_engine.py (2+ GPUs, one rank per GPU)."""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--scheme", default="hogwild")
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--striped", action="store_true")
    ap.add_argument("--async_push", action="store_true")
    ap.add_argument("--gpu0", action="store_true")
    args = ap.parse_args()
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    import dtg_b200  # noqa: F401
    from dtg_b200.models.mlp_native import NativeMLP
    from dtg_b200.parallel.engines import SchemeConfig
    from dtg_b200.parallel.fabric import Fabric
    from dtg_b200.parallel.gpu_engines import ASYNC, ELASTIC, WINDOW, GpuWorker

    fabric = Fabric.from_torch_distributed(dev)
    asyncish = args.scheme in ASYNC or args.scheme in WINDOW or args.scheme in ELASTIC or \
        args.scheme == "dynsgd"
    nsh = world if asyncish and not args.striped and not args.gpu0 else 1
    model = NativeMLP([784, 512, 256, 10], 512, dev, num_shards=nsh,
                      sharding="balanced" if nsh > 1 else "round_robin")
    cfg = SchemeConfig(scheme=args.scheme, lr=0.02, local_lr=0.02, update_window=2,
                       worker_lrs=[0.03, 0.01] + [0.01] * (world - 2) if args.scheme == "ssgd_lr" else None)
    w = GpuWorker(model, cfg, fabric, ps_ranks=list(range(nsh)), striped_ps=args.striped,
                  async_push=True if args.async_push else None)
    g = torch.Generator().manual_seed(100 + rank)
    x, y = model.synthetic_batch(generator=g)
    losses = []
    for _ in range(args.steps):
        losses.append(float(w.step(x, y).item()))
    torch.cuda.synchronize()
    dist.barrier()
    ok = losses[-1] < losses[0]
    info = {"rank": rank, "first": losses[0], "last": losses[-1], "status": w.status()}
    if not asyncish:
        info["nvls"] = bool(w.sync.nvls)
        # synchronous replicas must hold bit-identical parameters
        mine = w.sync.p32.clone()
        ref = mine.clone()
        dist.broadcast(ref, 0)
        info["replicas_identical"] = bool(torch.equal(mine, ref))
        ok = ok and info["replicas_identical"] and w.status() == 0
        info["global_step"] = w.global_step()
        ok = ok and info["global_step"] == args.steps
    else:
        dist.barrier()
        if args.striped:
            # one flat ps striped over both GPUs: every rank must read back the same bytes,
            # and a checkpoint gather must see them too
            mine = torch.empty_like(model.p32)
            w.store.pull(mine, None)
            ref = mine.clone()
            dist.broadcast(ref, 0)
            info["striped_identical"] = bool(torch.equal(mine, ref))
            named = w.store.gather_named()
            k0 = model.layout.order[0]
            info["gather_ok"] = bool(torch.equal(
                named[k0].reshape(-1), model.layout.unpack(mine.cpu())[k0].reshape(-1)))
            ok = ok and info["striped_identical"] and info["gather_ok"]
        info["global_step"] = w.global_step()
        ok = ok and info["global_step"] == args.steps * world     # every push counted
    flags = [None] * world
    dist.all_gather_object(flags, (ok, info))
    if rank == 0:
        print("RESULT " + json.dumps({"ok": all(f[0] for f in flags), "ranks": [f[1] for f in flags]}))
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
