"""Device heartbeat + watchdog across ranks (parallel/heartbeat.py).

    torchrun --nproc-per-node N --master-addr 127.0.0.1 tools/heartbeat_check.py

Every rank trains a small MLP under Hogwild for a few steps and beats; the last rank then stops
(its process stays up, as a hung worker would) while the others go on.  Rank 0's watchdog, reading
the heartbeat words through peer memory, has to name exactly that rank, and the asynchronous
scheme has to keep running without it.
"""
import json
import os
import sys
import time

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    import dtg_b200  # noqa: F401
    from dtg_b200.models.mlp_native import NativeMLP
    from dtg_b200.parallel.engines import SchemeConfig
    from dtg_b200.parallel.fabric import Fabric
    from dtg_b200.parallel.gpu_engines import GpuWorker
    from dtg_b200.parallel.heartbeat import DeviceHeartbeat, Watchdog
    fabric = Fabric.from_torch_distributed(dev)
    model = NativeMLP([784, 512, 10], 256, dev)
    w = GpuWorker(model, SchemeConfig(scheme="hogwild", lr=0.01), fabric, ps_ranks=[0])
    hb = DeviceHeartbeat(fabric)
    wd = Watchdog(hb.read, timeout_secs=1.0)
    x, y = model.synthetic_batch()
    quiet = world - 1
    for phase, steps in (("all", 20), ("one quiet", 20)):
        t_end = time.time() + 2.5
        while time.time() < t_end:
            if not (phase == "one quiet" and rank == quiet and world > 1):
                w.step(x, y)
                hb.beat()
            torch.cuda.synchronize()
            time.sleep(0.05)
            if rank == 0:
                wd.poll()
        torch.cuda.synchronize()
        dist.barrier()                 # nobody marks itself done while rank 0 still judges
        if rank == 0:
            want = [] if phase == "all" or world == 1 else [quiet]
            ok = wd.stale == want
            print("HEARTBEAT " + json.dumps({"phase": phase, "world": world, "stale": wd.stale,
                                             "expected": want, "beats": hb.read(), "ok": ok}),
                  flush=True)
            assert ok
    hb.mark_done()
    torch.cuda.synchronize()
    dist.barrier()
    if rank == 0:
        wd.poll()
        print("HEARTBEAT " + json.dumps({"phase": "done", "done": wd.done, "stale": wd.stale}), flush=True)
        assert wd.done == list(range(world)) and wd.stale == []
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
