"""The B200 path end to end: one process per GPU (torchrun), a scheme engine over NVLink peer
memory, the TF-style session lifecycle (hooks, StopAtStep, chief checkpoints by variable name,
resume) and the input pipeline.

  python examples/gpu_train.py --scheme hogwild --last_step 200                      # one GPU
  python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 \\
      examples/gpu_train.py --scheme ssgd --model gpt2_tiny --last_step 100 --checkpoint_dir logdir
"""
import argparse
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import dtg_b200 as dtg  # noqa: E402,F401
from dtg_b200.parallel.engines import SchemeConfig  # noqa: E402
from dtg_b200.parallel.fabric import Fabric  # noqa: E402
from dtg_b200.parallel.gpu_engines import ASYNC, ELASTIC, WINDOW, GpuWorker  # noqa: E402
from dtg_b200.parallel.gpu_session import GpuTrainingSession  # noqa: E402
from dtg_b200.session.hooks import StepCounterHook  # noqa: E402
from dtg_b200.utils.input_pipeline import DeviceFeeder, ResultReader  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--scheme", default="hogwild")
ap.add_argument("--model", default="mlp", choices=["mlp", "gpt2_tiny"])
ap.add_argument("--last_step", type=int, default=200)
ap.add_argument("--lr", type=float, default=0.02)
ap.add_argument("--checkpoint_dir", default=None)
ap.add_argument("--profile_dir", default=None,
                help="chrome-trace timelines + per-phase GB/s and roofline fractions (utils/trace.py)")
args = ap.parse_args()

world = int(os.environ.get("WORLD_SIZE", "1"))
local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    import torch.distributed as dist
    dist.init_process_group("nccl", device_id=dev)       # rendezvous only; no collective on the data path
    fabric = Fabric.from_torch_distributed(dev)
else:
    fabric = Fabric.single(dev)

asyncish = args.scheme in ASYNC or args.scheme in WINDOW or args.scheme in ELASTIC or args.scheme == "dynsgd"
striped = asyncish and world > 1 and fabric.striped_supported()
if args.model == "mlp":
    from dtg_b200.models.mlp_native import NativeMLP
    model = NativeMLP([784, 1024, 1024, 10], 1024, dev)
else:
    from dtg_b200.models.transformer import tiny_transformer
    from dtg_b200.models.transformer_native import NativeTransformer
    model = NativeTransformer(tiny_transformer().cfg, 8, dev)
cfg = SchemeConfig(scheme=args.scheme, lr=args.lr, local_lr=args.lr, update_window=2)
worker = GpuWorker(model, cfg, fabric, ps_ranks=[0], striped_ps=striped)

# a few pinned host batches stand in for a data loader
g = torch.Generator().manual_seed(fabric.rank)
batches = [tuple(t.cpu().pin_memory() for t in model.synthetic_batch(generator=g)) for _ in range(4)]
feeder, reader = DeviceFeeder(dev), ResultReader(dev)
hooks = [StepCounterHook(every_n_steps=50)] if worker.is_chief else []
if args.profile_dir:
    from dtg_b200.utils.trace import ProfilerHook
    hooks.append(ProfilerHook(worker.tracer, save_steps=50,
                              output_dir=os.path.join(args.profile_dir, "rank%d" % fabric.rank)))
with GpuTrainingSession(worker, last_step=args.last_step, checkpoint_dir=args.checkpoint_dir,
                        save_checkpoint_secs=5 if args.checkpoint_dir else None, hooks=hooks) as sess:
    if sess.restored_from:
        print("restored", sess.restored_from)
    i = 0
    feeder.put(*batches[0])
    while not sess.should_stop():
        feeder.put(*batches[(i + 1) % len(batches)])
        x, y = feeder.get()
        reader.push(sess.step(x, y))
        feeder.release()
        if reader.pending() > 1:
            loss = float(reader.pop()[0])
            if i % 50 == 0 and worker.is_chief:
                print("step %d loss %.4f global_step %d" % (i, loss, worker.global_step()))
        i += 1
print("worker %d done after %d local steps, global_step %d" % (fabric.rank, i, worker.global_step()))
if world > 1:
    dist.barrier()
    dist.destroy_process_group()
