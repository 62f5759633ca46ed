#!/usr/bin/env python
"""Headline benchmark: samples/sec of a training scheme on N B200s of one node.

    python bench.py --gpus 1 --steps 20 --warmup 5
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 \
        --master-port 29500 bench.py --gpus 8 --steps 20 --warmup 5

Default config = BASELINE.json config 2: Hogwild async SGD on the wide MLP
(784-4096x4-10, 53.6 M parameters), batch 8192 per worker, bf16 compute with fp32
master parameters on the ps shards, synthetic data, random init.  One worker per
GPU; the ps lives in one virtual range striped over the same GPUs' HBM
(``--ps_layout sharded`` keeps whole variables per GPU, ``gpu0`` puts the whole ps
on GPU 0 as config 2 literally says).  Weak scaling: per-GPU batch is
fixed.

Arms: ``--impl native`` (this framework, sm_100a kernels), ``--impl torch`` (same
algorithm with torch matmul/cuBLAS + NCCL: the measured baseline of BASELINE.md),
``--impl reference`` (the unmodified TF-1 reference: cannot run here, prints why).
The native arm, unless ``--no_baseline``, then times the torch arm on the same ranks with the
same steps / warm-up; ``vs_baseline`` is the ratio of those two measurements.

Prints ONE JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "samples/sec"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="native", choices=["native", "torch", "reference"])
    ap.add_argument("--scheme", default="hogwild",
                    choices=["hogwild", "downpour", "downpour_easy", "adag", "ssgd", "ssgd_lr",
                             "sdag", "tower_avg", "dynsgd", "aeasgd", "aeamsgd"])
    ap.add_argument("--model", default="wide_mlp",
                    choices=["wide_mlp", "mnist_mlp", "gpt2_medium", "bert_large", "tiny_gpt",
                             "resnet50", "tiny_resnet"])
    ap.add_argument("--batch", type=int, default=0,
                    help="per-worker batch (default 8192 rows for the MLPs, 8 sequences for GPT-2 "
                         "medium, 16 for BERT-large)")
    ap.add_argument("--lr", type=float, default=1e-3)
    ap.add_argument("--update_window", type=int, default=8)
    ap.add_argument("--ps_layout", default="striped",
                    choices=["sharded", "gpu0", "striped", "gpu0_ps_only"],
                    help="striped: one virtual range, 2 MB stripes rotating over all GPUs (CUDA VMM); "
                         "sharded: whole variables per GPU; gpu0: everything on rank 0 (which also "
                         "trains); gpu0_ps_only: BASELINE config 2 to the letter -- GPU0 holds the ps "
                         "and trains nothing, the other N-1 GPUs are the workers")
    ap.add_argument("--no_baseline", action="store_true",
                    help="native arm only: skip the same-run torch+NCCL arm (vs_baseline = null)")
    ap.add_argument("--post_ln", action="store_true",
                    help="transformer configs: BERT's post-LN block ordering + embedding LayerNorm "
                         "(both arms) instead of the pre-LN block")
    ap.add_argument("--device_init", action="store_true",
                    help="draw the initial parameters on the GPU (ops/csrc/init_kernels.cu)")
    ap.add_argument("--no_fused_push", action="store_true")
    ap.add_argument("--async_push", default="auto", choices=["auto", "0", "1"],
                    help="Hogwild: store gradients locally and push them layer by layer on a side "
                         "stream (auto: on when the whole ps sits behind one GPU's NVLink port)")
    ap.add_argument("--no_overlap", action="store_true")
    ap.add_argument("--no_prefetch_pull", action="store_true")
    ap.add_argument("--skip_e2e", action="store_true")
    return ap.parse_args()


class ClockSampler:
    """nvidia-smi clocks/throttle reasons sampled DURING the timed region."""

    Q = ("timestamp,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.gpu, self.proc, self.lines = gpu_index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.Q,
                 "--format=csv,noheader,nounits", "-lms", "50"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def mark(self, name: str):
        setattr(self, "t_" + name, time.time())

    @staticmethod
    def _ts(s: str) -> float:
        import datetime
        try:
            return datetime.datetime.strptime(s.strip(), "%Y/%m/%d %H:%M:%S.%f").timestamp()
        except ValueError:
            return float("nan")

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        t0, t1 = getattr(self, "t_begin", 0.0), getattr(self, "t_end", float("inf"))
        for l in self.lines:
            f = [x.strip() for x in l.split(",")]
            if len(f) < 9:
                continue
            ts = self._ts(f[0])
            if ts == ts and not (t0 - 0.05 <= ts <= t1 + 0.05):
                continue                      # keep only samples taken inside the timed region
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for n, v in zip(names, f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": statistics.median(sm) if sm else None,
                "sm_max_mhz": max(mx) if mx else None, "reasons": sorted(reasons),
                "samples": len(sm), "sm_mhz_min": min(sm) if sm else None}


def reference_arm(args):
    """The reference is eleven TF-1.x scripts (Python 2.7, tensorflow>=1.2,
    README.md:63-64).  There is no tensorflow wheel for this interpreter in the
    offline wheelhouse and no setup.py/pyproject in the reference tree, so neither
    the install nor the import can succeed; see DESIGN.md."""
    why = "reference is TF 1.x graph scripts (python2/tensorflow>=1.2); no tensorflow wheel offline, " \
          "reference tree has no setup.py -- pip install fails (DESIGN.md)"
    try:
        sys.path.insert(0, os.path.join(ROOT, "baseline", "_ref"))
        import tensorflow as tf  # noqa: F401
        if not hasattr(tf, "train") or not hasattr(tf.train, "Supervisor"):
            why = "tensorflow present but lacks the TF-1 tf.train.Supervisor API the scripts need"
    except Exception as e:
        why = why + " [import tensorflow: %s]" % type(e).__name__
    print(json.dumps({"impl": "reference", "unavailable": why}))
    return 0


MODEL_NAMES = {"wide_mlp": "MLP 784-4096x4-10 (53.6M params)",
               "mnist_mlp": "MLP 784-100-10",
               "gpt2_medium": "GPT-2 medium (24x1024, 355M params)",
               "bert_large": "BERT-large dims (24x1024, 335M params; pre-LN GPT-style blocks, "
                             "bidirectional attention, no token-type embedding / MLM head)",
               "tiny_gpt": "tiny GPT (test)",
               "resnet50": "ResNet-50 v1.5, 224x224 (25.6M params)",
               "tiny_resnet": "tiny ResNet (test)"}
WINDOWED = ("downpour", "downpour_easy", "adag", "sdag", "dynsgd", "aeasgd", "aeamsgd")


class Env:
    """Process-wide state shared by the arms of one run (device, rank, NCCL group)."""

    def __init__(self, args):
        import torch
        import torch.distributed as dist
        self.torch, self.dist = torch, dist
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.rank = int(os.environ.get("RANK", "0"))
        self.local_rank = int(os.environ.get("LOCAL_RANK", "0"))
        if self.world != args.gpus and self.world > 1:
            args.gpus = self.world
        torch.cuda.set_device(self.local_rank)
        self.dev = torch.device("cuda", self.local_rank)
        if self.world > 1:
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            dist.init_process_group("nccl", device_id=self.dev)

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def max_over_ranks(self, v: float) -> float:
        if self.world == 1:
            return v
        t = self.torch.tensor([v], device=self.dev, dtype=self.torch.float64)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())


def model_setup(args):
    """Shapes of the named config (no framework objects yet)."""
    import dtg_b200  # noqa: F401
    is_tf = args.model in ("gpt2_medium", "bert_large", "tiny_gpt")
    is_rn = args.model in ("resnet50", "tiny_resnet")
    sizes = [784, 4096, 4096, 4096, 4096, 10] if args.model == "wide_mlp" else [784, 100, 10]
    B = args.batch or {"gpt2_medium": 8, "bert_large": 16, "tiny_gpt": 4, "resnet50": 256,
                       "tiny_resnet": 32}.get(args.model, 8192)
    rcfg = tcfg = None
    if is_rn:
        from dtg_b200.models import resnet as rnm
        rcfg = (rnm.resnet50() if args.model == "resnet50" else rnm.tiny_resnet()).cfg
    if is_tf:
        from dtg_b200.models import transformer as tfm
        tcfg = {"gpt2_medium": tfm.gpt2_medium, "bert_large": tfm.bert_large,
                "tiny_gpt": tfm.tiny_transformer}[args.model]().cfg
        if args.post_ln:
            # BERT's own block ordering (LN after each residual sum + embedding LayerNorm);
            # same parameter count and FLOPs as the pre-LN rows
            tcfg.post_ln = True
    return is_tf, is_rn, sizes, B, rcfg, tcfg


def worker_lrs(args, world):
    # per-worker learning rates (config 4): worker k trains at lr * (1 + k/8)
    return [args.lr * (1.0 + k / 8.0) for k in range(world)] if args.scheme == "ssgd_lr" else None


def build_native(args, env, setup):
    is_tf, is_rn, sizes, B, rcfg, tcfg = setup
    from dtg_b200 import ops
    from dtg_b200.models.mlp_native import NativeMLP
    from dtg_b200.parallel.engines import SchemeConfig
    from dtg_b200.parallel.fabric import Fabric
    from dtg_b200.parallel.gpu_engines import ASYNC, ELASTIC, WINDOW, GpuWorker
    world, dev = env.world, env.dev
    fabric = Fabric.from_torch_distributed(dev) if world > 1 else Fabric.single(dev)
    asyncish = (args.scheme in ASYNC or args.scheme in WINDOW or args.scheme in ELASTIC
                or args.scheme == "dynsgd")
    sharded = args.ps_layout == "sharded" and asyncish
    striped = args.ps_layout == "striped" and asyncish and world > 1
    if striped and not fabric.striped_supported():      # no VMM here: whole variables per GPU
        striped, sharded, args.ps_layout = False, True, "sharded"
    nsh = world if sharded else 1
    kw = dict(num_shards=nsh, sharding="balanced" if nsh > 1 else "round_robin")
    if is_tf:
        from dtg_b200.models.transformer_native import NativeTransformer
        model = NativeTransformer(tcfg, B, dev, **kw)
    elif is_rn:
        from dtg_b200.models.resnet_native import NativeResNet
        model = NativeResNet(rcfg, B, dev, **kw)
    else:
        model = NativeMLP(sizes, B, dev, **kw)
    cfg = SchemeConfig(scheme=args.scheme, lr=args.lr, update_window=args.update_window,
                       local_lr=args.lr, worker_lrs=worker_lrs(args, world))
    worker = GpuWorker(model, cfg, fabric, ps_ranks=list(range(nsh)) if sharded else [0] * nsh,
                       fused_push=not args.no_fused_push, overlap=not args.no_overlap,
                       prefetch_pull=not args.no_prefetch_pull, striped_ps=striped,
                       async_push={"auto": None, "0": False, "1": True}[args.async_push],
                       device_init=bool(getattr(args, "device_init", False)))
    nvls = bool(getattr(worker.sync, "nvls", False)) if worker.sync is not None else False
    return worker, asyncish, nvls, ops.launch_count, ops.reset_launch_count, model.layout.total


def build_torch(args, env, setup):
    is_tf, is_rn, sizes, B, rcfg, tcfg = setup
    sys.path.insert(0, os.path.join(ROOT, "baseline"))
    from torch_nccl import TorchBaseline, TorchTransformerBaseline, torch_resnet50_baseline
    world, rank, dev = env.world, env.rank, env.dev
    if is_tf:
        worker = TorchTransformerBaseline(tcfg, B, dev, scheme=args.scheme, lr=args.lr,
                                          world=world, rank=rank, worker_lrs=worker_lrs(args, world))
    elif is_rn:
        if args.model != "resnet50":
            raise SystemExit("torch arm implements resnet50 only")
        worker = torch_resnet50_baseline(B, dev, scheme=args.scheme, lr=args.lr,
                                         update_window=args.update_window, world=world, rank=rank)
    else:
        worker = TorchBaseline(sizes, B, dev, scheme=args.scheme, lr=args.lr,
                               update_window=args.update_window, world=world, rank=rank)
    return worker, True, False, (lambda: 0), (lambda: None), worker.st.total


def synthetic_batches(args, env, setup):
    """Pinned host batches; a windowed scheme gets T stacked micro-batches per step (every local
    step trains on fresh data, and the end-to-end arm copies all of them)."""
    torch = env.torch
    is_tf, is_rn, sizes, B, rcfg, tcfg = setup
    T = args.update_window if args.scheme in WINDOWED else 1
    g = torch.Generator().manual_seed(1234 + env.rank)
    n_batches = 2 if (is_rn or T > 1) else 4
    rows = B * T
    if is_tf:
        host_x = [torch.randint(0, tcfg.vocab, (rows, tcfg.seq), generator=g).pin_memory()
                  for _ in range(n_batches)]
        host_y = [torch.randint(0, tcfg.vocab, (rows, tcfg.seq), generator=g).pin_memory()
                  for _ in range(n_batches)]
    elif is_rn:
        # host keeps the 3 real channels; the 5 zero channels that make rows 16-byte
        # multiples are added on the device
        host_x = [torch.randn(rows, rcfg.image, rcfg.image, 3, generator=g).pin_memory()
                  for _ in range(n_batches)]
        host_y = [torch.randint(0, rcfg.classes, (rows,), generator=g).pin_memory()
                  for _ in range(n_batches)]
    else:
        host_x = [torch.randn(rows, sizes[0], generator=g).pin_memory() for _ in range(n_batches)]
        host_y = [torch.randint(0, sizes[-1], (rows,), generator=g).pin_memory()
                  for _ in range(n_batches)]
    return host_x, host_y, B * T


def time_arm(args, env, step_fn, count, reset, host_x, host_y, samples_per_step, active_ranks):
    """Warm-up, K device-timed steps (max over ranks, clocks sampled during the region), then the
    end-to-end pass (pinned H2D every step + loss D2H every step)."""
    torch = env.torch
    dev, rank = env.dev, env.rank
    n_batches = len(host_x)
    dev_x = [t.to(dev) for t in host_x]
    dev_y = [t.to(dev) for t in host_y]
    for i in range(args.warmup):
        step_fn(dev_x[i % n_batches], dev_y[i % n_batches])
    env.barrier()
    sampler = ClockSampler(env.local_rank)
    if rank == 0:
        sampler.start()
        time.sleep(0.3)                       # let nvidia-smi reach its sampling loop
    reset()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    env.barrier()
    sampler.mark("begin")
    e0.record()
    loss = None
    for i in range(args.steps):
        loss = step_fn(dev_x[i % n_batches], dev_y[i % n_batches])
    e1.record()
    env.barrier()
    sampler.mark("end")
    launches = count()
    ms = env.max_over_ranks(e0.elapsed_time(e1))
    clocks = sampler.stop() if rank == 0 else None
    final_loss = float(loss.float().reshape(-1)[0].item()) if loss is not None else None
    value = active_ranks * samples_per_step * args.steps / (ms / 1e3)
    res = {"value": value, "ms_per_step": ms / args.steps, "clocks": clocks,
           "gpu_launches": launches, "final_loss": final_loss, "e2e": None}
    del dev_x, dev_y
    if args.skip_e2e:
        return res
    # the framework's input pipeline (utils/input_pipeline.py): batch i+1 is staged from
    # pinned memory on a copy stream while step i computes; every step's loss is copied to
    # pinned memory and read by the host one step later.  Both arms use it.
    from dtg_b200.utils.input_pipeline import DeviceFeeder, ResultReader
    feeder, reader = DeviceFeeder(dev), ResultReader(dev, depth=2)
    for i in range(2):
        feeder.put(host_x[i % n_batches], host_y[i % n_batches])
        xin, yin = feeder.get()
        step_fn(xin, yin)
        feeder.release()
    torch.cuda.synchronize()
    env.barrier()
    s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    env.barrier()
    s0.record()
    feeder.put(host_x[0], host_y[0])
    n_read = 0
    for i in range(args.steps):
        if i + 1 < args.steps:
            feeder.put(host_x[(i + 1) % n_batches], host_y[(i + 1) % n_batches])
        xin, yin = feeder.get()
        l = step_fn(xin, yin)
        feeder.release()
        if l is not None:
            reader.push(l)
            if reader.pending() > 1:                    # the user reads step i-1's loss
                _ = float(reader.pop()[0])
                n_read += 1
    while reader.pending():
        _ = float(reader.pop()[0])
        n_read += 1
    s1.record()
    env.barrier()
    ms2 = env.max_over_ranks(s0.elapsed_time(s1))
    res["e2e"] = {"value": active_ranks * samples_per_step * args.steps / (ms2 / 1e3),
                  "unit": METRIC,
                  "h2d_bytes_per_step": host_x[0].numel() * host_x[0].element_size() +
                  host_y[0].numel() * host_y[0].element_size(),
                  "d2h_bytes_per_step": 4, "ms_per_step": ms2 / args.steps,
                  "loss_reads": n_read}
    return res


def main():
    args = parse_args()
    if args.impl == "reference":
        return reference_arm(args)
    env = Env(args)
    world, rank = env.world, env.rank
    setup = model_setup(args)
    is_tf, is_rn, sizes, B, rcfg, tcfg = setup
    host_x, host_y, samples_per_step = synthetic_batches(args, env, setup)

    ps_only = args.ps_layout == "gpu0_ps_only"          # config 2 to the letter: GPU0 = ps, N-1 workers
    if ps_only and world < 2:
        raise SystemExit("--ps_layout gpu0_ps_only needs at least 2 GPUs (1 ps + workers)")
    arms = {}
    order = [args.impl] + (["torch"] if args.impl == "native" and not args.no_baseline else [])
    for impl in order:
        layout_saved = args.ps_layout
        if ps_only:
            args.ps_layout = "gpu0"
        worker, asyncish, nvls, count, reset, n_params = (build_native if impl == "native"
                                                          else build_torch)(args, env, setup)
        args.ps_layout = layout_saved
        active = world
        step_fn = worker.step
        if ps_only and impl == "native":
            active = world - 1
            if rank == 0:                      # the ps GPU hosts the shard and trains nothing
                step_fn = lambda x, y: None    # noqa: E731
        arms[impl] = time_arm(args, env, step_fn, count, reset, host_x, host_y, samples_per_step,
                              active)
        arms[impl].update(asyncish=asyncish, nvls=nvls, n_params=n_params, active=active)
        del worker
        env.torch.cuda.empty_cache()

    if rank == 0:
        main_arm = arms[args.impl]
        if args.impl == "native" and not main_arm["asyncish"]:
            where = "fp32 master sharded over %d GPUs, fused reduce-scatter + update + all-gather%s" % (
                world, " (NVLS multimem)" if main_arm["nvls"] else "")
        elif ps_only:
            where = "on GPU0, which trains nothing (1 ps GPU + %d worker GPUs)" % (world - 1)
        elif args.ps_layout == "sharded":
            where = "sharded x%d (colocated)" % world
        elif args.ps_layout == "striped" and world > 1:
            where = "striped over %d GPUs (2 MB stripes, one virtual range)" % world
        else:
            where = "on GPU0"
        par = "%s dp%d, one worker per GPU, ps %s" % (args.scheme, main_arm["active"], where)
        base = arms.get("torch") if args.impl == "native" else None
        # NVLink traffic of the ps data plane, averaged over the step (fp32 pull + fp32 push)
        nvlink = None
        if args.impl == "native" and main_arm["asyncish"] and world > 1:
            pbytes = main_arm["n_params"] * 4.0
            per_step = 1.0 / (main_arm["ms_per_step"] / 1e3)
            T = args.update_window if args.scheme in WINDOWED else 1
            if args.ps_layout in ("gpu0", "gpu0_ps_only"):
                remote_workers = world - 1
                nvlink = {"ps_port_in_GBps": remote_workers * pbytes * per_step / 1e9,
                          "ps_port_out_GBps": remote_workers * pbytes * per_step / 1e9,
                          "roofline_GBps_per_direction": 900.0,
                          "note": "all workers push into / pull from GPU0's one NVLink port"}
            else:
                frac = (world - 1) / world
                nvlink = {"per_gpu_in_GBps": frac * pbytes * per_step / 1e9,
                          "per_gpu_out_GBps": frac * pbytes * per_step / 1e9,
                          "roofline_GBps_per_direction": 900.0}
            nvlink["pushes_per_step"] = 1
            nvlink["local_steps_per_push"] = T
        out = {
            "metric": METRIC, "value": main_arm["value"], "unit": METRIC, "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": main_arm["ms_per_step"],
            "higher_is_better": True, "scaling": "weak",
            # measured in THIS run on the same ranks (torch + NCCL + cuBLAS arm, run right after
            # the native arm with the same steps / warm-up); the reference publishes no number
            "vs_baseline": (main_arm["value"] / base["value"]) if base else None,
            "dtype": "bf16", "data": "synthetic (randn inputs, uniform labels), random-init weights",
            "impl": args.impl,
            "config": {"model": MODEL_NAMES[args.model] if not (is_tf and args.post_ln) else
                       MODEL_NAMES[args.model].replace("pre-LN GPT-style blocks",
                                                       "post-LN BERT blocks + embedding LayerNorm"),
                       "scheme": args.scheme,
                       "global_batch": B * main_arm["active"], "per_worker_batch": B,
                       "seq_len": tcfg.seq if is_tf else None,
                       "tokens_per_sec": main_arm["value"] * tcfg.seq if is_tf else None,
                       "parallelism": par, "lr": args.lr, "ps_layout": args.ps_layout,
                       "update_window": args.update_window if samples_per_step != B else None,
                       "micro_batches_per_step": samples_per_step // B,
                       "l2": "working set per step (fp32 ps shard + fp32/bf16 mirrors + activations "
                             "> 600 MB) exceeds the 126 MB L2; no explicit flush",
                       "master_dtype": "fp32", "nvlink": nvlink},
            "clocks": main_arm["clocks"], "e2e": main_arm["e2e"],
            "gpu_launches": main_arm["gpu_launches"], "final_loss": main_arm["final_loss"],
        }
        if base:
            out["baseline"] = {"impl": "torch+NCCL+cuBLAS (baseline/torch_nccl.py), same run, same ranks",
                               "value": base["value"], "ms_per_step": base["ms_per_step"],
                               "clocks": base["clocks"], "final_loss": base["final_loss"],
                               "e2e": base["e2e"],
                               "e2e_ratio": (main_arm["e2e"]["value"] / base["e2e"]["value"])
                               if main_arm["e2e"] and base["e2e"] else None}
        print(json.dumps(out))
    if world > 1:
        env.dist.barrier()
        env.dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
