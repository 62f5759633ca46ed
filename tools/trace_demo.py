"""Phase timelines of real configs (utils/trace.py): runs a few steps of the named bench configs
with the tracer on and writes, per config and rank, a chrome-trace timeline and the per-phase
summary (ms, GB/s against the 900 GB/s NVLink port, TFLOP/s against MEASURED_PEAKS.json).

    torchrun --nproc-per-node N --master-addr 127.0.0.1 tools/trace_demo.py [--out gpurun_out/trace]
"""
import argparse
import json
import os
import sys
import types

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

CONFIGS = {
    "hogwild_wide_mlp": dict(model="wide_mlp", scheme="hogwild", batch=8192, lr=1e-3, steps=20),
    "adag_wide_mlp": dict(model="wide_mlp", scheme="adag", batch=8192, lr=1e-3, steps=6),
    "ssgd_lr_gpt2_medium": dict(model="gpt2_medium", scheme="ssgd_lr", batch=8, lr=1e-3, steps=6),
}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default="gpurun_out/trace")
    ap.add_argument("--configs", default=",".join(CONFIGS))
    a = ap.parse_args()
    env = bench.Env(types.SimpleNamespace(gpus=int(os.environ.get("WORLD_SIZE", "1"))))
    from dtg_b200.utils.trace import PhaseTracer
    for name in a.configs.split(","):
        cfg = dict(CONFIGS[name])
        steps = cfg.pop("steps")
        args = types.SimpleNamespace(gpus=env.world, steps=steps, warmup=0, impl="native",
                                     no_fused_push=False, no_overlap=False, no_prefetch_pull=False,
                                     skip_e2e=True, ps_layout="striped", no_baseline=True,
                                     async_push="auto", update_window=8, post_ln=False, **cfg)
        setup = bench.model_setup(args)
        host_x, host_y, _ = bench.synthetic_batches(args, env, setup)
        xs = [t.to(env.dev) for t in host_x]
        ys = [t.to(env.dev) for t in host_y]
        worker, *_ = bench.build_native(args, env, setup)
        worker.tracer = PhaseTracer(enabled=True, device=env.dev, rank=env.rank)
        for i in range(steps):
            worker.step(xs[i % len(xs)], ys[i % len(ys)])
        env.barrier()
        tr = worker.tracer
        d = os.path.join(a.out, name)
        tr.write_chrome_trace(os.path.join(d, "timeline-rank%d.json" % env.rank), skip_steps=2)
        summ = tr.summary(skip_steps=2)
        if env.rank == 0:
            os.makedirs(d, exist_ok=True)
            with open(os.path.join(d, "phases-rank0.json"), "w") as f:
                json.dump({"config": name, "world": env.world, "phases": summ}, f, indent=1)
            print("TRACE " + json.dumps({"config": name, "world": env.world, "phases": {
                k: {q: (round(v[q], 4) if isinstance(v[q], float) else v[q]) for q in v
                    if q in ("calls", "ms_mean", "GBps", "fraction_of_roofline", "TFLOPs",
                             "fraction_of_bf16_peak", "lane")} for k, v in summ.items()}}), flush=True)
        del worker, xs, ys
        torch.cuda.empty_cache()
        env.barrier()
    if env.world > 1:
        env.dist.destroy_process_group()


if __name__ == "__main__":
    main()
