"""``python -m dtg_b200`` -- run one role of a cluster, or launch all of them.

Reference launch style (kept verbatim)::

    python -m dtg_b200 --scheme hogwild --job_name "ps"     --task_index 0 &
    python -m dtg_b200 --scheme hogwild --job_name "worker" --task_index 0 &
    python -m dtg_b200 --scheme hogwild --job_name "worker" --task_index 1 &

(Hogwild/run.sh:2-4).  Without ``--job_name`` the command spawns every role of
``--cluster`` itself and waits for them.
"""
from __future__ import annotations

import os
import sys

import torch


def build_model(name: str):
    from .models.toy import ToyModel
    from .models import mlp
    if name == "toy":
        return ToyModel("zeros")
    if name == "toy_random":
        return ToyModel("truncated_normal")
    if name == "mnist_mlp":
        return mlp.mnist_mlp()
    if name == "wide_mlp":
        return mlp.wide_mlp()
    if name == "gpt2_medium":
        from .models.transformer import gpt2_medium
        return gpt2_medium()
    if name == "bert_large":
        from .models.transformer import bert_large
        return bert_large()
    if name == "resnet50":
        from .models.resnet import resnet50
        return resnet50()
    raise ValueError(name)


def main(argv=None) -> int:
    from .cluster import Server
    from .parallel.engines import SchemeConfig, WorkerEngine
    from .utils.flags import parse_cluster, parse_flags
    from .utils.launcher import launch_cluster, wait_all

    FLAGS, _ = parse_flags(argv)
    cluster = parse_cluster(FLAGS.cluster, FLAGS.base_port)

    if not FLAGS.job_name:
        args = [a for a in (argv if argv is not None else sys.argv[1:])]
        procs = launch_cluster(["-m", "dtg_b200"] + args, cluster,
                               pin_gpus=FLAGS.scheme == "multigpu_async")
        codes = wait_all(procs)
        return 0 if all(c == 0 for c in codes) else 1

    if FLAGS.job_name == "ps":
        server = Server(cluster, job_name="ps", task_index=FLAGS.task_index)
        server.join(linger=FLAGS.linger)
        return 0

    server = Server(cluster, job_name="worker", task_index=FLAGS.task_index)
    n_workers = cluster.num_tasks("worker")
    device = FLAGS.device
    if device == "auto":
        device = "cuda" if torch.cuda.is_available() and server.gpu_index is not None else "cpu"
    if device == "cuda":
        torch.cuda.set_device(server.gpu_index or 0)
    local_lr = FLAGS.local_lr
    if local_lr is not None and len(local_lr) == 1:
        local_lr = local_lr[0]
    cfg = SchemeConfig(
        scheme=FLAGS.scheme, lr=FLAGS.lr, local_lr=local_lr, update_window=FLAGS.update_window,
        elastic_alpha=FLAGS.elastic_alpha, momentum=FLAGS.momentum,
        replicas_to_aggregate=FLAGS.replicas_to_aggregate, worker_lrs=FLAGS.worker_lrs,
        last_step=FLAGS.last_step, max_iters=FLAGS.max_iters, checkpoint_dir=FLAGS.checkpoint_dir,
        save_checkpoint_secs=FLAGS.save_checkpoint_secs, ps_sharding=FLAGS.ps_sharding)
    model = build_model(FLAGS.model)
    eng = WorkerEngine(model, cfg, task_index=FLAGS.task_index, n_workers=n_workers,
                       device=device, batch_size=FLAGS.batch_size)
    if FLAGS.inject_fault:
        who, at = FLAGS.inject_fault.split("@")
        if who == "worker:%d" % FLAGS.task_index:
            orig, n = eng.train_op._fn, [0]

            def faulty(ctx):
                n[0] += 1
                if n[0] >= int(at):
                    os._exit(17)
                return orig(ctx)
            eng.train_op._fn = faulty
    print("Starting training on worker %d" % FLAGS.task_index, flush=True)
    hooks = []
    if FLAGS.profile_dir:
        from .utils.trace import ProfilerHook
        hooks.append(ProfilerHook(eng.tracer, save_steps=FLAGS.profile_steps,
                                  output_dir=os.path.join(FLAGS.profile_dir,
                                                          "worker%d" % FLAGS.task_index)))
    res = eng.run(server.target, log=lambda s: print(s, flush=True), log_every=FLAGS.log_every,
                  pace_secs=FLAGS.pace_secs, extra_hooks=hooks)
    print("Done", FLAGS.task_index, res, flush=True)
    print("Session from worker %d closed cleanly" % FLAGS.task_index, flush=True)
    server.stop()
    return 0


if __name__ == "__main__":
    sys.exit(main())
