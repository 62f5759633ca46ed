"""Synchronous SGD with SyncReplicas-style N-of-M aggregation.
(reference: Synchronous-SGD/ssgd.py; `--per_worker_lr`:
Synchronous-SGD-different-learning-rates/ssgd.py; `--window T`: SDAG, i.e. T local steps
whose mean gradient is aggregated synchronously, SDAG/README.md:3.)"""
import time
from common import cluster, dtg, flags, serve_if_ps, toy_graph


def extra(p):
    p.add_argument("--per_worker_lr", action="store_true")
    p.add_argument("--window", type=int, default=0)


FLAGS = flags(extra)
spec = cluster(FLAGS)
if not serve_if_ps(FLAGS, spec):
    k, chief, N = FLAGS.task_index, FLAGS.task_index == 0, FLAGS.workers
    server = dtg.Server(spec, job_name="worker", task_index=k)
    with dtg.device(dtg.replica_device_setter(ps_tasks=1, worker_device="/job:worker/task:%d/cpu:0" % k)):
        a, b, c, loss = toy_graph()
        global_step = dtg.Variable(0, dtype=dtg.framework.torch.int64, trainable=False,
                                   name="global_step")
        base_lr = 1.0 if FLAGS.per_worker_lr else 1e-4
        sync = dtg.train.SyncReplicasOptimizer(dtg.train.GradientDescentOptimizer(base_lr),
                                               replicas_to_aggregate=N, total_num_replicas=N)
        gv = sync.compute_gradients(loss)
        if FLAGS.per_worker_lr:               # scale on the producer; the mean becomes sum_k lr_k g_k
            lr_k = 0.1 if k == 0 else 1e-4
            gv = [((lr_k / base_lr) * N * g, v) for g, v in gv]
        step = sync.apply_gradients(gv, global_step=global_step)
    if FLAGS.window:
        # SDAG through the scheme engine: T local SGD steps, window mean, sync aggregate
        from dtg_b200.models.toy import ToyModel
        from dtg_b200.parallel.engines import SchemeConfig, WorkerEngine
        dtg.reset_default_graph()
        eng = WorkerEngine(ToyModel("zeros"), SchemeConfig(scheme="sdag", update_window=FLAGS.window,
                                                           last_step=10), k, N)
        print(eng.run(server.target, log=print, log_every=1))
    else:
        hooks = [sync.make_session_run_hook(chief), dtg.train.StopAtStepHook(last_step=10)]
        sess = dtg.train.MonitoredTrainingSession(master=server.target, is_chief=chief, hooks=hooks,
                                                  stop_grace_period_secs=10)
        print("Starting training on worker %d" % k)
        while not sess.should_stop():
            out = sess.run([step, c, global_step])
            if out is not None:
                print(out[1].tolist(), "step: ", int(out[2]), "worker: ", k)
            if not FLAGS.fast:
                time.sleep(2 if chief else 1)
        print("Done", k)
        sess.close()
        global_step._host().mark_done("worker:%d" % k)
        print("Session from worker %d closed cleanly" % k)
    server.stop()
