"""The schemes the reference lists as TODO (README.md:40-42), on the engine API:
  --variant aeasgd   asynchronous elastic averaging SGD: every tau local steps the worker and
                     the centre variable on the ps pull each other, e = alpha (x - centre)
  --variant aeamsgd  the same with Nesterov momentum on the worker
  --variant dynsgd   staleness-aware async SGD: a window pushed with weight 1/staleness
Launch like every other example: ./run.sh elastic_averaging.py --variant aeasgd --fast"""
import time
from common import cluster, dtg, flags, serve_if_ps
from dtg_b200.models.toy import ToyModel
from dtg_b200.parallel.engines import SchemeConfig, WorkerEngine

FLAGS = flags(lambda p: p.add_argument("--variant", default="aeasgd",
                                       choices=["aeasgd", "aeamsgd", "dynsgd"]))
spec = cluster(FLAGS)
if not serve_if_ps(FLAGS, spec):
    k = FLAGS.task_index
    server = dtg.Server(spec, job_name="worker", task_index=k)
    cfg = SchemeConfig(scheme=FLAGS.variant, lr=1e-3, update_window=4, last_step=40)
    eng = WorkerEngine(ToyModel("zeros"), cfg, task_index=k, n_workers=FLAGS.workers)
    print("Starting training on worker %d" % k)
    with eng.session(master=server.target) as sess:
        while not sess.should_stop():
            _, gs = sess.run([eng.train_op, eng.global_step])
            extra = " staleness: %d" % eng.last_staleness if FLAGS.variant == "dynsgd" else ""
            print(eng.output().tolist(), "global step: %d" % gs, "worker: %d" % k + extra)
            if not FLAGS.fast:
                time.sleep(0.5)
    eng.global_step._host().mark_done("worker:%d" % k)
    print("Session from worker %d closed cleanly" % k)
    server.stop()
