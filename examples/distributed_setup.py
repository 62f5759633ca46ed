"""1 ps + 1 worker with MonitoredTrainingSession (reference: Distributed-Setup/dist_setup.py).
`--supervisor` uses Supervisor(logdir, save_model_secs=30) instead (dist_setup_sup.py)."""
import os
import time
from common import cluster, dtg, flags, serve_if_ps, toy_graph

FLAGS = flags(lambda p: p.add_argument("--supervisor", action="store_true"))
FLAGS.workers = 1
spec = cluster(FLAGS)
if not serve_if_ps(FLAGS, spec):
    chief = FLAGS.task_index == 0
    server = dtg.Server(spec, job_name="worker", task_index=FLAGS.task_index)
    a, b, c, loss = toy_graph(init="truncated_normal")       # shared vars land on the ps
    step = dtg.train.GradientDescentOptimizer(1e-4).minimize(loss)
    if FLAGS.supervisor:
        sv = dtg.train.Supervisor(logdir=os.getcwd() + "/logdir", is_chief=chief, save_model_secs=30)
        sess, stop = sv.prepare_or_wait_for_session(server.target), sv.should_stop
    else:
        sess = dtg.train.MonitoredTrainingSession(master=server.target, is_chief=chief)
        stop = sess.should_stop
    for i in range(1000):
        if stop():
            break
        sess.run(step)
        if i % 10 == 0:
            print(sess.run(c).tolist())
            if not FLAGS.fast:
                time.sleep(0.1)
    if FLAGS.supervisor:
        sv.stop()
    a._host().mark_done("worker:%d" % FLAGS.task_index)
    server.stop()
