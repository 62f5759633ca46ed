"""Hogwild!: every worker runs SGD against the shared ps variables with no locks and no
barrier (reference: Hogwild/Hogwild.py).  `--gpu` pins worker k to GPU k with the ps
hidden from GPUs, i.e. Multiple-GPUs-Single-Machine/dist_mult_gpu_sing_mach.{py,sh}."""
import os
import time
from common import cluster, dtg, flags, serve_if_ps, toy_graph

FLAGS = flags(lambda p: p.add_argument("--gpu", action="store_true"))
spec = cluster(FLAGS)
if not serve_if_ps(FLAGS, spec):
    chief = FLAGS.task_index == 0
    config = dtg.ConfigProto(allow_soft_placement=True, gpu_options=dtg.GPUOptions(
        allow_growth=True, visible_device_list=str(FLAGS.task_index) if FLAGS.gpu else ""))
    server = dtg.Server(spec, job_name="worker", task_index=FLAGS.task_index, config=config)
    with dtg.device("/gpu:0" if FLAGS.gpu else "/cpu:0"):
        a, b, c, loss = toy_graph(init="truncated_normal")
        step = dtg.train.GradientDescentOptimizer(1e-4).minimize(loss)   # use_locking=False
    sv = dtg.train.Supervisor(logdir=os.getcwd() + "/logdir", is_chief=chief, save_model_secs=30)
    sess = sv.prepare_or_wait_for_session(server.target, config=config)
    for i in range(1000):
        if sv.should_stop():
            break
        sess.run(step)
        if i % 10 == 0:
            print(sess.run(c).tolist())
        if not FLAGS.fast:
            time.sleep(0.1)
    sv.stop()
    a._host().mark_done("worker:%d" % FLAGS.task_index)
    server.stop()
