"""DOWNPOUR / DOWNPOUR-Easy / ADAG with the low-level API: local replicas in
LOCAL_VARIABLES, global copies 'g/<name>' on the ps, a communication window of T local
steps, push of the window (sum + Adagrad, or mean + SGD), then pull.
(reference: DOWNPOUR/DOWNPOUR.py, DOWNPOUR-Easy/DOWNPOUR.py, ADAG/ADAG.py)

Unlike the TF graphs, each of the T gradient evaluations here sees the parameters
left by the previous local step (true local steps)."""
import time
from common import cluster, dtg, flags, serve_if_ps, toy_graph
from dtg_b200.parallel import local_global as lg

FLAGS = flags(lambda p: p.add_argument("--variant", default="downpour",
                                       choices=["downpour", "easy", "adag"]))
spec = cluster(FLAGS)
if not serve_if_ps(FLAGS, spec):
    k, chief = FLAGS.task_index, FLAGS.task_index == 0
    server = dtg.Server(spec, job_name="worker", task_index=k)
    lr, T = 1e-4, 3
    with dtg.device("/job:worker/task:%d" % k):
        a, b, c, loss = toy_graph(collections=[dtg.GraphKeys.LOCAL_VARIABLES])
        local_step = dtg.Variable(0, dtype=dtg.framework.torch.int64, trainable=False,
                                  name="local_step", collections=["local_non_trainable"])
        lopt = {"downpour": dtg.train.AdagradOptimizer(lr),
                "easy": dtg.train.GradientDescentOptimizer(lr * k),
                "adag": dtg.train.GradientDescentOptimizer(lr)}[FLAGS.variant]
        window, prev = [], None
        for t in range(T):
            with dtg.control_dependencies([prev] if prev is not None else []):
                gv = lopt.compute_gradients(loss, var_list=[a, b])
            window.append([g for g, _ in gv])
            with dtg.control_dependencies([g for g, _ in gv]):
                prev = lopt.apply_gradients(gv, global_step=local_step)
        with dtg.control_dependencies([prev]):
            reduce = dtg.reduce_mean if FLAGS.variant == "adag" else dtg.reduce_sum
            G = reduce(dtg.stack(window), axis=0)
        slots = lg.add_global_variables_to_local_collection()     # local Adagrad slots
        lg.clear_global_collection()
    with dtg.device(dtg.replica_device_setter(ps_tasks=spec.num_tasks("ps"),
                                              worker_device="/job:worker/task:%d" % k)):
        global_step = dtg.Variable(0, dtype=dtg.framework.torch.int64, trainable=False,
                                   name="global_step")
        gopt = dtg.train.GradientDescentOptimizer(lr) if FLAGS.variant == "adag" \
            else dtg.train.AdagradOptimizer(lr)
        l2g, g2l = lg.create_global_variables(slots)
        push = gopt.apply_gradients([(G[i], l2g[v]) for i, v in enumerate([a, b])],
                                    global_step=global_step)
        with dtg.control_dependencies([push]):
            pull = lg.assign_global_to_local(g2l)
        first_pull = lg.assign_global_to_local(g2l)
        seed = lg.assign_local_to_global(l2g)
        init = dtg.global_variables_initializer()
        init_local = dtg.variables_initializer(
            dtg.local_variables() + dtg.get_collection("local_non_trainable"))
    hooks = [dtg.train.StopAtStepHook(last_step=40 if FLAGS.variant == "adag" else 60)]
    sess = dtg.train.MonitoredTrainingSession(
        master=server.target, is_chief=chief, hooks=hooks,
        scaffold=dtg.train.Scaffold(init_op=init, local_init_op=[init_local]),
        save_checkpoint_secs=1, checkpoint_dir="logdir")
    if chief:
        sess.run(seed)
        time.sleep(0 if FLAGS.fast else 10)
    print("Starting training on worker %d" % k)
    sess.run(first_pull)
    while not sess.should_stop():
        _, _, r, gs, ls = sess.run([push, pull, c, global_step, local_step])
        print(r.tolist(), "global step: %d" % gs, "worker: %d" % k, "local step: %d" % ls)
        if not FLAGS.fast:
            time.sleep(1)
    print("Done", k)
    sess.close()
    global_step._host().mark_done("worker:%d" % k)
    print("Session from worker %d closed cleanly" % k)
    server.stop()
