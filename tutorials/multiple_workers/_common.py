import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import dtg_b200 as dtg


def build(task_index):
    cluster_spec = dtg.ClusterSpec({"ps": ["localhost:2222"],
                                    "worker": ["localhost:2223", "localhost:2224"]})
    server = dtg.Server(cluster_spec, job_name="worker", task_index=task_index)
    with dtg.device("/job:worker/task:%d" % task_index):
        a = dtg.Variable([0.0], name="a", collections=[dtg.GraphKeys.LOCAL_VARIABLES])
        loss = dtg.reduce_sum(dtg.abs(a - dtg.constant(100.0, shape=[1])))
        opt = dtg.train.GradientDescentOptimizer(0.1)
        grads = opt.compute_gradients(loss, var_list=[a])
    with dtg.device("/job:ps/task:0"):
        ga = dtg.Variable(dtg.truncated_normal([1]), name="g/a")     # same name => same ps variable
    global_update = dtg.assign_sub(ga, 0.1 * grads[0][0])
    return dtg, server, a, ga, global_update
