"""Shared bits of the example programs: the two reference flags and the localhost
cluster (1 ps + N workers on ports 2222..., Hogwild/Hogwild.py:20-23)."""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import dtg_b200 as dtg  # noqa: E402


def flags(extra=None):
    p = argparse.ArgumentParser()
    p.add_argument("--job_name", type=str, default="", help="One of 'ps', 'worker'")
    p.add_argument("--task_index", type=int, default=0, help="Index of task within the job")
    p.add_argument("--workers", type=int, default=2)
    p.add_argument("--base_port", type=int, default=2222)
    p.add_argument("--fast", action="store_true", help="skip the 'so we can observe training' sleeps")
    if extra:
        extra(p)
    return p.parse_known_args()[0]


def cluster(FLAGS, ps=1):
    return dtg.ClusterSpec("ps=%d,worker=%d" % (ps, FLAGS.workers), base_port=FLAGS.base_port)


def serve_if_ps(FLAGS, spec):
    """ps role: host the variables until every worker reports done."""
    if FLAGS.job_name == "ps":
        dtg.Server(spec, job_name="ps", task_index=FLAGS.task_index).join()
        return True
    return False


def toy_graph(collections=None, init="zeros"):
    """a, b in R^2; c = a + b; loss = mean((c - 100)^2)."""
    mk = (lambda: dtg.zeros([2])) if init == "zeros" else (lambda: dtg.truncated_normal([2]))
    a = dtg.Variable(mk(), collections=collections)
    b = dtg.Variable(mk(), collections=collections)
    c = a + b
    loss = dtg.reduce_mean(dtg.square(c - dtg.constant(100.0, shape=[2])))
    return a, b, c, loss
