"""Single-process baseline (reference: Non-Distributed_Setup.py): SGD lr 1e-4 on the
toy problem, 1000 iterations, print c every 10."""
import time
from common import dtg, flags, toy_graph

FLAGS = flags()
with dtg.device("/cpu:0"):
    a, b, c, loss = toy_graph(init="truncated_normal")
    step = dtg.train.GradientDescentOptimizer(1e-4).minimize(loss)
sess = dtg.train.Supervisor().prepare_or_wait_for_session()
for i in range(1000):
    sess.run(step)
    if i % 10 == 0:
        print(sess.run(c).tolist())
        if not FLAGS.fast:
            time.sleep(0.1)
