"""Local, then global, variables.  (Basics-Tutorial/Local-then-Global-Variables.ipynb)

Start `parameter_server.py` first, then run this as the worker."""
# %%
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import dtg_b200 as dtg

cluster_spec = dtg.ClusterSpec({"ps": ["localhost:2222"], "worker": ["localhost:2223"]})
server = dtg.Server(cluster_spec, job_name="worker", task_index=0)

# %% A local variable on the worker, its loss and two ways of applying the same gradient
with dtg.device("/job:worker/task:0"):
    a = dtg.Variable([0.0], name="a", collections=[dtg.GraphKeys.LOCAL_VARIABLES])
    target = dtg.constant(100.0, shape=[1])
    loss = dtg.reduce_sum(dtg.abs(a - target))
    opt = dtg.train.GradientDescentOptimizer(0.1)
    grads = opt.compute_gradients(loss, var_list=[a])
    local_update = opt.apply_gradients(grads)

# %% The global copy lives on the parameter server, under the name 'g/a'
with dtg.device("/job:ps/task:0"):
    ga = dtg.Variable(dtg.truncated_normal([1]), name="g/a")

# %% An op that writes a variable is placed WITH the variable, whatever scope builds it
with dtg.device("/job:worker/task:0"):
    global_update = dtg.assign_sub(ga, 0.1 * grads[0][0])
    init_global = ga.initializer
    init_local = dtg.local_variables_initializer()
for name, node in [("a", a), ("loss", loss), ("local_update", local_update), ("g/a", ga),
                   ("global_update", global_update), ("init_global", init_global)]:
    print("%-14s %s" % (name, node.device))

# %%
sess = dtg.Session(server.target)
sess.run([init_local, init_global])
print("a   =", sess.run(a).tolist(), " g/a =", sess.run(ga).tolist())
sess.run(local_update)                       # only the worker's copy moves
print("a   =", sess.run(a).tolist(), " g/a =", sess.run(ga).tolist())
sess.run(global_update)                      # only the ps copy moves (by +0.1)
print("a   =", sess.run(a).tolist(), " g/a =", sess.run(ga).tolist())
ga._host().mark_done("worker:0")
server.stop()
