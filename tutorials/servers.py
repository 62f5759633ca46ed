"""Servers: what a task is, how to talk to it.  (Basics-Tutorial/Servers.ipynb)"""
# %%
import os, sys, socket
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import dtg_b200 as dtg

# %% A cluster with a single task, started in this process
cluster = dtg.ClusterSpec({"worker": ["localhost:2222"]})
server = dtg.Server(cluster, job_name="worker", task_index=0)
print(server.target)                      # grpc://localhost:2222

# %% The task is really listening on that port (the notebook checks this with lsof)
with socket.create_connection(("127.0.0.1", 2222), timeout=2):
    print("port 2222 is accepting connections")

# %% The server's view of the cluster
print(server.server_def)

# %% A session is opened against a target; fetches are evaluated through that task
with dtg.Session(server.target) as sess:
    v = dtg.Variable([1.0, 2.0], name="v")
    sess.run(v.initializer)
    print(sess.run(v))
server.stop()

# %% A throw-away single-process cluster on an ephemeral port
local = dtg.Server.create_local_server()
print(local.target)
with dtg.Session(local.target) as sess:
    for d in sess.list_devices():
        print(d)                          # /job:local/replica:0/task:0/device:CPU:0 ...
local.stop()
