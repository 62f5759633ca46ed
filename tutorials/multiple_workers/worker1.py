"""Worker 1 = the chief: it initialises the global variable.
(Basics-Tutorial/Multiple-Workers/Local-then-Global-Variables-Worker1.ipynb)"""
import time
from _common import build

dtg, server, a, ga, global_update = build(0)
sess = dtg.Session(server.target)
sess.run([dtg.local_variables_initializer(), ga.initializer])      # init_local + init_global
print("W1 g/a after init:        ", sess.run(ga).tolist())
sess.run(global_update)
print("W1 g/a after my update:   ", sess.run(ga).tolist())
print("now run worker2.py; waiting 15 s ...")
time.sleep(15)
print("W1 g/a after W2's update: ", sess.run(ga).tolist(), "(moved although W1 did nothing)")
ga._host().mark_done("worker:0")
server.stop()
