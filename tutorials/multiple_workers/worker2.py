"""Worker 2: only its local variables are initialised here; g/a already lives on the ps.
(Basics-Tutorial/Multiple-Workers/Local-then-Global-Variables-Worker2.ipynb)"""
from _common import build

dtg, server, a, ga, global_update = build(1)
sess = dtg.Session(server.target)
sess.run(dtg.local_variables_initializer())                         # init_local only
print("W2 g/a as left by W1:     ", sess.run(ga).tolist())
sess.run(global_update)
print("W2 g/a after my update:   ", sess.run(ga).tolist())
ga._host().mark_done("worker:1")
server.stop()
