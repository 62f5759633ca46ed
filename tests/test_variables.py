"""L2 / L5': naming, collections, placement, local<->global helpers and the
notebooks' recorded traces."""
import pytest
import torch


def test_tf_auto_naming(dtg):
    a = dtg.Variable(dtg.zeros([2]))
    b = dtg.Variable(dtg.zeros([2]))
    c = dtg.Variable(dtg.zeros([2]))
    gs = dtg.Variable(0, dtype=torch.int32, trainable=False, name="global_step")
    assert [a.op.name, b.op.name, c.op.name] == ["Variable", "Variable_1", "Variable_2"]
    assert a.name == "Variable:0"
    assert gs.op.name == "global_step"
    assert dtg.train.get_global_step() is gs


def test_collections(dtg):
    a = dtg.Variable(dtg.zeros([2]), collections=[dtg.GraphKeys.LOCAL_VARIABLES])
    ls = dtg.Variable(0, dtype=torch.int32, trainable=False, name="local_step",
                      collections=["local_non_trainable"])
    g = dtg.Variable(dtg.zeros([2]))
    assert dtg.local_variables() == [a]
    assert dtg.global_variables() == [g]
    assert dtg.get_collection("local_non_trainable") == [ls]
    assert a in dtg.trainable_variables() and ls not in dtg.trainable_variables()


def test_replica_device_setter_round_robin(dtg, ports):
    p = ports(4)
    spec = {"ps": ["localhost:%d" % p[0], "localhost:%d" % p[1], "localhost:%d" % p[2]],
            "worker": ["localhost:%d" % p[3]]}
    wk = dtg.Server(spec, "worker", 0)
    try:
        with dtg.device(dtg.replica_device_setter(ps_tasks=3,
                                                  worker_device="/job:worker/task:0")):
            vs = [dtg.Variable(dtg.zeros([1])) for _ in range(5)]
            c = vs[0] + vs[1]
        assert [v.device for v in vs] == [
            "/job:ps/replica:0/task:%d/device:CPU:0" % (i % 3) for i in range(5)]
        assert c.device == "/job:worker/replica:0/task:0/device:CPU:0"
        # inner explicit scope overrides the setter (DOWNPOUR.py:202)
        with dtg.device(dtg.replica_device_setter(ps_tasks=3)):
            with dtg.device("/job:ps/task:0"):
                pinned = [dtg.Variable(dtg.zeros([1])) for _ in range(3)]
        assert all(v.device.startswith("/job:ps/replica:0/task:0") for v in pinned)
    finally:
        wk.stop()


def test_size_balanced_sharding(dtg):
    from dtg_b200.placement import flat_shards, shard_variables
    sizes = [("a", 100), ("b", 10), ("c", 10), ("d", 80), ("e", 5)]
    rr = shard_variables(sizes, 2, "round_robin")
    assert rr == {"a": 0, "b": 1, "c": 0, "d": 1, "e": 0}
    bal = shard_variables(sizes, 2, "balanced")
    load = [0, 0]
    for n, s in sizes:
        load[bal[n]] += s
    assert abs(load[0] - load[1]) <= 10
    sh = flat_shards(1003, 4, align_elems=4)
    assert sh[0][0] == 0 and sh[-1][1] == 1003
    assert all(b % 4 == 0 for b, _ in sh)
    assert all(sh[i][1] == sh[i + 1][0] for i in range(3))


def test_global_name_rule_and_helpers(dtg):
    from dtg_b200.parallel import local_global as lg
    with dtg.device(None):
        a = dtg.Variable(dtg.zeros([2]), collections=[dtg.GraphKeys.LOCAL_VARIABLES])
        b = dtg.Variable(dtg.zeros([2]), collections=[dtg.GraphKeys.LOCAL_VARIABLES])
    l2g, g2l = lg.create_global_variables()
    assert sorted(v.op.name for v in l2g.values()) == ["g/Variable", "g/Variable_1"]
    assert lg.get_global_variable_by_name("g/Variable:0") is l2g[a]
    dtg.global_variables_initializer().run()
    dtg.local_variables_initializer().run()
    a.assign([1.0, 2.0])
    lg.assign_local_to_global(l2g).run()
    assert l2g[a].read_value().tolist() == [1.0, 2.0]
    l2g[b].assign([5.0, 6.0])
    lg.assign_global_to_local(g2l).run()
    assert b.read_value().tolist() == [5.0, 6.0]


def test_downpour_slot_collection_hack(dtg):
    """DOWNPOUR/DOWNPOUR.py:80-84,216-233: the local Adagrad's slots are born in
    GLOBAL_VARIABLES and get moved to LOCAL_VARIABLES by hand."""
    from dtg_b200.parallel import local_global as lg
    a = dtg.Variable(dtg.zeros([2]), collections=[dtg.GraphKeys.LOCAL_VARIABLES])
    loss = dtg.reduce_mean(dtg.square(a - 100.0))
    lopt = dtg.train.AdagradOptimizer(1e-4)
    gv = lopt.compute_gradients(loss, var_list=dtg.local_variables())
    lopt.apply_gradients(gv)
    assert [v.op.name for v in dtg.global_variables()] == ["Variable/Adagrad"]
    moved = lg.add_global_variables_to_local_collection()
    lg.clear_global_collection()
    assert dtg.global_variables() == []
    assert [v.op.name for v in moved] == ["Variable/Adagrad"]
    assert moved[0] in dtg.local_variables()
    l2g, _ = lg.create_global_variables(moved)
    assert [v.op.name for v in l2g.values()] == ["g/Variable"]


def test_notebook_trace_local_then_global(dtg, ports):
    """Basics-Tutorial/Local-then-Global-Variables.ipynb:103-134,164-171,222-291:
    same gradient updates the local copy or the ps copy; update ops are
    colocated with the variable they write."""
    p0, p1 = ports(2)
    spec = {"ps": ["localhost:%d" % p0], "worker": ["localhost:%d" % p1]}
    ps = dtg.Server(spec, "ps", 0)
    wk = dtg.Server(spec, "worker", 0)
    try:
        with dtg.device("/job:worker/task:0"):
            a = dtg.Variable([0.0], name="a", collections=[dtg.GraphKeys.LOCAL_VARIABLES])
            loss = dtg.reduce_sum(dtg.abs(a - 100.0))
            opt = dtg.train.GradientDescentOptimizer(0.1)
            grads = opt.compute_gradients(loss, var_list=[a])
            local_update = opt.apply_gradients(grads)
        with dtg.device("/job:ps/task:0"):
            ga = dtg.Variable(torch.tensor([-1.26032162]), name="g/a")
        with dtg.device("/job:worker/task:0"):
            global_update = dtg.assign_sub(ga, 0.1 * grads[0][0])
        assert a.device == "/job:worker/replica:0/task:0/device:CPU:0"
        assert ga.device == "/job:ps/replica:0/task:0/device:CPU:0"
        assert global_update.device == ga.device          # colocated with the variable
        assert ga.initializer.device == ga.device
        sess = dtg.Session(wk.target)
        sess.run([dtg.global_variables_initializer(), dtg.local_variables_initializer()])
        sess.run(local_update)
        assert sess.run(a).item() == pytest.approx(0.1)                   # :253
        assert sess.run(ga).item() == pytest.approx(-1.26032162)          # unchanged :222
        sess.run(global_update)
        assert sess.run(ga).item() == pytest.approx(-1.16032159, abs=1e-6)  # :291
    finally:
        wk.stop()
        ps.stop()


def test_notebook_trace_two_workers_share_ps_state(dtg, ports):
    """Multiple-Workers notebooks: W1 -1.17584229 -> -1.07584226, W2 continues to
    -0.97584224 and W1 re-reads it (Worker1.ipynb:224,286,313; Worker2.ipynb:217,286).
    Two worker 'kernels' are emulated by two RPC clients of the same ps."""
    (p0,) = ports(1)
    ps = dtg.Server({"ps": ["localhost:%d" % p0]}, "ps", 0)
    try:
        from dtg_b200.parallel.rpc import HostClient
        w1, w2 = HostClient("localhost", p0), HostClient("localhost", p0)
        w1.assign("g/a", torch.tensor([-1.17584229]))           # chief runs init_global
        assert w2.uninitialized(["g/a"]) == []                   # W2 only inits locals
        g = torch.tensor([-1.0])                                 # d|a-100|/da
        w1.apply_sgd(["g/a"], [g], 0.1)
        assert w1.read("g/a").item() == pytest.approx(-1.07584226, abs=1e-6)
        assert w2.read("g/a").item() == pytest.approx(-1.07584226, abs=1e-6)
        w2.apply_sgd(["g/a"], [g], 0.1)
        assert w1.read("g/a").item() == pytest.approx(-0.97584224, abs=1e-6)
        w1.close(); w2.close()
    finally:
        ps.stop()


def test_control_dependencies_and_fetch_order(dtg):
    v = dtg.Variable([0.0])
    v.initializer.run()
    inc = dtg.assign_add(v, [1.0])
    with dtg.control_dependencies([inc]):
        read_after = dtg.tensor(lambda: v.read_value().clone())
    sess = dtg.Session()
    out = sess.run([read_after, v])
    assert out[0].item() == 1.0 and out[1].item() == 1.0
    # an op listed twice in one run still executes once
    sess.run([inc, inc])
    assert v.read_value().item() == 2.0


def test_uninitialized_read_raises(dtg):
    v = dtg.Variable([1.0])
    with pytest.raises(KeyError):
        v.read_value()
    assert dtg.report_uninitialized_variables().eval() == ["Variable"]
