"""The lazy graph layer the example scripts are written against (framework.py): math nodes,
gradients, assignment ops, grouping, name scopes, nested device scopes, seeded initialisers."""
import pytest
import torch


@pytest.fixture
def sess(dtg):
    srv = dtg.Server.create_local_server()
    s = dtg.Session(srv.target)
    yield s
    s.close()
    srv.stop()


def test_math_nodes_and_gradients(dtg, sess):
    b = dtg.Variable([1.0, -2.0, 3.0], name="b")
    y = dtg.reduce_sum(dtg.square(b)) + dtg.reduce_mean(dtg.abs(b))
    s = dtg.stack([b, b * 2.0])
    sess.run(dtg.global_variables_initializer())
    assert float(sess.run(y)) == pytest.approx(14.0 + 2.0)
    assert sess.run(s).tolist() == [[1.0, -2.0, 3.0], [2.0, -4.0, 6.0]]
    assert sess.run(dtg.reduce_sum(s, axis=0)).tolist() == [3.0, -6.0, 9.0]
    assert sess.run(dtg.reduce_mean(s, axis=0)).tolist() == [1.5, -3.0, 4.5]
    # d/db (sum b^2 + mean |b|) = 2b + sign(b)/3
    opt = dtg.train.GradientDescentOptimizer(0.1)
    (g, v), = opt.compute_gradients(y, var_list=[b])
    assert v is b
    assert sess.run(g).tolist() == pytest.approx([2 + 1 / 3, -4 - 1 / 3, 6 + 1 / 3])
    sess.run(opt.apply_gradients([(g, v)]))
    assert sess.run(b).tolist() == pytest.approx([1 - 0.1 * (2 + 1 / 3), -2 + 0.1 * (4 + 1 / 3),
                                                  3 - 0.1 * (6 + 1 / 3)])


def test_assign_ops_and_group(dtg, sess):
    b = dtg.Variable([1.0, -2.0, 3.0], name="b")
    c = dtg.Variable(dtg.zeros([3]), name="c")
    sess.run(dtg.global_variables_initializer())
    assert sess.run(dtg.assign_add(b, dtg.constant([1.0, 1.0, 1.0]))).tolist() == [2.0, -1.0, 4.0]
    assert sess.run(dtg.assign_sub(b, dtg.constant(1.0, shape=[3]))).tolist() == [1.0, -2.0, 3.0]
    assert sess.run(dtg.assign(b, [0.0, 0.0, 0.0])).tolist() == [0.0, 0.0, 0.0]
    both = dtg.group(dtg.assign_add(b, dtg.constant(1.0, shape=[3])),
                     dtg.assign_add(c, dtg.constant(2.0, shape=[3])))
    sess.run(both)                                           # one fetch runs both updates
    assert [t.tolist() for t in sess.run([b, c])] == [[1.0] * 3, [2.0] * 3]
    assert sess.run(dtg.no_op()) is None


def test_name_scopes_and_unique_names(dtg):
    a = dtg.Variable(dtg.zeros([1]), name="a")
    a2 = dtg.Variable(dtg.zeros([1]), name="a")
    with dtg.name_scope("outer"):
        c = dtg.Variable(dtg.zeros([1]), name="c")
        with dtg.name_scope("inner"):
            d = dtg.Variable(dtg.zeros([1]), name="c")
        e = dtg.Variable(dtg.zeros([1]), name="c")
    assert [v.op.name for v in (a, a2, c, d, e)] == ["a", "a_1", "outer/c", "outer/inner/c", "outer/c_1"]
    g = dtg.get_variable("gv", shape=[2], initializer=dtg.zeros([2]))
    assert g.op.name == "gv" and g in dtg.global_variables()


def test_nested_device_scopes_inner_wins(dtg):
    """Basics-Tutorial/Local-then-Global-Variables.ipynb:104-134: the innermost scope places a
    variable; ops that update a variable sit with the variable."""
    with dtg.device("/job:worker/task:0"):
        with dtg.device("/job:ps/task:0"):
            e = dtg.Variable([0.0], name="e")
        f = dtg.Variable([0.0], name="f")
        upd = dtg.assign_add(e, dtg.constant([1.0]))
    assert e.device == "/job:ps/replica:0/task:0/device:CPU:0"
    assert f.device == "/job:worker/replica:0/task:0/device:CPU:0"
    assert upd.device == e.device


def test_seeded_truncated_normal(dtg, sess):
    """Non-Distributed_Setup.py:13-14: tf.truncated_normal -- sigma 1, redrawn beyond 2 sigma."""
    dtg.set_random_seed(5)
    a = dtg.Variable(dtg.truncated_normal([4000]), name="a")
    sess.run(a.initializer)
    v = sess.run(a)
    assert float(v.abs().max()) <= 2.0 and abs(float(v.std()) - 0.8796) < 0.03
    first = v.clone()
    dtg.reset()
    srv = dtg.Server.create_local_server()
    dtg.set_random_seed(5)
    a2 = dtg.Variable(dtg.truncated_normal([4000]), name="a")
    with dtg.Session(srv.target) as s2:
        s2.run(a2.initializer)
        assert torch.equal(s2.run(a2), first)                  # same seed, same draw
    srv.stop()


def test_log_device_placement(dtg, capsys):
    """ConfigProto(log_device_placement=True) (the reference sets it False,
    DOWNPOUR/DOWNPOUR.py:21): each op is reported once, with the device it was placed on."""
    srv = dtg.Server.create_local_server()
    with dtg.device("/job:local/task:0"):
        b = dtg.Variable([1.0, 2.0], name="b")
        y = dtg.reduce_sum(dtg.square(b))
    with dtg.Session(srv.target, config=dtg.ConfigProto(log_device_placement=True)) as sess:
        sess.run(b.initializer)
        capsys.readouterr()
        sess.run(y)
        first = capsys.readouterr().out
        sess.run(y)
        second = capsys.readouterr().out
    lines = [l for l in first.splitlines() if l]
    assert lines and all("/job:local/replica:0/task:0/device:CPU:0" in l for l in lines)
    assert any(l.startswith(y.op.name + ": (") for l in lines)
    assert second == ""                                      # once per op and session
    with dtg.Session(srv.target) as quiet:
        quiet.run(y)
        assert capsys.readouterr().out == ""
    srv.stop()
