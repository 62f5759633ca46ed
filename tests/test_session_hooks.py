"""L4: hook protocol, StopAtStep, checkpoints by name, restore-on-start,
SyncReplicas token/stale-drop state machine."""
import os
import time

import pytest
import torch


def _toy_graph(dtg):
    a = dtg.Variable(dtg.zeros([2])); b = dtg.Variable(dtg.zeros([2]))
    gs = dtg.Variable(0, dtype=torch.int64, trainable=False, name="global_step")
    loss = dtg.reduce_mean(dtg.square(a + b - 100.0))
    opt = dtg.train.GradientDescentOptimizer(1e-4).minimize(loss, global_step=gs)
    return a, b, gs, loss, opt


def test_hook_call_order(dtg):
    a, b, gs, loss, opt = _toy_graph(dtg)
    calls = []

    class Rec(dtg.train.SessionRunHook):
        def begin(self): calls.append("begin")
        def after_create_session(self, s, c): calls.append("after_create_session")
        def before_run(self, rc):
            calls.append("before_run")
            return dtg.train.SessionRunArgs(gs)
        def after_run(self, rc, rv): calls.append("after_run:%d" % int(rv.results))
        def end(self, s): calls.append("end")

    with dtg.train.MonitoredTrainingSession(hooks=[Rec()]) as sess:
        sess.run(opt)
        sess.run(opt)
    assert calls == ["begin", "after_create_session", "before_run", "after_run:1",
                     "before_run", "after_run:2", "end"]


def test_stop_at_step_last_step_and_num_steps(dtg):
    a, b, gs, loss, opt = _toy_graph(dtg)
    sess = dtg.train.MonitoredTrainingSession(hooks=[dtg.train.StopAtStepHook(last_step=5)])
    n = 0
    while not sess.should_stop():
        sess.run(opt); n += 1
    assert n == 5 and int(gs.read_value()) == 5
    sess.close()
    sess = dtg.train.MonitoredTrainingSession(hooks=[dtg.train.StopAtStepHook(num_steps=3)],
                                              scaffold=dtg.train.Scaffold(init_op=dtg.no_op()))
    n = 0
    while not sess.should_stop():
        sess.run(opt); n += 1
    assert n == 3 and int(gs.read_value()) == 8
    with pytest.raises(ValueError):
        dtg.train.StopAtStepHook()


def test_checkpoint_roundtrip_by_name(dtg, tmp_path):
    a, b, gs, loss, opt = _toy_graph(dtg)
    dtg.Session().run(dtg.global_variables_initializer())
    a.assign([1.5, -2.5]); gs.assign(7)
    saver = dtg.train.Saver()
    prefix = saver.save(None, str(tmp_path / "model.ckpt"), global_step=gs)
    assert prefix.endswith("model.ckpt-7")
    assert dtg.train.latest_checkpoint(str(tmp_path)) == prefix
    state = open(tmp_path / "checkpoint").read()
    assert 'model_checkpoint_path: "model.ckpt-7"' in state
    data = dtg.train.load_checkpoint(prefix)
    assert sorted(data) == ["Variable", "Variable_1", "global_step"]
    assert data["Variable"].tolist() == [1.5, -2.5] and int(data["global_step"]) == 7
    a.assign([0.0, 0.0]); gs.assign(0)
    assert sorted(saver.restore(None, prefix)) == ["Variable", "Variable_1", "global_step"]
    assert a.read_value().tolist() == [1.5, -2.5] and int(gs.read_value()) == 7


def test_saver_max_to_keep(dtg, tmp_path):
    a, b, gs, loss, opt = _toy_graph(dtg)
    dtg.Session().run(dtg.global_variables_initializer())
    saver = dtg.train.Saver(max_to_keep=2)
    for s in (1, 2, 3):
        saver.save(None, str(tmp_path / "model.ckpt"), global_step=s)
    files = sorted(f for f in os.listdir(tmp_path) if f.endswith(".dtg"))
    assert files == ["model.ckpt-2.dtg", "model.ckpt-3.dtg"]
    st = dtg.train.get_checkpoint_state(str(tmp_path))
    assert len(st["all_model_checkpoint_paths"]) == 2


def test_monitored_session_saves_and_restores_on_start(dtg, tmp_path):
    """DOWNPOUR/DOWNPOUR.py:121-127: checkpoint_dir + save_checkpoint_secs; a new
    chief session resumes from the latest checkpoint (optimizer slots included)."""
    from dtg_b200.models.toy import ToyModel
    from dtg_b200.parallel.engines import SchemeConfig, WorkerEngine
    d = str(tmp_path / "logdir")
    cfg = SchemeConfig(scheme="downpour", checkpoint_dir=d, save_checkpoint_secs=0.01)
    eng = WorkerEngine(ToyModel("zeros"), cfg, 0, 1)
    sess = eng.session()
    for _ in range(3):
        sess.run(eng.train_op); time.sleep(0.02)
    c3 = eng.output().clone()
    sess.close()
    ck = dtg.train.latest_checkpoint(d)
    assert ck is not None and ck.endswith("model.ckpt-3")
    names = sorted(dtg.train.load_checkpoint(ck))
    assert names == ["g/Variable", "g/Variable/Adagrad", "g/Variable_1",
                     "g/Variable_1/Adagrad", "global_step"]
    # "restart the chief": fresh runtime + graph, same directory
    dtg.reset()
    eng2 = WorkerEngine(ToyModel("zeros"), cfg, 0, 1)
    sess2 = eng2.session()
    assert eng2.store.read_step() == 3
    assert torch.allclose(eng2.output(), c3)
    acc = eng2.store._opts[("adagrad", 1e-4)]._slots["g/Variable/Adagrad"].read_value()
    assert (acc > 0.1).all()
    sess2.close()


def test_supervisor_saves_on_timer_and_restores(dtg, tmp_path):
    a, b, gs, loss, opt = _toy_graph(dtg)
    d = str(tmp_path / "logdir")
    sv = dtg.train.Supervisor(logdir=d, is_chief=True, save_model_secs=0.05)
    sess = sv.prepare_or_wait_for_session()
    for _ in range(4):
        sess.run(opt)
    time.sleep(0.2)
    assert not sv.should_stop()
    sv.stop()
    assert sv.should_stop()
    assert dtg.train.latest_checkpoint(d).endswith("model.ckpt-4")
    dtg.reset()
    a, b, gs, loss, opt = _toy_graph(dtg)
    sess = dtg.train.Supervisor(logdir=d, save_model_secs=0).prepare_or_wait_for_session()
    assert int(gs.read_value()) == 4
    assert sess.run(a)[0].item() == pytest.approx(0.5 * 0.07997600, rel=1e-4)


def test_non_chief_waits_until_chief_initialised(dtg, ports):
    """WorkerSessionCreator: poll ready_op, then local init."""
    import threading
    (p0, p1, p2) = ports(3)
    spec = {"ps": ["localhost:%d" % p0], "worker": ["localhost:%d" % p1, "localhost:%d" % p2]}
    ps = dtg.Server(spec, "ps", 0)
    wk = dtg.Server(spec, "worker", 1)
    try:
        with dtg.device("/job:ps/task:0"):
            w = dtg.Variable([3.0], name="w")
        loc = dtg.Variable([0.0], name="loc", collections=[dtg.GraphKeys.LOCAL_VARIABLES])
        result = {}

        def worker():
            s = dtg.train.MonitoredTrainingSession(master=wk.target, is_chief=False,
                                                   recovery_wait_secs=0.02)
            result["t"] = time.time()
            result["w"] = s.run(w).item()
            s.close()
        t = threading.Thread(target=worker, daemon=True); t.start()
        time.sleep(0.3)
        assert "t" not in result                     # still waiting for the chief
        t_init = time.time()
        ps.host.assign("w", torch.tensor([3.0]))     # the chief's init_op lands on the ps
        t.join(timeout=5)
        assert result["w"] == 3.0 and result["t"] >= t_init
        assert loc.read_value().item() == 0.0        # local_init_op ran
    finally:
        wk.stop(); ps.stop()


def test_sync_replicas_stale_drop_and_tokens(dtg):
    """[TF] ConditionalAccumulator + token queue on the variable's host."""
    from dtg_b200.parallel.rpc import VariableHost
    h = VariableHost()
    h.acc_create("acc", [2])
    assert h.acc_apply_grad(["acc"], [torch.tensor([2.0, 4.0])], 0) == [True]
    assert h.acc_take_grad(["acc"], 2, timeout=0.05) is None          # needs N=2
    assert h.acc_apply_grad(["acc"], [torch.tensor([4.0, 8.0])], 0) == [True]
    (mean,) = h.acc_take_grad(["acc"], 2, timeout=1.0)
    assert mean.tolist() == [3.0, 6.0]                                 # mean, not sum
    assert h.acc_info("acc")["global_step"] == 1
    assert h.acc_apply_grad(["acc"], [torch.tensor([1.0, 1.0])], 0) == [False]   # stale
    assert h.acc_info("acc")["dropped"] == 1 and h.acc_info("acc")["count"] == 0
    assert h.acc_apply_grad(["acc"], [torch.tensor([1.0, 1.0])], 1) == [True]
    h.queue_enqueue_many("q", [5, 5])
    assert h.queue_dequeue("q", 0.1) == 5 and h.queue_size("q") == 1
    assert h.queue_dequeue("q", 0.1) == 5
    assert h.queue_dequeue("q", 0.05) is None                          # blocks, times out
    h.queue_close("q")
    assert h.queue_dequeue("q", 0.05) is None


def test_acc_take_grad_is_all_or_nothing(dtg):
    """A take over several accumulators that times out on a later one must not have drained
    the earlier ones (their sum reset + global_step bump would drop the next pushes as stale
    and starve the token queue)."""
    from dtg_b200.parallel.rpc import VariableHost
    h = VariableHost()
    h.acc_create("a", [1]); h.acc_create("b", [1])
    for _ in range(2):
        h.acc_apply_grad(["a"], [torch.tensor([1.0])], 0)
    h.acc_apply_grad(["b"], [torch.tensor([5.0])], 0)                 # b has 1 of 2
    assert h.acc_take_grad(["a", "b"], 2, timeout=0.05) is None
    assert h.acc_info("a") == {"count": 2, "global_step": 0, "dropped": 0}   # untouched
    assert h.acc_apply_grad(["b"], [torch.tensor([7.0])], 0) == [True]       # not stale
    ga, gb = h.acc_take_grad(["a", "b"], 2, timeout=1.0)
    assert ga.tolist() == [1.0] and gb.tolist() == [6.0]
    assert h.acc_info("a")["global_step"] == 1 and h.acc_info("b")["global_step"] == 1


def test_sync_replicas_n_of_m_with_backup_worker(dtg):
    """replicas_to_aggregate=2 of total 3: the step proceeds with the first two
    gradients; the third (stale) is dropped; 3 tokens are released."""
    a = dtg.Variable(dtg.zeros([2]))
    gs = dtg.Variable(0, dtype=torch.int64, trainable=False, name="global_step")
    opt = dtg.train.SyncReplicasOptimizer(dtg.train.GradientDescentOptimizer(1.0),
                                          replicas_to_aggregate=2, total_num_replicas=3)
    feed = {"g": torch.zeros(2)}
    train = opt.apply_gradients([(dtg.tensor(lambda: feed["g"]), a)], global_step=gs)
    dtg.Session().run([dtg.global_variables_initializer(), dtg.local_variables_initializer(),
                       opt.chief_init_op])
    host = a._host()
    host.acc_apply_grad(opt._acc_names, [torch.tensor([1.0, 1.0])], 0)
    host.acc_apply_grad(opt._acc_names, [torch.tensor([3.0, 3.0])], 0)
    assert opt._aggregate_once(timeout=1.0)
    assert a.read_value().tolist() == [-2.0, -2.0] and int(gs.read_value()) == 1
    assert host.queue_size(opt.TOKEN_QUEUE) == 3
    assert host.acc_apply_grad(opt._acc_names, [torch.tensor([9.0, 9.0])], 0) == [False]
    with pytest.raises(ValueError):
        dtg.train.SyncReplicasOptimizer(dtg.train.GradientDescentOptimizer(1.0), 2, 2) \
            .get_init_tokens_op()


def test_logging_nan_stepcounter_finalops_feedfn_hooks(dtg, tmp_path):
    a, b, gs, loss, opt = _toy_graph(dtg)
    lines = []
    log = dtg.train.LoggingTensorHook({"loss": loss}, every_n_iter=2, output=lines.append)
    sc = dtg.train.StepCounterHook(every_n_steps=2, output=lines.append, batch_size=8)
    fin = dtg.train.FinalOpsHook({"a": a})
    with dtg.train.MonitoredTrainingSession(hooks=[log, sc, fin]) as sess:
        for _ in range(5):
            sess.run(opt)
    assert len(log.logged) == 3 and "loss" in lines[0]
    assert len(sc.rates) >= 1 and any("samples/sec" in l for l in lines)
    assert fin.final_ops_values["a"][0].item() > 0
    # NaN
    bad = dtg.tensor(lambda: torch.tensor(float("nan")))
    with pytest.raises(dtg.train.NanLossDuringTrainingError):
        with dtg.train.MonitoredTrainingSession(
                hooks=[dtg.train.NanTensorHook(bad)],
                scaffold=dtg.train.Scaffold(init_op=dtg.no_op())) as sess:
            sess.run(opt)
    soft = dtg.train.NanTensorHook(bad, fail_on_nan_loss=False)
    sess = dtg.train.MonitoredTrainingSession(hooks=[soft],
                                              scaffold=dtg.train.Scaffold(init_op=dtg.no_op()))
    sess.run(opt)
    assert sess.should_stop()
    sess.close()
    # FeedFn overrides a variable's value for the run
    ff = dtg.train.FeedFnHook(lambda: {a: torch.tensor([50.0, 50.0])})
    with dtg.train.MonitoredTrainingSession(hooks=[ff],
                                            scaffold=dtg.train.Scaffold(init_op=dtg.no_op())) as s:
        assert s.run(a).tolist() == [50.0, 50.0]


def test_global_step_waiter_hook(dtg):
    import threading
    a, b, gs, loss, opt = _toy_graph(dtg)
    dtg.Session().run(dtg.global_variables_initializer())
    hook = dtg.train.GlobalStepWaiterHook(wait_until_step=2, poll_secs=0.01)
    sess = dtg.train.MonitoredTrainingSession(hooks=[hook],
                                              scaffold=dtg.train.Scaffold(init_op=dtg.no_op()))
    threading.Timer(0.15, lambda: gs.assign(2)).start()
    t0 = time.time()
    sess.run(opt)
    assert time.time() - t0 >= 0.1 and int(gs.read_value()) == 3
    sess.close()


def test_session_creators(dtg, tmp_path):
    """README.md:77 "Session Creator -- Factory for creating a session?": TF's
    ``MonitoredSession(session_creator=ChiefSessionCreator(...))`` form; the chief creator
    restores the latest checkpoint, the worker creator waits for an initialised model."""
    a, b, gs, loss, opt = _toy_graph(dtg)
    d = str(tmp_path / "ck")
    sc = dtg.train.Scaffold()
    creator = dtg.train.ChiefSessionCreator(scaffold=sc, checkpoint_dir=d)
    hook = dtg.train.CheckpointSaverHook(d, save_steps=1, scaffold=sc)
    with dtg.train.MonitoredSession(session_creator=creator, hooks=[hook]) as sess:
        for _ in range(3):
            sess.run(opt)
        c3 = sess.run(loss)
    assert not creator.restored
    assert dtg.train.latest_checkpoint(d).endswith("model.ckpt-3")
    # a worker creator finds the model ready straight away (same process, same variables)
    w = dtg.train.WorkerSessionCreator(scaffold=dtg.train.Scaffold(), max_wait_secs=5)
    ws = w.create_session()
    assert int(ws.run(gs)) == 3
    ws.close()
    # positional creator (TF's first argument) + restore by explicit file
    dtg.reset()
    a, b, gs, loss, opt = _toy_graph(dtg)
    c2 = dtg.train.ChiefSessionCreator(checkpoint_filename_with_path=dtg.train.latest_checkpoint(d))
    with dtg.train.MonitoredSession(c2) as sess:
        assert c2.restored and int(sess.run(gs)) == 3
        assert abs(float(sess.run(loss)) - float(c3)) < 1e-6
    with pytest.raises(NotImplementedError):
        dtg.train.SessionCreator().create_session()


def test_supervisor_step_counter_service(dtg, tmp_path):
    """[TF] Supervisor's standard services on the chief: besides the timed saver, a thread that
    logs global_step/sec every save_summaries_secs (Hogwild/Hogwild.py:47-49 runs with the
    defaults: 600 s / 120 s)."""
    import json
    a, b, gs, loss, opt = _toy_graph(dtg)
    d = str(tmp_path / "logdir")
    sv = dtg.train.Supervisor(logdir=d, is_chief=True, save_model_secs=0, save_summaries_secs=0.05)
    sess = sv.prepare_or_wait_for_session()
    t_end = time.time() + 0.5
    n = 0
    while time.time() < t_end:
        sess.run(opt); n += 1
        time.sleep(0.005)
    sv.stop()
    rows = [json.loads(l) for l in open(os.path.join(d, "metrics.jsonl"))]
    assert len(rows) >= 2 and sv.step_rates
    assert all(r["global_step/sec"] > 0 for r in rows[1:]) and rows[-1]["step"] <= n
    # non-chief: no services, nothing written
    dtg.reset()
    a, b, gs, loss, opt = _toy_graph(dtg)
    d2 = str(tmp_path / "other")
    sv2 = dtg.train.Supervisor(logdir=d2, is_chief=False, save_summaries_secs=0.01)
    assert sv2.start_standard_services(None) == [] and not os.path.exists(d2)


def test_saver_name_map_last_checkpoints_and_shape_check(dtg, tmp_path):
    """[TF] Saver(var_list={name_in_checkpoint: var}): read a DOWNPOUR-style checkpoint (ps copies
    named g/<var>) into differently named variables; last_checkpoints; shape mismatches refuse."""
    ga = dtg.Variable([1.0, 2.0], name="g/a")
    gs = dtg.Variable(7, dtype=torch.int64, trainable=False, name="global_step")
    d = str(tmp_path / "ck")
    sv = dtg.train.Saver(max_to_keep=2)
    with dtg.Session() as sess:
        sess.run(dtg.global_variables_initializer())
        p1 = sv.save(sess, os.path.join(d, "model.ckpt"), global_step=gs)
        sess.run(dtg.assign(ga, [3.0, 4.0]))
        sess.run(dtg.assign(gs, 8))
        p2 = sv.save(sess, os.path.join(d, "model.ckpt"), global_step=gs)
        sess.run(dtg.assign(gs, 9))
        p3 = sv.save(sess, os.path.join(d, "model.ckpt"), global_step=gs)
    assert sv.last_checkpoints == [p2, p3] and p1.endswith("-7") and not os.path.exists(p1 + ".dtg")
    dtg.reset()
    a = dtg.Variable(dtg.zeros([2]), name="a")                  # same tensor, other name
    with dtg.Session() as sess:
        sess.run(dtg.global_variables_initializer())
        assert dtg.train.Saver().restore(sess, p3) == []        # by-name restore finds nothing
        assert dtg.train.Saver({"g/a": a}).restore(sess, p3) == ["g/a"]
        assert sess.run(a).tolist() == [3.0, 4.0]
        # and back out under a third name
        p4 = dtg.train.Saver({"renamed": a}).save(sess, os.path.join(d, "other.ckpt"))
    assert sorted(dtg.train.load_checkpoint(p4)) == ["renamed"]
    dtg.reset()
    bad = dtg.Variable(dtg.zeros([3]), name="g/a")
    with dtg.Session() as sess:
        sess.run(dtg.global_variables_initializer())
        with pytest.raises(ValueError, match="shape"):
            dtg.train.Saver().restore(sess, p3)
