"""Closed-form toy oracles: dL/da_i = dL/db_i = c_i - 100, so
plain SGD gives e_{k+1} = (1 - 2*lr) e_k with e = c - 100."""
import pytest
import torch

from dtg_b200.models.toy import ToyModel
from dtg_b200.parallel.engines import SchemeConfig, WorkerEngine


def run_rounds(dtg, scheme, n, **kw):
    eng = WorkerEngine(ToyModel("zeros"), SchemeConfig(scheme=scheme, **kw), 0, 1)
    sess = eng.session()
    out = []
    for _ in range(n):
        sess.run(eng.train_op)
        out.append(eng.output()[0].item())
    step = eng.store.read_step()
    sess.close()
    return out, step, eng


def test_non_distributed_sgd_closed_form(dtg):
    a = dtg.Variable(dtg.zeros([2])); b = dtg.Variable(dtg.zeros([2]))
    c = a + b
    loss = dtg.reduce_mean(dtg.square(c - dtg.constant(100., shape=[2])))
    opt = dtg.train.GradientDescentOptimizer(1e-4).minimize(loss)
    sess = dtg.train.Supervisor().prepare_or_wait_for_session()
    for i in range(1000):
        sess.run(opt)
        if i == 9:
            assert sess.run(c)[0].item() == pytest.approx(0.19982007, rel=1e-5)
    assert sess.run(c)[0].item() == pytest.approx(18.128567, rel=1e-4)


def test_hogwild_single_worker_equals_sgd(dtg):
    out, step, _ = run_rounds(dtg, "hogwild", 10)
    assert out[-1] == pytest.approx(0.19982007, rel=1e-5)
    assert step == 10


def test_adag_true_local_steps(dtg):
    out, step, _ = run_rounds(dtg, "adag", 3)
    assert out == pytest.approx([0.019996, 0.039988, 0.059976], rel=1e-4)
    assert step == 3


def test_downpour_local_and_global_adagrad(dtg):
    out, step, eng = run_rounds(dtg, "downpour", 3)
    assert out == pytest.approx([1.9999986e-4, 3.414211e-4, 4.5689102e-4], rel=1e-4)
    names = sorted(v.op.name for v in dtg.global_variables())
    assert names == ["g/Variable", "g/Variable/Adagrad", "g/Variable_1",
                     "g/Variable_1/Adagrad", "global_step"]
    assert int(eng.local_step_var.read_value()) == 9             # T per round


def test_downpour_easy_chief_local_lr_is_zero(dtg):
    """DOWNPOUR-Easy/DOWNPOUR.py:53: local lr = lr*task_index -> 0 on the chief, so
    the three window grads are identical: G = 3*(-100), Adagrad on ps."""
    out, _, _ = run_rounds(dtg, "downpour_easy", 1)
    g = -300.0
    expect = 2 * (1e-4 * (-g) / (0.1 + g * g) ** 0.5)
    assert out[0] == pytest.approx(expect, rel=1e-5)


def test_ssgd_single_replica_lockstep(dtg):
    out, step, _ = run_rounds(dtg, "ssgd", 10, num_tokens=0)
    assert step == 10
    assert out[-1] == pytest.approx(0.19982007, rel=1e-5)


def test_ssgd_per_worker_lr_effective_rate(dtg):
    """Workers (0.1, 1e-4), base 1.0: w -= sum_k lr_k g_k -> eta_eff = 0.1001 with
    identical grads, c_10 = 100*(1-(1-2*0.1001)^10) = 89.2894.  One process plays
    both replicas by pushing twice per step."""
    eng = WorkerEngine(ToyModel("zeros"), SchemeConfig(scheme="ssgd_lr", num_tokens=0), 0, 2)
    sess = eng.session()
    store, sync = eng.store, eng.store.sync_optimizer
    sync.stop_queue_runner(close_queue=False)          # drive the chief by hand
    for _ in range(10):
        params = store.pull()
        _, g = eng.model.forward_backward(params)
        for k, lr_k in enumerate(eng.cfg.worker_lrs):
            scale = (lr_k / eng.cfg.lr) * 2
            for n in store.param_names:
                v = store.vars[n]
                v._host().acc_apply_grad([sync._acc_names[store.param_names.index(n)]],
                                         [g[n] * scale], store.read_step())
        assert sync._aggregate_once(timeout=1.0)
    assert eng.output()[0].item() == pytest.approx(89.2894, rel=1e-4)
    sess.close()


def test_sdag_window_mean_then_sync(dtg):
    """T=5 local SGD steps from w: grads e_t = (1-2lr)^t e_0; ps applies their mean."""
    out, step, _ = run_rounds(dtg, "sdag", 1, num_tokens=0)
    lr, e = 1e-4, -100.0
    gs = []
    for _ in range(5):
        gs.append(e)
        e = e * (1 - 2 * lr)
    expect = 2 * (-lr * sum(gs) / 5)
    assert out[0] == pytest.approx(expect, rel=1e-5)
    assert step == 1


def test_adagrad_slot_initial_value_and_update(dtg):
    v = dtg.Variable([1.0, 2.0])
    opt = dtg.train.AdagradOptimizer(0.5)
    step = opt.apply_gradients([(torch.tensor([3.0, 4.0]), v)])
    dtg.global_variables_initializer().run()
    slot = opt.get_slot(v, "Adagrad")
    assert slot.op.name == "Variable/Adagrad" and slot.device == v.device
    assert slot.read_value().tolist() == pytest.approx([0.1, 0.1])
    step.run()
    acc = torch.tensor([0.1 + 9.0, 0.1 + 16.0])
    assert slot.read_value().tolist() == pytest.approx(acc.tolist())
    expect = torch.tensor([1.0, 2.0]) - 0.5 * torch.tensor([3.0, 4.0]) / acc.sqrt()
    assert v.read_value().tolist() == pytest.approx(expect.tolist())


def test_compute_then_apply_same_grads_to_local_and_global(dtg):
    """compute_gradients/apply_gradients split (DOWNPOUR/DOWNPOUR.py:68-75,96-98)."""
    a = dtg.Variable([0.0], collections=[dtg.GraphKeys.LOCAL_VARIABLES])
    ga = dtg.Variable([10.0], name="g/a")
    gs = dtg.train.get_or_create_global_step()
    loss = dtg.reduce_sum(dtg.square(a - 1.0))
    opt = dtg.train.GradientDescentOptimizer(0.25)
    gv = opt.compute_gradients(loss, var_list=[a])
    local = opt.apply_gradients(gv)
    glob = opt.apply_gradients([(gv[0][0], ga)], global_step=gs)
    dtg.Session().run([dtg.global_variables_initializer(), dtg.local_variables_initializer()])
    dtg.Session().run([glob, local])          # one run: both see grad at a=0 -> -2
    assert a.read_value().item() == pytest.approx(0.5)
    assert ga.read_value().item() == pytest.approx(10.5)
    assert int(gs.read_value()) == 1


# ---- the reference's TODO schemes (README.md:40-42): DynSGD, AEASGD, AEAMSGD ----------------
def test_dynsgd_single_worker_is_sequential_sgd(dtg):
    """No concurrent pushes -> staleness 1 -> a round equals T sequential SGD steps."""
    out, step, eng = run_rounds(dtg, "dynsgd", 3)
    r = 1 - 2e-4
    assert out == pytest.approx([100 * (1 - r ** (4 * (k + 1))) for k in range(3)], rel=1e-4)
    assert step == 3 and eng.last_staleness == 1


def test_dynsgd_scales_by_staleness(dtg):
    """Two rounds of worker A land between worker B's pull and push: B's window is applied with
    weight 1/3 (staleness-aware instead of SyncReplicas' drop)."""
    a = WorkerEngine(ToyModel("zeros"), SchemeConfig(scheme="dynsgd"), 0, 2)
    b = WorkerEngine(ToyModel("zeros"), SchemeConfig(scheme="dynsgd"), 1, 2, store=a.store)
    sa, sb = a.session(), b.session()
    orig = b._window

    def window_with_interference(rule, lr):
        G = orig(rule, lr)                       # B computed on the parameters it pulled (c = 0)
        sa.run(a.train_op); sa.run(a.train_op)   # meanwhile A pushes twice
        return G
    b._window = window_with_interference
    before = None
    sb.run(b.train_op)
    assert b.last_staleness == 3
    r = 1 - 2e-4
    c_after_a = 100 * (1 - r ** 8)               # two staleness-1 rounds of A
    G_b = -100 * (1 - r ** 4) / (2e-4)           # B's window sum per variable, from c = 0
    expect = c_after_a - 2 * 1e-4 * G_b / 3
    assert a.output()[0].item() == pytest.approx(expect, rel=1e-4)
    sa.close(); sb.close()


def _elastic_oracle(rounds, alpha, tau, lr, delta=None):
    """Scalar replay of EASGD on the toy problem (a = b by symmetry; c = 2a)."""
    xa, ca, v = 0.0, 0.0, 0.0
    out = []
    for _ in range(rounds):
        e = alpha * (xa - ca)
        xa -= e
        ca += e
        for _ in range(tau):
            if delta is None:
                xa -= lr * (2 * xa - 100)
            else:
                g = 2 * (xa + delta * v) - 100
                v = delta * v - lr * g
                xa += v
        out.append(2 * ca)
    return out


def test_aeasgd_elastic_exchange(dtg):
    out, step, eng = run_rounds(dtg, "aeasgd", 4, lr=1e-3)
    assert eng.cfg.elastic_alpha == pytest.approx(0.9) and eng.cfg.update_window == 4
    assert out == pytest.approx(_elastic_oracle(4, 0.9, 4, 1e-3), rel=1e-4, abs=1e-6)
    assert out[0] == 0.0 and out[1] > 0          # the centre moves only through the elastic term
    assert step == 4


def test_aeamsgd_nesterov_local_steps(dtg):
    out, _, _ = run_rounds(dtg, "aeamsgd", 4, lr=1e-3, elastic_alpha=0.5, momentum=0.9,
                           update_window=3)
    assert out == pytest.approx(_elastic_oracle(4, 0.5, 3, 1e-3, delta=0.9), rel=1e-4, abs=1e-6)
