"""Native model + scheme engines on one GPU, and 2-GPU runs through torchrun."""
import json
import os
import subprocess
import sys

import pytest
import torch

from conftest import ROOT, free_ports

pytestmark = pytest.mark.gpu


def rel_err(out, ref):
    out, ref = out.float(), ref.float()
    return ((out - ref).abs().max() / (ref.abs().max() + 1e-6)).item()


def _setup(sizes=(784, 512, 256, 10), B=512):
    import dtg_b200  # noqa: F401
    from dtg_b200.models.mlp import MLP
    from dtg_b200.models.mlp_native import NativeMLP
    torch.manual_seed(0)
    native = NativeMLP(list(sizes), B, "cuda")
    ref = MLP(list(sizes))
    params = {k: v.cuda() for k, v in ref.init_params(seed=3).items()}
    for k in params:
        if k.endswith("bias"):
            params[k] = torch.randn_like(params[k]) * 0.1
    x, y = native.synthetic_batch()
    return native, ref, params, x, y


def test_native_mlp_gradients_match_fp32_reference():
    native, ref, params, x, y = _setup()
    # reference on bf16-rounded weights/inputs so only accumulation order differs
    pr = {k: (v.bfloat16().float() if k.endswith("kernel") else v) for k, v in params.items()}
    loss_ref, g_ref = ref.forward_backward(pr, (x.bfloat16().float(), y))
    loss, g = native.forward_backward(params, (x, y))
    assert abs(loss.item() - loss_ref.item()) < 2e-2 * abs(loss_ref.item())
    for k in g_ref:
        assert rel_err(g[k], g_ref[k]) < 1e-1, k     # dZ is rounded to bf16 between layers
    # padded rows of the last layer stay exactly zero
    last = native.layout.view(native.g32, "dense_2/kernel")
    assert last[10:].abs().max() == 0


def test_native_mlp_trains():
    native, ref, params, x, y = _setup()
    native.load_params(params)
    first = native.train_step_local(x, y, 0.05)
    for _ in range(30):
        last = native.train_step_local(x, y, 0.05)
    assert last < first * 0.9


@pytest.mark.parametrize("scheme", ["hogwild", "downpour", "downpour_easy", "adag", "ssgd",
                                    "ssgd_lr", "sdag", "tower_avg", "dynsgd", "aeasgd", "aeamsgd"])
def test_gpu_worker_single_rank_all_schemes(scheme):
    from dtg_b200.models.mlp_native import NativeMLP
    from dtg_b200.parallel.engines import SchemeConfig
    from dtg_b200.parallel.fabric import Fabric
    from dtg_b200.parallel.gpu_engines import GpuWorker
    torch.manual_seed(0)
    model = NativeMLP([784, 256, 128, 10], 256, "cuda")
    cfg = SchemeConfig(scheme=scheme, lr=0.05, local_lr=0.05, update_window=2,
                       worker_lrs=[0.05] if scheme == "ssgd_lr" else None)
    w = GpuWorker(model, cfg, Fabric.single())
    x, y = model.synthetic_batch()
    losses = [float(w.step(x, y).item()) for _ in range(25)]
    assert losses[-1] < losses[0] * 0.95, losses
    assert w.global_step() == 25
    assert w.status() == 0
    state = w.named_state()
    assert "dense/kernel" in state and state["dense_2/kernel"].shape == (10, 128)
    assert int(state["global_step"]) == 25


def test_fused_push_equals_separate_push():
    """wgrad epilogue red.add(-lr*dW) into the ps shard == materialise grads then push."""
    from dtg_b200.models.mlp_native import NativeMLP
    from dtg_b200.parallel.engines import SchemeConfig
    from dtg_b200.parallel.fabric import Fabric
    from dtg_b200.parallel.gpu_engines import GpuWorker
    outs = []
    for fused, async_push in ((True, False), (False, False), (False, True)):
        torch.manual_seed(0)
        model = NativeMLP([784, 256, 128, 10], 256, "cuda")
        w = GpuWorker(model, SchemeConfig(scheme="hogwild", lr=0.05), Fabric.single(),
                      fused_push=fused, async_push=async_push)
        assert w.async_push == async_push
        g = torch.Generator().manual_seed(5)
        x, y = model.synthetic_batch(generator=g)
        for _ in range(3):
            w.step(x, y)
        torch.cuda.synchronize()
        outs.append(w.named_state())
    # the layer-by-layer side-stream push (single-port ps) lands the same sums; its pull for
    # step t+1 may miss step t's push (Hogwild staleness), so it is compared after ONE step below
    for k in outs[0]:
        assert rel_err(outs[0][k], outs[1][k]) < 1e-5, k
    one = []
    for async_push in (False, True):
        torch.manual_seed(0)
        model = NativeMLP([784, 256, 128, 10], 256, "cuda")
        w = GpuWorker(model, SchemeConfig(scheme="hogwild", lr=0.05), Fabric.single(),
                      fused_push=True, async_push=async_push)
        x, y = model.synthetic_batch(generator=torch.Generator().manual_seed(5))
        w.step(x, y)
        torch.cuda.synchronize()
        one.append(w.named_state())
    for k in one[0]:
        assert rel_err(one[0][k], one[1][k]) < 1e-5, k


def test_sync_world1_equals_local_sgd():
    from dtg_b200.parallel.fabric import Fabric
    from dtg_b200.parallel.peer_store import PeerSync
    torch.manual_seed(0)
    n = 100_003
    sy = PeerSync(Fabric.single(), n, lr=0.1, adagrad=True)
    p0 = torch.randn(n, device="cuda")
    sy.load_master(p0)
    g = torch.randn(n, device="cuda")
    sy.g32.copy_(g)
    assert sy.step() == 1
    torch.cuda.synchronize()
    acc = 0.1 + g * g
    ref = p0 - 0.1 * g / acc.sqrt()
    assert int(sy.status.item()) == 0
    assert rel_err(sy.p32, ref) < 1e-6 and rel_err(sy.p16, ref.bfloat16()) < 1e-2


@pytest.mark.multigpu
@pytest.mark.parametrize("scheme", ["hogwild", "adag", "ssgd", "ssgd_lr", "dynsgd", "aeasgd"])
def test_two_gpu_worker(scheme):
    port = free_ports(1)[0]
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
           "--master-addr", "127.0.0.1", "--master-port", str(port),
           os.path.join(ROOT, "tests", "gpu_multi_worker.py"), "--scheme", scheme]
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True,
                       timeout=300, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-3000:]
    line = [l for l in r.stdout.splitlines() if l.startswith("RESULT ")][-1]
    res = json.loads(line[7:])
    assert res["ok"], res


@pytest.mark.multigpu
@pytest.mark.parametrize("scheme", ["hogwild", "downpour"])
def test_two_gpu_striped_ps(scheme):
    """ps in one virtual range striped over both GPUs' HBM (CUDA VMM, fds over unix sockets)."""
    port = free_ports(1)[0]
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
           "--master-addr", "127.0.0.1", "--master-port", str(port),
           os.path.join(ROOT, "tests", "gpu_multi_worker.py"), "--scheme", scheme, "--striped"]
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True,
                       timeout=300, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-3000:]
    line = [l for l in r.stdout.splitlines() if l.startswith("RESULT ")][-1]
    res = json.loads(line[7:])
    assert res["ok"], res


@pytest.mark.multigpu
def test_two_gpu_async_push_single_port_ps():
    """ps on GPU 0 only, gradients stored locally and pushed layer by layer on a side stream:
    every push is counted and the loss goes down."""
    port = free_ports(1)[0]
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
           "--master-addr", "127.0.0.1", "--master-port", str(port),
           os.path.join(ROOT, "tests", "gpu_multi_worker.py"), "--scheme", "hogwild", "--gpu0",
           "--async_push"]
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True,
                       timeout=300, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-3000:]
    line = [l for l in r.stdout.splitlines() if l.startswith("RESULT ")][-1]
    res = json.loads(line[7:])
    assert res["ok"], res


@pytest.mark.multigpu
def test_two_gpu_nvls_sync_step():
    """SSGD with the gradient / parameter buffers bound to an NVSwitch multicast object:
    multimem.ld_reduce + multimem.st must leave bit-identical replicas (forced on at 2 ranks)."""
    port = free_ports(1)[0]
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
           "--master-addr", "127.0.0.1", "--master-port", str(port),
           os.path.join(ROOT, "tests", "gpu_multi_worker.py"), "--scheme", "ssgd"]
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True,
                       timeout=300, cwd=ROOT, env=dict(os.environ, DTG_NVLS="1"))
    assert r.returncode == 0, r.stdout[-3000:]
    line = [l for l in r.stdout.splitlines() if l.startswith("RESULT ")][-1]
    res = json.loads(line[7:])
    assert res["ok"], res
    assert all(rk.get("nvls") for rk in res["ranks"]), res


@pytest.mark.multigpu
def test_two_gpu_heartbeat_watchdog_names_the_quiet_rank():
    """parallel/heartbeat.py across ranks: rank 0 reads every rank's heartbeat word through peer
    memory; when the last rank stops stepping, the watchdog names exactly that rank and the
    asynchronous scheme keeps running without it (tools/heartbeat_check.py)."""
    port = free_ports(1)[0]
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
           "--master-addr", "127.0.0.1", "--master-port", str(port),
           os.path.join(ROOT, "tools", "heartbeat_check.py")]
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True,
                       timeout=300, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-3000:]
    rows = [json.loads(l[10:]) for l in r.stdout.splitlines() if l.startswith("HEARTBEAT ")]
    assert [x["phase"] for x in rows] == ["all", "one quiet", "done"]
    assert rows[0]["stale"] == [] and rows[1]["stale"] == [1] and rows[2]["done"] == [0, 1]
    assert rows[1]["beats"][0][0] > rows[0]["beats"][0][0]        # rank 0 went on training


def test_bench_json_contract():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--steps", "3",
                        "--warmup", "3", "--batch", "1024"], stdout=subprocess.PIPE,
                       stderr=subprocess.STDOUT, text=True, timeout=300, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-3000:]
    out = json.loads(r.stdout.strip().splitlines()[-1])
    for k in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step",
              "higher_is_better", "scaling", "vs_baseline", "dtype", "data", "config", "clocks",
              "e2e", "gpu_launches"):
        assert k in out, k
    assert out["value"] > 0 and out["gpu_launches"] > 0 and out["e2e"]["value"] > 0
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference"],
                       stdout=subprocess.PIPE, text=True, timeout=120, cwd=ROOT)
    ref = json.loads(r.stdout.strip().splitlines()[-1])
    assert ref["impl"] == "reference" and "unavailable" in ref


@pytest.mark.parametrize("scheme", ["hogwild", "ssgd"])
def test_gpu_session_hooks_checkpoint_and_resume(tmp_path, scheme):
    """L4 on device: StopAtStep on the ps-resident global_step, chief checkpoints keyed by
    TF variable name, restart resumes from the latest checkpoint."""
    import dtg_b200 as dtg
    from dtg_b200.models.mlp_native import NativeMLP
    from dtg_b200.parallel.engines import SchemeConfig
    from dtg_b200.parallel.fabric import Fabric
    from dtg_b200.parallel.gpu_engines import GpuWorker
    from dtg_b200.parallel.gpu_session import GpuTrainingSession
    d = str(tmp_path / "logdir")

    def make():
        torch.manual_seed(0)
        model = NativeMLP([784, 128, 10], 128, "cuda")
        return model, GpuWorker(model, SchemeConfig(scheme=scheme, lr=0.05), Fabric.single())

    model, w = make()
    x, y = model.synthetic_batch()
    n = 0
    with GpuTrainingSession(w, last_step=6, checkpoint_dir=d, save_checkpoint_steps=2) as sess:
        while not sess.should_stop():
            sess.step(x, y); n += 1
    assert n == 6 and w.global_step() == 6
    ck = dtg.train.latest_checkpoint(d)
    assert ck.endswith("model.ckpt-6")
    data = dtg.train.load_checkpoint(ck)
    assert sorted(data) == ["dense/bias", "dense/kernel", "dense_1/bias", "dense_1/kernel",
                            "global_step"]
    assert data["dense_1/kernel"].shape == (10, 128) and int(data["global_step"]) == 6
    final = w.named_state()
    # "restart": fresh worker (fresh random init), same directory
    model2, w2 = make()
    sess2 = GpuTrainingSession(w2, last_step=8, checkpoint_dir=d, save_checkpoint_steps=100)
    assert sess2.restored_from == ck and w2.global_step() == 6
    st2 = w2.named_state()
    for k in final:
        assert rel_err(st2[k], final[k]) < 1e-6, k
    m = 0
    while not sess2.should_stop():
        sess2.step(x, y); m += 1
    sess2.close()
    assert m == 2 and w2.global_step() == 8


def test_elastic_and_dynsgd_kernels_match_reference():
    """Data-plane kernels of the reference's TODO schemes vs plain torch fp32."""
    import dtg_b200  # noqa: F401
    from dtg_b200.ops import elastic
    torch.manual_seed(4)
    n = 100003 * 4
    centre = torch.randn(n, device="cuda"); x = torch.randn(n, device="cuda")
    v = torch.randn(n, device="cuda") * 0.1; g = torch.randn(n, device="cuda")
    x16 = torch.empty(n, device="cuda", dtype=torch.bfloat16)
    c0, x0, v0 = centre.clone(), x.clone(), v.clone()
    elastic.easgd_elastic(centre, x, x16, 0.3, velocity=v, delta=0.9)
    e = 0.3 * (x0 - c0)
    assert rel_err(x, x0 - e) < 1e-6 and rel_err(centre, c0 + e) < 1e-6
    assert rel_err(x16, (x0 - e) + 0.9 * v0) < 1e-2
    x1 = x.clone()
    elastic.momentum_step(x, v, g, x16, 0.01, 0.9)
    vn = 0.9 * v0 - 0.01 * g
    assert rel_err(v, vn) < 1e-6 and rel_err(x, x1 + vn) < 1e-6
    assert rel_err(x16, x1 + vn + 0.9 * vn) < 1e-2
    w = torch.zeros(n, device="cuda")
    step = torch.tensor([7], device="cuda"); seen = torch.tensor([5], device="cuda")
    st = torch.zeros(1, dtype=torch.int64, device="cuda")
    elastic.push_axpy_dyn(w, g, -0.5, step, seen, st)
    assert int(st.item()) == 3 and rel_err(w, -0.5 / 3 * g) < 1e-6


def test_input_pipeline_feeder_and_reader():
    """Batches arrive in order and intact while later ones are being staged; results come back
    one step late without a device-wide synchronize."""
    import dtg_b200  # noqa: F401
    from dtg_b200.utils.input_pipeline import DeviceFeeder, ResultReader
    g = torch.Generator().manual_seed(0)
    hx = [torch.randn(512, 784, generator=g).pin_memory() for _ in range(5)]
    hy = [torch.randint(0, 10, (512,), generator=g).pin_memory() for _ in range(5)]
    feeder, reader = DeviceFeeder("cuda"), ResultReader("cuda", depth=2)
    feeder.put(hx[0], hy[0])
    sums = []
    for i in range(5):
        if i + 1 < 5:
            feeder.put(hx[i + 1], hy[i + 1])
        x, y = feeder.get()
        out = (x.double().sum() + y.sum()).float()      # "the step"
        feeder.release()
        reader.push(out)
        if reader.pending() > 1:
            sums.append(float(reader.pop()[0]))
    while reader.pending():
        sums.append(float(reader.pop()[0]))
    ref = [float(hx[i].double().sum() + hy[i].sum()) for i in range(5)]
    assert sums == pytest.approx(ref, rel=1e-5)
    with pytest.raises(RuntimeError):
        feeder.get()


@pytest.mark.parametrize("scheme", ["hogwild", "adag", "ssgd"])
def test_tracer_times_phases_on_device(scheme, tmp_path):
    """utils/trace.py on CUDA events + NVTX: phases land in the summary with device times, the
    prefetched pull on its own lane, and tracing does not change the result."""
    from dtg_b200.models.mlp_native import NativeMLP
    from dtg_b200.parallel.engines import SchemeConfig
    from dtg_b200.parallel.fabric import Fabric
    from dtg_b200.parallel.gpu_engines import GpuWorker
    from dtg_b200.utils.trace import PhaseTracer
    losses = {}
    for traced in (False, True):
        torch.manual_seed(0)
        model = NativeMLP([784, 256, 128, 10], 256, "cuda")
        tr = PhaseTracer(enabled=traced, device="cuda")
        w = GpuWorker(model, SchemeConfig(scheme=scheme, lr=0.05, local_lr=0.05, update_window=2),
                      Fabric.single(), tracer=tr)
        x, y = model.synthetic_batch()
        for _ in range(6):
            loss = w.step(x, y)
        losses[traced] = float(loss.item())
        if not traced:
            assert tr.collect() == []
            continue
        s = tr.summary(skip_steps=1)
        want = {"hogwild": {"forward", "pull(prefetch)", "backward+push"},
                "adag": {"local_steps", "push", "pull"},
                "ssgd": {"fwd+bwd"}}[scheme]
        assert want <= set(s), (want, set(s))
        for name in want:
            assert s[name]["calls"] == 5 and 0 < s[name]["ms_mean"] < 50
        if scheme == "hogwild":
            assert s["pull(prefetch)"]["lane"] == "pull"
        doc = json.load(open(tr.write_chrome_trace(str(tmp_path / "timeline.json"))))
        assert sum(1 for e in doc["traceEvents"] if e["ph"] == "X") >= 5 * len(want)
    if scheme != "hogwild":          # Hogwild's prefetched pull races its own push by design
        assert losses[True] == losses[False]


def test_device_heartbeat_and_watchdog_in_session():
    """parallel/heartbeat.py, device half: the heartbeat word advances once per step on the
    training stream, the watchdog sees it, and ``end()`` marks the rank done."""
    from dtg_b200.models.mlp_native import NativeMLP
    from dtg_b200.parallel.engines import SchemeConfig
    from dtg_b200.parallel.fabric import Fabric
    from dtg_b200.parallel.gpu_engines import GpuWorker
    from dtg_b200.parallel.gpu_session import GpuTrainingSession
    from dtg_b200.parallel.heartbeat import Watchdog
    torch.manual_seed(0)
    model = NativeMLP([784, 128, 10], 128, "cuda")
    w = GpuWorker(model, SchemeConfig(scheme="adag", lr=0.05, update_window=2), Fabric.single())
    x, y = model.synthetic_batch()
    with GpuTrainingSession(w, last_step=5, watchdog_secs=30.0) as sess:
        hb = sess.heartbeat
        assert hb.read() == [(0, 0)]
        while not sess.should_stop():
            sess.step(x, y)
        torch.cuda.synchronize()
        assert hb.read() == [(5, 0)]
        assert sess.watchdog.poll() == [] and sess.watchdog.alive() == [0]
        # a rank that stops beating goes stale after the time-out
        t = [0.0]
        wd = Watchdog(hb.read, timeout_secs=1.0, clock=lambda: t[0])
        assert wd.poll() == []
        t[0] = 5.0
        assert wd.poll() == [0]
        hb.beat()
        torch.cuda.synchronize()
        assert wd.poll() == []
    torch.cuda.synchronize()
    assert hb.read() == [(6, 1)]
