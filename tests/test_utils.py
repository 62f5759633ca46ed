"""Flags, launcher, runtime options and the host-side liveness / summary hooks."""
import json
import os
import sys
import time

import pytest
import torch

from conftest import ROOT, free_ports


def test_reference_flags_and_cluster_parsing(dtg):
    """Hogwild/Hogwild.py:59-75: --job_name / --task_index through parse_known_args."""
    from dtg_b200.utils.flags import parse_cluster, parse_flags
    FLAGS, unparsed = parse_flags(["--job_name", "worker", "--task_index", "1", "--unknown", "x"])
    assert FLAGS.job_name == "worker" and FLAGS.task_index == 1 and unparsed == ["--unknown", "x"]
    FLAGS, _ = parse_flags([], scheme_flags=False)
    assert FLAGS.job_name == "" and FLAGS.task_index == 0 and not hasattr(FLAGS, "scheme")
    c = parse_cluster("ps=1,worker=2")                 # the reference's ports: 2222, 2223, 2224
    assert c.as_dict() == {"ps": ["localhost:2222"], "worker": ["localhost:2223", "localhost:2224"]}
    c = parse_cluster('{"ps": ["h:1"], "worker": ["h:2", "h:3", "h:4"]}')
    assert c.num_tasks("worker") == 3 and c.task_address("worker", 2) == "h:4"
    FLAGS, _ = parse_flags(["--local_lr", "0.1,0.2", "--worker_lrs", "0.1,1e-4", "--scheme", "ssgd_lr"])
    assert FLAGS.local_lr == [0.1, 0.2] and FLAGS.worker_lrs == [0.1, 1e-4]


def test_config_proto_and_gpu_options(dtg):
    """Multiple-GPUs-Single-Machine/dist_mult_gpu_sing_mach.py:45-49."""
    cfg = dtg.ConfigProto(allow_soft_placement=True,
                          gpu_options=dtg.GPUOptions(allow_growth=True, allocator_type="BFC",
                                                     visible_device_list="1"))
    assert cfg.gpu_options.visible_device_list == "1" and cfg.allow_soft_placement
    assert dtg.ConfigProto().log_device_placement is False        # DOWNPOUR/DOWNPOUR.py:21
    sess = dtg.Session("", config=cfg) if dtg.runtime.current_server(required=False) else None
    assert sess is None or sess.config is cfg


@pytest.mark.timeout(120)
def test_launch_cluster_starts_every_role_like_run_sh(dtg, tmp_path):
    """Hogwild/run.sh:2-4 as one call; pin_gpus reproduces dist_mult_gpu_sing_mach.sh:2-7."""
    from dtg_b200.utils.launcher import launch_cluster, wait_all
    script = tmp_path / "role.py"
    script.write_text(
        "import argparse, os, json\n"
        "p = argparse.ArgumentParser(); p.add_argument('--job_name'); p.add_argument('--task_index', type=int)\n"
        "a, _ = p.parse_known_args()\n"
        "print(json.dumps({'job': a.job_name, 'task': a.task_index,\n"
        "                  'cvd': os.environ.get('CUDA_VISIBLE_DEVICES'), 'x': os.environ.get('X')}))\n")
    ports = free_ports(3)
    spec = dtg.ClusterSpec({"ps": ["localhost:%d" % ports[0]],
                            "worker": ["localhost:%d" % p for p in ports[1:]]})
    out = open(tmp_path / "out.txt", "w")
    procs = launch_cluster([str(script)], spec, pin_gpus=True, env={"X": "1"}, stdout=out)
    assert wait_all(procs, timeout=60) == [0, 0, 0]
    out.close()
    rows = [json.loads(l) for l in open(tmp_path / "out.txt") if l.startswith("{")]
    by = {(r["job"], r["task"]): r for r in rows}
    assert set(by) == {("ps", 0), ("worker", 0), ("worker", 1)}
    assert by[("ps", 0)]["cvd"] == "-1"                              # ps hidden from the GPUs
    assert by[("worker", 0)]["cvd"] == "0" and by[("worker", 1)]["cvd"] == "1"
    assert all(r["x"] == "1" for r in rows)
    # a role that hangs is killed at the deadline
    hang = tmp_path / "hang.py"
    hang.write_text("import time\ntime.sleep(60)\n")
    one = dtg.ClusterSpec({"worker": ["localhost:%d" % ports[0]]})
    t0 = time.time()
    assert wait_all(launch_cluster([str(hang)], one), timeout=1.0) == [-9]
    assert time.time() - t0 < 20


def test_host_heartbeat_hook_and_stale_tasks(dtg):
    """Host-side liveness (parallel/rpc.py): a task that stops beating shows up in stale_tasks."""
    from dtg_b200.parallel.rpc import VariableHost
    from dtg_b200.session.hooks import HeartbeatHook
    host = VariableHost()
    h0 = HeartbeatHook("worker:0", lambda: host, min_interval=0.0)
    h1 = HeartbeatHook("worker:1", lambda: host, min_interval=0.0)
    h0.before_run(None); h1.before_run(None)
    assert host.stale_tasks(10.0) == []
    time.sleep(0.15)
    h0.before_run(None)                                  # worker:1 went quiet
    assert host.stale_tasks(0.1) == ["worker:1"]
    # min_interval throttles the control-plane traffic
    slow = HeartbeatHook("worker:2", lambda: host, min_interval=60.0)
    slow.before_run(None)
    t = host.heartbeats["worker:2"]
    slow.before_run(None)
    assert host.heartbeats["worker:2"] == t


def test_summary_saver_hook_writes_scalars(dtg, tmp_path):
    a = dtg.Variable(dtg.zeros([2])); b = dtg.Variable(dtg.zeros([2]))
    gs = dtg.Variable(0, dtype=torch.int64, trainable=False, name="global_step")
    loss = dtg.reduce_mean(dtg.square(a + b - 100.0))
    opt = dtg.train.GradientDescentOptimizer(1e-4).minimize(loss, global_step=gs)
    d = str(tmp_path / "logdir")
    # scaffold.summary_op + checkpoint_dir: MonitoredTrainingSession adds the hook itself
    sc = dtg.train.Scaffold(summary_op=loss)
    with dtg.train.MonitoredTrainingSession(is_chief=True, checkpoint_dir=d, scaffold=sc,
                                            save_summaries_steps=2, save_checkpoint_secs=None,
                                            log_step_count_steps=None) as sess:
        for _ in range(6):
            sess.run(opt)
    rows = [json.loads(l) for l in open(os.path.join(d, "summaries.jsonl"))]
    assert [r["step"] for r in rows] == [1, 3, 5]          # global_step after the run it fired in
    assert rows[0]["value"] <= 10000.0 and rows[-1]["value"] < rows[0]["value"]


@pytest.mark.timeout(180)
def test_c_abi_header_and_examples_build(tmp_path):
    """include/dtg_b200.h is valid C99 and C++; the C programs under examples/c_abi build against
    libdtg_b200.so, and the one that needs no GPU runs (CRC-32C check value, schedule replay)."""
    import shutil
    import subprocess
    cc = shutil.which("gcc") or shutil.which("cc")
    if cc is None:
        pytest.skip("no C compiler")
    hdr = os.path.join(ROOT, "include", "dtg_b200.h")
    for lang, std in (("c", "-std=c99"), ("c++", "-std=c++11")):
        r = subprocess.run([cc, "-fsyntax-only", "-x", lang, std, "-Wall", "-Wextra", "-pedantic", hdr],
                           stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        assert r.returncode == 0 and not r.stdout.strip(), r.stdout
    lib = os.path.join(ROOT, "distributed-tensorflow-guide_b200", "ops", "libdtg_b200.so")
    if not os.path.exists(lib):
        pytest.skip("libdtg_b200.so not built")
    exe = str(tmp_path / "host_check")
    src = os.path.join(ROOT, "examples", "c_abi", "host_check.c")
    r = subprocess.run([cc, "-std=c99", "-Wall", "-Wextra", "-pedantic", "-I", os.path.join(ROOT, "include"),
                        src, "-L", os.path.dirname(lib), "-ldtg_b200",
                        "-Wl,-rpath," + os.path.dirname(lib), "-o", exe],
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    assert r.returncode == 0, r.stdout
    r = subprocess.run([exe], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=60)
    assert r.returncode == 0, r.stdout
    assert "crc32c E3069283" in r.stdout and "bijection" in r.stdout and "status 1" in r.stdout
    # the device-side trainer compiles wherever the CUDA headers are (it runs on a GPU box only)
    cuda = os.environ.get("CUDA_HOME", "/usr/local/cuda")
    if os.path.exists(os.path.join(cuda, "include", "cuda_runtime.h")):
        r = subprocess.run([cc, "-std=c99", "-fsyntax-only", "-I", os.path.join(ROOT, "include"),
                            "-I", os.path.join(cuda, "include"),
                            os.path.join(ROOT, "examples", "c_abi", "train_step.c")],
                           stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        assert r.returncode == 0, r.stdout


def test_fabric_host_rendezvous_exchange_and_barrier(dtg):
    """Fabric.from_host's control plane (no torchrun): an all-gather of small objects and a real
    host barrier over the ps key/value board -- the ordering the GPU fabric relies on for
    seed-before-pull and unmap-before-free."""
    import threading
    from dtg_b200.parallel.fabric import Fabric
    from dtg_b200.parallel.rpc import VariableHost
    host, world = VariableHost(), 3
    got, order, errs = {}, [], []
    lock = threading.Lock()

    def rank_fn(r):
        try:
            exchange, barrier = Fabric.host_rendezvous(host, r, world, tag="t")
            got[(r, 0)] = exchange({"rank": r, "ptr": 1000 + r})
            got[(r, 1)] = exchange([r] * 2)                       # rounds do not mix
            time.sleep(0.05 * r)                                  # ranks arrive at different times
            with lock:
                order.append(("before", r))
            barrier()
            with lock:
                order.append(("after", r))
            barrier()                                             # epochs do not mix either
        except Exception as e:                                    # pragma: no cover
            errs.append(e)
    ts = [threading.Thread(target=rank_fn, args=(r,)) for r in range(world)]
    for t in ts:
        t.start()
    for t in ts:
        t.join(timeout=30)
    assert not errs and not any(t.is_alive() for t in ts)
    for r in range(world):
        assert got[(r, 0)] == [{"rank": q, "ptr": 1000 + q} for q in range(world)]
        assert got[(r, 1)] == [[q, q] for q in range(world)]
    # nobody leaves the barrier before everybody has reached it
    assert [k for k, _ in order[:world]] == ["before"] * world
    # a barrier that not every rank reaches times out instead of hanging
    _, lonely = Fabric.host_rendezvous(host, 0, 2, tag="lonely")
    with pytest.raises(TimeoutError):
        lonely(timeout=0.2)
