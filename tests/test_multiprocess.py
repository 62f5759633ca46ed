"""Integration: real OS processes on localhost ports, launched the reference's way
(`--job_name J --task_index T`, Hogwild/run.sh:2-4).  CPU only."""
import json
import os
import re
import subprocess
import sys
import time

import pytest

from conftest import ROOT, free_ports


def launch(args, n_ps=1, n_workers=2, timeout=120, extra_env=None):
    base = free_ports(1)[0]
    # consecutive ports from a free base; retry on the rare collision
    cluster = {"ps": ["localhost:%d" % p for p in free_ports(n_ps)],
               "worker": ["localhost:%d" % p for p in free_ports(n_workers)]}
    env = dict(os.environ, PYTHONPATH=ROOT, OMP_NUM_THREADS="1")
    env.update(extra_env or {})
    common = [sys.executable, "-m", "dtg_b200", "--cluster", json.dumps(cluster),
              "--device", "cpu"] + args
    procs, outs = [], []
    for job, n in (("ps", n_ps), ("worker", n_workers)):
        for i in range(n):
            procs.append((job, i, subprocess.Popen(
                common + ["--job_name", job, "--task_index", str(i)], env=env, cwd=ROOT,
                stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    deadline = time.time() + timeout
    for job, i, p in procs:
        try:
            out, _ = p.communicate(timeout=max(1, deadline - time.time()))
        except subprocess.TimeoutExpired:
            p.kill()
            out, _ = p.communicate()
            out += "\n<<TIMEOUT>>"
        outs.append((job, i, p.returncode, out))
    return outs


def final_c(out):
    m = re.findall(r"^\[([-\d.e]+), [-\d.e]+\] global step: (\d+)", out, flags=re.M)
    return (float(m[-1][0]), int(m[-1][1])) if m else None


def assert_clean(outs, allow=()):
    for job, i, code, out in outs:
        if (job, i) in allow:
            continue
        assert code == 0, "%s:%d exited %s\n%s" % (job, i, code, out[-2000:])
        if job == "worker":
            assert "closed cleanly" in out, out[-2000:]


@pytest.mark.timeout(180)
def test_hogwild_two_workers_total_pushes():
    outs = launch(["--scheme", "hogwild", "--max_iters", "50", "--log_every", "50"])
    assert_clean(outs)
    workers = [o for o in outs if o[0] == "worker"]
    cs = [final_c(o[3]) for o in workers]
    # global_step == total pushes; c approaches 100 monotonically (never overshoots
    # for lr=1e-4): after k unlocked pushes from zero c <= 100*(1-(1-2e-4)^k)
    assert max(c[1] for c in cs) == 100
    assert 0.9 < max(c[0] for c in cs) <= 100 * (1 - (1 - 2e-4) ** 100) + 1e-3


@pytest.mark.timeout(180)
def test_ssgd_reaches_last_step_and_stops():
    outs = launch(["--scheme", "ssgd", "--last_step", "10", "--log_every", "1"])
    assert_clean(outs)
    best = max(final_c(o[3]) for o in outs if o[0] == "worker" and final_c(o[3]))
    assert best[1] == 10
    assert best[0] == pytest.approx(0.19982007, rel=1e-3)


@pytest.mark.timeout(180)
def test_downpour_and_adag_stop_hook_and_checkpoint(tmp_path):
    d = str(tmp_path / "logdir")
    outs = launch(["--scheme", "adag", "--last_step", "12", "--log_every", "1",
                   "--checkpoint_dir", d, "--save_checkpoint_secs", "0.01"])
    assert_clean(outs)
    sys.path.insert(0, ROOT)
    import dtg_b200 as dtg
    ck = dtg.train.latest_checkpoint(d)
    assert ck is not None
    data = dtg.train.load_checkpoint(ck)
    assert sorted(data) == ["g/Variable", "g/Variable_1", "global_step"]
    assert int(data["global_step"]) >= 12
    assert 0 < data["g/Variable"][0].item() < 1.0


@pytest.mark.timeout(240)
def test_config1_mnist_mlp_ssgd_1ps_2workers():
    """BASELINE.json config 1: MLP 784-100-10 as Synchronous-SGD, 1 ps + 2 workers."""
    outs = launch(["--scheme", "ssgd", "--model", "mnist_mlp", "--batch_size", "32",
                   "--lr", "0.05", "--last_step", "30", "--log_every", "10"])
    assert_clean(outs)
    for job, i, code, out in outs:
        if job == "worker":
            m = re.search(r"'global_step': (\d+).*'loss': ([\d.]+)", out)
            assert m and int(m.group(1)) >= 30
            assert float(m.group(2)) < 3.5          # random labels: ~ln(10) + noise


@pytest.mark.timeout(240)
def test_two_ps_tasks_shard_variables():
    outs = launch(["--scheme", "downpour", "--last_step", "6", "--log_every", "1"], n_ps=2)
    assert_clean(outs)
    assert any(final_c(o[3]) and final_c(o[3])[1] >= 6 for o in outs if o[0] == "worker")


@pytest.mark.timeout(240)
def test_async_survives_worker_death():
    """Fault test: worker 1 dies at its 5th round; Hogwild worker 0 keeps going and
    finishes; the ps is told only one worker is done, so we bound it with --max_iters
    on the survivor and kill the ps afterwards (the reference's `pkill python`)."""
    outs = launch(["--scheme", "hogwild", "--max_iters", "40", "--log_every", "40",
                   "--inject_fault", "worker:1@5"], timeout=25)
    by = {(j, i): (c, o) for j, i, c, o in outs}
    assert by[("worker", 1)][0] == 17
    assert by[("worker", 0)][0] == 0 and "closed cleanly" in by[("worker", 0)][1]
    c, step = final_c(by[("worker", 0)][1])
    assert 40 <= step <= 45


@pytest.mark.timeout(240)
@pytest.mark.parametrize("scheme", ["dynsgd", "aeasgd", "aeamsgd"])
def test_reference_todo_schemes_on_a_real_cluster(scheme):
    """The reference's TODO list (README.md:40-42) as 1 ps + 2 worker processes: the shared
    state on the ps moves toward the optimum and the step counter counts every round."""
    outs = launch(["--scheme", scheme, "--last_step", "20", "--log_every", "1", "--lr", "0.01"])
    assert_clean(outs)
    best = max(final_c(o[3]) for o in outs if o[0] == "worker" and final_c(o[3]))
    assert best[1] >= 20
    # closer to the optimum (c = 100) than the start (c = 0); Nesterov momentum may overshoot it
    assert abs(best[0] - 100.0) < 99.0, best


@pytest.mark.timeout(180)
def test_profile_dir_writes_timelines_per_worker(tmp_path):
    """--profile_dir: every worker leaves chrome-trace timelines and a per-phase summary."""
    d = str(tmp_path / "prof")
    outs = launch(["--scheme", "adag", "--update_window", "3", "--last_step", "6",
                   "--max_iters", "12", "--profile_dir", d, "--profile_steps", "2"])
    assert_clean(outs)
    for k in (0, 1):
        names = os.listdir(os.path.join(d, "worker%d" % k))
        assert "phases.jsonl" in names and any(n.startswith("timeline-") for n in names)
        rows = [json.loads(l) for l in open(os.path.join(d, "worker%d" % k, "phases.jsonl"))]
        assert {"local_steps", "push", "pull"} <= {r["phase"] for r in rows}
        assert all(r["rank"] == k and r["ms_total"] > 0 for r in rows)
