"""Every example program runs to completion the way the reference's run.sh starts its scripts
(1 ps + 2 workers as separate processes on localhost), CPU only."""
import os
import subprocess
import sys

import pytest

from conftest import ROOT, free_ports

EX = os.path.join(ROOT, "examples")
CASES = ["distributed_setup.py", "distributed_setup.py --supervisor", "hogwild.py",
         "downpour.py --variant downpour", "downpour.py --variant easy", "downpour.py --variant adag",
         "ssgd.py", "ssgd.py --per_worker_lr", "ssgd.py --window 5",
         "elastic_averaging.py --variant aeasgd", "elastic_averaging.py --variant aeamsgd",
         "elastic_averaging.py --variant dynsgd"]


def _free_base(n=3):
    """A base port with n consecutive free ports after it."""
    import socket
    for base in free_ports(20):
        ok = True
        for p in range(base, base + n):
            s = socket.socket()
            try:
                s.bind(("127.0.0.1", p))
            except OSError:
                ok = False
            finally:
                s.close()
        if ok:
            return base
    raise RuntimeError("no consecutive free ports")


@pytest.mark.timeout(300)
@pytest.mark.parametrize("case", CASES)
def test_example_runs_clean(case, tmp_path):
    env = dict(os.environ, PYTHONPATH=ROOT, OMP_NUM_THREADS="1")
    if case.startswith("distributed_setup.py"):        # 1 ps + 1 worker (Distributed-Setup/run.sh)
        cmd = ["bash", os.path.join(EX, "run_setup.sh")] + case.split()[1:]
    else:
        cmd = ["bash", os.path.join(EX, "run.sh"), os.path.join(EX, case.split()[0])] + case.split()[1:]
    cmd += ["--fast", "--base_port", str(_free_base())]
    r = subprocess.run(cmd, cwd=str(tmp_path), env=env, stdout=subprocess.PIPE,
                       stderr=subprocess.STDOUT, text=True, timeout=280)
    assert r.returncode == 0, r.stdout[-3000:]
    assert "Traceback" not in r.stdout, r.stdout[-3000:]
    assert r.stdout.strip(), "no output"


def test_non_distributed_example(tmp_path):
    r = subprocess.run([sys.executable, os.path.join(EX, "non_distributed_setup.py"), "--fast"],
                       cwd=str(tmp_path), env=dict(os.environ, PYTHONPATH=ROOT),
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=120)
    assert r.returncode == 0 and "Traceback" not in r.stdout, r.stdout[-2000:]


# ---- tutorials (the notebooks): fixed ports 2222-2224 like the reference ----------------------
TUT = os.path.join(ROOT, "tutorials")


def _ports_free(ports):
    import socket
    for p in ports:
        s = socket.socket()
        try:
            s.bind(("127.0.0.1", p))
        except OSError:
            return False
        finally:
            s.close()
    return True


def _run(path, timeout=120):
    return subprocess.run([sys.executable, path], cwd=os.path.dirname(path),
                          env=dict(os.environ, PYTHONPATH=ROOT), stdout=subprocess.PIPE,
                          stderr=subprocess.STDOUT, text=True, timeout=timeout)


@pytest.mark.timeout(200)
def test_tutorial_servers_and_local_then_global():
    if not _ports_free([2222, 2223]):
        pytest.skip("ports 2222/2223 are in use")
    r = _run(os.path.join(TUT, "servers.py"))
    assert r.returncode == 0 and "port 2222 is accepting connections" in r.stdout, r.stdout[-2000:]
    ps = subprocess.Popen([sys.executable, os.path.join(TUT, "parameter_server.py")], cwd=TUT,
                          env=dict(os.environ, PYTHONPATH=ROOT), stdout=subprocess.PIPE,
                          stderr=subprocess.STDOUT, text=True)
    try:
        r = _run(os.path.join(TUT, "local_then_global_variables.py"))
        assert r.returncode == 0 and "Traceback" not in r.stdout, r.stdout[-2000:]
        # ops that update a variable run where the variable lives (Servers/Local-then-Global notebooks)
        assert "/job:ps/replica:0/task:0" in r.stdout and "/job:worker/replica:0/task:0" in r.stdout
        ps.wait(timeout=60)                       # the ps leaves once its worker reported done
        assert ps.returncode == 0
    finally:
        if ps.poll() is None:
            ps.kill()


def test_tutorial_notebooks_match_their_scripts():
    """tutorials/notebooks/*.ipynb (recorded outputs, the reference's form) are generated from the
    `# %%` scripts by tools/make_notebooks.py; their cells must not drift from the scripts."""
    import json
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "make_notebooks.py"), "--check"],
                       cwd=ROOT, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=60)
    assert r.returncode == 0, r.stdout
    nb = json.load(open(os.path.join(ROOT, "tutorials", "notebooks", "local_then_global_variables.ipynb")))
    assert nb["nbformat"] == 4
    text = "".join("".join(o["text"]) for c in nb["cells"] if c["cell_type"] == "code"
                   for o in c["outputs"])
    # the golden trace of the reference's notebook: update ops sit with the variable on the ps,
    # a local update moves `a` only, a global update moves `g/a` only
    assert "global_update  /job:ps/replica:0/task:0" in text
    assert "a   = [0.10000000149011612]" in text
