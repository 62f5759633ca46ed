"""Build the tutorials as Jupyter notebooks WITH recorded outputs, the form the reference ships
them in (`/root/reference/Basics-Tutorial/*.ipynb`: one kernel per task, outputs kept in the file as
golden traces).  The `# %%` scripts under `tutorials/` stay the source of truth; this tool

  * splits each script into cells (`# %% title` -> a markdown heading + a code cell; the module
    docstring -> the notebook's first markdown cell),
  * runs every cell in one namespace with the other roles of that walkthrough started as real
    processes beside it (ps first, like the reference's instructions), capturing each cell's
    stdout, and
  * writes `tutorials/notebooks/<name>.ipynb` (nbformat 4).

    python tools/make_notebooks.py            # regenerate (needs ports 2222-2224 free)
    python tools/make_notebooks.py --check    # cells of the committed notebooks == the scripts' cells

No Jupyter installation is needed to build or check; any Jupyter opens the result.
"""
import contextlib
import io
import json
import os
import re
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TUT = os.path.join(ROOT, "tutorials")
OUT = os.path.join(TUT, "notebooks")

# notebook -> (script, companions started before it: (script, delay before the next start))
PLAN = [
    ("servers", "servers.py", []),
    ("parameter_server", "parameter_server.py", None),              # a companion: recorded as it exits
    ("local_then_global_variables", "local_then_global_variables.py", [("parameter_server.py", 1.0)]),
    ("multiple_workers_parameter_server", "multiple_workers/parameter_server.py", None),
    ("multiple_workers_worker1", "multiple_workers/worker1.py",
     [("multiple_workers/parameter_server.py", 1.0)]),
    ("multiple_workers_worker2", "multiple_workers/worker2.py", None),   # started while worker1 waits
]


def split_cells(src: str):
    """[(kind, text)] with kind in {"markdown", "code"}."""
    cells = []
    m = re.match(r'\s*"""(.*?)"""\s*\n', src, re.S)
    if m:
        cells.append(("markdown", m.group(1).strip()))
        src = src[m.end():]
    parts = re.split(r"(?m)^# %%(.*)$", src)
    head = parts[0].strip("\n")
    if head.strip():
        cells.append(("code", head))
    for i in range(1, len(parts), 2):
        title, body = parts[i].strip(), parts[i + 1].strip("\n")
        if title:
            cells.append(("markdown", "### " + title))
        if body.strip():
            cells.append(("code", body))
    return cells


def notebook(cells, outputs):
    nb_cells, k = [], 0
    for kind, text in cells:
        lines = text.split("\n")
        source = [l + "\n" for l in lines[:-1]] + [lines[-1]]
        if kind == "markdown":
            nb_cells.append({"cell_type": "markdown", "metadata": {}, "source": source})
            continue
        out = outputs[k] if k < len(outputs) else ""
        k += 1
        nb_cells.append({
            "cell_type": "code", "execution_count": k, "metadata": {}, "source": source,
            "outputs": ([{"name": "stdout", "output_type": "stream",
                          "text": [l + "\n" for l in out.rstrip("\n").split("\n")]}] if out.strip() else [])})
    return {"cells": nb_cells, "nbformat": 4, "nbformat_minor": 5,
            "metadata": {"kernelspec": {"display_name": "Python 3", "language": "python", "name": "python3"},
                         "language_info": {"name": "python"}}}


def exec_cells(script: str):
    """Child mode: run the script's code cells one by one, print a JSON list of their stdout."""
    path = os.path.join(TUT, script)
    os.chdir(os.path.dirname(path))
    sys.path.insert(0, os.path.dirname(path))
    ns = {"__name__": "__main__", "__file__": path}
    outs = []
    for kind, text in split_cells(open(path).read()):
        if kind != "code":
            continue
        buf = io.StringIO()
        with contextlib.redirect_stdout(buf):
            exec(compile(text, path, "exec"), ns)
        outs.append(buf.getvalue())
    sys.__stdout__.write("\n@@CELLS@@" + json.dumps(outs) + "\n")
    sys.__stdout__.flush()


def run_child(script: str) -> subprocess.Popen:
    return subprocess.Popen([sys.executable, os.path.abspath(__file__), "--exec", script], cwd=ROOT,
                            env=dict(os.environ, PYTHONPATH=ROOT), stdout=subprocess.PIPE,
                            stderr=subprocess.STDOUT, text=True)


def outputs_of(proc: subprocess.Popen, timeout=120):
    out, _ = proc.communicate(timeout=timeout)
    m = re.search(r"@@CELLS@@(.*)", out)
    if proc.returncode != 0 or not m:
        raise RuntimeError("tutorial failed:\n" + out[-3000:])
    return json.loads(m.group(1))


def build_all():
    os.makedirs(OUT, exist_ok=True)
    recorded = {}
    # 1. servers (alone)
    recorded["servers.py"] = outputs_of(run_child("servers.py"))
    # 2. one ps + one worker
    ps = run_child("parameter_server.py")
    time.sleep(1.5)
    recorded["local_then_global_variables.py"] = outputs_of(run_child("local_then_global_variables.py"))
    recorded["parameter_server.py"] = outputs_of(ps)
    # 3. one ps + two workers, worker 2 started while worker 1 waits for it
    ps = run_child("multiple_workers/parameter_server.py")
    time.sleep(1.5)
    w1 = run_child("multiple_workers/worker1.py")
    time.sleep(6.0)
    recorded["multiple_workers/worker2.py"] = outputs_of(run_child("multiple_workers/worker2.py"))
    recorded["multiple_workers/worker1.py"] = outputs_of(w1)
    recorded["multiple_workers/parameter_server.py"] = outputs_of(ps)
    for name, script, _ in PLAN:
        cells = split_cells(open(os.path.join(TUT, script)).read())
        with open(os.path.join(OUT, name + ".ipynb"), "w") as f:
            json.dump(notebook(cells, recorded[script]), f, indent=1)
            f.write("\n")
        print("wrote tutorials/notebooks/%s.ipynb (%d cells)" % (name, len(cells)))


def check_all() -> int:
    bad = 0
    for name, script, _ in PLAN:
        cells = split_cells(open(os.path.join(TUT, script)).read())
        p = os.path.join(OUT, name + ".ipynb")
        try:
            nb = json.load(open(p))
        except (OSError, ValueError) as e:
            print("MISSING/INVALID %s: %s" % (p, e))
            bad += 1
            continue
        got = [(c["cell_type"], "".join(c["source"])) for c in nb["cells"]]
        if nb.get("nbformat") != 4 or got != cells:
            print("STALE %s: regenerate with tools/make_notebooks.py" % p)
            bad += 1
    return bad


if __name__ == "__main__":
    if "--exec" in sys.argv:
        exec_cells(sys.argv[sys.argv.index("--exec") + 1])
    elif "--check" in sys.argv:
        sys.exit(1 if check_all() else 0)
    else:
        build_all()
