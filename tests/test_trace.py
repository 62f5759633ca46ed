"""Tracing: phase ranges, summaries with roofline fractions, chrome-trace
timelines and the ProfilerHook -- on the host clock here, on CUDA events in test_gpu_engine.py."""
import json
import os

import pytest


def _fake_clock():
    t = [0.0]

    def clock():
        t[0] += 0.001          # every reading advances 1 ms: a range lasts exactly 1 ms
        return t[0]
    return clock


def test_disabled_tracer_records_nothing(dtg):
    from dtg_b200.utils.trace import PhaseTracer
    tr = PhaseTracer(enabled=False)
    r1, r2 = tr.range("a"), tr.range("b", nbytes=10, kind="link")
    assert r1 is r2                       # one shared no-op context, nothing allocated
    with r1:
        pass
    tr.step_done()
    assert tr.collect() == [] and tr.summary() == {} and tr.step == 0


def test_summary_rates_and_roofline_fractions(dtg):
    from dtg_b200.utils.trace import NVLINK_PORT_GBPS, PhaseTracer
    peaks = {"hbm_gbs": 5000.0, "bf16_tflops": 1000.0, "bf16_tflops_sustained": 800.0, "source": "t"}
    tr = PhaseTracer(clock=_fake_clock(), peaks=peaks)
    for _ in range(4):
        with tr.range("pull", nbytes=450_000_000, kind="link"):       # 450 MB in 1 ms
            pass
        with tr.range("cast", nbytes=1_000_000_000, kind="hbm"):
            pass
        with tr.range("fwd+bwd", flops=4e11):                          # 0.4 TFLOP in 1 ms
            pass
        tr.step_done()
    s = tr.summary(skip_steps=1)
    assert s["pull"]["calls"] == 3 and abs(s["pull"]["ms_mean"] - 1.0) < 1e-6
    assert abs(s["pull"]["GBps"] - 450.0) < 1e-3
    assert abs(s["pull"]["fraction_of_roofline"] - 450.0 / NVLINK_PORT_GBPS) < 1e-6
    assert abs(s["cast"]["fraction_of_roofline"] - 1000.0 / 5000.0) < 1e-6
    assert abs(s["fwd+bwd"]["TFLOPs"] - 400.0) < 1e-3
    assert abs(s["fwd+bwd"]["fraction_of_bf16_peak"] - 0.5) < 1e-6


def test_chrome_trace_and_jsonl(dtg, tmp_path):
    from dtg_b200.utils.trace import PhaseTracer
    tr = PhaseTracer(clock=_fake_clock(), rank=3)
    with tr.range("forward"):
        pass
    with tr.range("pull(prefetch)", nbytes=1000, kind="link", lane="pull"):
        pass
    tr.step_done()
    p = tr.write_chrome_trace(str(tmp_path / "t" / "timeline.json"))
    doc = json.load(open(p))
    xs = [e for e in doc["traceEvents"] if e["ph"] == "X"]
    assert [e["name"] for e in xs] == ["forward", "pull(prefetch)"]
    assert all(e["pid"] == 3 for e in xs) and xs[0]["tid"] != xs[1]["tid"]
    assert xs[1]["ts"] > xs[0]["ts"] and abs(xs[0]["dur"] - 1000.0) < 1e-3
    lanes = {e["args"]["name"] for e in doc["traceEvents"] if e["name"] == "thread_name"}
    assert lanes == {"stream:main", "stream:pull"}
    j = tr.write_jsonl(str(tmp_path / "phases.jsonl"), scheme="hogwild")
    rows = [json.loads(l) for l in open(j)]
    assert {r["phase"] for r in rows} == {"forward", "pull(prefetch)"}
    assert all(r["scheme"] == "hogwild" and r["rank"] == 3 for r in rows)


def test_profiler_hook_on_host_engine(dtg, tmp_path):
    """tf.train.ProfilerHook's job: timeline-<step>.json every N runs, plus phases.jsonl."""
    from dtg_b200.models.toy import ToyModel
    from dtg_b200.parallel.engines import SchemeConfig, WorkerEngine
    from dtg_b200.utils.trace import ProfilerHook
    eng = WorkerEngine(ToyModel("zeros"), SchemeConfig(scheme="adag", update_window=3), 0, 1)
    assert not eng.tracer.enabled
    hook = ProfilerHook(eng.tracer, save_steps=2, output_dir=str(tmp_path))
    with eng.session(extra_hooks=[hook]) as sess:
        for _ in range(5):
            sess.run(eng.train_op)
    names = sorted(os.listdir(tmp_path))
    assert "phases.jsonl" in names
    assert [n for n in names if n.startswith("timeline-")] == \
        ["timeline-1.json", "timeline-3.json", "timeline-5.json"]
    doc = json.load(open(tmp_path / "timeline-3.json"))
    got = {e["name"] for e in doc["traceEvents"] if e["ph"] == "X"}
    assert got == {"local_steps", "push", "pull"}
    with pytest.raises(ValueError):
        ProfilerHook(eng.tracer)
