"""Source-level guards on the kernels.

TMA multicast loads here read only this GPU's memory, since on B200 a TMA multicast load from
another GPU's memory hung the GPU.  The general GEMM entry points cannot know whose memory their
operands live in, so no kernel of this library issues a TMA load with cluster multicast at all;
`tcgen05.commit ... multicast::cluster` (an mbarrier arrive, not a load) and NVLS `multimem`
instructions are different things and are used."""
import os
import re
import shutil
import subprocess

import pytest

from conftest import ROOT

CSRC = os.path.join(ROOT, "distributed-tensorflow-guide_b200", "ops", "csrc")
LIB = os.path.join(ROOT, "distributed-tensorflow-guide_b200", "ops", "libdtg_b200.so")


def _asm_statements(text):
    """Every inline-PTX string of a source file, with adjacent string literals joined."""
    for m in re.finditer(r"asm\s*(?:volatile)?\s*\((.*?)\)\s*;", text, re.S):
        yield "".join(re.findall(r'"((?:[^"\\]|\\.)*)"', m.group(1)))


def test_no_tma_load_uses_cluster_multicast():
    offenders = []
    for fn in sorted(os.listdir(CSRC)):
        if not fn.endswith((".cu", ".cuh")):
            continue
        for stmt in _asm_statements(open(os.path.join(CSRC, fn)).read()):
            if "cp.async.bulk" in stmt and "multicast" in stmt:
                offenders.append((fn, stmt[:120]))
    assert not offenders, offenders
    # the allowed multicast is still there: the slot-free arrive of the pair kernels
    gemm = open(os.path.join(CSRC, "gemm_sm100.cu")).read()
    assert any("tcgen05.commit" in s and "multicast::cluster" in s for s in _asm_statements(gemm))


def _check_sass(sass):
    loads = set(re.findall(r"\b(UTMALDG[.\w]*|UBLKCP[.\w]*)", sass))
    assert loads, "no TMA loads found at all -- wrong library?"
    assert not [m for m in loads if "MULTICAST" in m or ".MC" in m], loads
    # what the design relies on is present
    assert any(m.startswith("UTMALDG.4D.IM2COL") for m in loads)          # implicit-GEMM convolution
    for needle in ("UTCHMMA.2CTA", "UTMASTG.2D", "UTMAREDG.2D.ADD", "LDTM.x32", "LDGMC.E.ADD.F32"):
        assert needle in sass, needle


def test_committed_sass_excerpts_have_no_multicast_tma_loads():
    """profiles/sass/ (tools/sass_evidence.py): the per-kernel excerpts of the shipped listing."""
    d = os.path.join(ROOT, "profiles", "sass")
    text = "\n".join(open(os.path.join(d, f)).read() for f in sorted(os.listdir(d)) if f.endswith(".sass"))
    _check_sass(text)


@pytest.mark.skipif(os.environ.get("DTG_SLOW_TESTS") != "1",
                    reason="full cuobjdump of the library takes ~40 s; set DTG_SLOW_TESTS=1")
def test_built_library_has_no_multicast_tma_loads():
    if not os.path.exists(LIB) or shutil.which("cuobjdump") is None:
        pytest.skip("library or cuobjdump not available")
    _check_sass(subprocess.run(["cuobjdump", "-sass", LIB], stdout=subprocess.PIPE, text=True,
                               timeout=300).stdout)
