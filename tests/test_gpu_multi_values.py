"""Value-level multi-GPU tests (2, 4 and 8 ranks, with and without NVLS): the fused
reduce-scatter + update + all-gather kernel against the torch.distributed result, N-of-M with a
late replica, a dedicated ps rank, a rank that never arrives, Hogwild push sums (sharded and
striped ps), the wgrad-GEMM epilogue pushing into a peer's HBM, Adagrad-on-ps.

One torchrun launch per (world, nvls) runs every case (tests/gpu_value_worker.py); the
parametrised tests below read its per-case verdicts, so a failure names the case."""
import functools
import json
import os
import subprocess
import sys

import pytest

from conftest import free_ports

pytestmark = [pytest.mark.gpu, pytest.mark.multigpu]
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

SYNC_CASES = ["ssgd_mean", "ssgd_adagrad", "ssgd_lr_scaled", "tower_avg_buckets", "sdag_window_scale",
              "n_of_m_late_rank_dropped", "ps_rank_contributes_nothing",
              "absent_rank_times_out_without_hanging"]
STORE_CASES = ["hogwild_push_sum_sharded", "hogwild_push_sum_striped",
               "wgrad_gemm_push_into_peer_shard", "adagrad_push_one_writer_at_a_time"]


def _ngpu():
    try:
        import torch
        return torch.cuda.device_count() if torch.cuda.is_available() else 0
    except Exception:
        return 0


@functools.lru_cache(maxsize=None)
def launch(world: int, nvls: int) -> dict:
    port = free_ports(1)[0]
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world),
           "--master-addr", "127.0.0.1", "--master-port", str(port),
           os.path.join(ROOT, "tests", "gpu_value_worker.py")]
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600,
                       cwd=ROOT, env=dict(os.environ, DTG_NVLS=str(nvls)))
    lines = [l for l in r.stdout.splitlines() if l.startswith("RESULT ")]
    if r.returncode != 0 or not lines:
        return {"__error__": r.stdout[-3000:]}
    return json.loads(lines[-1][7:])


def check(world, nvls, case):
    if _ngpu() < world:
        pytest.skip("needs %d GPUs" % world)
    res = launch(world, nvls)
    assert "__error__" not in res, res["__error__"]
    assert case in res, "case did not run: %s" % sorted(res)
    assert res[case]["ok"], json.dumps(res[case])[:3000]


@pytest.mark.parametrize("case", SYNC_CASES)
@pytest.mark.parametrize("nvls", [0, 1])
@pytest.mark.parametrize("world", [2, 4, 8])
def test_sync_values(world, nvls, case):
    check(world, nvls, case)
    if nvls and case == "ssgd_mean":
        assert all(r["nvls"] for r in launch(world, nvls)[case]["ranks"]), "NVLS path was not taken"


@pytest.mark.parametrize("case", STORE_CASES)
@pytest.mark.parametrize("world", [2, 4, 8])
def test_store_values(world, case):
    check(world, 0, case)
