"""The allocation-free variant of the synchronous aggregation kernel, called the way a
C/C++ trainer would call it: straight through the C ABI (ctypes), with nothing but raw
device pointers.

Some users of ``libdtg_b200.so`` cannot let the library's caller allocate scratch for
it: a trainer that gave all of HBM to its own arena at start-up, an asynchronous worker
loop that cannot afford the device-wide synchronisation hidden in ``cudaMalloc``, or a
host with a fixed memory budget.  ``dtg_sync_rs_update_ag_noalloc`` keeps the kernel's
scratch (the "who contributes" word and the last-block counter) in an array inside the
library instead.  This test checks that it produces what ``dtg_sync_rs_update_ag``
produces with caller-provided scratch, on two calls in a row (the second call finds the
first call's leftovers in the shared array), and that a bucket index beyond the built-in
limit is refused.

Collected last on purpose (``zz``): if it fails under ``-x`` every other test has run.
"""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu

N = 200_003            # not a multiple of 4: exercises the scalar tail too
BLOCKS = 16


def rel_err(out, ref):
    out, ref = out.float(), ref.float()
    return ((out - ref).abs().max() / (ref.abs().max() + 1e-6)).item()


def _lib():
    import dtg_b200  # noqa: F401
    from dtg_b200 import ops
    lib = ops.lib()
    vp, i32 = ctypes.c_void_p, ctypes.c_int
    lib.dtg_sync_rs_update_ag_noalloc.argtypes = [vp, i32, i32, vp]
    lib.dtg_sync_rs_update_ag_noalloc.restype = i32
    lib.dtg_sync_noalloc_max_buckets.restype = i32
    return lib


def _run_two_steps(lib, use_noalloc: bool, bucket: int = 3):
    """Two consecutive aggregation steps of a one-rank job (world = 1): Adagrad on the
    master shard, parameters broadcast to the bf16 and fp32 mirrors."""
    from dtg_b200.parallel.peer_store import SyncParams
    dev = "cuda"
    g = torch.Generator(device="cpu").manual_seed(7)
    master = torch.randn(N, generator=g).to(dev)
    slot = torch.full((N,), 0.1, device=dev)
    grad = torch.zeros(N, device=dev)
    p32 = torch.zeros(N, device=dev)
    p16 = torch.zeros(N, device=dev, dtype=torch.bfloat16)
    flags = torch.zeros(1024, device=dev, dtype=torch.int32)
    status = torch.zeros(1, device=dev, dtype=torch.int32)
    # scratch the existing entry point needs from its caller
    decision = torch.zeros(1, device=dev, dtype=torch.int64)
    counter = torch.zeros(1, device=dev, dtype=torch.int32)

    p = SyncParams()
    for q in range(8):
        p.grad[q] = p.param_f32[q] = p.param_bf16[q] = p.flags[q] = None
    p.grad[0], p.param_f32[0] = grad.data_ptr(), p32.data_ptr()
    p.param_bf16[0], p.flags[0] = p16.data_ptr(), flags.data_ptr()
    p.mc_grad = p.mc_param_f32 = p.mc_param_bf16 = None
    p.master, p.slot = master.data_ptr(), slot.data_ptr()
    p.status = status.data_ptr()
    if use_noalloc:
        p.decision = p.cta_counter = None          # the library supplies these
    else:
        p.decision, p.cta_counter = decision.data_ptr(), counter.data_ptr()
    p.shard_begin, p.shard_end = 0, N
    p.timeout_cycles = int(5e9)
    p.world, p.rank = 1, 0
    p.ready_base, p.done_base = 64, 72
    p.replicas_to_aggregate = 1
    p.lr, p.gscale = 0.05, 1.0
    p.adagrad, p.contributes, p.owner_mask = 1, 1, 1
    stream = torch.cuda.current_stream().cuda_stream
    outs = []
    for step in (1, 2):
        grad.copy_(torch.randn(N, generator=g).to(dev))
        p.step = step
        if use_noalloc:
            rc = lib.dtg_sync_rs_update_ag_noalloc(ctypes.byref(p), bucket, BLOCKS, stream)
        else:
            rc = lib.dtg_sync_rs_update_ag(ctypes.byref(p), BLOCKS, stream)
        assert rc == 0, lib.dtg_last_error().decode()
        torch.cuda.synchronize()
        assert int(status.item()) == 0
        outs.append((master.clone(), slot.clone(), p32.clone(), p16.clone()))
    return outs


def test_noalloc_entry_matches_existing_entry_on_two_calls():
    lib = _lib()
    ref = _run_two_steps(lib, use_noalloc=False)
    got = _run_two_steps(lib, use_noalloc=True)
    for step, (r, o) in enumerate(zip(ref, got), 1):
        for name, a, b in zip(("master", "slot", "p32", "p16"), r, o):
            assert rel_err(b, a) < 1e-6, (step, name)
    # the update really happened (second step differs from the first)
    assert rel_err(got[1][0], got[0][0]) > 1e-4


def test_noalloc_entry_rejects_bucket_beyond_builtin_scratch():
    from dtg_b200.parallel.peer_store import SyncParams
    lib = _lib()
    limit = lib.dtg_sync_noalloc_max_buckets()
    assert limit >= 1
    p = SyncParams()
    p.world, p.rank = 1, 0
    status = torch.zeros(1, device="cuda", dtype=torch.int32)
    p.status = status.data_ptr()
    rc = lib.dtg_sync_rs_update_ag_noalloc(ctypes.byref(p), limit, BLOCKS,
                                           torch.cuda.current_stream().cuda_stream)
    assert rc == 4                                   # ERR_TOO_LARGE, nothing launched
    assert "outside the built-in scratch" in lib.dtg_last_error().decode()
