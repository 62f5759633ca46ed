"""torchrun worker for tests/test_gpu_multi_values.py: VALUE-level checks of the multi-GPU data
plane (2, 4 or 8 ranks, one per GPU).  Every case computes what the result must be with
``torch.distributed`` (NCCL all_reduce / all_gather of the same inputs) and compares the
kernels' output against it to fp32 tolerance.

Contract being checked: SyncReplicas aggregation = mean over the replicas that arrived, late
ones dropped (Synchronous-SGD/README.md:3, ssgd.py:51-55); per-worker learning rates as a
producer-side scale (Synchronous-SGD-different-learning-rates/ssgd.py:59-64); Hogwild pushes
= lock-free sums (Hogwild/Hogwild.py:44).

One launch runs every case and prints ``RESULT {case: {...}}``; the absent-rank case runs last
because it leaves the signal pads of its buckets behind the others'.
"""
import json
import os
import sys
import time
import traceback

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def rel_err(out, ref):
    out, ref = out.float(), ref.float()
    return ((out - ref).abs().max() / (ref.abs().max() + 1e-12)).item()


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    import dtg_b200  # noqa: F401
    from dtg_b200 import ops
    from dtg_b200.parallel.fabric import Fabric
    from dtg_b200.parallel.flat import FlatLayout
    from dtg_b200.parallel.peer_store import PeerStore, PeerSync

    fabric = Fabric.from_torch_distributed(dev)
    results = {}
    step_base = [0]

    def gather_all(t):
        out = [torch.empty_like(t) for _ in range(world)]
        dist.all_gather(out, t)
        return out

    def all_ok(ok: bool, info: dict, name: str):
        flags = [None] * world
        dist.all_gather_object(flags, (bool(ok), info))
        results[name] = {"ok": all(f[0] for f in flags), "ranks": [f[1] for f in flags]}

    def new_sync(n, **kw):
        sy = PeerSync(fabric, n, first_step=step_base[0], **kw)
        step_base[0] += 1000
        return sy

    def case(name):
        def deco(fn):
            try:
                fn(name)
            except Exception as e:           # every rank must keep going (collectives ahead)
                results[name] = {"ok": False, "error": "%s: %s" % (type(e).__name__, e),
                                 "trace": traceback.format_exc()[-800:]}
            torch.cuda.synchronize()
            dist.barrier()
            return fn
        return deco

    g = torch.Generator(device=dev).manual_seed(1000 + rank)
    n = 1_000_003 + 5          # odd tail: exercises the scalar remainder of the last shard

    # ---------------------------------------------------------------- synchronous aggregation
    def sync_case(name, adagrad, scales=None, bounds=None, steps=3, gscale=1.0, lr=0.05):
        sy = new_sync(n, lr=lr, adagrad=adagrad, bucket_bounds=bounds)
        p0 = torch.randn(n, device=dev, generator=torch.Generator(device=dev).manual_seed(7))
        sy.load_master(p0)
        ref = p0.clone()
        acc = torch.full_like(p0, 0.1)
        worst = 0.0
        for _ in range(steps):
            gk = torch.randn(n, device=dev, generator=g)
            if scales is not None:           # per-worker lr: (lr_k / lr_base) * N folded by the producer
                gk = gk * scales[rank]
            sy.g32.copy_(gk)
            tot = gk.clone()
            dist.all_reduce(tot)                         # NCCL: what the result must equal
            mean = tot * (gscale / world)
            if adagrad:
                acc += mean * mean
                ref -= lr * mean / acc.sqrt()
            else:
                ref -= lr * mean
            torch.cuda.synchronize()
            dist.barrier()
            sy.step(gscale=gscale)
            torch.cuda.synchronize()
            worst = max(worst, rel_err(sy.p32, ref), 0.0)
            e16 = rel_err(sy.p16, ref.bfloat16())
        ident = gather_all(sy.p32)
        same = all(torch.equal(ident[0], t) for t in ident)
        ok = worst < 2e-6 and e16 < 1e-2 and same and int(sy.status.item()) == 0 and \
            sy.step_no % 1000 == steps
        all_ok(ok, {"rank": rank, "err": worst, "err16": e16, "identical": same,
                    "status": int(sy.status.item()), "nvls": bool(sy.nvls)}, name)

    @case("ssgd_mean")
    def _(name):
        sync_case(name, adagrad=False)

    @case("ssgd_adagrad")
    def _(name):
        sync_case(name, adagrad=True)

    @case("ssgd_lr_scaled")
    def _(name):
        lrs = [0.1] + [1e-4 * (k + 1) for k in range(1, world)]      # ssgd.py:59-62 style
        sync_case(name, adagrad=False, scales=[l / 1.0 * world for l in lrs], lr=1.0)

    @case("tower_avg_buckets")
    def _(name):
        cuts = [0, 4 * 1000, 4 * 50_000, 4 * 50_001, 4 * 200_000, n]      # uneven, one tiny bucket
        sync_case(name, adagrad=False, bounds=cuts)

    @case("sdag_window_scale")
    def _(name):
        sync_case(name, adagrad=False, gscale=1.0 / 5.0)                  # mean over a window of 5

    # ---------------------------------------------------------------- N of M: a late replica
    @case("n_of_m_late_rank_dropped")
    def _(name):
        N = world - 1            # with 2 ranks N = 1: each owner takes whoever is first (itself)
        sy = new_sync(n, lr=0.05, replicas_to_aggregate=N)
        p0 = torch.randn(n, device=dev, generator=torch.Generator(device=dev).manual_seed(8))
        sy.load_master(p0)
        gk = torch.randn(n, device=dev, generator=g)
        sy.g32.copy_(gk)
        gs = gather_all(gk)
        late = world - 1
        torch.cuda.synchronize()
        dist.barrier()
        if rank == late:
            torch.cuda._sleep(int(0.25 * 1.9e9))          # ~0.25 s behind everybody else
        sy.step()
        torch.cuda.synchronize()
        full = sum(gs) / world                                      # nobody dropped
        early = sum(gs[k] for k in range(world) if k != late) / N   # the late one dropped
        cand_full, cand_early = p0 - 0.05 * full, p0 - 0.05 * early
        # the late rank's own shard is decided when it finally runs: everyone is there by then
        from dtg_b200.placement import flat_shards
        sh = flat_shards(n, world, 4)
        verdict = []
        for owner, (lo, hi) in enumerate(sh):
            if hi <= lo:
                continue
            got = sy.p32[lo:hi]
            ef, ee = rel_err(got, cand_full[lo:hi]), rel_err(got, cand_early[lo:hi])
            verdict.append((owner, "early" if ee < 2e-6 else ("full" if ef < 2e-6 else "WRONG")))
        ident = gather_all(sy.p32)
        same = all(torch.equal(ident[0], t) for t in ident)
        dropped = all(v == "early" for o, v in verdict if o != late)
        ok = same and dropped and all(v != "WRONG" for _, v in verdict) and int(sy.status.item()) == 0
        all_ok(ok, {"rank": rank, "verdict": verdict, "identical": same,
                    "status": int(sy.status.item())}, name)

    # ---------------------------------------------------------------- dedicated ps rank
    @case("ps_rank_contributes_nothing")
    def _(name):
        sy = new_sync(n, lr=0.05, replicas_to_aggregate=world - 1, contributes=(rank != 0))
        p0 = torch.randn(n, device=dev, generator=torch.Generator(device=dev).manual_seed(9))
        sy.load_master(p0)
        gk = torch.randn(n, device=dev, generator=g)
        sy.g32.copy_(gk if rank != 0 else torch.full_like(gk, 1e6))   # poison: must never be read
        gs = gather_all(gk)
        torch.cuda.synchronize()
        dist.barrier()
        sy.step()
        torch.cuda.synchronize()
        ref = p0 - 0.05 * sum(gs[1:]) / (world - 1)
        err = rel_err(sy.p32, ref)
        all_ok(err < 2e-6 and int(sy.status.item()) == 0,
               {"rank": rank, "err": err, "status": int(sy.status.item())}, name)

    # ---------------------------------------------------------------- Hogwild: lock-free sums
    def hogwild_case(name, striped):
        specs = []
        for i in range(world):
            specs += [("w%d" % i, (300, 1003), (300, 1003)), ("b%d" % i, (1003,), (1003,))]
        layout = FlatLayout(specs, num_shards=1 if striped else world)
        if striped and not fabric.striped_supported():
            results[name] = {"ok": True, "skipped": "no CUDA VMM"}
            return
        st = PeerStore(fabric, layout, None if striped else list(range(layout.num_shards)),
                       striped=striped)
        total = layout.total
        seed = torch.randn(total, device=dev, generator=torch.Generator(device=dev).manual_seed(11))
        if rank == 0:
            st.seed(seed)
        torch.cuda.synchronize()
        dist.barrier()
        gk = torch.randn(total, device=dev, generator=g)
        K = 5
        for _ in range(K):                       # no barrier between pushes: every rank races
            st.push_sgd(gk, -0.01)
            st.bump_step()
        torch.cuda.synchronize()
        dist.barrier()
        tot = gk.clone()
        dist.all_reduce(tot)
        ref = seed - 0.01 * K * tot
        got = torch.empty(total, device=dev)
        st.pull(got, None)
        torch.cuda.synchronize()
        err = rel_err(got, ref)
        steps = st.read_step()
        all_ok(err < 1e-5 and steps == K * world, {"rank": rank, "err": err, "steps": steps}, name)

    @case("hogwild_push_sum_sharded")
    def _(name):
        hogwild_case(name, False)

    @case("hogwild_push_sum_striped")
    def _(name):
        hogwild_case(name, True)

    @case("wgrad_gemm_push_into_peer_shard")
    def _(name):
        """The fused path itself: every rank's wgrad GEMM epilogue red.adds -lr*dW into a shard
        that lives on ANOTHER GPU (rank+1), all ranks at once."""
        M, N_, K = 512, 768, 1024
        buf = fabric.alloc("wgrad_target", M * N_ * 4)
        mine = buf.local(torch.float32, 0, M * N_)
        mine.fill_(1.0)
        torch.cuda.synchronize()
        dist.barrier()
        target = buf.tensor((rank + 1) % world, torch.float32, 0, M * N_).view(M, N_)
        A = torch.randn(K, M, device=dev, generator=g).bfloat16()        # dY^T given MN-major
        B = torch.randn(K, N_, device=dev, generator=g).bfloat16()
        ops.gemm(A, B, target, a_mn=True, b_mn=True, alpha=-0.01, mode=ops.EPI_F32_RED_COAL)
        torch.cuda.synchronize()
        dist.barrier()
        # what landed in MY buffer came from rank-1
        As, Bs = gather_all(A), gather_all(B)
        src = (rank - 1) % world
        ref = 1.0 - 0.01 * (As[src].float().t() @ Bs[src].float())
        err = rel_err(mine.view(M, N_), ref)
        all_ok(err < 1e-2, {"rank": rank, "err": err}, name)

    @case("adagrad_push_one_writer_at_a_time")
    def _(name):
        layout = FlatLayout([("w", (513, 1001), (513, 1001))], num_shards=1)
        st = PeerStore(fabric, layout, [0], adagrad=True)
        total = layout.total
        seed = torch.randn(total, device=dev, generator=torch.Generator(device=dev).manual_seed(12))
        if rank == 0:
            st.seed(seed)
        torch.cuda.synchronize()
        dist.barrier()
        gk = torch.randn(total, device=dev, generator=g)
        gs = gather_all(gk)
        for turn in range(world):                # sequential: the result has one right answer
            if turn == rank:
                st.push_adagrad(gk, 0.01)
                torch.cuda.synchronize()
            dist.barrier()
        ref, acc = seed.clone(), torch.full_like(seed, 0.1)
        for k in range(world):
            acc += gs[k] * gs[k]
            ref -= 0.01 * gs[k] / acc.sqrt()
        got = torch.empty(total, device=dev)
        st.pull(got, None)
        torch.cuda.synchronize()
        err = rel_err(got, ref)
        all_ok(err < 1e-5, {"rank": rank, "err": err}, name)

    # ---------------------------------------------------------------- a rank that never arrives
    @case("absent_rank_times_out_without_hanging")
    def _(name):
        old = fabric.timeout_cycles
        fabric.timeout_cycles = int(0.2 * 1.9e9)
        sy = new_sync(4096 * world, lr=0.05)
        fabric.timeout_cycles = old
        sy.load_master(torch.zeros(4096 * world, device=dev))
        sy.g32.fill_(1.0)
        torch.cuda.synchronize()
        dist.barrier()
        t0 = time.time()
        if rank != world - 1:                     # the last rank never launches its kernel
            sy.step()
        torch.cuda.synchronize()
        dt = time.time() - t0
        status = int(sy.status.item())
        ok = (status != 0 and dt < 10.0) if rank != world - 1 else True
        all_ok(ok, {"rank": rank, "status": status, "seconds": round(dt, 3)}, name)

    if rank == 0:
        print("RESULT " + json.dumps(results))
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
