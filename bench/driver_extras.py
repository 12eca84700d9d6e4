"""What bench.py adds to its JSON line at N >= 2, so that the real multi-GPU product runs under the driver's eyes:

correctness_matrix()  every dense collective in its one-shot / two-shot / NVLS forms for float and double, the
                      fused ThreadCommSlave collectives, allreduce_map, K12 (tcgen05 GEMM + reduce-scatter) and
                      K13 (GBDT histogram + reduce-scatter) on the REAL peer-mapped path, each against a
                      sequential torch reference that every rank computes from seeds. Returns pass / fail counts.
dense_sweep()         BASELINE.json's headline table: allreduce / reduce-scatter / allgather bus bandwidth (nccl-tests
                      convention) for float and double from 1 KiB to 1 GiB on symmetric tensors (zero copy), on
                      ordinary torch tensors (staged through the symmetric window) and with NCCL in the same
                      process on the same GPUs; plus the one-shot allreduce latency.

Device-timed (CUDA events on the launching stream, after warm-up), max over ranks through the library's own host
allreduce. Buffers of small sizes rotate through a pool larger than L2.
"""
import time

import torch

from ytk_mp4j_b200 import CommUtils, Operands, Operators
from ytk_mp4j_b200 import _native as N


def _seq_sum(parts):
    acc = parts[0].clone()
    for x in parts[1:]:
        acc = acc + x
    return acc


def correctness_matrix(pcs, plane, rank, world):
    """pcs: ProcessCommSlave with its GPU plane up. Every rank runs the same cases; a case passes only if it
    passes on every rank (host allreduce of the verdicts)."""
    results = []
    lop, LMIN = Operands.LONG_OPERAND(), Operators.Long.MIN

    def case(name, fn):
        ok = True
        try:
            ok = bool(fn())
            torch.cuda.synchronize()
            plane.check()
        except Exception as e:  # noqa: BLE001 - a failing case must not end the bench
            ok = False
            name = "%s (%s)" % (name, str(e)[:80])
        ok = bool(pcs.allreduce(1 if ok else 0, lop, LMIN))
        results.append((name, ok))

    def gen(seed, n, dtype):
        g = torch.Generator().manual_seed(seed)
        return torch.randn(n, generator=g, dtype=dtype)

    algos = [("oneshot", N.ALGO_ONESHOT), ("twoshot", N.ALGO_TWOSHOT)] + ([("nvls", N.ALGO_NVLS)] if plane.has_multicast else [])
    for dtype, operand, SUM in ((torch.float32, Operands.FLOAT_OPERAND(), Operators.Float.SUM),
                                (torch.float64, Operands.DOUBLE_OPERAND(), Operators.Double.SUM)):
        tname = "float" if dtype == torch.float32 else "double"
        for aname, algo in algos:
            for n, f in ((1000, 0), (50001, 3)) if aname == "oneshot" else ((50001, 3), ((2 << 20) + 7, 0)):
                def allreduce_case(n=n, f=f, algo=algo, dtype=dtype, operand=operand, SUM=SUM):
                    parts = [gen(1000 * q + n, n, dtype) for q in range(world)]
                    want = _seq_sum([p_[f:] for p_ in parts])  # ascending rank order, like the kernels
                    x = parts[rank].cuda()
                    plane.allreduce(x, operand, SUM, f, n, algo=algo)
                    got = x.cpu()
                    exact = torch.equal(got[f:], want)
                    # NVLS adds inside the switch (order not ours): tolerance instead of bits. A one-shot request
                    # beyond the slot size falls to AUTO, which may pick NVLS too.
                    close = torch.allclose(got[f:], want, rtol=1e-5 if dtype == torch.float32 else 1e-12, atol=1e-5)
                    return (exact or ("nvls" in plane.last() and close)) and torch.equal(got[:f], parts[rank][:f])
                case("allreduce %s %s n=%d" % (tname, aname, n), allreduce_case)
        # reduce-scatter / allgather / reduce / broadcast / gather / scatter, two-shot and NVLS
        n = 300007
        froms, tos = CommUtils.createProcessArrayFroms(n, world), CommUtils.createProcessArrayTos(n, world)
        counts = [t - f for f, t in zip(froms, tos)]
        root = world - 1
        for aname, algo in algos[1:]:
            def rs_case(algo=algo, dtype=dtype, operand=operand, SUM=SUM):
                parts = [gen(7 + q, n, dtype) for q in range(world)]
                want = _seq_sum(parts)
                x = parts[rank].cuda()
                plane.reduce_scatter(x, operand, SUM, 0, counts, algo=algo)
                got = x.cpu()[froms[rank]:tos[rank]]
                w = want[froms[rank]:tos[rank]]
                return torch.equal(got, w) or (algo == N.ALGO_NVLS and torch.allclose(got, w, rtol=1e-5, atol=1e-5))
            case("reduce_scatter %s %s" % (tname, aname), rs_case)

            def ag_case(algo=algo, dtype=dtype, operand=operand):
                parts = [gen(17 + q, n, dtype) for q in range(world)]
                want = torch.cat([parts[q][froms[q]:tos[q]] for q in range(world)])
                x = parts[rank].cuda()
                plane.allgather(x, operand, froms, tos, algo=algo)
                return torch.equal(x.cpu(), want)
            case("allgather %s %s" % (tname, aname), ag_case)

            def bcast_case(algo=algo, dtype=dtype, operand=operand):
                parts = [gen(27 + q, n, dtype) for q in range(world)]
                x = parts[rank].cuda()
                plane.broadcast(x, operand, 5, n - 5, root, algo=algo)
                want = parts[rank].clone()
                want[5:n - 5] = parts[root][5:n - 5]
                return torch.equal(x.cpu(), want)
            case("broadcast %s %s" % (tname, aname), bcast_case)

            def reduce_case(algo=algo, dtype=dtype, operand=operand, SUM=SUM):
                parts = [gen(37 + q, n, dtype) for q in range(world)]
                x = parts[rank].cuda()
                plane.reduce(x, operand, SUM, 0, n, root, algo=algo)
                if rank != root:
                    return True
                want = _seq_sum(parts)
                return torch.equal(x.cpu(), want) or (algo == N.ALGO_NVLS and torch.allclose(x.cpu(), want, rtol=1e-5, atol=1e-5))
            case("reduce %s %s" % (tname, aname), reduce_case)

        def gather_scatter_case(dtype=dtype, operand=operand):
            parts = [gen(47 + q, n, dtype) for q in range(world)]
            x = parts[rank].cuda()
            plane.gather(x, operand, froms, tos, root)
            ok = True
            if rank == root:
                ok = torch.equal(x.cpu(), torch.cat([parts[q][froms[q]:tos[q]] for q in range(world)]))
            y = parts[rank].cuda()
            plane.scatter(y, operand, froms, tos, 0)
            return ok and torch.equal(y.cpu()[froms[rank]:tos[rank]], parts[0][froms[rank]:tos[rank]])
        case("gather+scatter %s" % tname, gather_scatter_case)

    # ---- fused ThreadCommSlave collectives on the real path (T = 4 buffers per rank)
    T, n = 4, 200003
    fop, FSUM = Operands.FLOAT_OPERAND(), Operators.Float.SUM

    def thread_parts():
        return [[gen(100 * q + t, n, torch.float32) for t in range(T)] for q in range(world)]

    def thread_total(parts):
        return _seq_sum([_seq_sum(row) for row in parts])

    def thread_allreduce_case():
        parts = thread_parts()
        bufs = [x.cuda() for x in parts[rank]]
        plane.thread_allreduce(bufs, fop, FSUM, 0, n - 3)  # odd length: generic fallback kernel
        want = thread_total(parts)[:n - 3]
        nvls = "nvls" in plane.last()  # the switch adds in its own order: tolerance instead of bits
        return all(torch.equal(b.cpu()[:n - 3], want) or (nvls and torch.allclose(b.cpu()[:n - 3], want, rtol=1e-5, atol=1e-5))
                   for b in bufs)
    case("thread allreduce (fallback kernel)", thread_allreduce_case)

    def thread_allreduce_fused_case():
        m = 3 << 20
        parts = [[gen(100 * q + t + 9, m, torch.float32) for t in range(T)] for q in range(world)]
        bufs = [x.cuda() for x in parts[rank]]
        plane.thread_allreduce(bufs, fop, FSUM, 0, m)
        got = bufs[T - 1].cpu()
        want = thread_total(parts)
        fused = "fused-" in plane.last()
        exact = torch.equal(got, want)
        return fused and (exact or ("nvls" in plane.last() and torch.allclose(got, want, rtol=1e-5, atol=1e-5)))
    case("thread allreduce (fused kernel)", thread_allreduce_fused_case)

    w = (n - 10) // (world * T)
    slices = [(q, t, 5 + (q * T + t) * w, 5 + (q * T + t + 1) * w - (q + t) % 2) for q in range(world) for t in range(T)]

    def thread_rs_case():
        parts = thread_parts()
        bufs = [x.cuda() for x in parts[rank]]
        plane.thread_segments("thread_reduce_scatter", bufs, fop, FSUM, [(lo, hi, -1, -1, q, t) for q, t, lo, hi in slices])
        want = thread_total(parts)
        return all(torch.equal(bufs[t].cpu()[lo:hi], want[lo:hi]) for q, t, lo, hi in slices if q == rank)
    case("thread reduce_scatter (one launch)", thread_rs_case)

    def thread_ag_case():
        parts = thread_parts()
        bufs = [x.cuda() for x in parts[rank]]
        plane.thread_segments("thread_allgather", bufs, fop, None, [(lo, hi, q, t, -1, -1) for q, t, lo, hi in slices])
        return all(torch.equal(b.cpu()[lo:hi], parts[q][t][lo:hi]) for b in bufs for q, t, lo, hi in slices)
    case("thread allgather (one launch)", thread_ag_case)

    def thread_bcast_reduce_case():
        parts = thread_parts()
        bufs = [x.cuda() for x in parts[rank]]
        plane.thread_segments("thread_broadcast", bufs, fop, None, [(3, n - 3, world - 1, 1, -1, -1)])
        ok = all(torch.equal(b.cpu()[3:n - 3], parts[world - 1][1][3:n - 3]) for b in bufs)
        bufs = [x.cuda() for x in parts[rank]]
        plane.thread_segments("thread_reduce", bufs, fop, FSUM, [(0, n, -1, -1, 0, T - 1)])
        if rank == 0:
            ok = ok and torch.equal(bufs[T - 1].cpu(), thread_total(parts))
        return ok
    case("thread broadcast + reduce (one launch each)", thread_bcast_reduce_case)

    # ---- sparse maps
    def map_case():
        big = 100000
        sets = [torch.randperm(400000, generator=torch.Generator().manual_seed(100 + q))[:big] for q in range(world)]
        cnt = torch.zeros(400000, dtype=torch.float32)
        for sq in sets:
            cnt[sq] += 1.0
        k, v = plane.allreduce_map(sets[rank].to(torch.int64).cuda(), torch.ones(big, device="cuda"), FSUM)
        return k.numel() == int((cnt > 0).sum()) and torch.equal(v.cpu(), cnt[k.cpu()])
    case("allreduce_map", map_case)

    # ---- K12 and K13
    def k12_case():
        F, C, K = 1024, 128, 2048
        g = torch.Generator().manual_seed(5)
        xts = [(torch.randn(F, K, generator=g) * 0.5).to(torch.bfloat16) for _ in range(world)]
        rts = [(torch.randn(C, K, generator=g) * 0.5).to(torch.bfloat16) for _ in range(world)]
        shard, fpr = plane.lr_xtr_reduce_scatter(xts[rank].cuda(), rts[rank].cuda())
        torch.cuda.synchronize()
        want = sum(x.float() @ r.float().t() for x, r in zip(xts, rts))
        lo, hi = rank * fpr, min(F, (rank + 1) * fpr)
        err = (shard[:hi - lo].cpu() - want[lo:hi]).abs().max().item()
        return err <= 2e-3 * want.abs().max().item() + 1e-3
    case("K12 tcgen05 X^T R + reduce-scatter", k12_case)

    def k13_case():
        F, rows = 64, 20000
        g = torch.Generator().manual_seed(9)
        want = torch.zeros(F, 256, 2, dtype=torch.float64)
        mine = None
        for q in range(world):
            b = torch.randint(0, 256, (F, rows), generator=g, dtype=torch.uint8)
            gr = torch.randint(-8, 8, (rows,), generator=g).float()
            he = torch.randint(0, 4, (rows,), generator=g).float()
            idx = b.long()
            for f in range(F):
                want[f, :, 0].index_add_(0, idx[f], gr.double())
                want[f, :, 1].index_add_(0, idx[f], he.double())
            if q == rank:
                mine = (b, gr, he)
        shard, fpr = plane.gbdt_hist_reduce_scatter(mine[0].cuda(), mine[1].cuda(), mine[2].cuda())
        torch.cuda.synchronize()
        lo, hi = rank * fpr, min(F, (rank + 1) * fpr)
        return hi <= lo or torch.equal(shard[:hi - lo].cpu().double(), want[lo:hi])
    case("K13 GBDT histogram + reduce-scatter", k13_case)

    failed = [nm for nm, ok in results if not ok]
    return {"cases": len(results), "passed": len(results) - len(failed), "failed": failed}


def dense_sweep(pcs, plane, rank, world, max_bytes=1 << 30, with_nccl=True, budget_s=120.0):
    """Bus bandwidth table of BASELINE.json's headline metric. Returns {"sizes": [...], "float": {...}, "double": {...}}
    with, per collective, GB/s lists for ours on symmetric tensors (zero copy, AUTO), ours on ordinary tensors
    (staged), and NCCL; entries are None where the time budget ran out."""
    import torch.distributed as dist
    t_start = time.time()
    dop, DMAX = Operands.DOUBLE_OPERAND(), Operators.Double.MAX
    nccl_ok = False
    if with_nccl:
        try:
            if not dist.is_initialized():
                dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", torch.cuda.current_device()))
            nccl_ok = True
        except Exception:  # noqa: BLE001
            nccl_ok = False
    sizes = []
    s = 1024
    while s <= max_bytes:
        sizes.append(s)
        if s == (256 << 20) and max_bytes >= (512 << 20):
            sizes.append(512 << 20)
        s *= 4
    pool_bytes = max(512 << 20, max_bytes)
    out = {"sizes_bytes": sizes, "convention": "nccl-tests bus bandwidth, GB/s; allreduce x2(p-1)/p, others x(p-1)/p",
           "nccl": "torch.distributed NCCL in the same process" if nccl_ok else "unavailable"}

    def timed(fn, nbytes):
        iters = 20 if nbytes <= (16 << 20) else (8 if nbytes <= (256 << 20) else 4)
        warm = 3 if nbytes <= (16 << 20) else 1
        nreg = max(1, min(64, (256 << 20) // max(nbytes, 1)))
        for i in range(warm):
            fn(i % nreg)
        torch.cuda.synchronize()
        pcs.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(iters):
            fn(i % nreg)
        e1.record()
        torch.cuda.synchronize()
        return pcs.allreduce(e0.elapsed_time(e1) * 1e3 / iters, dop, DMAX)  # us, max over ranks

    for tname, tdt, es, operand, SUM in (("float", torch.float32, 4, Operands.FLOAT_OPERAND(), Operators.Float.SUM),
                                         ("double", torch.float64, 8, Operands.DOUBLE_OPERAND(), Operators.Double.SUM)):
        sym = pcs.empty(pool_bytes // es, operand)
        sym.fill_(1e-30)
        plain = torch.full((pool_bytes // es,), 1e-30, dtype=tdt, device="cuda")
        table = {c: {"symmetric": [], "ordinary": [], "nccl": [], "picked": []} for c in ("allreduce", "reduce_scatter", "allgather")}
        table["allreduce_oneshot_latency_us"] = None
        for size in sizes:
            n = size // es
            counts = [n // world] * (world - 1) + [n - (n // world) * (world - 1)]
            froms = [sum(counts[:q]) for q in range(world)]
            tos = [f + c for f, c in zip(froms, counts)]
            stride = ((n + 63) // 64) * 64
            m = n // world
            over = (time.time() - t_start) > budget_s
            over = bool(pcs.allreduce(1.0 if over else 0.0, dop, DMAX))  # same decision on every rank
            for coll in ("allreduce", "reduce_scatter", "allgather"):
                factor = (2 * (world - 1) / world) if coll == "allreduce" else (world - 1) / world

                def ours(pool):
                    def fn(i):
                        x = pool[i * stride:i * stride + n]
                        if coll == "allreduce":
                            plane.allreduce(x, operand, SUM, 0, n)
                        elif coll == "reduce_scatter":
                            plane.reduce_scatter(x, operand, SUM, 0, counts)
                        else:
                            plane.allgather(x, operand, froms, tos)
                    return fn

                def nccl(i):
                    x = plain[i * stride:i * stride + n]
                    if coll == "allreduce":
                        dist.all_reduce(x)
                    elif coll == "reduce_scatter":
                        dist.reduce_scatter_tensor(x[rank * m:(rank + 1) * m], x[:m * world])
                    else:
                        dist.all_gather_into_tensor(x[:m * world], x[rank * m:(rank + 1) * m])
                row = table[coll]
                if over or n < world:
                    for k in ("symmetric", "ordinary", "nccl", "picked"):
                        row[k].append(None)
                    continue
                us = timed(ours(sym), size)
                row["symmetric"].append(round(size / us / 1e3 * factor, 2))
                row["picked"].append(plane.last().split("algo=")[1].split(" ")[0])
                if coll == "allreduce" and size == (32 << 10):
                    pass
                us = timed(ours(plain), size)
                row["ordinary"].append(round(size / us / 1e3 * factor, 2))
                if nccl_ok and m > 0:
                    us = timed(nccl, size)
                    row["nccl"].append(round(size / us / 1e3 * factor, 2))
                else:
                    row["nccl"].append(None)
            if size == 4096 and not over:
                n4 = 4096 // es

                def one(i):
                    plane.allreduce(sym[i * 1024:i * 1024 + n4], operand, SUM, 0, n4, algo=N.ALGO_ONESHOT)
                lat = {"ours_oneshot_4KiB": round(timed(one, 4096), 2)}
                if nccl_ok:
                    lat["nccl_4KiB"] = round(timed(lambda i: dist.all_reduce(plain[i * 1024:i * 1024 + n4]), 4096), 2)
                table["allreduce_oneshot_latency_us"] = lat
        out[tname] = table
        plane.free(sym)
        del plain
        torch.cuda.empty_cache()
    plane.check()
    return out


def consumers(pcs, plane, rank, world, steps=8):
    """BASELINE.json's remaining metrics ("sparse map keys/s; LR step/s") and configs 4-5, through the models
    that ship with the package (ytk_mp4j_b200/models): device-timed whole steps, max over ranks.
      dense LR   : 65536 rows x 8192 features x 128 classes per GPU, bf16; step = forward GEMM, softmax residual,
                   fused tcgen05 X^T R + reduce-scatter (K12), sharded update, allgather of the weights
      sparse LR  : 10M hashed features, 262144 rows x 8 non-zeros per GPU and step; the gradient travels as a map
                   through map_reduce_local + allreduce_map (K10/K11)
      GBDT       : 1M rows (split over the GPUs) x 2048 features x 256 bins: histogram build fused with the
                   reduce-scatter (K13) + split scan + MAX_LOC allreduce
    """
    from ytk_mp4j_b200.models.gbdt import HistGBDT
    from ytk_mp4j_b200.models.lr import DenseLR, SparseLR
    dop, DMAX = Operands.DOUBLE_OPERAND(), Operators.Double.MAX
    out = {}

    def timed(fn, k=steps, warm=2):
        for _ in range(warm):
            fn()
        torch.cuda.synchronize()
        pcs.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(k):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return pcs.allreduce(e0.elapsed_time(e1) / k, dop, DMAX)  # ms per step

    g = torch.Generator(device="cuda").manual_seed(rank)
    # ---- dense LR
    F, Cn, rows = 8192, 128, 65536
    model = DenseLR(pcs, F, Cn, lr=0.1)
    X = torch.randn(rows, F, device="cuda", generator=g)
    y = torch.randint(0, Cn, (rows,), device="cuda", generator=g)
    model.set_data(X, y)
    del X
    ms = timed(model.step)
    out["dense_lr"] = {"steps_per_s": round(1e3 / ms, 2), "ms_per_step": round(ms, 3),
                       "config": "%d rows x %d features x %d classes per GPU, bf16, N=%d" % (rows, F, Cn, world)}
    del model
    torch.cuda.empty_cache()
    # ---- sparse LR + the map collective alone
    Fs, n, nnz = 10_000_000, 262144, 8
    sm = SparseLR(pcs, Fs, lr=0.1, max_nnz=n * nnz, rows_total=n * world)
    cols = torch.randint(0, Fs, (n * nnz,), device="cuda", generator=g)
    rws = torch.arange(n, device="cuda").repeat_interleave(nnz)
    vals = torch.ones(n * nnz, device="cuda")
    ys = (torch.rand(n, device="cuda", generator=g) > 0.5).float()
    ms = timed(lambda: sm.step(rws, cols, vals, ys), k=max(3, steps // 2))
    out["sparse_lr"] = {"steps_per_s": round(1e3 / ms, 2), "ms_per_step": round(ms, 3),
                        "config": "10M hashed features, %d rows x %d nnz per GPU and step, N=%d" % (n, nnz, world)}
    keys = cols[: 2_000_000].contiguous()
    kv = torch.ones(2_000_000, device="cuda")
    if world > 1:
        ms = timed(lambda: plane.allreduce_map(keys, kv, Operators.Float.SUM, max_n=2_000_000,
                                               out_cap=min(2_000_000 * world, 10_000_000), copy=False),
                   k=max(3, steps // 2))
        out["allreduce_map"] = {"Gkeys_per_s": round(world * 2_000_000 / (ms * 1e-3) / 1e9, 2), "ms": round(ms, 3),
                                "config": "2M (int64 key, fp32) pairs per GPU drawn from 10M ids, N=%d" % world}
    else:
        ms = timed(lambda: plane.map_reduce_local(keys, kv, N.SUM), k=max(3, steps // 2))
        out["map_reduce_local"] = {"Gkeys_per_s": round(2_000_000 / (ms * 1e-3) / 1e9, 2), "ms": round(ms, 3),
                                   "config": "2M (int64 key, fp32) pairs drawn from 10M ids, one GPU"}
    del sm, cols, rws, vals
    torch.cuda.empty_cache()
    # ---- GBDT split search
    Fg, rows_g = 2048, 1_000_000 // world
    gb = HistGBDT(pcs, Fg)
    gb.set_data(torch.randint(0, 256, (Fg, rows_g), device="cuda", generator=g, dtype=torch.uint8))
    gr = torch.randn(rows_g, device="cuda", generator=g)
    he = torch.rand(rows_g, device="cuda", generator=g) + 0.1
    ms = timed(lambda: gb.best_split(gr, he), k=max(3, steps // 2))
    out["gbdt_best_split"] = {"ms": round(ms, 3), "splits_per_s": round(1e3 / ms, 1),
                              "config": "1M rows / %d GPUs x 2048 features x 256 bins" % world}
    del gb
    torch.cuda.empty_cache()
    plane.check()
    return out
