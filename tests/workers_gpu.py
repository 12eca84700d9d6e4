"""GPU worker bodies for tests/test_gpu_multiproc.py (public API end to end on CUDA tensors)."""
import numpy as np
import torch

from ytk_mp4j_b200 import CommUtils, Operands, Operators
from ytk_mp4j_b200 import _native as N


def process_gpu(comm):
    p, r = comm.getSlaveNum(), comm.getRank()
    torch.cuda.set_device(r % torch.cuda.device_count())
    comm.init_gpu(device=r % torch.cuda.device_count(), heap_bytes=512 << 20, staging_bytes=32 << 20)
    info = {"desc": comm.gpu.describe(), "ok": False}
    fop, dop, iop = Operands.FLOAT_OPERAND(), Operands.DOUBLE_OPERAND(), Operands.INT_OPERAND()
    for n in (5, 1000, 100003, (4 << 20) + 3):
        for algo in (N.ALGO_AUTO, N.ALGO_ONESHOT, N.ALGO_TWOSHOT) + ((N.ALGO_NVLS,) if comm.gpu.has_multicast else ()):
            x = torch.full((n,), float(r + 1), device="cuda")
            comm.gpu.allreduce(x, fop, Operators.Float.SUM, 1, n, algo=algo)
            torch.cuda.synchronize()
            assert torch.all(x[1:] == p * (p + 1) / 2) and x[0] == r + 1, (n, algo, comm.gpu.last())
    # zero-copy symmetric tensor
    n = 1 << 20
    x = comm.empty(n, dop)
    x.fill_(r)
    comm.allreduceArray(x, dop, Operators.Double.MAX, 0, n)
    torch.cuda.synchronize()
    assert torch.all(x == p - 1) and "staged=0" in comm.gpu.last()
    # the other six collectives, reference value patterns
    n = 100003
    froms, tos = CommUtils.createProcessArrayFroms(n, p), CommUtils.createProcessArrayTos(n, p)
    owner = torch.empty(n, dtype=torch.int32)
    for q in range(p):
        owner[froms[q]:tos[q]] = q
    owner = owner.cuda()
    root = p - 1
    x = torch.full((n,), -1, dtype=torch.int32, device="cuda")
    x[froms[r]:tos[r]] = r
    comm.allgatherArray(x, iop, froms, tos)
    assert torch.equal(x, owner)
    x = torch.full((n,), -1, dtype=torch.int32, device="cuda")
    x[froms[r]:tos[r]] = r
    comm.gatherArray(x, iop, froms, tos, root)
    if r == root:
        assert torch.equal(x, owner)
    x = owner.clone() if r == root else torch.full((n,), -1, dtype=torch.int32, device="cuda")
    comm.scatterArray(x, iop, froms, tos, root)
    assert torch.all(x[froms[r]:tos[r]] == r)
    x = owner.clone()
    comm.reduceScatterArray(x, iop, Operators.Int.SUM, 0, [t - f for f, t in zip(froms, tos)])
    assert torch.all(x[froms[r]:tos[r]] == r * p)
    x = torch.full((n,), 1 if r == root else -1, dtype=torch.int32, device="cuda")
    comm.broadcastArray(x, iop, 0, n, root)
    assert torch.all(x == 1)
    x = torch.ones(n, dtype=torch.int32, device="cuda")
    comm.reduceArray(x, iop, Operators.Int.SUM, 0, n, root)
    if r == root:
        assert torch.all(x == p)
    x = torch.ones(7, dtype=torch.float32, device="cuda")
    comm.allreduceArrayRpc(x, fop, Operators.Float.SUM)
    assert torch.all(x == p)
    # custom device operator through the public API: NVRTC-compiled into the same fused kernel,
    # evaluated in ascending rank order on every rank (non-commutative on purpose)
    halfsum = Operators.device_expr("a * 0.5f + b", host_fn=lambda a, b: a * 0.5 + b)
    x = torch.full((100003,), float(r + 1), device="cuda")
    comm.allreduceArray(x, fop, halfsum, 0, 100003)
    want = np.float32(1.0)
    for q in range(1, p):
        want = np.float32(want * np.float32(0.5)) + np.float32(q + 1)
    assert torch.all(x == float(want)), (x[:3], want)
    # sparse map allreduce on tensors: shared keys + one private key per rank
    # (ProcessAllReduceCheck.java:86-116), twice to exercise buffer reuse
    obj = 20000
    for rep in range(2):
        keys = torch.cat([torch.arange(obj, dtype=torch.int64), torch.tensor([-(r + 1)], dtype=torch.int64)]).cuda()
        vals = torch.ones(obj + 1, dtype=torch.float32, device="cuda")
        k, v = comm.gpu.allreduce_map(keys, vals, Operators.Float.SUM)
        assert k.numel() == obj + p
        d = dict(zip(k.tolist(), v.tolist()))
        assert all(d[i] == p for i in range(0, obj, 97)) and all(d[-(q + 1)] == 1 for q in range(p))
    # larger, randomly overlapping key sets: forces buffer re-creation (the allocation race found at
    # 8 GPUs) and checks the result against a reference every rank can compute from the seeds
    big = 300000
    sets = [torch.randperm(1000000, generator=torch.Generator().manual_seed(100 + q))[:big] for q in range(p)]
    cnt = torch.zeros(1000000, dtype=torch.float32)
    for sq in sets:
        cnt[sq] += 1.0
    k, v = comm.gpu.allreduce_map(sets[r].to(torch.int64).cuda(), torch.ones(big, device="cuda"), Operators.Float.SUM)
    assert k.numel() == int((cnt > 0).sum()), (k.numel(), int((cnt > 0).sum()))
    assert torch.equal(v.cpu(), cnt[k.cpu()])
    k, v = comm.gpu.reduce_scatter_map(keys, vals, Operators.Float.SUM)
    tot = comm.allreduce(int(k.numel()), Operands.LONG_OPERAND(), Operators.Long.SUM)
    assert tot == obj + p
    k, v = comm.gpu.allgather_map(torch.full((3,), r, dtype=torch.int64, device="cuda"),
                                  torch.full((3,), float(r), device="cuda"))
    assert k.tolist() == [q for q in range(p) for _ in range(3)]
    # the same results when the gather step goes through the multicast mapping (one multimem.st lands in every
    # rank's output; default only beyond 2 ranks): unaligned offsets (3 and big + odd counts), 4- and 8-byte values
    comm.gpu.set_option("map_nvls", 2)
    for dt in (torch.float32, torch.float64):
        k, v = comm.gpu.allreduce_map(sets[r].to(torch.int64).cuda(), torch.ones(big, device="cuda", dtype=dt),
                                      Operators.Float.SUM if dt == torch.float32 else Operators.Double.SUM)
        assert k.numel() == int((cnt > 0).sum()), (k.numel(), int((cnt > 0).sum()))
        assert torch.equal(v.cpu().float(), cnt[k.cpu()])
        k, v = comm.gpu.allgather_map(torch.full((3 + r,), r, dtype=torch.int64, device="cuda"),
                                      torch.full((3 + r,), float(r), device="cuda", dtype=dt))
        assert k.tolist() == [q for q in range(p) for _ in range(3 + q)] and v.tolist() == [float(x) for x in k.tolist()]
    comm.gpu.set_option("map_nvls", 1)
    # ---- Map<String, number> through the PUBLIC API with the GPU plane up: String keys are bridged to 64-bit
    # keys (comm/map_bridge.py), values reduce on the GPU, result comes back as a dict (PCS:2053, 1299, 666)
    objn = 3000
    for rep in range(2):  # second call: every string is already known, no string traffic
        m = {"f%d" % i: 1.0 for i in range(objn)}
        m["only-%d" % r] = float(r + 1)
        out = comm.allreduceMap(m, dop, Operators.Double.SUM)
        assert "allreduce_map" in comm.gpu.last(), comm.gpu.last()
        assert len(out) == objn + p and all(out["f%d" % i] == p for i in range(0, objn, 37))
        assert all(out["only-%d" % q] == q + 1 for q in range(p))
    lists = [{"k%d" % j: 1 for j in range(q * 100, (q + 1) * 100)} for q in range(p)]
    out = comm.reduceScatterMap(lists, Operands.LONG_OPERAND(), Operators.Long.SUM)
    assert "reduce_scatter_map" in comm.gpu.last(), comm.gpu.last()
    assert sorted(out) == sorted("k%d" % j for j in range(r * 100, (r + 1) * 100)) and all(v == p for v in out.values())
    outs = comm.allgatherMap({"g%d" % j: float(j) for j in range(r * 10, (r + 1) * 10)}, fop)
    assert len(outs) == p and all(outs[q] == {"g%d" % j: float(j) for j in range(q * 10, (q + 1) * 10)} for q in range(p))
    torch.cuda.synchronize()
    comm.gpu.check()
    comm.barrier()
    info["ok"] = True
    return info


def thread_gpu(comm, tid):
    p, r, T = comm.getSlaveNum(), comm.getRank(), comm.getThreadNum()
    dev = r % torch.cuda.device_count()
    torch.cuda.set_device(dev)
    if tid == 0:
        comm.init_gpu(device=dev, heap_bytes=512 << 20, staging_bytes=32 << 20)
    comm.threadBarrier()
    fop = Operands.FLOAT_OPERAND()
    for n in (1000, (2 << 20) + 4):
        x = torch.full((n,), float(r * T + tid + 1), device="cuda")
        torch.cuda.synchronize()
        comm.allreduceArray(x, fop, Operators.Float.SUM, 0, n)
        torch.cuda.synchronize()
        total = p * T * (p * T + 1) / 2
        assert torch.all(x == total), (n, x[:4])
    # custom device operator on the thread path: thread-id-ordered fold, inter-GPU kernel, fan-out
    halfsum = Operators.device_expr("a * 0.5f + b")
    x = torch.full((5001,), float(r * T + tid + 1), device="cuda")
    torch.cuda.synchronize()
    comm.allreduceArray(x, fop, halfsum, 0, 5001)
    torch.cuda.synchronize()
    per_rank = []
    for q in range(p):
        acc = np.float32(q * T + 1)
        for t in range(1, T):
            acc = np.float32(acc * np.float32(0.5)) + np.float32(q * T + t + 1)
        per_rank.append(acc)
    want = per_rank[0]
    for q in range(1, p):
        want = np.float32(want * np.float32(0.5)) + per_rank[q]
    assert torch.all(x == float(want)), (x[:3], want)
    # reduce to (root, rootThread) through the generic thread path on CUDA tensors
    n = 4099
    x = torch.ones(n, device="cuda")
    torch.cuda.synchronize()
    comm.reduceArray(x, fop, Operators.Float.SUM, 0, n, p - 1, T - 1)
    torch.cuda.synchronize()
    if r == p - 1 and tid == T - 1:
        assert torch.all(x == p * T)
    # ---- the other thread collectives on CUDA tensors: ONE fused launch each (thread_seg.cu). Position-
    # dependent values, so a slice that lands in the wrong place or comes from the wrong thread shows.
    n, P, me = 50021, p * T, r * T + tid
    w = (n - 3) // P
    fr = [[(q * T + t) * w + 3 for t in range(T)] for q in range(p)]
    to = [[(q * T + t + 1) * w + 3 - (1 if (q + t) % 2 else 0) for t in range(T)] for q in range(p)]
    idx = torch.arange(n, device="cuda", dtype=torch.float32)

    def pat(k):
        return idx % 7 + 1000.0 * k

    def sync_call(fn, *a):
        torch.cuda.synchronize()
        fn(*a)
        torch.cuda.synchronize()
    x = torch.full((n,), -1.0, device="cuda")
    x[fr[r][tid]:to[r][tid]] = pat(me)[fr[r][tid]:to[r][tid]]
    sync_call(comm.allgatherArray, x, fop, fr, to)
    for q in range(p):
        for t in range(T):
            assert torch.equal(x[fr[q][t]:to[q][t]], pat(q * T + t)[fr[q][t]:to[q][t]]), ("allgather", q, t)
    if tid == 0:
        assert "fused-segments" in comm.gpu.last(), comm.gpu.last()
    root = (p - 1, T - 1)
    x = torch.full((n,), -1.0, device="cuda")
    x[fr[r][tid]:to[r][tid]] = pat(me)[fr[r][tid]:to[r][tid]]
    sync_call(comm.gatherArray, x, fop, fr, to, root[0], root[1])
    if (r, tid) == root:
        for q in range(p):
            for t in range(T):
                assert torch.equal(x[fr[q][t]:to[q][t]], pat(q * T + t)[fr[q][t]:to[q][t]]), ("gather", q, t)
    root = (0, T - 1)
    x = pat(77) if (r, tid) == root else torch.full((n,), -1.0, device="cuda")
    sync_call(comm.scatterArray, x, fop, fr, to, root[0], root[1])
    assert torch.equal(x[fr[r][tid]:to[r][tid]], pat(77)[fr[r][tid]:to[r][tid]]), "scatter"
    if (r, tid) != root:
        assert torch.all(x[:fr[r][tid]] == -1.0) and torch.all(x[to[r][tid]:] == -1.0)
    root = (p - 1, 0)
    x = pat(5) if (r, tid) == root else torch.full((n,), -1.0, device="cuda")
    sync_call(comm.broadcastArray, x, fop, 7, n - 9, root[0], root[1])
    assert torch.equal(x[7:n - 9], pat(5)[7:n - 9]), "broadcast"
    counts = [[to[q][t] - fr[q][t] for t in range(T)] for q in range(p)]
    x = pat(me)
    sync_call(comm.reduceScatterArray, x, fop, Operators.Float.SUM, 3, counts)
    off = 3 + sum(sum(c) for c in counts[:r]) + sum(counts[r][:tid])
    want = (idx % 7) * P + 1000.0 * (P * (P - 1) / 2)
    assert torch.equal(x[off:off + counts[r][tid]], want[off:off + counts[r][tid]]), "reduce_scatter"
    # ---- thread-level allreduceMap: the T thread maps of every process go to the GPU unmerged
    m = {"w%d" % i: 1 for i in range(500)}
    m["t-%d" % me] = me + 1
    out = comm.allreduceMap(m, Operands.LONG_OPERAND(), Operators.Long.SUM)
    assert len(out) == 500 + P and out["w7"] == P and all(out["t-%d" % g] == g + 1 for g in range(P))
    if tid == 0 and p > 1:
        assert "allreduce_map" in comm.gpu.last(), comm.gpu.last()
    # ---- pinned host tensors: streamed through the GPU plane (H2D || fused kernel || D2H pipeline)
    n = (3 << 20) + 5
    h = torch.full((n,), float(me + 1)).pin_memory()
    comm.allreduceArray(h, fop, Operators.Float.SUM, 0, n)
    assert torch.all(h == P * (P + 1) / 2), h[:4]
    if tid == 0:
        assert "host-pipeline" in comm.gpu.last(), comm.gpu.last()
    comm.barrier()
    return True


def check_matrix_cuda(comm, tid=None):
    """The CommCheckTool matrix with device=cuda: every array collective x every primitive type on CUDA tensors,
    process level (tid None) or thread level."""
    from ytk_mp4j_b200.check import checks
    r = comm.getRank()
    dev = r % torch.cuda.device_count()
    torch.cuda.set_device(dev)
    if tid is None or tid == 0:
        comm.init_gpu(device=dev, heap_bytes=512 << 20, staging_bytes=32 << 20)
    if tid is not None:
        comm.threadBarrier()
    checks.run_all(comm, tid, 20011, 50, 1, False, False, device="cuda")
    torch.cuda.synchronize()
    if tid is None or tid == 0:
        comm.gpu.check()
    return True


def models_worker(comm):
    """The two ML consumers end to end on the GPU plane (reference README.md:268-278): data-parallel logistic
    regression (dense: fused X^T R GEMM + reduce-scatter; sparse: gradient as a map through allreduce_map) against
    a single-process fp32 PyTorch run on the concatenated data, and the GBDT split search against a CPU scan."""
    from ytk_mp4j_b200.models.gbdt import HistGBDT
    from ytk_mp4j_b200.models.lr import DenseLR, SparseLR
    p, r = comm.getSlaveNum(), comm.getRank()
    dev = r % torch.cuda.device_count()
    torch.cuda.set_device(dev)
    comm.init_gpu(device=dev, heap_bytes=512 << 20, staging_bytes=32 << 20)
    out = {}
    # ---- dense LR: every rank can rebuild every rank's shard from the seeds
    F, Cn, rows, steps, lr = 256, 16, 2048, 12, 0.5
    g = torch.Generator().manual_seed(7)
    Wtrue = torch.randn(F, Cn, generator=g)
    Xs = [torch.randn(rows, F, generator=g) for _ in range(p)]
    ys = [(X @ Wtrue + 0.1 * torch.randn(rows, Cn, generator=g)).argmax(1) for X in Xs]
    model = DenseLR(comm, F, Cn, lr=lr)
    model.set_data(Xs[r].cuda(), ys[r].cuda())
    ours = []
    for _ in range(steps):
        ours.append(model.loss(model.step()))
    # reference: full-batch gradient descent in fp32 on all the data (same bf16-rounded inputs)
    Xall = torch.cat(Xs).to(torch.bfloat16).float()
    yall = torch.cat(ys)
    W = torch.zeros(F, Cn)
    ref = []
    for _ in range(steps):
        logp = torch.log_softmax(Xall @ W, 1)
        ref.append(float(-logp.gather(1, yall[:, None]).sum() / yall.numel()))
        R = logp.exp()
        R[torch.arange(yall.numel()), yall] -= 1.0
        W -= lr / yall.numel() * (Xall.t() @ R)
    out["dense_ours"], out["dense_ref"] = ours, ref
    out["dense_ok"] = all(abs(a - b) <= 0.02 * abs(b) + 1e-3 for a, b in zip(ours, ref)) and ours[-1] < 0.8 * ours[0]
    # ---- sparse LR: binary, hashed features, gradient through allreduce_map
    Fs, n, nnz_row, steps = 5000, 3000, 8, 8
    data = []
    for q in range(p):
        gq = torch.Generator().manual_seed(100 + q)
        cols = torch.randint(0, Fs, (n * nnz_row,), generator=gq)
        rws = torch.arange(n).repeat_interleave(nnz_row)
        vals = torch.ones(n * nnz_row)
        wtrue = torch.randn(Fs, generator=torch.Generator().manual_seed(5))
        y = (torch.zeros(n).index_add_(0, rws, vals * wtrue[cols]) > 0).float()
        data.append((rws, cols, vals, y))
    # odd ranks tell the model its sizes up front? no: the sizes must agree, so all ranks do, every other test run
    sm = SparseLR(comm, Fs, lr=0.5, max_nnz=n * nnz_row, rows_total=n * p) if p % 2 == 0 else SparseLR(comm, Fs, lr=0.5)
    rws, cols, vals, y = (t.cuda() for t in data[r])
    s_ours = []
    for _ in range(steps):
        ls = sm.step(rws, cols, vals, y)
        s_ours.append(comm.allreduce(float(ls), Operands.DOUBLE_OPERAND(), Operators.Double.SUM) / (n * p))
    w = torch.zeros(Fs)
    s_ref = []
    for _ in range(steps):
        gsum, lsum = torch.zeros(Fs), 0.0
        for rws_, cols_, vals_, y_ in data:
            z = torch.zeros(n).index_add_(0, rws_, vals_ * w[cols_])
            lsum += float(torch.nn.functional.binary_cross_entropy_with_logits(z, y_, reduction="sum"))
            gsum.index_add_(0, cols_, vals_ * (torch.sigmoid(z) - y_)[rws_])
        s_ref.append(lsum / (n * p))
        w -= 0.5 / (n * p) * gsum
    out["sparse_ours"], out["sparse_ref"] = s_ours, s_ref
    out["sparse_ok"] = all(abs(a - b) <= 1e-4 * abs(b) + 1e-5 for a, b in zip(s_ours, s_ref)) and s_ours[-1] < s_ours[0]
    # ---- GBDT: best split of one node over all ranks' rows, real-valued gradients
    Fg, rws_n = 24, 5000
    bins_all, g_all, h_all = [], [], []
    for q in range(p):
        gq = torch.Generator().manual_seed(200 + q)
        bins_all.append(torch.randint(0, 256, (Fg, rws_n), generator=gq, dtype=torch.uint8))
        g_all.append(torch.randn(rws_n, generator=gq))
        h_all.append(torch.rand(rws_n, generator=gq) + 0.1)
    gb = HistGBDT(comm, Fg)
    gb.set_data(bins_all[r].cuda())
    gain, feat, b = gb.best_split(g_all[r].cuda(), h_all[r].cuda())
    hist = torch.zeros(Fg, 256, 2, dtype=torch.float64)
    for q in range(p):
        idx = bins_all[q].long()
        for f in range(Fg):
            hist[f, :, 0].index_add_(0, idx[f], g_all[q].double())
            hist[f, :, 1].index_add_(0, idx[f], h_all[q].double())
    gl, hl = hist[..., 0].cumsum(1), hist[..., 1].cumsum(1)
    gt, ht = gl[:, -1:], hl[:, -1:]
    gn = gl * gl / (hl + 1.0) + (gt - gl) ** 2 / (ht - hl + 1.0) - gt * gt / (ht + 1.0)
    gn[:, -1] = -1.0
    best = float(gn.max())
    out["gbdt"] = (gain, feat, b, best, float(gn[feat, b]))
    # the fused histogram is fixed-point: the chosen split must be the best one up to quantisation
    out["gbdt_ok"] = abs(gain - best) <= 2e-3 * abs(best) + 1e-3 and float(gn[feat, b]) >= best - 2e-3 * abs(best) - 1e-3
    torch.cuda.synchronize()
    comm.gpu.check()
    comm.barrier()
    return out


def staged_process(comm):
    """CUDA tensors in a job WITHOUT a GPU plane (MP4J_GPU_PLANE=0 stands in for ranks on several machines): every
    array collective goes device -> host plane -> device and gives the results of the host path."""
    p, r = comm.getSlaveNum(), comm.getRank()
    torch.cuda.set_device(0)
    assert comm._gpu_plane_blocker() and comm._cuda_plane() is None
    fop, dop = Operands.FLOAT_OPERAND(), Operands.DOUBLE_OPERAND()
    n = 10007
    x = torch.full((n,), float(r + 1), device="cuda")
    comm.allreduceArray(x, fop, Operators.Float.SUM, 3, n - 2)
    want = p * (p + 1) / 2
    assert torch.all(x[3:n - 2] == want) and x[0] == r + 1 and x[n - 1] == r + 1
    froms, tos = CommUtils.createProcessArrayFroms(n, p), CommUtils.createProcessArrayTos(n, p)
    counts = [t - f for f, t in zip(froms, tos)]
    y = torch.full((n,), float(r + 1), device="cuda", dtype=torch.float64)
    comm.reduceScatterArray(y, dop, Operators.Double.MAX, 0, counts)
    assert torch.all(y[froms[r]:tos[r]] == p)
    comm.allgatherArray(y, dop, froms, tos)
    assert torch.all(y == p)
    z = torch.full((n,), float(r), device="cuda")
    comm.broadcastArray(z, fop, 5, n - 5, p - 1)
    assert torch.all(z[5:n - 5] == p - 1) and z[0] == r
    g = torch.full((n,), float(r), device="cuda")
    comm.gatherArray(g, fop, froms, tos, 0)
    if r == 0:
        assert all(torch.all(g[froms[q]:tos[q]] == q) for q in range(p))
    s = torch.arange(n, device="cuda", dtype=torch.float32) if r == 1 else torch.zeros(n, device="cuda")
    comm.scatterArray(s, fop, froms, tos, 1)
    assert torch.equal(s[froms[r]:tos[r]].cpu(), torch.arange(froms[r], tos[r], dtype=torch.float32))
    w = torch.full((n,), float(r + 1), device="cuda")
    comm.reduceArray(w, fop, Operators.Float.SUM, 0, n, p - 1)
    if r == p - 1:
        assert torch.all(w == want)
    try:
        comm.init_gpu()
        raise AssertionError("init_gpu must refuse")
    except Exception as e:
        assert "GPU plane cannot be set up" in str(e), e
    return {"ok": True}


def staged_thread(comm, tid):
    """Thread-level collectives on CUDA tensors without a GPU plane: thread step with torch ops, process step staged."""
    p, r, T = comm.getSlaveNum(), comm.getRank(), comm.getThreadNum()
    torch.cuda.set_device(0)
    fop = Operands.FLOAT_OPERAND()
    n = 6001
    x = torch.full((n,), float(r * T + tid + 1), device="cuda")
    comm.allreduceArray(x, fop, Operators.Float.SUM, 1, n)
    total = p * T * (p * T + 1) / 2
    assert comm._cuda_plane() is None
    assert torch.all(x[1:] == total) and x[0] == r * T + tid + 1, x[:3]
    b = torch.full((n,), float(r * T + tid), device="cuda")
    comm.broadcastArray(b, fop, 0, n, p - 1, T - 1)
    assert torch.all(b == (p - 1) * T + T - 1)
    m = torch.full((n,), float(r * T + tid), device="cuda")
    comm.reduceArray(m, fop, Operators.Float.MAX, 0, n, 0, 1 % T)
    if r == 0 and tid == 1 % T:
        assert torch.all(m == p * T - 1)
    return True
