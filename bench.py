#!/usr/bin/env python
"""Headline benchmark (driver contract: see the task statement / DESIGN.md "bench").

Metric (BASELINE.json: "allreduce bus GB/s"): config "ThreadCommSlave N procs x 8 threads, float SUM
allreduce 64M elems (thread-level + GPU-level fusion)". Per GPU: T = 8 thread buffers of 64 Mi
fp32 (256 MiB each, 2 GiB per GPU - far larger than the 126 MB L2, so no flush is needed between
iterations); one step = ThreadCommSlave.allreduceArray over all N*T participants: ONE fused
kernel per GPU (T-way register reduce -> inter-GPU reduce-scatter/allgather over NVLink -> fan-out
to the T buffers).

  busbw_per_rank = (S / t) * 2(P-1)/P,  S = 256 MiB, P = N*T participants     (nccl-tests convention)
  value          = N * busbw_per_rank   (whole-job aggregate; weak scaling: per-GPU work is fixed)
  vs_baseline    = value / (N * 0.1028 GB/s)  - the best per-rank allreduce bus bandwidth anywhere in
                   the reference's published tables (README.md:347, 1 GbE), scaled to N ranks.

`value` is device-timed (CUDA events around each step on the launching stream, summed, max over
ranks). `e2e` goes through the public API with T Python threads per process and includes, every
step, the host->device copy of all T inputs from pinned memory and the device->host read of the
result array.
"""
import argparse
import json
import os
import statistics
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

PUBLISHED_BEST_BUSBW_GBPS = 0.1028  # BASELINE.md §3 (README.md:347)


def reference_arm():
    # The reference is a Maven/Java project: `pip install /root/reference` fails ("Neither
    # 'setup.py' nor 'pyproject.toml' found") and the image has no JDK/Maven. See DESIGN.md.
    print(json.dumps({"impl": "reference",
                      "unavailable": "reference is pure Java (Maven); no JDK/Maven/pyproject on this image, "
                                     "pip install of /root/reference is not possible"}))
    return 0


class ClockSampler:
    """SM clocks + throttle reasons of the job's GPUs DURING the timed region, by the profiling recipe's own
    tool: one `nvidia-smi --query-gpu=... -lms 5` child process (rank 0 only, all N GPUs), started before
    the warm-up, killed after the timed steps; only samples whose timestamp lies inside the timed region
    count. In-process NVML polling is not used: from a thread per rank, or even synchronously between
    launches, it stalled the launch loop of whichever rank was asking, and every other rank's kernel then
    waited for that rank inside its timed step (8 GPUs: 0.82 -> 0.89-1.04 ms per step)."""

    FIELDS = ("timestamp,index,clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, n_gpus, enabled=True):
        self.proc, self.path = None, None
        if not enabled or os.environ.get("MP4J_BENCH_NOSAMPLER"):
            return
        import subprocess
        vis = os.environ.get("CUDA_VISIBLE_DEVICES")
        ids = vis.split(",")[:n_gpus] if vis else [str(i) for i in range(n_gpus)]
        fd, self.path = tempfile.mkstemp(prefix="mp4j_clocks_", suffix=".csv")
        os.close(fd)
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "--query-gpu=" + self.FIELDS, "--format=csv,noheader,nounits",
                                          "-lms", "5", "-i", ",".join(ids)], stdout=open(self.path, "w"),
                                         stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def wait_started(self, timeout_s=5.0):
        """Blocks until the child has written its first sample (its start-up takes a few hundred ms)."""
        if self.proc is None:
            return
        t0 = time.time()
        while time.time() - t0 < timeout_s:
            try:
                if os.path.getsize(self.path) > 0:
                    return
            except OSError:
                pass
            time.sleep(0.01)

    def finish(self, t_begin, t_end):
        """t_begin / t_end: time.time() at the start / end of the timed region."""
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0, "tool": "nvidia-smi -lms 5 (rank 0, all GPUs)"}
        if self.proc is None:
            out["tool"] = "none"
            return out
        import datetime
        self.proc.terminate()
        try:
            self.proc.wait(timeout=3)
        except Exception:
            self.proc.kill()
        inside, around, reasons, maxs = [], [], set(), []
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        for line in open(self.path):
            f = [x.strip() for x in line.split(",")]
            if len(f) < 8:
                continue
            try:
                ts = datetime.datetime.strptime(f[0], "%Y/%m/%d %H:%M:%S.%f").timestamp()
                mhz, mx = float(f[2]), float(f[3])
            except ValueError:
                continue
            maxs.append(mx)
            if t_begin - 0.003 <= ts <= t_end + 0.003:
                inside.append(mhz)
                for nm, v in zip(names, f[4:8]):
                    if v.lower().startswith("active"):
                        reasons.add(nm)
            elif t_begin - 0.05 <= ts <= t_end + 0.05:
                around.append(mhz)
        try:
            os.unlink(self.path)
        except OSError:
            pass
        use = inside if inside else around
        out.update({"sm_mhz": statistics.median(use) if use else None, "sm_max_mhz": max(maxs) if maxs else None,
                    "reasons": sorted(reasons), "samples": len(inside),
                    "window": "timed region" if inside else "within 50 ms of the timed region (under the same load)"})
        return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--elems", type=int, default=64 << 20)
    ap.add_argument("--threads", type=int, default=8)
    ap.add_argument("--e2e-steps", type=int, default=3)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--e2e-fanout", action="store_true", help="e2e: write the result into all T host arrays")
    ap.add_argument("--no-extras", action="store_true",
                    help="N >= 2: skip the correctness matrix and the dense float/double sweep vs NCCL")
    ap.add_argument("--sweep-max-bytes", type=int, default=1 << 30)
    args = ap.parse_args()
    if args.impl == "reference":
        return reference_arm()

    import torch
    from ytk_mp4j_b200 import CommMaster, Operands, Operators, ThreadCommSlave
    from ytk_mp4j_b200 import _native as NV

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    addr = os.environ.get("MASTER_ADDR", "127.0.0.1")
    base_port = int(os.environ.get("MASTER_PORT", "29500"))
    W = max(args.warmup, 3)
    K, T, n = args.steps, args.threads, args.elems
    torch.cuda.set_device(local_rank)

    master = None
    if rank == 0:
        port = base_port + 17 if world > 1 else 0
        master = CommMaster(world, port, quiet=True, work_dir=tempfile.gettempdir()).start()
        port = master.port
    else:
        port = base_port + 17
    comm = ThreadCommSlave("bench", T, addr, port, rank_hint=rank, host="127.0.0.1")
    pcs = comm.processCommSlave
    extras = world > 1 and not args.no_extras
    # the dense sweep needs a symmetric pool of the largest message next to the plane's own regions
    plane = comm.init_gpu(device=local_rank, heap_bytes=(args.sweep_max_bytes + (1 << 30)) if extras else None)
    fop, SUM = Operands.FLOAT_OPERAND(), Operators.Float.SUM
    P = world * T
    S = n * 4

    bufs = [torch.full((n,), 1e-30 * (t + 1 + rank * T), dtype=torch.float32, device="cuda") for t in range(T)]
    # ---- correctness gate before timing: known pattern through the same kernel
    chk = [torch.full((4099,), float(rank * T + t + 1), device="cuda") for t in range(T)]
    plane.thread_allreduce(chk, fop, SUM, 0, 4099)
    torch.cuda.synchronize()
    if not os.environ.get("MP4J_BENCH_NOCHECK"):
        assert all(bool(torch.all(c == P * (P + 1) / 2)) for c in chk), "allreduce self-check failed"

    def step():
        plane.thread_allreduce(bufs, fop, SUM, 0, n)

    sampler = ClockSampler(world, enabled=(rank == 0))
    sampler.wait_started()
    pcs.barrier()
    for _ in range(W):
        step()
    torch.cuda.synchronize()
    pcs.barrier()
    starts = [torch.cuda.Event(enable_timing=True) for _ in range(K)]
    ends = [torch.cuda.Event(enable_timing=True) for _ in range(K)]
    torch.cuda.synchronize()
    t_begin = time.time()
    t0 = time.perf_counter()
    for i in range(K):
        starts[i].record()
        step()
        ends[i].record()
    torch.cuda.synchronize()
    wall_ms = (time.perf_counter() - t0) * 1e3
    t_end = time.time()
    pcs.barrier()
    clocks = sampler.finish(t_begin, t_end)
    plane.check()
    dev_ms = sum(s.elapsed_time(e) for s, e in zip(starts, ends))
    algo_desc = plane.last()
    # max over ranks through our own host allreduce (double MAX)
    dop, DMAX = Operands.DOUBLE_OPERAND(), Operators.Double.MAX
    dev_ms = pcs.allreduce(dev_ms, dop, DMAX)
    wall_ms = pcs.allreduce(wall_ms, dop, DMAX)
    ms_per_step = dev_ms / K
    busbw = (S / (ms_per_step * 1e-3)) * 2 * (P - 1) / P / 1e9
    value = world * busbw
    # HBM traffic of one step on one GPU: read T*S + write T*S (+ staging at N > 1)
    hbm_bytes = 2 * T * S + (0 if world == 1 else 3 * S)
    hbm_gbs = hbm_bytes / (ms_per_step * 1e-3) / 1e9

    # ---- end to end through the public API: T python threads call ThreadCommSlave.allreduceArray on PINNED
    # HOST tensors. The library streams them through the GPU (chunked H2D || fused kernel || D2H, one
    # cudaMemcpyAsync per buffer and chunk). Every step moves the T input arrays (T*S bytes) host->device and
    # reads the result array (S bytes) back into thread 0's host array (`resultThread=0`: the T threads of a
    # process share one result; `--e2e-fanout` writes it into all T host arrays instead, T*S bytes back).
    e2e = None
    if not args.no_e2e:
        Ke = max(1, args.e2e_steps)
        rt = None if args.e2e_fanout else 0
        host = [torch.full((n,), 1e-30, dtype=torch.float32).pin_memory() for _ in range(T)]
        e2e_times = []
        e2e_err = []

        def worker(t):
            try:
                comm.setThreadId(t)
                torch.cuda.set_device(local_rank)
                for it in range(Ke + 1):
                    comm.barrier()                                    # all ranks, all threads: step starts together
                    if t == 0:
                        e2e_times.append(time.perf_counter())
                    comm.allreduceArray(host[t], fop, SUM, 0, n, resultThread=rt)  # public API: host array in, result out
                    if t == 0:
                        e2e_times.append(time.perf_counter())
            except BaseException as e:  # noqa: BLE001 - reported on the main thread
                e2e_err.append(e)
        ths = [threading.Thread(target=worker, args=(t,)) for t in range(T)]
        for th in ths:
            th.start()
        for th in ths:
            th.join()
        if e2e_err:
            raise e2e_err[0]
        e2e_algo = plane.last()
        per = [(e2e_times[2 * i + 1] - e2e_times[2 * i]) * 1e3 for i in range(1, Ke + 1)]  # drop warm-up
        e2e_ms = pcs.allreduce(sum(per) / len(per), dop, DMAX)
        e2e_busbw = (S / (e2e_ms * 1e-3)) * 2 * (P - 1) / P / 1e9
        # result check: every thread's host array holds the same finite sum
        # result check: rank 0's result (thread 0's array; every thread's with --e2e-fanout) must equal the
        # sum the inputs imply: all inputs were 1e-30, and each step multiplies the live value by P once
        # for the threads that received the previous result
        chk_src = host if rt is None else [host[0]]
        e2e_ok = all(bool(torch.equal(chk_src[0][:4096], h[:4096])) for h in chk_src) and \
            bool(torch.isfinite(host[0][:4096]).all()) and float(host[0][0]) > 1e-30
        e2e = {"value": round(world * e2e_busbw, 3), "unit": "GB/s", "ms_per_step": round(e2e_ms, 3),
               "steps": Ke, "h2d_bytes_per_step": T * S, "d2h_bytes_per_step": (T if rt is None else 1) * S,
               "h2d_GBps_per_gpu": round(T * S / (e2e_ms * 1e-3) / 1e9, 2),
               "result_ok": e2e_ok, "path": e2e_algo,
               "api": "ThreadCommSlave.allreduceArray on pinned host tensors (T python threads); result in "
                      + ("all T host arrays" if rt is None else "thread 0's host array (resultThread=0)")}

    # ---- the other metrics BASELINE.json names (LR step/s, sparse map keys/s, GBDT), at every N
    consumers = None
    if not args.no_extras:
        sys.path.insert(0, os.path.join(ROOT, "bench"))
        import driver_extras
        del bufs
        bufs = None
        torch.cuda.empty_cache()
        try:
            consumers = driver_extras.consumers(pcs, plane, rank, world)
        except Exception as e:  # noqa: BLE001 - the headline number must still be printed
            consumers = {"error": repr(e)[:300]}
    # ---- N >= 2: the rest of the product on the real multi-GPU path, under the driver's eyes
    correctness = sweep = None
    if extras:
        torch.cuda.empty_cache()
        correctness = driver_extras.correctness_matrix(pcs, plane, rank, world)
        sweep = driver_extras.dense_sweep(pcs, plane, rank, world, max_bytes=args.sweep_max_bytes)
        try:
            import torch.distributed as dist
            if dist.is_initialized():
                dist.destroy_process_group()
        except Exception:  # noqa: BLE001
            pass

    pcs.barrier()
    comm.close(0)
    if master is not None:
        master.stop(timeout_s=10)
        master.close()
    if rank == 0:
        out = {
            "metric": "allreduce_bus_bandwidth_aggregate",
            "value": round(value, 3),
            "unit": "GB/s",
            "n_gpus": world,
            "steps": K,
            "warmup": W,
            "ms_per_step": round(ms_per_step, 4),
            "higher_is_better": True,
            "scaling": "weak",
            "vs_baseline": round(value / (world * PUBLISHED_BEST_BUSBW_GBPS), 1),
            "dtype": "fp32",
            "data": "synthetic",
            "impl": "ours",
            "config": {
                "model": "ThreadCommSlave allreduceArray float SUM (BASELINE.json config 3)",
                "procs": world, "threads_per_proc": T, "participants": P,
                "elems_per_array": n, "bytes_per_array": S, "global_batch": P, "seq_len": n,
                "parallelism": "dp%d x %d threads" % (world, T),
                "l2_policy": "inputs (2 GiB per GPU) larger than L2; no flush needed",
                "baseline_number": "0.1028 GB/s per-rank bus bandwidth (README.md:347, 1 GbE) x N ranks",
                "kernel": algo_desc,
            },
            "busbw_per_rank_GBps": round(busbw, 3),
            "hbm_GBps_per_gpu": round(hbm_gbs, 1),
            "wall_ms_per_step": round(wall_ms / K, 4),
            "clocks": clocks,
            "gpu_launches": K,
            "e2e": e2e,
            "consumers": consumers,
            "correctness_matrix": correctness,
            "dense_sweep": sweep,
        }
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
            out["hbm_frac_of_measured"] = round(hbm_gbs / peaks["hbm_gbs"], 3)
        except Exception:
            pass
        print(json.dumps(out))
    return 0


if __name__ == "__main__":
    sys.exit(main())
