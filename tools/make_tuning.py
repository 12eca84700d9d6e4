#!/usr/bin/env python
"""K0 autotuner, second half: turn a measured sweep (bench/sweep.py -> *.jsonl) into the table the
runtime consults for AUTO (tuning/b200_p<N>.json).

  torchrun ... bench/sweep.py --out gpurun_out/sweep_8.jsonl      # measure every algorithm
  python tools/make_tuning.py gpurun_out/sweep_8.jsonl            # -> tuning/b200_p8.json

Thresholds derived for allreduce (the only collective with three algorithms):
  oneshot_max_bytes : largest size up to which one-shot is within 3 % of the best algorithm
  nvls_min_bytes    : smallest size from which NVLS is within 3 % of the best for every larger size
  nvls_min_world    : p-1 if NVLS ever wins at this p, else p (AUTO uses NVLS only when world > it)
and, per family, because their crossovers are not allreduce's (at 8 GPUs and 1 GiB two-shot reduce-scatter /
allgather beat NVLS by 2 %):
  rs_nvls_min_bytes .. rs_nvls_max_bytes, ag_nvls_min_bytes .. ag_nvls_max_bytes : the window of sizes in which
  NVLS is within 1 % of two-shot on the reduce-scatter / allgather rows (max 0 = no upper end, min 2^62 = never)
"""
import json
import os
import sys


def nvls_window(all_rows, coll):
    """[min, max): the sizes for which AUTO should use NVLS for this family. max = the smallest size of the run
    of LARGEST sizes where two-shot is more than 1 % faster (0: no upper end); min = the smallest size from
    which NVLS is within 1 % of two-shot for every size below max (2^62: never)."""
    rows = [r for r in all_rows if r["impl"] == "ours" and r["coll"] == coll and r["algo"] in ("nvls", "twoshot")]
    sizes = sorted({r["bytes"] for r in rows})
    good = {}
    for s in sizes:
        cand = {r["algo"]: r["us"] for r in rows if r["bytes"] == s}
        good[s] = "nvls" in cand and cand["nvls"] <= 1.01 * min(cand.values())
    mx = 0
    top = list(reversed(sizes))
    i = 0
    while i < len(top) and not good[top[i]]:
        mx = top[i]
        i += 1
    th = None
    for s in top[i:]:
        if good[s]:
            th = s
        else:
            break
    if th is None:
        return 1 << 62, 0
    # the crossover lies between the last good size and the first bad one above it: take their geometric middle
    if mx:
        below = max(s for s in sizes if s < mx)
        mx = int((below * mx) ** 0.5)
    return th, mx


def main(path):
    rows = [json.loads(l) for l in open(path) if l.strip()]
    all_rows = rows
    rows = [r for r in rows if r["impl"] == "ours" and r["coll"] == "allreduce" and r["algo"] != "auto"]
    p = rows[0]["n_gpus"]
    sizes = sorted({r["bytes"] for r in rows})
    best = {}
    for s in sizes:
        cand = {r["algo"]: r["us"] for r in rows if r["bytes"] == s}
        best[s] = (min(cand.values()), cand)
    oneshot_max = 0
    for s in sizes:
        b, cand = best[s]
        if "oneshot" in cand and cand["oneshot"] <= 1.03 * b:
            oneshot_max = s
        elif "oneshot" in cand:
            break
    nvls_min = None
    for s in reversed(sizes):
        b, cand = best[s]
        if "nvls" in cand and cand["nvls"] <= 1.03 * b:
            nvls_min = s
        else:
            break
    table = {
        "device": "B200", "world": p, "source": os.path.basename(path),
        "oneshot_max_bytes": oneshot_max if (nvls_min is None or oneshot_max < nvls_min) else max(0, nvls_min // 2),
        "nvls_min_bytes": nvls_min if nvls_min is not None else 1 << 62,
        "nvls_min_world": p - 1 if nvls_min is not None else p,
        "rs_nvls_min_bytes": nvls_window(all_rows, "reduce_scatter")[0],
        "rs_nvls_max_bytes": nvls_window(all_rows, "reduce_scatter")[1],
        "ag_nvls_min_bytes": nvls_window(all_rows, "allgather")[0],
        "ag_nvls_max_bytes": nvls_window(all_rows, "allgather")[1],
        "allreduce_us": {str(s): best[s][1] for s in sizes},
    }
    out = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tuning", "b200_p%d.json" % p)
    with open(out, "w") as f:
        json.dump(table, f, indent=1, sort_keys=True)
    print(out, {k: table[k] for k in ("oneshot_max_bytes", "nvls_min_bytes", "nvls_min_world", "rs_nvls_min_bytes",
                                      "rs_nvls_max_bytes", "ag_nvls_min_bytes", "ag_nvls_max_bytes")})


if __name__ == "__main__":
    main(sys.argv[1])
