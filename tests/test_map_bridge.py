"""String keys <-> 64-bit keys of the GPU map collectives (ytk_mp4j_b200/comm/map_bridge.py): Java placement in the
low word, identical dictionaries on all ranks, collision policy."""
import numpy as np

import workers
from ytk_mp4j_b200.comm.map_bridge import keys64
from ytk_mp4j_b200.comm.process import java_string_hash
from ytk_mp4j_b200.parallel.local_cluster import run_cluster


def test_keys64_low_word_is_java_hashcode_and_never_reserved():
    strs = ["", "a", "hello", "0", "12345", "ключ", "鍵", "😀x", "f" * 100]
    ks = keys64(strs)
    for s, k in zip(strs, ks.tolist()):
        assert (k & 0xFFFFFFFF) == (java_string_hash(s) & 0xFFFFFFFF), s
        assert k != -(1 << 63)
    assert len(set(ks.tolist())) == len(strs)
    many = keys64(["feature_%d" % i for i in range(200000)])
    assert len(np.unique(many)) == 200000  # 64 bits: no collision among 200k keys


def test_bridge_dictionaries_agree_across_ranks():
    res, code, _ = run_cluster(3, workers.map_bridge_worker, (None,), timeout=120)
    assert code == 0
    for out in res:
        assert out == {"broken": False, "n": 200 + 3 * 52, "roundtrip": True, "mine": True, "stable": True, "tag": True}, out


def test_bridge_collision_breaks_every_rank_together():
    # 6-bit keys: collisions are certain; every rank must notice in the same call (then the host path is used)
    res, code, _ = run_cluster(3, workers.map_bridge_worker, (0x3F,), timeout=120)
    assert code == 0
    assert all(out["broken"] for out in res), res
