"""Map<String, number> collectives on the GPU: the binding-side bridge between String keys and the
64-bit keys of the sparse-map kernels (reference: allreduceMap PCS:2053-2081, reduceScatterMap PCS:1299-1327,
allgatherMap PCS:666-692 on Map<String,T>; kernels: csrc/gpu/map_kernels.cu).

Key of a string = (FNV-1a of its UTF-8 bytes) << 32 | String.hashCode(). The low word keeps the reference's
placement (owner = hashCode % slaveNum, partition "java32"); the high word is an independent hash, so two
different strings share a key only if BOTH 32-bit hashes agree.

The way back needs the strings: every rank keeps a dictionary key -> string and, on each call, allgathers
(host plane) the strings it has not announced before. All ranks see all announcements, so the dictionaries are
identical everywhere and grow only with NEW keys - in an iterative job (the same features every step) the
string traffic vanishes after the first steps and only (key, value) tensors move, over NVLink.

Collision policy: a key that arrives with two different strings (local input or announcement) marks the bridge
broken on every rank at the same call - all ranks parse the same announcements - and that call and all later
ones take the host implementation, which compares whole strings. Nothing is ever merged wrongly.
"""
import ctypes as C

import numpy as np

from .. import _native as N
from ..exception import Mp4jException, check


_TEST_MASK = None  # tests narrow the hash to provoke collisions


def keys64(strings):
    """64-bit keys of a list of str (native batch hash)."""
    n = len(strings)
    out = np.empty(n, dtype=np.int64)
    if n == 0:
        return out
    try:
        blob = "\0".join(strings).encode("utf-8")
    except TypeError:
        raise Mp4jException("map keys must be str (Map<String, V> in the reference)")
    if blob.count(b"\0") != n - 1:
        raise Mp4jException("map keys must not contain NUL characters")
    check(N.lib.mp4j_map_keys64(blob, len(blob), n, out.ctypes.data_as(N.c_i64p)))
    if _TEST_MASK is not None:
        out &= _TEST_MASK
    return out


class MapKeyBridge:
    MASK60 = (1 << 60) - 1

    def __init__(self, pcs):
        self._pcs = pcs
        self._rev = {}        # key64 -> str, identical on every rank
        self._rev60 = {}      # low 60 bits -> str: for keys whose top 4 bits carry a destination rank ("tagged")
        self.broken = False   # a collision was seen: host path from now on (same on every rank)

    def _learn(self, strings, ks):
        rev, rev60, m60 = self._rev, self._rev60, self.MASK60
        for s, k in zip(strings, ks):
            prev = rev.setdefault(k, s)
            if prev is not s and prev != s:
                self.broken = True
            prev = rev60.setdefault(k & m60, s)
            if prev is not s and prev != s:
                self.broken = True

    def encode(self, strings):
        """Keys of this rank's strings; announces the ones nobody has seen to every rank (collective)."""
        from ..operands import Operands
        ks = keys64(strings)
        kl = ks.tolist()
        rev = self._rev
        # new to everybody - or colliding with a known string: announced too, so that EVERY rank sees the collision
        fresh = [s for s, k in zip(strings, kl) if rev.get(k) != s]
        # local duplicates / collisions inside this call
        self._learn(strings, kl)
        blob = "\0".join(fresh).encode("utf-8")
        p, r = self._pcs.getSlaveNum(), self._pcs.getRank()
        sizes = np.zeros(2 * p, dtype=np.int64)
        sizes[2 * r], sizes[2 * r + 1] = len(blob), len(fresh)
        lop = Operands.LONG_OPERAND()
        self._pcs.allgatherArray(sizes, lop, [2 * q for q in range(p)], [2 * q + 2 for q in range(p)])
        offs = np.concatenate([[0], np.cumsum(sizes[0::2])]).astype(np.int64)
        if int(offs[-1]) > 0:
            buf = np.zeros(int(offs[-1]), dtype=np.int8)
            buf[offs[r]:offs[r + 1]] = np.frombuffer(blob, dtype=np.int8)
            self._pcs.allgatherArray(buf, Operands.BYTE_OPERAND(), offs[:-1].tolist(), offs[1:].tolist())
            raw = buf.tobytes()
            for q in range(p):
                if q == r or sizes[2 * q + 1] == 0:
                    continue
                theirs = raw[offs[q]:offs[q + 1]].decode("utf-8").split("\0")
                self._learn(theirs, keys64(theirs).tolist())
        return ks

    def decode(self, ks, tagged=False):
        try:
            if tagged:
                rev, m60 = self._rev60, self.MASK60
                return [rev[k & m60] for k in ks]
            rev = self._rev
            return [rev[k] for k in ks]
        except KeyError:
            raise Mp4jException("GPU map result carries a key nobody announced (internal error)")

    @staticmethod
    def tag(ks, dest):
        """keys whose top 4 bits name the destination rank (partition "tagged")."""
        u = ks.astype(np.uint64) & np.uint64(MapKeyBridge.MASK60)
        return (u | (np.uint64(dest) << np.uint64(60))).astype(np.int64)
