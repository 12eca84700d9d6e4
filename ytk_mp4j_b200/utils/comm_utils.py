"""CommUtils: range validators and balanced partitioners (parity: J/utils/CommUtils.java:34-222).

Rule: avg = size // n; participants 0..n-2 get avg elements, the last one takes
the remainder. The same rule applies to the threads inside a process.
"""
from ..exception import Mp4jException


class CommUtils:
    @staticmethod
    def isfromsTosLegal(froms, tos, threadNum=None, size=None):
        """size: length of the array the ranges index; a range that ends beyond it is rejected here (the
        reference fails with ArrayIndexOutOfBoundsException; native code would read past the buffer)."""
        if threadNum is not None:  # 2-D form: [slaveNum][threadNum]
            for f, t in zip(froms, tos):
                CommUtils.isfromsTosLegal(f, t, size=size)
                if len(f) != threadNum:
                    raise Mp4jException("froms arrays must be array[slaveNum][threadNum]")
                if len(t) != threadNum:
                    raise Mp4jException("tos arrays must be array[slaveNum][threadNum]")
            CommUtils.isfromsTosLegal(CommUtils.getProcessFroms(froms), CommUtils.getProcessTos(tos))
            return
        if len(froms) != len(tos):
            raise Mp4jException("froms, tos's length must be equal!")
        for i in range(len(froms)):
            if froms[i] < 0:
                raise Mp4jException("froms[i] must be >= 0!")
            if tos[i] < 0:
                raise Mp4jException("tos[i] must be >= 0!")
            if froms[i] > tos[i]:
                raise Mp4jException("froms[i] must be <= tos[i]!")
            if i >= 1 and froms[i] < tos[i - 1]:
                raise Mp4jException("froms[i] must be >= to[i - 1]")
            if size is not None and tos[i] > size:
                raise Mp4jException("tos[i] = %d is beyond the array length %d" % (tos[i], size))

    @staticmethod
    def isFromCountsLegal(from_, counts, size=None):
        if from_ < 0:
            raise Mp4jException("from must be >= 0!")
        total = 0
        for c in counts:
            if isinstance(c, (list, tuple)):
                for cc in c:
                    if cc < 0:
                        raise Mp4jException("counts[i][j] must be >= 0!")
                    total += cc
            elif c < 0:
                raise Mp4jException("counts[i] must be >= 0!")
            else:
                total += c
        if size is not None and from_ + total > size:
            raise Mp4jException("from + sum(counts) = %d is beyond the array length %d" % (from_ + total, size))

    @staticmethod
    def isFromToLegal(from_, to, size=None):
        if from_ < 0:
            raise Mp4jException("from must be >= 0!")
        if to < 0:
            raise Mp4jException("to must be >= 0!")
        if from_ > to:
            raise Mp4jException("from must be <= to!")
        if size is not None and to > size:
            raise Mp4jException("to = %d is beyond the array length %d" % (to, size))

    @staticmethod
    def getFromsFromCount(from_, counts, slaveNum):
        froms = []
        for i in range(slaveNum):
            froms.append(from_)
            from_ += counts[i]
        return froms

    @staticmethod
    def getTosFromCount(from_, counts, slaveNum):
        tos = []
        for i in range(slaveNum):
            from_ += counts[i]
            tos.append(from_)
        return tos

    @staticmethod
    def getProcessFroms(froms):
        return [f[0] for f in froms]

    @staticmethod
    def getProcessTos(tos):
        return [t[-1] for t in tos]

    @staticmethod
    def createProcessArrayFroms(size, slaveNum):
        avg = size // slaveNum
        return [r * avg for r in range(slaveNum)]

    @staticmethod
    def createProcessArrayTos(size, slaveNum):
        avg = size // slaveNum
        return [size if r == slaveNum - 1 else (r + 1) * avg for r in range(slaveNum)]

    @staticmethod
    def createThreadArrayFroms(size, slaveNum, threadNum):
        pf = CommUtils.createProcessArrayFroms(size, slaveNum)
        pt = CommUtils.createProcessArrayTos(size, slaveNum)
        return [[x + pf[r] for x in CommUtils.createProcessArrayFroms(pt[r] - pf[r], threadNum)]
                for r in range(slaveNum)]

    @staticmethod
    def createThreadArrayTos(size, slaveNum, threadNum):
        pf = CommUtils.createProcessArrayFroms(size, slaveNum)
        pt = CommUtils.createProcessArrayTos(size, slaveNum)
        return [[x + pf[r] for x in CommUtils.createProcessArrayTos(pt[r] - pf[r], threadNum)]
                for r in range(slaveNum)]

    @staticmethod
    def main(argv=None):
        """Self-check in the spirit of CommUtils.main (J/utils/CommUtils.java:224-238): prints the balanced
        partitions of a few sizes and asserts they tile the range."""
        for size, p, t in ((10, 3, 2), (1000, 8, 8), (7, 8, 1)):
            pf, pt = CommUtils.createProcessArrayFroms(size, p), CommUtils.createProcessArrayTos(size, p)
            assert pf[0] == 0 and pt[-1] == size and all(pt[i] == pf[i + 1] for i in range(p - 1))
            tf, tt = CommUtils.createThreadArrayFroms(size, p, t), CommUtils.createThreadArrayTos(size, p, t)
            CommUtils.isfromsTosLegal(tf, tt, t)  # raises on an illegal partition
            print("size=%d slaves=%d threads=%d froms=%s tos=%s" % (size, p, t, pf, pt))
        return 0

    # helper shared by the communicators (PCS:1741-1756)
    @staticmethod
    def splitRange(from_, to, n):
        avg = (to - from_) // n
        froms = [from_ + r * avg for r in range(n)]
        tos = [to if r == n - 1 else from_ + (r + 1) * avg for r in range(n)]
        return froms, tos
