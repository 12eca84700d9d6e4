"""Data-parallel GBDT split finding on top of the communicator (ytk-learn style, the reference's
second named consumer, R/README.md:278): every rank holds a row shard of the binned feature matrix;
per boosting node the (grad, hess) histograms of all features are built and reduce-scattered in ONE
fused kernel (K13), every rank scans the features it owns for its best split, and the global best
split is an allreduce with the packed FLOAT_MAX_LOC operator (Operators.java:56-88) - gain in the
float half, feature*256+bin in the int half.
"""
import torch

from .. import Operands, Operators


class HistGBDT:
    def __init__(self, comm, n_features, reg_lambda=1.0):
        self.comm, self.F, self.lam = comm, n_features, reg_lambda
        self.plane = comm.init_gpu()
        self.p, self.r = comm.getSlaveNum(), comm.getRank()

    def set_data(self, bins_t):
        """bins_t: uint8 [F, rows] (feature-major, 256 bins); rows are padded to a multiple of 16."""
        F, rows = bins_t.shape
        pitch = (rows + 15) // 16 * 16
        buf = torch.zeros(F, pitch, dtype=torch.uint8, device=bins_t.device)
        buf[:, :rows] = bins_t
        self.bins_t, self.rows = buf[:, :rows], rows

    def best_split(self, grad, hess):
        """One node: returns (gain, feature, bin) of the best split over all ranks' rows."""
        shard, fpr = self.plane.gbdt_hist_reduce_scatter(self.bins_t, grad, hess)   # fused build + reduce-scatter
        lo = self.r * fpr
        nown = max(0, min(self.F, lo + fpr) - lo)
        best = Operators.Double.compositeDouble(-1.0, 0)
        if nown > 0:
            h = shard[:nown]
            gl, hl = h[..., 0].cumsum(1), h[..., 1].cumsum(1)
            gt, ht = gl[:, -1:], hl[:, -1:]
            gr, hr = gt - gl, ht - hl
            gain = gl * gl / (hl + self.lam) + gr * gr / (hr + self.lam) - gt * gt / (ht + self.lam)
            gain[:, -1] = -1.0  # splitting after the last bin separates nothing
            flat = torch.argmax(gain)
            f, b = int(flat) // 256, int(flat) % 256
            best = Operators.Double.compositeDouble(float(gain[f, b]), (lo + f) * 256 + b)
        best = self.comm.allreduce(best, Operands.DOUBLE_OPERAND(), Operators.Double.FLOAT_MAX_LOC)
        loc = Operators.Double.getIntLoc(best)
        return Operators.Double.getFloatVal(best), loc // 256, loc % 256
