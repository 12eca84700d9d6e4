"""Data-parallel multinomial logistic regression on top of the communicator - the consumer the
reference names first (ytk-learn LR / L-BFGS: instance counts, gradients, R/README.md:268-278).

Per step, on every rank (one GPU each):
  logits = X @ W                       bf16 GEMM (cuBLAS: a plain library GEMM)
  R      = softmax(logits) - onehot(y) one fused kernel (lr_residual.cu), written transposed as bf16 [C, rows]
  G_r    = reduce-scatter over ranks of X^T R    ONE fused kernel: tcgen05 GEMM whose output tiles
                                                  are added into the owner rank's shard (K12)
  W_r   -= lr * G_r / N                 sharded update: rank r owns rows [r*fpr, (r+1)*fpr) of W
  W      = allgather(W_r)               our allgather kernel on the symmetric weight tensor
The weights stay sharded between the reduce-scatter and the allgather, exactly the shape the
reference gets from reduceScatterArray + allgatherArray (PCS:1174, 620). X is kept twice, row-major
for the forward pass and feature-major (Xt) for the gradient, so both GEMMs are K-major.

Sparse variant (`SparseLR`): hashed features, gradient as a sparse map reduced with allreduce_map.
"""
import torch

from .. import Operands, Operators
from .. import _native as N


class DenseLR:
    def __init__(self, comm, n_features, n_classes, lr=0.1, dtype=torch.bfloat16):
        self.comm = comm
        self.p, self.r = comm.getSlaveNum(), comm.getRank()
        self.F, self.C, self.lr = n_features, n_classes, lr
        self.plane = comm.init_gpu()
        self.fpr = (n_features + self.p - 1) // self.p
        # full weights live in the symmetric heap so the allgather is zero-copy
        self.W = comm.empty(self.fpr * self.p * n_classes, Operands.FLOAT_OPERAND()).view(self.fpr * self.p, n_classes)
        self.W.zero_()
        self.dtype = dtype

    def set_data(self, X, y):
        """X: [rows, F] float tensor on this rank's GPU, y: [rows] int64 labels."""
        rows = X.shape[0]
        pad = (-rows) % 8  # TMA row pitch of Xt must be a multiple of 16 bytes
        if pad:
            X = torch.cat([X, X.new_zeros(pad, X.shape[1])])
            y = torch.cat([y, y.new_full((pad,), -1)])
        self.X = X.to(self.dtype).contiguous()
        self.Xt = self.X.t().contiguous()
        self.y = y
        self.valid = (y >= 0)
        self.n_local = int(self.valid.sum())
        self.n_total = self.comm.allreduce(self.n_local, Operands.LONG_OPERAND(), Operators.Long.SUM)
        self._rt = torch.empty(self.C, self.X.shape[0], dtype=torch.bfloat16, device=self.X.device)
        self._loss = torch.zeros(1, dtype=torch.float32, device=self.X.device)

    def step(self):
        from ..parallel.gpu_plane import lr_softmax_residual
        W = self.W[: self.F]
        logits = self.X @ W.to(self.dtype)                               # plain library GEMM (cuBLAS), bf16
        # softmax, one-hot, mask, loss, transpose and cast in ONE kernel, written K-major for the next one
        self._loss.zero_()
        Rt, loss_local = lr_softmax_residual(logits, self.y, self._rt, self._loss)
        shard, fpr = self.plane.lr_xtr_reduce_scatter(self.Xt, Rt)     # fused GEMM + reduce-scatter
        lo = self.r * fpr
        self.W[lo:lo + fpr].add_(shard, alpha=-self.lr / self.n_total)  # sharded update
        flat = self.W.view(-1)
        froms = [q * fpr * self.C for q in range(self.p)]
        tos = [(q + 1) * fpr * self.C for q in range(self.p)]
        self.comm.allgatherArray(flat, Operands.FLOAT_OPERAND(), froms, tos)
        return loss_local

    def loss(self, loss_local):
        total = self.comm.allreduce(float(loss_local), Operands.DOUBLE_OPERAND(), Operators.Double.SUM)
        return total / self.n_total


class SparseLR:
    """Binary LR over hashed sparse features: the gradient of a minibatch touches few of the
    n_features weights, so it travels as a sparse map (keys = feature ids) through allreduce_map
    (the GPU form of the reference's allreduceMap, PCS:2053)."""

    def __init__(self, comm, n_features, lr=0.1, max_nnz=None, rows_total=None):
        """max_nnz: an upper bound of the non-zeros of any rank's minibatch; rows_total: rows of a minibatch over all
        ranks. Both must be the same on every rank. When given, a step involves no host communication (the map
        collective gets its sizes from them); when not, they are agreed through host allreduces every step."""
        self.comm, self.lr, self.F = comm, lr, n_features
        self.plane = comm.init_gpu()
        self.w = torch.zeros(n_features, device="cuda")
        self.max_nnz, self.rows_total = max_nnz, rows_total

    def step(self, rows, cols, vals, y):
        """COO minibatch: rows/cols int64, vals float32, y float32 in {0,1} of length n_rows."""
        n = y.numel()
        z = torch.zeros(n, device="cuda").index_add_(0, rows, vals * self.w[cols])
        r = torch.sigmoid(z) - y
        g_vals = vals * r[rows]
        keys, g = self.plane.map_reduce_local(cols, g_vals, N.SUM)       # local pre-aggregation by feature
        if self.max_nnz is not None:                                     # sparse allreduce over ranks
            cap = min(self.max_nnz * self.comm.getSlaveNum(), self.F)
            keys, g = self.plane.allreduce_map(keys, g, Operators.Float.SUM, max_n=self.max_nnz, out_cap=cap,
                                               copy=False)
        else:
            keys, g = self.plane.allreduce_map(keys, g, Operators.Float.SUM, copy=False)
        n_total = self.rows_total if self.rows_total is not None else \
            self.comm.allreduce(n, Operands.LONG_OPERAND(), Operators.Long.SUM)
        self.w.index_add_(0, keys, g, alpha=-self.lr / n_total)
        return torch.nn.functional.binary_cross_entropy_with_logits(z, y, reduction="sum")
