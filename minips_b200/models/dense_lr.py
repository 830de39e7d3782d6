"""Dense logistic regression on the parameter server, forward on the tensor cores.

The reference's LR application handles sparse libsvm rows; a dense feature matrix turns its
iteration (lr_example.cpp:260,293-315) into   logits = X.w  ->  coef = alpha (y - sigmoid)  ->
grad = coef^T X  ->  Add  ->  Clock.   Device path: `mps_dense_lr_forward_tc` (tcgen05 + TMEM +
TMA, weight pull fused into the prologue, sigmoid/loss in the TMEM epilogue), `mps_dense_lr_grad`
and the dense slab push.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch

from ..ops import _lib
from ..parallel.device_table import KVClientTable


class DenseLRTrainer:
    def __init__(self, client: KVClientTable, x: torch.Tensor, label: torch.Tensor, alpha: float = 0.1):
        assert x.dtype == torch.bfloat16 and x.is_contiguous() and x.shape[1] % 8 == 0
        self.client, self.x, self.label, self.alpha = client, x, label.float().contiguous(), float(alpha)
        self.table = client.table
        self.D = x.shape[1]
        assert self.table.num_keys == self.D and self.table.dtype == torch.float32 and self.table.k == 1
        self.dev, self.lib = client.dev, client.lib
        dp = (self.D + 127) // 128 * 128
        self.w16 = torch.zeros(16 * dp, dtype=torch.bfloat16, device=self.dev)
        # one-pass kernels (X crosses HBM once), D up to 1024, whole iteration at 1 Mi x 1024 on one B200:
        # "rows" (default; one warp per row, bulk-copy ring, nothing resident) 0.349 ms = 0.94 of the HBM copy
        # rate; "stream" (TMA ring + CUDA cores, 64-row tile resident) 0.64 ms; "mma" (both products on tcgen05,
        # the second with the tile as an MN-major operand) 0.66 ms; "twopass" = tcgen05 forward + gradient pass
        # 0.75 ms; PyTorch (two cuBLAS GEMVs) 0.72 ms
        import os
        self.mode = os.environ.get("MINIPS_DENSE_LR", "rows") if dp <= 1024 else "twopass"
        self.fused = self.mode in ("rows", "stream", "mma")
        self.w32 = torch.empty(self.D, dtype=torch.float32, device=self.dev)
        self.sync = torch.zeros(4, dtype=torch.int32, device=self.dev)
        self.grad = torch.zeros(self.D, dtype=torch.float32, device=self.dev)
        self._coef: Optional[torch.Tensor] = None

    def forward(self, start: int, batch: int, loss: Optional[torch.Tensor] = None,
        weights: Optional[torch.Tensor] = None):
        """coef = alpha (y - sigmoid(X w)) for rows [start, start+batch); returns (coef, loss[2])."""
        xb = self.x[start:start + batch]
        n = xb.shape[0]
        if self._coef is None or self._coef.numel() != n:
            self._coef = torch.empty(n, dtype=torch.float32, device=self.dev)
        if loss is None:
            loss = torch.zeros(2, dtype=torch.float32, device=self.dev)
        t = self.table
        c = self.client
        view = clk = wptr = None
        if weights is None:
            c._local_gate()
            view, clk = C.byref(t.view), C.byref(t.clock_view())
        else:
            wptr = C.c_void_p(weights.data_ptr())
        _lib.check(self.lib.mps_dense_lr_forward_tc(
            view, clk, c.progress, wptr, C.c_void_p(xb.data_ptr()),
            C.c_void_p(self.label[start:start + batch].data_ptr()),
            n, self.D, self.alpha, C.c_void_p(self.w16.data_ptr()), C.c_void_p(self.w32.data_ptr()),
            C.c_void_p(self.sync.data_ptr()), C.c_void_p(self._coef.data_ptr()), C.c_void_p(loss.data_ptr()), 0,
            c._sp()),
            "dense_lr_forward_tc")
        return self._coef, loss

    def step(self, start: int, batch: int, loss: Optional[torch.Tensor] = None) -> torch.Tensor:
        """One iteration: Get (fused) -> forward -> gradient -> Add -> Clock.  With D <= 1024 forward and
        gradient are ONE kernel that reads X once (dense_lr_rows.cu by default; MINIPS_DENSE_LR picks another)."""
        if self.fused:
            xb = self.x[start:start + batch]
            if loss is None:
                loss = torch.zeros(2, dtype=torch.float32, device=self.dev)
            t, c = self.table, self.client
            c._local_gate()
            if self.mode in ("rows", "stream"):
                fn = self.lib.mps_dense_lr_step_rows if self.mode == "rows" else self.lib.mps_dense_lr_step_stream
                _lib.check(fn(
                    C.byref(t.view), C.byref(t.clock_view()), c.progress, None, C.c_void_p(xb.data_ptr()),
                    C.c_void_p(self.label[start:start + batch].data_ptr()), xb.shape[0], self.D, self.alpha,
                    C.c_void_p(self.w32.data_ptr()), C.c_void_p(self.grad.data_ptr()), C.c_void_p(loss.data_ptr()), 0,
                    c._sp()), "dense_lr_step_" + self.mode)
            else:
                _lib.check(self.lib.mps_dense_lr_step_tc(
                    C.byref(t.view), C.byref(t.clock_view()), c.progress, None, C.c_void_p(xb.data_ptr()),
                    C.c_void_p(self.label[start:start + batch].data_ptr()), xb.shape[0], self.D, self.alpha,
                    C.c_void_p(self.w16.data_ptr()), C.c_void_p(self.w32.data_ptr()),
                    C.c_void_p(self.sync.data_ptr()), C.c_void_p(self.grad.data_ptr()), C.c_void_p(loss.data_ptr()),
                    0, c._sp()), "dense_lr_step_tc")
            self.client.add_all(self.grad, zero_src=True)
            self.client.clock()
            return loss
        coef, loss = self.forward(start, batch, loss)
        xb = self.x[start:start + batch]
        _lib.check(self.lib.mps_dense_lr_grad(C.c_void_p(xb.data_ptr()), C.c_void_p(coef.data_ptr()), xb.shape[0],
                                              self.D, C.c_void_p(self.grad.data_ptr()), self.client._sp()),
                   "dense_lr_grad")
        self.client.add_all(self.grad, zero_src=True)
        self.client.clock()
        return loss


def oracle_step(w: torch.Tensor, x: torch.Tensor, y: torch.Tensor, alpha: float):
    """fp32 reference on the bf16-rounded inputs: returns (delta, logloss_sum, n_correct)."""
    xf = x.float()
    z = xf @ w.to(torch.bfloat16).float()
    p = torch.sigmoid(z)
    coef = alpha * (y - p)
    loss = (torch.clamp(z, min=0) + torch.log1p(torch.exp(-z.abs())) - y * z).sum().item()
    return coef @ xf, loss, ((p > 0.5) == (y > 0.5)).float().sum().item()
