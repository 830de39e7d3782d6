"""Mini-batch k-means on the parameter server (reference apps/kmeans/kmeans.cpp:154-297,
apps/kmeans/kmeans_helper.hpp).

Two tables, like the reference (kmeans.cpp:117,143-150): table 1 holds the K*D centroid
coordinates (row-major), table 2 the K per-cluster counts.  Per iteration every worker pulls
both, assigns its batch, and pushes centroid / count deltas, then clocks both tables.
Device path: the assignment is the tcgen05 GEMM of csrc/kernels/kmeans_tc.cu with the centroid
pull fused into its B-operand fill; partial sums, normalisation and the dense push are
csrc/kernels/kmeans.cu + push_pull.cu.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Tuple

import torch

from ..ops import _lib
from ..parallel.device_table import KVClientTable
from ..parallel.types import ModelType
from ..utils.tracing import traced


def make_blobs(n: int, d: int, k_true: int, device="cuda", seed: int = 0, spread: float = 0.15,
               dtype=torch.bfloat16, chunk: int = 1 << 22) -> torch.Tensor:
    """Synthetic dense points: k_true Gaussian blobs, generated directly in device memory."""
    g = torch.Generator(device=device).manual_seed(seed)
    means = torch.randn(k_true, d, device=device, generator=torch.Generator(device=device).manual_seed(4242))
    out = torch.empty(n, d, dtype=dtype, device=device)
    for s in range(0, n, chunk):
        m = min(chunk, n - s)
        which = torch.randint(0, k_true, (m,), device=device, generator=g)
        out[s:s + m] = (means[which] + spread * torch.randn(m, d, device=device, generator=g)).to(dtype)
    return out


def padded_shape(K: int, D: int) -> Tuple[int, int]:
    """The tcgen05 tile shapes the kernel runs: K rounded up to a multiple of 128, D to 64 or 128."""
    if D > 128:
        raise ValueError(f"device k-means holds a point in one 128-wide operand row: num_dims={D} > 128 "
                         "(use the host backend, build/bin/kmeans)")
    return (K + 127) // 128 * 128, (64 if D <= 64 else 128)


def pad_points(x: torch.Tensor, dtype=torch.bfloat16) -> torch.Tensor:
    """[n, D] points -> [n, Dp] bf16 with zero columns (Dp = 64 or 128)."""
    n, D = x.shape
    _, Dp = padded_shape(1, D)
    if D == Dp and x.dtype == dtype:
        return x.contiguous()
    out = torch.zeros(n, Dp, dtype=dtype, device=x.device)
    out[:, :D] = x.to(dtype)
    return out


def csr_to_points(csr, num_dims: int) -> torch.Tensor:
    """Densify libsvm samples (reference kmeans.cpp:154-160 keeps them sparse and walks K*D per point;
    the tensor-core kernel wants rows): [n, Dp] bf16, feature ids outside [0, num_dims) dropped."""
    n = csr.n_rows
    _, Dp = padded_shape(1, num_dims)
    dev = csr.val.device
    lens = (csr.row_ptr[1:] - csr.row_ptr[:-1])
    row_of = torch.repeat_interleave(torch.arange(n, device=dev), lens)
    base = int(csr.row_ptr[0])
    col = csr.col[base:base + int(lens.sum())].to(torch.int64) & 0xFFFFFFFF
    val = csr.val[base:base + int(lens.sum())]
    keep = col < num_dims
    out = torch.zeros(n, Dp, dtype=torch.float32, device=dev)
    out.index_put_((row_of[keep], col[keep]), val[keep], accumulate=True)
    return out.to(torch.bfloat16)


class KMeansAssigner:
    """Scratch + launcher of the tensor-core assignment kernel.  K / D are the operand (padded) shape;
    K_valid / D_valid the real problem inside it (see padded_shape)."""

    def __init__(self, K: int, D: int, device, K_valid: int = 0, D_valid: int = 0):
        assert K % 128 == 0 and D in (64, 128), "tcgen05 path: K multiple of 128, D in {64, 128}"
        self.K, self.D, self.dev = K, D, device
        self.Kv, self.Dv = (K_valid or K), (D_valid or D)
        self.lib = _lib.load()
        # B-operand scratch: [K, D] bf16(-2c) | [K, 64] ||c||^2 as a 3-way bf16 split | [128, 64] ones
        self.c_bf16 = torch.empty(K * D + K * 64 + 128 * 64, dtype=torch.bfloat16, device=device)
        self.cnorm = torch.empty(K, dtype=torch.float32, device=device)
        self.c_f32 = torch.empty(K * D, dtype=torch.float32, device=device)
        self.sync = torch.zeros(4, dtype=torch.int32, device=device)

    def assign(self, x: torch.Tensor, centers: Optional[torch.Tensor] = None, client: Optional[KVClientTable] = None,
               out: Optional[Tuple[torch.Tensor, torch.Tensor]] = None, max_ctas: int = 0, stream=None):
        """x: [n, D] bf16.  Centres from `client`'s table (pull fused in-kernel, consistency gate
        applied) or from the dense fp32 tensor `centers` [K, D]."""
        assert x.dtype == torch.bfloat16 and x.is_contiguous() and x.shape[1] == self.D
        n = x.shape[0]
        if out is None:
            out = (torch.empty(n, dtype=torch.int32, device=self.dev), torch.empty(n, dtype=torch.float32,
                device=self.dev))
        assign, dist = out
        view = clk = None
        progress = 0
        cptr = None
        if client is not None:
            t = client.table
            assert t.num_keys == self.Kv * self.Dv and t.k == 1 and t.dtype == torch.float32
            client._local_gate()
            view, clk, progress = C.byref(t.view), C.byref(t.clock_view()), client.progress
        else:
            assert centers is not None and centers.dtype == torch.float32 and centers.numel() == self.Kv * self.Dv
            cptr = C.c_void_p(centers.data_ptr())
        _lib.check(self.lib.mps_kmeans_assign_tc(
            view, clk, progress, cptr, C.c_void_p(x.data_ptr()), n, self.K, self.D, self.Kv, self.Dv,
            C.c_void_p(self.c_bf16.data_ptr()), C.c_void_p(self.cnorm.data_ptr()), C.c_void_p(self.c_f32.data_ptr()),
            C.c_void_p(self.sync.data_ptr()), C.c_void_p(assign.data_ptr()), C.c_void_p(dist.data_ptr()), max_ctas,
            _lib.stream_ptr(stream)), "kmeans_assign_tc")
        return assign, dist

    def assign_ref(self, x: torch.Tensor, stream=None):
        """CUDA-core differential kernel against the bf16 centres of the LAST assign() call."""
        n = x.shape[0]
        assign = torch.empty(n, dtype=torch.int32, device=self.dev)
        dist = torch.empty(n, dtype=torch.float32, device=self.dev)
        _lib.check(self.lib.mps_kmeans_assign_ref(
            C.c_void_p(x.data_ptr()), C.c_void_p(self.c_bf16.data_ptr()), C.c_void_p(self.cnorm.data_ptr()), n,
            self.K, self.D, C.c_void_p(assign.data_ptr()), C.c_void_p(dist.data_ptr()), _lib.stream_ptr(stream)),
            "kmeans_assign_ref")
        return assign, dist


class KMeansTrainer:
    """One worker's view of the two tables + its local points.  K clusters of D coordinates, any K and any
    D <= 128 (reference defaults: K=5, num_dims=123, scripts/kmeans.py:17-43): the tables hold exactly K*D and
    K keys; the kernels run the padded operand shape.  x: [n, Dp] bf16 (pad_points / csr_to_points)."""

    def __init__(self, centers_client: KVClientTable, counts_client: KVClientTable, x: torch.Tensor, K: int,
                 alpha: float = 1.0, D: Optional[int] = None):
        self.tc, self.tn = centers_client, counts_client
        self.x = x
        self.Kv, self.Dv = K, (D if D is not None else x.shape[1])
        self.K, self.D = padded_shape(self.Kv, self.Dv)
        assert x.shape[1] == self.D, f"points must be padded to {self.D} columns (pad_points)"
        self.alpha = float(alpha)
        self.dev = centers_client.dev
        self.lib = centers_client.lib
        K, D = self.K, self.D
        self.assigner = KMeansAssigner(K, D, self.dev, self.Kv, self.Dv)
        self.sums = torch.zeros(K * self.D, dtype=torch.float32, device=self.dev)
        self.cnt = torch.zeros(K, dtype=torch.float32, device=self.dev)
        self.delta_c = torch.empty(K * self.D, dtype=torch.float32, device=self.dev)
        self.delta_n = torch.empty(K, dtype=torch.float32, device=self.dev)
        self.counts_old = torch.zeros(K, dtype=torch.float32, device=self.dev)
        self._out = None
        self._scratch = None
        self.allow_fused = True      # False forces the call-by-call path (tests, debugging)

    @traced("kmeans.step")
    def step(self, start: int, batch: int, sse: Optional[torch.Tensor] = None) -> torch.Tensor:
        """One iteration on rows [start, start+batch).  Returns the 1-element SSE tensor."""
        xb = self.x[start:start + batch]
        n = xb.shape[0]
        if self._out is None or self._out[0].numel() != n:
            self._out = (torch.empty(n, dtype=torch.int32, device=self.dev),
                         torch.empty(n, dtype=torch.float32, device=self.dev))
        if sse is None:
            sse = torch.zeros(1, dtype=torch.float32, device=self.dev)
        if self._fusable():
            return self._step_fused(xb, n, sse)
        # generic path (several local workers share the tables, or non-f32 tables):
        # Get of both tables: counts as a slab pull, centres inside the GEMM kernel
        self.tn.get_all(out=self.counts_old[:self.Kv])
        assign, dist = self.assigner.assign(xb, client=self.tc, out=self._out)
        self.sums.zero_()
        self.cnt.zero_()
        sp = self.tc._sp()
        _lib.check(self.lib.mps_kmeans_accumulate_sorted(
            C.c_void_p(xb.data_ptr()), C.c_void_p(assign.data_ptr()), C.c_void_p(dist.data_ptr()), n, self.K, self.D,
            C.c_void_p(self.sums.data_ptr()), C.c_void_p(self.cnt.data_ptr()), C.c_void_p(sse.data_ptr()),
            C.c_void_p(self._sort_scratch(n).data_ptr()), sp), "kmeans_accumulate_sorted")
        _lib.check(self.lib.mps_kmeans_delta(
            C.c_void_p(self.assigner.c_f32.data_ptr()), C.c_void_p(self.counts_old.data_ptr()),
            C.c_void_p(self.sums.data_ptr()), C.c_void_p(self.cnt.data_ptr()), self.K, self.D, self.alpha,
            C.c_void_p(self.delta_c.data_ptr()), C.c_void_p(self.delta_n.data_ptr()), sp), "kmeans_delta")
        # Add + Clock on both tables (kmeans.cpp:269-274); the tables hold the real K x D only
        if (self.Kv, self.Dv) != (self.K, self.D):
            self.tc.add_all(self.delta_c.view(self.K, self.D)[:self.Kv, :self.Dv].contiguous().view(-1))
            self.tn.add_all(self.delta_n[:self.Kv].contiguous())
        else:
            self.tc.add_all(self.delta_c)
            self.tn.add_all(self.delta_n)
        self.tc.clock()
        self.tn.clock()
        return sse

    def _sort_scratch(self, n: int) -> torch.Tensor:
        """Counting-sort scratch of mps_kmeans_accumulate_sorted: hist [K] (zero), base [K+1], cursor [K], perm [n]."""
        need = 3 * self.K + 1 + n
        if self._scratch is None or self._scratch.numel() < need:
            self._scratch = torch.zeros(need, dtype=torch.int32, device=self.dev)
        return self._scratch

    def _fusable(self) -> bool:
        tc, tn = self.tc.table, self.tn.table
        return (self.allow_fused and tc.dtype == torch.float32
                and tn.dtype == torch.float32 and tc.hview is None and tn.hview is None
                and tc.model_type == tn.model_type and self.tc.progress == self.tn.progress
                and self.tc.stream is self.tn.stream)

    def _step_fused(self, xb: torch.Tensor, n: int, sse: torch.Tensor) -> torch.Tensor:
        """The whole iteration as ONE host call (mps_kmeans_step): assign with the centre pull fused ->
        accumulate -> normalise + push of both tables fused -> Clock of both tables.  `sums` / `cnt`
        are cleared by the kernel that consumes them, so nothing else is launched."""
        tc, tn = self.tc.table, self.tn.table
        a = self.assigner
        assign, dist = self._out
        progress = self.tc.progress
        s = self.tc.stream or torch.cuda.current_stream()
        # several workers on this GPU: their clocks this step depends on are ordered on my stream by
        # events, never waited for inside a kernel (kernels of different streams need not co-run)
        self.tc._local_gate()
        self.tn._local_gate()
        with tc.lock, tn.lock:
            # who folds (BSP): the last local worker to clock, decided and published atomically like
            # KVClientTable.clock; it waits for the other local workers' adds of this clock
            fold = 0
            if tc.model_type == ModelType.BSP:
                others = [w for w in tc.local_workers if w != self.tc.worker_id]
                if all(tc.local_clock[w] >= progress + 1 for w in others):
                    fold = 1
                    for t in (tc, tn):
                        for w in others:
                            if w in t.local_events:
                                s.wait_event(t.local_events[w])
            _lib.check(self.lib.mps_kmeans_step(
                C.byref(tc.view), C.byref(tc.clock_view()), C.byref(tn.view), C.byref(tn.clock_view()),
                self.tc.worker_id, progress, fold, C.c_void_p(xb.data_ptr()), n, self.K, self.D, self.Kv, self.Dv,
                self.alpha, C.c_void_p(a.c_bf16.data_ptr()), C.c_void_p(a.cnorm.data_ptr()),
                C.c_void_p(a.c_f32.data_ptr()),
                C.c_void_p(a.sync.data_ptr()), C.c_void_p(assign.data_ptr()), C.c_void_p(dist.data_ptr()),
                C.c_void_p(self.sums.data_ptr()), C.c_void_p(self.cnt.data_ptr()), C.c_void_p(sse.data_ptr()),
                C.c_void_p(self._sort_scratch(n).data_ptr()), self.tc._sp()), "kmeans_step")
            ev = torch.cuda.Event()
            ev.record(s)
            for t, cl in ((tc, self.tc), (tn, self.tn)):
                t.local_clock[cl.worker_id] = progress + 1
                t.local_events[cl.worker_id] = ev
                t.write_epoch += 1
                if fold and t.own_count > 0:
                    t.fold_clock = progress + 1
                t.lock.notify_all()
                cl.progress = progress + 1
        return sse


# ------------------------------------------------------------------ initialisation (kmeans_helper.hpp)
def init_random(x: torch.Tensor, K: int, seed: int = 0) -> torch.Tensor:
    """K distinct random points (kmeans_helper.hpp:22-42)."""
    g = torch.Generator(device=x.device).manual_seed(seed)
    idx = torch.randperm(x.shape[0], device=x.device, generator=g)[:K]
    return x[idx].float()


def _min_sqdist(x: torch.Tensor, centers: torch.Tensor, chunk: int = 1 << 16) -> torch.Tensor:
    """min_c ||x - c||^2 for every point.  On the device this is the tcgen05 assignment kernel itself
    (the seeding loops of kmeans_helper.hpp:69-207 are this operation over and over); the PyTorch
    expression below is the CPU / oversized-K fallback and the oracle."""
    K, D = centers.shape
    if x.is_cuda and x.dtype == torch.bfloat16 and D <= 128 and K <= 8192 and x.shape[0] > 0:
        Kp, Dp = padded_shape(K, D)
        xp = x if (x.shape[1] == Dp and x.is_contiguous()) else pad_points(x[:, :D])
        km = KMeansAssigner(Kp, Dp, x.device, K, D)
        _, dist = km.assign(xp, centers=centers.float().contiguous())
        xf = xp.float()
        return (dist + (xf * xf).sum(1)).clamp_min(0)
    return _min_sqdist_torch(x, centers, chunk)


def _min_sqdist_torch(x: torch.Tensor, centers: torch.Tensor, chunk: int = 1 << 16) -> torch.Tensor:
    """The same quantity in plain PyTorch fp32 (oracle of the tests)."""
    out = torch.empty(x.shape[0], device=x.device)
    cf = centers.float()
    cn = (cf * cf).sum(1)
    for s in range(0, x.shape[0], chunk):
        xb = x[s:s + chunk].float()
        d = (xb * xb).sum(1, keepdim=True) - 2 * xb @ cf.T + cn
        out[s:s + chunk] = d.min(1).values.clamp_min(0)
    return out


def init_kmeanspp(x: torch.Tensor, K: int, seed: int = 0, sample: int = 1 << 18) -> torch.Tensor:
    """k-means++ D^2 sampling (kmeans_helper.hpp:69-109) on a uniform subsample of the points."""
    g = torch.Generator(device=x.device).manual_seed(seed)
    if x.shape[0] > sample:
        x = x[torch.randperm(x.shape[0], device=x.device, generator=g)[:sample]]
    n = x.shape[0]
    centers = x[torch.randint(0, n, (1,), device=x.device, generator=g)].float()
    d2 = _min_sqdist(x, centers)
    for _ in range(1, K):
        probs = d2 / d2.sum().clamp_min(1e-30)
        nxt = torch.multinomial(probs, 1, generator=g)
        c = x[nxt].float()
        centers = torch.cat([centers, c])
        xf = x.float()
        d2 = torch.minimum(d2, ((xf - c) ** 2).sum(1))
    return centers


def init_kmeans_parallel(x: torch.Tensor, K: int, seed: int = 0, rounds: int = 5, oversample: float = 2.0,
                         sample: int = 1 << 18) -> torch.Tensor:
    """k-means|| (kmeans_helper.hpp:112-207): a few rounds of oversampled D^2 draws, then the
    candidates are reduced to K by weighted k-means++."""
    g = torch.Generator(device=x.device).manual_seed(seed)
    if x.shape[0] > sample:
        x = x[torch.randperm(x.shape[0], device=x.device, generator=g)[:sample]]
    n = x.shape[0]
    cand = x[torch.randint(0, n, (1,), device=x.device, generator=g)].float()
    for _ in range(rounds):
        d2 = _min_sqdist(x, cand)
        p = (oversample * K * d2 / d2.sum().clamp_min(1e-30)).clamp_max(1.0)
        pick = torch.rand(n, device=x.device, generator=g) < p
        if pick.any():
            cand = torch.cat([cand, x[pick].float()])
    if cand.shape[0] <= K:
        extra = init_random(x, K - cand.shape[0] + 1, seed + 1)
        cand = torch.cat([cand, extra])[:max(K, cand.shape[0])]
    # weight = number of points closest to each candidate, then k-means++ on the candidates
    cf = cand
    xf_n = (cf * cf).sum(1)
    w = torch.zeros(cand.shape[0], device=x.device)
    for s in range(0, n, 1 << 16):
        xb = x[s:s + (1 << 16)].float()
        a = ((xb * xb).sum(1, keepdim=True) - 2 * xb @ cf.T + xf_n).argmin(1)
        w.index_add_(0, a, torch.ones_like(a, dtype=torch.float32))
    centers = cand[torch.multinomial(w / w.sum(), 1, generator=g)]
    d2 = ((cand - centers[0]) ** 2).sum(1)
    for _ in range(1, K):
        probs = w * d2
        nxt = torch.multinomial(probs / probs.sum().clamp_min(1e-30), 1, generator=g)
        centers = torch.cat([centers, cand[nxt]])
        d2 = torch.minimum(d2, ((cand - cand[nxt]) ** 2).sum(1))
    return centers


INIT_MODES = {"random": init_random, "kmeans++": init_kmeanspp, "kmeans_parallel": init_kmeans_parallel}


# ------------------------------------------------------------------ fp32 PyTorch oracle
def oracle_assign(x: torch.Tensor, centers: torch.Tensor):
    """Exact assignment on the bf16-rounded inputs in fp32: (assign, dist_min_without_xnorm, full matrix)."""
    xf = x.float()
    cf = centers.to(torch.bfloat16).float()
    d = (cf * cf).sum(1)[None, :] - 2.0 * xf @ cf.T
    return d.argmin(1), d.min(1).values, d


def oracle_step(centers: torch.Tensor, counts: torch.Tensor, batches, alpha: float):
    """Batched mini-batch update of ALL workers' batches against the same centres (BSP).
    centers [K, D] fp32, counts [K]; returns (new_centers, new_counts, sse)."""
    K, D = centers.shape
    dc = torch.zeros_like(centers)
    dn = torch.zeros_like(counts)
    sse = 0.0
    for xb in batches:
        a, dmin, _ = oracle_assign(xb, centers)
        xf = xb.float()
        sums = torch.zeros_like(centers).index_add_(0, a, xf)
        m = torch.zeros_like(counts).index_add_(0, a, torch.ones(xb.shape[0], device=xb.device))
        upd = alpha * (sums - m[:, None] * centers) / (counts + m).clamp_min(1e-30)[:, None]
        dc += torch.where(m[:, None] > 0, upd, torch.zeros_like(upd))
        dn += m
        sse += float(((xf * xf).sum(1) + dmin).clamp_min(0).sum())
    return centers + dc, counts + dn, sse
