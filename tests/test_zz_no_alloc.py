"""mps_lr_train_dense_noalloc — the training kernel for callers that cannot allocate.

Some users of libminips_kernels.so cannot give the dense LR kernel the two scratch vectors it
needs: their own memory pool already owns the GPU's free memory, or they sit in a loop where
an allocation (and the device-wide sync that comes with it) is not acceptable, or they are a
bare C program.  For them the library carries a fixed 8 MiB scratch array of its own.  This
test drives that entry point the way such a caller would — straight through the C ABI with
ctypes, no wrapper — and checks it against the ordinary entry point (caller-provided scratch)
on identical tables and data, twice in a row: the second call runs on the array as the first
call left it, which is what every later call in a real process sees.

The file name sorts last on purpose: if this addition ever breaks, a `pytest -x` run has
already exercised everything else.
"""
import ctypes as C

import pytest
import torch

from minips_b200.parallel.types import MLTask, ModelType, Node, WorkerAlloc

pytestmark = pytest.mark.gpu

MAX_KEYS = 1 << 20  # MPS_NOALLOC_MAX_KEYS in include/minips_kernels.h


def _entry(lib):
    from minips_b200.ops import _lib
    fn = lib.mps_lr_train_dense_noalloc
    P = C.POINTER
    fn.argtypes = [P(_lib.TableView), P(_lib.ClockView), C.c_int, C.c_int, P(_lib.Csr), C.c_void_p, C.c_int,
                   C.c_int64, C.c_float, C.c_void_p, C.c_int, C.c_void_p]
    fn.restype = C.c_int
    return fn


def test_noalloc_matches_regular_entry_point_twice():
    from minips_b200.models.lr import SparseLRTrainer, make_synthetic
    from minips_b200.ops import _lib
    from minips_b200.parallel.engine import Engine

    node = Node(0, "localhost", 0)
    engine = Engine(node, [node], device=0)
    engine.start_everything(num_server_threads_per_node=8)
    try:
        num_dims, n_rows, batch, alpha = 300_007, 30_000, 8192, 0.05
        data = make_synthetic(n_rows, num_dims, 40, seed=11)
        tid_ref = engine.create_table(num_dims=num_dims, model_type=ModelType.BSP)
        tid_new = engine.create_table(num_dims=num_dims, model_type=ModelType.BSP)
        tid_big = engine.create_table(num_dims=MAX_KEYS + 8, model_type=ModelType.ASP)
        starts = [torch.tensor([0, 9000, 21_000], dtype=torch.int64, device="cuda"),
                  torch.tensor([100, 20_000], dtype=torch.int64, device="cuda")]
        out = {"ref_w": [], "new_w": [], "ref_loss": [], "new_loss": []}

        def fn(info):
            lib = _lib.load()
            entry = _entry(lib)
            ref = SparseLRTrainer(info.create_kv_client_table(tid_ref), data, alpha)
            client = info.create_kv_client_table(tid_new)
            table = engine.tables[tid_new]
            csr = data.c_struct()
            clock = 0
            for st in starts:  # two calls in a row on each entry point
                out["ref_loss"].append(ref.train_dense(st, batch).cpu())
                out["ref_w"].append(ref.client.get_all().cpu())
                n = st.numel()
                loss = torch.zeros(n, 2, dtype=torch.float32, device="cuda")
                cv = table.clock_view()
                rc = entry(C.byref(table.view), C.byref(cv), client.worker_id, clock, C.byref(csr),
                           C.c_void_p(st.data_ptr()), n, batch, alpha, C.c_void_p(loss.data_ptr()), 0,
                           _lib.stream_ptr())
                assert rc == 0, lib.mps_error_string(rc)
                clock += n
                client.progress = clock  # the kernel advanced the device-side clock itself
                out["new_loss"].append(loss.cpu())
                out["new_w"].append(client.get_all().cpu())
            # a table beyond the built-in scratch is refused, nothing is launched
            big = engine.tables[tid_big]
            cvb = big.clock_view()
            loss = torch.zeros(1, 2, dtype=torch.float32, device="cuda")
            rc = entry(C.byref(big.view), C.byref(cvb), 0, 0, C.byref(csr), C.c_void_p(starts[0].data_ptr()), 1,
                       batch, alpha, C.c_void_p(loss.data_ptr()), 0, _lib.stream_ptr())
            out["big_rc"] = rc
            torch.cuda.synchronize()
            out["big_loss"] = loss.cpu()

        task = MLTask()
        task.set_worker_alloc([WorkerAlloc(0, 1)])
        task.set_tables([tid_ref, tid_new, tid_big])
        task.set_lambda(fn)
        engine.run(task)
    finally:
        engine.stop_everything()

    for call in range(2):
        torch.testing.assert_close(out["new_w"][call], out["ref_w"][call], rtol=2e-4, atol=2e-5)
        torch.testing.assert_close(out["new_loss"][call][:, 0], out["ref_loss"][call][:, 0], rtol=2e-3, atol=1e-2)
        assert (out["new_loss"][call][:, 1] - out["ref_loss"][call][:, 1]).abs().max() <= 2
        assert out["ref_w"][call].abs().max() > 0
    assert out["big_rc"] == 1  # cudaErrorInvalidValue
    assert float(out["big_loss"].abs().sum()) == 0.0
