"""H2D copy bandwidth from pinned host memory: torch's pin_memory() vs cudaHostAlloc write-combined,
one large copy vs the loader's four-array batch (sizes of one LR step)."""
import ctypes as C
import sys

import torch

rt = C.CDLL("libcudart.so")
rt.cudaHostAlloc.argtypes = [C.POINTER(C.c_void_p), C.c_size_t, C.c_uint]


def wc_pinned(nbytes):
    p = C.c_void_p()
    rc = rt.cudaHostAlloc(C.byref(p), nbytes, 4)  # cudaHostAllocWriteCombined
    assert rc == 0, rc
    buf = (C.c_uint8 * nbytes).from_address(p.value)
    return torch.frombuffer(buf, dtype=torch.uint8)


def bw(host_parts, dev_parts, iters=50):
    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        for _ in range(5):
            for h, d in zip(host_parts, dev_parts):
                d.copy_(h, non_blocking=True)
        st.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(st)
        for _ in range(iters):
            for h, d in zip(host_parts, dev_parts):
                d.copy_(h, non_blocking=True)
        e1.record(st)
        st.synchronize()
    nbytes = sum(h.numel() * h.element_size() for h in host_parts)
    return nbytes * iters / (e0.elapsed_time(e1) * 1e-3) / 1e9


torch.cuda.init()
torch.zeros(1, device="cuda")
sizes = [65536, 6_553_600, 10_485_760, 65536]  # row_len, packed ids, values, labels of one step
for name, alloc in (("pin_memory", lambda n: torch.empty(n, dtype=torch.uint8).pin_memory()), ("write-combined",
    wc_pinned)):
    host = [alloc(n) for n in sizes]
    for h in host:
        h.fill_(1)
    dev = [torch.empty(n, dtype=torch.uint8, device="cuda") for n in sizes]
    one_h, one_d = alloc(sum(sizes)), torch.empty(sum(sizes), dtype=torch.uint8, device="cuda")
    one_h.fill_(1)
    big_h, big_d = alloc(256 << 20), torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    big_h.fill_(1)
    print(f"{name:15s} pinned={host[0].is_pinned()}  4 arrays/step: {bw(host, dev):6.1f} GB/s   one 17 MB copy: "
          f"{bw([one_h], [one_d]):6.1f} GB/s   one 256 MB copy: {bw([big_h], [big_d], 10):6.1f} GB/s")
sys.exit(0)
