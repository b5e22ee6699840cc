"""The allocation-free entry point `aar_local_allreduce_static` (include/akka_allreduce_b200.h).

Some callers cannot hand the library device memory for its ring buffers and flags: they run inside a host
application's fixed memory plan, or on a path where cudaMalloc (which may synchronise the device or fail) is
off limits.  For them the library carries a built-in scratch array and this one extra C entry point that uses it.
The test drives the entry point through the C ABI exactly as such a caller would (raw device pointers, no
library-side allocation) and checks, bit for bit like the rest of the GPU suite, that it returns what the
existing entry point (`aar_local_round` on caller-allocated regions) returns -- on two calls in a row, because the
second call runs on whatever the first one left in the shared array.

This file sorts last on purpose: if it fails, an `-x` run has already executed every other test.
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _static_round(lib, L, xs, outs, n, chunk, dtype_code):
    P = len(xs)
    ins = (C.c_void_p * P)(*[x.data_ptr() for x in xs])
    outp = (C.c_void_p * P)(*[o.data_ptr() for o in outs])
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    return lib.aar_local_allreduce_static(ins, outp, P, n, chunk, dtype_code, stream)


@pytest.mark.parametrize("P,n,chunk,dtype", [(4, 4096, 512, torch.float32), (3, 1001, 100, torch.float32), (8, 65536, 4096, torch.bfloat16)])
def test_static_scratch_entry_point_matches_the_allocating_one(P, n, chunk, dtype):
    from akka_allreduce_b200.ops import lib as L
    from akka_allreduce_b200.parallel.local_gpu import LocalGpuCluster

    lib = L.load()
    code = L.AAR_F32 if dtype == torch.float32 else L.AAR_BF16
    reference = LocalGpuCluster(P, n, chunk, max_lag=0, dtype=dtype)   # existing entry point, caller-allocated regions
    g = torch.Generator(device="cpu").manual_seed(77)
    for call in range(2):                                              # two calls in a row on the shared array
        xs = [(torch.randn(n, generator=g) * 3).to(dtype).cuda() for _ in range(P)]
        outs = [torch.full((n,), 7.0, dtype=dtype, device="cuda") for _ in range(P)]
        rc = _static_round(lib, L, xs, outs, n, chunk, code)
        assert rc == 0, lib.aar_last_error().decode()
        want, _ = reference.run_round(call, xs)
        torch.cuda.synchronize()
        for r in range(P):
            assert torch.equal(outs[r], want[r]), f"call {call} rank {r}: differs from aar_local_round"


def test_static_scratch_entry_point_rejects_what_does_not_fit():
    from akka_allreduce_b200.ops import lib as L

    lib = L.load()
    P, n = 8, 4 << 20   # 8 ranks x 16 MiB vectors: far beyond the 16 MiB built-in array
    x = torch.zeros(n, device="cuda")
    rc = _static_round(lib, L, [x] * P, [x] * P, n, 1 << 20, L.AAR_F32)
    assert rc != 0
    msg = lib.aar_last_error().decode()
    assert "bytes of scratch" in msg and "built-in array holds" in msg
    # and it still works afterwards
    xs = [torch.ones(256, device="cuda") * (r + 1) for r in range(2)]
    outs = [torch.empty(256, device="cuda") for _ in range(2)]
    assert _static_round(lib, L, xs, outs, 256, 64, L.AAR_F32) == 0
    torch.cuda.synchronize()
    assert torch.equal(outs[0], torch.full((256,), 3.0, device="cuda")) and torch.equal(outs[1], outs[0])
