"""The public API on one GPU: AllreduceMaster + one AllreduceWorker(backend="cuda"), with the round loop in Python and in C++
(csrc/core/device_loop.cpp, launching through aar_engine_enqueue_ex without the interpreter).  DataWrapper.scala:3-7,
AllreduceWorker.scala:84-104, :253-268."""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("native_loop", [False, True])
@pytest.mark.parametrize("dtype", ["fp32", "bf16"])
def test_master_and_one_cuda_worker_end_to_end(native_loop, dtype):
    from akka_allreduce_b200 import AllreduceMaster, AllreduceWorker, AllReduceInput

    n, chunk, rounds = 1 << 18, 1 << 14, 16
    td = torch.float32 if dtype == "fp32" else torch.bfloat16
    master = AllreduceMaster(1, 1.0, 1.0, 1.0, 2, n, rounds - 1, chunk, pipeline_depth=2)
    master.start(60)
    base = torch.arange(n, device="cuda:0", dtype=torch.float32) % 251
    seen, bad = [], []

    def source(req):
        return AllReduceInput((base + req.iteration).to(td))

    def sink(out):
        want = (base + out.iteration).to(td)
        if not torch.equal(out.data, want) or int(out.count.min()) != 1:
            bad.append(out.iteration)
        seen.append(out.iteration)

    w = AllreduceWorker(source, sink, backend="cuda", device=0, dtype=dtype, session=f"t_api_{dtype}_{int(native_loop)}", native_loop=native_loop)
    assert w.run("127.0.0.1", master.port, timeout_sec=60)
    assert master.join(30)
    m = w.metrics()
    assert sorted(seen) == list(range(rounds)) and not bad
    assert m["rounds_delivered"] == rounds and m["max_scattered"] == rounds - 1 and m["launches"] == rounds
    if native_loop:
        assert w._direct, "the C++ loop must launch through the engine's C entry point, not through a Python thunk"
    w.close()
