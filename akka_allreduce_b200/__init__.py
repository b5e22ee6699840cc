"""B200-native threshold allreduce: an AllreduceMaster that admits workers and drives numbered rounds, AllreduceWorkers
that pull a dataSize buffer from a caller source and push the reduced result to a caller sink, maxChunkSize chunking,
thAllreduce / thReduce / thComplete thresholds and a maxLag window of in-flight rounds -- the capabilities of
mike199515/akka-allreduce-1, with the data plane as hand-written sm_100a kernels over NVLink.

    from akka_allreduce_b200 import AllreduceMaster, AllreduceWorker, AllReduceInput, AllReduceInputRequest, AllReduceOutput
"""
from .api import AllReduceInput, AllReduceInputRequest, AllReduceOutput, DataSink, DataSource  # noqa: F401


def __getattr__(name):
    # Lazy: the native modules are built in-tree; importing the package must not require them to exist yet
    # (the build helper lives inside the package).
    if name in ("AllreduceMaster", "load_settings"):
        from . import master

        return getattr(master, name)
    if name == "AllreduceWorker":
        from .worker import AllreduceWorker

        return AllreduceWorker
    if name == "_core":
        import importlib

        return importlib.import_module("._core", __name__)
    raise AttributeError(name)


__all__ = ["AllreduceMaster", "AllreduceWorker", "AllReduceInput", "AllReduceInputRequest", "AllReduceOutput", "load_settings"]
