"""Caller-facing records (parity with DataWrapper.scala:3-7).

    data_source: AllReduceInputRequest -> AllReduceInput
    data_sink:   AllReduceOutput -> None

`data` is a NumPy array on the CPU data plane and a CUDA tensor (fp32 or bf16) on the device data plane.
`count` really carries the per-chunk contributor counts (owner-major, chunk-minor; 0 = chunk never arrived);
the reference declares the field but always passes Array(0) (AllreduceWorker.scala:191).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Any, Callable


@dataclass
class AllReduceInputRequest:
    iteration: int
    # Device plane, optional (AllreduceWorker(zero_copy_input=True) on the in-switch path): a CUDA tensor of dataSize elements
    # inside the engine's staging area.  A source that writes its data there and returns that same tensor as
    # AllReduceInput.data saves one device-side copy; any other tensor is still accepted.  None everywhere else.
    buffer: Any = None


@dataclass
class AllReduceInput:
    data: Any
    # Device plane, optional: the tensor the reduced vector is written to.  None = the worker's own output buffer of the round's
    # ring slot; the SAME tensor as `data` = in place (gradient buckets: no copy on either side of the allreduce).
    out: Any = None


@dataclass
class AllReduceOutput:
    data: Any
    count: Any
    iteration: int


DataSource = Callable[[AllReduceInputRequest], AllReduceInput]
DataSink = Callable[[AllReduceOutput], None]
