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


@dataclass
class AllReduceInput:
    data: Any


@dataclass
class AllReduceOutput:
    data: Any
    count: Any
    iteration: int


DataSource = Callable[[AllReduceInputRequest], AllReduceInput]
DataSink = Callable[[AllReduceOutput], None]
