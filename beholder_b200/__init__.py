"""beholder_b200 — B200-native implementation of tritonmedia/beholder's progress-event
consume-and-aggregate path (index.js:127-155) behind the reference's handler interface.

Layout: csrc/ (CUDA kernels + C-ABI -> libbeholder_gpu.so), _lib (ctypes binding),
engine (context wrapper), handler (mirror of amqp.listen / rmsg.ack / prom Counter),
schema (the assumed api.TelemetryProgress descriptor), synth (benchmark streams),
dist (multi-GPU window exchange), windows (watermark-driven window close), amqp (raw AMQP 0-9-1 delivery tap).
"""
from . import _lib
from ._lib import (BeholderError, MODE_REFERENCE, MODE_WINDOWED, OP_COUNT, OP_SUM, OP_MIN, OP_MAX, OP_HIST,
                   EV_OK, EV_DROP_DECODE, EV_DROP_ENUM, EV_DROP_NAN)
from .engine import Aggregator, PinnedBuffer, RECORD_DTYPE
from .handler import BeholderGPU, StubAMQP, StubRmsg
from .schema import Schema, Field, telemetry_progress, telemetry_status, encode
from . import metrics
from . import amqp
from .amqp import AmqpTap
from .synth import SynthSpec, zipf_cdf
from .windows import WindowStream

__all__ = [
    "Aggregator", "AmqpTap", "BeholderGPU", "BeholderError", "PinnedBuffer", "RECORD_DTYPE", "Schema", "Field",
    "StubAMQP", "StubRmsg", "SynthSpec", "WindowStream", "encode", "telemetry_progress", "zipf_cdf",
]
