"""Closing policy of the windowed extension (the reference's counter is cumulative and has no
windows, index.js:30-34): tumbling windows are exported and recycled as the stream's watermark passes them, so a
stream of any length runs in a ring of a few window tables."""
from __future__ import annotations

from typing import Callable, Optional


class WindowStream:
    """Feeds batches to an Aggregator and closes windows behind the watermark.

    `on_close(window_id, stream)` must export the window wherever its rows are wanted and recycle its table (e.g.
    `export_window_device` + `close_window_device`, or a multi-GPU `close_window`).  `watermark_ms` passed with a
    batch promises that every LATER event has a timestamp at or above it (arrival time minus the largest
    out-of-order delay); windows that end at or before it can no longer change."""

    def __init__(self, agg, window_ms: int, on_close: Callable[[int, int], None]):
        self.agg, self.window_ms, self.on_close = agg, window_ms, on_close
        self.next_open: Optional[int] = None     # lowest window id not yet closed
        self.closed = 0

    def process_device(self, d_payload, d_offsets, d_ts, n: int, *, first_ts_ms: int, watermark_ms: int, stream: int = 0) -> None:
        if self.next_open is None:
            self.next_open = first_ts_ms // self.window_ms
        self.agg.process_device(d_payload, d_offsets, d_ts, n, stream=stream)
        self.advance(watermark_ms, stream)

    def advance(self, watermark_ms: int, stream: int = 0) -> None:
        done = watermark_ms // self.window_ms     # windows below `done` are complete
        while self.next_open is not None and self.next_open < done:
            self.on_close(self.next_open, stream)
            self.next_open += 1
            self.closed += 1

    def finish(self, last_ts_ms: int, stream: int = 0) -> None:
        """end of stream: close everything up to and including the window of the last timestamp"""
        self.advance((last_ts_ms // self.window_ms + 1) * self.window_ms, stream)
