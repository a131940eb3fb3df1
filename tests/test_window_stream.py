"""Host logic of the watermark-driven window close (beholder_b200.windows.WindowStream) with a stand-in
aggregator: which windows are closed, when, and in which order.  No GPU."""
from beholder_b200.windows import WindowStream


class FakeAgg:
    def __init__(self):
        self.batches = []

    def process_device(self, p, o, t, n, stream=0):
        self.batches.append(n)


def test_windows_close_behind_the_watermark_in_order_and_once():
    closed = []
    ws = WindowStream(FakeAgg(), 10_000, lambda w, st: closed.append(w))
    t0 = 1_700_000_003_000                      # inside window 170000000
    # batches of 4 s of stream time, 2 s of allowed lateness
    for k in range(10):
        first = t0 + 4000 * k - 2000
        ws.process_device(None, None, None, 100, first_ts_ms=first, watermark_ms=t0 + 4000 * (k + 1) - 2000)
    w0 = (t0 - 2000) // 10_000
    assert closed == list(range(w0, (t0 + 40_000 - 2000) // 10_000))          # everything strictly below the watermark's window
    assert ws.closed == len(closed) and ws.agg.batches == [100] * 10
    ws.finish(t0 + 40_000 + 2000)
    assert closed == list(range(w0, (t0 + 42_000) // 10_000 + 1)) and len(set(closed)) == len(closed)


def test_a_watermark_that_does_not_move_closes_nothing():
    closed = []
    ws = WindowStream(FakeAgg(), 1000, lambda w, st: closed.append(w))
    ws.process_device(None, None, None, 1, first_ts_ms=5500, watermark_ms=5500)
    ws.advance(5999)
    assert closed == []
    ws.advance(6000)
    assert closed == [5]
