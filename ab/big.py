import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import beholder_b200 as bh
from beholder_b200 import _lib as L
from beholder_b200.synth import SynthSpec, generate_device
st = torch.cuda.Stream(); torch.cuda.set_stream(st); s = st.cuda_stream
out = {}
for payload, n in ((128, 6553600), (256, 3276800), (512, 1638400), (1024, 819200)):
    spec = SynthSpec(n_keys=1000, rate_per_ms=500, payload_bytes=payload)
    p, o, t, total = generate_device(spec, 0, n)
    g = bh.Aggregator(bh.telemetry_progress(), mode=L.MODE_WINDOWED, table_capacity=2048, window_ring=32, message_bytes_hint=payload)
    for _ in range(3): g.process_device(p, o, t, n, stream=s)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10): g.process_device(p, o, t, n, stream=s)
    e1.record(); torch.cuda.synchronize()
    out[payload] = round(e0.elapsed_time(e1) / 10, 4)
print(os.environ.get("BEHOLDER_GPU_LIB", "head"), out)
