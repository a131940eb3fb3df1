import sys, time, types, torch
sys.path.insert(0, "/root/repo")
import bench
args = types.SimpleNamespace(c3_events=200_000_000)
st = torch.cuda.Stream(); torch.cuda.set_stream(st)
orig = bench.many_key_pass
def timed(*a, **k):
    t0 = time.perf_counter(); r = orig(*a, **k); t1 = time.perf_counter()
    torch.cuda.synchronize(); t2 = time.perf_counter()
    print("host enqueue ms", (t1 - t0) * 1e3, "until GPU done ms", (t2 - t0) * 1e3)
    return r
bench.many_key_pass = timed
for i in range(2):
    out = bench.c3_leg(args, 0, st.cuda_stream, 6560.0)
    print(out["ms"], out["value"], out["windows_closed"])
