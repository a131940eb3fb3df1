O=gpurun_out
run() { echo -n "$* : "; python tools/c3bench.py --slabs 16 "$@" | python -c "import sys,json; r=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('ms/12.5M %.4f full %d' % (r['ms_per_12.5M'], r['counters']['dropped_table_full']))"; }
run --no-hist --rate 100000; run --rate 100000; run --int --no-hist --keys 10000000 --rate 100000
ncu --set full --import-source on --clock-control none -k regex:k_part_fold -s 1 -c 1 -f -o $O/fold_nohist python tools/c3bench.py --slabs 12 --no-hist --rate 100000 > $O/fold_nohist.log 2>&1
ncu --set full --import-source on --clock-control none -k regex:k_part_fold -s 1 -c 1 -f -o $O/fold_hist python tools/c3bench.py --slabs 12 --rate 100000 > $O/fold_hist.log 2>&1
