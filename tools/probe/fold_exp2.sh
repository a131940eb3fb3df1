run() { echo -n "$ENVV $* : "; env $ENVV python tools/c3bench.py --slabs 16 "$@" | python -c "import sys,json; r=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('ms/12.5M %.4f full %d' % (r['ms_per_12.5M'], r['counters']['dropped_table_full']))"; }
ENVV="BEHOLDER_FOLD=1"; run --int --no-hist --keys 10000000; run --no-hist --keys 10000000; run --no-hist; run
ENVV="BEHOLDER_FOLD=0"; run --no-hist --keys 10000000; run --keys 10000000
ENVV="BEHOLDER_FOLD=1"; run --keys 10000000
