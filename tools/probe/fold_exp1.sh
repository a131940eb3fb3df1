run() { echo -n "$* : "; env $ENVV python tools/c3bench.py --slabs 16 "$@" | python -c "import sys,json; r=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('ms/12.5M %.4f full %d' % (r['ms_per_12.5M'], r['counters']['dropped_table_full']))"; }
for f in 0 1; do
ENVV="BEHOLDER_FOLD=$f BEHOLDER_PART_LG=12" ; echo "== $ENVV"; run --int --no-hist --keys 10000000
ENVV="BEHOLDER_FOLD=$f BEHOLDER_PART_LG=11" ; echo "== $ENVV"; run
ENVV="BEHOLDER_FOLD=$f" ; echo "== $ENVV one window"; run --rate 100000; run --no-hist --rate 100000; run --int --no-hist --keys 10000000 --rate 100000
ENVV="BEHOLDER_FOLD=$f BEHOLDER_PART_LG=12" ; echo "== $ENVV one window"; run --int --no-hist --keys 10000000 --rate 100000
ENVV="BEHOLDER_FOLD=$f BEHOLDER_PART_LG=11" ; echo "== $ENVV one window"; run --rate 100000
done
