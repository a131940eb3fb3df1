#!/bin/bash
# A/B of the two fold kernels (BEHOLDER_FOLD=0: hashed front table + HBM atomics; 1: slice mirror) on the many-key shapes.
#   bash tools/foldbench.sh [extra c3bench args]
for cfg in "" "--zipf" "--no-hist" "--int --no-hist" "--int --no-hist --keys 10000000"; do
  for f in 0 1; do
    echo -n "fold=$f $cfg $* : "
    BEHOLDER_FOLD=$f python tools/c3bench.py $cfg --slabs 16 "$@" | python -c "import sys,json; r=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('ms/12.5M %.4f  ev/s %.3e  full %d win %d' % (r['ms_per_12.5M'], r['events_per_s'], r['counters']['dropped_table_full'], r['counters']['dropped_window']))"
  done
done
