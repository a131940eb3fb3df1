O=gpurun_out
ncu --set full --import-source on --clock-control none -k regex:k_part_fold -s 1 -c 1 -f -o $O/fold_10m python tools/c3bench.py --slabs 12 --int --no-hist --keys 10000000 > $O/fold_10m.log 2>&1
BEHOLDER_FOLD=0 ncu --set full --import-source on --clock-control none -k regex:k_part_drain -s 1 -c 1 -f -o $O/drain_10m python tools/c3bench.py --slabs 12 --int --no-hist --keys 10000000 > $O/drain_10m.log 2>&1
