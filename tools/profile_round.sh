#!/bin/bash
# ncu evidence for profiles/ (run on the GPU box through gpurun, one GPU; every command has run without ncu first).
#   bash tools/profile_round.sh r02
R=${1:-r02}; O=gpurun_out
mkdir -p $O
# 1. launch list of the bench command (kernel shares of the step)
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/${R}_launches.csv python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu --no-extra > $O/${R}_ncu_bench.log 2>&1
python tools/launch_summary.py $O/${R}_launches.csv > $O/${R}_launch_list_summary.txt
# 2. the dominant kernel, full set
ncu --set full --import-source on --clock-control none -k regex:k_process_raw -s 6 -c 1 -f -o $O/${R}_fused python tools/kbench.py > $O/${R}_ncu_fused.log 2>&1
# 3. the many-key fold kernel, full set (C3 shape)
ncu --set full --import-source on --clock-control none -k regex:k_part_drain -s 2 -c 1 -f -o $O/${R}_drain python tools/c3bench.py --slabs 12 > $O/${R}_ncu_drain.log 2>&1
# 3b. the slice-mirror fold, full set (10 M keys, 2^24 slots: the shape that selects it)
ncu --set full --import-source on --clock-control none -k regex:k_part_fold -s 1 -c 1 -f -o $O/${R}_fold python tools/c3bench.py --slabs 12 --int --no-hist --keys 10000000 > $O/${R}_ncu_fold.log 2>&1
# 4. DRAM traffic of every kernel of a C3 pass
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:'^k_(process|part|export|close)' --csv --log-file $O/${R}_c3_traffic.csv python tools/c3bench.py --slabs 16 > $O/${R}_ncu_c3.log 2>&1
ls -la $O | tail -12
