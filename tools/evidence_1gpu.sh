#!/bin/bash
# Everything the round's 1-GPU evidence under profiles/ comes from, in one gpurun call.  usage: tools/evidence_1gpu.sh <tag>
# (run from the repository root on a GPU box; writes gpurun_out/<tag>_*).
set -u
T=${1:-r02h}
O=gpurun_out
mkdir -p $O
python __graft_entry__.py smoke > $O/${T}_smoke.log 2>&1; tail -4 $O/${T}_smoke.log
python -m pytest tests -m gpu -q 2>&1 | tail -4 > $O/${T}_gputest_1gpu.log; cat $O/${T}_gputest_1gpu.log
python bench.py --steps 20 --warmup 5 2> $O/${T}_bench_stderr.log | tail -1 > $O/${T}_bench_1gpu.json
python bench.py --impl reference --steps 2 --warmup 1 2>> $O/${T}_bench_stderr.log | tail -1 > $O/${T}_bench_reference.json
python tools/conv_table.py > $O/${T}_conv_table.log 2>&1
B200_CUDA_GRAPHS=0 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file $O/${T}_launches.csv \
  python bench.py --steps 2 --warmup 3 --no-bert --no-vgg --no-cpu-baseline > $O/${T}_ncu_launches.log 2>&1
python tools/launch_summary.py $O/${T}_launches.csv 3 > $O/${T}_launch_summary.txt 2>&1; head -12 $O/${T}_launch_summary.txt
ncu --set full --clock-control none --profile-from-start off -o /tmp/${T}_named python tools/named_launches.py > $O/${T}_named_ncu.log 2>&1
ncu -i /tmp/${T}_named.ncu-rep --page raw --csv > $O/${T}_named_raw.csv 2>/dev/null
python tools/ncu_named_summary.py $O/${T}_named_raw.csv $O/named_launches_labels.json $O/${T}_ncu_named_launches.json > $O/${T}_ncu_named_launches.txt 2>&1
head -45 $O/${T}_ncu_named_launches.txt
