#!/bin/bash
# Round-end evidence on one GPU: full GPU test-suite, bench line, per-launch ncu list of one step, what-if table.
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python bench.py --steps 60 --warmup 5 | tail -1 | tee gpurun_out/bench_n1_final.json | cut -c1-400
for s in 128 448; do python bench.py --steps 40 --warmup 5 --size $s --no-e2e | tail -1 > gpurun_out/bench_n1_size$s.json; done
python bench.py --steps 3 --warmup 3 --no-e2e > /dev/null 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_v12.csv python bench.py --steps 3 --warmup 3 --no-e2e > gpurun_out/ncu_v12.log 2>&1
python scripts/step_list.py gpurun_out/launches_v12.csv -v > gpurun_out/step_launches_v12.txt 2>&1; tail -30 gpurun_out/step_launches_v12.txt
bash scripts/whatif.sh 2>&1 | tee gpurun_out/whatif_v12.txt
