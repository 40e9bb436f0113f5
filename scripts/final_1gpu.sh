#!/bin/bash
# Round-end evidence on one GPU: full GPU test-suite, bench lines (256 / 128 / 448), activation bytes of the schedule,
# per-launch ncu list of one step, full ncu captures of the two depthwise kernels (CSV exports only).
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python bench.py --steps 60 --warmup 5 | tail -1 | tee gpurun_out/bench_n1_final.json | cut -c1-300
for s in 128 448; do python bench.py --steps 40 --warmup 5 --size $s --no-e2e | tail -1 > gpurun_out/bench_n1_size$s.json; done
python scripts/step_bytes.py > gpurun_out/step_bytes.log 2>&1; tail -3 gpurun_out/step_bytes.log
python bench.py --steps 3 --warmup 3 --no-e2e > /dev/null 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_v13.csv python bench.py --steps 3 --warmup 3 --no-e2e > gpurun_out/ncu_v13.log 2>&1
python scripts/step_list.py gpurun_out/launches_v13.csv -v > gpurun_out/step_launches_v13.txt 2>&1; tail -28 gpurun_out/step_launches_v13.txt
bash scripts/ncu_many.sh < scripts/ncu_dw.txt
for t in dwbwd2_128 dwfwd_128; do python scripts/ncu_src_top.py gpurun_out/prof_${t}_src.csv gpurun_out/prof_${t}_raw.csv > gpurun_out/top_$t.txt 2>&1; done
rm -f gpurun_out/prof_*_src.csv
