#!/bin/bash
# Multi-GPU evidence in one gpurun call (usage: scripts/multigpu_round.sh NGPU [full]): cross-GPU tests at this world size,
# the aggregation sweep (correctness on the whole arena + aligned timing, BASELINE config 5), kernel variants, bench.py.
N=${1:-2}
FULL=${2:-full}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
echo "== tests (world $N)"
timeout 400 python -m pytest tests/test_gpu_federation.py -m gpu -x -q -k "across_gpus and -$N" 2>&1 | tail -3
SIZES="1,8.233,64,256,1024,4096"
echo "== agg_sweep P2P (default kernel selection)"
timeout 600 $TR --master-port 29611 bench/agg_sweep.py --sizes-mb $SIZES --iters 8 --out gpurun_out/agg_sweep_${N}gpu.json 2>&1 | grep -E "^\{|AGG_SWEEP" | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print(d['payload'], d['strategy'], d['mb'], round(d['ms_best'], 4), round(d['roofline_frac_900'], 3), d['ok'])
    else: print(l.strip())"
if [ "$FULL" = "full" ]; then
echo "== agg_sweep NVLS"
CRACKFL_NVLS=1 timeout 600 $TR --master-port 29612 bench/agg_sweep.py --sizes-mb 8.233,64,256,1024 --iters 8 --payloads fp32 --out gpurun_out/agg_sweep_${N}gpu_nvls.json 2>&1 | grep -E "^\{|AGG_SWEEP" | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print('nvls', d['strategy'], d['mb'], round(d['ms_best'], 4), round(d['roofline_frac_900'], 3), d['ok'])
    else: print(l.strip())"
for v in "CRACKFL_AGG_TMA=0" "CRACKFL_AGG_TMA=1"; do
  echo "== points $v"
  env $v timeout 300 $TR --master-port 29613 bench/agg_points.py --sizes-mb 8.233,64,256,1024 --tag "$v" --out gpurun_out/agg_points_${N}gpu_${v##*=}.json 2>&1 | grep "^{" | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l); print(d['mb'], d['ms_best'], d['ms_median'], round(d['roofline_frac_900'], 3))"
done
echo "== latency"
timeout 200 $TR --master-port 29614 bench/agg_latency.py 2>&1 | grep "^{" | tee gpurun_out/agg_latency_${N}gpu.json
fi
echo "== bench"
timeout 300 $TR --master-port 29615 bench.py --gpus $N --steps 20 --warmup 5 2>/dev/null | tail -1 | tee gpurun_out/bench_${N}gpu.json
