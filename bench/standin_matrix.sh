#!/bin/bash
# Labelled comparison arms on ONE B200: this framework vs the same graph in stock PyTorch/cuDNN
# (eager, CUDA-graph replayed, torch.compile) in bf16 and fp32, at 128^2 / 256^2 / 448^2, B = 16.  One JSON line per arm.
OUT=${1:-gpurun_out/standin_matrix.jsonl}
mkdir -p "$(dirname "$OUT")"; : > "$OUT"
for SIZE in 128 256 448; do
  python bench.py --size $SIZE --steps 40 --warmup 5 2>/dev/null | tail -1 >> "$OUT"
  for ARM in "eager bf16" "graph bf16" "graph fp32" "eager fp32" "compile bf16"; do
    set -- $ARM
    timeout 600 python bench.py --impl standin --standin-mode $1 --standin-dtype $2 --size $SIZE --steps 40 --warmup 5 2>gpurun_out/standin_err.log | tail -1 >> "$OUT" \
      || echo "{\"impl\": \"standin\", \"mode\": \"$1\", \"dtype\": \"$2\", \"size\": $SIZE, \"failed\": true}" >> "$OUT"
  done
done
python - "$OUT" <<'PY'
import json, sys
for l in open(sys.argv[1]):
    try:
        d = json.loads(l)
    except Exception:
        continue
    if "value" not in d:
        print(d); continue
    print(f"{d['config']['image']:>8s} {d['label'][:60]:60s} {d['value']:10.0f} img/s  e2e {d['e2e']['value'] if d.get('e2e') else 0:10.0f}  {d['ms_per_step']:.3f} ms")
PY
