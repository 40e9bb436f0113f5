for k in "" weight_prep colsum bn_act head_bwd adam_step dec_combine down2sum entry_im2col pool_add pool_bwd bn_bwd_apply bn_bwd_reduce dw3x3_bwd dw3x3_tma wgrad relu_bwd_add; do
  CRACKFL_SKIP_KERNELS=$k python bench.py --steps 40 --warmup 5 --no-e2e 2>&1 | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('skip=$k', round(d['ms_per_step'],4), d['gpu_launches']//40)"
done
