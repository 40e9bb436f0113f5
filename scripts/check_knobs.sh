#!/bin/bash
# Engine-vs-oracle tests under every fallback / tuning knob (run on a B200: gpurun -- bash scripts/check_knobs.sh)
for e in "CRACKFL_SIDE_STREAM=0" "CRACKFL_PDL=0" \
         "CRACKFL_DW_MERGED=0 CRACKFL_ENTRY_TC=0 CRACKFL_WGRAD_SPLIT=0 CRACKFL_POOL_BWD_FUSED=0 CRACKFL_HEAD_FUSED=0" \
         "CRACKFL_DW_TMA=0 CRACKFL_FUSE_FINALIZE=0" "CRACKFL_WGRAD_DEFERRED=1 CRACKFL_WGRAD_SPLIT=0" \
         "CRACKFL_WGRAD_CLUSTER=4 CRACKFL_WGRAD_SPLIT=0" "CRACKFL_TC_HALO=0" "CRACKFL_TC_HALO=2" "CRACKFL_TC_RESIDENT=0" \
         "CRACKFL_GEMM_STATS=0" "CRACKFL_DW_BN_FUSED=0"; do
  echo "$e: $(env $e timeout 300 python -m pytest tests/test_gpu_kernels.py -x -q -k engine 2>&1 | tail -1)"
done
