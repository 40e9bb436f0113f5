for v in "X=1" "CRACKFL_ADAM_EARLY=0" "CRACKFL_DW_FWD=2" "CRACKFL_DW_BWD=3" "CRACKFL_PREP_FLAT=0" "X=2"; do
  echo "== $v"; env $v timeout 200 python -m pytest tests/test_gpu_federation.py -m gpu -x -q -k three_rounds 2>&1 | grep -E "AssertionError|passed|failed" | cut -c1-400
done
