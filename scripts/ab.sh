# usage: scripts/ab.sh "ENV1=.. ENV2=.." "ENV=.." ...   one bench line per argument
for v in "$@"; do
  r=$(env $v python bench.py --steps 60 --warmup 5 --no-e2e 2>&1 | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(d['ms_per_step'],4))")
  echo "$v -> $r"
done
