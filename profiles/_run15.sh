set -u
echo "all gpu tests (smem tail default)"; timeout 900 python -m pytest tests -x -q -m gpu 2>&1 | tail -3
for cfg in "0 6" "1 6" "1 5"; do
  set -- $cfg
  echo "SMEM=$1 LEVEL=$2"
  MG_TAIL_SMEM=$1 MG_TAIL_LEVEL=$2 timeout 300 python profiles/time_coarse.py --top 8 | python -c "
import json,sys
d=json.loads(sys.stdin.read())['us_per_subcycle']
print({k:(v['tail'], v['tail_graph']) for k,v in d.items()})"
  MG_TAIL_SMEM=$1 MG_TAIL_LEVEL=$2 timeout 300 python profiles/time_cycle.py | cut -c1-140
done
