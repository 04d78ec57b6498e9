set -u
echo "parity ctas=16 level=7"; MG_TAIL_CTAS=16 MG_TAIL_LEVEL=7 timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -m gpu 2>&1 | tail -3
echo "parity ctas=8 level=6 solo=3"; MG_TAIL_CTAS=8 MG_TAIL_LEVEL=6 MG_TAIL_SOLO=3 timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_domain.py -x -q -m gpu 2>&1 | tail -3
for cfg in "1 6 4" "8 6 4" "16 6 4" "16 7 4" "16 7 3" "16 7 5" "16 8 4" "8 7 4"; do
  set -- $cfg
  echo "CTAS=$1 LEVEL=$2 SOLO=$3"
  MG_TAIL_CTAS=$1 MG_TAIL_LEVEL=$2 MG_TAIL_SOLO=$3 timeout 300 python profiles/time_coarse.py --top 9 | python -c "
import json,sys
d=json.loads(sys.stdin.read())['us_per_subcycle']
print({k:(v['tail'], v['tail_graph']) for k,v in d.items()})"
  MG_TAIL_CTAS=$1 MG_TAIL_LEVEL=$2 MG_TAIL_SOLO=$3 timeout 300 python profiles/time_cycle.py | cut -c1-140
done
