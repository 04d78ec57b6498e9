set -u
echo "all gpu tests"; timeout 900 python -m pytest tests -x -q -m gpu 2>&1 | tail -3
timeout 300 python profiles/time_coarse.py --top 8 | python -c "
import json,sys
d=json.loads(sys.stdin.read())['us_per_subcycle']
print({k:(v['tail'], v['tail_graph']) for k,v in d.items()})"
timeout 300 python profiles/time_cycle.py | cut -c1-140
