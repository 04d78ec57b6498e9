set -u
bash profiles/_run17.sh
cd evostencils_b200/csrc && cp libmgengine_timing.so libmgengine.so && cd ../.. && python profiles/_dbg_tail.py | tail -38 | awk '{print $2, $4, $6, $8, $12}' | paste - - - - | head -12
