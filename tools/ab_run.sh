#!/bin/bash
# Development aid: same-box A/B of tools/_ab/*.so on the C2 bench line.  usage: tools/ab_run.sh name1 name2 ... [-- bench args]
names=(); while [ $# -gt 0 ] && [ "$1" != "--" ]; do names+=("$1"); shift; done; [ "$1" == "--" ] && shift
for rep in 1 2; do for n in "${names[@]}"; do
  r=$(B200KM_LIB=$PWD/tools/_ab/$n.so timeout 300 python bench.py --no-e2e --no-cpu --no-also --no-parity --steps 20 --warmup 3 "$@" 2>/dev/null | tail -1 | python -c "import sys,json; j=json.loads(sys.stdin.read()); print(round(j['ms_per_step'],4), round(j['roofline']['frac'],4), j['config']['result'].get('centroid_checksum'))")
  echo "$n: $r"
done; done
