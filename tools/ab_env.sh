#!/bin/bash
# A/B of one build with / without an environment switch: tools/ab_env.sh VAR [bench args]
V=$1; shift
for r in 1 2; do
  for v in on off; do
    if [ $v = on ]; then unset $V; else export $V=1; fi
    python bench.py --steps 200 --warmup 5 --no-cpu --settle-max 0 "$@" 2>&1 | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$V set' if '$v'=='off' else '$V unset', round(d['ms_per_step'],4), round(d['windows']['transient']['ms_per_step'],4) if 'windows' in d else '', {k:round(v['ms'],4) for k,v in d.get('kernels',{}).items()})"
  done
done
