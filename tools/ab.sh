#!/bin/bash
# A/B of two builds of the library on one box: tools/ab.sh [bench args]; expects tools/scratch/libA.so and libB.so
for r in 1 2; do
  for v in A B; do
    cp tools/scratch/lib$v.so chrono_wisc_b200/libchrono_b200.so
    python bench.py --steps 200 --warmup 5 --no-cpu --settle-max 0 "$@" 2>&1 | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$v', round(d['ms_per_step'],4), round(d['windows']['transient']['ms_per_step'],4) if 'windows' in d else '', {k:round(v['ms'],4) for k,v in d.get('kernels',{}).items()})"
  done
done
