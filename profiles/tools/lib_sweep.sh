#!/bin/bash
# Runs bench.py on C2 and a 50k-fragment C4 with each experiment build of the library given as arguments (HC_LIBHCGPU).
# usage: profiles/tools/lib_sweep.sh haploclique_b200/libhcgpu.so haploclique_b200/libhcgpu_v1.so ...
for lib in "$@"; do
  for w in "C2 0" "C4 50000" "C5 0"; do
    set -- $w
    extra=""; [ "$2" != "0" ] && extra="--n-frag $2"
    HC_LIBHCGPU=$PWD/$lib python bench.py --workload $1 $extra --steps 5 --no-also --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$lib', '$1', 'kernel_ms %.2f' % d['ms_per_step'], 'e2e_ms %.2f' % d['e2e']['ms_per_step'], 'frac %.3f' % d['roofline']['frac'], 'edges', d['pairs']['edges'])"
  done
done
