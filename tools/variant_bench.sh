#!/bin/bash
# usage: tools/variant_bench.sh <workload> lib1.so lib2.so ...   (prints value / count-kernel ms per variant)
wl=$1; shift
for so in "$@"; do
  CHARON_GPU_LIB=$PWD/$so python bench.py --workload $wl --steps 6 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$so', '$wl', round(d['value']/1e6,3), 'Mreads/s  count_ms', round(d['roofline']['avg_launch_ms'],3))"
done
