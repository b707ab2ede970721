#!/bin/bash
# usage: tools/pack_ab.sh  — e2e and host packing time per step under a few packer settings (run on the GPU box)
run() { echo -n "$* : "; env "$@" python bench.py --steps 6 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); e=d['e2e']; print('e2e', round(e['value']/1e6,2), 'pack_ms', round(e['host_pack_ms_per_step'],2), 'serial', round(e['serial_cg_classify_batch_value']/1e6,2))"; }
lscpu | grep -E "Model name|Thread|Core|Socket|NUMA node0" | tr -s ' ' | tr '\n' ';'; echo
run CHARON_PACK_NT=1
run CHARON_PACK_NT=0
run CHARON_PACK_NT=1 CHARON_NO_AVX512=1
run CHARON_PACK_NT=0 CHARON_NO_AVX512=1
