#!/bin/bash
# where the -t team's time goes without zlib (CHARON_PREP_PROFILE=1): usage tools/prep_probe.sh <n_reads> <threads>
set -e
N=${1:-50000}; T=${2:-1}
D=$(mktemp -d /tmp/charon_prep.XXXX)
python - "$D" "$N" <<'PY'
import sys, numpy as np
sys.path.insert(0, ".")
from tools import synth
d, n = sys.argv[1], int(sys.argv[2])
refs = [synth.reference(1000 + b, 5_000_000) for b in range(2)]
for b, name in enumerate(["host", "microbial"]):
    with open(f"{d}/{name}.fa", "wb") as f:
        f.write(f">{name}\n".encode()); f.write(refs[b].tobytes()); f.write(b"\n")
open(f"{d}/input.tsv", "w").write(f"{d}/host.fa\thost\n{d}/microbial.fa\tmicrobial\n")
L = 8000
bases, src = synth.reads(42, refs, [0.45, 0.45], n, L, 0.02)
q = np.full(L, ord("5"), np.uint8)
with open(f"{d}/reads.fq", "wb") as f:
    rows = bases.reshape(n, L)
    for i in range(n):
        f.write(b"@r%d_%d\n" % (i, src[i])); f.write(rows[i].tobytes()); f.write(b"\n+\n"); f.write(q.tobytes()); f.write(b"\n")
PY
charon_b200/charon index $D/input.tsv -p $D/demo --log $D/i.log
for FLAG in --gpu_compression ""; do for TT in $T 16; do
echo "== -t $TT $FLAG"
CHARON_PREP_PROFILE=1 charon_b200/charon dehost $D/reads.fq --db $D/demo.idx -t $TT -v $FLAG --log $D/d.log > $D/out.tsv
grep -E "Read loop|prepare_batch" $D/d.log | tail -2
done; done
rm -rf $D
