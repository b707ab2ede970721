#!/bin/bash
# End-to-end demo of the host program on a GPU box: charon index -> charon dehost on synthetic data, wall-clock timed,
# with the oracle's TSV of the same run diffed for byte identity.  usage: tools/cli_demo.sh <ref_bp_per_category> <n_reads> <threads>
set -e
BP=${1:-50000000}; N=${2:-50000}; T=${3:-16}
D=$(mktemp -d /tmp/charon_demo.XXXX)
python - "$D" "$BP" "$N" <<'PY'
import sys, numpy as np
sys.path.insert(0, ".")
from tools import synth
d, bp, n = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
refs = [synth.reference(1000 + b, bp) for b in range(2)]
for b, name in enumerate(["host", "microbial"]):
    with open(f"{d}/{name}.fa", "wb") as f:
        f.write(f">{name}\n".encode())
        r = refs[b]
        pad = (-len(r)) % 80
        body = np.concatenate([r, np.full(pad, ord("A"), np.uint8)]).reshape(-1, 80)
        lines = np.concatenate([body, np.full((len(body), 1), 10, np.uint8)], axis=1).reshape(-1)
        f.write(lines[:len(lines) - pad - 1].tobytes() if pad else lines.tobytes())
        f.write(b"\n")
open(f"{d}/input.tsv", "w").write(f"{d}/host.fa\thost\n{d}/microbial.fa\tmicrobial\n")
L = 8000
bases, src = synth.reads(42, refs, [0.45, 0.45], n, L, 0.02)
q = np.full(L, ord("5"), np.uint8)
with open(f"{d}/reads.fq", "wb") as f:
    rows = bases.reshape(n, L)
    for i in range(n):
        f.write(b"@r%d_%d\n" % (i, src[i])); f.write(rows[i].tobytes()); f.write(b"\n+\n"); f.write(q.tobytes()); f.write(b"\n")
PY
ls -la $D | head
t0=$(date +%s.%N); charon_b200/charon index $D/input.tsv -p $D/demo --log $D/i.log; t1=$(date +%s.%N); echo "charon index wall $(python -c "print(round($t1 - $t0, 2))") s"
t0=$(date +%s.%N); charon_b200/charon dehost $D/reads.fq --db $D/demo.idx -t $T -v --log $D/d.log > $D/out.tsv; t1=$(date +%s.%N); echo "charon dehost -t $T wall $(python -c "print(round($t1 - $t0, 2))") s for $N reads"
grep "Read loop" $D/d.log
if [ "${SKIP_T1:-0}" != "1" ]; then
t0=$(date +%s.%N); charon_b200/charon dehost $D/reads.fq --db $D/demo.idx -t 1 -v --log $D/d1.log > $D/out1.tsv; t1=$(date +%s.%N); echo "charon dehost -t 1 wall $(python -c "print(round($t1 - $t0, 2))") s for $N reads"
grep "Read loop" $D/d1.log
else cp $D/out.tsv $D/out1.tsv; fi
cmp $D/out.tsv $D/out1.tsv && echo "TSV identical for -t 1 and -t $T"
# the same with the compression ratios computed on the GPU (bit-exact zlib level-6 size): the -t team no longer runs zlib
for TT in $T 4 1; do
t0=$(date +%s.%N); charon_b200/charon dehost $D/reads.fq --db $D/demo.idx -t $TT -v --gpu_compression --log $D/dg.log > $D/outg.tsv; t1=$(date +%s.%N); echo "charon dehost -t $TT --gpu_compression wall $(python -c "print(round($t1 - $t0, 2))") s for $N reads"
grep "Read loop" $D/dg.log
cmp $D/out.tsv $D/outg.tsv && echo "TSV identical with --gpu_compression (-t $TT)"
done
wc -l $D/out.tsv; head -2 $D/out.tsv; cut -f1,3 $D/out.tsv | sort | uniq -c
rm -rf $D
