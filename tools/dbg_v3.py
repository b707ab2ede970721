import sys, random, collections
sys.path.insert(0,'.')
import numpy as np, oracle
from tests import common
from tests.test_gpu_parity import _tie_heavy_cases
idx, refs = common.build_oracle_index([300_000, 300_000], [0, 1], ["host", "microbial"], seed=11)
gpu = common.gpu_from_oracle(idx)
s=[c for c in _tie_heavy_cases() if len(c)==8160][0]
def ok(t):
    return collections.Counter(int(x) for x in gpu.debug_minimisers(t)) == collections.Counter(int(x) for x in oracle.minimisers(t))
t=s[4450:7173]
# shrink from the end
while len(t)>60 and not ok(t[:-1]): t=t[:-1]
while len(t)>60 and not ok(t[1:]): t=t[1:]
while len(t)>60 and not ok(t[:-1]): t=t[:-1]
print("minimal", len(t), "nkm", len(t)-18, "nwin", len(t)-40)
print(t)
v=[int(x) for x in oracle.kmer_values(t)]
print("v low16:", [x&0xFFFF for x in v])
want=[int(x)&0xFFFF for x in oracle.minimisers(t)]
got=sorted(int(x)&0xFFFF for x in gpu.debug_minimisers(t))
print("want(sorted low16)", sorted(want)); print("got", got)
