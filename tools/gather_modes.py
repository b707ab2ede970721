"""ncu helper: launch the random-gather microbenchmark once per load flavour on a 1 GB filter (2 bins x 4e9 rows / ...)."""
import sys, numpy as np
sys.path.insert(0, ".")
import charon_b200
rows = 3_400_000_000
# empty filter of the human-scale size (content does not matter for traffic): build from sd_vector with zero ones
g = charon_b200.GpuClassifier(k=19, w=41, bins=2, bin_size=rows, hash_funs=3, bin_to_cat=[0, 1], ncat=2,
                              sd=(np.zeros(1, np.uint64), 0, np.zeros(1, np.uint64), 1, 0))
n = 60_000_000
for mode in range(1, 9):
    ms = g.bench_gather(n, mode=mode, reps=3)
    print(f"mode {mode}: {ms:.3f} ms  {32 * n / ms / 1e6:.0f} GB/s sector-terms  {n / ms / 1e6:.1f} G gathers/s", flush=True)
