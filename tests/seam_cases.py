"""Deterministic tile-seam inputs for the minimiser kernels.

Two k-mers `a < b` with EQUAL canonical value, d <= 22 k-mers apart, forced to be the strict minimum of every window that
holds either of them, are planted so that `a` sits at a chosen k-mer position near the end of a kernel tile (713 windows
for the warp-per-read kernel, 2921 for the CTA-per-read kernel).  seqan3's minimiser (oracle.hpp:95-123) tracks
`a` (strictly smaller newcomer), ignores `b` (equal), and EMITS again when `a` falls out of the window (recomputation picks
`b`) although the window minimum has not changed.  A kernel that decides emissions from values alone misses that emission
unless it notices the tie; the case that got through in round 1 has a and b inside the last window of one tile and the
departure of a in the next tile.

Equal canonical values come in two flavours:
  * repeat: the same 19-mer d bases later (d < 19: a stretch of period d);
  * twin:   a 19-mer followed d bases later by its reverse complement (the stretch of 19+d bases is a reverse-complement
            palindrome; for even d its centre base is N, the only self-complementary dna5 letter).
"""
import random

SEED = 0x8F3F73B5CF1C9ADE
K = 19
W = 23                      # window in k-mers (w - k + 1)
RANK = {"A": 0, "C": 1, "G": 2, "N": 3, "T": 4}
COMP = {"A": "T", "C": "G", "G": "C", "T": "A", "N": "N"}
P18 = 5 ** (K - 1)


def revcomp(s):
    return "".join(COMP[c] for c in reversed(s))


def canon(kmer):
    f = 0
    for c in kmer:
        f = f * 5 + RANK[c]
    r = 0
    for c in reversed(kmer):
        r = r * 5 + RANK[COMP[c]]
    return min(f ^ SEED, r ^ SEED)


_SAMPLE = None


def quantile(v):
    """Fraction of random ACGT 19-mers whose canonical value lies below v (estimated on a fixed sample)."""
    global _SAMPLE
    if _SAMPLE is None:
        r = random.Random(12345)
        _SAMPLE = sorted(canon("".join(r.choice("ACGT") for _ in range(K))) for _ in range(3000))
    import bisect
    return bisect.bisect_left(_SAMPLE, v) / len(_SAMPLE)


def _stretch(rng, d, twin, tries=400):
    """19+d bases whose k-mers at offsets 0 and d have the same canonical value, that value being smaller than the value of
    every k-mer in between; among `tries` candidates the one with the smallest pair value (the flanks are grown to fit
    whatever value comes out, so it need not be small)."""
    best = None
    n = K + d
    for _ in range(tries):
        if twin:
            half = [rng.choice("ACGT") for _ in range(n // 2)]
            mid = ["N"] if n % 2 else []
            s = "".join(half + mid + [COMP[c] for c in reversed(half)])
        else:
            unit = "".join(rng.choice("ACGT") for _ in range(d))
            s = (unit * (n // d + 1))[:n]
        v = canon(s[:K])
        assert canon(s[d:d + K]) == v
        if any(canon(s[j:j + K]) <= v for j in range(1, d)):
            continue
        if best is None or v < best[0]:
            best = (v, s)
            if quantile(v) < 0.05:
                break
    return best


def _grow_right(rng, s, bound, n_more, budget=1500):
    """Append n_more bases to s so that every k-mer ending in the appended part has a canonical value > bound (depth-first
    with backtracking: each new base fixes exactly one new k-mer).  None when `budget` k-mer evaluations did not suffice."""
    s = list(s)
    start = len(s)
    choices = []
    while len(s) - start < n_more:
        opts = list("ACGT")
        rng.shuffle(opts)
        choices.append(opts)
        s.append(None)
        while True:
            while choices and not choices[-1]:
                choices.pop()
                s.pop()
            if not choices:
                return None
            s[-1] = choices[-1].pop()
            budget -= 1
            if budget < 0:
                return None
            if len(s) < K or canon("".join(s[-K:])) > bound:
                break
    return "".join(s)


def planted(rng, d, twin, flank=72):
    """(sequence, offset of a's first base): flank + stretch + flank, every k-mer other than a and b above the pair."""
    got = _stretch(rng, d, twin)
    if got is None:
        return None
    v, st = got
    if quantile(v) > 0.6:      # e.g. d = 1: only poly-A/T and poly-C/G exist and neither needs to be small; such pairs cannot be
        return None            # made the window minimum by any flank worth searching for
    right = _grow_right(rng, st, v, flank)
    left = _grow_right(rng, revcomp(st), v, flank)           # the same on the other strand = the left flank
    if right is None or left is None:
        return None
    seq = revcomp(left[len(st):]) + right
    return seq, flank


def sweep(tile_windows, total_len, seed, positions=None, ds=range(1, 23), cores_per_kind=3):
    """[(d, twin, pos_a, sequence)] for the seam after the first tile of `tile_windows` windows: that tile's last window is
    number tile_windows - 1 and holds k-mers tile_windows - 1 .. tile_windows + 21."""
    rng = random.Random(seed)
    last = tile_windows - 1
    if positions is None:
        positions = range(last - W - 2, last + W + 3)
    rs = lambda n: "".join(rng.choices("ACGT", k=n))
    out = []
    for d in ds:
        for twin in (False, True):
            cores = []
            for _ in range(cores_per_kind * 8):      # a stretch whose overlapping neighbours cannot be lifted above it is redrawn
                c = planted(rng, d, twin)
                if c is not None:
                    cores.append(c)
                    if len(cores) == cores_per_kind:
                        break
            if not cores:
                continue
            for i, pa in enumerate(positions):
                core, off = cores[i % len(cores)]
                pre = pa - off
                assert pre >= 0 and total_len >= pre + len(core)
                out.append((d, twin, pa, rs(pre) + core + rs(total_len - pre - len(core))))
    return out
