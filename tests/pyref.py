"""Third, independent (pure-Python, from-scratch, non-rolling) statement of SURVEY Appendix A.1-A.3, used to
cross-check the C++ oracle on small inputs.  Deliberately written differently from oracle.hpp: k-mer hashes are
recomputed from scratch per k-mer and the window minimiser re-scans the window instead of carrying a deque."""
SEED = 0x8F3F73B5CF1C9ADE
M64 = (1 << 64) - 1
RANK5 = {**{c: 3 for c in map(chr, range(256))}, "A": 0, "a": 0, "C": 1, "c": 1, "G": 2, "g": 2, "T": 4, "t": 4, "U": 4, "u": 4}
COMP5 = [4, 2, 1, 3, 0]
RANK4 = {"A": 0, "C": 1, "G": 2, "T": 3}
COMP4 = [3, 2, 1, 0]
IBF_SEEDS = [13572355802537770549, 13043817825332782213, 10650232656628343401, 16499269484942379435, 4893150838803335377]


def kmer_values(seq, k, seed=SEED, sigma=5):
    rank, comp = (RANK5, COMP5) if sigma == 5 else (RANK4, COMP4)
    r = [rank[c] for c in seq]
    out = []
    for i in range(len(r) - k + 1):
        f = 0
        for j in range(k):
            f = f * sigma + r[i + j]
        rc = 0
        for j in range(k):            # reverse complement read left to right = original right to left, complemented
            rc = rc * sigma + comp[r[i + k - 1 - j]]
        out.append(min(f ^ seed, rc ^ seed))
    return out


def minimisers(seq, k, w, seed=SEED, sigma=5):
    v = kmer_values(seq, k, seed, sigma)
    n = len(v)
    if n == 0:
        return []
    W = min(w - k + 1, n)

    def rightmost(lo):
        best = lo
        for i in range(lo, lo + W):
            if v[i] <= v[best]:
                best = i
        return best

    pos = rightmost(0)
    out = [v[pos]]
    for start in range(1, n - W + 1):
        new = start + W - 1
        if pos < start:
            pos = rightmost(start)
            out.append(v[pos])
        elif v[new] < v[pos]:
            pos = new
            out.append(v[pos])
    return out


def ibf_row(h, i, bin_size):
    shift = 64 - bin_size.bit_length()
    h = (h * IBF_SEEDS[i]) & M64
    h ^= h >> shift
    h = (h * 11400714819323198485) & M64
    return (h * bin_size) >> 64
