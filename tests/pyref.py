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


class TrainingTrace:
    """Third statement of the training state machine that `--extract` runs on the first reads, written from the reference
    text alone (include/classify_stats.hpp: TrainingData :34-114, Model::train / train_kde :341-368,
    StatsModel::add_read_to_training_data / train_model_at / check_if_ready / force_ready :463-533,535-578) and deliberately
    shaped differently from oracle.hpp and host/stats_model.hpp (dict records instead of classes, no early returns)."""

    def __init__(self, ncat, lo_hi_threshold, num_reads_to_fit):
        self.thr, self.cap, self.ready = lo_hi_threshold, num_reads_to_fit, False
        self.cat = [dict(pos=[], neg=[], pos_done=False, neg_done=False, done=False, model_ready=False, fit_pos=None, fit_neg=None)
                    for _ in range(ncat)]

    def _status(self, c):                       # TrainingData::check_status
        c["pos_done"] = c["pos_done"] or len(c["pos"]) >= self.cap
        c["neg_done"] = c["neg_done"] or len(c["neg"]) >= self.cap
        c["done"] = c["done"] or (c["pos_done"] and c["neg_done"])

    def _train(self, c):                        # Model::train with dist == "kde": fit what is complete, then clear the data
        if c["pos_done"]:
            c["fit_pos"] = list(c["pos"])
        if c["neg_done"]:
            c["fit_neg"] = list(c["neg"])
        c["model_ready"] = True
        c["pos"], c["neg"] = [], []

    def _offer(self, i, value, positive):       # add_pos / add_neg, then train_model_at when the data just became complete
        c = self.cat[i]
        room = len(c["pos" if positive else "neg"]) < self.cap
        if room and (positive or value > 0):
            c["pos" if positive else "neg"].append(value)
        else:
            self._status(c)
        if c["done"] and not c["model_ready"]:
            self._train(c)
            self.ready = self.ready or all(x["model_ready"] for x in self.cat)

    def add(self, props):
        best, best_val, above = None, 0.0, 0
        for i, v in enumerate(props):
            above += 1 if v > self.thr else 0
            if v == best_val:
                best = None
            elif v > best_val:
                best, best_val = i, v
        to_pos = best is not None and above == 1
        if to_pos:
            self._offer(best, props[best], True)
        if to_pos or above == 0:
            for i, v in enumerate(props):
                if i != best:
                    self._offer(i, v, False)
        return self.ready

    def force_ready(self):
        for c in self.cat:
            if not c["model_ready"]:
                self._train(c)
        self.ready = True

    def state(self):
        return (self.ready, [(len(c["pos"]), len(c["neg"]), c["model_ready"]) for c in self.cat])
