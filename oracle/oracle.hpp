// =============================================================================
// oracle/oracle.hpp — CPU ORACLE for Charon's read-classification hot path.
//
// THIS IS TEST INFRASTRUCTURE, NOT PRODUCT CODE.  Only tests/, bench.py's
// cpu_baseline / --impl reference leg and __graft_entry__.smoke() may build,
// link or call it.  The product (charon_b200/csrc, libcharon_gpu.so, the
// `charon` host binary) never includes this header.
//
// PARITY STATUS: *parity unpinned*.  The reference (rmcolq/charon) ships no tests
// or golden vectors and all hot-path arithmetic lives in third-party code that
// is not vendored under /root/reference and cannot be fetched here:
//   seqan3 @ 30bdf8d0a5d59b79342b472504b95ae50c33da6d  (cmake/package-lock.cmake:5-13)
//   sdsl-lite v3 + cereal (via seqan3), kthohr/stats 3.4.0, kthohr/gcem 1.18.0,
//   mapbox/gzip-hpp @ 7546b35 (+ zlib).
// What follows restates (a) the *published algorithms* of those libraries (SURVEY.md
// Appendix A) and (b) charon's own glue, citing the reference file:line of every
// call site.  The anchors that DO exist (seqan3's minimiser_hash unit-test vector,
// hand-computable k-mer hashes, IBF seed closed forms) are checked in
// tests/test_oracle_kat.py.
// =============================================================================
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <deque>
#include <iomanip>
#include <limits>
#include <sstream>
#include <string>
#include <unordered_map>
#include <utility>
#include <vector>

namespace orc {

// ----------------------------------------------------------------------------
// A.1 alphabet: seqan3::dna5 ranks A0 C1 G2 N3 T4 (my_traits, reference
// include/utils.hpp:17-19).  char -> dna5: a/A c/C g/G t/T u/U(->T), anything else -> N.
// ----------------------------------------------------------------------------
inline uint8_t dna5_rank(unsigned char c) {
    switch (c) {
        case 'A': case 'a': return 0;
        case 'C': case 'c': return 1;
        case 'G': case 'g': return 2;
        case 'T': case 't': case 'U': case 'u': return 4;
        default: return 3;
    }
}
inline char dna5_char(uint8_t r) { return "ACGNT"[r]; }
inline uint8_t dna5_complement(uint8_t r) { static const uint8_t c[5] = {4, 2, 1, 3, 0}; return c[r]; }
// legal input alphabet of sequence_file_input_default_traits_dna is dna15 (IUPAC), case-insensitive.
inline bool dna15_legal(unsigned char c) {
    static const char *ok = "ACGTURYSWKMBDHVNacgturyswkmbdhvn";
    return c != 0 && std::strchr(ok, c) != nullptr;
}

constexpr uint64_t MINIMISER_SEED = 0x8F3F73B5CF1C9ADEULL;  // seqan3::seed default (A.2)

// ----------------------------------------------------------------------------
// A.1/A.2  seqan3::views::minimiser_hash(shape{ungapped{k}}, window_size{w})
// call sites: reference src/dehost_main.cpp:317-318,367 ; src/classify_main.cpp:123-124,174 ;
//             src/index_main.cpp:120-121,144.
// `ranks` = dna5 ranks of the sequence.  Output: the emitted minimiser values in order
// (duplicates kept — each one is one bulk_contains query and one num_hashes_ tick).
// Sigma is a parameter only so that the seqan3 unit-test vector (dna4) can be replayed.
// ----------------------------------------------------------------------------
inline void kmer_values(const uint8_t *ranks, size_t len, unsigned k, uint64_t seed, unsigned sigma,
                        const uint8_t *comp, std::vector<uint64_t> &out) {
    out.clear();
    if (k == 0 || len < k) return;
    size_t n = len - k + 1;
    out.resize(n);
    uint64_t top = 1;  // sigma^(k-1)
    for (unsigned i = 1; i < k; ++i) top *= sigma;
    uint64_t f = 0, r = 0;
    // first k-mer: forward = sum rank[j]*sigma^(k-1-j); reverse complement = sum comp(rank[j])*sigma^j
    uint64_t pw = 1;
    for (unsigned j = 0; j < k; ++j) {
        f = f * sigma + ranks[j];
        r += uint64_t(comp[ranks[j]]) * pw;
        pw *= sigma;
    }
    out[0] = std::min(f ^ seed, r ^ seed);
    for (size_t i = 1; i < n; ++i) {
        f = (f - uint64_t(ranks[i - 1]) * top) * sigma + ranks[i + k - 1];
        r = (r - comp[ranks[i - 1]]) / sigma + uint64_t(comp[ranks[i + k - 1]]) * top;
        out[i] = std::min(f ^ seed, r ^ seed);
    }
}

// The stateful window-minimiser of seqan3::detail::minimiser_view (A.2), stated literally
// with a deque like the library: rightmost minimum on (re)computation, strict '<' for a new
// value, emission every time the tracked element changes.
inline void minimiser_stream(const std::vector<uint64_t> &v, size_t window_values, std::vector<uint64_t> &out) {
    out.clear();
    size_t n = v.size();
    if (n == 0 || window_values == 0) return;
    size_t W = std::min(window_values, n);
    std::deque<uint64_t> win(v.begin(), v.begin() + W);
    auto rightmost_min = [&](uint64_t &val, size_t &off) {
        val = win[0]; off = 0;
        for (size_t i = 1; i < win.size(); ++i)
            if (win[i] <= val) { val = win[i]; off = i; }   // min_element with less_equal -> rightmost
    };
    uint64_t cur; size_t off;
    rightmost_min(cur, off);
    out.push_back(cur);
    for (size_t t = W; t < n; ++t) {
        uint64_t nv = v[t];
        win.pop_front();
        win.push_back(nv);
        if (off == 0) {
            rightmost_min(cur, off);
            out.push_back(cur);
        } else if (nv < cur) {
            cur = nv; off = win.size() - 1;
            out.push_back(cur);
        } else {
            --off;
        }
    }
}

inline void minimiser_hash(const uint8_t *ranks, size_t len, unsigned k, unsigned w, std::vector<uint64_t> &out,
                           uint64_t seed = MINIMISER_SEED) {
    static const uint8_t comp5[5] = {4, 2, 1, 3, 0};
    std::vector<uint64_t> v;
    kmer_values(ranks, len, k, seed, 5, comp5, v);
    minimiser_stream(v, size_t(w) - k + 1, out);
}

// ----------------------------------------------------------------------------
// A.3  seqan3::interleaved_bloom_filter, queried through membership_agent::bulk_contains
// call sites: reference include/index.hpp:26,110-112 ; src/dehost_main.cpp:321,368 ;
//             src/index_main.cpp:243-245,254 (emplace).
// The reference queries the sd_vector-compressed form; the oracle holds the same bits plain.
// ----------------------------------------------------------------------------
static const uint64_t IBF_SEEDS[5] = {13572355802537770549ULL, 13043817825332782213ULL, 10650232656628343401ULL,
                                      16499269484942379435ULL, 4893150838803335377ULL};
struct Ibf {
    uint64_t bins = 0, technical_bins = 0, bin_size = 0, hash_shift = 0, bin_words = 0, hash_funs = 0;
    std::vector<uint64_t> data;  // technical_bins * bin_size bits, bit i = bit (i%64) of word i/64

    Ibf() = default;
    Ibf(uint64_t bins_, uint64_t bin_size_, uint64_t hash_funs_) { init(bins_, bin_size_, hash_funs_); }
    void init(uint64_t bins_, uint64_t bin_size_, uint64_t hash_funs_) {
        bins = bins_; bin_size = bin_size_; hash_funs = hash_funs_;
        hash_shift = bin_size ? uint64_t(__builtin_clzll(bin_size)) : 64;
        bin_words = (bins + 63) >> 6;
        technical_bins = bin_words << 6;
        data.assign(bin_words * bin_size, 0);
    }
    inline uint64_t row_of(uint64_t h, uint64_t seed) const {  // hash_and_fit without the *technical_bins
        h *= seed;
        h ^= h >> hash_shift;
        h *= 11400714819323198485ULL;
        return uint64_t((static_cast<unsigned __int128>(h) * static_cast<unsigned __int128>(bin_size)) >> 64);
    }
    void emplace(uint64_t value, uint64_t bin) {
        for (uint64_t i = 0; i < hash_funs; ++i) {
            uint64_t bit = row_of(value, IBF_SEEDS[i]) * technical_bins + bin;
            __atomic_fetch_or(&data[bit >> 6], uint64_t(1) << (bit & 63), __ATOMIC_RELAXED);
        }
    }
    // result: bin_words u64, bit b of word b/64 = membership in bin b
    inline void bulk_contains(uint64_t value, uint64_t *result) const {
        for (uint64_t b = 0; b < bin_words; ++b) result[b] = ~uint64_t(0);
        for (uint64_t i = 0; i < hash_funs; ++i) {
            const uint64_t *row = &data[row_of(value, IBF_SEEDS[i]) * bin_words];
            for (uint64_t b = 0; b < bin_words; ++b) result[b] &= row[b];
        }
    }
};

// ----------------------------------------------------------------------------
// InputSummary (reference include/input_summary.hpp:17-62)
// ----------------------------------------------------------------------------
struct Summary {
    uint8_t num_bins = 0;
    std::vector<std::string> categories;
    std::vector<std::pair<std::string, uint8_t>> filepath_to_bin;
    std::unordered_map<uint8_t, std::string> bin_to_category;
    uint8_t num_categories() const { return uint8_t(categories.size()); }
    uint8_t category_index(const std::string &c) const {
        for (size_t i = 0; i < categories.size(); ++i) if (categories[i] == c) return uint8_t(i);
        return 255;
    }
    uint8_t host_category_index() const { return std::min(category_index("human"), category_index("host")); }
    std::string category_name(uint8_t idx) const { return idx > categories.size() ? std::string() : categories.at(idx); }
};
struct Stats {  // reference include/input_stats.hpp:17-19
    uint32_t num_files = 0;
    std::unordered_map<uint8_t, uint64_t> records_per_bin, hashes_per_bin;
};
struct Index {  // reference include/index.hpp:21-26
    uint8_t window_size = 0, kmer_size = 0;
    double max_fpr = 0;
    Summary summary;
    Stats stats;
    Ibf ibf;
};

// ----------------------------------------------------------------------------
// Options: union of DehostArguments / ClassifyArguments thresholds
// (reference include/dehost_arguments.hpp:22-43, include/classify_arguments.hpp:16-39)
// ----------------------------------------------------------------------------
struct Options {
    bool dehost = true;                 // dehost -> call_host ; classify -> call_category
    std::string dist = "kde";
    float lo_hi_threshold = 0.15f;
    uint16_t num_reads_to_fit = 5000;
    float min_quality = 15.0f;
    uint32_t min_length = 140;
    float min_compression = 0.0f;
    uint8_t confidence_threshold = 7;
    float confidence_probability_threshold = 0.0f;
    float host_unique_prop_lo_threshold = 0.05f;
    float min_proportion_difference = 0.04f;
    float min_prob_difference = 0.0f;
    uint8_t min_hits = 0;               // uninitialised in the reference (classify_stats.hpp:406); treated as 0
    bool run_extract = false;           // only effect modelled: training-cache capacity (result.hpp:80-86)
    static Options dehost_defaults() { return Options(); }
    static Options classify_defaults() {
        Options o; o.dehost = false; o.dist = "beta"; o.min_quality = 10.0f; o.min_compression = 0.15f;
        o.confidence_threshold = 2; o.min_proportion_difference = 0.0f; return o;
    }
};

// ----------------------------------------------------------------------------
// Model arithmetic (A.5).  reference include/classify_stats.hpp:210-254 (KDEParams),
// :370-389 (Model::prob), stats::dexp via kthohr/stats (float instantiation).
// ----------------------------------------------------------------------------
extern const float KDE_DEFAULT_POS[500];
extern const float KDE_DEFAULT_NEG[500];

struct Kde {
    std::vector<float> dataset;
    float h = 0;
    Kde() = default;
    Kde(const float *d, size_t n, float h_) : dataset(d, d + n), h(h_) { std::sort(dataset.begin(), dataset.end()); }
    void fit(const std::vector<float> &t) { dataset = t; }                       // :234-240 (h unchanged, NOT re-sorted)
    float K(const float &x) const {                                               // :242-244
        return std::exp(-std::pow(x, 2) / 2) / std::sqrt(2 * 3.141592653589793238463);
    }
    float prob(const float &x) const {                                            // :246-252
        float total_sum = 0;
        for (const auto &xi : dataset) total_sum += K((x - xi) / h);
        return total_sum / (h * dataset.size());
    }
};

// stats::dexp(float x, int rate): float arithmetic, exp(log(rate) - rate*x); x<0 -> 0; NaN -> NaN (A.5)
inline float dexp_f(float x, float rate) {
    if (std::isnan(x)) return std::numeric_limits<float>::quiet_NaN();
    if (x < 0.0f) return 0.0f;
    return std::exp(std::log(rate) - rate * x);
}
// stats::dbeta / stats::dgamma (float): log-density then exp, lgamma from libm (gcem's series unavailable:
// classify probabilities are best-effort).
inline float dbeta_f(float x, float a, float b) {
    if (std::isnan(x)) return std::numeric_limits<float>::quiet_NaN();
    if (x < 0.0f || x > 1.0f) return 0.0f;
    float lbeta = std::lgamma(a) + std::lgamma(b) - std::lgamma(a + b);
    return std::exp(-lbeta + (a - 1.0f) * std::log(x) + (b - 1.0f) * std::log(1.0f - x));
}
inline float dgamma_f(float x, float shape, float scale) {
    if (std::isnan(x)) return std::numeric_limits<float>::quiet_NaN();
    if (x < 0.0f) return 0.0f;
    return std::exp(-std::lgamma(shape) - shape * std::log(scale) + (shape - 1.0f) * std::log(x) - x / scale);
}

struct Model {  // reference classify_stats.hpp:261-393
    bool ready = false;
    std::string dist = "kde";
    float g_pos[3] = {25, 0, 0.02f}, g_neg[3] = {10, 0, 0.005f};   // shape, loc, scale
    float b_pos[2] = {6, 4}, b_neg[2] = {6, 40};
    Kde k_pos{KDE_DEFAULT_POS, 500, 0.1f}, k_neg{KDE_DEFAULT_NEG, 500, 0.001f};
    // returns ProbPair.pos (the only member used downstream, read_entry.hpp:277)
    double prob_pos(const float &x) const {
        const float p_err = dexp_f(x, 300);
        float p_pos = 0, p_neg = 0;
        if (dist == "kde") { p_pos = k_pos.prob(x); p_neg = k_neg.prob(x); }
        else if (dist == "gamma") { p_pos = dgamma_f(x - g_pos[1], g_pos[0], g_pos[2]); p_neg = dgamma_f(x - g_neg[1], g_neg[0], g_neg[2]); }
        else if (dist == "beta") { p_pos = dbeta_f(x, b_pos[0], b_pos[1]); p_neg = dbeta_f(x, b_neg[0], b_neg[1]); }
        if (x == 1) p_pos = 1;
        const float total = p_err + p_pos + p_neg;
        return p_pos / total;
    }
};

struct TrainingData {  // classify_stats.hpp:34-114
    bool complete = false, pos_complete = false, neg_complete = false;
    uint16_t num_reads_to_fit = 0;
    std::vector<float> pos, neg;
    bool check_status() {
        if (pos.size() >= num_reads_to_fit) pos_complete = true;
        if (neg.size() >= num_reads_to_fit) neg_complete = true;
        if (pos_complete && neg_complete) complete = true;
        return complete;
    }
    bool add_pos(float v) { if (pos.size() < num_reads_to_fit) pos.push_back(v); else check_status(); return complete; }
    bool add_neg(float v) { if (neg.size() < num_reads_to_fit && v > 0) neg.push_back(v); else check_status(); return complete; }
};

// mean / variance helpers and the moment fits (classify_stats.hpp:20-31,127-141,171-190)
template <class V> inline double mean_d(const V &v) { double s = 0.0; for (auto x : v) s += x; return v.empty() ? 0 : s / double(v.size()); }
inline double variance_d(const std::vector<float> &v, double mean) {   // the reference's lambda takes a FLOAT accumulator
    double acc = 0.0;
    for (float val : v) { const float a = float(acc); acc = double(a) + ((double(val) - mean) * (double(val) - mean)); }
    return v.size() <= 1 ? 0 : acc / double(v.size() - 1);
}
inline void fit_gamma(float *g, const std::vector<float> &data) {
    const double mu = mean_d(data), ln_mu = std::log(mu);
    std::vector<float> logs(data.size());
    for (size_t i = 0; i < data.size(); ++i) logs[i] = std::log(data[i]);
    const double s = ln_mu - mean_d(logs);
    g[0] = float((3 - s + std::sqrt((s - 3) * (s - 3) + 24 * s)) / (12 * s));
    g[2] = float(mu / g[0]);
}
inline void fit_beta(float *b, const std::vector<float> &data) {
    const double mu = mean_d(data), var = variance_d(data, mu);
    b[0] = float(mu * ((mu * (1 - mu) / var) - 1));
    b[1] = float((1 - mu) * ((mu * (1 - mu) / var) - 1));
}

struct StatsModel {  // classify_stats.hpp:395-584
    bool ready_ = false;
    Options opt;
    std::vector<TrainingData> training;
    std::vector<Model> models;
    StatsModel() = default;
    StatsModel(const Options &o, size_t ncat) : opt(o) {
        for (size_t i = 0; i < ncat; ++i) {
            Model m; m.dist = o.dist; models.push_back(m);
            TrainingData t; t.num_reads_to_fit = o.num_reads_to_fit; training.push_back(t);
        }
    }
    void train(size_t i) {  // Model::train :358-368 ; only the KDE fit changes parameters we evaluate exactly
        Model &m = models[i]; TrainingData &t = training[i];
        if (m.dist == "kde") {
            if (t.pos_complete) m.k_pos.fit(t.pos);
            if (t.neg_complete) m.k_neg.fit(t.neg);
        }
        else if (m.dist == "beta") {   // train_beta :315-339, BetaParams::fit :171-190
            if (t.pos_complete) fit_beta(m.b_pos, t.pos);
            if (t.neg_complete) fit_beta(m.b_neg, t.neg);
        } else {                       // train_gamma :289-313, GammaParams::fit :127-136, fit_loc :138-141
            if (t.pos_complete) fit_gamma(m.g_pos, t.pos);
            if (t.neg_complete) fit_gamma(m.g_neg, t.neg);
            else m.g_neg[1] = float(mean_d(t.neg) - double(m.g_neg[0] * m.g_neg[2]));   // NB: also on the default path (empty data -> loc = -shape*scale)
        }
        m.ready = true; t.pos.clear(); t.neg.clear();
    }
    void force_ready() { for (size_t i = 0; i < models.size(); ++i) if (!models[i].ready) train(i); ready_ = true; }
    void check_if_ready() { if (ready_) return; for (auto &m : models) if (!m.ready) return; ready_ = true; }
    void train_model_at(size_t i) { train(i); check_if_ready(); }
    bool add_read_to_training_data(const std::vector<float> &props) {  // :535-578
        uint8_t pos_i = 255; double max_val = 0.0; int above = 0;
        for (uint8_t i = 0; i < props.size(); ++i) {
            const float &val = props[i];
            if (val > opt.lo_hi_threshold) above += 1;
            if (val == max_val) pos_i = 255;
            else if (val > max_val) { pos_i = i; max_val = val; }
        }
        bool add_pos = (pos_i != 255 && above == 1);
        bool add_neg = add_pos || (above == 0);
        if (add_pos) { bool r = training.at(pos_i).add_pos(props[pos_i]); if (r && !models[pos_i].ready) train_model_at(pos_i); }
        if (add_neg) for (uint8_t i = 0; i < props.size(); ++i) if (i != pos_i) {
            bool r = training.at(i).add_neg(props[i]); if (r && !models[i].ready) train_model_at(i);
        }
        return ready_;
    }
};

// ----------------------------------------------------------------------------
// ReadEntry (reference include/read_entry.hpp:16-338)
// ----------------------------------------------------------------------------
struct ReadEntry {
    std::string read_id;
    uint32_t length = 0;
    float mean_quality = 0, compression = 0;
    uint32_t num_hashes = 0;
    size_t bin_words = 1;
    std::vector<uint64_t> bits;  // num_hashes * bin_words
    std::vector<uint32_t> counts, unique_counts;
    std::vector<float> proportions, unique_proportions;
    std::vector<double> probabilities;
    uint8_t call = 255, confidence_score = 0;

    ReadEntry() = default;
    ReadEntry(const std::string &id, uint32_t len, float mq, float cr, const Summary &s, size_t bw)
        : read_id(id), length(len), mean_quality(mq), compression(cr), bin_words(bw),
          counts(s.num_categories(), 0), unique_counts(s.num_categories(), 0),
          proportions(s.num_categories(), 0), unique_proportions(s.num_categories(), 0),
          probabilities(s.num_categories(), 1) {}

    void update_entry(const uint64_t *entry) { bits.insert(bits.end(), entry, entry + bin_words); num_hashes += 1; }  // :86-90

    void get_counts(const Summary &s) {  // :92-138
        std::vector<uint64_t> total(s.num_bins, 0);
        for (uint32_t h = 0; h < num_hashes; ++h)
            for (unsigned b = 0; b < s.num_bins; ++b)
                total[b] += (bits[h * bin_words + (b >> 6)] >> (b & 63)) & 1;
        std::vector<uint8_t> best(s.num_categories(), 255);
        for (unsigned b = 0; b < total.size(); ++b) {
            uint8_t idx = s.category_index(s.bin_to_category.at(uint8_t(b)));
            if (best.at(idx) == 255 || total[b] > total[best[idx]]) { best.at(idx) = uint8_t(b); counts.at(idx) = uint32_t(total[b]); }
        }
        std::vector<uint8_t> found;
        for (uint32_t h = 0; h < num_hashes; ++h) {
            found.clear();
            for (unsigned c = 0; c < best.size(); ++c) {
                unsigned b = best[c];
                if ((bits[h * bin_words + (b >> 6)] >> (b & 63)) & 1) found.push_back(uint8_t(c));
            }
            if (found.size() == 1) unique_counts.at(found.front()) += 1;
        }
    }
    void get_proportions() {  // :140-150
        for (size_t i = 0; i < proportions.size(); ++i) {
            proportions[i] = static_cast<float>(counts[i]) / static_cast<float>(num_hashes);
            unique_proportions[i] = static_cast<float>(unique_counts[i]) / static_cast<float>(num_hashes);
        }
    }
    void post_process(const Summary &s) { get_counts(s); get_proportions(); }

    void apply_model(const StatsModel &m) {  // :271-279
        for (size_t i = 0; i < unique_proportions.size(); ++i) probabilities[i] *= m.models.at(i).prob_pos(unique_proportions[i]);
    }
    void call_category(const StatsModel &m) {  // :157-216
        const Options &o = m.opt;
        uint8_t first = 0, second = 1;
        if (unique_counts.at(second) > unique_counts.at(first)) std::swap(first, second);
        for (size_t i = 2; i < unique_counts.size(); ++i)
            if (unique_counts[i] > unique_counts[second]) {
                second = uint8_t(i);
                if (unique_counts[second] > unique_counts[first]) std::swap(first, second);
            }
        uint32_t raw = unique_counts[first] - unique_counts[second];
        confidence_score = raw > 255 ? 255 : uint8_t(raw);
        if (mean_quality < o.min_quality) return;
        if (length < o.min_length) return;
        if (compression < o.min_compression) return;
        if (probabilities[second] == 0 && probabilities[first] > 0) call = first;
        else if (confidence_score > int8_t(o.confidence_threshold) && probabilities[first] > probabilities[second]) call = first;
        uint32_t fc = counts[first], sc = counts[second];
        if (sc > fc || fc - sc < o.min_hits) call = 255;
        float fp = proportions[first], sp = proportions[second];
        if (sp > fp || fp - sp < o.min_proportion_difference) call = 255;
    }
    void call_host(const StatsModel &m, uint8_t host_index) {  // :218-269
        const Options &o = m.opt;
        const uint8_t other = 1 - host_index;
        double hu = unique_proportions.at(host_index), ou = unique_proportions.at(other);
        double hp = probabilities.at(host_index), op = probabilities.at(other);
        uint8_t first = host_index, second = other;
        if (hu < ou) { first = other; second = host_index; }
        uint32_t raw = unique_counts[first] - unique_counts[second];
        confidence_score = raw > 255 ? 255 : uint8_t(raw);
        if (confidence_score < int8_t(o.confidence_threshold)) return;
        if (mean_quality < o.min_quality) return;
        if (length < o.min_length) return;
        if (compression < o.min_compression) return;
        if (hu > ou && hu - ou > o.min_proportion_difference && hp > op && hp - op > o.min_prob_difference &&
            std::max(hp * confidence_score, static_cast<double>(confidence_score)) >= o.confidence_probability_threshold)
            call = host_index;
        else if (hu < o.host_unique_prop_lo_threshold && hu < ou && ou - hu > o.min_proportion_difference && hp < op &&
                 op - hp > o.min_prob_difference &&
                 std::max(op * confidence_score, static_cast<double>(confidence_score)) >= o.confidence_probability_threshold)
            call = other;
    }
    // print_assignment_result :322-337 (std::ostream, precision 6, trailing space, newline)
    std::string tsv_row(const Summary &s) const {
        std::ostringstream os;
        os << (call == 255 ? "U" : "C") << "\t";
        os.precision(6);
        os << read_id << "\t" << s.category_name(call) << "\t" << length << "\t" << num_hashes << "\t" << mean_quality
           << "\t" << +confidence_score << "\t" << compression << "\t";
        for (size_t i = 0; i < s.num_categories(); ++i)
            os << s.categories.at(i) << ":" << counts.at(i) << ":" << proportions.at(i) << ":" << unique_proportions.at(i) << ":"
               << probabilities.at(i) << " ";
        os << "\n";
        return os.str();
    }
};

// ----------------------------------------------------------------------------
// Result: training/cache state machine + output order (reference include/result.hpp:59-213),
// at -t 1 semantics (input order; first read dropped when the cache has capacity 0: Appendix B-1).
// ----------------------------------------------------------------------------
struct Result {
    Summary summary;
    StatsModel model;
    std::vector<ReadEntry> cache;
    size_t cache_capacity = 0;
    std::vector<uint64_t> classified_counts;
    uint64_t unclassified = 0;
    std::string out;               // TSV text
    std::vector<ReadEntry> rows;   // classified entries in output order (for structured comparison)
    bool keep_rows = true, keep_text = true;

    Result(const Options &o, const Summary &s) : summary(s), model(o, s.num_categories()), classified_counts(s.num_categories(), 0) {
        if (o.run_extract) cache_capacity = size_t(o.num_reads_to_fit) * s.num_categories() * 4;  // result.hpp:84
    }
    uint8_t classify_read(ReadEntry &e) {  // :97-116
        e.apply_model(model);
        if (model.opt.dehost) e.call_host(model, summary.host_category_index()); else e.call_category(model);
        if (keep_text) out += e.tsv_row(summary);
        if (e.call < 255) classified_counts[e.call] += 1; else unclassified += 1;
        if (keep_rows) rows.push_back(e);
        return e.call;
    }
    void classify_cache() { for (auto e : cache) classify_read(e); cache.clear(); }  // :181-198
    void add_read(ReadEntry &e) {  // :130-153
        if (model.ready_) { classify_read(e); return; }
        bool done = false;
        if (cache.size() < cache_capacity) { cache.push_back(e); done = model.add_read_to_training_data(e.unique_proportions); }
        else { model.force_ready(); done = true; }   // NB: `e` itself is dropped here (Appendix B-1)
        if (done) classify_cache();
    }
    void complete() { classify_cache(); }  // :200-202
};

// split(id, " ")[0]  (reference src/utils.cpp:9-20, src/dehost_main.cpp:345)
inline std::string first_token(const std::string &id) { size_t p = id.find(' '); return p == std::string::npos ? id : id.substr(0, p); }

// gzip::compress size / length (reference src/utils.cpp:114-124): zlib deflateInit2(level -1, Z_DEFLATED,
// 15+16 gzip wrapper, memLevel 8, default strategy), one-shot Z_FINISH.  Defined in oracle.cpp (needs zlib).
float compression_ratio(const char *seq, size_t n);

// One read through the loop body of dehost_reads / classify_reads (reference src/dehost_main.cpp:343-373).
// Up to two segments (paired: minimisers of mate 1 then mate 2 into one entry, :458-465).
inline ReadEntry process_read(const Index &idx, const std::string &id, const char *s1, size_t n1, const char *s2, size_t n2,
                              float mean_quality, float compression) {
    ReadEntry e(first_token(id), uint32_t(n1 + n2), mean_quality, compression, idx.summary, idx.ibf.bin_words);
    std::vector<uint8_t> ranks;
    std::vector<uint64_t> mins;
    std::vector<uint64_t> res(idx.ibf.bin_words);
    for (int seg = 0; seg < 2; ++seg) {
        const char *s = seg ? s2 : s1; size_t n = seg ? n2 : n1;
        if (!s || !n) continue;
        ranks.resize(n);
        for (size_t i = 0; i < n; ++i) ranks[i] = dna5_rank((unsigned char)s[i]);
        minimiser_hash(ranks.data(), n, idx.kmer_size, idx.window_size, mins);
        for (uint64_t v : mins) { idx.ibf.bulk_contains(v, res.data()); e.update_entry(res.data()); }
    }
    e.post_process(idx.summary);
    return e;
}

// Index file I/O in the reference's cereal-binary layout; oracle.cpp.
void write_index(const Index &idx, const std::string &path);
bool read_index(Index &idx, const std::string &path, std::string &err);

// `charon index` restated (reference src/index_main.cpp:75-263, src/utils.cpp:75-102): one bin per
// (sequence set, category) pair, distinct minimiser hashes per bin, bin_size from the largest bin.
size_t bin_size_in_bits(uint64_t num_elements, unsigned num_hash = 3, double max_fpr = 0.01,
                        uint64_t max_bits = uint64_t(std::numeric_limits<uint32_t>::max()) - 2);
Index build_index(const std::vector<std::pair<std::string, std::string>> &path_and_category_unused,
                  const std::vector<std::vector<std::string>> &bin_sequences, const std::vector<std::string> &bin_category,
                  const std::vector<std::string> &category_order, unsigned k, unsigned w, int threads);

}  // namespace orc
