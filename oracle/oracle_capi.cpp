// oracle/oracle_capi.cpp — TEST INFRASTRUCTURE (see oracle.hpp).  extern "C" surface used by
// tests/, bench.py (cpu_baseline / --impl reference) and __graft_entry__.smoke() through ctypes.
#include <chrono>
#include <cstdio>
#ifdef _OPENMP
#include <omp.h>
#endif
#include "oracle.hpp"

using namespace orc;

extern "C" {

struct orc_options {           // mirrors orc::Options (dehost_arguments.hpp / classify_arguments.hpp)
    int32_t dehost;            // 1 dehost (call_host), 0 classify (call_category)
    int32_t dist;              // 0 kde, 1 beta, 2 gamma
    float lo_hi_threshold;
    uint32_t num_reads_to_fit;
    float min_quality;
    uint32_t min_length;
    float min_compression;
    int32_t confidence_threshold;
    float confidence_probability_threshold;
    float host_unique_prop_lo_threshold;
    float min_proportion_difference;
    float min_prob_difference;
    uint32_t min_hits;
    int32_t run_extract;
};

static Options to_opt(const orc_options *o) {
    Options r;
    r.dehost = o->dehost != 0;
    r.dist = o->dist == 0 ? "kde" : (o->dist == 1 ? "beta" : "gamma");
    r.lo_hi_threshold = o->lo_hi_threshold; r.num_reads_to_fit = uint16_t(o->num_reads_to_fit);
    r.min_quality = o->min_quality; r.min_length = o->min_length; r.min_compression = o->min_compression;
    r.confidence_threshold = uint8_t(o->confidence_threshold);
    r.confidence_probability_threshold = o->confidence_probability_threshold;
    r.host_unique_prop_lo_threshold = o->host_unique_prop_lo_threshold;
    r.min_proportion_difference = o->min_proportion_difference; r.min_prob_difference = o->min_prob_difference;
    r.min_hits = uint8_t(o->min_hits); r.run_extract = o->run_extract != 0;
    return r;
}

void orc_default_options(int dehost, orc_options *o) {
    Options d = dehost ? Options::dehost_defaults() : Options::classify_defaults();
    o->dehost = d.dehost; o->dist = d.dist == "kde" ? 0 : (d.dist == "beta" ? 1 : 2);
    o->lo_hi_threshold = d.lo_hi_threshold; o->num_reads_to_fit = d.num_reads_to_fit; o->min_quality = d.min_quality;
    o->min_length = d.min_length; o->min_compression = d.min_compression; o->confidence_threshold = d.confidence_threshold;
    o->confidence_probability_threshold = d.confidence_probability_threshold;
    o->host_unique_prop_lo_threshold = d.host_unique_prop_lo_threshold;
    o->min_proportion_difference = d.min_proportion_difference; o->min_prob_difference = d.min_prob_difference;
    o->min_hits = d.min_hits; o->run_extract = 0;
}

// ---- stage-level functions -------------------------------------------------------------------
// minimiser_hash of an ASCII sequence.  sigma 5: dna5 (product alphabet); sigma 4: dna4 with ranks
// A0 C1 G2 T3 (only to replay the seqan3 unit-test vector).  Returns the number of emitted values.
uint64_t orc_minimisers(const char *seq, uint64_t len, uint32_t k, uint32_t w, uint64_t seed, uint32_t sigma,
                        uint64_t *out, uint64_t cap) {
    std::vector<uint8_t> ranks(len);
    std::vector<uint64_t> v, m;
    if (sigma == 4) {
        static const uint8_t comp4[4] = {3, 2, 1, 0};
        for (uint64_t i = 0; i < len; ++i) { uint8_t r = dna5_rank((unsigned char)seq[i]); ranks[i] = r == 4 ? 3 : (r == 3 ? 0 : r); }
        kmer_values(ranks.data(), len, k, seed, 4, comp4, v);
    } else {
        static const uint8_t comp5[5] = {4, 2, 1, 3, 0};
        for (uint64_t i = 0; i < len; ++i) ranks[i] = dna5_rank((unsigned char)seq[i]);
        kmer_values(ranks.data(), len, k, seed, 5, comp5, v);
    }
    minimiser_stream(v, size_t(w) - k + 1, m);
    for (uint64_t i = 0; i < m.size() && i < cap; ++i) out[i] = m[i];
    return m.size();
}
// canonical seeded k-mer values (one per k-mer), dna5
uint64_t orc_kmer_values(const char *seq, uint64_t len, uint32_t k, uint64_t seed, uint64_t *out, uint64_t cap) {
    static const uint8_t comp5[5] = {4, 2, 1, 3, 0};
    std::vector<uint8_t> ranks(len);
    for (uint64_t i = 0; i < len; ++i) ranks[i] = dna5_rank((unsigned char)seq[i]);
    std::vector<uint64_t> v;
    kmer_values(ranks.data(), len, k, seed, 5, comp5, v);
    for (uint64_t i = 0; i < v.size() && i < cap; ++i) out[i] = v[i];
    return v.size();
}
uint64_t orc_ibf_seed(uint32_t i) { return IBF_SEEDS[i]; }
float orc_compression_ratio(const char *seq, uint64_t n) { return compression_ratio(seq, n); }
double orc_model_prob(const orc_options *o, float x) { Model m; m.dist = to_opt(o).dist; return m.prob_pos(x); }
uint64_t orc_bin_size_in_bits(uint64_t n) { return bin_size_in_bits(n); }
void orc_kde_defaults(float *pos_sorted, float *neg_sorted) {
    Model m;
    std::copy(m.k_pos.dataset.begin(), m.k_pos.dataset.end(), pos_sorted);
    std::copy(m.k_neg.dataset.begin(), m.k_neg.dataset.end(), neg_sorted);
}

// ---- index handle ----------------------------------------------------------------------------
struct orc_index { Index idx; std::vector<uint8_t> bin_to_cat; std::string err; };

static void finish(orc_index *h) {
    h->bin_to_cat.assign(h->idx.summary.num_bins, 255);
    for (auto &p : h->idx.summary.bin_to_category) if (p.first < h->bin_to_cat.size()) h->bin_to_cat[p.first] = h->idx.summary.category_index(p.second);
}

// Build from in-memory references.  refs = concatenated sequences, ref_off[nrec+1]; rec_bin[nrec] = bin of
// each record; bin_cat[nbins] = index into cat_names (a '\n'-joined list giving the category ORDER).
orc_index *orc_index_build(const char *refs, const uint64_t *ref_off, const uint32_t *rec_bin, uint32_t nrec, uint32_t nbins,
                           const uint32_t *bin_cat, const char *cat_names, uint32_t k, uint32_t w, int threads) {
    std::vector<std::string> cats;
    { std::string s(cat_names), cur; for (char c : s) { if (c == '\n') { cats.push_back(cur); cur.clear(); } else cur += c; } if (!cur.empty()) cats.push_back(cur); }
    std::vector<std::vector<std::string>> bins(nbins);
    for (uint32_t r = 0; r < nrec; ++r) bins[rec_bin[r]].emplace_back(refs + ref_off[r], refs + ref_off[r + 1]);
    std::vector<std::string> bc(nbins);
    for (uint32_t b = 0; b < nbins; ++b) bc[b] = cats[bin_cat[b]];
    orc_index *h = new orc_index;
    h->idx = build_index({}, bins, bc, cats, k, w, threads > 0 ? threads : 1);
    finish(h);
    return h;
}
orc_index *orc_index_load(const char *path) {
    orc_index *h = new orc_index;
    if (!read_index(h->idx, path, h->err)) { std::fprintf(stderr, "orc_index_load: %s\n", h->err.c_str()); delete h; return nullptr; }
    finish(h);
    return h;
}
int orc_index_save(orc_index *h, const char *path) { try { write_index(h->idx, path); } catch (std::exception &e) { std::fprintf(stderr, "%s\n", e.what()); return -1; } return 0; }
void orc_index_free(orc_index *h) { delete h; }
// header: k, w, bins, technical_bins, bin_size, hash_shift, bin_words, hash_funs, ncat
void orc_index_header(orc_index *h, uint64_t *out9) {
    const Ibf &f = h->idx.ibf;
    uint64_t v[9] = {h->idx.kmer_size, h->idx.window_size, f.bins, f.technical_bins, f.bin_size, f.hash_shift, f.bin_words, f.hash_funs,
                     h->idx.summary.num_categories()};
    std::copy(v, v + 9, out9);
}
const uint64_t *orc_index_rows(orc_index *h) { return h->idx.ibf.data.data(); }
const uint8_t *orc_index_bin_to_cat(orc_index *h) { return h->bin_to_cat.data(); }
int orc_index_category_name(orc_index *h, uint32_t i, char *buf, uint32_t cap) {
    const std::string &s = h->idx.summary.categories.at(i);
    if (s.size() + 1 > cap) return -1;
    std::memcpy(buf, s.c_str(), s.size() + 1);
    return int(s.size());
}
int orc_index_host_category(orc_index *h) { return h->idx.summary.host_category_index(); }
void orc_index_query(orc_index *h, const uint64_t *values, uint64_t n, uint64_t *masks) {
    for (uint64_t i = 0; i < n; ++i) h->idx.ibf.bulk_contains(values[i], masks + i * h->idx.ibf.bin_words);
}

// ---- batch path: loop body of dehost_reads for n_reads reads, OpenMP over reads like the reference's
// `omp parallel for` over a chunk (src/dehost_main.cpp:341).  Model fixed (default datasets) and applied
// to every read (no first-read drop here: this is the per-read function, the drop lives in Result).
// with_gzip: also compute the gzip ratio per read (cost model of the reference loop); when 0 the given
// `compression` array is used.  Returns wall seconds.
double orc_classify_batch(orc_index *h, const orc_options *o, const char *bases, const uint64_t *seg_off, const uint32_t *read_seg,
                          uint32_t n_reads, const float *mean_q, const float *compression, int with_gzip, int threads,
                          uint32_t *num_hashes, uint32_t *counts, uint32_t *unique_counts, double *prob, uint8_t *call,
                          uint8_t *confidence, float *compression_out) {
    Options opt = to_opt(o);
    const Index &idx = h->idx;
    size_t ncat = idx.summary.num_categories();
    StatsModel model(opt, ncat);
    model.force_ready();
    uint8_t host_idx = idx.summary.host_category_index();
    auto t0 = std::chrono::steady_clock::now();
#pragma omp parallel for schedule(dynamic, 16) num_threads(threads > 0 ? threads : 1)
    for (uint32_t r = 0; r < n_reads; ++r) {
        uint32_t s0 = read_seg ? read_seg[r] : r, s1 = read_seg ? read_seg[r + 1] : r + 1;
        const char *a = nullptr, *b = nullptr; size_t na = 0, nb = 0;
        if (s1 > s0) { a = bases + seg_off[s0]; na = seg_off[s0 + 1] - seg_off[s0]; }
        if (s1 > s0 + 1) { b = bases + seg_off[s0 + 1]; nb = seg_off[s0 + 2] - seg_off[s0 + 1]; }
        float cr = compression ? compression[r] : 0.0f;
        if (with_gzip && na + nb) {
            std::string up(na + nb, 'N');
            for (size_t i = 0; i < na; ++i) up[i] = dna5_char(dna5_rank((unsigned char)a[i]));
            for (size_t i = 0; i < nb; ++i) up[na + i] = dna5_char(dna5_rank((unsigned char)b[i]));
            cr = compression_ratio(up.data(), up.size());
        }
        if (compression_out) compression_out[r] = cr;
        ReadEntry e = process_read(idx, "", a, na, b, nb, mean_q ? mean_q[r] : 0.0f, cr);
        e.apply_model(model);
        if (opt.dehost) e.call_host(model, host_idx); else e.call_category(model);
        num_hashes[r] = e.num_hashes;
        for (size_t c = 0; c < ncat; ++c) {
            counts[r * ncat + c] = e.counts[c]; unique_counts[r * ncat + c] = e.unique_counts[c];
            if (prob) prob[r * ncat + c] = e.probabilities[c];
        }
        if (call) call[r] = e.call;
        if (confidence) confidence[r] = e.confidence_score;
    }
    return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
}

// Whole `charon dehost|classify` single-end / paired run at -t 1 semantics over in-memory records:
// Result state machine + TSV text (include/result.hpp, read_entry.hpp:322-337).  ids = '\n'-joined headers.
// Returns a malloc'd NUL-terminated TSV (free with orc_free).
char *orc_run_tsv(orc_index *h, const orc_options *o, const char *ids, const char *bases, const uint64_t *seg_off,
                  const uint32_t *read_seg, uint32_t n_reads, const float *mean_q, const float *compression) {
    Options opt = to_opt(o);
    const Index &idx = h->idx;
    Result res(opt, idx.summary);
    res.keep_rows = false;
    const char *p = ids;
    for (uint32_t r = 0; r < n_reads; ++r) {
        const char *q = std::strchr(p, '\n');
        std::string id = q ? std::string(p, q) : std::string(p);
        p = q ? q + 1 : p + id.size();
        uint32_t s0 = read_seg ? read_seg[r] : r, s1 = read_seg ? read_seg[r + 1] : r + 1;
        const char *a = nullptr, *b = nullptr; size_t na = 0, nb = 0;
        if (s1 > s0) { a = bases + seg_off[s0]; na = seg_off[s0 + 1] - seg_off[s0]; }
        if (s1 > s0 + 1) { b = bases + seg_off[s0 + 1]; nb = seg_off[s0 + 2] - seg_off[s0 + 1]; }
        if (na + nb == 0) continue;   // zero-length reads are skipped (src/dehost_main.cpp:351-354)
        ReadEntry e = process_read(idx, id, a, na, b, nb, mean_q[r], compression[r]);
        res.add_read(e);
    }
    res.complete();
    char *out = static_cast<char *>(std::malloc(res.out.size() + 1));
    std::memcpy(out, res.out.c_str(), res.out.size() + 1);
    return out;
}
// The reference program's own shape over an in-memory FASTQ text (src/dehost_main.cpp:314-381): records parsed serially,
// chunks of `chunk` reads (default 100, include/dehost_arguments.hpp:22), `omp parallel for` over a chunk with the per-read
// scalars (id token, length, mean quality :355-360, gzip ratio of the upper-case dna5 string :361-363), the hot path
// (:366-373) and, inside ONE critical section (:374), model + call + TSV row.  The TSV goes to `tsv_path` (tmpfs file or
// /dev/null; NULL = discarded).  stage_s[5] = wall seconds of {parse, whole loop}, then CPU seconds summed over threads of
// {gzip ratio, minimiser + query + count, model + call + print}.  Returns the number of reads seen, <0 on error.
long long orc_run_like_reference(orc_index *h, const orc_options *o, const char *fastq, uint64_t fastq_len, int threads, uint32_t chunk,
                                 const char *tsv_path, double *stage_s) {
    using clk = std::chrono::steady_clock;
    auto secs = [](clk::time_point a, clk::time_point b) { return std::chrono::duration<double>(b - a).count(); };
    Options opt = to_opt(o);
    const Index &idx = h->idx;
    Result res(opt, idx.summary);
    res.keep_rows = false;
    struct Rec { const char *id; size_t id_len; const char *seq; size_t seq_len; const char *qual; };
    std::vector<Rec> recs;
    const auto t0 = clk::now();
    {   // FASTQ, four lines per record (what sequence_file_input sees in the synthetic files)
        const char *p = fastq, *end = fastq + fastq_len;
        auto line = [&](const char *&b, size_t &n) -> bool {
            if (p >= end) return false;
            const char *q = static_cast<const char *>(std::memchr(p, '\n', size_t(end - p)));
            if (!q) q = end;
            b = p; n = size_t(q - p); p = q < end ? q + 1 : end;
            return true;
        };
        for (;;) {
            Rec r{}; const char *plus; size_t nplus, nq;
            if (!line(r.id, r.id_len)) break;
            if (!r.id_len || r.id[0] != '@') return -1;
            ++r.id; --r.id_len;
            if (!line(r.seq, r.seq_len) || !line(plus, nplus) || !line(r.qual, nq) || nq != r.seq_len) return -1;
            recs.push_back(r);
        }
    }
    const auto t1 = clk::now();
    FILE *out = tsv_path ? std::fopen(tsv_path, "w") : nullptr;
    if (tsv_path && !out) return -2;
    double cpu_gzip = 0, cpu_hot = 0, cpu_model = 0;
    const uint8_t host_idx = idx.summary.host_category_index();
    (void)host_idx;
    if (!chunk) chunk = 100;
    for (size_t c0 = 0; c0 < recs.size(); c0 += chunk) {
        const size_t c1 = std::min(recs.size(), c0 + size_t(chunk));
#pragma omp parallel for schedule(static) num_threads(threads > 0 ? threads : 1) reduction(+ : cpu_gzip, cpu_hot)
        for (size_t i = c0; i < c1; ++i) {
            const Rec &r = recs[i];
            if (r.seq_len == 0) continue;                                  // :351-354
            const auto a = clk::now();
            std::string id(r.id, r.id_len);
            uint64_t qsum = 0;
            for (size_t j = 0; j < r.seq_len; ++j) qsum += uint64_t((unsigned char)r.qual[j] - 33);
            const float mean_q = float(qsum / r.seq_len);                  // integer division, then float (:355-360)
            std::string up(r.seq_len, 'N');
            for (size_t j = 0; j < r.seq_len; ++j) up[j] = dna5_char(dna5_rank((unsigned char)r.seq[j]));
            const float cr = compression_ratio(up.data(), up.size());
            const auto b = clk::now();
            ReadEntry e = process_read(idx, id, r.seq, r.seq_len, nullptr, 0, mean_q, cr);
            const auto c = clk::now();
            cpu_gzip += secs(a, b); cpu_hot += secs(b, c);
#pragma omp critical(add_read_to_results)
            {
                const auto d = clk::now();
                const size_t before = res.out.size();
                res.add_read(e);
                if (res.out.size() > before) {
                    if (out) std::fwrite(res.out.data(), 1, res.out.size(), out);
                    res.out.clear();
                }
                cpu_model += secs(d, clk::now());
            }
        }
    }
    res.complete();
    if (out) { std::fwrite(res.out.data(), 1, res.out.size(), out); std::fclose(out); }
    const auto t2 = clk::now();
    if (stage_s) { stage_s[0] = secs(t0, t1); stage_s[1] = secs(t1, t2); stage_s[2] = cpu_gzip; stage_s[3] = cpu_hot; stage_s[4] = cpu_model; }
    return (long long)recs.size();
}
void orc_free(void *p) { std::free(p); }
int orc_max_threads() {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

}  // extern "C"
