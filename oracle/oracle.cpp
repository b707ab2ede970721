// oracle/oracle.cpp — TEST INFRASTRUCTURE (see oracle.hpp header).  Parity unpinned.
// Index file I/O, gzip ratio, index construction.
#include "oracle.hpp"

#include <zlib.h>

#include <cstdio>
#include <fstream>
#include <stdexcept>
#ifdef _OPENMP
#include <omp.h>
#endif

#include "../charon_b200/csrc/kde_defaults.inc"

namespace orc {

const float KDE_DEFAULT_POS[500] = {CHARON_KDE_DEFAULT_POS};
const float KDE_DEFAULT_NEG[500] = {CHARON_KDE_DEFAULT_NEG};

// reference src/utils.cpp:114-124 -> gzip::compress(ptr, size) of mapbox/gzip-hpp@7546b35:
// deflateInit2(Z_DEFAULT_COMPRESSION, Z_DEFLATED, 15|16, 8, Z_DEFAULT_STRATEGY), deflate(Z_FINISH).
float compression_ratio(const char *seq, size_t n) {
    z_stream zs;
    std::memset(&zs, 0, sizeof(zs));
    if (deflateInit2(&zs, Z_DEFAULT_COMPRESSION, Z_DEFLATED, 15 | 16, 8, Z_DEFAULT_STRATEGY) != Z_OK)
        throw std::runtime_error("deflate init failed");
    std::vector<unsigned char> buf(deflateBound(&zs, uLong(n)) + 64);
    zs.next_in = reinterpret_cast<Bytef *>(const_cast<char *>(seq));
    zs.avail_in = uInt(n);
    zs.next_out = buf.data();
    zs.avail_out = uInt(buf.size());
    int rc = deflate(&zs, Z_FINISH);
    size_t out = zs.total_out;
    deflateEnd(&zs);
    if (rc != Z_STREAM_END) throw std::runtime_error("deflate failed");
    return float(static_cast<double>(out) / static_cast<double>(n));
}

// reference src/utils.cpp:75-90
size_t bin_size_in_bits(uint64_t num_elements, unsigned num_hash, double max_fpr, uint64_t max_bits) {
    double const numerator{-static_cast<double>(num_elements * num_hash)};
    double const denominator{std::log(1 - std::exp(std::log(max_fpr) / num_hash))};
    double const result{std::ceil(numerator / denominator)};
    if (result > double(max_bits)) return size_t(max_bits);
    return size_t(result);
}

// ---------------------------------------------------------------------------------------------
// cereal BinaryOutputArchive layout (little-endian, no header) — SURVEY Appendix A.4
// ---------------------------------------------------------------------------------------------
namespace {
struct Out {
    std::ofstream f;
    template <class T> void put(const T &v) { f.write(reinterpret_cast<const char *>(&v), sizeof(T)); }
    void str(const std::string &s) { put<uint64_t>(s.size()); f.write(s.data(), std::streamsize(s.size())); }
};
struct In {
    std::ifstream f;
    bool ok = true;
    template <class T> T get() { T v{}; f.read(reinterpret_cast<char *>(&v), sizeof(T)); if (!f) ok = false; return v; }
    std::string str() { uint64_t n = get<uint64_t>(); if (!ok || n > (1u << 20)) { ok = false; return ""; } std::string s(n, 0); f.read(&s[0], std::streamsize(n)); if (!f) ok = false; return s; }
};
inline unsigned hi_bit(uint64_t x) { return x ? 63u - unsigned(__builtin_clzll(x)) : 0u; }

// sdsl::select_support_mcl<b,1> serialisation restated from memory (sdsl-lite v3 select_support_mcl.hpp):
// m_arg_cnt, m_logn, m_logn2, m_logn4, then if m_arg_cnt>0: m_superblock (int_vector<0>), and per
// superblock either a long-superblock int_vector or a miniblock int_vector.  Our loader never reads
// these (decode needs only low/high); we emit a structurally plausible but *unverified* encoding.
void put_int_vector(Out &o, uint8_t width, const std::vector<uint64_t> &vals) {
    o.put<uint8_t>(width);
    o.put<float>(1.5f);
    uint64_t bits = uint64_t(vals.size()) * width;
    o.put<uint64_t>(bits);
    std::vector<uint64_t> words((bits + 63) / 64, 0);
    for (size_t i = 0; i < vals.size(); ++i) {
        uint64_t p = uint64_t(i) * width, v = width == 64 ? vals[i] : (vals[i] & ((uint64_t(1) << width) - 1));
        words[p >> 6] |= v << (p & 63);
        if ((p & 63) + width > 64) words[(p >> 6) + 1] |= v >> (64 - (p & 63));
    }
    o.f.write(reinterpret_cast<const char *>(words.data()), std::streamsize(words.size() * 8));
}
void put_select_support(Out &o, const std::vector<uint64_t> &high_words, uint64_t high_bits, int bit) {
    std::vector<uint64_t> pos;
    for (uint64_t i = 0; i < high_bits; ++i)
        if (int((high_words[i >> 6] >> (i & 63)) & 1) == bit) pos.push_back(i);
    uint64_t arg_cnt = pos.size();
    o.put<uint64_t>(arg_cnt);
    uint8_t logn = uint8_t(hi_bit(high_bits ? high_bits : 1) + 1);
    uint8_t logn2 = uint8_t(uint32_t(logn) * logn), logn4 = uint8_t(uint32_t(logn2) * logn2);
    o.put<uint8_t>(logn); o.put<uint8_t>(logn2); o.put<uint8_t>(logn4);
    if (!arg_cnt) return;
    uint64_t sb = (arg_cnt + 4095) >> 12;
    std::vector<uint64_t> super(sb);
    for (uint64_t i = 0; i < sb; ++i) super[i] = pos[i << 12];
    put_int_vector(o, logn, super);
    for (uint64_t i = 0; i < sb; ++i) {
        uint64_t lo = i << 12, hi = std::min<uint64_t>(arg_cnt, lo + 4096);
        uint64_t span = pos[hi - 1] - pos[lo];
        bool is_long = span > uint64_t(logn4 ? logn4 : 1);
        o.put<uint8_t>(is_long ? 0 : 1);  // "has miniblock" discriminator
        std::vector<uint64_t> v;
        if (is_long) { for (uint64_t j = lo; j < lo + 4096; ++j) v.push_back(j < hi ? pos[j] : 0); put_int_vector(o, logn, v); }
        else { for (uint64_t j = lo; j < hi; j += 64) v.push_back(pos[j] - pos[lo]); v.resize(64, 0); put_int_vector(o, uint8_t(hi_bit(span ? span : 1) + 1), v); }
    }
}
}  // namespace

void write_index(const Index &idx, const std::string &path) {
    Out o; o.f.open(path, std::ios::binary);
    if (!o.f) throw std::runtime_error("cannot open " + path);
    o.put<uint8_t>(idx.window_size); o.put<uint8_t>(idx.kmer_size); o.put<double>(idx.max_fpr);  // index.hpp:125-127
    const Summary &s = idx.summary;                                                                 // input_summary.hpp:67-70
    o.put<uint8_t>(s.num_bins);
    o.put<uint64_t>(s.categories.size()); for (auto &c : s.categories) o.str(c);
    o.put<uint64_t>(s.filepath_to_bin.size()); for (auto &p : s.filepath_to_bin) { o.str(p.first); o.put<uint8_t>(p.second); }
    o.put<uint64_t>(s.bin_to_category.size()); for (auto &p : s.bin_to_category) { o.put<uint8_t>(p.first); o.str(p.second); }
    o.put<uint32_t>(idx.stats.num_files);                                                           // input_stats.hpp:54-56
    o.put<uint64_t>(idx.stats.records_per_bin.size()); for (auto &p : idx.stats.records_per_bin) { o.put<uint8_t>(p.first); o.put<uint64_t>(p.second); }
    o.put<uint64_t>(idx.stats.hashes_per_bin.size()); for (auto &p : idx.stats.hashes_per_bin) { o.put<uint8_t>(p.first); o.put<uint64_t>(p.second); }
    const Ibf &f = idx.ibf;                                                                         // index.hpp:130
    o.put<uint64_t>(f.bins); o.put<uint64_t>(f.technical_bins); o.put<uint64_t>(f.bin_size);
    o.put<uint64_t>(f.hash_shift); o.put<uint64_t>(f.bin_words); o.put<uint64_t>(f.hash_funs);
    // sd_vector(bit_vector) construction: n = size, m = ones; wl = logn - logm (floor logs, +1 if pow2 mismatch)
    uint64_t n = f.technical_bins * f.bin_size, m = 0;
    for (uint64_t w : f.data) m += uint64_t(__builtin_popcountll(w));
    uint8_t logm = uint8_t(hi_bit(m) + 1), logn = uint8_t(hi_bit(n) + 1);
    if (logm == logn) logm = uint8_t(logm - 1);
    uint8_t wl = uint8_t(logn - logm);
    o.put<uint64_t>(n); o.put<uint8_t>(wl);
    // m_low: m values of wl bits ; m_high: m + 2^logm bits, bit (pos>>wl)+j set for j-th one
    o.put<uint8_t>(wl); o.put<float>(1.5f); o.put<uint64_t>(m * wl);
    std::vector<uint64_t> low((m * wl + 63) / 64, 0);
    uint64_t high_bits = m + (uint64_t(1) << logm);
    std::vector<uint64_t> high((high_bits + 63) / 64, 0);
    uint64_t j = 0, lowmask = wl ? ((wl == 64) ? ~uint64_t(0) : ((uint64_t(1) << wl) - 1)) : 0;
    for (uint64_t wi = 0; wi < f.data.size(); ++wi) {
        uint64_t w = f.data[wi];
        while (w) {
            uint64_t pos = wi * 64 + uint64_t(__builtin_ctzll(w)); w &= w - 1;
            if (wl) {
                uint64_t p = j * wl, v = pos & lowmask;
                low[p >> 6] |= v << (p & 63);
                if ((p & 63) + wl > 64) low[(p >> 6) + 1] |= v >> (64 - (p & 63));
            }
            uint64_t hb = (pos >> wl) + j;
            high[hb >> 6] |= uint64_t(1) << (hb & 63);
            ++j;
        }
    }
    o.f.write(reinterpret_cast<const char *>(low.data()), std::streamsize(low.size() * 8));
    o.put<uint8_t>(1); o.put<float>(1.5f); o.put<uint64_t>(high_bits);
    o.f.write(reinterpret_cast<const char *>(high.data()), std::streamsize(high.size() * 8));
    put_select_support(o, high, high_bits, 1);
    put_select_support(o, high, high_bits, 0);
    if (!o.f) throw std::runtime_error("write failed: " + path);
}

bool read_index(Index &idx, const std::string &path, std::string &err) {
    In in; in.f.open(path, std::ios::binary);
    if (!in.f) { err = "cannot open " + path; return false; }
    idx = Index();
    idx.window_size = in.get<uint8_t>(); idx.kmer_size = in.get<uint8_t>(); idx.max_fpr = in.get<double>();
    Summary &s = idx.summary;
    s.num_bins = in.get<uint8_t>();
    uint64_t n = in.get<uint64_t>(); if (n > 256) { err = "bad category count"; return false; }
    for (uint64_t i = 0; i < n; ++i) s.categories.push_back(in.str());
    n = in.get<uint64_t>(); if (n > 65536) { err = "bad file count"; return false; }
    for (uint64_t i = 0; i < n; ++i) { std::string p = in.str(); uint8_t b = in.get<uint8_t>(); s.filepath_to_bin.emplace_back(p, b); }
    n = in.get<uint64_t>(); if (n > 256) { err = "bad bin map"; return false; }
    for (uint64_t i = 0; i < n; ++i) { uint8_t b = in.get<uint8_t>(); s.bin_to_category[b] = in.str(); }
    idx.stats.num_files = in.get<uint32_t>();
    n = in.get<uint64_t>(); if (n > 256) { err = "bad stats"; return false; }
    for (uint64_t i = 0; i < n; ++i) { uint8_t b = in.get<uint8_t>(); idx.stats.records_per_bin[b] = in.get<uint64_t>(); }
    n = in.get<uint64_t>(); if (n > 256) { err = "bad stats"; return false; }
    for (uint64_t i = 0; i < n; ++i) { uint8_t b = in.get<uint8_t>(); idx.stats.hashes_per_bin[b] = in.get<uint64_t>(); }
    Ibf &f = idx.ibf;
    f.bins = in.get<uint64_t>(); f.technical_bins = in.get<uint64_t>(); f.bin_size = in.get<uint64_t>();
    f.hash_shift = in.get<uint64_t>(); f.bin_words = in.get<uint64_t>(); f.hash_funs = in.get<uint64_t>();
    if (!in.ok || f.technical_bins != f.bin_words * 64 || f.bins > f.technical_bins) { err = "bad ibf header"; return false; }
    uint64_t size = in.get<uint64_t>(); uint8_t wl = in.get<uint8_t>();
    if (size != f.technical_bins * f.bin_size) { err = "sd_vector size mismatch"; return false; }
    uint8_t lw = in.get<uint8_t>(); (void)in.get<float>(); uint64_t lbits = in.get<uint64_t>();
    if (!in.ok || lw != wl) { err = "m_low width mismatch"; return false; }
    std::vector<uint64_t> low((lbits + 63) / 64);
    in.f.read(reinterpret_cast<char *>(low.data()), std::streamsize(low.size() * 8));
    uint8_t hw = in.get<uint8_t>(); (void)in.get<float>(); uint64_t hbits = in.get<uint64_t>();
    if (!in.ok || hw != 1) { err = "m_high width mismatch"; return false; }
    std::vector<uint64_t> high((hbits + 63) / 64);
    in.f.read(reinterpret_cast<char *>(high.data()), std::streamsize(high.size() * 8));
    if (!in.f) { err = "truncated"; return false; }
    uint64_t ones = wl ? lbits / wl : 0;
    f.data.assign(f.bin_words * f.bin_size, 0);
    uint64_t j = 0;
    for (uint64_t wi = 0; wi < high.size(); ++wi) {
        uint64_t w = high[wi];
        while (w) {
            uint64_t q = wi * 64 + uint64_t(__builtin_ctzll(w)); w &= w - 1;
            uint64_t lo = 0;
            if (wl) { uint64_t p = j * wl; lo = low[p >> 6] >> (p & 63); if ((p & 63) + wl > 64) lo |= low[(p >> 6) + 1] << (64 - (p & 63)); lo &= (wl == 64 ? ~uint64_t(0) : ((uint64_t(1) << wl) - 1)); }
            uint64_t pos = ((q - j) << wl) | lo;
            if (pos >= size) { err = "position out of range"; return false; }
            f.data[pos >> 6] |= uint64_t(1) << (pos & 63);
            ++j;
        }
    }
    if (wl && j != ones) { err = "ones count mismatch"; return false; }
    return true;
}

// ---------------------------------------------------------------------------------------------
// charon index (reference src/index_main.cpp:118-263): distinct minimisers per bin, bin_size from
// the largest bin (bin_size_in_bits), emplace everything.  `bin_sequences[b]` = records of bin b.
// ---------------------------------------------------------------------------------------------
Index build_index(const std::vector<std::pair<std::string, std::string>> &, const std::vector<std::vector<std::string>> &bin_sequences,
                  const std::vector<std::string> &bin_category, const std::vector<std::string> &category_order, unsigned k,
                  unsigned w, int threads) {
    Index idx;
    idx.window_size = uint8_t(w); idx.kmer_size = uint8_t(k); idx.max_fpr = 0.01;
    size_t nb = bin_sequences.size();
    idx.summary.num_bins = uint8_t(nb);
    idx.summary.categories = category_order;
    std::vector<std::vector<uint64_t>> hashes(nb);
    (void)threads;
    // records are hashed independently (the reference parallelises over files, :129); per-bin distinct set afterwards
    std::vector<std::pair<size_t, size_t>> flat;
    for (size_t b = 0; b < nb; ++b) for (size_t r = 0; r < bin_sequences[b].size(); ++r) flat.emplace_back(b, r);
    std::vector<std::vector<uint64_t>> per_rec(flat.size());
#pragma omp parallel for schedule(dynamic) num_threads(threads)
    for (size_t f = 0; f < flat.size(); ++f) {
        const std::string &rec = bin_sequences[flat[f].first][flat[f].second];
        std::vector<uint8_t> ranks(rec.size());
        for (size_t i = 0; i < rec.size(); ++i) ranks[i] = dna5_rank((unsigned char)rec[i]);
        minimiser_hash(ranks.data(), ranks.size(), k, w, per_rec[f]);
        std::sort(per_rec[f].begin(), per_rec[f].end());
        per_rec[f].erase(std::unique(per_rec[f].begin(), per_rec[f].end()), per_rec[f].end());
    }
#pragma omp parallel for schedule(dynamic) num_threads(threads)
    for (size_t b = 0; b < nb; ++b) {
        std::vector<uint64_t> &hs = hashes[b];
        for (size_t f = 0; f < flat.size(); ++f) if (flat[f].first == b) { hs.insert(hs.end(), per_rec[f].begin(), per_rec[f].end()); std::vector<uint64_t>().swap(per_rec[f]); }
        std::sort(hs.begin(), hs.end());
        hs.erase(std::unique(hs.begin(), hs.end()), hs.end());  // unordered_set<uint64_t> semantics (:141-149)
    }
    uint64_t max_hashes = 0;
    for (size_t b = 0; b < nb; ++b) {
        idx.summary.bin_to_category[uint8_t(b)] = bin_category[b];
        idx.summary.filepath_to_bin.emplace_back("/synthetic/bin" + std::to_string(b) + ".fa", uint8_t(b));
        idx.stats.records_per_bin[uint8_t(b)] = bin_sequences[b].size();
        idx.stats.hashes_per_bin[uint8_t(b)] = hashes[b].size();
        idx.stats.num_files += 1;
        max_hashes = std::max<uint64_t>(max_hashes, hashes[b].size());
    }
    idx.ibf.init(nb, bin_size_in_bits(max_hashes), 3);   // index_arguments.hpp:20-22, index_main.cpp:240-245
    for (size_t b = 0; b < nb; ++b) {
        const std::vector<uint64_t> &hs = hashes[b];
#pragma omp parallel for schedule(static) num_threads(threads)
        for (size_t i = 0; i < hs.size(); ++i) idx.ibf.emplace(hs[i], b);   // atomic OR inside
    }
    return idx;
}

}  // namespace orc
