// charon_b200/host/index_file.hpp — reader of the `.idx` file written by `charon index`
// (reference include/index.hpp:122-138 Index::serialize, include/input_summary.hpp:64-78,
// include/input_stats.hpp:50-60, src/load_index.cpp:8-15; byte layout: SURVEY Appendix A.4).
//
// The file is a cereal BinaryOutputArchive: native-endian, no header, no versioning.  The Interleaved
// Bloom Filter payload is an sdsl::sd_vector (Elias-Fano); this loader does NOT decode it on the host:
// it keeps m_low / m_high as raw 64-bit words and hands them to the GPU (CG_FILTER_SD_VECTOR), which
// decodes them straight into the compacted HBM layout.  The two select supports that follow in the file
// are not needed for decoding and are skipped.
#pragma once
#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <limits>
#include <map>
#include <string>
#include <utility>
#include <vector>

namespace charon_host {

struct IndexFile {
    uint8_t window_size = 0, kmer_size = 0;   // index.hpp:21-22
    double max_fpr = 0;                       // index.hpp:23
    // InputSummary (input_summary.hpp:16-20)
    uint8_t num_bins = 0;
    std::vector<std::string> categories;
    std::vector<std::pair<std::string, uint8_t>> filepath_to_bin;
    std::map<uint8_t, std::string> bin_to_category;
    // InputStats (input_stats.hpp:17-19)
    uint32_t num_files = 0;
    std::map<uint8_t, uint64_t> records_per_bin, hashes_per_bin;
    // seqan3::interleaved_bloom_filter header
    uint64_t bins = 0, technical_bins = 0, bin_size = 0, hash_shift = 0, bin_words = 0, hash_funs = 0;
    // sdsl::sd_vector parts
    uint64_t sd_size = 0;
    uint8_t sd_wl = 0;
    std::vector<uint64_t> sd_low, sd_high;
    uint64_t sd_low_bits = 0, sd_high_bits = 0;

    uint8_t num_categories() const { return uint8_t(categories.size()); }
    // InputSummary::category_index (input_summary.hpp:39-45)
    uint8_t category_index(const std::string &c) const {
        for (size_t i = 0; i < categories.size(); ++i) if (categories[i] == c) return uint8_t(i);
        return std::numeric_limits<uint8_t>::max();
    }
    // InputSummary::host_category_index (input_summary.hpp:47-55)
    uint8_t host_category_index() const { return std::min(category_index("human"), category_index("host")); }
    // InputSummary::category_name (input_summary.hpp:57-62)
    std::string category_name(uint8_t i) const { return i >= categories.size() ? std::string() : categories[i]; }
    // category index of every bin (what ReadEntry::get_counts looks up per bin, read_entry.hpp:104-107)
    bool bin_to_cat(std::vector<uint8_t> &out, std::string &err) const {
        out.assign(bins, 0);
        for (uint64_t b = 0; b < bins; ++b) {
            auto it = bin_to_category.find(uint8_t(b));
            if (it == bin_to_category.end()) { err = "index: bin " + std::to_string(b) + " has no category"; return false; }
            const uint8_t c = category_index(it->second);
            if (c == 255) { err = "index: bin category '" + it->second + "' is not in the category list"; return false; }
            out[b] = c;
        }
        return true;
    }
};

namespace detail {
struct Reader {
    FILE *f = nullptr;
    bool ok = true;
    ~Reader() { if (f) std::fclose(f); }
    template <class T> T get() { T v{}; if (std::fread(&v, sizeof(T), 1, f) != 1) ok = false; return v; }
    std::string str() {
        const uint64_t n = get<uint64_t>();
        if (!ok || n > (1u << 20)) { ok = false; return std::string(); }
        std::string s(size_t(n), '\0');
        if (n && std::fread(&s[0], 1, size_t(n), f) != n) ok = false;
        return s;
    }
    bool words(std::vector<uint64_t> &v, uint64_t bits) {
        v.assign(size_t((bits + 63) / 64), 0);
        if (!v.empty() && std::fread(v.data(), 8, v.size(), f) != v.size()) ok = false;
        return ok;
    }
};
}  // namespace detail

// load_index (src/load_index.cpp:8-15).  Every structural invariant is checked; a file that fails one
// is rejected with a message instead of being half-trusted.
inline bool load_index_file(const std::string &path, IndexFile &x, std::string &err) {
    detail::Reader in;
    in.f = std::fopen(path.c_str(), "rb");
    if (!in.f) { err = "cannot open index file " + path; return false; }
    x = IndexFile();
    x.window_size = in.get<uint8_t>(); x.kmer_size = in.get<uint8_t>(); x.max_fpr = in.get<double>();
    x.num_bins = in.get<uint8_t>();
    uint64_t n = in.get<uint64_t>();
    if (!in.ok || n == 0 || n > 256) { err = "index: bad category count"; return false; }
    for (uint64_t i = 0; i < n; ++i) x.categories.push_back(in.str());
    n = in.get<uint64_t>();
    if (!in.ok || n > (1u << 24)) { err = "index: bad file list"; return false; }
    for (uint64_t i = 0; i < n && in.ok; ++i) { std::string p = in.str(); const uint8_t b = in.get<uint8_t>(); x.filepath_to_bin.emplace_back(std::move(p), b); }
    n = in.get<uint64_t>();
    if (!in.ok || n > 256) { err = "index: bad bin->category map"; return false; }
    for (uint64_t i = 0; i < n && in.ok; ++i) { const uint8_t b = in.get<uint8_t>(); x.bin_to_category[b] = in.str(); }
    x.num_files = in.get<uint32_t>();
    n = in.get<uint64_t>();
    if (!in.ok || n > 256) { err = "index: bad records_per_bin"; return false; }
    for (uint64_t i = 0; i < n && in.ok; ++i) { const uint8_t b = in.get<uint8_t>(); x.records_per_bin[b] = in.get<uint64_t>(); }
    n = in.get<uint64_t>();
    if (!in.ok || n > 256) { err = "index: bad hashes_per_bin"; return false; }
    for (uint64_t i = 0; i < n && in.ok; ++i) { const uint8_t b = in.get<uint8_t>(); x.hashes_per_bin[b] = in.get<uint64_t>(); }
    x.bins = in.get<uint64_t>(); x.technical_bins = in.get<uint64_t>(); x.bin_size = in.get<uint64_t>();
    x.hash_shift = in.get<uint64_t>(); x.bin_words = in.get<uint64_t>(); x.hash_funs = in.get<uint64_t>();
    if (!in.ok) { err = "index: truncated header"; return false; }
    if (x.bins == 0 || x.bins > 256 || x.bin_words != (x.bins + 63) / 64 || x.technical_bins != x.bin_words * 64) { err = "index: IBF header inconsistent (bins/bin_words/technical_bins)"; return false; }
    if (x.bin_size == 0 || x.hash_shift != uint64_t(__builtin_clzll(x.bin_size))) { err = "index: IBF hash_shift != countl_zero(bin_size)"; return false; }
    if (x.hash_funs == 0 || x.hash_funs > 5) { err = "index: IBF hash_funs out of range"; return false; }
    if (x.kmer_size == 0 || x.window_size < x.kmer_size) { err = "index: window_size < kmer_size"; return false; }
    x.sd_size = in.get<uint64_t>(); x.sd_wl = in.get<uint8_t>();
    if (!in.ok || x.sd_size != x.technical_bins * x.bin_size) { err = "index: sd_vector size != technical_bins * bin_size"; return false; }
    const uint8_t lw = in.get<uint8_t>(); (void)in.get<float>(); x.sd_low_bits = in.get<uint64_t>();
    if (!in.ok || lw != x.sd_wl || x.sd_wl > 64) { err = "index: sd_vector m_low width != wl"; return false; }
    if (!in.words(x.sd_low, x.sd_low_bits)) { err = "index: truncated m_low"; return false; }
    const uint8_t hw = in.get<uint8_t>(); (void)in.get<float>(); x.sd_high_bits = in.get<uint64_t>();
    if (!in.ok || hw != 1) { err = "index: sd_vector m_high is not a bit vector"; return false; }
    if (!in.words(x.sd_high, x.sd_high_bits)) { err = "index: truncated m_high"; return false; }
    return true;
}

}  // namespace charon_host
