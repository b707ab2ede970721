// charon_b200/host/index_writer.hpp — writer of the `.idx` file (`charon index`; reference include/store_index.hpp:12-17,
// Index::serialize include/index.hpp:122-138).  Same cereal-binary layout the loader in index_file.hpp reads.
//
// The sd_vector parts (m_low / m_high) come from the GPU encoder (cg_export_sd_vector).  The two sdsl select supports that
// close the stream are rebuilt on the host from m_high.  Their byte layout is restated from memory of sdsl-lite v3
// (select_support_mcl.hpp) and could NOT be checked against the real library on this image: our own loader never reads them
// (decoding needs only m_low / m_high), a real `charon` build might — this is the one unverified part of the file.
#pragma once
#include <cstdint>
#include <cstdio>
#include <functional>
#include <stdexcept>
#include <string>
#include <vector>

#include "index_file.hpp"

namespace charon_host {

namespace detail {
struct Writer {
    FILE *f = nullptr;
    explicit Writer(const std::string &path) { f = std::fopen(path.c_str(), "wb"); if (!f) throw std::runtime_error("cannot open " + path + " for writing"); }
    ~Writer() { if (f) std::fclose(f); }
    template <class T> void put(const T &v) { if (std::fwrite(&v, sizeof(T), 1, f) != 1) throw std::runtime_error("index write failed"); }
    void str(const std::string &s) { put<uint64_t>(s.size()); if (!s.empty() && std::fwrite(s.data(), 1, s.size(), f) != s.size()) throw std::runtime_error("index write failed"); }
    void words(const uint64_t *w, uint64_t n) { if (n && std::fwrite(w, 8, n, f) != n) throw std::runtime_error("index write failed"); }
    // sdsl::int_vector<0>: u8 width, f32 growth factor, u64 size in bits, packed words
    void int_vector(uint8_t width, const std::vector<uint64_t> &vals) {
        put<uint8_t>(width); put<float>(1.5f);
        const uint64_t bits = uint64_t(vals.size()) * width;
        put<uint64_t>(bits);
        std::vector<uint64_t> w((bits + 63) / 64, 0);
        for (size_t i = 0; i < vals.size(); ++i) {
            const uint64_t p = uint64_t(i) * width, v = width >= 64 ? vals[i] : (vals[i] & ((uint64_t(1) << width) - 1));
            w[p >> 6] |= v << (p & 63);
            if ((p & 63) + width > 64) w[(p >> 6) + 1] |= v >> (64 - (p & 63));
        }
        words(w.data(), w.size());
    }
};
inline unsigned hi_bit(uint64_t x) { return x ? 63u - unsigned(__builtin_clzll(x)) : 0u; }

// select_support_mcl<bit, 1> over m_high: superblocks of 4096 occurrences; a superblock spanning more than log^4(n) bits
// stores every position ("long"), otherwise one offset per 64 occurrences ("miniblock").
inline void select_support(Writer &o, const std::vector<uint64_t> &high, uint64_t high_bits, int bit) {
    uint64_t cnt = 0;
    for (uint64_t i = 0; i < high.size(); ++i) cnt += uint64_t(__builtin_popcountll(bit ? high[i] : ~high[i]));
    if (!bit) cnt -= uint64_t(high.size()) * 64 - high_bits;   // padding bits of the last word are zeros, not members
    o.put<uint64_t>(cnt);
    const uint8_t logn = uint8_t(hi_bit(high_bits ? high_bits : 1) + 1);
    const uint8_t logn2 = uint8_t(uint32_t(logn) * logn), logn4 = uint8_t(uint32_t(logn2) * logn2);
    o.put<uint8_t>(logn); o.put<uint8_t>(logn2); o.put<uint8_t>(logn4);
    if (!cnt) return;
    const uint64_t sb = (cnt + 4095) >> 12;
    // two streaming passes over m_high: superblock table first, then one superblock (<= 4096 positions) at a time
    auto for_each_pos = [&](const std::function<void(uint64_t)> &fn) {
        for (uint64_t wi = 0; wi < high.size(); ++wi) {
            uint64_t w = bit ? high[wi] : ~high[wi];
            while (w) {
                const uint64_t pos = wi * 64 + uint64_t(__builtin_ctzll(w));
                w &= w - 1;
                if (pos >= high_bits) return;
                fn(pos);
            }
        }
    };
    std::vector<uint64_t> super;
    super.reserve(sb);
    uint64_t seen = 0;
    for_each_pos([&](uint64_t pos) { if ((seen++ & 4095) == 0) super.push_back(pos); });
    o.int_vector(logn, super);
    std::vector<uint64_t> block, v;
    block.reserve(4096);
    auto flush = [&]() {
        if (block.empty()) return;
        const uint64_t span = block.back() - block.front();
        const bool is_long = span > uint64_t(logn4 ? logn4 : 1);
        o.put<uint8_t>(is_long ? 0 : 1);
        v.clear();
        if (is_long) { v = block; v.resize(4096, 0); o.int_vector(logn, v); }
        else { for (size_t j = 0; j < block.size(); j += 64) v.push_back(block[j] - block.front()); v.resize(64, 0); o.int_vector(uint8_t(hi_bit(span ? span : 1) + 1), v); }
        block.clear();
    };
    for_each_pos([&](uint64_t pos) { block.push_back(pos); if (block.size() == 4096) flush(); });
    flush();
}
}  // namespace detail

// x: everything but the sd_vector payload words (taken from low / high).  Writes Index::serialize's field order.
inline void write_index_file(const std::string &path, const IndexFile &x, const std::vector<uint64_t> &low, const std::vector<uint64_t> &high) {
    detail::Writer o(path);
    o.put<uint8_t>(x.window_size); o.put<uint8_t>(x.kmer_size); o.put<double>(x.max_fpr);                 // index.hpp:125-127
    o.put<uint8_t>(x.num_bins);                                                                            // input_summary.hpp:67-70
    o.put<uint64_t>(x.categories.size()); for (const auto &c : x.categories) o.str(c);
    o.put<uint64_t>(x.filepath_to_bin.size()); for (const auto &p : x.filepath_to_bin) { o.str(p.first); o.put<uint8_t>(p.second); }
    o.put<uint64_t>(x.bin_to_category.size()); for (const auto &p : x.bin_to_category) { o.put<uint8_t>(p.first); o.str(p.second); }
    o.put<uint32_t>(x.num_files);                                                                          // input_stats.hpp:54-56
    o.put<uint64_t>(x.records_per_bin.size()); for (const auto &p : x.records_per_bin) { o.put<uint8_t>(p.first); o.put<uint64_t>(p.second); }
    o.put<uint64_t>(x.hashes_per_bin.size()); for (const auto &p : x.hashes_per_bin) { o.put<uint8_t>(p.first); o.put<uint64_t>(p.second); }
    o.put<uint64_t>(x.bins); o.put<uint64_t>(x.technical_bins); o.put<uint64_t>(x.bin_size);              // seqan3 IBF header
    o.put<uint64_t>(x.hash_shift); o.put<uint64_t>(x.bin_words); o.put<uint64_t>(x.hash_funs);
    o.put<uint64_t>(x.sd_size); o.put<uint8_t>(x.sd_wl);                                                   // sdsl::sd_vector
    o.put<uint8_t>(x.sd_wl); o.put<float>(1.5f); o.put<uint64_t>(x.sd_low_bits);
    o.words(low.data(), (x.sd_low_bits + 63) / 64);
    o.put<uint8_t>(1); o.put<float>(1.5f); o.put<uint64_t>(x.sd_high_bits);
    o.words(high.data(), (x.sd_high_bits + 63) / 64);
    detail::select_support(o, high, x.sd_high_bits, 1);
    detail::select_support(o, high, x.sd_high_bits, 0);
}

}  // namespace charon_host
