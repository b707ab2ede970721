// charon_b200/csrc/deflate_size.cuh — the SIZE of the gzip stream zlib would write for a read, without writing it.
//
// Reference: get_compression_ratio (src/utils.cpp:114-124) = gzip::compress of mapbox/gzip-hpp = zlib deflateInit2(
// Z_DEFAULT_COMPRESSION (level 6), Z_DEFLATED, 15 + 16, memLevel 8, Z_DEFAULT_STRATEGY) + one deflate(Z_FINISH); the ratio
// compressed_size / length gates every call (read_entry.hpp:176, 236) and is printed in every row.  zlib is a third-party
// dependency that is not in /root/reference; the algorithm restated here is zlib's deflate.c / trees.c (1.2.11 .. 1.3.1 agree
// on everything that decides the size at level 6):
//   * deflate_slow with configuration_table[6] = {good 8, lazy 16, nice 128, chain 128}: hash chains over 3-byte strings, lazy
//     match evaluation, TOO_FAR = 4096 for length-3 matches, position 0 never a match candidate (NIL), MAX_DIST = 32768 - 262;
//   * a block is closed after lit_bufsize - 1 = 16383 symbols and at the end of the input;
//   * per block build_tree (heap with depth tie-break) / gen_bitlen (overflow repair) / scan_tree / build_bl_tree give opt_len
//     and static_len; the block is stored if stored_len + 4 <= min(opt, static) bytes, else static if static <= opt, else dynamic;
//   * 10 bytes of gzip header, 8 of trailer.
// The uppercase dna5 string only holds A, C, G, T, N, so zlib's 15-bit hash ((b0 << 10) ^ (b1 << 5) ^ b2) & 0x7fff is
// collision-free on it and a hash chain is exactly "the earlier positions that start with the same three letters, most recent
// first".  That is what makes the match search parallel: the positions of a read are counting-sorted by their 3-letter code
// once (occ[], idx[]), and the up-to-128 chain candidates of a position are a contiguous run of occ[] a warp examines 32 at a
// time.  Everything sequential (the lazy-match state machine, the trees) is written once, DS_HD, and compiled for the device
// (gzip_size_kernel) and, for tests only, for the host (tests/deflate_size_host.cpp) where it is checked against the real
// zlib of the box; the product never runs the host build.
// Longer inputs than the 64 KB window: zlib slides the window by 32 KB whenever the lookahead runs low.  In absolute positions
// that changes nothing for the match search (a candidate further back than MAX_DIST is refused anyway, and the slot that
// becomes NIL after a slide is always further back than that); what it does change is that a block that began before the
// current window base can no longer be emitted as a stored block (zlib passes buf = NULL), which gzip_size tracks.
// Inputs longer than DSZ_MAX_LEN (16 M letters) are not handled: the caller gets DSZ_UNSUPPORTED.
#pragma once
#include <stdint.h>

#ifdef __CUDACC__
#define DS_HD __host__ __device__ __forceinline__
#else
#define DS_HD inline
#endif

namespace dsz {

DS_HD int msb32(uint32_t x) {     // index of the highest set bit, x != 0
#ifdef __CUDA_ARCH__
    return 31 - __clz(int(x));
#else
    return 31 - __builtin_clz(x);
#endif
}

constexpr int L_CODES = 286, D_CODES = 30, BL_CODES = 19, HEAP_SIZE = 2 * L_CODES + 1, MAX_BITS = 15, MAX_BL_BITS = 7;
constexpr int END_BLOCK = 256, REP_3_6 = 16, REPZ_3_10 = 17, REPZ_11_138 = 18;
constexpr int MIN_MATCH = 3, MAX_MATCH = 258, MIN_LOOKAHEAD = MAX_MATCH + MIN_MATCH + 1, W_SIZE = 32768;
constexpr uint32_t MAX_DIST = W_SIZE - MIN_LOOKAHEAD, TOO_FAR = 4096;
constexpr uint32_t SYMS_PER_BLOCK = 16384 - 1;          // lit_bufsize - 1 with memLevel 8
constexpr uint32_t GOOD_LENGTH = 8, MAX_LAZY = 16, NICE_LENGTH = 128, MAX_CHAIN = 128;
constexpr uint32_t DSZ_MAX_LEN = (1u << 24) - 1;        // positions are kept in 24 bits (gzip_size_kernel.cuh)
constexpr uint32_t DSZ_UNSUPPORTED = 0xFFFFFFFFu;
constexpr int N_LETTER_CODES = 125;                      // 5^3 three-letter strings over A C G T N

// length - 3 in [0, 255] -> length code 0..28 (trees.c tr_static_init: _length_code), and its extra bits
DS_HD int length_code(uint32_t lc) {
    if (lc < 8) return int(lc);
    if (lc == 255) return 28;
    const int e = msb32(lc) - 2;
    return 4 * (e + 1) + int((lc - (4u << e)) >> e);
}
DS_HD int extra_lbits(int code) { return (code < 8 || code == 28) ? 0 : (code >> 2) - 1; }
// distance - 1 in [0, 32767] -> distance code 0..29, and its extra bits
DS_HD int dist_code(uint32_t d) {
    if (d < 4) return int(d);
    const int e = msb32(d) - 1;
    return 2 * (e + 1) + int((d >> e) & 1u);
}
DS_HD int extra_dbits(int code) { return code < 4 ? 0 : (code >> 1) - 1; }
DS_HD int extra_blbits(int code) { return code < 16 ? 0 : (code == 16 ? 2 : (code == 17 ? 3 : 7)); }
DS_HD int static_llen(int n) { return n < 144 ? 8 : (n < 256 ? 9 : (n < 280 ? 7 : 8)); }
DS_HD int bl_order(int i) {
    // {16, 17, 18, 0, 8, 7, 9, 6, 10, 5, 11, 4, 12, 3, 13, 2, 14, 1, 15}
    if (i < 3) return 16 + i;
    if (i == 3) return 0;
    const int k = i - 4;            // 8, 7, 9, 6, 10, 5, ...
    return (k & 1) ? 7 - (k >> 1) : 8 + (k >> 1);
}

// Work area of one stream's trees (one per warp on the device).  zlib keeps Freq/Code and Dad/Len in unions; here they
// are separate arrays, which changes nothing as long as the read/write order of gen_bitlen is kept.
struct Trees {
    uint16_t lfreq[HEAP_SIZE], dfreq[2 * D_CODES + 1], blfreq[2 * BL_CODES + 1];
    uint8_t llen[HEAP_SIZE + 1], dlen[2 * D_CODES + 2], bllen[2 * BL_CODES + 1];    // +1: the guard scan_tree writes at max_code + 1 (0xff here, 0xffff in zlib: no length is either)
    uint16_t dad[HEAP_SIZE];
    uint16_t heap[HEAP_SIZE];
    uint8_t depth[HEAP_SIZE];
    uint16_t bl_count[MAX_BITS + 1];
    int heap_len, heap_max;
    uint32_t opt_len, static_len;      // bits
    uint32_t syms;                     // symbols tallied in the current block
};

DS_HD void init_block(Trees &T) {
    for (int n = 0; n < L_CODES; ++n) T.lfreq[n] = 0;
    for (int n = 0; n < D_CODES; ++n) T.dfreq[n] = 0;
    for (int n = 0; n < BL_CODES; ++n) T.blfreq[n] = 0;
    T.lfreq[END_BLOCK] = 1;
    T.opt_len = T.static_len = 0;
    T.syms = 0;
}

DS_HD bool smaller(const uint16_t *freq, const uint8_t *depth, int n, int m) {
    return freq[n] < freq[m] || (freq[n] == freq[m] && depth[n] <= depth[m]);
}
DS_HD void pqdownheap(Trees &T, const uint16_t *freq, int k) {
    const int v = T.heap[k];
    int j = k << 1;
    while (j <= T.heap_len) {
        if (j < T.heap_len && smaller(freq, T.depth, T.heap[j + 1], T.heap[j])) j++;
        if (smaller(freq, T.depth, v, T.heap[j])) break;
        T.heap[k] = T.heap[j];
        k = j;
        j <<= 1;
    }
    T.heap[k] = uint16_t(v);
}

// which: 0 literal/length tree, 1 distance tree, 2 bit-length tree
template <int WHICH>
DS_HD int build_tree(Trees &T) {
    uint16_t *freq = WHICH == 0 ? T.lfreq : (WHICH == 1 ? T.dfreq : T.blfreq);
    uint8_t *len = WHICH == 0 ? T.llen : (WHICH == 1 ? T.dlen : T.bllen);
    constexpr int elems = WHICH == 0 ? L_CODES : (WHICH == 1 ? D_CODES : BL_CODES);
    constexpr int max_length = WHICH == 2 ? MAX_BL_BITS : MAX_BITS;
    constexpr int base = WHICH == 0 ? 257 : 0;
    auto xbits_of = [](int n) -> int { return WHICH == 0 ? (n >= base ? extra_lbits(n - base) : 0) : (WHICH == 1 ? extra_dbits(n) : extra_blbits(n)); };
    auto static_len_of = [](int n) -> int { return WHICH == 0 ? static_llen(n) : 5; };
    int max_code = -1;
    T.heap_len = 0; T.heap_max = HEAP_SIZE;
    for (int n = 0; n < elems; ++n) {
        if (freq[n] != 0) { T.heap[++T.heap_len] = uint16_t(max_code = n); T.depth[n] = 0; }
        else len[n] = 0;
    }
    while (T.heap_len < 2) {
        const int node = T.heap[++T.heap_len] = uint16_t(max_code < 2 ? ++max_code : 0);
        freq[node] = 1; T.depth[node] = 0; T.opt_len--;
        if (WHICH != 2) T.static_len -= uint32_t(static_len_of(node));
    }
    for (int n = T.heap_len / 2; n >= 1; --n) pqdownheap(T, freq, n);
    int node = elems;
    do {
        const int n = T.heap[1];
        T.heap[1] = T.heap[T.heap_len--];
        pqdownheap(T, freq, 1);
        const int m = T.heap[1];
        T.heap[--T.heap_max] = uint16_t(n);
        T.heap[--T.heap_max] = uint16_t(m);
        freq[node] = uint16_t(freq[n] + freq[m]);
        T.depth[node] = uint8_t((T.depth[n] >= T.depth[m] ? T.depth[n] : T.depth[m]) + 1);
        T.dad[n] = T.dad[m] = uint16_t(node);
        T.heap[1] = uint16_t(node++);
        pqdownheap(T, freq, 1);
    } while (T.heap_len >= 2);
    T.heap[--T.heap_max] = T.heap[1];
    // gen_bitlen
    for (int b = 0; b <= MAX_BITS; ++b) T.bl_count[b] = 0;
    int overflow = 0, h;
    len[T.heap[T.heap_max]] = 0;
    for (h = T.heap_max + 1; h < HEAP_SIZE; ++h) {
        const int n = T.heap[h];
        int bits = len[T.dad[n]] + 1;
        if (bits > max_length) { bits = max_length; overflow++; }
        len[n] = uint8_t(bits);
        if (n > max_code) continue;
        T.bl_count[bits]++;
        const int xb = xbits_of(n);
        const uint32_t f = freq[n];
        T.opt_len += f * uint32_t(bits + xb);
        if (WHICH != 2) T.static_len += f * uint32_t(static_len_of(n) + xb);
    }
    if (overflow > 0) {
        do {
            int bits = max_length - 1;
            while (T.bl_count[bits] == 0) bits--;
            T.bl_count[bits]--;
            T.bl_count[bits + 1] = uint16_t(T.bl_count[bits + 1] + 2);
            T.bl_count[max_length]--;
            overflow -= 2;
        } while (overflow > 0);
        for (int bits = max_length; bits != 0; --bits) {
            int n = T.bl_count[bits];
            while (n != 0) {
                const int m = T.heap[--h];
                if (m > max_code) continue;
                if (len[m] != unsigned(bits)) {
                    T.opt_len += (uint32_t(bits) - uint32_t(len[m])) * uint32_t(freq[m]);
                    len[m] = uint8_t(bits);
                }
                n--;
            }
        }
    }
    return max_code;
}

DS_HD void scan_tree(Trees &T, uint8_t *len, int max_code) {
    int prevlen = -1, nextlen = len[0], count = 0, max_count = 7, min_count = 4;
    if (nextlen == 0) { max_count = 138; min_count = 3; }
    len[max_code + 1] = 0xFF;
    for (int n = 0; n <= max_code; ++n) {
        const int curlen = nextlen;
        nextlen = len[n + 1];
        if (++count < max_count && curlen == nextlen) continue;
        else if (count < min_count) T.blfreq[curlen] = uint16_t(T.blfreq[curlen] + count);
        else if (curlen != 0) { if (curlen != prevlen) T.blfreq[curlen]++; T.blfreq[REP_3_6]++; }
        else if (count <= 10) T.blfreq[REPZ_3_10]++;
        else T.blfreq[REPZ_11_138]++;
        count = 0; prevlen = curlen;
        if (nextlen == 0) { max_count = 138; min_count = 3; }
        else if (curlen == nextlen) { max_count = 6; min_count = 3; }
        else { max_count = 7; min_count = 4; }
    }
}

// _tr_flush_block: bits this block adds to the stream whose length so far is `bits` (stored blocks align to a byte first)
DS_HD uint64_t flush_block(Trees &T, uint64_t bits, uint32_t stored_len, bool have_buf) {
    const int lmax = build_tree<0>(T);
    const int dmax = build_tree<1>(T);
    scan_tree(T, T.llen, lmax);
    scan_tree(T, T.dlen, dmax);
    build_tree<2>(T);
    int max_blindex = BL_CODES - 1;
    for (; max_blindex >= 3; --max_blindex)
        if (T.bllen[bl_order(max_blindex)] != 0) break;
    T.opt_len += 3u * (uint32_t(max_blindex) + 1u) + 5 + 5 + 4;
    uint32_t opt_lenb = (T.opt_len + 3 + 7) >> 3;
    const uint32_t static_lenb = (T.static_len + 3 + 7) >> 3;
    if (static_lenb <= opt_lenb) opt_lenb = static_lenb;
    if (stored_len + 4 <= opt_lenb && have_buf) {
        bits += 3;
        bits = (bits + 7) & ~uint64_t(7);
        bits += 32 + uint64_t(stored_len) * 8;
    } else if (static_lenb == opt_lenb) {
        bits += 3 + uint64_t(T.static_len);
    } else {
        bits += 3 + uint64_t(T.opt_len);
    }
    init_block(T);
    return bits;
}

DS_HD void tally_lit(Trees &T, uint8_t c) { T.lfreq[c]++; T.syms++; }
DS_HD void tally_dist(Trees &T, uint32_t dist, uint32_t lc) {   // dist = match distance, lc = match length - MIN_MATCH
    T.lfreq[length_code(lc) + 256 + 1]++;
    T.dfreq[dist_code(dist - 1)]++;
    T.syms++;
}

// A C G T N (the only letters of an uppercase dna5 string) -> 0..4; three letters -> 0..124
DS_HD uint32_t letter_code(uint8_t c) { return c == 'A' ? 0u : (c == 'C' ? 1u : (c == 'G' ? 2u : (c == 'T' ? 3u : 4u))); }

// The lazy-match state machine of deflate_slow.  M supplies the match search, S receives the symbols:
//   bool M.head(p, prev_length, q)             most recent earlier position with the same three letters (false: none);
//                                              prev_length = the value longest() would get, a hint for what to fetch
//   uint32_t M.longest(p, prev_length, start)  zlib's longest_match at position p (start: in = previous match start, out = new)
//   uint8_t M.at(p)                            letter at position p
//   S.lit(c), S.dist(distance, length - 3)     _tr_tally;  uint64_t S.flush(bits, stored_len, have_buf)  _tr_flush_block
// (on the device every lane of a warp runs this with identical values; only S decides which lane touches the trees).
// Returns the size in bytes of the gzip stream.
template <class M, class S>
DS_HD uint32_t gzip_size(M &m, S &sink, uint32_t n) {
    if (n > DSZ_MAX_LEN) return DSZ_UNSUPPORTED;
    uint64_t bits = 0;
    uint32_t syms = 0;
    uint32_t strstart = 0, block_start = 0, match_length = MIN_MATCH - 1, prev_length = MIN_MATCH - 1, match_start = 0, prev_match = 0;
    bool match_available = false;
    uint32_t base = 0, filled = n < 2u * W_SIZE ? n : 2u * W_SIZE;     // window base and end of the data read into it (fill_window)
    while (strstart < n) {
        if (filled - strstart < uint32_t(MIN_LOOKAHEAD) && filled < n) {
            do {
                if (strstart - base >= uint32_t(W_SIZE) + MAX_DIST) base += W_SIZE;
                const uint32_t more = 2u * W_SIZE - (filled - base), left = n - filled;
                filled += more < left ? more : left;
            } while (filled - strstart < uint32_t(MIN_LOOKAHEAD) && filled < n);
        }
        const uint32_t lookahead = n - strstart;   // matches never reach the end of the filled window while input is left (262 > 258)
        uint32_t hash_head = 0;
        const bool has_head = lookahead >= uint32_t(MIN_MATCH) && m.head(strstart, match_length, hash_head) && hash_head != 0;
        prev_length = match_length; prev_match = match_start;
        match_length = MIN_MATCH - 1;
        if (has_head && prev_length < MAX_LAZY && strstart - hash_head <= MAX_DIST) {
            match_length = m.longest(strstart, prev_length, match_start);
            if (match_length == uint32_t(MIN_MATCH) && strstart - match_start > TOO_FAR) match_length = MIN_MATCH - 1;
        }
        if (prev_length >= uint32_t(MIN_MATCH) && match_length <= prev_length) {
            sink.dist(strstart - 1 - prev_match, prev_length - MIN_MATCH);
            strstart += prev_length - 1;          // the strings in between are inserted by construction (occ[] holds every position)
            match_available = false;
            match_length = MIN_MATCH - 1;
            if (++syms == SYMS_PER_BLOCK) { bits = sink.flush(bits, strstart - block_start, block_start >= base); block_start = strstart; syms = 0; }
        } else if (match_available) {
            sink.lit(m.at(strstart - 1));
            if (++syms == SYMS_PER_BLOCK) { bits = sink.flush(bits, strstart - block_start, block_start >= base); block_start = strstart; syms = 0; }
            strstart++;
        } else {
            match_available = true;
            strstart++;
        }
    }
    if (match_available) sink.lit(m.at(strstart - 1));
    bits = sink.flush(bits, strstart - block_start, block_start >= base);     // the last block (Z_FINISH), also for an empty input
    return 10u + uint32_t((bits + 7) >> 3) + 8u;
}

// the plain sink: one thread owns the trees
struct TreeSink {
    Trees &T;
    DS_HD void lit(uint8_t c) { tally_lit(T, c); }
    DS_HD void dist(uint32_t d, uint32_t lc) { tally_dist(T, d, lc); }
    DS_HD uint64_t flush(uint64_t bits, uint32_t stored_len, bool have_buf) { return flush_block(T, bits, stored_len, have_buf); }
};

}  // namespace dsz
