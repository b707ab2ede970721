// charon_b200/csrc/gzip_size_kernel.cuh — the compression-ratio scalar on the device ("compression-ratio computation
// (the real e2e bottleneck)"; reference src/utils.cpp:114-124 + src/dehost_main.cpp:359-361: ratio = size of gzip::compress(
// uppercase dna5 string) / length).  deflate_size.cuh holds the algorithm; here: one WARP per read.
//   1. text: the read's segments, as the uppercase dna5 letters A C G T N, one byte each, in the warp's scratch (global memory);
//   2. counting sort of the positions by their three-letter code (125 buckets, stable): occ[] = positions bucket by bucket in
//      ascending order, idx[p] = where p sits in occ[]; a zlib hash chain from p = occ[idx[p] - 1], occ[idx[p] - 2], ...;
//   3. deflate_slow, every lane running the same state machine; longest_match examines 32 chain candidates per round, one
//      per lane, and reproduces zlib's sequential rules (strictly longer replaces, stop at nice_length) with a prefix maximum;
//   4. lane 0 owns the Huffman trees (shared memory) and closes the blocks.
#pragma once
#include <cstdint>
#include "deflate_size.cuh"

namespace cgk {

struct GzParams {
    const uint8_t *bases;                 // staged batch, device format: 0 ASCII, 2 two bits per base
    uint32_t format;
    const uint64_t *seg_off, *seg_len;    // seg_len may be NULL
    const uint32_t *read_seg;             // may be NULL
    const unsigned long long *exc_pos; const uint32_t *exc_seg_off; unsigned long long n_exc;
    uint32_t n_reads;
    uint32_t *work_counter;
    uint8_t *scratch; uint64_t scratch_stride; uint32_t cap;   // per warp: text[cap + 64] | idx u32[cap] | occ u16 / u32[cap]; scratch_stride = 9 cap + 64
    uint32_t prefetch;                    // fetch the predicted next position's chain candidates one step early
    uint32_t *size_out;                   // gzip size per read (DSZ_UNSUPPORTED: longer than DSZ_MAX_LEN)
    float *ratio_out;                     // float(double(size) / double(length)); may be NULL
};

constexpr int GZ_WARPS = 8;

// unaligned 8 letters at text + x (text is 4-byte aligned; 12 readable bytes follow every position a compare can reach)
__device__ __forceinline__ uint64_t load8(const uint8_t *text, uint32_t x) {
    const uint32_t *w = reinterpret_cast<const uint32_t *>(text + (x & ~3u));
    const uint32_t w0 = w[0], w1 = w[1], w2 = w[2], sh = (x & 3u) * 8u;
    return uint64_t(__funnelshift_r(w0, w1, sh)) | (uint64_t(__funnelshift_r(w1, w2, sh)) << 32);
}

template <class OccT>     // uint16_t while every position of the batch's longest read fits, else uint32_t
struct WarpMatcher {
    const uint8_t *text;             // shared memory when the read fits there, else the warp's global scratch
    const OccT *occ; const uint32_t *idx;   // idx[p] = index of p in occ[] (24 bits) | min(rank of p inside its bucket, 255) << 24
    uint32_t n, np, lane;
    // register window over idx[wb .. wb + 64): lane i holds idx[wb + i] and idx[wb + 32 + i]; the upper half is fetched one
    // half ahead of its use
    uint32_t wb, iwA, iwB;
    // chain candidates (lane i: entries i, i + 32, i + 64, i + 96 of the chain) of the position being searched, and of the
    // position predicted to be searched next (fetched one step early: their latency hides behind this step)
    uint32_t cq[4], nc;
    uint32_t pq[4], prank, ppos;
    bool prefetch;
    __device__ __forceinline__ uint32_t ld_idx(uint32_t x) const { return x < np ? idx[x] : 0u; }
    __device__ __forceinline__ void advance(uint32_t p) {      // p never decreases
        if (p - wb >= 64u) { wb = p; iwA = ld_idx(wb + lane); iwB = ld_idx(wb + 32u + lane); }
        else if (p - wb >= 32u) { iwA = iwB; wb += 32u; iwB = ld_idx(wb + 32u + lane); }
    }
    __device__ __forceinline__ uint32_t entry(uint32_t x) const {   // wb <= x < wb + 64
        const uint32_t d = x - wb;
        const uint32_t a = __shfl_sync(0xFFFFFFFFu, iwA, int(d & 31u)), b = __shfl_sync(0xFFFFFFFFu, iwB, int(d & 31u));
        return d < 32u ? a : b;
    }
    __device__ __forceinline__ void fetch(uint32_t e, uint32_t (&q)[4], uint32_t &rank) const {   // up to MAX_CHAIN candidates
        const uint32_t k = e & 0xFFFFFFu;
        rank = e >> 24;                                   // saturated at 255: only min(rank, MAX_CHAIN = 128) matters
        const uint32_t m = rank < dsz::MAX_CHAIN ? rank : dsz::MAX_CHAIN;
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            const uint32_t j = uint32_t(r) * 32u + lane;
            q[r] = 0u;
            if (uint32_t(r) * 32u < m && j < m) q[r] = uint32_t(occ[k - 1u - j]);
        }
    }
    __device__ __forceinline__ uint8_t at(uint32_t q) const { return text[q]; }
    __device__ __forceinline__ bool head(uint32_t p, uint32_t prev_length, uint32_t &q) {
        advance(p);
        uint32_t rank;
        if (ppos == p) {
#pragma unroll
            for (int r = 0; r < 4; ++r) cq[r] = pq[r];
            rank = prank;
        } else {
            fetch(entry(p), cq, rank);
        }
        const uint32_t chain = prev_length >= dsz::GOOD_LENGTH ? dsz::MAX_CHAIN >> 2 : dsz::MAX_CHAIN;
        nc = rank < chain ? rank : chain;
        // the position searched next: with a match of prev_length pending at p - 1, p + prev_length - 1 unless this position
        // finds a longer one (lazy evaluation); without one, always p + 1
        const uint32_t t = prev_length >= uint32_t(dsz::MIN_MATCH) ? p + prev_length - 1u : p + 1u;
        ppos = 0xFFFFFFFFu;
        if (prefetch && t < np && t - wb < 64u) { fetch(entry(t), pq, prank); ppos = t; }
        if (rank == 0) return false;
        q = __shfl_sync(0xFFFFFFFFu, cq[0], 0);
        return true;
    }
    __device__ __forceinline__ uint32_t longest(uint32_t p, uint32_t prev_length, uint32_t &match_start) {
        const uint32_t lookahead = n - p;
        const uint32_t nice = dsz::NICE_LENGTH < lookahead ? dsz::NICE_LENGTH : lookahead;
        const uint32_t cap = uint32_t(dsz::MAX_MATCH) < lookahead ? uint32_t(dsz::MAX_MATCH) : lookahead;
        const uint32_t limit = p > dsz::MAX_DIST ? p - dsz::MAX_DIST : 0u;
        uint32_t best = prev_length;
        if (best >= cap) return best <= lookahead ? best : lookahead;      // no candidate can be strictly longer
        // zlib's quick reject, made exact for a whole round: a candidate can only replace the best so far if it is longer than
        // prev_length, i.e. if it also agrees at offset prev_length.  One letter per candidate decides that for three in four.
        const uint32_t pb = text[p + best];
        const uint64_t b0 = load8(text, p);
        bool valid[4], pass[4];
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            valid[r] = false; pass[r] = false;
            if (uint32_t(r) * 32u < nc) {                        // warp-uniform: most chains need one or two rounds
                const uint32_t j = uint32_t(r) * 32u + lane;
                valid[r] = j < nc && (j == 0 || cq[r] > limit);  // the chain ends at the first candidate that is too far (or is position 0)
                pass[r] = valid[r] && uint32_t(text[cq[r] + best]) == pb;
            }
        }
        uint64_t a0[4];
#pragma unroll
        for (int r = 0; r < 4; ++r) { a0[r] = 0ull; if (uint32_t(r) * 32u < nc && pass[r]) a0[r] = load8(text, cq[r]); }
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            if (uint32_t(r) * 32u >= nc) break;
            const uint32_t j = uint32_t(r) * 32u + lane, q = cq[r];
            uint32_t len = 0;
            if (pass[r]) {
                uint64_t x = a0[r] ^ b0;
                while (x == 0) {
                    len += 8;
                    if (len >= cap) break;
                    x = load8(text, q + len) ^ load8(text, p + len);
                }
                if (x) len += uint32_t(__ffsll((long long)x) - 1) >> 3;
                len = len < cap ? len : cap;
            }
            // zlib walks the candidates in order: one replaces the best so far only if strictly longer, and the walk stops at
            // the first such candidate that reaches nice_length
            const uint32_t any = __ballot_sync(0xFFFFFFFFu, len > best);
            if (any) {
                uint32_t incl = len;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) { const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, incl, o); if (lane >= uint32_t(o)) incl = incl > t ? incl : t; }
                uint32_t excl = __shfl_up_sync(0xFFFFFFFFu, incl, 1);
                if (lane == 0) excl = 0;
                const uint32_t before = excl > best ? excl : best;
                const bool improving = valid[r] && len > before;
                const uint32_t brk = __ballot_sync(0xFFFFFFFFu, improving && len >= nice);
                const uint32_t upto = brk ? uint32_t(__ffs(int(brk)) - 1) : 31u;
                const uint32_t cand = (valid[r] && lane <= upto) ? len : 0u;
                const uint32_t mx = __reduce_max_sync(0xFFFFFFFFu, cand);
                if (mx > best) {
                    const uint32_t who = uint32_t(__ffs(int(__ballot_sync(0xFFFFFFFFu, cand == mx))) - 1);
                    match_start = __shfl_sync(0xFFFFFFFFu, q, int(who));
                    best = mx;
                }
                if (brk) break;
            }
            if (__ballot_sync(0xFFFFFFFFu, !valid[r] && j < nc)) break;
        }
        return best <= lookahead ? best : lookahead;
    }
};

// lane 0 owns the trees; every lane learns the stream length after a block
struct WarpSink {
    dsz::Trees &T; uint32_t lane;
    __device__ __forceinline__ void lit(uint8_t c) { if (lane == 0) dsz::tally_lit(T, c); }
    __device__ __forceinline__ void dist(uint32_t d, uint32_t lc) { if (lane == 0) dsz::tally_dist(T, d, lc); }
    __device__ __forceinline__ uint64_t flush(uint64_t bits, uint32_t stored_len, bool have_buf) {
        if (lane == 0) bits = dsz::flush_block(T, bits, stored_len, have_buf);
        __syncwarp();
        return __shfl_sync(0xFFFFFFFFu, bits, 0);
    }
};

constexpr size_t GZ_TREES_BYTES = (sizeof(dsz::Trees) + 15) / 16 * 16;
constexpr uint32_t GZ_SMEM_TEXT = 8192 + 64;     // reads of up to 8192 letters keep their text in shared memory
constexpr size_t GZ_WARP_SMEM = GZ_TREES_BYTES + (dsz::N_LETTER_CODES + 3) * 4 + GZ_SMEM_TEXT;
constexpr size_t GZ_SMEM = GZ_WARPS * GZ_WARP_SMEM;

template <class OccT>
__global__ void __launch_bounds__(GZ_WARPS * 32) gzip_size_kernel(const GzParams P) {
    extern __shared__ __align__(16) unsigned char gz_smem[];
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    dsz::Trees *const s_trees = reinterpret_cast<dsz::Trees *>(gz_smem + size_t(warp) * GZ_WARP_SMEM);
    uint32_t *const s_start = reinterpret_cast<uint32_t *>(gz_smem + size_t(warp) * GZ_WARP_SMEM + GZ_TREES_BYTES);
    uint8_t *const text_s = reinterpret_cast<uint8_t *>(s_start + dsz::N_LETTER_CODES + 3);
    const uint64_t slot = uint64_t(blockIdx.x) * GZ_WARPS + warp;
    const uint32_t cap = P.cap;
    uint8_t *const text_g = P.scratch + slot * P.scratch_stride;
    uint32_t *const idx = reinterpret_cast<uint32_t *>(text_g + cap + 64);
    OccT *const occ = reinterpret_cast<OccT *>(idx + cap);
    uint32_t *const start = s_start;
    dsz::Trees &T = *s_trees;
    for (;;) {
        uint32_t r = 0;
        if (lane == 0) r = atomicAdd(P.work_counter, 1u);
        r = __shfl_sync(0xFFFFFFFFu, r, 0);
        if (r >= P.n_reads) break;
        const uint32_t seg0 = P.read_seg ? P.read_seg[r] : r, seg1 = P.read_seg ? P.read_seg[r + 1] : r + 1;
        uint64_t total = 0;
        for (uint32_t s = seg0; s < seg1; ++s) total += P.seg_len ? P.seg_len[s] : P.seg_off[s + 1] - P.seg_off[s];
        if (total > dsz::DSZ_MAX_LEN || total > cap) {
            if (lane == 0) { P.size_out[r] = dsz::DSZ_UNSUPPORTED; if (P.ratio_out) P.ratio_out[r] = __int_as_float(0x7FC00000); }
            continue;
        }
        const uint32_t n = uint32_t(total);
        uint8_t *const text = n + 64u <= GZ_SMEM_TEXT ? text_s : text_g;
        // 1. the uppercase dna5 string
        uint32_t at = 0;
        for (uint32_t s = seg0; s < seg1; ++s) {
            const uint64_t off = P.seg_off[s];
            const uint32_t len = uint32_t(P.seg_len ? P.seg_len[s] : P.seg_off[s + 1] - off);
            if (P.format == 2) {
                for (uint32_t i = lane; i < len; i += 32) {
                    const uint64_t g = off + i;
                    const uint32_t two = (P.bases[g >> 2] >> ((g & 3u) * 2u)) & 3u;
                    text[at + i] = uint8_t(0x54474341u >> (two * 8u));      // "ACGT"
                }
                if (P.n_exc) {
                    __syncwarp();
                    for (uint32_t e = P.exc_seg_off[s] + lane; e < P.exc_seg_off[s + 1]; e += 32) text[at + uint32_t(P.exc_pos[e] - off)] = 'N';
                }
            } else {
                for (uint32_t i = lane; i < len; i += 32) {
                    const uint32_t rank = base_code(P.bases[off + i]) & 15u;      // dna5 rank A0 C1 G2 N3 T4 (IUPAC -> N, lower case folded, u -> T)
                    text[at + i] = uint8_t(rank == 0 ? 'A' : (rank == 1 ? 'C' : (rank == 2 ? 'G' : (rank == 4 ? 'T' : 'N'))));
                }
            }
            at += len;
        }
        if (lane < 8) reinterpret_cast<uint32_t *>(text + ((n + 3u) & ~3u))[lane] = 0u;   // 32 zero bytes behind the text (the scratch still holds the previous read)
        if (lane < 4 && n + lane < ((n + 3u) & ~3u)) text[n + lane] = 0;
        __syncwarp();
        // 2. stable counting sort of the positions 0 .. n-3 by three-letter code
        for (uint32_t b = lane; b < uint32_t(dsz::N_LETTER_CODES + 1); b += 32) start[b] = 0;
        __syncwarp();
        const uint32_t np = n >= 3 ? n - 2 : 0;
        auto code3 = [&](uint32_t p) -> uint32_t { return dsz::letter_code(text[p]) * 25u + dsz::letter_code(text[p + 1]) * 5u + dsz::letter_code(text[p + 2]); };
        for (uint32_t p = lane; p < np; p += 32) atomicAdd(&start[code3(p) + 1], 1u);
        __syncwarp();
        {   // exclusive prefix over 125 counts -> start[b]; start[b + 1] - start[b] = bucket size
            uint32_t c[4], sum = 0;
#pragma unroll
            for (int t = 0; t < 4; ++t) { const uint32_t b = lane * 4 + t; c[t] = b < uint32_t(dsz::N_LETTER_CODES) ? start[b + 1] : 0u; sum += c[t]; }
            uint32_t incl = sum;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, incl, o); if (lane >= uint32_t(o)) incl += t; }
            uint32_t run = incl - sum;
            __syncwarp();
#pragma unroll
            for (int t = 0; t < 4; ++t) { const uint32_t b = lane * 4 + t; if (b < uint32_t(dsz::N_LETTER_CODES)) start[b] = run; run += c[t]; }
            if (lane == 31) start[dsz::N_LETTER_CODES] = incl;
        }
        __syncwarp();
        // fill: chunks of 32 consecutive positions; the lanes of one bucket take consecutive slots in lane (= position) order,
        // so occ[] stays ascending inside every bucket.  The running fill counts cur[b] borrow the Trees area, which is not
        // in use before the state machine starts.
        uint32_t *const cur = reinterpret_cast<uint32_t *>(&T);
        for (uint32_t b = lane; b < uint32_t(dsz::N_LETTER_CODES); b += 32) cur[b] = start[b];
        __syncwarp();
        for (uint32_t p0 = 0; p0 < np; p0 += 32) {
            const uint32_t p = p0 + lane;
            const bool act = p < np;
            const uint32_t b = act ? code3(p) : 0xFFFFu;
            const uint32_t same = __match_any_sync(0xFFFFFFFFu, b);
            const uint32_t rank = __popc(same & ((1u << lane) - 1u));
            const uint32_t leader = uint32_t(__ffs(int(same)) - 1);
            uint32_t basep = 0;
            if (act && lane == leader) { basep = cur[b]; cur[b] = basep + uint32_t(__popc(same)); }
            basep = __shfl_sync(0xFFFFFFFFu, basep, int(leader));
            if (act) { const uint32_t k = basep + rank, rk = k - start[b]; idx[p] = k | ((rk < 255u ? rk : 255u) << 24); occ[k] = OccT(p); }
            __syncwarp();
        }
        __syncwarp();
        // 3 + 4. deflate_slow
        if (lane == 0) dsz::init_block(T);
        __syncwarp();
        WarpSink sink{T, lane};
        WarpMatcher<OccT> M{text, occ, idx, n, np, lane, 0u - 64u, 0u, 0u, {0u, 0u, 0u, 0u}, 0u, {0u, 0u, 0u, 0u}, 0u, 0xFFFFFFFFu, P.prefetch != 0};   // wb: nothing windowed yet
        const uint32_t size = dsz::gzip_size(M, sink, n);
        if (lane == 0) {
            P.size_out[r] = size;
            if (P.ratio_out) P.ratio_out[r] = float(double(size) / double(n));
        }
        __syncwarp();
    }
}

}  // namespace cgk
