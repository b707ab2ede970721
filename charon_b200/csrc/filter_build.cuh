// charon_b200/csrc/filter_build.cuh — one-off kernels that lay the Interleaved Bloom Filter out in HBM.
//
// HBM layout ("column-compacted rows"): row r of the IBF (one row = the `bins` membership bits selected by one
// hash value, seqan3 layout bit index = row * technical_bins + bin) is stored in RB = 2^row_shift
// bits, RB = smallest power of two >= bins, at bit offset r * RB, LSB-first in 32-bit words (RB <= 32) or in
// RB/64 consecutive 64-bit words (RB >= 64).  A 2-bin index therefore costs 2 bits per row instead of the 64 of the
// plain seqan3 layout: 32x smaller, so bacterial-scale filters sit in the 126 MB L2 and a human-scale one is < 1 GB.
// Rows never straddle a 32-byte sector, so one hash function = one sector.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace cgk {

// plain rows chunk (host layout, bin_words u64 per row) -> compact.  rows [row0, row0 + nrows), row0 % 32 == 0.
__global__ void compact_rows_kernel(const uint64_t *__restrict__ plain, uint64_t row0, uint64_t nrows, uint32_t bin_words,
                                    uint32_t row_shift, void *filter) {
    const uint32_t rb = 1u << row_shift;
    if (rb <= 32) {
        const uint32_t per = 32u / rb;                       // rows per output word
        const uint64_t nwords = (nrows + per - 1) / per;
        for (uint64_t wi = blockIdx.x * uint64_t(blockDim.x) + threadIdx.x; wi < nwords; wi += uint64_t(gridDim.x) * blockDim.x) {
            uint32_t out = 0;
            const uint32_t msk = rb >= 32 ? 0xFFFFFFFFu : ((1u << rb) - 1u);
            for (uint32_t j = 0; j < per; ++j) {
                const uint64_t rr = wi * per + j;
                if (rr < nrows) out |= (uint32_t(plain[rr * bin_words]) & msk) << (j * rb);
            }
            reinterpret_cast<uint32_t *>(filter)[row0 / per + wi] = out;
        }
    } else {
        const uint32_t ow = rb / 64;                          // output words per row (>= bin_words)
        const uint64_t n = nrows * ow;
        for (uint64_t i = blockIdx.x * uint64_t(blockDim.x) + threadIdx.x; i < n; i += uint64_t(gridDim.x) * blockDim.x) {
            const uint64_t rr = i / ow;
            const uint32_t j = uint32_t(i % ow);
            reinterpret_cast<uint64_t *>(filter)[(row0 + rr) * ow + j] = j < bin_words ? plain[rr * bin_words + j] : 0ull;
        }
    }
}

// compact -> plain rows (debug dump / tests)
__global__ void expand_rows_kernel(const void *filter, uint64_t row0, uint64_t nrows, uint32_t bin_words, uint32_t row_shift,
                                   uint64_t *__restrict__ plain) {
    const uint32_t rb = 1u << row_shift;
    for (uint64_t rr = blockIdx.x * uint64_t(blockDim.x) + threadIdx.x; rr < nrows; rr += uint64_t(gridDim.x) * blockDim.x) {
        const uint64_t row = row0 + rr;
        if (rb <= 32) {
            const uint64_t bit = row << row_shift;
            const uint32_t w = reinterpret_cast<const uint32_t *>(filter)[bit >> 5];
            const uint32_t msk = rb >= 32 ? 0xFFFFFFFFu : ((1u << rb) - 1u);
            plain[rr] = (w >> (uint32_t(bit) & 31u)) & msk;
        } else {
            const uint32_t ow = rb / 64;
            for (uint32_t j = 0; j < bin_words; ++j) plain[rr * bin_words + j] = reinterpret_cast<const uint64_t *>(filter)[row * ow + j];
        }
    }
}

// ---- sdsl::sd_vector (Elias-Fano) decode on the device (index file payload) ----------------------
__global__ void popcount_words_kernel(const uint64_t *__restrict__ words, uint64_t n, uint64_t *__restrict__ out) {
    for (uint64_t i = blockIdx.x * uint64_t(blockDim.x) + threadIdx.x; i < n; i += uint64_t(gridDim.x) * blockDim.x)
        out[i] = uint64_t(__popcll(words[i]));
}

struct SdDecodeParams {
    const uint64_t *high; uint64_t high_words; uint64_t high_bits;
    const uint64_t *low; uint32_t wl;
    const uint64_t *prefix;          // exclusive prefix popcount of high words
    uint64_t size_bits;              // technical_bins * bin_size
    uint32_t technical_bins, bins, row_shift;
    void *filter;
    unsigned long long *errors;      // [0] position out of range, [1] bin >= bins
    uint64_t row_begin, row_end;     // partitioned filter: only rows in [row_begin, row_end) are kept, stored relative to row_begin
};

__global__ void sd_decode_kernel(const SdDecodeParams P) {
    const uint32_t rb = 1u << P.row_shift;
    for (uint64_t wi = blockIdx.x * uint64_t(blockDim.x) + threadIdx.x; wi < P.high_words; wi += uint64_t(gridDim.x) * blockDim.x) {
        uint64_t w = P.high[wi];
        uint64_t j = P.prefix[wi];
        while (w) {
            const uint64_t q = wi * 64 + uint64_t(__ffsll((long long)w) - 1);
            w &= w - 1;
            if (q >= P.high_bits) break;
            uint64_t lo = 0;
            if (P.wl) {
                const uint64_t p = j * P.wl;
                lo = P.low[p >> 6] >> (p & 63);
                if ((p & 63) + P.wl > 64) lo |= P.low[(p >> 6) + 1] << (64 - (p & 63));
                if (P.wl < 64) lo &= (1ull << P.wl) - 1ull;
            }
            const uint64_t pos = ((q - j) << P.wl) | lo;
            ++j;
            if (pos >= P.size_bits) { atomicAdd(&P.errors[0], 1ull); continue; }
            uint64_t row = pos / P.technical_bins;
            const uint32_t bin = uint32_t(pos % P.technical_bins);
            if (bin >= P.bins) { atomicAdd(&P.errors[1], 1ull); continue; }
            if (row < P.row_begin || row >= P.row_end) continue;
            row -= P.row_begin;
            if (rb <= 32) {
                const uint64_t bit = (row << P.row_shift) + bin;
                atomicOr(reinterpret_cast<uint32_t *>(P.filter) + (bit >> 5), 1u << (uint32_t(bit) & 31u));
            } else {
                const uint64_t word = row * (rb / 64) + (bin >> 6);
                atomicOr(reinterpret_cast<unsigned long long *>(P.filter) + word, 1ull << (bin & 63u));
            }
        }
    }
}

// ---- compacted filter -> sdsl::sd_vector parts (what Index::serialize stores, include/index.hpp:50,130) ----
// compact bit c = row << row_shift | bin  <->  seqan3 position row * technical_bins + bin; the mapping is monotone, so the
// j-th set bit of the compact filter is the j-th one of the sd_vector.  prefix = exclusive popcount scan of the words.
__global__ void sd_encode_kernel(const uint64_t *__restrict__ filter, uint64_t nwords, const uint64_t *__restrict__ prefix, uint32_t row_shift,
                                 uint32_t technical_bins, uint32_t wl, unsigned long long *__restrict__ low, unsigned long long *__restrict__ high) {
    const uint64_t lowmask = wl >= 64 ? ~0ull : ((1ull << wl) - 1ull);
    for (uint64_t wi = blockIdx.x * uint64_t(blockDim.x) + threadIdx.x; wi < nwords; wi += uint64_t(gridDim.x) * blockDim.x) {
        uint64_t w = filter[wi];
        uint64_t j = prefix[wi];
        while (w) {
            const uint64_t c = wi * 64 + uint64_t(__ffsll((long long)w) - 1);
            w &= w - 1;
            const uint64_t row = c >> row_shift, bin = c & ((1ull << row_shift) - 1ull);
            const uint64_t pos = row * technical_bins + bin;
            if (wl) {
                const uint64_t p = j * wl, v = pos & lowmask;
                atomicOr(&low[p >> 6], v << (p & 63));
                if ((p & 63) + wl > 64) atomicOr(&low[(p >> 6) + 1], v >> (64 - (p & 63)));
            }
            const uint64_t hb = (pos >> wl) + j;
            atomicOr(&high[hb >> 6], 1ull << (hb & 63));
            ++j;
        }
    }
}

// seqan3 interleaved_bloom_filter::emplace(value, bin) for an array of values (reference src/index_main.cpp:252-254)
__global__ void ibf_insert_kernel(const uint64_t *__restrict__ values, uint64_t n, uint32_t bin, uint64_t bin_size, uint32_t hash_shift,
                                  uint32_t hash_funs, uint32_t row_shift, void *filter) {
    const uint32_t rb = 1u << row_shift;
    for (uint64_t i = blockIdx.x * uint64_t(blockDim.x) + threadIdx.x; i < n; i += uint64_t(gridDim.x) * blockDim.x) {
        const uint64_t v = values[i];
        for (uint32_t f = 0; f < hash_funs; ++f) {
            const uint64_t row = ibf_row(v, c_ibf_seeds[f], hash_shift, bin_size);
            if (rb <= 32) {
                const uint64_t bit = (row << row_shift) + bin;
                atomicOr(reinterpret_cast<uint32_t *>(filter) + (bit >> 5), 1u << (uint32_t(bit) & 31u));
            } else {
                atomicOr(reinterpret_cast<unsigned long long *>(filter) + row * (rb / 64) + (bin >> 6), 1ull << (bin & 63u));
            }
        }
    }
}

// membership masks for raw hash values (test hook cg_debug_query): expands to the seqan3 binning_bitvector layout
template <int MODE>
__global__ void query_kernel(const CountParams P, const uint64_t *__restrict__ values, uint64_t n, uint32_t bin_words,
                             uint64_t *__restrict__ masks) {
    for (uint64_t i = blockIdx.x * uint64_t(blockDim.x) + threadIdx.x; i < n; i += uint64_t(gridDim.x) * blockDim.x) {
        const auto m = bulk_contains<MODE>(P, values[i], l2_policy_evict_last());
        for (uint32_t j = 0; j < bin_words; ++j) masks[i * bin_words + j] = m.w[j < ModeTraits<MODE>::NW ? j : 0];
    }
}

}  // namespace cgk
