// charon_b200/csrc/host_pack.hpp — host side of the packed-bases transport (internal to libcharon_gpu.so).
//
// A batch handed to the C-ABI is ASCII (what the reference's FASTQ parser produces, src/dehost_main.cpp:335-339).  One base
// per byte makes the PCIe link the end-to-end bound (8 kb reads: 1 GB per 131 072 reads at ~52 GB/s = 20 ms against 6 ms of
// kernels), so before the copy the library packs the bases to 4 bits each — the dna5 rank the kernels would compute anyway
// (A0 C1 G2 N3 T4, U -> T, anything else -> N; seqan3 dna5 char_to_rank) — on a small thread pool, chunk by chunk, each
// chunk's H2D copy enqueued as soon as it is packed.  Base i of the concatenated batch lives in nibble i: byte i >> 1, low
// nibble for even i.  The kernels expand a nibble back to the same code byte the ASCII path produces (count_kernel_v2.cuh
// decode_packed16), so every downstream integer is unchanged.
#pragma once
#include <cstddef>
#include <cstdint>
#include <functional>
#include <vector>

namespace cgh {

// rank nibble of one ASCII character (scalar definition; the SIMD path is tested against it)
inline uint8_t rank_of(uint8_t c) {
    const uint8_t u = uint8_t(c & 0xDFu);
    return u == 'A' ? 0 : u == 'C' ? 1 : u == 'G' ? 2 : (u == 'T' || u == 'U') ? 4 : 3;
}

// dst nibble i = rank_of(src[i]) for i in [begin, end); begin must be even; bytes of dst outside the range are untouched
// except the last one when end is odd (its high nibble becomes 0).  Single-threaded.
void pack_range(const uint8_t *src, uint8_t *dst, uint64_t begin, uint64_t end);

// CG_BASES_2BIT writer for one run of bases (include/charon_gpu.h cg_pack_bases_2bit): dst gets (n + 3) / 4 bytes, base i =
// bits 2 * (i & 3) of dst[i >> 2], A0 C1 G2 T3 (U -> T); every other character is stored as 0 and its position pos0 + i is
// appended to exc (at most cap entries are stored).  Returns how many such characters the run holds.  Single-threaded.
uint64_t pack2_range(const uint8_t *src, uint64_t n, uint8_t *dst, uint64_t pos0, uint64_t *exc, uint64_t cap);

// Thread pool shared by all handles of the process.  threads() resolves CHARON_HOST_THREADS, else min(16, hardware).
int threads();
void set_threads(int n);   // takes effect before the first job, or re-creates the pool
// Packs [0, n) in blocks on the pool (the caller packs too); after every `chunk` bases (a multiple of 64) that are complete,
// and in order, on_chunk(begin_base, end_base) is called on the CALLER's thread (it enqueues that chunk's H2D copy).
void pack_chunked(const uint8_t *src, uint8_t *dst, uint64_t n, uint64_t chunk, const std::function<void(uint64_t, uint64_t)> &on_chunk);
// The same to two bits per base (pack2_range per block); exc receives the ascending positions of everything that is not A/C/G/T/U.
void pack2_chunked(const uint8_t *src, uint8_t *dst, uint64_t n, uint64_t chunk, const std::function<void(uint64_t, uint64_t)> &on_chunk,
                   std::vector<uint64_t> &exc);

}  // namespace cgh
