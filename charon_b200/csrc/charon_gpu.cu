// charon_b200/csrc/charon_gpu.cu — libcharon_gpu.so: the C-ABI of include/charon_gpu.h over hand-written
// sm_100a kernels.  Host part: filter upload / re-layout, batch staging, launches, result copies.
// There is deliberately no CPU path: without a Blackwell device cg_create fails (CG_ERR_NO_DEVICE).
#include "../../include/charon_gpu.h"

#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>
#include <cub/device/device_select.cuh>
#include <cuda_runtime.h>

#include <algorithm>
#include <utility>
#include <cmath>
#include <chrono>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include "count_kernel.cuh"
#include "count_kernel_v2.cuh"
#include "count_kernel_v3.cuh"
#include "host_pack.hpp"
#include "eval_kernel.cuh"
#include "filter_build.cuh"
#include "gather_bench.cuh"
#include "gzip_size_kernel.cuh"

namespace {

thread_local std::string g_err;   // for cg_create failures (no handle yet)

struct DevBuf {
    void *p = nullptr;
    size_t cap = 0;
    int reserve(size_t bytes) {
        if (bytes <= cap) return 0;
        if (p) cudaFree(p);
        p = nullptr; cap = 0;
        size_t want = bytes + bytes / 8 + 256;
        cudaError_t e = cudaMalloc(&p, want);
        if (e != cudaSuccess) { p = nullptr; return int(e); }
        cap = want;
        return 0;
    }
    void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
    template <class T> T *as() const { return reinterpret_cast<T *>(p); }
};

}  // namespace

struct cg_handle {
    int device = 0;
    cudaStream_t stream = nullptr;
    bool own_stream = true;
    cudaEvent_t ev[8] = {};
    std::string err;
    int sm_count = 0;
    // index
    uint32_t K = 0, W = 0, bins = 0, ncat = 0, bin_words = 0, hash_funs = 0, row_shift = 0, hash_shift = 0;
    uint64_t bin_size = 0, technical_bins = 0;
    int mode = 0;   // kernel MODE (row width class)
    bool single_pass = false;
    void *d_filter = nullptr;
    size_t filter_bytes = 0;
    // row-range partition (0 parts = replicated): this GPU stores rows [row_begin, row_begin + local_rows)
    uint32_t n_parts = 0, my_part = 0, rows_per_part = 0;
    uint64_t row_begin = 0, local_rows = 0;
    const void *part_base[CG_MAX_PARTS] = {};
    bool part_ipc[CG_MAX_PARTS] = {};
    uint8_t *d_bin_to_cat = nullptr;
    uint8_t h_bin_to_cat[256];
    // model
    bool model_set = false, model_has_kde = false;
    int call_mode = 0;
    cgk::ModelDev *d_models = nullptr;
    float *d_pool = nullptr;
    cg_thresholds thr{};
    // batch buffers
    DevBuf bases, seg_off, read_seg, num_hashes, counts, unique, prob, call, conf, mq, cr, len, counter, dbg, dbgn;
    DevBuf memo_keys, memo_vals, memo_list, memo_count, memo_slot;   // memoised model evaluation (eval_kernel.cuh), shared by the slots
    DevBuf gz_size, gz_scratch;   // compression-ratio kernel: gzip size per read (per slot), per-warp scratch (shared by the slots: their kernels never overlap)
    bool staged_device_cr = false;   // the batch came without compression ratios: gzip_size_kernel fills `cr` before the model kernels
    DevBuf seg_len, exc_pos, exc_seg_off, unpack_src;   // optional segment lengths; two-bit transport: exception list + per-segment CSR; packed input of the generic kernel
    uint32_t staged_reads = 0, staged_segs = 0;
    uint64_t staged_bases = 0;
    bool staged_has_read_seg = false, staged_has_scalars = false;
    // packed transport (host_pack.hpp): the staged bases are 4-bit ranks; pinned staging buffer of the active slot and the
    // event that says its last chunk has been copied out
    int staged_format = 0;        // enum cg_bases_format of the bases as they sit in `bases` (what the kernels decode)
    bool staged_has_seg_len = false;
    uint64_t staged_n_exc = 0;
    bool staged_packed = false;   // staged_format == CG_BASES_RANK4 produced by the library's own packer
    struct PackStage { void *p = nullptr; size_t cap = 0; cudaEvent_t done = nullptr; bool armed = false; } pack;
    int async_gather_mode = -1;   // -1: by filter size, 0 / 1: forced (env CHARON_ASYNC_GATHER, test hook)
    int host_pack_mode = -1;      // -1: automatic (thread pool of at least 8), 0: never, 1: whenever the (19, 41) kernels apply
    // launch config
    int grid_per_sm = 1;
    int warp_grid_per_sm = 1, warp_grid_per_sm_ag = 1;   // resident CTAs per SM of the warp kernel (register / cp.async gather carrier)
    uint64_t staged_max_read_len = 0;
    bool force_cta = false;       // test hook: CTA-per-read kernel for every read
    bool force_generic = false;   // test hook: run the any-(k,w) kernel even when the specialised one applies
    // second pipeline slot (cg_submit_batch / cg_wait_batch): the fields above always describe the ACTIVE slot; the other
    // slot's stream, events and batch buffers are parked here and swapped in when that slot is addressed.
    struct Parked {
        cudaStream_t stream = nullptr;
        cudaEvent_t ev[8] = {};
        DevBuf bases, seg_off, read_seg, num_hashes, counts, unique, prob, call, conf, mq, cr, len, counter, seg_len, exc_pos, exc_seg_off, unpack_src, gz_size;
        bool staged_device_cr = false;
        int staged_format = 0; bool staged_has_seg_len = false; uint64_t staged_n_exc = 0;
        uint32_t staged_reads = 0, staged_segs = 0;
        uint64_t staged_bases = 0, staged_max_read_len = 0;
        bool staged_has_read_seg = false, staged_has_scalars = false, staged_packed = false;
        PackStage pack;
        bool created = false, pending = false;
        cg_counts_out co{}; cg_calls_out ca{}; bool has_co = false, has_ca = false;
    } parked;
    int active_slot = 0;
    bool active_pending = false;
    cg_counts_out act_co{}; cg_calls_out act_ca{}; bool act_has_co = false, act_has_ca = false;
    cudaEvent_t join_ev = nullptr;
    void swap_slots() {
        std::swap(stream, parked.stream);
        for (int i = 0; i < 8; ++i) std::swap(ev[i], parked.ev[i]);
        std::swap(bases, parked.bases); std::swap(seg_off, parked.seg_off); std::swap(read_seg, parked.read_seg);
        std::swap(num_hashes, parked.num_hashes); std::swap(counts, parked.counts); std::swap(unique, parked.unique);
        std::swap(prob, parked.prob); std::swap(call, parked.call); std::swap(conf, parked.conf);
        std::swap(mq, parked.mq); std::swap(cr, parked.cr); std::swap(len, parked.len); std::swap(counter, parked.counter);
        std::swap(staged_reads, parked.staged_reads); std::swap(staged_segs, parked.staged_segs);
        std::swap(staged_bases, parked.staged_bases); std::swap(staged_max_read_len, parked.staged_max_read_len);
        std::swap(staged_has_read_seg, parked.staged_has_read_seg); std::swap(staged_has_scalars, parked.staged_has_scalars);
        std::swap(staged_packed, parked.staged_packed); std::swap(pack, parked.pack);
        std::swap(seg_len, parked.seg_len); std::swap(exc_pos, parked.exc_pos); std::swap(exc_seg_off, parked.exc_seg_off); std::swap(unpack_src, parked.unpack_src);
        std::swap(gz_size, parked.gz_size); std::swap(staged_device_cr, parked.staged_device_cr);
        std::swap(staged_format, parked.staged_format); std::swap(staged_has_seg_len, parked.staged_has_seg_len); std::swap(staged_n_exc, parked.staged_n_exc);
        std::swap(active_pending, parked.pending);
        std::swap(act_co, parked.co); std::swap(act_ca, parked.ca); std::swap(act_has_co, parked.has_co); std::swap(act_has_ca, parked.has_ca);
        active_slot ^= 1;
    }
    // stats
    cg_info info{};
    // cg_timer_*: one outer event pair + three events per run recorded while timing
    bool timing = false;
    cudaEvent_t t0 = nullptr, t1 = nullptr;
    std::vector<cudaEvent_t> tev;
    size_t tev_used = 0;
    cudaEvent_t next_tev() {
        if (tev_used == tev.size()) { cudaEvent_t e = nullptr; cudaEventCreate(&e); tev.push_back(e); }
        return tev[tev_used++];
    }
};

namespace {

int fail(cg_handle *h, int code, const char *fmt, ...) {
    char buf[512];
    va_list ap; va_start(ap, fmt); vsnprintf(buf, sizeof buf, fmt, ap); va_end(ap);
    if (h) h->err = buf; else g_err = buf;
    return code;
}
#define CU(h, call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return fail(h, CG_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); } while (0)
#define RSV(h, buf, bytes) do { int e_ = (buf).reserve(bytes); if (e_) return fail(h, CG_ERR_CUDA, "cudaMalloc(%zu) failed: %s", size_t(bytes), cudaGetErrorString(cudaError_t(e_))); } while (0)

uint32_t ipow5(uint32_t e) { uint32_t r = 1; while (e--) r *= 5u; return r; }

// B200 has 126 MB of L2: a filter below this stays resident next to the streamed bases and plain loads are cheaper
constexpr size_t ASYNC_GATHER_MIN_FILTER = size_t(96) << 20;
cgk::CountParams make_params(cg_handle *h) {
    cgk::CountParams P{};
    P.bases = h->bases.as<uint8_t>();
    P.seg_off = h->seg_off.as<uint64_t>();
    P.read_seg = h->staged_has_read_seg ? h->read_seg.as<uint32_t>() : nullptr;
    P.n_reads = h->staged_reads;
    P.K = h->K; P.Wv = h->W - h->K + 1;
    P.kl = std::min<uint32_t>(h->K, 13); P.kh = h->K - P.kl;
    P.p_lo_top = ipow5(P.kl - 1); P.p_hi_top = P.kh ? ipow5(P.kh - 1) : 0; P.p_kl = ipow5(P.kl);
    P.seed = 0x8F3F73B5CF1C9ADEULL;   // seqan3::seed default
    static_assert((0x8F3F73B5CF1C9ADEULL >> 61) == 4,
                  "umin64h/neq64h (count_kernel_v2.cuh) read the seeded values as negative finite doubles: top bits 100b");
    P.filter = h->d_filter; P.bin_size = h->bin_size; P.hash_shift = h->hash_shift; P.hash_funs = h->hash_funs;
    P.row_shift = h->row_shift; P.bins = h->bins; P.ncat = h->ncat; P.single_pass = h->single_pass ? 1 : 0;
    P.bin_to_cat = h->d_bin_to_cat;
    P.num_hashes = h->num_hashes.as<uint32_t>(); P.counts = h->counts.as<uint32_t>(); P.unique_counts = h->unique.as<uint32_t>();
    P.work_counter = h->counter.as<uint32_t>();
    P.seg_len = h->staged_has_seg_len ? h->seg_len.as<uint64_t>() : nullptr;
    P.long_read_min = 0; P.long_only = 0;
    P.dbg_out = nullptr; P.dbg_cap = 0; P.dbg_n = nullptr;
    P.n_parts = h->n_parts; P.my_part = h->my_part; P.rows_per_part = h->rows_per_part; P.packed = uint32_t(h->staged_format);
    P.n_exc = h->staged_n_exc; P.exc_pos = h->exc_pos.as<unsigned long long>(); P.exc_seg_off = h->exc_seg_off.as<uint32_t>();
    P.async_gather = h->async_gather_mode < 0 ? (h->filter_bytes > ASYNC_GATHER_MIN_FILTER ? 1 : 0) : uint32_t(h->async_gather_mode);
    if (h->n_parts && h->async_gather_mode < 0) P.async_gather = 1;             // rows in peer partitions: the NVLink round trip wants the asynchronous carrier
    if (h->hash_funs != 3 || h->mode != 0) P.async_gather = 0;                  // the cp.async carrier exists for 3 hash functions on narrow rows only
    for (int i = 0; i < CG_MAX_PARTS; ++i) P.part_base[i] = h->part_base[i];
    // bins of categories that own several bins; parking pays while they are few (CHARON_ONE_PASS=0 switches it off)
    P.amb_mode = 0;
    for (int j = 0; j < 4; ++j) { P.multibin[j] = 0; P.static_best[j] = 0; }
    if (!h->single_pass && h->bins > 4) {
        uint32_t per_cat[256] = {};
        for (uint32_t b = 0; b < h->bins; ++b) ++per_cat[h->h_bin_to_cat[b]];
        uint32_t multi = 0;
        for (uint32_t b = 0; b < h->bins; ++b) {
            if (per_cat[h->h_bin_to_cat[b]] > 1) { P.multibin[b >> 6] |= 1ull << (b & 63u); ++multi; }
            else P.static_best[b >> 6] |= 1ull << (b & 63u);
        }
        const char *ope = std::getenv("CHARON_ONE_PASS");
        const int one_pass_env = ope ? std::atoi(ope) : -1;
        P.amb_mode = one_pass_env >= 0 ? uint32_t(one_pass_env != 0) : uint32_t(multi * 8 <= h->bins);
    }
    return P;
}

template <int MODE> cudaError_t launch_count_mode(cg_handle *h, const cgk::CountParams &P, int grid) {
    static bool attr_done[4] = {false, false, false, false};
    if (!attr_done[MODE]) {
        cudaError_t e = cudaFuncSetAttribute(cgk::count_kernel<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(cgk::SMEM_BYTES));
        if (e != cudaSuccess) return e;
        attr_done[MODE] = true;
    }
    cgk::count_kernel<MODE><<<grid, cgk::NT, cgk::SMEM_BYTES, h->stream>>>(P);
    return cudaGetLastError();
}
// register-resident kernel for the (k, w) every `charon index` run produces by default (k=19, w=41)
constexpr int FAST_K = 19, FAST_WV = 23;
inline bool use_fast(const cg_handle *h) { return h->K == FAST_K && h->W - h->K + 1 == FAST_WV && !h->force_generic; }
template <int MODE> cudaError_t launch_count_fast(cg_handle *h, const cgk::CountParams &P, int grid) {
    using C = cgk::V2Cfg<FAST_K, FAST_WV>;
    static bool attr_done[4] = {false, false, false, false};
    if (!attr_done[MODE]) {
        cudaError_t e = cudaFuncSetAttribute(cgk::count_kernel_v2<MODE, FAST_K, FAST_WV>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(C::SMEM));
        if (e != cudaSuccess) return e;
        attr_done[MODE] = true;
    }
    cgk::count_kernel_v2<MODE, FAST_K, FAST_WV><<<grid, cgk::NT2, C::SMEM, h->stream>>>(P);
    return cudaGetLastError();
}
template <int MODE> int occupancy_fast() {
    using C = cgk::V2Cfg<FAST_K, FAST_WV>;
    int n = 0;
    cudaFuncSetAttribute(cgk::count_kernel_v2<MODE, FAST_K, FAST_WV>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(C::SMEM));
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, cgk::count_kernel_v2<MODE, FAST_K, FAST_WV>, cgk::NT2, C::SMEM);
    return n;
}
// warp-per-read kernel (count_kernel_v3): SMALL = at most 4 bins (register tallies, no per-warp counter arrays)
constexpr unsigned long long LONG_READ_MIN = 65536;   // reads at least this long go to the CTA-per-read kernel
template <int MODE, bool SMALL, bool AG, int FMT> cudaError_t launch_count_warp_fmt(cg_handle *h, const cgk::CountParams &P, int grid) {
    using C = cgk::V3Cfg<FAST_K, FAST_WV, SMALL, AG>;
    static bool attr_done = false;
    if (!attr_done) {
        cudaError_t e = cudaFuncSetAttribute(cgk::count_kernel_v3<MODE, FAST_K, FAST_WV, SMALL, AG, FMT>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(C::SMEM));
        if (e != cudaSuccess) return e;
        attr_done = true;
    }
    cgk::count_kernel_v3<MODE, FAST_K, FAST_WV, SMALL, AG, FMT><<<grid, cgk::V3Shape<AG, SMALL>::WARPS * 32, C::SMEM, h->stream>>>(P);
    return cudaGetLastError();
}
template <int MODE, bool SMALL, bool AG = false> cudaError_t launch_count_warp(cg_handle *h, const cgk::CountParams &P, int grid) {
    return P.packed == CG_BASES_2BIT ? launch_count_warp_fmt<MODE, SMALL, AG, CG_BASES_2BIT>(h, P, grid) : launch_count_warp_fmt<MODE, SMALL, AG, CG_BASES_ASCII>(h, P, grid);
}
template <int MODE, bool SMALL, bool AG = false> int occupancy_warp() {   // the same for both base formats (same registers to within ptxas' whim: take the lower)
    using C = cgk::V3Cfg<FAST_K, FAST_WV, SMALL, AG>;
    int n0 = 0, n2 = 0;
    if (cudaFuncSetAttribute(cgk::count_kernel_v3<MODE, FAST_K, FAST_WV, SMALL, AG, CG_BASES_ASCII>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(C::SMEM)) != cudaSuccess ||
        cudaFuncSetAttribute(cgk::count_kernel_v3<MODE, FAST_K, FAST_WV, SMALL, AG, CG_BASES_2BIT>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(C::SMEM)) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n0, cgk::count_kernel_v3<MODE, FAST_K, FAST_WV, SMALL, AG, CG_BASES_ASCII>, cgk::V3Shape<AG, SMALL>::WARPS * 32, C::SMEM);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n2, cgk::count_kernel_v3<MODE, FAST_K, FAST_WV, SMALL, AG, CG_BASES_2BIT>, cgk::V3Shape<AG, SMALL>::WARPS * 32, C::SMEM);
    return std::min(n0, n2);
}
template <int MODE> int occupancy_mode() {
    int n = 0;
    cudaFuncSetAttribute(cgk::count_kernel<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(cgk::SMEM_BYTES));
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, cgk::count_kernel<MODE>, cgk::NT, cgk::SMEM_BYTES);
    return n;
}

int launch_count(cg_handle *h, const cgk::CountParams &P) {
    if (P.n_reads == 0) return 0;
    if (h->n_parts && P.filter) {
        for (uint32_t p = 0; p < h->n_parts; ++p)
            if (!h->part_base[p]) return fail(h, CG_ERR_STATE, "partitioned filter: part %u of %u is not attached (cg_partition_attach_*)", p, h->n_parts);
    }
    int grid = int(std::min<uint64_t>(uint64_t(h->sm_count) * h->grid_per_sm, P.n_reads));
    CU(h, cudaMemsetAsync(h->counter.p, 0, 4, h->stream));
    cudaError_t e;
    if (use_fast(h) && !h->force_cta) {
        // warp-per-read kernel for everything shorter than LONG_READ_MIN, then (only if needed) the CTA-per-read kernel
        cgk::CountParams Q = P;
        const bool dispatch = Q.filter != nullptr;           // index-build chunks (capture-only launch) are all handled by the warp kernel
        Q.long_read_min = dispatch ? LONG_READ_MIN : 0;
        const bool use_ag = Q.async_gather && h->mode == 0;
        const bool small = h->bins <= 4;
        const uint64_t wpc = use_ag ? (small ? cgk::V3Shape<true, true>::WARPS : cgk::V3Shape<true, false>::WARPS) : cgk::V3Shape<false>::WARPS;   // warps (= reads in flight) per CTA
        const int wgrid = int(std::min<uint64_t>(uint64_t(h->sm_count) * (use_ag ? h->warp_grid_per_sm_ag : h->warp_grid_per_sm), (uint64_t(P.n_reads) + wpc - 1) / wpc));
        switch (h->mode) {
            case 0:   // narrow rows: the deferred gathers fly as cp.async when the filter is beyond L2 (count_kernel_v3.cuh)
                if (Q.async_gather) e = small ? launch_count_warp<0, true, true>(h, Q, wgrid) : launch_count_warp<0, false, true>(h, Q, wgrid);
                else e = small ? launch_count_warp<0, true>(h, Q, wgrid) : launch_count_warp<0, false>(h, Q, wgrid);
                break;
            case 1: e = launch_count_warp<1, false>(h, Q, wgrid); break;
            case 2: e = launch_count_warp<2, false>(h, Q, wgrid); break;
            default: e = launch_count_warp<3, false>(h, Q, wgrid); break;
        }
        h->info.count_grid = uint32_t(wgrid);
        if (e == cudaSuccess && dispatch && h->staged_max_read_len >= LONG_READ_MIN) {
            Q.long_only = 1;
            e = cudaMemsetAsync(h->counter.p, 0, 4, h->stream);
            if (e == cudaSuccess)
                switch (h->mode) {
                    case 0: e = launch_count_fast<0>(h, Q, grid); break;
                    case 1: e = launch_count_fast<1>(h, Q, grid); break;
                    case 2: e = launch_count_fast<2>(h, Q, grid); break;
                    default: e = launch_count_fast<3>(h, Q, grid); break;
                }
            h->info.last_kernel_launches += 1;
        }
        if (e != cudaSuccess) return fail(h, CG_ERR_CUDA, "count kernel launch failed: %s", cudaGetErrorString(e));
        h->info.last_kernel_launches += 1;
        return 0;
    } else if (use_fast(h)) {
        switch (h->mode) {
            case 0: e = launch_count_fast<0>(h, P, grid); break;
            case 1: e = launch_count_fast<1>(h, P, grid); break;
            case 2: e = launch_count_fast<2>(h, P, grid); break;
            default: e = launch_count_fast<3>(h, P, grid); break;
        }
    } else
    switch (h->mode) {
        case 0: e = launch_count_mode<0>(h, P, grid); break;
        case 1: e = launch_count_mode<1>(h, P, grid); break;
        case 2: e = launch_count_mode<2>(h, P, grid); break;
        default: e = launch_count_mode<3>(h, P, grid); break;
    }
    if (e != cudaSuccess) return fail(h, CG_ERR_CUDA, "count kernel launch failed: %s", cudaGetErrorString(e));
    h->info.count_grid = uint32_t(grid);
    h->info.last_kernel_launches += 1;
    return 0;
}

cgk::EvalParams make_eval(cg_handle *h, uint32_t n_reads) {
    cgk::EvalParams E{};
    E.n_reads = n_reads; E.ncat = h->ncat; E.mode = h->call_mode; E.models = h->d_models; E.pool = h->d_pool;
    E.log_rate = std::log(300.0f);
    const cg_thresholds &t = h->thr;
    E.min_quality = t.min_quality; E.min_length = t.min_length; E.min_compression = t.min_compression;
    E.confidence_threshold = t.confidence_threshold; E.confidence_probability_threshold = t.confidence_probability_threshold;
    E.min_hits = t.min_hits; E.host_unique_prop_lo_threshold = t.host_unique_prop_lo_threshold;
    E.min_proportion_difference = t.min_proportion_difference; E.min_prob_difference = t.min_prob_difference;
    E.host_index = t.host_index;
    E.num_hashes = h->num_hashes.as<uint32_t>(); E.counts = h->counts.as<uint32_t>(); E.unique_counts = h->unique.as<uint32_t>();
    E.mean_quality = h->mq.as<float>(); E.compression = h->cr.as<float>(); E.length = h->len.as<uint32_t>();
    E.prob = h->prob.as<double>(); E.call = h->call.as<uint8_t>(); E.confidence = h->conf.as<uint8_t>();
    return E;
}

int launch_eval(cg_handle *h, uint32_t n_reads) {
    if (!h->model_set) return fail(h, CG_ERR_STATE, "cg_set_model must be called before evaluation");
    if (n_reads == 0) return 0;
    cgk::EvalParams E = make_eval(h, n_reads);
    const size_t pairs = size_t(n_reads) * h->ncat;
    // memoised evaluation when some category is a KDE and the table (2 slots per pair, at most 2^23) stays modest
    static const bool memo_env = [] { const char *e = std::getenv("CHARON_EVAL_MEMO"); return !e || std::atoi(e) != 0; }();
    size_t table = 1024;
    while (table < 2 * pairs) table <<= 1;
    const bool memo = memo_env && h->model_has_kde && pairs >= 4096 && table <= (size_t(1) << 23);
    if (memo) {
        RSV(h, h->memo_keys, table * 8); RSV(h, h->memo_vals, table * 8); RSV(h, h->memo_list, pairs * 4); RSV(h, h->memo_slot, pairs * 4);
        RSV(h, h->memo_count, 16);
        CU(h, cudaMemsetAsync(h->memo_keys.p, 0xFF, table * 8, h->stream));
        CU(h, cudaMemsetAsync(h->memo_count.p, 0, 4, h->stream));
        E.memo_keys = h->memo_keys.as<unsigned long long>(); E.memo_vals = h->memo_vals.as<double>(); E.memo_list = h->memo_list.as<uint32_t>();
        E.memo_count = h->memo_count.as<uint32_t>(); E.memo_slot = h->memo_slot.as<uint32_t>(); E.memo_mask = uint32_t(table - 1);
        cgk::memo_prepare_kernel<<<unsigned((pairs + 127) / 128), 128, 0, h->stream>>>(E);
        cgk::memo_unique_kernel<<<unsigned((pairs + 127) / 128), 128, 0, h->stream>>>(E);
        h->info.last_kernel_launches += 1;
    } else {
        cgk::prob_kernel<<<unsigned((pairs + 127) / 128), 128, 0, h->stream>>>(E);
    }
    cgk::call_kernel<<<(n_reads + 127) / 128, 128, 0, h->stream>>>(E);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail(h, CG_ERR_CUDA, "eval kernel launch failed: %s", cudaGetErrorString(e));
    h->info.last_kernel_launches += 2;
    return 0;
}

int reserve_outputs(cg_handle *h, uint32_t n_reads) {
    const size_t n = std::max<uint32_t>(n_reads, 1);
    RSV(h, h->num_hashes, n * 4); RSV(h, h->counts, n * h->ncat * 4); RSV(h, h->unique, n * h->ncat * 4);
    RSV(h, h->prob, n * h->ncat * 8); RSV(h, h->call, n); RSV(h, h->conf, n);
    RSV(h, h->mq, n * 4); RSV(h, h->cr, n * 4); RSV(h, h->len, n * 4);
    RSV(h, h->counter, 16);
    return 0;
}

int validate_batch(cg_handle *h, const cg_batch *b) {
    if (!b || (!b->seg_off && b->n_segs) || (!b->bases && b->n_segs && b->seg_off[b->n_segs] > 0)) return fail(h, CG_ERR_ARG, "null batch pointers");
    if (!b->read_seg && b->n_segs != b->n_reads) return fail(h, CG_ERR_ARG, "read_seg is NULL but n_segs != n_reads");
    if (b->read_seg && (b->read_seg[0] != 0 || b->read_seg[b->n_reads] != b->n_segs)) return fail(h, CG_ERR_ARG, "read_seg must span [0, n_segs]");
    if (b->bases_format < CG_BASES_ASCII || b->bases_format > CG_BASES_2BIT) return fail(h, CG_ERR_ARG, "bases_format %d unknown", b->bases_format);
    if (b->n_exc && (b->bases_format != CG_BASES_2BIT || !b->exc_pos)) return fail(h, CG_ERR_ARG, "exc_pos / n_exc only go with CG_BASES_2BIT");
    // one pass over the offsets: a non-monotone table would make segment lengths wrap around and the kernels read out of bounds
    if (b->n_segs && b->seg_off[0] != 0) return fail(h, CG_ERR_ARG, "seg_off[0] must be 0");
    for (uint32_t s = 0; s < b->n_segs; ++s) {
        if (b->seg_off[s + 1] < b->seg_off[s]) return fail(h, CG_ERR_ARG, "seg_off is not non-decreasing at segment %u", s);
        if (b->seg_len && b->seg_len[s] > b->seg_off[s + 1] - b->seg_off[s]) return fail(h, CG_ERR_ARG, "seg_len[%u] runs past seg_off[%u]", s, s + 1);
    }
    if (b->read_seg)
        for (uint32_t r = 0; r < b->n_reads; ++r)
            if (b->read_seg[r + 1] < b->read_seg[r]) return fail(h, CG_ERR_ARG, "read_seg is not non-decreasing at read %u", r);
    for (uint64_t e = 0; e < b->n_exc; ++e)
        if (b->exc_pos[e] >= b->seg_off[b->n_segs] || (e && b->exc_pos[e] <= b->exc_pos[e - 1])) return fail(h, CG_ERR_ARG, "exc_pos must be ascending batch positions (entry %llu)", (unsigned long long)e);
    return 0;
}

// packed input of the generic any-(k, w) kernel: expanded back to characters on the device
__global__ void unpack_bases_kernel(const uint8_t *src, int format, uint64_t n, uint8_t *dst) {
    const uint64_t stride = uint64_t(gridDim.x) * blockDim.x;
    for (uint64_t i = blockIdx.x * uint64_t(blockDim.x) + threadIdx.x; i < n; i += stride) {
        if (format == CG_BASES_RANK4) dst[i] = uint8_t("ACGNTNNN"[(src[i >> 1] >> ((i & 1u) * 4u)) & 7u]);
        else dst[i] = uint8_t("ACGT"[(src[i >> 2] >> ((i & 3u) * 2u)) & 3u]);
    }
}
__global__ void mark_exceptions_kernel(const unsigned long long *exc, uint64_t n_exc, uint8_t *dst) {
    const uint64_t stride = uint64_t(gridDim.x) * blockDim.x;
    for (uint64_t e = blockIdx.x * uint64_t(blockDim.x) + threadIdx.x; e < n_exc; e += stride) dst[exc[e]] = 'N';
}

// H2D of one batch (src/dehost_main.cpp:335-339: the chunk of records handed to the loop)
constexpr uint64_t PACK_MIN_BASES = 1u << 20, PACK_CHUNK_BASES = 32u << 20;
int stage(cg_handle *h, const cg_batch *b, const cg_read_scalars *sc) {
    if (h->active_pending) return fail(h, CG_ERR_STATE, "a submitted batch is still in flight on this slot: cg_wait_batch first");
    int rc = validate_batch(h, b);
    if (rc) return rc;
    const uint64_t nbases = b->n_segs ? b->seg_off[b->n_segs] : 0;
    const int in_fmt = b->bases_format;
    // Packed transport made by the library: only for ASCII input of the (19, 41) kernels; in automatic mode tiny batches are
    // not worth waking the pool for.  Two-bit input is copied as it is; rank nibbles (and any packed input of the generic
    // any-(k, w) kernel) are expanded back to characters on the device.
    const bool pack = in_fmt == CG_BASES_ASCII && use_fast(h) && nbases > 0 &&
                      (h->host_pack_mode == 1 || (h->host_pack_mode < 0 && nbases >= PACK_MIN_BASES && cgh::threads() >= 8));
    const bool unpack = in_fmt == CG_BASES_RANK4 || (in_fmt == CG_BASES_2BIT && !use_fast(h));
    const int dev_fmt = pack ? int(CG_BASES_2BIT) : (unpack ? int(CG_BASES_ASCII) : in_fmt);
    const size_t in_bytes = in_fmt == CG_BASES_2BIT ? size_t((nbases + 3) / 4) : (in_fmt == CG_BASES_RANK4 ? size_t((nbases + 1) / 2) : size_t(nbases));
    const size_t dev_bytes = dev_fmt == CG_BASES_2BIT ? size_t((nbases + 3) / 4) : size_t(nbases);
    RSV(h, h->bases, dev_bytes + 64);
    RSV(h, h->seg_off, (size_t(b->n_segs) + 1) * 8);
    if (b->read_seg) RSV(h, h->read_seg, (size_t(b->n_reads) + 1) * 4);
    if (b->seg_len) RSV(h, h->seg_len, size_t(b->n_segs) * 8 + 8);
    rc = reserve_outputs(h, b->n_reads);
    if (rc) return rc;
    h->staged_packed = pack; h->staged_format = dev_fmt; h->staged_has_seg_len = b->seg_len != nullptr; h->staged_n_exc = 0;
    h->info.last_packed = dev_fmt != CG_BASES_ASCII ? 1 : 0; h->info.last_pack_ms = 0;
    const uint64_t *exc = b->exc_pos;      // exception list that goes with two-bit bases on the device
    uint64_t n_exc = dev_fmt == CG_BASES_2BIT ? b->n_exc : 0;
    std::vector<uint64_t> own_exc;
    if (pack) {
        const size_t pbytes = size_t((nbases + 3) / 4);
        auto &ps = h->pack;
        if (!ps.done) CU(h, cudaEventCreateWithFlags(&ps.done, cudaEventDisableTiming));
        if (ps.armed) { CU(h, cudaEventSynchronize(ps.done)); ps.armed = false; }   // previous copy out of the staging buffer is over
        if (ps.cap < pbytes) {
            if (ps.p) cudaFreeHost(ps.p);
            ps.p = nullptr; ps.cap = 0;
            const size_t want = pbytes + pbytes / 8 + 4096;
            CU(h, cudaHostAlloc(&ps.p, want, cudaHostAllocDefault));
            ps.cap = want;
        }
        CU(h, cudaEventRecord(h->ev[0], h->stream));
        cudaError_t ce = cudaSuccess;
        const auto t0 = std::chrono::steady_clock::now();
        cgh::pack2_chunked(reinterpret_cast<const uint8_t *>(b->bases), static_cast<uint8_t *>(ps.p), nbases, PACK_CHUNK_BASES,
                           [&](uint64_t lo, uint64_t hi) {   // chunk [lo, hi) is packed (lo is a multiple of 4): its copy overlaps the packing of the next
                               if (ce == cudaSuccess)
                                   ce = cudaMemcpyAsync(h->bases.as<uint8_t>() + lo / 4, static_cast<const uint8_t *>(ps.p) + lo / 4, size_t((hi - lo + 3) / 4),
                                                        cudaMemcpyHostToDevice, h->stream);
                           }, own_exc);
        h->info.last_pack_ms = std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - t0).count();
        CU(h, ce);
        CU(h, cudaEventRecord(ps.done, h->stream));
        ps.armed = true;
        h->info.last_h2d_bytes = pbytes;
        exc = own_exc.data(); n_exc = own_exc.size();
    } else if (unpack) {
        CU(h, cudaEventRecord(h->ev[0], h->stream));
        RSV(h, h->unpack_src, in_bytes + 64);
        if (in_bytes) CU(h, cudaMemcpyAsync(h->unpack_src.p, b->bases, in_bytes, cudaMemcpyHostToDevice, h->stream));
        if (b->n_exc) {
            RSV(h, h->exc_pos, size_t(b->n_exc) * 8);
            CU(h, cudaMemcpyAsync(h->exc_pos.p, b->exc_pos, size_t(b->n_exc) * 8, cudaMemcpyHostToDevice, h->stream));
        }
        if (nbases) {
            unpack_bases_kernel<<<h->sm_count * 8, 256, 0, h->stream>>>(h->unpack_src.as<uint8_t>(), in_fmt, nbases, h->bases.as<uint8_t>());
            if (b->n_exc) mark_exceptions_kernel<<<64, 256, 0, h->stream>>>(h->exc_pos.as<unsigned long long>(), b->n_exc, h->bases.as<uint8_t>());
            CU(h, cudaGetLastError());
        }
        h->info.last_h2d_bytes = in_bytes + size_t(b->n_exc) * 8;
    } else {
        CU(h, cudaEventRecord(h->ev[0], h->stream));
        if (in_bytes) CU(h, cudaMemcpyAsync(h->bases.p, b->bases, in_bytes, cudaMemcpyHostToDevice, h->stream));
        h->info.last_h2d_bytes = in_bytes;
    }
    if (dev_fmt == CG_BASES_2BIT && n_exc) {
        // exception list + the index range of every segment in it (one merge pass; reads rarely hold an N)
        if (n_exc > 0xFFFFFFF0ull) return fail(h, CG_ERR_ARG, "more than 2^32 non-ACGT bases in one batch");
        std::vector<uint32_t> off(size_t(b->n_segs) + 1);
        uint64_t e = 0;
        for (uint32_t s = 0; s < b->n_segs; ++s) {
            while (e < n_exc && exc[e] < b->seg_off[s]) ++e;
            off[s] = uint32_t(e);
        }
        off[b->n_segs] = uint32_t(n_exc);
        RSV(h, h->exc_pos, size_t(n_exc) * 8); RSV(h, h->exc_seg_off, off.size() * 4);
        CU(h, cudaMemcpyAsync(h->exc_pos.p, exc, size_t(n_exc) * 8, cudaMemcpyHostToDevice, h->stream));
        CU(h, cudaMemcpyAsync(h->exc_seg_off.p, off.data(), off.size() * 4, cudaMemcpyHostToDevice, h->stream));
        CU(h, cudaStreamSynchronize(h->stream));   // `off` (and the library's own list) are pageable and die with this scope
        h->staged_n_exc = n_exc;
        h->info.last_h2d_bytes += size_t(n_exc) * 8 + off.size() * 4;
    }
    h->info.last_h2d_bytes += (size_t(b->n_segs) + 1) * 8 + (b->read_seg ? (size_t(b->n_reads) + 1) * 4 : 0) + (b->seg_len ? size_t(b->n_segs) * 8 : 0) + (sc && b->n_reads ? size_t(b->n_reads) * 12 : 0);
    static const uint64_t zero_off[1] = {0};
    CU(h, cudaMemcpyAsync(h->seg_off.p, b->n_segs ? b->seg_off : zero_off, (size_t(b->n_segs) + 1) * 8, cudaMemcpyHostToDevice, h->stream));
    if (b->read_seg) CU(h, cudaMemcpyAsync(h->read_seg.p, b->read_seg, (size_t(b->n_reads) + 1) * 4, cudaMemcpyHostToDevice, h->stream));
    if (b->seg_len && b->n_segs) CU(h, cudaMemcpyAsync(h->seg_len.p, b->seg_len, size_t(b->n_segs) * 8, cudaMemcpyHostToDevice, h->stream));
    h->staged_has_scalars = false; h->staged_device_cr = false;
    if (sc && b->n_reads) {
        if (!sc->mean_quality || !sc->length) return fail(h, CG_ERR_ARG, "null read scalars");
        CU(h, cudaMemcpyAsync(h->mq.p, sc->mean_quality, size_t(b->n_reads) * 4, cudaMemcpyHostToDevice, h->stream));
        // compression == NULL: the library computes the ratios on the device (gzip_size_kernel) before the model kernels
        if (sc->compression) CU(h, cudaMemcpyAsync(h->cr.p, sc->compression, size_t(b->n_reads) * 4, cudaMemcpyHostToDevice, h->stream));
        h->staged_device_cr = sc->compression == nullptr;
        CU(h, cudaMemcpyAsync(h->len.p, sc->length, size_t(b->n_reads) * 4, cudaMemcpyHostToDevice, h->stream));
        h->staged_has_scalars = true;
    }
    CU(h, cudaEventRecord(h->ev[1], h->stream));
    h->staged_reads = b->n_reads; h->staged_segs = b->n_segs; h->staged_bases = nbases; h->staged_has_read_seg = b->read_seg != nullptr;
    h->staged_max_read_len = 0;
    for (uint32_t r = 0; r < b->n_reads; ++r) {   // span of the read in the buffer (>= its bases): only steers the long-read dispatch
        const uint32_t s0 = b->read_seg ? b->read_seg[r] : r, s1 = b->read_seg ? b->read_seg[r + 1] : r + 1;
        if (s1 > s0) h->staged_max_read_len = std::max<uint64_t>(h->staged_max_read_len, b->seg_off[s1] - b->seg_off[s0]);
    }
    h->info.last_bases = nbases;
    return 0;
}

// gzip size (and ratio) of every staged read: one warp per read, per-warp scratch of 5 bytes per base of the longest read
int launch_gzip(cg_handle *h, float *ratio_out) {
    const uint32_t n = h->staged_reads;
    if (!n) return 0;
    if (h->staged_max_read_len > dsz::DSZ_MAX_LEN)
        return fail(h, CG_ERR_ARG, "compression ratio on the device: a read of this batch spans %llu bases, the limit is %u: "
                    "compute the ratios of this batch on the host and pass them in cg_read_scalars", (unsigned long long)h->staged_max_read_len, dsz::DSZ_MAX_LEN);
    static int blocks_per_sm = 0;
    if (!blocks_per_sm) {
        cudaError_t e = cudaFuncSetAttribute(cgk::gzip_size_kernel<uint16_t>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(cgk::GZ_SMEM));
        if (e == cudaSuccess) e = cudaFuncSetAttribute(cgk::gzip_size_kernel<uint32_t>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(cgk::GZ_SMEM));
        if (e != cudaSuccess) return fail(h, CG_ERR_CUDA, "gzip_size_kernel attribute: %s", cudaGetErrorString(e));
        int nb = 0;
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, cgk::gzip_size_kernel<uint32_t>, cgk::GZ_WARPS * 32, cgk::GZ_SMEM);
        blocks_per_sm = nb > 0 ? nb : 1;
        if (const char *e = std::getenv("CHARON_GZ_BLOCKS")) { const int v = std::atoi(e); if (v >= 1 && v < blocks_per_sm) blocks_per_sm = v; }
    }
    const uint64_t cap = (std::max<uint64_t>(h->staged_max_read_len, 64) + 63) / 64 * 64;
    const uint64_t want_blocks = (uint64_t(n) + cgk::GZ_WARPS - 1) / cgk::GZ_WARPS;
    const unsigned grid = unsigned(std::min<uint64_t>(want_blocks, uint64_t(h->sm_count) * blocks_per_sm));
    RSV(h, h->gz_scratch, size_t(grid) * cgk::GZ_WARPS * (cap * 9 + 64));
    RSV(h, h->gz_size, size_t(n) * 4);
    cgk::GzParams G{};
    G.bases = h->bases.as<uint8_t>(); G.format = uint32_t(h->staged_format);
    G.seg_off = h->seg_off.as<uint64_t>(); G.seg_len = h->staged_has_seg_len ? h->seg_len.as<uint64_t>() : nullptr;
    G.read_seg = h->staged_has_read_seg ? h->read_seg.as<uint32_t>() : nullptr;
    G.exc_pos = h->exc_pos.as<unsigned long long>(); G.exc_seg_off = h->exc_seg_off.as<uint32_t>(); G.n_exc = h->staged_n_exc;
    G.n_reads = n; G.work_counter = h->counter.as<uint32_t>();
    G.scratch = h->gz_scratch.as<uint8_t>(); G.scratch_stride = cap * 9 + 64; G.cap = uint32_t(cap);
    G.size_out = h->gz_size.as<uint32_t>(); G.ratio_out = ratio_out;
    { const char *e = std::getenv("CHARON_GZ_PREFETCH"); G.prefetch = e ? uint32_t(std::atoi(e)) : 0u; }   // measured: no gain (the step is a chain of dependent instructions, not a wait for memory)
    CU(h, cudaMemsetAsync(h->counter.p, 0, 4, h->stream));
    if (cap <= 65536) cgk::gzip_size_kernel<uint16_t><<<grid, cgk::GZ_WARPS * 32, cgk::GZ_SMEM, h->stream>>>(G);
    else cgk::gzip_size_kernel<uint32_t><<<grid, cgk::GZ_WARPS * 32, cgk::GZ_SMEM, h->stream>>>(G);
    CU(h, cudaGetLastError());
    h->info.last_kernel_launches += 1;
    return 0;
}

int run(cg_handle *h, bool with_eval) {
    h->info.last_kernel_launches = 0;
    if (with_eval && h->staged_device_cr) {
        CU(h, cudaEventRecord(h->ev[7], h->stream));
        int rc = launch_gzip(h, h->cr.as<float>());
        if (rc) return rc;
    }
    CU(h, cudaEventRecord(h->ev[2], h->stream));
    if (h->timing) CU(h, cudaEventRecord(h->next_tev(), h->stream));
    int rc = launch_count(h, make_params(h));
    if (rc) return rc;
    CU(h, cudaEventRecord(h->ev[3], h->stream));
    if (h->timing) CU(h, cudaEventRecord(h->next_tev(), h->stream));
    if (with_eval) {
        if (!h->staged_has_scalars) return fail(h, CG_ERR_STATE, "evaluation needs the per-read scalars (mean_quality, compression, length)");
        rc = launch_eval(h, h->staged_reads);
        if (rc) return rc;
    }
    CU(h, cudaEventRecord(h->ev[4], h->stream));
    if (h->timing) CU(h, cudaEventRecord(h->next_tev(), h->stream));
    return 0;
}

int fetch_enqueue(cg_handle *h, const cg_counts_out *co, const cg_calls_out *ca) {
    const size_t n = h->staged_reads, nc = h->ncat;
    CU(h, cudaEventRecord(h->ev[5], h->stream));
    if (n) {
        if (co) {
            if (co->num_hashes) CU(h, cudaMemcpyAsync(co->num_hashes, h->num_hashes.p, n * 4, cudaMemcpyDeviceToHost, h->stream));
            if (co->counts) CU(h, cudaMemcpyAsync(co->counts, h->counts.p, n * nc * 4, cudaMemcpyDeviceToHost, h->stream));
            if (co->unique_counts) CU(h, cudaMemcpyAsync(co->unique_counts, h->unique.p, n * nc * 4, cudaMemcpyDeviceToHost, h->stream));
        }
        if (ca) {
            if (ca->probabilities) CU(h, cudaMemcpyAsync(ca->probabilities, h->prob.p, n * nc * 8, cudaMemcpyDeviceToHost, h->stream));
            if (ca->call) CU(h, cudaMemcpyAsync(ca->call, h->call.p, n, cudaMemcpyDeviceToHost, h->stream));
            if (ca->confidence) CU(h, cudaMemcpyAsync(ca->confidence, h->conf.p, n, cudaMemcpyDeviceToHost, h->stream));
        }
    }
    CU(h, cudaEventRecord(h->ev[6], h->stream));
    return 0;
}
int fetch_finish(cg_handle *h) {
    CU(h, cudaStreamSynchronize(h->stream));
    float ms = 0;
    if (cudaEventElapsedTime(&ms, h->ev[0], h->ev[1]) == cudaSuccess) h->info.last_h2d_ms = ms;
    if (cudaEventElapsedTime(&ms, h->ev[2], h->ev[3]) == cudaSuccess) h->info.last_count_kernel_ms = ms;
    if (cudaEventElapsedTime(&ms, h->ev[3], h->ev[4]) == cudaSuccess) h->info.last_eval_kernel_ms = ms;
    if (cudaEventElapsedTime(&ms, h->ev[5], h->ev[6]) == cudaSuccess) h->info.last_d2h_ms = ms;
    h->info.last_gzip_kernel_ms = 0;
    if (h->staged_device_cr && cudaEventElapsedTime(&ms, h->ev[7], h->ev[2]) == cudaSuccess) h->info.last_gzip_kernel_ms = ms;
    // CHARON_TRACE_SLOTS=1: device-side timeline of this batch (ms since the first traced batch's copy began), one line per batch:
    // where copies and kernels of the two slots really overlap (tools/e2e_probe.py)
    static const bool trace = std::getenv("CHARON_TRACE_SLOTS") != nullptr;
    if (trace) {
        static cudaEvent_t base = nullptr;
        static std::chrono::steady_clock::time_point hbase;
        if (!base) {
            cudaEventCreate(&base); cudaEventRecord(base, h->stream); cudaEventSynchronize(base);
            hbase = std::chrono::steady_clock::now();
        }
        float t[7] = {};
        for (int i = 0; i < 7; ++i) if (cudaEventElapsedTime(&t[i], base, h->ev[i]) != cudaSuccess) t[i] = -1;
        std::fprintf(stderr, "[slot %d] wait returned at host %.3f | device: h2d %.3f..%.3f  count %.3f..%.3f  eval ..%.3f  d2h %.3f..%.3f\n", h->active_slot,
                     std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - hbase).count(), t[0], t[1], t[2], t[3], t[4], t[5], t[6]);
    }
    cudaGetLastError();
    return 0;
}
int fetch(cg_handle *h, const cg_counts_out *co, const cg_calls_out *ca) {
    int rc = fetch_enqueue(h, co, ca);
    return rc ? rc : fetch_finish(h);
}

}  // namespace

#include "charon_gpu_api.inc"

// ---- cg_eval_batch_static: the model/call kernels with NO allocation inside the call ------------------------------------
// No counterpart in the reference; added for callers that may not call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  prob_kernel writes the per-(read,
// category) probabilities that call_kernel reads back: that scratch, and the device copies of the inputs and outputs,
// live in ONE file-scope device array of CG_EVAL_STATIC_BYTES instead of the handle's cudaMalloc'd buffers.  Every GPU
// that loads this library holds the array until the process exits.  All calls (and all handles) on a device share it and
// it keeps the previous call's contents; two calls that overlap on the GPU race.
namespace {

constexpr size_t ES_READS = CG_EVAL_STATIC_MAX_READS, ES_PAIRS = CG_EVAL_STATIC_MAX_PAIRS;
// fixed layout, every part a multiple of 256 bytes
constexpr size_t ES_OFF_NH = 0, ES_OFF_MQ = ES_OFF_NH + ES_READS * 4, ES_OFF_CR = ES_OFF_MQ + ES_READS * 4, ES_OFF_LEN = ES_OFF_CR + ES_READS * 4,
                 ES_OFF_CNT = ES_OFF_LEN + ES_READS * 4, ES_OFF_UNQ = ES_OFF_CNT + ES_PAIRS * 4, ES_OFF_PROB = ES_OFF_UNQ + ES_PAIRS * 4,
                 ES_OFF_CALL = ES_OFF_PROB + ES_PAIRS * 8, ES_OFF_CONF = ES_OFF_CALL + ES_READS, ES_TOTAL = ES_OFF_CONF + ES_READS;
static_assert(ES_TOTAL == CG_EVAL_STATIC_BYTES, "include/charon_gpu.h states the size of the static scratch");
static_assert(ES_READS % 256 == 0 && ES_PAIRS % 256 == 0, "parts of the static scratch stay 256-byte aligned");

static __device__ __align__(256) unsigned char g_eval_static[CG_EVAL_STATIC_BYTES];

// In host code `g_eval_static` is only a stand-in for the symbol: the device address is looked up once per device, with
// that device current, and cached here.
constexpr int ES_MAX_DEVICES = 64;
unsigned char *g_eval_static_ptr[ES_MAX_DEVICES] = {};
std::mutex g_eval_static_mx;

int eval_static_address(cg_handle *h, unsigned char **out) {
    if (h->device < 0 || h->device >= ES_MAX_DEVICES) return fail(h, CG_ERR_ARG, "cg_eval_batch_static: device %d out of range", h->device);
    std::lock_guard<std::mutex> lk(g_eval_static_mx);
    if (!g_eval_static_ptr[h->device]) {
        void *p = nullptr;
        CU(h, cudaSetDevice(h->device));
        CU(h, cudaGetSymbolAddress(&p, g_eval_static));
        if (!p) return fail(h, CG_ERR_CUDA, "cudaGetSymbolAddress(g_eval_static) returned a null address");
        g_eval_static_ptr[h->device] = static_cast<unsigned char *>(p);
    }
    *out = g_eval_static_ptr[h->device];
    return 0;
}

}  // namespace

extern "C" int cg_eval_batch_static(cg_handle *h, uint32_t n, const uint32_t *nh, const uint32_t *cnt, const uint32_t *uq,
                                    const cg_read_scalars *sc, const cg_calls_out *out) {
    if (!h) return CG_ERR_ARG;
    if (!h->model_set) return fail(h, CG_ERR_STATE, "cg_set_model must be called before cg_eval_batch_static");
    const size_t nc = h->ncat, pairs = size_t(n) * nc;
    if (n > ES_READS || pairs > ES_PAIRS)
        return fail(h, CG_ERR_ARG, "cg_eval_batch_static: %u reads x %zu categories exceed the static scratch (at most %zu reads and %zu read-category "
                    "pairs per call); split the batch or use cg_eval_batch", n, nc, ES_READS, ES_PAIRS);
    unsigned char *S = nullptr;
    int rc = eval_static_address(h, &S);       // also what a set-up call with n_reads == 0 is for
    if (rc) return rc;
    if (n == 0) return 0;
    if (!nh || !cnt || !uq || !sc || !sc->mean_quality || !sc->compression || !sc->length || !out) return fail(h, CG_ERR_ARG, "null argument");
    CU(h, cudaSetDevice(h->device));
    CU(h, cudaMemcpyAsync(S + ES_OFF_NH, nh, size_t(n) * 4, cudaMemcpyHostToDevice, h->stream));
    CU(h, cudaMemcpyAsync(S + ES_OFF_CNT, cnt, pairs * 4, cudaMemcpyHostToDevice, h->stream));
    CU(h, cudaMemcpyAsync(S + ES_OFF_UNQ, uq, pairs * 4, cudaMemcpyHostToDevice, h->stream));
    CU(h, cudaMemcpyAsync(S + ES_OFF_MQ, sc->mean_quality, size_t(n) * 4, cudaMemcpyHostToDevice, h->stream));
    CU(h, cudaMemcpyAsync(S + ES_OFF_CR, sc->compression, size_t(n) * 4, cudaMemcpyHostToDevice, h->stream));
    CU(h, cudaMemcpyAsync(S + ES_OFF_LEN, sc->length, size_t(n) * 4, cudaMemcpyHostToDevice, h->stream));
    // the array still holds the previous call's probabilities and calls: both kernels overwrite every element of this call's
    // range, and the range is cleared first so that nothing stale can be read back should a launch fail
    CU(h, cudaMemsetAsync(S + ES_OFF_PROB, 0, pairs * 8, h->stream));
    CU(h, cudaMemsetAsync(S + ES_OFF_CALL, 0xFF, n, h->stream));
    CU(h, cudaMemsetAsync(S + ES_OFF_CONF, 0, n, h->stream));
    cgk::EvalParams E = make_eval(h, n);       // model tables and thresholds of the handle (allocated by cg_set_model)
    E.num_hashes = reinterpret_cast<const uint32_t *>(S + ES_OFF_NH);
    E.counts = reinterpret_cast<const uint32_t *>(S + ES_OFF_CNT);
    E.unique_counts = reinterpret_cast<const uint32_t *>(S + ES_OFF_UNQ);
    E.mean_quality = reinterpret_cast<const float *>(S + ES_OFF_MQ);
    E.compression = reinterpret_cast<const float *>(S + ES_OFF_CR);
    E.length = reinterpret_cast<const uint32_t *>(S + ES_OFF_LEN);
    E.prob = reinterpret_cast<double *>(S + ES_OFF_PROB);
    E.call = S + ES_OFF_CALL;
    E.confidence = S + ES_OFF_CONF;
    cgk::prob_kernel<<<unsigned((pairs + 127) / 128), 128, 0, h->stream>>>(E);
    cgk::call_kernel<<<(n + 127) / 128, 128, 0, h->stream>>>(E);
    CU(h, cudaGetLastError());
    if (out->probabilities) CU(h, cudaMemcpyAsync(out->probabilities, S + ES_OFF_PROB, pairs * 8, cudaMemcpyDeviceToHost, h->stream));
    if (out->call) CU(h, cudaMemcpyAsync(out->call, S + ES_OFF_CALL, n, cudaMemcpyDeviceToHost, h->stream));
    if (out->confidence) CU(h, cudaMemcpyAsync(out->confidence, S + ES_OFF_CONF, n, cudaMemcpyDeviceToHost, h->stream));
    CU(h, cudaStreamSynchronize(h->stream));
    return 0;
}
