// charon_b200/csrc/eval_kernel.cuh — model evaluation + call logic (sm_100a).
// Replaces ReadEntry::get_proportions / apply_model / call_host / call_category
// (reference include/read_entry.hpp:140-150,157-291) and Model::prob / KDEParams::prob
// (include/classify_stats.hpp:242-252,370-389).
//
// Arithmetic contract: proportions are float divisions of the integer counts; the KDE is a
// *sequential float* sum in dataset order of float(exp(-z^2/2 in double)/sqrt(2 pi)); terms with |z| > 15 are
// exactly +0.0f after the float rounding and are skipped (adding +0.0f never changes the running sum).
// No fused multiply-add may be formed across the reference's separate float operations: explicit _rn intrinsics.
#pragma once
#include <cuda_runtime.h>
#include <math_constants.h>
#include <stdint.h>

#include "kde_exp_table.inc"

namespace cgk {

struct ModelDev {            // one per category, device-resident
    int32_t dist;            // 0 kde, 1 beta, 2 gamma
    uint32_t pos_off, npos;  // offsets into the shared float pool
    uint32_t neg_off, nneg;
    float h_pos, h_neg;
    float beta_pos[2], beta_neg[2];
    float gamma_pos[3], gamma_neg[3];
    uint32_t pos_sorted, neg_sorted;   // dataset ascending (the KDEParams ctor sorts; a fitted dataset keeps training order)
};

struct EvalParams {
    uint32_t n_reads, ncat;
    int32_t mode;            // 0 dehost (call_host), 1 classify (call_category)
    const ModelDev *models;
    const float *pool;
    float log_rate;          // logf(300.f) computed on the host exactly as the reference's float log
    // thresholds
    float min_quality; uint32_t min_length; float min_compression; int32_t confidence_threshold;
    float confidence_probability_threshold; uint32_t min_hits; float host_unique_prop_lo_threshold;
    float min_proportion_difference, min_prob_difference; uint32_t host_index;
    // inputs
    const uint32_t *num_hashes, *counts, *unique_counts;
    const float *mean_quality, *compression; const uint32_t *length;
    // outputs
    double *prob; uint8_t *call; uint8_t *confidence;
    // Memoised evaluation (memo_keys != NULL): the probability of a (read, category) pair depends only on (category,
    // unique count, num_hashes); a batch of 131 072 reads holds about 3x fewer distinct triples than pairs, and a KDE is 500+
    // sequential terms.  memo_prepare_kernel files every pair under its triple in an open-addressing table (the first
    // arrival appends the slot to memo_list), memo_unique_kernel evaluates each distinct triple once, call_kernel copies the
    // value into prob[].  Which thread arrives first does not matter: the value is a pure function of the key.
    unsigned long long *memo_keys; double *memo_vals; uint32_t *memo_list, *memo_count, *memo_slot; uint32_t memo_mask;
};
constexpr unsigned long long MEMO_EMPTY = ~0ull;     // a key's top byte is the category (<= 254)
constexpr uint32_t MEMO_NONE = 0xFFFFFFFFu;
constexpr uint32_t MEMO_FIELD_BITS = 28;             // unique count and num_hashes each below 2^28

// One KDE term: float(exp(-z^2 / 2) / sqrt(2 pi)) exactly as the reference rounds it, without libm's exp and without the
// double division (kde_fast_exp.inc; proven identical on every float of [-15, 15] by tests/kde_exp_exhaustive.c).
__constant__ double c_exp_t64[64] = CG_EXP_T64_INIT;
#define CG_KDE_FN __device__ __forceinline__
#define CG_KDE_TPARAM , const double *__restrict__ cg_t64
#define CG_KDE_TABLE(j) cg_t64[j]
#define CG_KDE_MUL(a, b) __dmul_rn(a, b)
#define CG_KDE_ADD(a, b) __dadd_rn(a, b)
#define CG_KDE_BITS(d) ((unsigned long long)__double_as_longlong(d))
#define CG_KDE_FROM_BITS(b) __longlong_as_double((long long)(b))
#include "kde_fast_exp.inc"

__device__ __forceinline__ float kde_z(const float *__restrict__ ds, uint32_t i, float h, float x) { return __fdiv_rn(__fsub_rn(x, ds[i]), h); }

// `sorted`: ds ascending, so z_i = (x - ds[i]) / h is non-increasing in i (IEEE subtraction and division by h > 0 are monotone)
// and the terms that survive the |z| <= 15 test form one contiguous range, found by two binary searches with the very same
// float expression; everything outside contributes exactly +0.0f.  The range is then summed in dataset order as before.
__device__ __forceinline__ float kde_prob(const float *__restrict__ ds, uint32_t n, float h, float x, bool sorted, const double *__restrict__ t64) {
    float total = 0.0f;
    uint32_t i0 = 0, i1 = n;
    if (sorted && !isnan(x) && h > 0.0f) {
        uint32_t lo = 0, hi = n;                 // first i with z_i <= 15
        while (lo < hi) { const uint32_t mid = (lo + hi) >> 1; if (kde_z(ds, mid, h, x) > 15.0f) lo = mid + 1; else hi = mid; }
        i0 = lo; hi = n;                         // first i with z_i < -15
        while (lo < hi) { const uint32_t mid = (lo + hi) >> 1; if (kde_z(ds, mid, h, x) >= -15.0f) lo = mid + 1; else hi = mid; }
        i1 = lo;
    }
    // Branch-free and unrolled: the terms are independent (≈25 dependent operations each), only the float additions form a
    // chain, so four terms in flight hide most of a term's latency.  A term with |z| > 15 is exactly +0.0f; adding +0.0f
    // leaves the running sum unchanged bit for bit (the sum starts at +0.0f and never turns into -0.0f: every term is >= +0
    // or NaN), so "skip" and "add zero" are the same thing.  exp(NaN) = NaN poisons the sum, as in the reference.
#pragma unroll 4
    for (uint32_t i = i0; i < i1; ++i) {
        const float z = kde_z(ds, i, h, x);
        const float t = kde_term(z, t64);                                // garbage outside [-15, 15] (never out of bounds): replaced below
        total = __fadd_rn(total, fabsf(z) > 15.0f ? 0.0f : (z != z ? z : t));
    }
    return __fdiv_rn(total, __fmul_rn(h, float(n)));                    // NaN z falls through: exp(NaN) = NaN
}

__device__ __forceinline__ float dexp300(float x, float log_rate) {     // stats::dexp(float, 300)
    if (isnan(x)) return CUDART_NAN_F;
    if (x < 0.0f) return 0.0f;
    const float a = __fsub_rn(log_rate, __fmul_rn(300.0f, x));
    return float(exp(double(a)));                                        // correctly rounded float exp
}
__device__ __forceinline__ float dbeta_dev(float x, float a, float b) {
    if (isnan(x)) return CUDART_NAN_F;
    if (x < 0.0f || x > 1.0f) return 0.0f;
    const float lbeta = __fsub_rn(__fadd_rn(lgammaf(a), lgammaf(b)), lgammaf(__fadd_rn(a, b)));
    const float t = __fadd_rn(__fadd_rn(-lbeta, __fmul_rn(__fsub_rn(a, 1.0f), logf(x))), __fmul_rn(__fsub_rn(b, 1.0f), logf(__fsub_rn(1.0f, x))));
    return expf(t);
}
__device__ __forceinline__ float dgamma_dev(float x, float shape, float scale) {
    if (isnan(x)) return CUDART_NAN_F;
    if (x < 0.0f) return 0.0f;
    float t = __fsub_rn(-lgammaf(shape), __fmul_rn(shape, logf(scale)));
    t = __fadd_rn(t, __fmul_rn(__fsub_rn(shape, 1.0f), logf(x)));
    t = __fsub_rn(t, __fdiv_rn(x, scale));
    return expf(t);
}

// probabilities_[i] = 1.0 * Model::prob(unique_proportions_[i]).pos for category c with the given counts
__device__ __forceinline__ double pair_prob(const EvalParams &P, uint32_t c, uint32_t unique, uint32_t num_hashes, const double *s_t64) {
    const float x = __fdiv_rn(float(unique), float(num_hashes));   // read_entry.hpp:145
    const ModelDev m = P.models[c];
    const float p_err = dexp300(x, P.log_rate);
    float p_pos, p_neg;
    if (m.dist == 0) {
        p_pos = kde_prob(P.pool + m.pos_off, m.npos, m.h_pos, x, m.pos_sorted != 0, s_t64);
        p_neg = kde_prob(P.pool + m.neg_off, m.nneg, m.h_neg, x, m.neg_sorted != 0, s_t64);
    } else if (m.dist == 2) {
        p_pos = dgamma_dev(__fsub_rn(x, m.gamma_pos[1]), m.gamma_pos[0], m.gamma_pos[2]);
        p_neg = dgamma_dev(__fsub_rn(x, m.gamma_neg[1]), m.gamma_neg[0], m.gamma_neg[2]);
    } else {
        p_pos = dbeta_dev(x, m.beta_pos[0], m.beta_pos[1]);
        p_neg = dbeta_dev(x, m.beta_neg[0], m.beta_neg[1]);
    }
    if (x == 1.0f) p_pos = 1.0f;
    const float total = __fadd_rn(__fadd_rn(p_err, p_pos), p_neg);
    return double(__fdiv_rn(p_pos, total));
}

// one thread per (read, category)
__global__ void prob_kernel(const EvalParams P) {
    __shared__ double s_t64[64];   // 2^(j/64): per-lane indices differ, so shared memory, not the constant cache
    if (threadIdx.x < 64) s_t64[threadIdx.x] = c_exp_t64[threadIdx.x];
    __syncthreads();
    const size_t idx = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (idx >= size_t(P.n_reads) * P.ncat) return;
    const uint32_t r = uint32_t(idx / P.ncat), c = uint32_t(idx % P.ncat);
    P.prob[idx] = pair_prob(P, c, P.unique_counts[idx], P.num_hashes[r], s_t64);
}

// memoised variant, step 1: one thread per (read, category).  Pairs of a KDE category are filed under their triple; the cheap
// closed-form densities (beta, gamma) and counts beyond the key's fields are evaluated on the spot.
__global__ void memo_prepare_kernel(const EvalParams P) {
    __shared__ double s_t64[64];
    if (threadIdx.x < 64) s_t64[threadIdx.x] = c_exp_t64[threadIdx.x];
    __syncthreads();
    const size_t idx = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (idx >= size_t(P.n_reads) * P.ncat) return;
    const uint32_t r = uint32_t(idx / P.ncat), c = uint32_t(idx % P.ncat);
    const uint32_t u = P.unique_counts[idx], nh = P.num_hashes[r];
    if (P.models[c].dist != 0 || (u >> MEMO_FIELD_BITS) || (nh >> MEMO_FIELD_BITS)) {
        P.prob[idx] = pair_prob(P, c, u, nh, s_t64);
        P.memo_slot[idx] = MEMO_NONE;
        return;
    }
    const unsigned long long key = ((unsigned long long)c << 56) | ((unsigned long long)u << MEMO_FIELD_BITS) | nh;
    unsigned long long x = key * 0x9E3779B97F4A7C15ull;
    uint32_t h = uint32_t(x >> 32) & P.memo_mask;
    for (;;) {
        const unsigned long long old = atomicCAS(&P.memo_keys[h], MEMO_EMPTY, key);
        if (old == MEMO_EMPTY) { P.memo_list[atomicAdd(P.memo_count, 1u)] = h; break; }
        if (old == key) break;
        h = (h + 1u) & P.memo_mask;
    }
    P.memo_slot[idx] = h;
}
// step 2: one thread per DISTINCT triple (the grid covers the worst case; threads beyond the count leave)
__global__ void memo_unique_kernel(const EvalParams P) {
    __shared__ double s_t64[64];
    if (threadIdx.x < 64) s_t64[threadIdx.x] = c_exp_t64[threadIdx.x];
    __syncthreads();
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= *P.memo_count) return;
    const uint32_t h = P.memo_list[t];
    const unsigned long long key = P.memo_keys[h];
    const uint32_t field = (1u << MEMO_FIELD_BITS) - 1u;
    P.memo_vals[h] = pair_prob(P, uint32_t(key >> 56), uint32_t(key >> MEMO_FIELD_BITS) & field, uint32_t(key) & field, s_t64);
}

// one thread per read: call_host (mode 0) / call_category (mode 1)
__global__ void call_kernel(const EvalParams P) {
    const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= P.n_reads) return;
    const uint32_t nc = P.ncat;
    if (P.memo_keys)     // step 3 of the memoised evaluation: the pair's probability is its triple's
        for (uint32_t c = 0; c < nc; ++c) {
            const uint32_t slot = P.memo_slot[size_t(r) * nc + c];
            if (slot != MEMO_NONE) P.prob[size_t(r) * nc + c] = P.memo_vals[slot];
        }
    const uint32_t *uc = P.unique_counts + size_t(r) * nc;
    const uint32_t *ct = P.counts + size_t(r) * nc;
    const double *pr = P.prob + size_t(r) * nc;
    const float nh = float(P.num_hashes[r]);
    const float mq = P.mean_quality[r], cr = P.compression[r];
    const uint32_t len = P.length[r];
    uint8_t call = 255, conf = 0;
    const int thr = int(int8_t(P.confidence_threshold));
    if (P.mode == 0) {   // read_entry.hpp:218-269
        const uint32_t host = P.host_index, other = 1u - host;
        const double hu = double(__fdiv_rn(float(uc[host]), nh)), ou = double(__fdiv_rn(float(uc[other]), nh));
        const double hp = pr[host], op = pr[other];
        uint32_t first = host, second = other;
        if (hu < ou) { first = other; second = host; }
        const uint32_t raw = uc[first] - uc[second];   // unsigned wrap like the reference when NaN ordering misleads
        conf = raw > 255u ? 255 : uint8_t(raw);
        bool gate = !(int(conf) < thr) && !(mq < P.min_quality) && !(len < P.min_length) && !(cr < P.min_compression);
        if (gate) {
            const double dc = double(conf);
            if (hu > ou && hu - ou > double(P.min_proportion_difference) && hp > op && hp - op > double(P.min_prob_difference) &&
                fmax(hp * dc, dc) >= double(P.confidence_probability_threshold))
                call = uint8_t(host);
            else if (hu < double(P.host_unique_prop_lo_threshold) && hu < ou && ou - hu > double(P.min_proportion_difference) && hp < op &&
                     op - hp > double(P.min_prob_difference) && fmax(op * dc, dc) >= double(P.confidence_probability_threshold))
                call = uint8_t(other);
        }
    } else {             // read_entry.hpp:157-216
        uint32_t first = 0, second = 1;
        if (uc[second] > uc[first]) { first = 1; second = 0; }
        for (uint32_t i = 2; i < nc; ++i)
            if (uc[i] > uc[second]) {
                second = i;
                if (uc[second] > uc[first]) { const uint32_t t = first; first = second; second = t; }
            }
        const uint32_t raw = uc[first] - uc[second];
        conf = raw > 255u ? 255 : uint8_t(raw);
        if (!(mq < P.min_quality) && !(len < P.min_length) && !(cr < P.min_compression)) {
            if (pr[second] == 0.0 && pr[first] > 0.0) call = uint8_t(first);
            else if (int(conf) > thr && pr[first] > pr[second]) call = uint8_t(first);
            const uint32_t fc = ct[first], sc = ct[second];
            if (sc > fc || fc - sc < P.min_hits) call = 255;
            const float fp = __fdiv_rn(float(fc), nh), sp = __fdiv_rn(float(sc), nh);
            if (sp > fp || __fsub_rn(fp, sp) < P.min_proportion_difference) call = 255;
        }
    }
    P.call[r] = call;
    P.confidence[r] = conf;
}

}  // namespace cgk
