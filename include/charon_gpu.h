/* =============================================================================
 * include/charon_gpu.h — C-ABI of libcharon_gpu.so (B200 / sm_100a)
 *
 * The reference (rmcolq/charon) has no plugin/FFI seam: its hot path is the body of
 * the OpenMP loop in dehost_reads / classify_reads.  This header is the seam a
 * maintainer would cut there (see INTEGRATION.md for the call-site patch).  Each
 * entry point cites the reference code it replaces (paths relative to the reference
 * root).  Plain pointers and sizes only; no C++/torch types cross the boundary.
 *
 * Conventions
 *   - every function returns 0 on success, a negative cg_status on error;
 *     cg_last_error(h) (or cg_last_error(NULL) for cg_create failures) returns a
 *     NUL-terminated description valid until the next call on the same thread.
 *   - one handle per GPU; calls on one handle must be serialised by the caller,
 *     different handles may be driven from different host threads.
 *   - host buffers are caller-owned (pinned memory recommended); device memory and
 *     streams are library-owned.  Outputs are indexed by input read position.
 *   - there is NO CPU fallback: without a usable sm_100 device cg_create fails.
 * ============================================================================= */
#ifndef CHARON_GPU_H
#define CHARON_GPU_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct cg_handle cg_handle;

enum cg_status {
    CG_OK = 0,
    CG_ERR_ARG = -1,      /* invalid argument / unsupported parameter (k > 26, bins > 256, ...) */
    CG_ERR_CUDA = -2,     /* CUDA runtime error (message has the cudaError string) */
    CG_ERR_NO_DEVICE = -3,/* no CUDA device / not an sm_100 part: the product has no CPU path */
    CG_ERR_FORMAT = -4,   /* index payload fails a structural invariant */
    CG_ERR_STATE = -5,    /* call order violated (e.g. eval before set_model) */
    CG_ERR_NOMEM = -6     /* filter does not fit the given memory budget in replicated mode */
};

/* How the Interleaved Bloom Filter payload is handed over (include/index.hpp:26,122-138). */
enum cg_filter_format {
    CG_FILTER_PLAIN_ROWS = 0, /* host bit matrix, row-major, bin_words u64 per row (seqan3 uncompressed layout) */
    CG_FILTER_SD_VECTOR = 1   /* sdsl::sd_vector parts exactly as stored in the .idx file; decoded on the GPU */
};

/* Index description produced by the host loader (replaces Index + ibf_ members,
 * include/index.hpp:21-26 ; IBF header fields as serialised by seqan3). */
typedef struct cg_index_desc {
    uint8_t kmer_size;       /* include/index.hpp:22 */
    uint8_t window_size;     /* include/index.hpp:21 */
    uint64_t bins, technical_bins, bin_size, hash_shift, bin_words, hash_funs;
    int32_t filter_format;   /* enum cg_filter_format */
    const uint64_t *rows;    /* PLAIN_ROWS: bin_size * bin_words words */
    const uint64_t *sd_low;  /* SD_VECTOR: m_low words */
    uint64_t sd_low_bits;
    const uint64_t *sd_high; /* SD_VECTOR: m_high words */
    uint64_t sd_high_bits;
    uint8_t sd_wl;
    uint16_t ncat;             /* InputSummary::num_categories(), include/input_summary.hpp:35-37 */
    const uint8_t *bin_to_cat; /* bins entries: category index of each bin (input_summary.hpp:19-20,39-45) */
} cg_index_desc;

enum cg_dist { CG_DIST_KDE = 0, CG_DIST_BETA = 1, CG_DIST_GAMMA = 2 };

/* One Model (include/classify_stats.hpp:261-393).  For KDE the datasets are given in the exact
 * order the reference sums them (KDEParams ctor sorts ascending, fit() keeps training order). */
typedef struct cg_category_model {
    int32_t dist;                 /* enum cg_dist */
    const float *pos; uint32_t npos; float h_pos;   /* k_pos: dataset, bandwidth (:270) */
    const float *neg; uint32_t nneg; float h_neg;   /* k_neg (:271) */
    float beta_pos[2], beta_neg[2];                 /* alpha, beta (:268-269) */
    float gamma_pos[3], gamma_neg[3];               /* shape, loc, scale (:266-267) */
} cg_category_model;

/* StatsModel thresholds (include/classify_stats.hpp:399-409). */
typedef struct cg_thresholds {
    float min_quality;
    uint32_t min_length;
    float min_compression;
    int8_t confidence_threshold;
    float confidence_probability_threshold;
    uint8_t min_hits;
    float host_unique_prop_lo_threshold;
    float min_proportion_difference;
    float min_prob_difference;
    uint8_t host_index;           /* InputSummary::host_category_index(), input_summary.hpp:47-55 */
} cg_thresholds;

enum cg_mode { CG_MODE_DEHOST = 0 /* ReadEntry::call_host */, CG_MODE_CLASSIFY = 1 /* ReadEntry::call_category */ };

/* How the bases of a batch are encoded.  The reference's parser hands the loop `std::vector<seqan3::dna5>` records
 * (include/utils.hpp:17-19, src/dehost_main.cpp:335-339): one rank per base.  ASCII is accepted for convenience; the two
 * packed formats are what a parser written for this library emits directly (host/fastx.hpp does), so that nothing has to
 * be re-encoded between the file and the PCIe copy.  Offsets (seg_off, seg_len, exc_pos) always count BASES. */
enum cg_bases_format {
    CG_BASES_ASCII = 0,  /* one character per base, any case, IUPAC; U -> T, anything but ACGTU -> N */
    CG_BASES_RANK4 = 1,  /* dna5 ranks (A0 C1 G2 N3 T4), 4 bits per base: base i = nibble (i & 1) of byte i >> 1 (cg_pack_bases) */
    CG_BASES_2BIT = 2    /* A0 C1 G2 T3, 2 bits per base: base i = bits 2*(i & 3).. of byte i >> 2; every base that is not
                            A/C/G/T/U is stored as 0 and listed in exc_pos (it hashes as N, like in the other formats)
                            (cg_pack_bases_2bit).  The buffer must be readable up to the next multiple of 16 bytes. */
};

/* A batch of reads.  Segment s holds bases [seg_off[s], seg_off[s] + len(s)) of the batch, len(s) = seg_len[s] when
 * seg_len is given, else seg_off[s + 1] - seg_off[s] (back-to-back segments).  With seg_len a producer may start every
 * segment where it likes (e.g. on a 64-base boundary so that reads are packed independently of each other): the bases
 * between the end of a segment and seg_off[s + 1] are padding and are never looked at.  Read r owns segments
 * read_seg[r] .. read_seg[r+1] (1 for single-end, 2 for paired: src/dehost_main.cpp:458-465); read_seg may be NULL: one
 * segment per read.  A zero-initialised tail (bases_format .. n_exc = 0) means back-to-back ASCII, as in round 1. */
typedef struct cg_batch {
    const char *bases;
    const uint64_t *seg_off;   /* n_segs + 1 entries, non-decreasing, seg_off[0] = 0; seg_off[n_segs] = bases in the buffer */
    const uint32_t *read_seg;  /* n_reads + 1 entries or NULL */
    uint32_t n_reads;
    uint32_t n_segs;
    int32_t bases_format;      /* enum cg_bases_format */
    const uint64_t *seg_len;   /* n_segs entries or NULL */
    const uint64_t *exc_pos;   /* CG_BASES_2BIT: ascending batch positions of the bases that are not A/C/G/T/U; NULL when n_exc = 0 */
    uint64_t n_exc;
} cg_batch;

/* Per-read integer outputs of ReadEntry::get_counts (include/read_entry.hpp:92-138). */
typedef struct cg_counts_out {
    uint32_t *num_hashes;      /* [n_reads]            read_entry.hpp:86-90 */
    uint32_t *counts;          /* [n_reads * ncat]     read_entry.hpp:102-114 */
    uint32_t *unique_counts;   /* [n_reads * ncat]     read_entry.hpp:122-136 */
} cg_counts_out;

/* Per-read host-side scalars feeding the call logic (src/dehost_main.cpp:345-363). */
/* compression may be NULL: the library then computes it on the device (see cg_compression_batch) */
typedef struct cg_read_scalars {
    const float *mean_quality;
    const float *compression;
    const uint32_t *length;
} cg_read_scalars;

/* Per-read outputs of apply_model + call_host / call_category (read_entry.hpp:157-291). */
typedef struct cg_calls_out {
    double *probabilities;     /* [n_reads * ncat] */
    uint8_t *call;             /* [n_reads], 255 = unclassified */
    uint8_t *confidence;       /* [n_reads] */
} cg_calls_out;

/* Reference sequences for building an index on the device (the `charon index` data path:
 * src/index_main.cpp:118-160 count_and_store_hashes, :238-263 build_index, src/utils.cpp:75-90 bin_size_in_bits). */
typedef struct cg_build_desc {
    uint8_t kmer_size, window_size;  /* include/index_arguments.hpp:16-17 (41 / 19) */
    uint32_t hash_funs;              /* include/index_arguments.hpp:21 (3) */
    double max_fpr;                  /* include/index_arguments.hpp:22 (0.01) */
    uint64_t max_bits;               /* include/index_arguments.hpp:20 (2^32 - 3); 0 = that default */
    const char *bases;               /* concatenated records (raw ASCII) */
    const uint64_t *rec_off;         /* n_records + 1 */
    const uint32_t *rec_bin;         /* bin of each record */
    uint32_t n_records;
    uint32_t bins;
    uint16_t ncat;
    const uint8_t *bin_to_cat;
    uint64_t force_max_hashes;       /* 0: size the filter for the largest bin's distinct hash count (src/index_main.cpp:239-240);
                                        else use this count (--optimize sums per-file counts of merged bins, :192-201) */
} cg_build_desc;

/* ---- lifetime -------------------------------------------------------------------------- */
/* Replaces load_index' in-memory IBF + Index::agent() (src/load_index.cpp:8-15, include/index.hpp:110-112):
 * uploads the filter, re-lays it out column-compacted in HBM.  mem_budget_bytes = 0: whole device. */
int cg_create(int device, const cg_index_desc *desc, size_t mem_budget_bytes, cg_handle **out);
/* Builds the filter ON the GPU from reference sequences: minimiser hashing with the classify kernels, distinct
 * hashes per bin (sort + unique), bin_size from the largest bin, atomic-OR insertion.  hashes_per_bin[bins] receives
 * InputStats::hashes_per_bin (include/input_stats.hpp:19).  The handle is then ready for classification; the filter
 * can be exported with cg_debug_dump_filter to write an .idx file. */
int cg_build_index(int device, const cg_build_desc *desc, size_t mem_budget_bytes, uint64_t *hashes_per_bin, cg_handle **out);
/* ---- filter partitioned by row (bit-address) range across GPUs (BASELINE config 4) ----------------
 * For an index whose filter must not be replicated (memory budget), GPU `part_rank` of `part_world` stores rows
 * [part_rank * rpp, (part_rank + 1) * rpp) only (rpp = ceil(bin_size / part_world) rounded up to 64).  Reads stay
 * sharded across GPUs; the count kernel fetches rows owned by other GPUs through NVLink/NVSwitch peer mappings of
 * their partitions, so bulk_contains (include/index.hpp:110-112) keeps its exact AND-of-rows semantics and
 * unique_counts (read_entry.hpp:122-136) need no per-minimiser exchange.  All parts must be attached before the
 * first batch.  Exchange of the 64-byte handles between processes is the caller's job (torch.distributed, MPI, ...). */
int cg_create_partitioned(int device, const cg_index_desc *desc, uint32_t part_rank, uint32_t part_world, cg_handle **out);
/* Turn a replicated handle (cg_create / cg_build_index) into part `part_rank` of `part_world`: frees the other rows. */
int cg_partition_from_replicated(cg_handle *h, uint32_t part_rank, uint32_t part_world);
/* cudaIpcMemHandle_t (64 bytes) of this GPU's partition, for a peer PROCESS. */
int cg_partition_export(cg_handle *h, void *ipc_handle_64);
int cg_partition_attach_ipc(cg_handle *h, uint32_t peer_part, const void *ipc_handle_64);
/* Peer handle living in the SAME process (one host thread per GPU): enables peer access and uses its pointer. */
int cg_partition_attach_peer(cg_handle *h, uint32_t peer_part, cg_handle *peer);

/* `charon index` output path (include/store_index.hpp:12-17, include/index.hpp:50): the filter Elias-Fano encoded ON the
 * device into the two arrays sdsl::sd_vector serialises (m_low: low_bits/wl values of wl bits; m_high: high_bits bits).
 * First call with low_out = high_out = NULL for the sizes, then with buffers of ceil(bits/64) words each. */
int cg_export_sd_vector(cg_handle *h, uint64_t *size_bits, uint8_t *wl, uint64_t *low_bits, uint64_t *high_bits,
                        uint64_t *low_out, uint64_t *high_out);

void cg_destroy(cg_handle *h);
const char *cg_last_error(const cg_handle *h);
/* Use an externally created cudaStream_t (e.g. torch's current stream) for all work; NULL = own stream. */
int cg_set_stream(cg_handle *h, void *cuda_stream);
int cg_sync(cg_handle *h);

/* Replaces StatsModel construction / Model::train results (include/classify_stats.hpp:428-473). */
int cg_set_model(cg_handle *h, const cg_category_model *per_category, const cg_thresholds *thr, int mode);

/* ---- hot path, host buffers in / host buffers out -------------------------------------- */
/* loop body src/dehost_main.cpp:366-373: minimiser_hash -> bulk_contains -> update_entry -> post_process */
int cg_count_batch(cg_handle *h, const cg_batch *batch, const cg_counts_out *out);
/* Result::classify_read -> ReadEntry::dehost / classify (include/result.hpp:97-102) on counts already known
 * (needed when the host trains the model between counting and calling, result.hpp:136-152). */
int cg_eval_batch(cg_handle *h, uint32_t n_reads, const uint32_t *num_hashes, const uint32_t *counts,
                  const uint32_t *unique_counts, const cg_read_scalars *scalars, const cg_calls_out *out);
/* count + eval fused (model fixed up front: the default no---extract invocation). */
int cg_classify_batch(cg_handle *h, const cg_batch *batch, const cg_read_scalars *scalars,
                      const cg_counts_out *counts_out, const cg_calls_out *calls_out);
/* cg_eval_batch for callers that may not allocate.  NO counterpart in the reference: added
 * (code under CUDA graph capture, tight serving loops and harnesses that forbid cudaMalloc).  Same arguments and
 * results as cg_eval_batch, but the call makes no cudaMalloc / cudaFree: the probabilities the model kernel hands to the
 * call kernel, and the device copies of inputs and outputs, live in one static device array of CG_EVAL_STATIC_BYTES
 * (1 343 488 bytes, held by every GPU that loads the library until the process exits).  Limits per call: n_reads <=
 * CG_EVAL_STATIC_MAX_READS and n_reads * categories <= CG_EVAL_STATIC_MAX_PAIRS; beyond them CG_ERR_ARG with the limits in
 * cg_last_error.  The array is shared by all handles of a device and keeps the last call's contents (this call's output
 * range is cleared first): calls that overlap on the GPU — two streams, two threads — race; serialise them.  The first
 * call on a device resolves the array's device address (cudaGetSymbolAddress): call once with n_reads == 0 at set-up. */
#define CG_EVAL_STATIC_MAX_READS 16384u
#define CG_EVAL_STATIC_MAX_PAIRS 65536u
#define CG_EVAL_STATIC_BYTES (CG_EVAL_STATIC_MAX_READS * 18u + CG_EVAL_STATIC_MAX_PAIRS * 16u)
int cg_eval_batch_static(cg_handle *h, uint32_t n_reads, const uint32_t *num_hashes, const uint32_t *counts,
                         const uint32_t *unique_counts, const cg_read_scalars *scalars, const cg_calls_out *out);

/* ---- per-read compression ratio on the device ------------------------------------------------------------------------
 * get_compression_ratio (reference src/utils.cpp:114-124, called at src/dehost_main.cpp:359-361 / 449-452): size of
 * gzip::compress(uppercase dna5 string of the read; both mates of a pair concatenated) = zlib level 6, gzip wrapper, memLevel 8,
 * divided by the length.  The size is computed bit-exactly without writing the stream (charon_b200/csrc/deflate_size.cuh).
 * gzip_size_out[n_reads] (may be NULL): bytes; ratio_out[n_reads] (may be NULL): float(double(size) / double(length)) as the
 * reference forms it.  Reads longer than CG_GZIP_MAX_LEN bases (16 M: positions are kept in 24 bits) get size 0xFFFFFFFF / ratio
 * NaN; compute those on the host.
 * The same kernel runs inside cg_classify_batch / cg_submit_batch / cg_stage_batch+cg_run_staged when
 * cg_read_scalars.compression is NULL; there a batch that holds a read longer than CG_GZIP_MAX_LEN is refused (CG_ERR_ARG),
 * and cg_fetch_compression copies the ratios of the slot's last batch back for the TSV row (after cg_wait_batch). */
#define CG_GZIP_MAX_LEN 16777215u
int cg_compression_batch(cg_handle *h, const cg_batch *batch, uint32_t *gzip_size_out, float *ratio_out);
int cg_fetch_compression(cg_handle *h, int slot, float *ratio_out);

/* ---- two-slot pipeline over the same path: cg_submit_batch enqueues H2D + kernels + D2H of one batch on the slot's
 * own stream and returns; cg_wait_batch blocks until that slot's outputs are in the caller's buffers.  With slots 0/1
 * used alternately the PCIe copy of batch i+1 overlaps the kernels of batch i (the chunk loop of
 * src/dehost_main.cpp:335-381 software-pipelined).  Host buffers must stay valid (and should be pinned) until the wait.
 * scalars == NULL: counts only (no model evaluation). */
int cg_submit_batch(cg_handle *h, int slot, const cg_batch *batch, const cg_read_scalars *scalars,
                    const cg_counts_out *counts_out, const cg_calls_out *calls_out);
int cg_wait_batch(cg_handle *h, int slot);
/* Make `slot` the one the staged calls below (and the synchronous calls above) operate on. */
int cg_select_slot(cg_handle *h, int slot);

/* ---- staged (device-resident) variant used to time the kernels without PCIe -------------- */
int cg_stage_batch(cg_handle *h, const cg_batch *batch, const cg_read_scalars *scalars);  /* H2D only */
int cg_run_staged(cg_handle *h, int with_eval);                                            /* kernels only, async */
int cg_fetch_staged(cg_handle *h, const cg_counts_out *counts_out, const cg_calls_out *calls_out); /* D2H + sync */

/* ---- pinned host memory + device-side timing (CUDA events on the handle's stream) ---------- */
/* Transport of the bases.  The batch a caller hands over is ASCII, one base per byte (what the reference's parser yields,
 * src/dehost_main.cpp:335-339); at 8 kb per read that makes PCIe the end-to-end bound.  When the k=19 / w=41 kernels
 * apply, the library can pack the bases to 4-bit dna5 ranks on a host thread pool, chunk by chunk, and copy the packed
 * chunks while the next ones are being packed; the kernels expand a nibble to the very code they derive from ASCII.
 * mode: -1 automatic (pool of >= 8 threads), 0 never, 1 always.  threads > 0 resizes the process-wide pool
 * (default: env CHARON_HOST_THREADS, else min(16, hardware threads)).  Env CHARON_HOST_PACK presets the mode of new handles. */
int cg_set_host_pack(cg_handle *h, int mode, int threads);
/* packed[i >> 1] nibble (i & 1) = dna5 rank of bases[i] (A0 C1 G2 N3 T4; U -> T, anything else -> N); (n + 1) / 2 bytes. Host only. */
int cg_pack_bases(const char *bases, uint64_t n, uint8_t *packed);
/* CG_BASES_2BIT writer for ONE run of bases (host only, single-threaded, SIMD; call it from as many threads as there are
 * readers).  packed receives (n + 3) / 4 bytes: base i of the run = bits 2*(i & 3) of packed[i >> 2], so the run must start
 * on a byte boundary of the batch buffer (pos0 % 4 == 0; pos0 = batch position of the run's first base).  Positions
 * (pos0 + i) of the bases that are not A/C/G/T/U are appended to exc_pos (capacity exc_cap entries) in ascending order;
 * *n_exc receives how many there were — when it exceeds exc_cap only the first exc_cap were stored.  This is the char ->
 * dna5 conversion of the reference's parser (include/utils.hpp:17-19) fused with the packing. */
int cg_pack_bases_2bit(const char *bases, uint64_t n, uint8_t *packed, uint64_t pos0, uint64_t *exc_pos, uint64_t exc_cap, uint64_t *n_exc);

int cg_host_alloc(size_t bytes, void **out);   /* cudaHostAlloc: pinned buffers for cg_batch / outputs */
void cg_host_free(void *p);
/* Everything enqueued on the handle between start and stop is timed with one event pair; additionally every
 * cg_run_staged / cg_classify_batch in between records per-launch events so the count kernel's own duration is known. */
int cg_timer_start(cg_handle *h);
int cg_timer_stop(cg_handle *h, float *total_ms, float *count_kernel_ms_sum, float *eval_kernel_ms_sum, uint32_t *n_runs);

/* SURVEY 8(d) denominator: random 32-byte-sector gather rate on this handle's filter, measured with nothing else in the
 * kernel.  mode 0: the query's own addressing (hash_funs rows per value, i.e. 96 B algorithmic per value for 3 hashes);
 * mode 1: one random 32-bit load per value (32 B sector each).  avg_ms = mean of `reps` launches after a warm-up. */
int cg_bench_gather(cg_handle *h, uint64_t n_values, int mode, int reps, float *avg_ms);

/* ---- introspection ----------------------------------------------------------------------- */
typedef struct cg_info {
    uint64_t bin_size;            /* IBF rows (bits per bin) */
    uint64_t filter_bytes;        /* compacted filter size in HBM */
    uint32_t row_bits;            /* compacted row width (power of two >= bins) */
    uint32_t single_pass;         /* 1 when every category owns exactly one bin (best bin static) */
    uint32_t sm_count;
    uint32_t count_grid, count_block, count_smem_bytes;
    uint64_t last_minimisers;     /* sum of num_hashes over the last counted batch */
    uint64_t last_bases;
    uint32_t last_kernel_launches;
    float last_count_kernel_ms;   /* CUDA-event time of the count kernel of the last batch */
    float last_eval_kernel_ms;
    float last_h2d_ms, last_d2h_ms;
    uint64_t last_h2d_bytes;      /* bytes the last staged batch moved host -> device (bases as sent + offsets + scalars) */
    float last_pack_ms;           /* host wall time of the 4-bit packing (its chunk copies overlap it) */
    uint32_t last_packed;         /* 1 when the last batch went over the link as 4-bit ranks */
    float last_gzip_kernel_ms;    /* CUDA-event time of the compression-ratio kernel of the last batch (0: ratios came from the host) */
} cg_info;
int cg_get_info(cg_handle *h, cg_info *out);

/* Test hooks (stage-level parity): emitted minimiser values of one sequence in emission order is not
 * defined on the GPU (compaction is unordered) — values are returned sorted ascending. */
int cg_debug_minimisers(cg_handle *h, const char *bases, uint64_t len, uint64_t *out, uint64_t cap, uint64_t *n_out);
/* bulk_contains for n hash values: masks[n * bin_words] in the seqan3 binning_bitvector word layout. */
int cg_debug_query(cg_handle *h, const uint64_t *values, uint64_t n, uint64_t *masks);
/* Run the generic any-(k,w) kernel even where the (k=19,w=41) specialisation applies (both are parity-tested). */
int cg_debug_force_generic(cg_handle *h, int on);
/* Copy the compacted filter back expanded to the plain row layout (bin_size * bin_words words). */
int cg_debug_dump_filter(cg_handle *h, uint64_t *rows_out);

#ifdef __cplusplus
}
#endif
#endif /* CHARON_GPU_H */
