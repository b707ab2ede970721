// charon_b200/host/charon_main.cpp — `charon dehost` / `charon classify` (C++14 host program over libcharon_gpu.so).
// Process boundary of the reference kept: CLI (src/main.cpp:40-62, src/dehost_main.cpp:208-312,479-548,
// src/classify_main.cpp:25-119,299-352), `.idx` input (src/load_index.cpp), TSV rows on stdout
// (include/read_entry.hpp:322-337), extracted reads, charon.log.  There is no CPU classification path:
// without a B200 the program stops with the library's error.
#include <omp.h>

#include <algorithm>
#include <chrono>
#include <condition_variable>
#include <deque>
#include <exception>
#include <iostream>
#include <mutex>

#include "cli.hpp"
#include "runner.hpp"

namespace charon_host {

// pinned host array (cudaHostAlloc through the library): what the copy engines read and write
template <class T> struct Pinned {
    T *p = nullptr;
    size_t cap = 0;
    void reserve(size_t n) {
        if (n <= cap) return;
        if (p) cg_host_free(p);
        p = nullptr; cap = 0;
        void *q = nullptr;
        const size_t want = n + n / 4 + 64;
        if (cg_host_alloc(want * sizeof(T), &q) != 0) throw std::runtime_error(std::string("pinned allocation failed: ") + cg_last_error(nullptr));
        p = static_cast<T *>(q); cap = want;
    }
    T &operator[](size_t i) { return p[i]; }
    Pinned() = default;
    Pinned(const Pinned &) = delete;
    Pinned &operator=(const Pinned &) = delete;
    ~Pinned() { if (p) cg_host_free(p); }
};

// One GPU's share of one batch, in the form the library takes it: CG_BASES_2BIT bases written while the per-read scalars
// are computed (every segment on a 64-base boundary), offsets, scalars, and the outputs the GPU copies back.  Two of
// these per GPU = the two pipeline slots of cg_submit_batch / cg_wait_batch.
struct SlotBuf {
    Pinned<uint8_t> packed;
    Pinned<uint64_t> seg_off, seg_len, exc;
    Pinned<uint32_t> read_seg, length, num_hashes, counts, unique_counts;
    Pinned<float> mq, cr;
    Pinned<double> prob;
    Pinned<uint8_t> call, conf;
    size_t r0 = 0, n_reads = 0, n_segs = 0, n_exc = 0;
    uint64_t total_bases = 0;
};

struct GpuSlice {   // one GPU: its handle and its two slots
    cg_handle *h = nullptr;
    SlotBuf slot[2];
};

// A batch of reads on its way through the pipeline: parsed by the reader thread, scalars + packing by the -t team,
// classified on the GPU(s) (slot = batch number & 1), printed by the writer thread.
struct Batch {
    std::vector<ReadRow> rows;
    std::vector<std::vector<uint64_t>> read_exc;   // per read: batch positions of its non-ACGT bases (almost always empty)
    uint64_t seq = 0;
    bool device_cr = false;                        // the compression ratios of this batch are computed on the GPU (--gpu_compression)
    bool paired_mismatch = false;                  // reader: the pair after the last row of this batch has differing ids
    std::string mismatch_text;
};

template <class T> class BoundedQueue {
public:
    explicit BoundedQueue(size_t cap) : cap_(cap) {}
    bool push(T &&v) {
        std::unique_lock<std::mutex> g(m_);
        not_full_.wait(g, [&] { return q_.size() < cap_ || closed_; });
        if (closed_) return false;
        q_.push_back(std::move(v));
        not_empty_.notify_one();
        return true;
    }
    bool pop(T &out) {
        std::unique_lock<std::mutex> g(m_);
        not_empty_.wait(g, [&] { return !q_.empty() || closed_; });
        if (q_.empty()) return false;
        out = std::move(q_.front());
        q_.pop_front();
        not_full_.notify_one();
        return true;
    }
    void close() { std::lock_guard<std::mutex> g(m_); closed_ = true; not_empty_.notify_all(); not_full_.notify_all(); }
private:
    std::mutex m_;
    std::condition_variable not_empty_, not_full_;
    std::deque<T> q_;
    size_t cap_;
    bool closed_ = false;
};

class Runner {
public:
    Runner(const Options &o, Logger &lg) : opt_(o), log_(lg) {}
    ~Runner() { for (auto &s : gpu_) if (s->h) cg_destroy(s->h); }

    int run() {
        std::string err;
        if (!load_index_file(opt_.db, idx_, err)) { CH_ERROR(log_, "%s", err.c_str()); std::fprintf(stderr, "charon: %s\n", err.c_str()); return 1; }
        ncat_ = idx_.num_categories();
        if (opt_.dehost) {
            host_index_ = idx_.host_category_index();
            if (host_index_ == 255) { CH_ERROR(log_, "Neither 'human' nor 'host' appear as categories in the index"); std::fprintf(stderr, "charon: neither 'human' nor 'host' is a category of this index\n"); return 1; }
            CH_INFO(log_, "Found host at index %u in the index categories", unsigned(host_index_));
        }
        if (!setup_extract()) return 1;
        if (opt_.dehost ? (opt_.dist != "gamma" && opt_.dist != "beta" && opt_.dist != "kde") : (opt_.dist != "gamma" && opt_.dist != "beta")) {
            CH_ERROR(log_, opt_.dehost ? "Supported distributions are [gamma , beta, kde]" : "Supported distributions are [gamma , beta]");
            return 1;   // the reference swallows this return value; so does main() below
        }
        std::vector<uint8_t> b2c;
        if (!idx_.bin_to_cat(b2c, err)) { CH_ERROR(log_, "%s", err.c_str()); std::fprintf(stderr, "charon: %s\n", err.c_str()); return 1; }
        cg_index_desc d{};
        d.kmer_size = idx_.kmer_size; d.window_size = idx_.window_size;
        d.bins = idx_.bins; d.technical_bins = idx_.technical_bins; d.bin_size = idx_.bin_size; d.hash_shift = idx_.hash_shift;
        d.bin_words = idx_.bin_words; d.hash_funs = idx_.hash_funs;
        d.filter_format = CG_FILTER_SD_VECTOR;
        d.sd_low = idx_.sd_low.data(); d.sd_low_bits = idx_.sd_low_bits; d.sd_high = idx_.sd_high.data(); d.sd_high_bits = idx_.sd_high_bits;
        d.sd_wl = idx_.sd_wl; d.ncat = ncat_; d.bin_to_cat = b2c.data();
        for (int g = 0; g < opt_.gpus; ++g) {
            std::unique_ptr<GpuSlice> s(new GpuSlice);
            const bool part = opt_.partition_filter && opt_.gpus > 1;
            if ((part ? cg_create_partitioned(g, &d, uint32_t(g), uint32_t(opt_.gpus), &s->h) : cg_create(g, &d, 0, &s->h)) != 0) {
                const std::string m = cg_last_error(nullptr);
                CH_ERROR(log_, "GPU %d: %s", g, m.c_str());
                std::fprintf(stderr, "charon: GPU %d: %s\n", g, m.c_str());
                return 1;
            }
            cg_set_host_pack(s->h, 0, 0);   // the reads reach the library already packed (prepare_batch)
            gpu_.push_back(std::move(s));
        }
        if (opt_.partition_filter && opt_.gpus > 1)
            for (int g = 0; g < opt_.gpus; ++g)
                for (int q = 0; q < opt_.gpus; ++q)
                    if (q != g && cg_partition_attach_peer(gpu_[size_t(g)]->h, uint32_t(q), gpu_[size_t(q)]->h) != 0) {
                        const std::string m = cg_last_error(gpu_[size_t(g)]->h);
                        CH_ERROR(log_, "GPU %d: %s", g, m.c_str());
                        std::fprintf(stderr, "charon: GPU %d: %s\n", g, m.c_str());
                        return 1;
                    }
        std::vector<uint64_t>().swap(idx_.sd_low); std::vector<uint64_t>().swap(idx_.sd_high);   // the filter now lives in HBM
        model_ = StatsModel(opt_, ncat_);
        if (opt_.run_extract) cache_capacity_ = size_t(opt_.num_reads_to_fit) * ncat_ * 4;   // result.hpp:84 (reserve == capacity)
        classified_counts_.assign(ncat_, 0);
        if (opt_.is_paired) CH_INFO(log_, "%s files %s and %s", opt_.dehost ? "Dehosting" : "Classifying", opt_.read_file.c_str(), opt_.read_file2.c_str());
        else CH_INFO(log_, "%s file %s", opt_.dehost ? "Dehosting" : "Classifying", opt_.read_file.c_str());
        const auto t_loop0 = std::chrono::steady_clock::now();
        run_pipeline();
        {
            const double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t_loop0).count();
            unsigned long long n = unclassified_;
            for (uint64_t c : classified_counts_) n += c;
            CH_INFO(log_, "Read loop: %llu reads in %.3f s (%.0f reads/s); stage wall times: reader %.3f s, scalars + packing (-t %u) %.3f s, waiting for the GPU %.3f s, writer %.3f s",
                    n, sec, sec > 0 ? double(n) / sec : 0.0, t_reader_, unsigned(opt_.threads), t_prepare_, t_gpu_wait_, t_writer_);
            if (prof_prep_) CH_INFO(log_, "prepare_batch, summed over threads: quality %.3f s (of which the phred sum %.3f s), packing %.3f s, freeing %.3f s", t_pq_, t_pz_, t_pp_, t_pf_);
        }
        CH_INFO(log_, "Results summary: ");
        for (size_t i = 0; i < ncat_; ++i) CH_INFO(log_, "%s :\t\t%llu", idx_.categories[i].c_str(), (unsigned long long)classified_counts_[i]);
        CH_INFO(log_, "unclassified :\t%llu", (unsigned long long)unclassified_);
        return 0;
    }

private:
    // *_main: extraction set-up (src/dehost_main.cpp:506-536, src/classify_main.cpp:321-345)
    bool setup_extract() {
        opt_.run_extract = !opt_.category_to_extract.empty();
        if (!opt_.run_extract) return true;
        const auto &cats = idx_.categories;
        if (opt_.category_to_extract != "all" && std::find(cats.begin(), cats.end(), opt_.category_to_extract) == cats.end()) {
            std::string options;
            for (const auto &c : cats) options += c + " ";
            CH_ERROR(log_, "Cannot extract %s, please chose one of [ all %s]", opt_.category_to_extract.c_str(), options.c_str());
            return false;
        }
        if (opt_.prefix.empty()) opt_.prefix = "charon";
        std::vector<std::string> to_extract;
        if (opt_.category_to_extract == "all") to_extract = cats; else to_extract.push_back(opt_.category_to_extract);
        const std::string ext = read_file_extension(opt_.read_file);
        for (const auto &c : to_extract) {
            const uint8_t ci = idx_.category_index(c);
            if (opt_.is_paired) { extract_paths_[ci].push_back(opt_.prefix + "_" + c + "_1" + ext + ".gz"); extract_paths_[ci].push_back(opt_.prefix + "_" + c + "_2" + ext + ".gz"); }
            else extract_paths_[ci].push_back(opt_.prefix + "_" + c + ext + ".gz");
        }
        return true;
    }
    void open_extract(bool fastq) {
        if (extract_open_) return;
        extract_open_ = true;
        for (auto &kv : extract_paths_)
            for (auto &p : kv.second) extract_[kv.first].emplace_back(new ExtractWriter(p, fastq));
    }

    void run_pipeline();
    void reader_thread(BoundedQueue<Batch> &parsed);
    void writer_thread(BoundedQueue<std::vector<ReadRow>> &done);
    void prepare_batch(Batch &b, int slot, bool allow_device_cr);
    bool prof_prep_ = std::getenv("CHARON_PREP_PROFILE") != nullptr;
    double t_pq_ = 0, t_pp_ = 0, t_pf_ = 0, t_pz_ = 0;
    void submit_batch(Batch &b, int slot, bool with_eval);
    void collect_batch(Batch &b, int slot, bool with_eval);
    void train_batch(Batch &b, std::vector<ReadRow> &to_print);
    void eval_rows(std::vector<ReadRow *> &rows);
    void add_read(ReadRow &row, std::vector<ReadRow *> &pending, std::vector<ReadRow> &moved_cache);
    void emit(ReadRow &row);
    void flush_out() { if (!out_.empty()) { std::fwrite(out_.data(), 1, out_.size(), stdout); out_.clear(); } std::fflush(stdout); }
    void upload_model() { for (auto &s : gpu_) if (model_.upload(s->h, host_index_) != 0) throw std::runtime_error(std::string("cg_set_model: ") + cg_last_error(s->h)); }

    Options opt_;
    Logger &log_;
    IndexFile idx_;
    size_t ncat_ = 0;
    uint8_t host_index_ = 0;
    std::vector<std::unique_ptr<GpuSlice>> gpu_;
    StatsModel model_;
    bool model_uploaded_ = false;
    std::vector<ReadRow> cache_;
    size_t cache_capacity_ = 0;
    std::vector<uint64_t> classified_counts_;
    uint64_t unclassified_ = 0;
    std::map<uint8_t, std::vector<std::string>> extract_paths_;
    std::map<uint8_t, std::vector<std::unique_ptr<ExtractWriter>>> extract_;
    bool extract_open_ = false, input_fastq_ = true;
    std::string out_;
    std::exception_ptr reader_err_, writer_err_;
    double t_prepare_ = 0, t_gpu_wait_ = 0, t_reader_ = 0, t_writer_ = 0;   // stage wall times (charon.log, -v)
};

}  // namespace charon_host

#include "index_main.inc"
#include "runner_impl.inc"
