// charon_b200/host/runner.hpp — the read loop of `charon dehost` / `charon classify` with its body on B200.
//
// Mirrors dehost_reads / dehost_paired_reads / classify_reads / classify_paired_reads
// (src/dehost_main.cpp:314-477, src/classify_main.cpp:121-296) and Result (include/result.hpp:59-213):
//   host : chunk ingest, read id / length / mean quality / gzip ratio per read (:345-363), the training
//          cache + state machine (only active with --extract), TSV rows (read_entry.hpp:322-337),
//          extraction writers
//   GPU  : minimiser_hash -> bulk_contains -> update_entry -> post_process -> apply_model -> call_*
//          through libcharon_gpu.so (include/charon_gpu.h)
// Rows are written in input order and the first read that reaches Result::add_read without --extract is
// dropped, i.e. the reference's behaviour at -t 1.
#pragma once
#include <zlib.h>

#include <cstdarg>
#include <cstdio>
#include <ctime>
#include <map>
#include <memory>
#include <string>
#include <thread>
#include <vector>

#include "../../include/charon_gpu.h"
#include "fastx.hpp"
#include "index_file.hpp"
#include "stats_model.hpp"

namespace charon_host {

// ---- minimal file logger (the reference logs through plog into charon.log; content is informational) ----
class Logger {
public:
    void open(const std::string &path, int verbosity) { f_ = std::fopen(path.c_str(), "a"); level_ = verbosity; }
    ~Logger() { if (f_) std::fclose(f_); }
    void log(int level, const char *tag, const char *fmt, ...) {
        if (!f_ || level > level_) return;
        char ts[32];
        std::time_t t = std::time(nullptr);
        std::tm tmv; localtime_r(&t, &tmv);
        std::strftime(ts, sizeof ts, "%Y-%m-%d %H:%M:%S", &tmv);
        std::fprintf(f_, "%s %-5s ", ts, tag);
        va_list ap; va_start(ap, fmt); std::vfprintf(f_, fmt, ap); va_end(ap);
        std::fputc('\n', f_);
        std::fflush(f_);
    }
private:
    FILE *f_ = nullptr;
    int level_ = 0;
};
#define CH_INFO(lg, ...) (lg).log(0, "INFO", __VA_ARGS__)
#define CH_WARN(lg, ...) (lg).log(0, "WARN", __VA_ARGS__)
#define CH_ERROR(lg, ...) (lg).log(0, "ERROR", __VA_ARGS__)
#define CH_DEBUG(lg, ...) (lg).log(1, "DEBUG", __VA_ARGS__)

// sum of the bytes of a string (SSE2: sixteen at a time through psadbw)
inline uint64_t byte_sum(const std::string &q) {
    const unsigned char *p = reinterpret_cast<const unsigned char *>(q.data());
    const size_t n = q.size();
    size_t i = 0;
    uint64_t total = 0;
#if defined(__SSE2__)
    __m128i acc = _mm_setzero_si128();
    const __m128i zero = _mm_setzero_si128();
    for (; i + 16 <= n; i += 16) acc = _mm_add_epi64(acc, _mm_sad_epu8(_mm_loadu_si128(reinterpret_cast<const __m128i *>(p + i)), zero));
    alignas(16) uint64_t lanes[2];
    _mm_store_si128(reinterpret_cast<__m128i *>(lanes), acc);
    total = lanes[0] + lanes[1];
#endif
    for (; i < n; ++i) total += p[i];
    return total;
}

// get_compression_ratio (src/utils.cpp:114-124): gzip::compress of mapbox/gzip-hpp = zlib deflateInit2(
// Z_DEFAULT_COMPRESSION, Z_DEFLATED, 15 + 16 (gzip wrapper), memLevel 8, Z_DEFAULT_STRATEGY), one Z_FINISH pass.
class GzipRatio {
public:
    GzipRatio() { std::memset(&zs_, 0, sizeof zs_); ok_ = deflateInit2(&zs_, Z_DEFAULT_COMPRESSION, Z_DEFLATED, 15 | 16, 8, Z_DEFAULT_STRATEGY) == Z_OK; }
    ~GzipRatio() { if (ok_) deflateEnd(&zs_); }
    GzipRatio(const GzipRatio &) = delete;
    float operator()(const char *s1, size_t n1, const char *s2 = nullptr, size_t n2 = 0) {
        if (!ok_) throw std::runtime_error("zlib deflateInit2 failed");
        deflateReset(&zs_);
        const size_t n = n1 + n2;
        buf_.resize(deflateBound(&zs_, uLong(n)) + 64);
        zs_.next_out = buf_.data(); zs_.avail_out = uInt(buf_.size());
        zs_.next_in = reinterpret_cast<Bytef *>(const_cast<char *>(s1)); zs_.avail_in = uInt(n1);
        int rc = deflate(&zs_, n2 ? Z_NO_FLUSH : Z_FINISH);
        if (n2) {
            zs_.next_in = reinterpret_cast<Bytef *>(const_cast<char *>(s2)); zs_.avail_in = uInt(n2);
            rc = deflate(&zs_, Z_FINISH);
        }
        if (rc != Z_STREAM_END) throw std::runtime_error("zlib deflate failed");
        return float(static_cast<double>(zs_.total_out) / static_cast<double>(n));
    }
private:
    z_stream zs_;
    bool ok_ = false;
    std::vector<unsigned char> buf_;
};

// gzip-compressed FASTA/FASTQ writer for --extract (seqan3::sequence_file_output on "<prefix>_<category><ext>.gz",
// src/dehost_main.cpp:524-535, include/result.hpp:118-128).  FASTA lines are wrapped at 80 letters as seqan3 does.
class ExtractWriter {
public:
    ExtractWriter(const std::string &path, bool fastq) : fastq_(fastq) {
        gz_ = gzopen(path.c_str(), "wb");
        if (!gz_) throw std::runtime_error("cannot open " + path + " for writing");
    }
    ~ExtractWriter() { if (gz_) gzclose(gz_); }
    ExtractWriter(const ExtractWriter &) = delete;
    void push_back(const FastxRecord &r) {
        std::string s;
        s.reserve(r.id.size() + 2 * r.seq.size() + 16);
        if (fastq_) { s += '@'; s += r.id; s += '\n'; s += r.seq; s += "\n+\n"; s += r.qual; s += '\n'; }
        else {
            s += '>'; s += r.id; s += '\n';
            for (size_t i = 0; i < r.seq.size(); i += 80) { s.append(r.seq, i, 80); s += '\n'; }
        }
        gzwrite(gz_, s.data(), unsigned(s.size()));
    }
private:
    gzFile gz_ = nullptr;
    bool fastq_;
};

// One read on its way through the loop (ReadEntry's data members, include/read_entry.hpp:26-44).
struct ReadRow {
    std::string read_id;
    uint32_t length = 0, num_hashes = 0;
    float mean_quality = 0, compression = 0;
    std::vector<uint32_t> counts, unique_counts;
    std::vector<double> probabilities;
    uint8_t call = 255, confidence = 0;
    FastxRecord rec1, rec2;   // kept only while extraction may need them
    bool paired = false;
};

// "%g"-style text exactly as std::ostream prints float/double with precision(6) (read_entry.hpp:329)
inline void append_g(std::string &s, double v) { char b[40]; const int n = std::snprintf(b, sizeof b, "%.6g", v); s.append(b, size_t(n)); }
inline void append_u(std::string &s, unsigned long long v) { char b[24]; const int n = std::snprintf(b, sizeof b, "%llu", v); s.append(b, size_t(n)); }

// ReadEntry::print_assignment_result (include/read_entry.hpp:322-337); proportions as get_proportions (:140-150)
inline void append_tsv_row(std::string &out, const ReadRow &r, const IndexFile &idx) {
    out += (r.call == 255 ? "U\t" : "C\t");
    out += r.read_id; out += '\t';
    out += idx.category_name(r.call); out += '\t';
    append_u(out, r.length); out += '\t';
    append_u(out, r.num_hashes); out += '\t';
    append_g(out, r.mean_quality); out += '\t';
    append_u(out, r.confidence); out += '\t';
    append_g(out, r.compression); out += '\t';
    for (size_t i = 0; i < idx.categories.size(); ++i) {
        const float prop = static_cast<float>(r.counts[i]) / static_cast<float>(r.num_hashes);
        const float uprop = static_cast<float>(r.unique_counts[i]) / static_cast<float>(r.num_hashes);
        out += idx.categories[i]; out += ':';
        append_u(out, r.counts[i]); out += ':';
        append_g(out, prop); out += ':';
        append_g(out, uprop); out += ':';
        append_g(out, r.probabilities[i]); out += ' ';
    }
    out += '\n';
}

}  // namespace charon_host
