// charon_b200/host/fastx.hpp — FASTA/FASTQ (+gzip) record reader for the host side of the path.
// Replaces seqan3::sequence_file_input<my_traits> as used at src/dehost_main.cpp:324,335-339
// (traits include/utils.hpp:17-19: sequence alphabet dna5, legal input alphabet dna15, qualities phred94).
//   id   = whole header line after '@' / '>'
//   seq  = dna5 characters, upper case: A C G T N (U -> T, other IUPAC letters -> N); any other character
//          is an error, as in seqan3 (the legal alphabet is dna15)
//   qual = raw phred+33 characters (FASTQ) or empty (FASTA)
// The format is sniffed from the first byte rather than from the file extension; gzip is detected by zlib.
#pragma once
#include <zlib.h>
#if defined(__SSE2__)
#include <emmintrin.h>
#endif

#include <cstdint>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

namespace charon_host {

struct FastxRecord {
    std::string id, seq, qual;
};

class FastxReader {
public:
    explicit FastxReader(const std::string &path) : path_(path), buf_(1u << 20) {
        gz_ = gzopen(path.c_str(), "rb");
        if (!gz_) throw std::runtime_error("cannot open read file " + path);
        gzbuffer(gz_, 1u << 20);
        for (int c = 0; c < 256; ++c) lut_[c] = 0;
        const char *acgt = "ACGT";
        for (int i = 0; i < 4; ++i) { lut_[(unsigned char)acgt[i]] = acgt[i]; lut_[(unsigned char)(acgt[i] | 0x20)] = acgt[i]; }
        lut_[(unsigned char)'U'] = 'T'; lut_[(unsigned char)'u'] = 'T';
        for (const char *p = "NRYSWKMBDHVnryswkmbdhv"; *p; ++p) lut_[(unsigned char)*p] = 'N';
    }
    ~FastxReader() { if (gz_) gzclose(gz_); }
    FastxReader(const FastxReader &) = delete;
    FastxReader &operator=(const FastxReader &) = delete;

    bool is_fastq() const { return fastq_; }

    // Next record; false at end of file.  Throws std::runtime_error on malformed input.
    bool next(FastxRecord &r) {
        if (!started_) {
            started_ = true;
            have_line_ = read_line(line_);
            while (have_line_ && line_.empty()) have_line_ = read_line(line_);
            if (!have_line_) return false;
            if (line_[0] == '@') fastq_ = true;
            else if (line_[0] == '>') fastq_ = false;
            else throw std::runtime_error(path_ + ": neither FASTA ('>') nor FASTQ ('@')");
        }
        while (have_line_ && line_.empty()) have_line_ = read_line(line_);
        if (!have_line_) return false;
        const char lead = fastq_ ? '@' : '>';
        if (line_[0] != lead) throw std::runtime_error(path_ + ": record " + std::to_string(nrec_ + 1) + " does not start with '" + lead + "'");
        r.id.assign(line_, 1, std::string::npos);
        r.seq.clear(); r.qual.clear();
        if (fastq_) {
            // sequence lines up to the '+' separator, then as many quality characters as bases
            for (;;) {
                have_line_ = read_line(line_);
                if (!have_line_) throw std::runtime_error(path_ + ": truncated FASTQ record " + r.id);
                if (!line_.empty() && line_[0] == '+') break;
                append_seq(r, line_);
            }
            while (r.qual.size() < r.seq.size()) {
                have_line_ = read_line(line_);
                if (!have_line_) throw std::runtime_error(path_ + ": truncated qualities in record " + r.id);
                r.qual += line_;
            }
            if (r.qual.size() != r.seq.size()) throw std::runtime_error(path_ + ": sequence and quality lengths differ in record " + r.id);
            if (!quality_ok(r.qual)) throw std::runtime_error(path_ + ": illegal quality character in record " + r.id);
            have_line_ = read_line(line_);
        } else {
            for (;;) {
                have_line_ = read_line(line_);
                if (!have_line_ || (!line_.empty() && line_[0] == '>')) break;
                append_seq(r, line_);
            }
        }
        ++nrec_;
        return true;
    }

private:
    // Length of the leading run of plain upper-case A/C/G/T in [p, p + n), rounded down to whole vectors: those letters need
    // neither translation nor rejection and are copied as they are (SSE2 is part of x86-64; elsewhere the run is empty and
    // every letter goes through the table).
    static bool quality_ok(const std::string &q) {   // phred94: every character in '!' .. '~'
        size_t i = 0;
        unsigned char lo = 255, hi = 0;
#if defined(__SSE2__)
        __m128i vlo = _mm_set1_epi8(char(0xFF)), vhi = _mm_setzero_si128();
        for (; i + 16 <= q.size(); i += 16) {
            const __m128i x = _mm_loadu_si128(reinterpret_cast<const __m128i *>(q.data() + i));
            vlo = _mm_min_epu8(vlo, x); vhi = _mm_max_epu8(vhi, x);
        }
        alignas(16) unsigned char a[16], b[16];
        _mm_store_si128(reinterpret_cast<__m128i *>(a), vlo); _mm_store_si128(reinterpret_cast<__m128i *>(b), vhi);
        for (int k = 0; k < 16; ++k) { if (a[k] < lo) lo = a[k]; if (b[k] > hi) hi = b[k]; }
#endif
        for (; i < q.size(); ++i) { const unsigned char c = static_cast<unsigned char>(q[i]); if (c < lo) lo = c; if (c > hi) hi = c; }
        return q.empty() || (lo >= 33 && hi <= 126);
    }
    static size_t plain_run(const char *p, size_t n) {
        size_t i = 0;
#if defined(__SSE2__)
        const __m128i A = _mm_set1_epi8('A'), Cc = _mm_set1_epi8('C'), G = _mm_set1_epi8('G'), T = _mm_set1_epi8('T');
        for (; i + 16 <= n; i += 16) {
            const __m128i x = _mm_loadu_si128(reinterpret_cast<const __m128i *>(p + i));
            const __m128i ok = _mm_or_si128(_mm_or_si128(_mm_cmpeq_epi8(x, A), _mm_cmpeq_epi8(x, Cc)), _mm_or_si128(_mm_cmpeq_epi8(x, G), _mm_cmpeq_epi8(x, T)));
            if (_mm_movemask_epi8(ok) != 0xFFFF) break;
        }
#endif
        return i;
    }
    void append_seq(FastxRecord &r, const std::string &l) {
        const size_t o = r.seq.size();
        r.seq.resize(o + l.size());
        size_t n = o, i = 0;
        while (i < l.size()) {
            if (n == o + i) {   // nothing skipped so far: source and destination offsets agree, copy the plain run in one go
                const size_t run = plain_run(l.data() + i, l.size() - i);
                if (run) { std::memcpy(&r.seq[0] + n, l.data() + i, run); n += run; i += run; if (i == l.size()) break; }
            }
            const unsigned char c = static_cast<unsigned char>(l[i++]);
            const char m = lut_[c];
            if (!m) {
                if (c == ' ' || c == '\t') continue;
                throw std::runtime_error(path_ + ": illegal sequence character '" + std::string(1, char(c)) + "' in record " + r.id);
            }
            r.seq[n++] = m;
        }
        r.seq.resize(n);
    }
    bool fill() {
        const int n = gzread(gz_, buf_.data(), unsigned(buf_.size()));
        if (n < 0) { int e = 0; throw std::runtime_error(path_ + ": " + gzerror(gz_, &e)); }
        pos_ = 0; len_ = size_t(n);
        return n > 0;
    }
    bool read_line(std::string &out) {
        out.clear();
        bool any = false;
        for (;;) {
            if (pos_ == len_ && !fill()) break;
            any = true;
            const char *s = buf_.data() + pos_;
            const char *nl = static_cast<const char *>(std::memchr(s, '\n', len_ - pos_));
            if (nl) { out.append(s, size_t(nl - s)); pos_ += size_t(nl - s) + 1; break; }
            out.append(s, len_ - pos_);
            pos_ = len_;
        }
        if (!out.empty() && out.back() == '\r') out.pop_back();
        return any;
    }

    std::string path_;
    gzFile gz_ = nullptr;
    std::vector<char> buf_;
    size_t pos_ = 0, len_ = 0;
    std::string line_;
    bool started_ = false, have_line_ = false, fastq_ = true;
    uint64_t nrec_ = 0;
    char lut_[256];
};

// get_extension (src/utils.cpp:126-133): extension of the read file with a trailing ".gz" removed first
inline std::string read_file_extension(const std::string &path) {
    auto ext_of = [](const std::string &p) {
        const size_t slash = p.find_last_of('/');
        const std::string name = slash == std::string::npos ? p : p.substr(slash + 1);
        const size_t dot = name.find_last_of('.');
        if (dot == std::string::npos || dot == 0) return std::string();
        return name.substr(dot);
    };
    std::string ext = ext_of(path);
    if (ext == ".gz") ext = ext_of(path.substr(0, path.size() - 3));
    return ext;
}

}  // namespace charon_host
