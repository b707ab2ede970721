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
            for (unsigned char q : r.qual) if (q < 33 || q > 126) throw std::runtime_error(path_ + ": illegal quality character in record " + r.id);
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
    void append_seq(FastxRecord &r, const std::string &l) {
        const size_t o = r.seq.size();
        r.seq.resize(o + l.size());
        size_t n = o;
        for (unsigned char c : l) {
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
