// mtx_parse.h — single-pass, multi-threaded reader of the vartrix alt.mtx / ref.mtx pair.
//
// Replaces the text passes of the reference (cellector/src/load_data.rs): `reader` (:240-251, .gz by
// extension), `consume_mtx_header` (:206-223: exactly three lines are skipped in each file, the dimensions come
// from the third line of the REF file, the nnz field is ignored), `read_mtx_lines` (:190-204: the two files are
// read in lock-step, coordinates come from the ALT line, the value is token 3 of each line, 1-based -> 0-based).
// The reference parses both files twice (three times with --vcf); here they are parsed once, by byte ranges on
// all host cores, straight into the caller's (pinned) COO arrays in file order.
#pragma once
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>
#include <zlib.h>

#include <algorithm>
#include <cstdint>
#include <cstring>
#include <functional>
#include <stdexcept>
#include <string>
#include <thread>
#include <vector>

namespace clxhost {

// Whole file in memory: mmap for plain files, inflated into a buffer for *.gz (multi-member, like MultiGzDecoder).
class FileBytes {
public:
    explicit FileBytes(const std::string& path) {
        const bool gz = path.size() >= 3 && path.compare(path.size() - 3, 3, ".gz") == 0;
        if (gz) {
            gzFile f = gzopen(path.c_str(), "rb");
            if (!f) throw std::runtime_error("couldn't open file " + path);
            gzbuffer(f, 1 << 20);
            std::vector<char> chunk(8 << 20);
            for (;;) {
                const int n = gzread(f, chunk.data(), (unsigned)chunk.size());
                if (n < 0) {
                    gzclose(f);
                    throw std::runtime_error("error while inflating " + path);
                }
                if (n == 0) break;
                owned_.insert(owned_.end(), chunk.data(), chunk.data() + n);
            }
            gzclose(f);
            data_ = owned_.data();
            size_ = owned_.size();
            return;
        }
        fd_ = open(path.c_str(), O_RDONLY);
        if (fd_ < 0) throw std::runtime_error("couldn't open file " + path);
        struct stat st;
        if (fstat(fd_, &st) != 0) throw std::runtime_error("couldn't stat file " + path);
        size_ = (size_t)st.st_size;
        if (size_) {
            void* p = mmap(nullptr, size_, PROT_READ, MAP_PRIVATE, fd_, 0);
            if (p == MAP_FAILED) throw std::runtime_error("couldn't mmap file " + path);
            madvise(p, size_, MADV_SEQUENTIAL | MADV_WILLNEED);
            map_ = p;
            data_ = (const char*)p;
        }
    }
    ~FileBytes() {
        if (map_) munmap(map_, size_);
        if (fd_ >= 0) close(fd_);
    }
    FileBytes(const FileBytes&) = delete;
    FileBytes& operator=(const FileBytes&) = delete;
    const char* data() const { return data_; }
    size_t size() const { return size_; }

private:
    const char* data_ = nullptr;
    size_t size_ = 0;
    void* map_ = nullptr;
    int fd_ = -1;
    std::vector<char> owned_;
};

// Lines of a text buffer the way Rust's BufRead::lines() yields them: split at '\n', a trailing "\r" dropped, no
// empty last line after a final '\n'.
inline void for_each_line(const char* p, size_t n, const std::function<void(const char*, size_t)>& fn) {
    size_t i = 0;
    while (i < n) {
        const char* nl = (const char*)memchr(p + i, '\n', n - i);
        size_t end = nl ? (size_t)(nl - p) : n;
        size_t len = end - i;
        if (len && p[i + len - 1] == '\r') --len;
        fn(p + i, len);
        i = end + 1;
    }
}

inline bool is_ws(char c) { return c == ' ' || c == '\t' || c == '\r' || c == '\n' || c == '\v' || c == '\f'; }

// next whitespace-separated token of [p, e) parsed like Rust's `parse::<usize>()`: optional '+', digits only
inline bool next_usize(const char*& p, const char* e, uint64_t* out) {
    while (p < e && is_ws(*p)) ++p;
    if (p >= e) return false;
    if (*p == '+') ++p;
    const char* d0 = p;
    uint64_t v = 0;
    while (p < e && *p >= '0' && *p <= '9') {
        const uint64_t nv = v * 10 + (uint64_t)(*p - '0');
        if (nv < v || v > UINT64_MAX / 10) return false;  // overflow: the reference panics
        v = nv;
        ++p;
    }
    if (p == d0) return false;
    if (p < e && !is_ws(*p)) return false;  // e.g. "3.0": parse::<usize> fails
    *out = v;
    return true;
}

struct MtxHeader {
    uint64_t n_loci = 0, n_cells = 0;
    size_t alt_body = 0, ref_body = 0;  // byte offset of the first data line
};

inline size_t skip_lines(const char* p, size_t n, int lines, size_t* third_line_start) {
    size_t i = 0;
    for (int k = 0; k < lines; ++k) {
        if (k == lines - 1 && third_line_start) *third_line_start = i;
        const char* nl = i < n ? (const char*)memchr(p + i, '\n', n - i) : nullptr;
        i = nl ? (size_t)(nl - p) + 1 : n;
    }
    return i;
}

inline MtxHeader read_mtx_header(const FileBytes& alt, const FileBytes& ref) {
    MtxHeader h;
    size_t l3 = 0;
    h.alt_body = skip_lines(alt.data(), alt.size(), 3, nullptr);
    h.ref_body = skip_lines(ref.data(), ref.size(), 3, &l3);
    const char* p = ref.data() + l3;
    const char* e = ref.data() + h.ref_body;
    if (!next_usize(p, e, &h.n_loci) || !next_usize(p, e, &h.n_cells))
        throw std::runtime_error("cannot parse the dimension line of the ref matrix market file");
    return h;
}

// byte offsets of the starts of lines 0, step, 2*step, ... (and the line count) of [p + begin, p + n)
struct LineIndex {
    uint64_t n_lines = 0;
    std::vector<size_t> chunk_start;     // byte offset where chunk k starts (aligned to a line start)
    std::vector<uint64_t> chunk_line0;   // index of the first line of chunk k
};

inline LineIndex index_lines(const char* p, size_t begin, size_t n, unsigned threads) {
    LineIndex ix;
    const size_t body = n > begin ? n - begin : 0;
    const unsigned T = (unsigned)std::max<size_t>(1, std::min<size_t>(threads, body / (1 << 16) + 1));
    std::vector<size_t> start(T + 1);
    for (unsigned k = 0; k <= T; ++k) start[k] = begin + body * k / T;
    // align chunk starts to the byte after a newline
    for (unsigned k = 1; k < T; ++k) {
        size_t s = start[k];
        if (s > begin && s < n && p[s - 1] != '\n') {
            const char* nl = (const char*)memchr(p + s, '\n', n - s);
            s = nl ? (size_t)(nl - p) + 1 : n;
        }
        start[k] = s;
    }
    start[T] = n;
    std::vector<uint64_t> cnt(T, 0);
    std::vector<std::thread> th;
    for (unsigned k = 0; k < T; ++k)
        th.emplace_back([&, k] {
            uint64_t c = 0;
            size_t i = start[k];
            const size_t e = start[k + 1];
            while (i < e) {
                const char* nl = (const char*)memchr(p + i, '\n', e - i);
                ++c;  // a last line without '\n' still counts
                if (!nl) break;
                i = (size_t)(nl - p) + 1;
            }
            cnt[k] = c;
        });
    for (auto& t : th) t.join();
    ix.chunk_start.assign(start.begin(), start.end() - 1);
    ix.chunk_line0.resize(T);
    uint64_t acc = 0;
    for (unsigned k = 0; k < T; ++k) {
        ix.chunk_line0[k] = acc;
        acc += cnt[k];
    }
    ix.n_lines = acc;
    return ix;
}

// byte offset of line `line` (0-based, < n_lines) of the body
inline size_t line_offset(const char* p, size_t n, const LineIndex& ix, uint64_t line) {
    size_t k = std::upper_bound(ix.chunk_line0.begin(), ix.chunk_line0.end(), line) - ix.chunk_line0.begin() - 1;
    size_t i = ix.chunk_start[k];
    for (uint64_t l = ix.chunk_line0[k]; l < line; ++l) {
        const char* nl = (const char*)memchr(p + i, '\n', n - i);
        i = nl ? (size_t)(nl - p) + 1 : n;
    }
    return i;
}

struct CooSink {  // caller-owned arrays of at least n_pairs elements
    uint32_t *locus, *cell, *ref, *alt;
};

// Number of (alt line, ref line) pairs: izip! stops at the shorter file.
struct MtxPairPlan {
    MtxHeader hdr;
    LineIndex alt_ix, ref_ix;
    uint64_t n_pairs = 0;
};

inline MtxPairPlan plan_mtx_pair(const FileBytes& alt, const FileBytes& ref, unsigned threads) {
    MtxPairPlan pl;
    pl.hdr = read_mtx_header(alt, ref);
    pl.alt_ix = index_lines(alt.data(), pl.hdr.alt_body, alt.size(), threads);
    pl.ref_ix = index_lines(ref.data(), pl.hdr.ref_body, ref.size(), threads);
    pl.n_pairs = std::min(pl.alt_ix.n_lines, pl.ref_ix.n_lines);
    return pl;
}

// Parses pair i into sink[i] for all i, in parallel over line ranges.  Throws on the inputs the reference
// panics on (missing token, non-integer value, locus or cell index 0).
inline void parse_mtx_pair(const FileBytes& alt, const FileBytes& ref, const MtxPairPlan& pl, CooSink sink, unsigned threads) {
    const uint64_t n = pl.n_pairs;
    if (n == 0) return;
    const unsigned T = (unsigned)std::max<uint64_t>(1, std::min<uint64_t>(threads, n / 4096 + 1));
    std::vector<std::string> errors(T);
    std::vector<std::thread> th;
    for (unsigned k = 0; k < T; ++k)
        th.emplace_back([&, k] {
            const uint64_t l0 = n * k / T, l1 = n * (k + 1) / T;
            if (l0 >= l1) return;
            const char* ap = alt.data();
            const char* rp = ref.data();
            size_t ai = line_offset(ap, alt.size(), pl.alt_ix, l0);
            size_t ri = line_offset(rp, ref.size(), pl.ref_ix, l0);
            for (uint64_t l = l0; l < l1; ++l) {
                const char* anl = (const char*)memchr(ap + ai, '\n', alt.size() - ai);
                const char* rnl = (const char*)memchr(rp + ri, '\n', ref.size() - ri);
                const char* ae = anl ? anl : ap + alt.size();
                const char* re = rnl ? rnl : rp + ref.size();
                const char* a = ap + ai;
                const char* r = rp + ri;
                uint64_t locus, cell, av, rv, skip;
                const bool ok = next_usize(a, ae, &locus) && next_usize(a, ae, &cell) && next_usize(a, ae, &av) &&
                                next_usize(r, re, &skip) && next_usize(r, re, &skip) && next_usize(r, re, &rv);
                if (!ok || locus == 0 || cell == 0 || locus > 0xFFFFFFFFull || cell > 0xFFFFFFFFull || av > 0xFFFFFFFFull ||
                    rv > 0xFFFFFFFFull) {
                    errors[k] = "malformed matrix market data line " + std::to_string(l + 4) +
                                " (expected `locus cell count`, 1-based unsigned integers)";
                    return;
                }
                sink.locus[l] = (uint32_t)(locus - 1);
                sink.cell[l] = (uint32_t)(cell - 1);
                sink.alt[l] = (uint32_t)av;
                sink.ref[l] = (uint32_t)rv;
                ai = (size_t)(ae - ap) + 1;
                ri = (size_t)(re - rp) + 1;
            }
        });
    for (auto& t : th) t.join();
    for (const auto& e : errors)
        if (!e.empty()) throw std::runtime_error(e);
}

}  // namespace clxhost
