#include "read_feeder.hpp"

#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>
#include <zlib.h>

#include <atomic>
#include <chrono>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <stdexcept>
#include <thread>

namespace mtg {
namespace gpu {

namespace {

constexpr size_t kRangeBytes = size_t(64) << 20;   // parse unit of a memory-mapped file

struct Batch {   // one worker's output buffer
    char *bases = nullptr;
    size_t cap = 0, used = 0, read_start = 0;
    std::vector<uint64_t> offsets{0};
    ReadFeeder::Alloc alloc;
    ReadFeeder::Free free_fn;
    const ReadFeeder::Sink *sink;
    size_t flush_at;
    bool want_sum;
    FeederStats st;

    void reserve(size_t extra) {
        if (used + extra <= cap) return;
        size_t ncap = cap ? cap : flush_at + (flush_at >> 2);
        while (ncap < used + extra) ncap *= 2;
        char *nb = static_cast<char *>(alloc(ncap));
        if (!nb) throw std::runtime_error("read feeder: cannot allocate a batch buffer");
        if (used) memcpy(nb, bases, used);
        if (bases) free_fn(bases);
        bases = nb;
        cap = ncap;
    }
    void append(const char *s, size_t n) {
        reserve(n);
        memcpy(bases + used, s, n);
        used += n;
    }
    void end_record() {
        if (used == read_start) return;   // empty record
        if (want_sum) {
            uint64_t h = 1469598103934665603ull;
            for (size_t i = read_start; i < used; ++i) h = (h ^ (unsigned char)bases[i]) * 1099511628211ull;
            st.checksum += h;
        }
        st.n_reads += 1;
        st.n_bases += used - read_start;
        offsets.push_back(used);
        read_start = used;
        if (used >= flush_at) flush();
    }
    void flush() {   // only between records
        if (offsets.size() > 1) {
            (*sink)(bases, offsets.data(), offsets.size() - 1);
            st.n_batches += 1;
        }
        used = read_start = 0;
        offsets.assign(1, 0);
    }
    ~Batch() { if (bases) free_fn(bases); }
};

// FASTA / FASTQ line state machine; fed with arbitrary byte blocks
struct RecordParser {
    Batch &out;
    int fmt = 0;        // 0 unknown, 1 FASTA, 2 FASTQ
    int fq_line = 0;
    std::string carry;  // unterminated tail of the previous block
    explicit RecordParser(Batch &b) : out(b) {}

    void line(const char *s, size_t n) {
        if (n && s[n - 1] == '\r') --n;
        if (!fmt) {
            if (!n) return;
            fmt = s[0] == '@' ? 2 : 1;
        }
        if (fmt == 1) {
            if (n && s[0] == '>') out.end_record();
            else if (n) out.append(s, n);
        } else {
            if (fq_line == 1) { out.append(s, n); out.end_record(); }
            fq_line = (fq_line + 1) & 3;
        }
    }
    void feed(const char *p, size_t n) {
        const char *end = p + n;
        if (!carry.empty()) {
            const char *nl = static_cast<const char *>(memchr(p, '\n', n));
            if (!nl) { carry.append(p, n); return; }
            carry.append(p, nl - p);
            line(carry.data(), carry.size());
            carry.clear();
            p = nl + 1;
        }
        while (p < end) {
            const char *nl = static_cast<const char *>(memchr(p, '\n', end - p));
            if (!nl) { carry.assign(p, end - p); return; }
            line(p, nl - p);
            p = nl + 1;
        }
    }
    void finish() {
        if (!carry.empty()) { line(carry.data(), carry.size()); carry.clear(); }
        if (fmt == 1) out.end_record();
    }
};

struct Unit {   // a whole gzip file, or a byte range of a mapped plain file
    int file = 0;
    const char *begin = nullptr, *end = nullptr;   // plain
    std::string path;                              // gzip
    int fmt = 0;
};

struct Mapped {
    const char *p = nullptr;
    size_t n = 0;
};

bool is_gzip(const std::string &path) {
    FILE *f = fopen(path.c_str(), "rb");
    if (!f) throw std::runtime_error("ERROR: cannot open " + path);
    unsigned char m[2] = {0, 0};
    size_t got = fread(m, 1, 2, f);
    fclose(f);
    return got == 2 && m[0] == 0x1f && m[1] == 0x8b;
}

// first record start at or after p (p itself if it is one); end if none
const char *next_record(const char *base, const char *p, const char *end, int fmt) {
    if (p <= base) return base;
    // move to a line start
    if (p[-1] != '\n') {
        p = static_cast<const char *>(memchr(p, '\n', end - p));
        if (!p) return end;
        ++p;
    }
    while (p < end) {
        if (fmt == 1) {
            if (*p == '>') return p;
        } else if (*p == '@') {
            // a record header is followed by the sequence line and a '+' line; a quality line that
            // starts with '@' is followed by a header and a SEQUENCE line (bases, never '+')
            const char *l1 = static_cast<const char *>(memchr(p, '\n', end - p));
            const char *l2 = l1 ? static_cast<const char *>(memchr(l1 + 1, '\n', end - l1 - 1)) : nullptr;
            if (!l2 || l2 + 1 >= end) return p;
            if (l2[1] == '+') return p;
        }
        p = static_cast<const char *>(memchr(p, '\n', end - p));
        if (!p) return end;
        ++p;
    }
    return end;
}

}  // namespace

ReadFeeder::ReadFeeder(Sink sink, size_t num_threads, size_t batch_bytes, Alloc alloc, Free free_fn, bool checksum)
    : sink_(std::move(sink)), threads_(num_threads ? num_threads : 1), batch_bytes_(batch_bytes),
      alloc_(alloc ? alloc : +[](size_t n) { return malloc(n); }),
      free_(free_fn ? free_fn : +[](void *p) { free(p); }), checksum_(checksum) {}

FeederStats ReadFeeder::run(const std::vector<std::string> &files) {
    auto t0 = std::chrono::steady_clock::now();
    std::vector<Unit> units;
    std::vector<Mapped> maps;
    FeederStats total;
    for (size_t f = 0; f < files.size(); ++f) {
        total.n_files += 1;
        if (is_gzip(files[f])) {
            Unit u;
            u.file = (int)f;
            u.path = files[f];
            units.push_back(u);
            struct stat sb;
            if (!stat(files[f].c_str(), &sb)) total.input_bytes += sb.st_size;
            continue;
        }
        int fd = open(files[f].c_str(), O_RDONLY);
        struct stat sb;
        if (fd < 0 || fstat(fd, &sb)) throw std::runtime_error("ERROR: cannot open " + files[f]);
        total.input_bytes += sb.st_size;
        if (sb.st_size == 0) { close(fd); continue; }
        void *m = mmap(nullptr, sb.st_size, PROT_READ, MAP_PRIVATE, fd, 0);
        close(fd);
        if (m == MAP_FAILED) throw std::runtime_error("ERROR: cannot map " + files[f]);
        madvise(m, sb.st_size, MADV_SEQUENTIAL);
        const char *base = static_cast<const char *>(m), *end = base + sb.st_size;
        maps.push_back({base, (size_t)sb.st_size});
        const char *q = base;
        while (q < end && (*q == '\n' || *q == '\r')) ++q;
        const int fmt = q < end && *q == '@' ? 2 : 1;
        const char *cur = base;
        while (cur < end) {
            const char *nominal = cur + kRangeBytes < end ? cur + kRangeBytes : end;
            const char *stop = nominal < end ? next_record(base, nominal, end, fmt) : end;
            Unit u;
            u.file = (int)f;
            u.begin = cur;
            u.end = stop;
            u.fmt = fmt;
            units.push_back(u);
            cur = stop;
        }
    }

    std::atomic<size_t> next{0};
    std::mutex mu;
    std::string error;
    auto worker = [&]() {
        Batch b;
        b.alloc = alloc_;
        b.free_fn = free_;
        b.sink = &sink_;
        b.flush_at = batch_bytes_;
        b.want_sum = checksum_;
        try {
            for (size_t i = next++; i < units.size(); i = next++) {
                const Unit &u = units[i];
                RecordParser rp(b);
                if (u.begin) {
                    rp.fmt = u.fmt;
                    rp.feed(u.begin, u.end - u.begin);
                } else {
                    gzFile g = gzopen(u.path.c_str(), "rb");
                    if (!g) throw std::runtime_error("ERROR: cannot open " + u.path);
                    gzbuffer(g, 1 << 20);
                    std::vector<char> buf(size_t(4) << 20);
                    int n;
                    while ((n = gzread(g, buf.data(), (unsigned)buf.size())) > 0) rp.feed(buf.data(), n);
                    const bool bad = n < 0;
                    gzclose(g);
                    if (bad) throw std::runtime_error("ERROR: corrupt gzip stream in " + u.path);
                }
                rp.finish();
            }
            b.flush();
        } catch (const std::exception &e) {
            std::lock_guard<std::mutex> lk(mu);
            if (error.empty()) error = e.what();
            next = units.size();
        }
        std::lock_guard<std::mutex> lk(mu);
        total.n_reads += b.st.n_reads;
        total.n_bases += b.st.n_bases;
        total.n_batches += b.st.n_batches;
        total.checksum += b.st.checksum;
    };
    const size_t nt = std::min(threads_, std::max<size_t>(units.size(), 1));
    std::vector<std::thread> th;
    for (size_t t = 1; t < nt; ++t) th.emplace_back(worker);
    worker();
    for (auto &t : th) t.join();
    for (auto &m : maps) munmap(const_cast<char *>(m.p), m.n);
    if (!error.empty()) throw std::runtime_error(error);
    total.seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    return total;
}

}  // namespace gpu
}  // namespace mtg
