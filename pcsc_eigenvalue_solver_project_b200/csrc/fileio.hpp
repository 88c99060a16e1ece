// fileio.hpp -- fast ingest of the reference's matrix files (host side).
//
// The reference reads its text format with `std::ifstream >> value` one token at a time and
// `sparse.insert(r, c)` per entry (src/reader/file_matrix_reader.hpp:33-66 dense, :84-132 sparse, :170-200
// entry point; format in README.md:87-171):
//     dense|sparse  rows cols  [nnz]  then  row-major values (dense)  or  "r c value" triples (sparse);
//     a complex value is two reals.
// Here the file is mapped, split at whitespace into per-thread chunks, tokens are counted (pass 1), and each
// thread then parses its tokens straight into their final slots by token index (pass 2, std::from_chars).
// Sparse triples are assembled into the arrays Eigen::SparseMatrix<Scalar> (ColMajor, int) would hold after
// makeCompressed(): CSC, rows ascending inside a column.  Error texts are the reference's.
// A second, binary container (magic + header + 64-byte aligned raw arrays) can be mapped and handed to
// eigsol_b200_matrix_create_csc / _dense without any parsing or copying.
#pragma once
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <algorithm>
#include <atomic>
#include <charconv>
#include <climits>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <stdexcept>
#include <string>
#include <thread>
#include <vector>

namespace eigsol_b200 {

struct IoError : std::runtime_error {
    using std::runtime_error::runtime_error;
};

struct MappedFile {
    const char* data = nullptr;
    size_t size = 0;
    int fd = -1;
    explicit MappedFile(const char* path) {
        fd = ::open(path, O_RDONLY);
        if (fd < 0) throw IoError(std::string("Impossible to open the file: ") + path);  // reader :172-175
        struct stat st;
        if (fstat(fd, &st) != 0) {
            ::close(fd);
            throw IoError(std::string("Impossible to open the file: ") + path);
        }
        size = size_t(st.st_size);
        if (size) {
            void* p = mmap(nullptr, size, PROT_READ, MAP_PRIVATE, fd, 0);
            if (p == MAP_FAILED) {
                ::close(fd);
                throw IoError(std::string("Impossible to open the file: ") + path);
            }
            data = static_cast<const char*>(p);
            madvise(p, size, MADV_SEQUENTIAL);
        }
    }
    ~MappedFile() {
        if (data) munmap(const_cast<char*>(data), size);
        if (fd >= 0) ::close(fd);
    }
    MappedFile(const MappedFile&) = delete;
    MappedFile& operator=(const MappedFile&) = delete;
};

inline bool is_space(char c) { return c == ' ' || c == '\n' || c == '\t' || c == '\r' || c == '\v' || c == '\f'; }

// Next token in [p, end): returns false at end of input.
inline bool next_token(const char*& p, const char* end, const char*& tb, const char*& te) {
    while (p < end && is_space(*p)) ++p;
    if (p >= end) return false;
    tb = p;
    while (p < end && !is_space(*p)) ++p;
    te = p;
    return true;
}

inline bool parse_int(const char* b, const char* e, long long& v) {
    if (b < e && *b == '+') ++b;
    if (b >= e) return false;
    const auto r = std::from_chars(b, e, v);
    return r.ec == std::errc() && r.ptr == e;
}

inline bool parse_real(const char* b, const char* e, double& v) {
    if (b < e && *b == '+') ++b;
    if (b >= e) return false;
    const auto r = std::from_chars(b, e, v);
    return r.ec == std::errc() && r.ptr == e;
}

// What a parsed file holds: exactly the arrays of Matrix::Dense<Scalar> (column-major) or of a compressed
// Matrix::Sparse<Scalar> (CSC, int32).  Complex values are interleaved (re, im).
struct FileMatrixHost {
    bool dense = true;
    int dtype = 0;  // 0 = f64, 1 = c128
    int64_t rows = 0, cols = 0, nnz = 0;
    std::vector<double> values;   // dense: rows*cols scalars column-major; sparse: nnz scalars
    std::vector<int32_t> outer;   // sparse: cols + 1
    std::vector<int32_t> inner;   // sparse: nnz
};

inline int io_threads(size_t bytes) {
    const char* e = std::getenv("EIGSOL_B200_IO_THREADS");
    int t = e && *e ? std::atoi(e) : int(std::thread::hardware_concurrency());
    t = std::max(1, std::min(t, 32));
    const size_t by_size = bytes / (size_t(1) << 20) + 1;  // at least ~1 MB per thread
    return int(std::min<size_t>(size_t(t), by_size));
}

// Runs f(t) on nthreads threads (t = 0 on the caller).
template <class F> inline void parallel_for_threads(int nthreads, F&& f) {
    std::vector<std::thread> pool;
    for (int t = 1; t < nthreads; ++t) pool.emplace_back([&f, t] { f(t); });
    f(0);
    for (auto& th : pool) th.join();
}


// -------------------------------------------------------------------------------------------------
// Text format.
// -------------------------------------------------------------------------------------------------
struct TokenChunks {
    std::vector<const char*> begin;   // nthreads + 1 chunk boundaries (at whitespace)
    std::vector<int64_t> first_token; // nthreads + 1: index of the first token of each chunk
};

inline TokenChunks count_tokens(const char* b, const char* e, int nthreads) {
    TokenChunks c;
    c.begin.resize(size_t(nthreads) + 1);
    c.first_token.assign(size_t(nthreads) + 1, 0);
    const size_t len = size_t(e - b);
    for (int t = 0; t <= nthreads; ++t) {
        const char* p = b + len * size_t(t) / size_t(nthreads);
        if (t > 0 && t < nthreads)
            while (p < e && !is_space(*p)) ++p;  // never split a token
        c.begin[size_t(t)] = t == nthreads ? e : p;
    }
    std::vector<int64_t> cnt(static_cast<size_t>(nthreads), 0);
    parallel_for_threads(nthreads, [&](int t) {
        const char* p = c.begin[size_t(t)];
        const char* end = c.begin[size_t(t) + 1];
        const char *tb, *te;
        int64_t k = 0;
        while (next_token(p, end, tb, te)) ++k;
        cnt[size_t(t)] = k;
    });
    for (int t = 0; t < nthreads; ++t) c.first_token[size_t(t) + 1] = c.first_token[size_t(t)] + cnt[size_t(t)];
    return c;
}

// dtype: the Scalar the caller reads the file as (the file itself does not say): 0 = double, 1 = complex<double>.
inline FileMatrixHost read_matrix_text(const char* path, int dtype) {
    MappedFile f(path);
    const char* p = f.data;
    const char* end = f.data + f.size;
    const char *tb, *te;
    FileMatrixHost m;
    m.dtype = dtype;
    if (!f.data || !next_token(p, end, tb, te)) throw IoError("Failed to read matrix storage type");        // :178-180
    const std::string kind(tb, te);
    if (kind == "dense") m.dense = true;
    else if (kind == "sparse") m.dense = false;
    else throw IoError("Unknown storage type: " + kind);                                                      // :155
    long long rows = 0, cols = 0;
    if (!next_token(p, end, tb, te) || !parse_int(tb, te, rows) || !next_token(p, end, tb, te) || !parse_int(tb, te, cols))
        throw IoError("Failed to read matrix dimensions");                                                     // :185-187
    if (rows <= 0 || cols <= 0) throw IoError("Matrix dimensions must be positive");                        // :188-190
    if (rows > INT32_MAX || cols > INT32_MAX) throw IoError("Matrix dimensions exceed int");
    m.rows = rows;
    m.cols = cols;
    const int per = dtype ? 2 : 1;  // reals per value

    if (m.dense) {
        const int64_t need = int64_t(rows) * int64_t(cols) * per;
        const int nt = io_threads(size_t(end - p));
        const TokenChunks ch = count_tokens(p, end, nt);
        m.values.assign(size_t(need), 0.0);
        std::atomic<int64_t> bad{INT64_MAX};
        parallel_for_threads(nt, [&](int t) {
            const char* q = ch.begin[size_t(t)];
            const char* qe = ch.begin[size_t(t) + 1];
            int64_t k = ch.first_token[size_t(t)];
            const char *b2, *e2;
            while (k < need && next_token(q, qe, b2, e2)) {
                double v;
                if (!parse_real(b2, e2, v)) {
                    int64_t cur = bad.load();
                    while (k < cur && !bad.compare_exchange_weak(cur, k)) {}
                    return;
                }
                // token k = value (r, c), part = k % per, row-major in the file -> column-major in memory
                const int64_t el = k / per, part = k % per;
                const int64_t r = el / cols, c = el % cols;
                m.values[size_t((c * rows + r) * per + part)] = v;
                ++k;
            }
        });
        const int64_t have = ch.first_token[size_t(nt)];
        if (bad.load() != INT64_MAX || have < need)
            throw IoError(dtype ? "Failed to read complex entry in dense matrix" : "Failed to read scalar entry in dense matrix");
        return m;
    }

    // sparse: nnz, then nnz entries "r c value"
    long long nnz = 0;
    if (!next_token(p, end, tb, te) || !parse_int(tb, te, nnz))
        throw IoError("Cannot read number of non-zero entries in the sparse matrix");                         // :91-93
    if (nnz <= 0) throw IoError("number of non-zero entries must be positive in a sparse matrix");            // :94-96
    if (nnz > INT32_MAX) throw IoError("number of non-zero entries exceeds int");
    m.nnz = nnz;
    const int per_entry = 2 + per;
    const int64_t need = int64_t(nnz) * per_entry;
    const int nt = io_threads(size_t(end - p));
    const TokenChunks ch = count_tokens(p, end, nt);
    const size_t nz = size_t(nnz);
    std::vector<int32_t> er(nz), ec(nz);
    std::vector<double> ev(nz * size_t(per));
    // first offending token (file order) and what was wrong with it: 1 = index unreadable, 2 = out of range,
    // 3 = value unreadable -- the reference reports whichever comes first in the file (:101-127)
    std::atomic<int64_t> bad{INT64_MAX};
    std::vector<int> why(static_cast<size_t>(nt), 0);
    std::vector<int64_t> where(static_cast<size_t>(nt), INT64_MAX);
    parallel_for_threads(nt, [&](int t) {
        const char* q = ch.begin[size_t(t)];
        const char* qe = ch.begin[size_t(t) + 1];
        int64_t k = ch.first_token[size_t(t)];
        const char *b2, *e2;
        auto fail = [&](int w) {
            why[size_t(t)] = w;
            where[size_t(t)] = k;
            int64_t cur = bad.load();
            while (k < cur && !bad.compare_exchange_weak(cur, k)) {}
        };
        while (k < need && next_token(q, qe, b2, e2)) {
            const int64_t entry = k / per_entry, field = k % per_entry;
            if (field < 2) {
                long long idx;
                if (!parse_int(b2, e2, idx)) return fail(1);
                const long long lim = field == 0 ? rows : cols;
                if (idx < 0 || idx >= lim) {
                    // the range test comes after BOTH indices were read (:108-110): attribute it to the column token
                    k = entry * per_entry + 1;
                    return fail(2);
                }
                (field == 0 ? er : ec)[size_t(entry)] = int32_t(idx);
            } else {
                double v;
                if (!parse_real(b2, e2, v)) return fail(3);
                ev[size_t(entry) * size_t(per) + size_t(field - 2)] = v;
            }
            ++k;
        }
    });
    const int64_t have = ch.first_token[size_t(nt)];
    int64_t first_bad = INT64_MAX;
    int reason = 0;
    for (int t = 0; t < nt; ++t)
        if (where[size_t(t)] < first_bad) {
            first_bad = where[size_t(t)];
            reason = why[size_t(t)];
        }
    if (have < need && have < first_bad) {  // the file ends first
        first_bad = have;
        reason = (have % per_entry) < 2 ? 1 : 3;
    }
    if (reason == 1) throw IoError("Error when trying to read indices in sparse matrix");
    if (reason == 2) throw IoError("Sparse indices out of range");
    if (reason == 3) throw IoError("Failed to read scalar entry in sparse matrix");

    // CSC with rows ascending inside a column (what insert() + makeCompressed() leave, :120-131)
    // counting sort by column, every phase split over the threads by COLUMN range: a thread scans all entries
    // (sequential reads) and handles the ones whose column is its own, so file order is kept inside a column
    // and no two threads touch the same counter
    const int nt2 = io_threads(size_t(nnz) * 16);
    m.outer.assign(size_t(cols) + 1, 0);
    parallel_for_threads(nt2, [&](int t) {
        const int64_t j0 = cols * t / nt2, j1 = cols * (t + 1) / nt2;
        for (int64_t k = 0; k < nnz; ++k) {
            const int32_t c = ec[size_t(k)];
            if (c >= j0 && c < j1) ++m.outer[size_t(c) + 1];
        }
    });
    for (int64_t j = 0; j < cols; ++j) m.outer[size_t(j) + 1] += m.outer[size_t(j)];
    std::vector<int32_t> cursor(m.outer.begin(), m.outer.end() - 1);
    std::vector<int32_t> order(nz);
    parallel_for_threads(nt2, [&](int t) {
        const int64_t j0 = cols * t / nt2, j1 = cols * (t + 1) / nt2;
        for (int64_t k = 0; k < nnz; ++k) {
            const int32_t c = ec[size_t(k)];
            if (c >= j0 && c < j1) order[size_t(cursor[size_t(c)]++)] = int32_t(k);
        }
    });
    std::atomic<int> dup{0};
    parallel_for_threads(nt2, [&](int t) {
        const int64_t j0 = cols * t / nt2, j1 = cols * (t + 1) / nt2;
        for (int64_t j = j0; j < j1; ++j) {
            int32_t* ob = order.data() + m.outer[size_t(j)];
            int32_t* oe = order.data() + m.outer[size_t(j) + 1];
            std::sort(ob, oe, [&](int32_t a, int32_t b) { return er[size_t(a)] < er[size_t(b)]; });
            for (int32_t* o = ob; o + 1 < oe; ++o)
                if (er[size_t(o[0])] == er[size_t(o[1])]) dup.store(1);
        }
    });
    if (dup.load()) throw IoError("Duplicate (row, column) entry in sparse matrix");  // the reference's insert() forbids it
    m.inner.resize(size_t(nnz));
    m.values.resize(size_t(nnz) * size_t(per));
    parallel_for_threads(nt2, [&](int t) {
        const int64_t q0 = nnz * t / nt2, q1 = nnz * (t + 1) / nt2;
        for (int64_t q = q0; q < q1; ++q) {
            const int32_t k = order[size_t(q)];
            m.inner[size_t(q)] = er[size_t(k)];
            for (int c2 = 0; c2 < per; ++c2)
                m.values[size_t(q) * size_t(per) + size_t(c2)] = ev[size_t(k) * size_t(per) + size_t(c2)];
        }
    });
    return m;
}

// -------------------------------------------------------------------------------------------------
// Binary container: 64-byte header, then the arrays, each 64-byte aligned.
//   sparse: outer int32[cols+1] | inner int32[nnz] | values scalar[nnz]        dense: values scalar[rows*cols]
// -------------------------------------------------------------------------------------------------
struct BinHeader {
    char magic[8];  // "EIGSOLB2"
    int32_t version, dtype, dense, reserved;
    int64_t rows, cols, nnz;
    int64_t pad[2];
};
static_assert(sizeof(BinHeader) == 64, "header is one cache line");

inline size_t align64(size_t x) { return (x + 63) & ~size_t(63); }

inline void write_matrix_bin(const char* path, int dtype, bool dense, int64_t rows, int64_t cols, int64_t nnz,
                             const int32_t* outer, const int32_t* inner, const void* values) {
    FILE* f = std::fopen(path, "wb");
    if (!f) throw IoError(std::string("Impossible to open the file: ") + path);
    BinHeader h;
    std::memset(&h, 0, sizeof h);
    std::memcpy(h.magic, "EIGSOLB2", 8);
    h.version = 1;
    h.dtype = dtype;
    h.dense = dense ? 1 : 0;
    h.rows = rows;
    h.cols = cols;
    h.nnz = dense ? rows * cols : nnz;
    const size_t s = dtype ? 16 : 8;
    static const char zeros[64] = {0};
    auto put = [&](const void* p, size_t n) {
        bool ok = n == 0 || std::fwrite(p, 1, n, f) == n;
        const size_t padn = align64(n) - n;
        ok = ok && (padn == 0 || std::fwrite(zeros, 1, padn, f) == padn);
        if (!ok) {
            std::fclose(f);
            throw IoError(std::string("Failed to write the file: ") + path);
        }
    };
    put(&h, sizeof h);
    if (!dense) {
        put(outer, sizeof(int32_t) * size_t(cols + 1));
        put(inner, sizeof(int32_t) * size_t(nnz));
    }
    put(values, s * size_t(h.nnz));
    if (std::fclose(f) != 0) throw IoError(std::string("Failed to write the file: ") + path);
}

// A mapped binary container: the pointers go straight into eigsol_b200_matrix_create_csc / _dense.
struct MappedMatrix {
    MappedFile file;
    BinHeader h;
    const int32_t* outer = nullptr;
    const int32_t* inner = nullptr;
    const void* values = nullptr;
    explicit MappedMatrix(const char* path) : file(path) {
        if (file.size < sizeof(BinHeader)) throw IoError("Not an EIGSOLB2 matrix file");
        std::memcpy(&h, file.data, sizeof h);
        if (std::memcmp(h.magic, "EIGSOLB2", 8) != 0 || h.version != 1 || (h.dtype != 0 && h.dtype != 1) || h.rows <= 0 ||
            h.cols <= 0 || h.nnz < 0)
            throw IoError("Not an EIGSOLB2 matrix file");
        // The header is untrusted input: bound every count before it sizes an offset, compare offsets with the real
        // file size (the last array may lack its alignment padding, never its payload), and for sparse files check
        // that the column pointer array agrees with the header -- the arrays go straight to the upload path, which
        // sizes its reads from outer[cols].
        if (h.dense != 0 && h.dense != 1) throw IoError("Not an EIGSOLB2 matrix file");
        if (h.rows > INT32_MAX || h.cols > INT32_MAX) throw IoError("Truncated/corrupt EIGSOLB2 matrix file");
        if (h.dense ? (h.nnz != h.rows * h.cols) : (h.nnz > INT32_MAX)) throw IoError("Truncated/corrupt EIGSOLB2 matrix file");
        const uint64_t s = h.dtype ? 16 : 8;
        const uint64_t size = file.size;
        uint64_t off = sizeof(BinHeader);
        auto take = [&](uint64_t bytes, bool last) {  // consumes one array; returns its offset
            const uint64_t at = off;
            if (bytes > size || at > size - bytes) throw IoError("Truncated/corrupt EIGSOLB2 matrix file");
            off = at + bytes;
            if (!last) {
                const uint64_t padded = (off + 63) & ~uint64_t(63);
                if (padded > size) throw IoError("Truncated/corrupt EIGSOLB2 matrix file");
                off = padded;
            }
            return at;
        };
        if (!h.dense) {
            outer = reinterpret_cast<const int32_t*>(file.data + take(sizeof(int32_t) * (uint64_t(h.cols) + 1), false));
            inner = reinterpret_cast<const int32_t*>(file.data + take(sizeof(int32_t) * uint64_t(h.nnz), false));
        }
        values = file.data + take(s * uint64_t(h.nnz), true);
        if (!h.dense && (outer[0] != 0 || int64_t(outer[h.cols]) != h.nnz))
            throw IoError("Truncated/corrupt EIGSOLB2 matrix file");
    }
};

}  // namespace eigsol_b200
