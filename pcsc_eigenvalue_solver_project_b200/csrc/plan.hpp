// plan.hpp -- host-side planning for the B200 power-iteration path (no CUDA in this file).
//
// Everything the runtime decides BEFORE a kernel is launched lives here so that it can be unit-tested
// on a CPU-only box through the C ABI (include/eigsol_b200.h, "host-side planning"):
//   * CSC -> CSR transposition of what Eigen::SparseMatrix<Scalar> holds
//     (reference: src/matrix/matrix.hpp:43-44, eigen/Eigen/src/SparseCore/SparseMatrix.h:171-193),
//     keeping columns ascending inside every row = the order the reference accumulates in
//     (eigen/Eigen/src/SparseCore/SparseDenseProduct.h:101-110);
//   * row-length statistics and the kernel choice made from them;
//   * the sliced (SELL) layout and the row blocks of the streaming CSR kernel;
//   * the nnz-balanced row partition and the halo (remote column) plan of the multi-GPU path.
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

namespace eigsol_b200 {

struct PlanError : std::runtime_error {
    using std::runtime_error::runtime_error;
};

// ---------------------------------------------------------------------------------------------
// CSR container on the host.  Scalars are stored as raw doubles: 1 per entry (f64) or 2 (c128).
// ---------------------------------------------------------------------------------------------
struct CsrHost {
    int dtype = 0;  // 0 = f64, 1 = c128
    int64_t rows = 0, cols = 0;
    std::vector<int32_t> rowptr;  // rows + 1
    std::vector<int32_t> colidx;  // nnz
    std::vector<double> vals;     // nnz * (1 + dtype)
    int64_t nnz() const { return rowptr.empty() ? 0 : rowptr.back(); }
    int spd() const { return dtype ? 2 : 1; }  // doubles per scalar
};

inline void validate_csc(int64_t rows, int64_t cols, const int32_t* outer, const int32_t* inner) {
    if (rows < 0 || cols < 0) throw PlanError("negative matrix dimension");
    if (rows > INT32_MAX || cols > INT32_MAX) throw PlanError("matrix dimension exceeds int32 indices");
    if (!outer) throw PlanError("null outer index array");
    if (outer[0] != 0) throw PlanError("CSC outer index must start at 0 (compressed storage expected)");
    for (int64_t j = 0; j < cols; ++j)
        if (outer[j + 1] < outer[j]) throw PlanError("CSC outer index not monotone");
    const int64_t nnz = outer[cols];
    if (nnz > 0 && !inner) throw PlanError("null inner index array");
    for (int64_t p = 0; p < nnz; ++p)
        if (inner[p] < 0 || inner[p] >= rows) throw PlanError("CSC inner index out of range");
}

// Counting-sort transposition, O(nnz + rows).  Columns are visited in ascending order, so entries
// land in each CSR row with ascending column index (and duplicates keep their input order).
inline void csc_to_csr_raw(int dtype, int64_t rows, int64_t cols, const int32_t* outer,
                           const int32_t* inner, const double* vals, int32_t* rowptr, int32_t* colidx,
                           double* vals_out) {
    const int64_t nnz = outer[cols];
    std::fill(rowptr, rowptr + rows + 1, 0);
    for (int64_t p = 0; p < nnz; ++p) ++rowptr[inner[p] + 1];
    for (int64_t i = 0; i < rows; ++i) rowptr[i + 1] += rowptr[i];
    std::vector<int32_t> cursor(rowptr, rowptr + rows);
    if (dtype == 0) {
        for (int64_t j = 0; j < cols; ++j)
            for (int32_t p = outer[j]; p < outer[j + 1]; ++p) {
                const int32_t q = cursor[inner[p]]++;
                colidx[q] = static_cast<int32_t>(j);
                vals_out[q] = vals[p];
            }
    } else {
        for (int64_t j = 0; j < cols; ++j)
            for (int32_t p = outer[j]; p < outer[j + 1]; ++p) {
                const int32_t q = cursor[inner[p]]++;
                colidx[q] = static_cast<int32_t>(j);
                vals_out[2 * int64_t(q)] = vals[2 * int64_t(p)];
                vals_out[2 * int64_t(q) + 1] = vals[2 * int64_t(p) + 1];
            }
    }
}

inline CsrHost csc_to_csr(int dtype, int64_t rows, int64_t cols, const int32_t* outer,
                          const int32_t* inner, const void* values) {
    validate_csc(rows, cols, outer, inner);
    CsrHost m;
    m.dtype = dtype;
    m.rows = rows;
    m.cols = cols;
    const int64_t nnz = outer[cols];
    m.rowptr.resize(size_t(rows) + 1);
    m.colidx.resize(size_t(nnz));
    m.vals.resize(size_t(nnz) * m.spd());
    csc_to_csr_raw(dtype, rows, cols, outer, inner, static_cast<const double*>(values), m.rowptr.data(),
                   m.colidx.data(), m.vals.data());
    return m;
}

inline void validate_csr_rows(int64_t n, int64_t row_count, const int32_t* rowptr, const int32_t* colidx) {
    if (n < 0 || row_count < 0 || row_count > n) throw PlanError("bad CSR row block dimensions");
    if (n > INT32_MAX) throw PlanError("matrix dimension exceeds int32 indices");
    if (!rowptr) throw PlanError("null rowptr");
    if (rowptr[0] != 0) throw PlanError("CSR rowptr must start at 0");
    for (int64_t i = 0; i < row_count; ++i)
        if (rowptr[i + 1] < rowptr[i]) throw PlanError("CSR rowptr not monotone");
    const int64_t nnz = rowptr[row_count];
    if (nnz > 0 && !colidx) throw PlanError("null colidx");
    for (int64_t p = 0; p < nnz; ++p)
        if (colidx[p] < 0 || colidx[p] >= n) throw PlanError("CSR column index out of range");
}

// ---------------------------------------------------------------------------------------------
// Row-length statistics and kernel choice.
// ---------------------------------------------------------------------------------------------
struct RowStats {
    int64_t rows = 0, nnz = 0;
    double mean = 0, max = 0, min = 0, cv = 0;
    double sell_fill = 1.0;  // padded entries / nnz for a 64-row slice (1.0 = no padding)
    double near_diag = 1.0;  // fraction of the entries with |column - row| <= 1 (the tridiagonal core every PDE stencil
                             // has); 1.0 when the column indices were not looked at
};

// colidx (optional): GLOBAL column indices of the rows, row0 = global index of the first row
inline RowStats row_stats(int64_t rows, const int32_t* rowptr, const int32_t* colidx = nullptr, int64_t row0 = 0) {
    RowStats s;
    s.rows = rows;
    if (rows == 0) return s;
    s.nnz = rowptr[rows];
    s.mean = double(s.nnz) / double(rows);
    double var = 0;
    int32_t mx = 0, mn = INT32_MAX;
    int64_t padded = 0;
    int32_t slice_max = 0;
    for (int64_t i = 0; i < rows; ++i) {
        const int32_t len = rowptr[i + 1] - rowptr[i];
        mx = std::max(mx, len);
        mn = std::min(mn, len);
        const double d = len - s.mean;
        var += d * d;
        slice_max = std::max(slice_max, len);
        if ((i & 63) == 63 || i == rows - 1) {
            padded += int64_t(slice_max) * 64;
            slice_max = 0;
        }
    }
    s.max = mx;
    s.min = mn;
    s.cv = s.mean > 0 ? std::sqrt(var / double(rows)) / s.mean : 0.0;
    s.sell_fill = s.nnz > 0 ? double(padded) / double(s.nnz) : 1.0;
    if (colidx && s.nnz > 0) {
        int64_t near = 0;
        for (int64_t i = 0; i < rows; ++i)
            for (int32_t p = rowptr[i]; p < rowptr[i + 1]; ++p) {
                const int64_t d = int64_t(colidx[p]) - (row0 + i);
                near += (d >= -1 && d <= 1) ? 1 : 0;
            }
        s.near_diag = double(near) / double(s.nnz);
    }
    return s;
}

// Kernel ids mirror include/eigsol_b200.h.
enum KernelId { K_AUTO = 0, K_CSR_SCALAR = 1, K_CSR_VECTOR = 2, K_CSR_STREAM = 3, K_SELL = 4, K_GEMV = 5, K_CSR_WINDOW = 6 };

struct KernelChoice {
    int kernel = K_CSR_SCALAR;
    int lanes = 0;  // CSR_VECTOR: lanes per row; SELL: rows per lane
};

// The dispatch rule (DESIGN.md "kernel choice"; every threshold is backed by a measured sweep on the B200,
// profiles/r02/variants_*.jsonl, and pinned by tests/test_host_plan.py::test_dispatch_rule_picks_the_measured_winner):
//   very short rows, no long outliers, AND
//     stencil-like (>= 25 % of the entries on the tridiagonal: 2D/3D Laplacians, 43-60 %)  -> CSR_SCALAR: the gathers of a
//         warp fall into a handful of lines, thread-per-row off the CSR arrays runs at 98 % of the copy bandwidth and
//         beats SELL by 7-15 %
//     or irregular enough that SELL would pad heavily (fill > 2.25, e.g. Poisson(4) rows)    -> CSR_SCALAR
//   short / medium rows, padding <= 2.5x -> SELL (R = 2 rows per lane for f64, 1 for c128): every stream coalesced;
//         wins from 4 regular entries per row up to 64 (random-offset bands: 8 % at 4, 46 % at 8, 20-60 % at 12..64)
//   short rows with long outliers        -> CSR_STREAM (long rows reduced by a whole CTA)
//   long rows                            -> CSR_VECTOR, lanes = pow2 <= mean clipped to [2, 32]
inline KernelChoice choose_kernel(int dtype, const RowStats& s) {
    KernelChoice c;
    if (s.rows == 0 || s.nnz == 0) {
        c.kernel = K_CSR_SCALAR;
        return c;
    }
    if (s.mean <= 8.0 && s.max <= 32.0 && (s.near_diag >= 0.25 || s.sell_fill > 2.25)) {
        c.kernel = K_CSR_SCALAR;
        return c;
    }
    if (s.mean <= 64.0 && s.sell_fill <= 2.5) {
        c.kernel = K_SELL;
        c.lanes = dtype ? 1 : 2;
        return c;
    }
    if (s.mean <= 24.0) {
        c.kernel = K_CSR_STREAM;
        return c;
    }
    c.kernel = K_CSR_VECTOR;
    int lanes = 2;
    while (lanes < 32 && lanes * 2 <= s.mean) lanes *= 2;
    c.lanes = lanes;
    return c;
}

// ---------------------------------------------------------------------------------------------
// SELL layout: slices of H = 32*R consecutive rows (R rows per lane).  Inside a slice entry j of the
// row handled by (lane l, sub-row r) sits at  slice_off + (j*32 + l)*R + r, so one warp instruction
// reads 32*R consecutive values / column indices.  Padding entries have column -1 and value 0.
// ---------------------------------------------------------------------------------------------
struct SellHost {
    int rows_per_lane = 1;
    int64_t rows = 0, nslices = 0;
    std::vector<int64_t> slice_off;  // nslices + 1, in entries
    std::vector<int32_t> colidx;
    std::vector<double> vals;
    int64_t padded() const { return slice_off.empty() ? 0 : slice_off.back(); }
};

inline SellHost build_sell(const CsrHost& a, int rows_per_lane) {
    SellHost s;
    const int R = rows_per_lane;
    const int H = 32 * R;
    s.rows_per_lane = R;
    s.rows = a.rows;
    s.nslices = (a.rows + H - 1) / H;
    s.slice_off.assign(size_t(s.nslices) + 1, 0);
    for (int64_t sl = 0; sl < s.nslices; ++sl) {
        int32_t w = 0;
        const int64_t r0 = sl * H, r1 = std::min<int64_t>(a.rows, r0 + H);
        for (int64_t i = r0; i < r1; ++i) w = std::max(w, a.rowptr[i + 1] - a.rowptr[i]);
        s.slice_off[sl + 1] = s.slice_off[sl] + int64_t(w) * H;
    }
    const int spd = a.spd();
    s.colidx.assign(size_t(s.padded()), -1);
    s.vals.assign(size_t(s.padded()) * spd, 0.0);
    for (int64_t sl = 0; sl < s.nslices; ++sl) {
        const int64_t base = s.slice_off[sl];
        const int64_t r0 = sl * H, r1 = std::min<int64_t>(a.rows, r0 + H);
        for (int64_t i = r0; i < r1; ++i) {
            const int64_t loc = i - r0;
            const int64_t lane = loc / R, sub = loc % R;
            for (int32_t p = a.rowptr[i], j = 0; p < a.rowptr[i + 1]; ++p, ++j) {
                const int64_t q = base + (int64_t(j) * 32 + lane) * R + sub;
                s.colidx[q] = a.colidx[p];
                for (int d = 0; d < spd; ++d) s.vals[q * spd + d] = a.vals[int64_t(p) * spd + d];
            }
        }
    }
    return s;
}

// ---------------------------------------------------------------------------------------------
// Row blocks for the streaming CSR kernel: consecutive rows whose entries fit one shared-memory
// tile (<= max_nnz entries and <= max_rows rows).  A single row longer than the tile gets a block of
// its own and is reduced by the whole CTA.
// ---------------------------------------------------------------------------------------------
inline std::vector<int32_t> build_row_blocks(int64_t rows, const int32_t* rowptr, int max_nnz, int max_rows) {
    std::vector<int32_t> rb;
    rb.push_back(0);
    int64_t i = 0;
    while (i < rows) {
        int64_t j = i;
        // entries are addressed relative to the 4-aligned start of the block, so reserve 3 slots
        const int64_t base = rowptr[i] & ~int64_t(3);
        while (j < rows && (j - i) < max_rows && (rowptr[j + 1] - base) <= max_nnz) ++j;
        if (j == i) j = i + 1;  // one long row
        rb.push_back(int32_t(j));
        i = j;
    }
    return rb;
}

// ---------------------------------------------------------------------------------------------
// Multi-GPU: nnz-balanced contiguous row partition, and the halo plan of one rank.
// ---------------------------------------------------------------------------------------------
inline std::vector<int64_t> partition_rows(int64_t rows, const int32_t* rowptr, int nranks) {
    if (nranks < 1) throw PlanError("nranks must be >= 1");
    std::vector<int64_t> rb(size_t(nranks) + 1, 0);
    const int64_t nnz = rows > 0 ? rowptr[rows] : 0;
    rb[nranks] = rows;
    for (int p = 1; p < nranks; ++p) {
        if (nnz == 0) {
            rb[p] = rows * p / nranks;
            continue;
        }
        // first row whose prefix reaches p/nranks of the entries
        const int64_t target = (nnz * p + nranks - 1) / nranks;
        const int32_t* it = std::lower_bound(rowptr, rowptr + rows + 1, target,
                                             [](int32_t v, int64_t t) { return int64_t(v) < t; });
        int64_t r = it - rowptr;
        r = std::min<int64_t>(std::max<int64_t>(r, rb[p - 1]), rows);
        rb[p] = r;
    }
    for (int p = 1; p <= nranks; ++p) rb[p] = std::max(rb[p], rb[p - 1]);
    return rb;
}

inline std::vector<int64_t> partition_even(int64_t rows, int nranks) {
    std::vector<int64_t> rb(size_t(nranks) + 1, 0);
    for (int p = 0; p <= nranks; ++p) rb[p] = rows * p / nranks;
    return rb;
}

struct HaloPlan {
    std::vector<int32_t> cols;        // sorted distinct remote global columns this rank reads
    std::vector<int64_t> peer_count;  // how many of them each peer owns (0 for self)
    std::vector<int64_t> peer_off;    // prefix of peer_count (nranks + 1)
};

inline int owner_of(const std::vector<int64_t>& row_begin, int64_t col) {
    const auto it = std::upper_bound(row_begin.begin(), row_begin.end(), col);
    return int(it - row_begin.begin()) - 1;
}

inline HaloPlan halo_plan(int64_t row_count, const int32_t* rowptr, const int32_t* colidx,
                          const std::vector<int64_t>& row_begin, int rank) {
    const int nranks = int(row_begin.size()) - 1;
    const int64_t lo = row_begin[rank], hi = row_begin[rank + 1];
    HaloPlan h;
    const int64_t nnz = row_count > 0 ? rowptr[row_count] : 0;
    for (int64_t p = 0; p < nnz; ++p) {
        const int32_t c = colidx[p];
        if (c < lo || c >= hi) h.cols.push_back(c);
    }
    std::sort(h.cols.begin(), h.cols.end());
    h.cols.erase(std::unique(h.cols.begin(), h.cols.end()), h.cols.end());
    h.peer_count.assign(size_t(nranks), 0);
    for (int32_t c : h.cols) ++h.peer_count[owner_of(row_begin, c)];
    h.peer_off.assign(size_t(nranks) + 1, 0);
    for (int p = 0; p < nranks; ++p) h.peer_off[p + 1] = h.peer_off[p] + h.peer_count[p];
    return h;
}

// Rewrites global column indices of a rank's rows into the rank's extended local numbering:
// [0, n_local) = owned entries, n_local_padded + k = k-th halo entry.
inline void localize_columns(int64_t nnz, int32_t* colidx, int64_t lo, int64_t hi, int64_t halo_base,
                             const std::vector<int32_t>& halo_cols) {
    for (int64_t p = 0; p < nnz; ++p) {
        const int32_t c = colidx[p];
        if (c >= lo && c < hi) {
            colidx[p] = int32_t(c - lo);
        } else {
            const auto it = std::lower_bound(halo_cols.begin(), halo_cols.end(), c);
            colidx[p] = int32_t(halo_base + (it - halo_cols.begin()));
        }
    }
}

// Algorithmic bytes of one matvec: everything touched exactly once.
inline int64_t alg_bytes_csr(int dtype, int64_t rows, int64_t nnz) {
    const int64_t ss = dtype ? 16 : 8;
    return nnz * (ss + 4) + 4 * (rows + 1) + 2 * ss * rows;
}
inline int64_t alg_bytes_dense(int dtype, int64_t rows, int64_t cols) {
    const int64_t ss = dtype ? 16 : 8;
    return ss * rows * cols + ss * (rows + cols);
}

}  // namespace eigsol_b200
