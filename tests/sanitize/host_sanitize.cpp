// host_sanitize.cpp -- TEST INFRASTRUCTURE: the host-side code of the B200 path under AddressSanitizer +
// UndefinedBehaviorSanitizer (compute-sanitizer is closed on the GPU pool, SURVEY.md section 5 asks for a host
// substitute).  Exercises, with g++ -fsanitize=address,undefined -fno-sanitize-recover=all:
//   * csrc/plan.hpp     CSC -> CSR, row statistics, kernel choice, SELL layout, row blocks, partition, halo plan,
//                       column localisation -- on regular, ragged, empty-row and single-long-row matrices;
//   * csrc/fileio.hpp   the text reader (dense / sparse, real / complex, shuffled entries, every error message) and the
//                       EIGSOLB2 container (round trip, truncated / corrupt headers);
//   * csrc/blockmath.hpp through tests/blockmath/sim.cpp (Cholesky, triangular inverse, Hessenberg + shifted QR) on
//                       identity, zero, Jordan, clustered and random blocks.
// Any sanitizer report aborts the process; tests/test_sanitizers.py builds and runs it.  Exit code 0 = clean.
#include <cmath>
#include <complex>
#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <random>
#include <string>
#include <vector>

#include "../../pcsc_eigenvalue_solver_project_b200/csrc/fileio.hpp"
#include "../../pcsc_eigenvalue_solver_project_b200/csrc/plan.hpp"

using namespace eigsol_b200;

extern "C" {
int bm_chol(const double* w, int k, double* r_out);
void bm_triinv(const double* r, int k, double* t_out);
int bm_eig(const double* h, int k, double* theta_out, double* s_out);
int bm_schur(const double* h, int k, double* t_out, double* z_out);
}

static int g_checks = 0;
#define REQUIRE(c)                                                              \
    do {                                                                        \
        ++g_checks;                                                             \
        if (!(c)) {                                                             \
            std::fprintf(stderr, "REQUIRE failed: %s (%s:%d)\n", #c, __FILE__, __LINE__); \
            std::exit(2);                                                       \
        }                                                                       \
    } while (0)

struct Csc {
    int64_t rows = 0, cols = 0;
    std::vector<int32_t> outer, inner;
    std::vector<double> vals;
};

// random CSC with `mean` entries per column, optional one dense row and some empty columns
static Csc random_csc(int64_t n, double mean, unsigned seed, int dtype, bool long_row) {
    std::mt19937_64 rng(seed);
    Csc m;
    m.rows = m.cols = n;
    m.outer.push_back(0);
    for (int64_t j = 0; j < n; ++j) {
        std::vector<int32_t> r;
        const int len = (j % 7 == 3) ? 0 : int(rng() % uint64_t(2 * mean + 1));
        for (int k = 0; k < len; ++k) r.push_back(int32_t(rng() % uint64_t(n)));
        if (long_row) r.push_back(int32_t(n / 2));
        std::sort(r.begin(), r.end());
        r.erase(std::unique(r.begin(), r.end()), r.end());
        for (int32_t i : r) {
            m.inner.push_back(i);
            for (int d = 0; d <= dtype; ++d) m.vals.push_back(double(rng() % 1000) / 100.0 - 5.0);
        }
        m.outer.push_back(int32_t(m.inner.size()));
    }
    return m;
}

static void plan_suite() {
    for (int dtype = 0; dtype <= 1; ++dtype) {
        for (int variant = 0; variant < 4; ++variant) {
            const int64_t n = variant == 3 ? 1 : 257 + 64 * variant;
            const Csc a = random_csc(n, variant == 1 ? 40.0 : 5.0, 17u + unsigned(variant), dtype, variant == 2);
            const CsrHost c = csc_to_csr(dtype, a.rows, a.cols, a.outer.data(), a.inner.data(), a.vals.data());
            REQUIRE(c.nnz() == int64_t(a.inner.size()));
            for (int64_t i = 0; i < c.rows; ++i)
                for (int32_t p = c.rowptr[size_t(i)] + 1; p < c.rowptr[size_t(i) + 1]; ++p) REQUIRE(c.colidx[size_t(p) - 1] < c.colidx[size_t(p)]);
            const RowStats st = row_stats(c.rows, c.rowptr.data());
            const KernelChoice ch = choose_kernel(dtype, st);
            REQUIRE(ch.kernel >= K_CSR_SCALAR && ch.kernel <= K_SELL);
            for (int R = 1; R <= 2; ++R) {
                const SellHost s = build_sell(c, R);
                int64_t live = 0;
                for (int32_t col : s.colidx) live += col >= 0 ? 1 : 0;
                REQUIRE(live == c.nnz());
            }
            const std::vector<int32_t> rb = build_row_blocks(c.rows, c.rowptr.data(), 64, 16);
            REQUIRE(rb.front() == 0 && rb.back() == c.rows);
            for (int P : {1, 2, 3, 8}) {
                const std::vector<int64_t> part = partition_rows(c.rows, c.rowptr.data(), P);
                REQUIRE(part.front() == 0 && part.back() == c.rows);
                for (int r = 0; r < P; ++r) {
                    const int64_t lo = part[size_t(r)], hi = part[size_t(r) + 1];
                    REQUIRE(lo <= hi);
                    std::vector<int32_t> rp(size_t(hi - lo) + 1);
                    for (int64_t i = lo; i <= hi; ++i) rp[size_t(i - lo)] = c.rowptr[size_t(i)] - c.rowptr[size_t(lo)];
                    std::vector<int32_t> ci(c.colidx.begin() + c.rowptr[size_t(lo)], c.colidx.begin() + c.rowptr[size_t(hi)]);
                    const HaloPlan h = halo_plan(hi - lo, rp.data(), ci.data(), part, r);
                    REQUIRE(h.peer_count[size_t(r)] == 0);
                    const int64_t pad = (hi - lo + 15) / 16 * 16;
                    localize_columns(int64_t(ci.size()), ci.data(), lo, hi, pad, h.cols);
                    for (int32_t v : ci) REQUIRE(v >= 0 && v < pad + int64_t(h.cols.size()));
                }
            }
        }
    }
    // malformed inputs are rejected, not read out of bounds
    const int32_t outer_bad[3] = {0, 2, 1}, inner_ok[2] = {0, 1}, inner_bad[2] = {0, 9};
    const int32_t outer_ok[3] = {0, 1, 2};
    const double v[2] = {1.0, 2.0};
    bool threw = false;
    try {
        csc_to_csr(0, 2, 2, outer_bad, inner_ok, v);
    } catch (const PlanError&) {
        threw = true;
    }
    REQUIRE(threw);
    threw = false;
    try {
        csc_to_csr(0, 2, 2, outer_ok, inner_bad, v);
    } catch (const PlanError&) {
        threw = true;
    }
    REQUIRE(threw);
}

static std::string tmp_path(const char* name) {
    const char* d = std::getenv("SANITIZE_TMP");
    return std::string(d ? d : "/tmp") + "/" + name;
}

static void expect_io_error(const std::string& path, int dtype, const char* what) {
    bool threw = false;
    try {
        read_matrix_text(path.c_str(), dtype);
    } catch (const IoError& e) {
        threw = std::string(e.what()).find(what) != std::string::npos;
        if (!threw) std::fprintf(stderr, "unexpected message: %s (wanted %s)\n", e.what(), what);
    }
    REQUIRE(threw);
}

static void fileio_suite() {
    // sparse text file with shuffled entries, real and complex
    for (int dtype = 0; dtype <= 1; ++dtype) {
        const int64_t n = 300;
        const Csc a = random_csc(n, 4.0, 99u, dtype, false);
        std::vector<size_t> order(a.inner.size());
        std::vector<int32_t> colof(a.inner.size());
        for (int64_t j = 0; j < n; ++j)
            for (int32_t p = a.outer[size_t(j)]; p < a.outer[size_t(j) + 1]; ++p) colof[size_t(p)] = int32_t(j);
        for (size_t i = 0; i < order.size(); ++i) order[i] = i;
        std::mt19937_64 rng(5);
        std::shuffle(order.begin(), order.end(), rng);
        const std::string p = tmp_path(dtype ? "san_sparse_c.txt" : "san_sparse_r.txt");
        {
            std::ofstream f(p);
            f.precision(17);
            f << "sparse\n" << n << " " << n << " " << a.inner.size() << "\n";
            for (size_t k : order) {
                f << a.inner[k] << " " << colof[k] << " " << a.vals[k * size_t(dtype + 1)];
                if (dtype) f << " " << a.vals[k * 2 + 1];
                f << "\n";
            }
        }
        const FileMatrixHost m = read_matrix_text(p.c_str(), dtype);
        REQUIRE(!m.dense && m.rows == n && m.cols == n && m.nnz == int64_t(a.inner.size()));
        REQUIRE(m.outer == a.outer && m.inner == a.inner && m.values == a.vals);
        const std::string b = tmp_path("san.bin");
        write_matrix_bin(b.c_str(), dtype, false, n, n, m.nnz, m.outer.data(), m.inner.data(), m.values.data());
        {
            const MappedMatrix mm(b.c_str());
            REQUIRE(mm.h.nnz == m.nnz && mm.outer[n] == m.nnz);
            REQUIRE(std::memcmp(mm.values, m.values.data(), m.values.size() * sizeof(double)) == 0);
        }
        // truncate the container at a few places: always the IO error
        std::ifstream in(b, std::ios::binary);
        const std::string all((std::istreambuf_iterator<char>(in)), std::istreambuf_iterator<char>());
        for (size_t cut : {size_t(10), size_t(64), size_t(100), all.size() / 2, all.size() - 70}) {
            const std::string t = tmp_path("san_cut.bin");
            std::ofstream(t, std::ios::binary).write(all.data(), std::streamsize(cut));
            bool threw = false;
            try {
                MappedMatrix mm(t.c_str());
            } catch (const IoError&) {
                threw = true;
            }
            REQUIRE(threw);
        }
    }
    // dense text
    {
        const std::string p = tmp_path("san_dense.txt");
        std::ofstream(p) << "dense\n2 3\n1 2 3\n4 5 6\n";
        const FileMatrixHost m = read_matrix_text(p.c_str(), 0);
        REQUIRE(m.dense && m.rows == 2 && m.cols == 3 && m.values[1] == 4.0 && m.values[4] == 3.0);
    }
    // the reference's error messages (src/reader/file_matrix_reader.hpp)
    expect_io_error(tmp_path("san_missing.txt"), 0, "Impossible to open the file");
    {
        const std::string p = tmp_path("san_bad1.txt");
        std::ofstream(p) << "banana\n2 2\n";
        expect_io_error(p, 0, "Unknown storage type");
        std::ofstream(p) << "sparse\n2 2 1\n5 0 1.0\n";
        expect_io_error(p, 0, "Sparse indices out of range");
        std::ofstream(p) << "dense\n2 2\n1 2 3\n";
        expect_io_error(p, 0, "Failed to read");
        std::ofstream(p) << "";
        expect_io_error(p, 0, "");
    }
}

static void blockmath_suite() {
    std::mt19937_64 rng(11);
    std::uniform_real_distribution<double> u(-1.0, 1.0);
    for (int k : {1, 2, 4, 8}) {
        const size_t kk = size_t(k) * size_t(k);
        // interleaved complex k x k, column-major (tests/blockmath/sim.cpp conventions)
        for (int variant = 0; variant < 5; ++variant) {
            std::vector<double> h(2 * kk, 0.0), w(2 * kk, 0.0), r(2 * kk, 0.0), t(2 * kk, 0.0), theta(2 * size_t(k), 0.0), s(2 * kk, 0.0), z(2 * kk, 0.0);
            for (int a = 0; a < k; ++a)
                for (int b = 0; b < k; ++b) {
                    double re = 0.0, im = 0.0;
                    if (variant == 0) re = a == b ? 1.0 : 0.0;                       // identity
                    else if (variant == 2) re = (b == a + 1) ? 1.0 : (a == b ? 2.0 : 0.0);  // Jordan block
                    else if (variant == 3) re = a == b ? 1.0 + 1e-13 * a : 0.0;      // clustered
                    else if (variant == 4) { re = u(rng); im = u(rng); }              // random complex
                    h[2 * (size_t(b) * k + a)] = re;                                  // variant 1: zero
                    h[2 * (size_t(b) * k + a) + 1] = im;
                }
            // W = H^H H + I is Hermitian positive definite
            for (int a = 0; a < k; ++a)
                for (int b = 0; b < k; ++b) {
                    std::complex<double> acc(a == b ? 1.0 : 0.0, 0.0);
                    for (int c = 0; c < k; ++c) {
                        const std::complex<double> x(h[2 * (size_t(a) * k + c)], h[2 * (size_t(a) * k + c) + 1]);
                        const std::complex<double> y(h[2 * (size_t(b) * k + c)], h[2 * (size_t(b) * k + c) + 1]);
                        acc += std::conj(x) * y;
                    }
                    w[2 * (size_t(b) * k + a)] = acc.real();
                    w[2 * (size_t(b) * k + a) + 1] = acc.imag();
                }
            REQUIRE(bm_chol(w.data(), k, r.data()) == 1);
            bm_triinv(r.data(), k, t.data());
            for (double v : t) REQUIRE(std::isfinite(v));
            bm_eig(h.data(), k, theta.data(), s.data());
            bm_schur(h.data(), k, t.data(), z.data());
            for (double v : theta) REQUIRE(std::isfinite(v));
        }
    }
}

int main() {
    plan_suite();
    fileio_suite();
    blockmath_suite();
    std::printf("host_sanitize: %d checks, no sanitizer report\n", g_checks);
    return 0;
}
