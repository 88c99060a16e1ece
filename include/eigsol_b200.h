/*
 * eigsol_b200.h -- C ABI of libeigsol_b200.so: the B200 (sm_100a) power-iteration path.
 *
 * This is the drop-in boundary for ONE path of the reference library
 * hugoheziyang/PCSC_Eigenvalue_Solver_Project: the loop inside
 *     EigSol::powerMethodImpl<Mat>(const Mat& A, const SolverOptions& opts)
 *         src/power_method/power_method.hpp:47-99
 * reached through
 *     EigSol::powerMethod<Scalar>(const Matrix& M, const SolverOptions& opts)
 *         src/power_method/power_method.hpp:135-148.
 * (All file:line citations in this header are relative to the reference checkout.)
 *
 * The reference has no FFI of its own (it is header-only C++ over Eigen), so the entry points below
 * are what a binding for this path has to carry: the raw arrays Eigen exposes for the two canonical
 * matrix types of src/matrix/matrix.hpp:39-44, the two fields of SolverOptions
 * (src/option/solver_option.hpp:14-20) and the four fields of EigenResult
 * (src/result/eigen_result.hpp:23-52).  INTEGRATION.md shows the C++ header shim and the ctypes stub.
 *
 * Conventions
 *  - plain pointers and sizes only; every pointer is HOST memory, borrowed for the duration of the
 *    call (the library copies to the device); no device pointer crosses this boundary.
 *  - dtype: EIGSOL_B200_F64 = double, EIGSOL_B200_C128 = std::complex<double> passed as interleaved
 *    (re, im) doubles -- the memory layout of std::complex<double>.
 *  - sparse input is what Eigen::SparseMatrix<Scalar> (ColMajor, int) holds in compressed mode:
 *    outerIndexPtr()/innerIndexPtr()/valuePtr() = CSC with int32 indices
 *    (eigen/Eigen/src/SparseCore/SparseMatrix.h:171-193).  The library transposes to CSR once at upload.
 *  - dense input is column-major with leading dimension lda (Eigen::Matrix<Scalar,Dynamic,Dynamic>).
 *  - x0 is the unit-norm start vector; the reference draws it inside the call
 *    (power_method.hpp:62, Vector<Scalar>::Random(n) then normalize()) -- the host shim does the same
 *    with the same Eigen expression and hands it over, so both sides start from the same vector.
 *  - every function returns an eigsol_b200_status; on failure eigsol_b200_last_error() holds a
 *    thread-local message.  There is NO CPU fallback: without a usable CUDA device every compute
 *    entry point fails with EIGSOL_B200_ERR_DEVICE.
 */
#ifndef EIGSOL_B200_H
#define EIGSOL_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define EIGSOL_B200_API __attribute__((visibility("default")))
#else
#define EIGSOL_B200_API
#endif

typedef enum eigsol_b200_status {
    EIGSOL_B200_OK = 0,
    EIGSOL_B200_ERR_INVALID = 1,    /* bad argument (null pointer, negative size, malformed CSC, ...) */
    EIGSOL_B200_ERR_NOT_SQUARE = 2, /* power_method.hpp:52-55  "powerMethod: matrix must be square"  */
    EIGSOL_B200_ERR_ZERO_SIZE = 3,  /* power_method.hpp:56-58  "powerMethod: matrix has zero size"   */
    EIGSOL_B200_ERR_DEVICE = 4,     /* CUDA / NCCL failure, or no device: "powerMethod: device error: ..." */
    EIGSOL_B200_ERR_STATE = 5,      /* call sequence error on a handle (step before begin, ...)          */
    EIGSOL_B200_ERR_IO = 6          /* matrix file cannot be opened / is malformed: the reader's messages  */
} eigsol_b200_status;

enum { EIGSOL_B200_F64 = 0, EIGSOL_B200_C128 = 1 };

/* SpMV / GEMV kernel variants (eigsol_b200_matrix_info.kernel, eigsol_b200_matrix_set_kernel). */
enum {
    EIGSOL_B200_KERNEL_AUTO = 0,       /* choose from row-length statistics (sparse) / GEMV (dense)    */
    EIGSOL_B200_KERNEL_CSR_SCALAR = 1, /* CSR, one thread per row                                      */
    EIGSOL_B200_KERNEL_CSR_VECTOR = 2, /* CSR, 2..32 lanes per row (32 = warp per row)                  */
    EIGSOL_B200_KERNEL_CSR_STREAM = 3, /* CSR, nnz streamed through shared memory, rows reduced there   */
    EIGSOL_B200_KERNEL_SELL = 4,       /* sliced layout built from the CSR (slice = 32 lanes x R rows)  */
    EIGSOL_B200_KERNEL_GEMV = 5,       /* dense column-major GEMV                                       */
    EIGSOL_B200_KERNEL_CSR_WINDOW = 6  /* CSR, thread per row, x window of each row block staged in shared
                                          memory by one TMA bulk copy (banded / stencil matrices)       */
};

typedef struct eigsol_b200_matrix eigsol_b200_matrix; /* device-resident matrix + iteration state */
typedef struct eigsol_b200_comm eigsol_b200_comm;     /* one rank of a multi-GPU group            */

typedef struct eigsol_b200_matrix_info {
    int32_t dtype;            /* EIGSOL_B200_F64 / _C128                                              */
    int32_t is_dense;
    int64_t n;                /* global order of the matrix                                           */
    int64_t nnz;              /* global stored entries (dense: n*n)                                   */
    int64_t local_rows;       /* rows owned by this rank (== n on one GPU)                            */
    int64_t local_nnz;
    int64_t halo;             /* remote x entries this rank receives per iteration                    */
    int32_t kernel;           /* EIGSOL_B200_KERNEL_* actually used                                   */
    int32_t lanes_per_row;    /* CSR_VECTOR: lanes per row; SELL: rows per lane; else 0               */
    int32_t grid;             /* CTAs per launch of the dominant kernel                               */
    int32_t block;            /* threads per CTA                                                      */
    double row_len_mean;      /* row-length statistics the choice was made from                       */
    double row_len_max;
    double row_len_cv;        /* coefficient of variation (stddev / mean)                             */
    int64_t alg_bytes;        /* algorithmic bytes of ONE matvec on this rank (SURVEY.md 8d formula)  */
    int64_t stored_bytes;     /* bytes the device layout actually streams per matvec on this rank     */
    int32_t nranks;
    int32_t rank;
    int32_t kernels_per_trip; /* kernel launches one power iteration issues on this rank              */
    int32_t exchange;         /* 0 = one GPU (fused finish), 1 = NCCL send/recv + all-reduce,
                                 2 = peer memory: one kernel does matvec + x exchange + scalar all-reduce */
} eigsol_b200_matrix_info;

/* ---- diagnostics ------------------------------------------------------------------------------ */

/* Message of the last failure on the calling thread ("" if none). */
EIGSOL_B200_API const char* eigsol_b200_last_error(void);

/* Number of CUDA devices visible (0 when there is no driver / no GPU; never fails). */
EIGSOL_B200_API int eigsol_b200_device_count(void);

/* Library version string. */
EIGSOL_B200_API const char* eigsol_b200_version(void);

/* ---- one-shot entry points: what EigSol::powerMethodImpl's loop body is replaced by ---------- */

/*
 * Sparse: replaces power_method.hpp:61-98 for Mat = Matrix::Sparse<Scalar> (matrix.hpp:43-44).
 *   rows, cols        A.rows(), A.cols()                     (checked as :52-58, in that order)
 *   outer[cols+1]     A.outerIndexPtr()   inner[nnz]  A.innerIndexPtr()   values[nnz]  A.valuePtr()
 *   x0[rows]          normalised start vector (:62)
 *   max_iterations    opts.maxIterations    tolerance   opts.tolerance   (solver_option.hpp:14-20)
 *   x_out[rows]       EigenResult::eigenvector    lambda_out = {re, im} of EigenResult::eigenvalue
 *   iterations_out    EigenResult::iterations     converged_out  EigenResult::converged (0/1)
 */
EIGSOL_B200_API int eigsol_b200_power_csc(int dtype, int64_t rows, int64_t cols, const int32_t* outer,
                                          const int32_t* inner, const void* values, const void* x0,
                                          int max_iterations, double tolerance, void* x_out,
                                          double lambda_out[2], int* iterations_out,
                                          int* converged_out);

/* Dense: replaces power_method.hpp:61-98 for Mat = Matrix::Dense<Scalar> (matrix.hpp:39-40);
 * a = A.data(), lda = A.outerStride(). */
EIGSOL_B200_API int eigsol_b200_power_dense(int dtype, int64_t rows, int64_t cols, const void* a,
                                            int64_t lda, const void* x0, int max_iterations,
                                            double tolerance, void* x_out, double lambda_out[2],
                                            int* iterations_out, int* converged_out);

/* The same two calls, but on success the device-resident matrix is handed to the caller (*keep_out; destroy it with
 * eigsol_b200_matrix_destroy) instead of being freed: the first solve on a Matrix and its upload are ONE call (the start
 * vector rides on the matrix upload), later solves call eigsol_b200_power_run on the kept handle.  This is what the
 * drop-in Matrix cache uses.  keep_out == NULL is the plain one-shot call. */
EIGSOL_B200_API int eigsol_b200_power_csc_keep(int dtype, int64_t rows, int64_t cols, const int32_t* outer,
                                               const int32_t* inner, const void* values, const void* x0,
                                               int max_iterations, double tolerance, void* x_out,
                                               double lambda_out[2], int* iterations_out, int* converged_out,
                                               eigsol_b200_matrix** keep_out);
EIGSOL_B200_API int eigsol_b200_power_dense_keep(int dtype, int64_t rows, int64_t cols, const void* a,
                                                 int64_t lda, const void* x0, int max_iterations,
                                                 double tolerance, void* x_out, double lambda_out[2],
                                                 int* iterations_out, int* converged_out,
                                                 eigsol_b200_matrix** keep_out);

/* ---- handle API: upload once, iterate many (device-resident Matrix; bench + repeated callers) -- */

/* comm == NULL: the calling process owns the whole solve.  By default that is ONE GPU (the calling thread's
 * current device).  With the environment variable EIGSOL_B200_NGPU=P (P <= 8, <= visible devices) the handle
 * becomes a SINGLE-PROCESS GROUP: the library row-shards the matrix over P devices (starting at the current one),
 * drives each from its own host thread, and runs the same fused matvec + halo-exchange + all-reduce kernel over
 * peer memory as a process-per-GPU group does -- the caller, e.g. the unmodified EigSol::powerMethod(M, opts), sees
 * one handle and global vectors.  Matrices with fewer than EIGSOL_B200_MIN_ROWS_PER_GPU (1024) rows per device use
 * fewer devices.  One-shot calls (eigsol_b200_power_csc / _dense) follow the same rule.
 * With a comm (one process per GPU), EVERY rank passes the same global matrix; the library keeps this rank's
 * nnz-balanced row block. */
EIGSOL_B200_API int eigsol_b200_matrix_create_csc(int dtype, int64_t rows, int64_t cols,
                                                  const int32_t* outer, const int32_t* inner,
                                                  const void* values, eigsol_b200_comm* comm,
                                                  eigsol_b200_matrix** out);

/* Row block [row_begin, row_begin+row_count) of a global n x n matrix, already in CSR with GLOBAL
 * column indices ascending within each row; rowptr[row_count+1] starts at 0.  Lets each rank of a
 * large run build only its own rows.  comm == NULL requires row_begin == 0 and row_count == n. */
EIGSOL_B200_API int eigsol_b200_matrix_create_csr_rows(int dtype, int64_t n, int64_t row_begin,
                                                       int64_t row_count, const int32_t* rowptr,
                                                       const int32_t* colidx, const void* values,
                                                       eigsol_b200_comm* comm,
                                                       eigsol_b200_matrix** out);

EIGSOL_B200_API int eigsol_b200_matrix_create_dense(int dtype, int64_t rows, int64_t cols,
                                                    const void* a, int64_t lda,
                                                    eigsol_b200_comm* comm, eigsol_b200_matrix** out);

EIGSOL_B200_API int eigsol_b200_matrix_destroy(eigsol_b200_matrix* m);

EIGSOL_B200_API int eigsol_b200_matrix_get_info(const eigsol_b200_matrix* m,
                                                eigsol_b200_matrix_info* info);

/* Force a kernel variant (benchmarks / parity tests of every variant).  lanes: CSR_VECTOR lanes per
 * row (2,4,8,16,32) or SELL rows per lane (1,2); 0 = default.  Rebuilds derived layouts if needed. */
EIGSOL_B200_API int eigsol_b200_matrix_set_kernel(eigsol_b200_matrix* m, int kernel, int lanes);

/* Full solve on a resident matrix: same semantics and outputs as the one-shot calls.  x0 and x_out
 * are GLOBAL length-n host vectors on every rank. */
EIGSOL_B200_API int eigsol_b200_power_run(eigsol_b200_matrix* m, const void* x0, int max_iterations,
                                          double tolerance, void* x_out, double lambda_out[2],
                                          int* iterations_out, int* converged_out);

/* Stepwise form of the same loop (device-resident between calls):
 *   begin : upload x0, reset the iteration state (max_iterations / tolerance as in power_run)
 *   step  : issue up to `trips` further power iterations; *ms_out (may be NULL) = device time of
 *           those launches measured with CUDA events on the launching stream (a multi-GPU rank first passes
 *           a device-side barrier of the ranks, so the timers of all ranks start together; a single-process
 *           group returns the slowest rank's time)
 *   end   : normalise and download the current eigenvector and the state                         */
EIGSOL_B200_API int eigsol_b200_power_begin(eigsol_b200_matrix* m, const void* x0, int max_iterations,
                                            double tolerance);
EIGSOL_B200_API int eigsol_b200_power_step(eigsol_b200_matrix* m, int trips, float* ms_out);
EIGSOL_B200_API int eigsol_b200_power_end(eigsol_b200_matrix* m, void* x_out, double lambda_out[2],
                                          int* iterations_out, int* converged_out);

/* One product y = A*x on the resident matrix (power_method.hpp:69 in isolation); x, y global host
 * vectors.  Used by the parity tests to check every kernel variant element-wise. */
EIGSOL_B200_API int eigsol_b200_matvec(eigsol_b200_matrix* m, const void* x, void* y);

/* dot_out = x^H (A x) (conjugated, eigen/Eigen/src/Core/InnerProduct.h:255), *norm2_out = ||A x||^2, from ONE
 * fused launch on the resident matrix: the Rayleigh-quotient step `x.dot(A * x)` of the reference's shifted
 * inverse method (src/power_method/shifted_inverse_power_solver.hpp:62), whose LU solve stays host code.
 * x is a GLOBAL length-n host vector on every rank; either output may be NULL. */
EIGSOL_B200_API int eigsol_b200_rayleigh(eigsol_b200_matrix* m, const void* x, double dot_out[2],
                                         double* norm2_out);

/* Block (simultaneous) power iteration: `block` vectors advance per pass over A (no
 * counterpart in the reference, whose powerMethod keeps one vector, src/power_method/power_method.hpp:68-96).
 *     Q = orth(U),  U <- A Q,  H = Q^H A Q (block x block),  Ritz values = eig(H), ordered by modulus
 * With block = 1 this is the reference's loop trip for trip (same usedIters, flag, eigenvalue, vector).  A trip
 * is one SpMM launch + one single-CTA control launch; convergence = the first `nev` Ritz values each pass
 * is_close_relative (src/core/tolerance.hpp:28-33) against the previous trip.
 *   block        1, 2, 4 or 8;  1 <= nev <= block
 *   x0           n x block start block, column-major (Eigen's default), full column rank
 *   x_out        n x block unit-norm Ritz vectors, column-major, or NULL.  dtype F64 with a complex-conjugate
 *                pair of Ritz values in columns j, j+1 returns Re and Im of the pair's vector there (dgeev style)
 *   lambda_out   2 * block doubles: (re, im) of each Ritz value
 *   breakdown_out  1 if the block lost rank (A Q or x0 rank deficient -- the block counterpart of ||y|| == 0,
 *                power_method.hpp:73-76); results are those of the last good trip
 *   ms_out       device time of the iteration loop
 * One GPU (comm == NULL matrices); any of the out pointers except lambda_out may be NULL. */
EIGSOL_B200_API int eigsol_b200_block_power_run(eigsol_b200_matrix* m, int block, int nev, const void* x0,
                                                int max_iterations, double tolerance, void* x_out,
                                                double* lambda_out, int* iterations_out, int* converged_out,
                                                int* breakdown_out, float* ms_out);

/* Total kernel launches this handle has issued (bench.py's gpu_launches claim). */
EIGSOL_B200_API int64_t eigsol_b200_matrix_launch_count(const eigsol_b200_matrix* m);

/* Matrix handles created by this process so far (every one is an upload of a whole matrix): lets a caller -- and the
 * tests of the drop-in Matrix cache -- see that repeated solves on one EigSol::Matrix upload once. */
EIGSOL_B200_API int64_t eigsol_b200_matrices_created(void);

/* Multi-GPU diagnostics (library built-in, enabled by EIGSOL_B200_PEER_STATS=1 at matrix creation; zeros otherwise):
 * where this rank's trips waited since the previous call, in nanoseconds of the GPU's global timer.
 *   out[0] control CTA: time between publishing its flag and holding every peer's flag, summed over launches
 *   out[1] launches counted in out[0]
 *   out[2] matvec warps that reached a halo unit before the owner's flag: time spun, summed     out[3] such warps
 *   out[4] longest single halo wait            out[5] matvec warps that reached a halo unit (waited or not) */
EIGSOL_B200_API int eigsol_b200_matrix_peer_stats(eigsol_b200_matrix* m, uint64_t out[8]);

/* Measurement only (nothing on the path calls it): HBM throughput of a read-dominated stream -- `planes`
 * (1, 4, 8 or 12) coalesced read planes of plane_bytes summed into one written plane, 128-bit loads -- in GB/s
 * counted over read + written bytes; best and mean of `reps` launches timed with CUDA events.  12 planes is the
 * 92 / 8 read/write mix of an fp64 CSR matvec; bench.py prints it as roofline.peak_read_mix. */
EIGSOL_B200_API int eigsol_b200_hbm_probe(int64_t plane_bytes, int planes, int reps, double* best_gbps,
                                          double* mean_gbps);

/* ---- multi-GPU, one PROCESS per GPU (torchrun / MPI): every process creates a comm and passes it to the
 * ---- matrix constructors.  (One process driving several GPUs needs no comm: see EIGSOL_B200_NGPU above.) ---- */

/* 128-byte NCCL unique id; rank 0 creates it, the host program broadcasts it. */
EIGSOL_B200_API int eigsol_b200_comm_unique_id(void* id128);
EIGSOL_B200_API int eigsol_b200_comm_create(int rank, int nranks, const void* id128, int device,
                                            eigsol_b200_comm** out);
EIGSOL_B200_API int eigsol_b200_comm_destroy(eigsol_b200_comm* comm);
/* Device-side barrier + max over ranks of a float (bench timing: max over ranks of event times). */
EIGSOL_B200_API int eigsol_b200_comm_max_float(eigsol_b200_comm* comm, float* inout);

/* ---- matrix files: the reference's text format read fast, and a mappable binary container (no GPU needed) ----
 *
 * eigsol_b200_read_matrix_text replaces EigSol::readMatrixFromFile<Scalar>(filename)
 * (src/reader/file_matrix_reader.hpp:170-200; dense branch :33-66, sparse branch :84-132; format README.md:87-171):
 * the file is mapped and parsed by all host threads; `dtype` is the Scalar it is read as (two reals per value
 * for EIGSOL_B200_C128).  On return the struct holds exactly what the reference's Matrix would own: dense ->
 * values[rows*cols] column-major (Matrix::Dense<Scalar>); sparse -> outer[cols+1], inner[nnz], values[nnz], rows
 * ascending inside a column (a compressed Matrix::Sparse<Scalar>).  Failures return EIGSOL_B200_ERR_IO with the
 * reference's message ("Impossible to open the file: ...", "Unknown storage type: ...", "Sparse indices out of
 * range", ...).  The arrays belong to the struct until eigsol_b200_file_matrix_release.
 * The binary container (EIGSOLB2: 64-byte header + 64-byte aligned arrays) is mapped, not parsed: its pointers
 * can go straight into eigsol_b200_matrix_create_csc / _dense. */
typedef struct eigsol_b200_file_matrix {
    int32_t dtype;
    int32_t is_dense;
    int64_t rows, cols, nnz;   /* nnz: stored entries (dense: rows*cols) */
    const int32_t* outer;      /* sparse only */
    const int32_t* inner;      /* sparse only */
    const void* values;
    void* owner;               /* internal */
} eigsol_b200_file_matrix;

EIGSOL_B200_API int eigsol_b200_read_matrix_text(const char* path, int dtype, eigsol_b200_file_matrix* out);
EIGSOL_B200_API int eigsol_b200_read_matrix_bin(const char* path, eigsol_b200_file_matrix* out);
EIGSOL_B200_API int eigsol_b200_write_matrix_bin(const char* path, int dtype, int is_dense, int64_t rows,
                                                 int64_t cols, int64_t nnz, const int32_t* outer,
                                                 const int32_t* inner, const void* values);
EIGSOL_B200_API int eigsol_b200_file_matrix_release(eigsol_b200_file_matrix* m);

/* ---- host-side planning, no GPU needed (unit-testable on CPU; used internally at upload) ------- */

/* CSC -> CSR with ascending columns inside each row (the reference's accumulation order,
 * eigen/Eigen/src/SparseCore/SparseDenseProduct.h:101-110).  rowptr[rows+1], colidx[nnz],
 * values_out[nnz] are caller-allocated. */
EIGSOL_B200_API int eigsol_b200_csc_to_csr(int dtype, int64_t rows, int64_t cols, const int32_t* outer,
                                           const int32_t* inner, const void* values, int32_t* rowptr,
                                           int32_t* colidx, void* values_out);

/* nnz-balanced contiguous row partition of a CSR matrix: row_begin[nranks+1]. */
EIGSOL_B200_API int eigsol_b200_partition_rows(int64_t rows, const int32_t* rowptr, int nranks,
                                               int64_t* row_begin);

/* Halo plan of one rank: the sorted distinct GLOBAL columns outside [row_begin[rank], row_begin[rank+1])
 * that the rank's rows reference.  Pass halo_cols = NULL to query the count; returns the count in
 * *halo_count.  peer_count[nranks] (may be NULL) = how many of them each peer owns. */
EIGSOL_B200_API int eigsol_b200_halo_plan(int64_t row_count, const int32_t* rowptr,
                                          const int32_t* colidx, int nranks, const int64_t* row_begin,
                                          int rank, int64_t* halo_count, int32_t* halo_cols,
                                          int64_t* peer_count);

/* Kernel choice (what KERNEL_AUTO does) from the statistics of a whole CSR matrix with global column indices: row
 * lengths (mean, max, padding of a sliced layout) and the share of entries on the tridiagonal, which tells PDE stencils
 * from random-offset bands.  colidx == NULL: row lengths only (the matrix is taken to be stencil-like). */
EIGSOL_B200_API int eigsol_b200_choose_kernel_csr(int dtype, int64_t rows, const int32_t* rowptr,
                                                  const int32_t* colidx, int* kernel_out, int* lanes_out);
/* The same from row lengths alone (= colidx NULL). */
EIGSOL_B200_API int eigsol_b200_choose_kernel(int dtype, int64_t rows, const int32_t* rowptr,
                                              int* kernel_out, int* lanes_out);

/* The start vector of power_method.hpp:62 -- Vector<Scalar>::Random(n) followed by normalize() --
 * restated on the host for callers that have no Eigen (the Python mirror): draws from libc rand()
 * exactly as eigen/Eigen/src/Core/RandomImpl.h:46-51,66-69,89-99 does (two 31-bit rand() words per
 * double, value in [-1,1); complex = imaginary part drawn first, as g++ evaluates the constructor
 * arguments of RandomImpl.h:256), then divides by the 2-norm (eigen/Eigen/src/Core/Dot.h:110-114).
 * seed >= 0: srand(seed) first; seed < 0: continue the caller's rand() stream. */
EIGSOL_B200_API int eigsol_b200_random_unit(int dtype, int64_t n, int seed, void* x_out);

#ifdef __cplusplus
}
#endif
#endif /* EIGSOL_B200_H */
