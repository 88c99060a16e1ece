// Test helper: reads a matrix file through the drop-in EigSol::readMatrixFromFile<Scalar> header and dumps
// what the resulting EigSol::Matrix holds, so tests/test_file_reader.py can compare it with the reference's
// reader (oracle/_ref) and with the Python wrapper.  usage: reader_dump <d|z|f> <file> <out.bin>
// Output: int64 {is_dense, rows, cols, nnz}; sparse: outer[cols+1], inner[nnz] (int32); values as doubles
// (re, im interleaved for z).  Exit code 3 + message on stderr when the reader throws.
#include <cstdint>
#include <cstdio>
#include <iostream>

#include "src/reader/file_matrix_reader.hpp"

template <typename S>
static void put_values(std::FILE* f, const S* v, int64_t n) {
    for (int64_t k = 0; k < n; ++k) {
        if constexpr (std::is_same_v<S, std::complex<double>>) {
            const double d[2] = {v[k].real(), v[k].imag()};
            std::fwrite(d, sizeof(double), 2, f);
        } else {
            const double d = static_cast<double>(v[k]);
            std::fwrite(&d, sizeof(double), 1, f);
        }
    }
}

template <typename S>
static int dump(const char* path, const char* out) {
    EigSol::Matrix m = EigSol::readMatrixFromFile<S>(path);
    std::FILE* f = std::fopen(out, "wb");
    if (!f) return 2;
    if (m.isDense()) {
        const auto& a = m.cast<EigSol::Matrix::Dense<S>>();
        const int64_t h[4] = {1, a.rows(), a.cols(), a.rows() * a.cols()};
        std::fwrite(h, sizeof(int64_t), 4, f);
        put_values(f, a.data(), h[3]);
    } else {
        const auto& a = m.cast<EigSol::Matrix::Sparse<S>>();
        if (!a.isCompressed()) return 4;
        const int64_t h[4] = {0, a.rows(), a.cols(), a.nonZeros()};
        std::fwrite(h, sizeof(int64_t), 4, f);
        std::fwrite(a.outerIndexPtr(), sizeof(int32_t), size_t(a.cols()) + 1, f);
        std::fwrite(a.innerIndexPtr(), sizeof(int32_t), size_t(a.nonZeros()), f);
        put_values(f, a.valuePtr(), h[3]);
    }
    std::fclose(f);
    return 0;
}

int main(int argc, char** argv) {
    if (argc != 4) return 1;
    try {
        switch (argv[1][0]) {
            case 'd': return dump<double>(argv[2], argv[3]);
            case 'z': return dump<std::complex<double>>(argv[2], argv[3]);
            case 'f': return dump<float>(argv[2], argv[3]);
        }
    } catch (const std::exception& e) {
        std::cerr << e.what() << "\n";
        return 3;
    }
    return 1;
}
