// probe.cuh -- HBM ceiling for a READ-DOMINATED stream with the matvec's read/write mix.
//
// MEASURED_PEAKS.json's hbm_gbs is a copy (50 % reads, 50 % writes).  One fp64 CSR matvec of the 256^3 Laplacian
// reads 1.61 GB and writes 0.13 GB (92 / 8).  This probe streams `planes` coalesced read planes into one written
// plane -- out[i] = sum_j in[j][i], 128-bit non-coherent loads, nothing else -- so that bench.py can print, beside
// roofline.peak, what a kernel with the matvec's mix can reach on the same box in the same run (peak_read_mix).
// Measurement infrastructure only: nothing on the power-iteration path calls it.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace eigsol_b200 {

template <int PLANES>
__global__ void __launch_bounds__(256) read_mix_kernel(const double2* __restrict__ in, int64_t n2, double2* __restrict__ out) {
    const int64_t stride = int64_t(gridDim.x) * blockDim.x;
    for (int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n2; i += stride) {
        double2 v[PLANES];
#pragma unroll
        for (int j = 0; j < PLANES; ++j) {
            const double* p = reinterpret_cast<const double*>(in + int64_t(j) * n2 + i);
            asm volatile("ld.global.nc.L1::no_allocate.v2.f64 {%0, %1}, [%2];" : "=d"(v[j].x), "=d"(v[j].y) : "l"(p));
        }
        double2 acc = v[0];
#pragma unroll
        for (int j = 1; j < PLANES; ++j) {
            acc.x += v[j].x;
            acc.y += v[j].y;
        }
        out[i] = acc;
    }
}

}  // namespace eigsol_b200
