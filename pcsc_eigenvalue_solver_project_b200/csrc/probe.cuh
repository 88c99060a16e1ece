// probe.cuh -- HBM ceiling for a READ-DOMINATED stream with the matvec's read/write mix.
//
// MEASURED_PEAKS.json's hbm_gbs is a copy (50 % reads, 50 % writes).  One fp64 CSR matvec of the 256^3 Laplacian
// reads 1.61 GB and writes 0.13 GB (92 / 8).  This probe streams `planes` coalesced read planes into one written
// plane -- out[i] = sum_j in[j][i], 256-bit non-coherent loads, nothing else -- so that bench.py can print, beside
// roofline.peak, what a kernel with the matvec's mix can reach on the same box in the same run (peak_read_mix).
// Measurement infrastructure only: nothing on the power-iteration path calls it.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace eigsol_b200 {

// 32 bytes per thread and plane (one 256-bit load: SASS LDG.E.ENL2.256), PLANES loads in flight before the first add
template <int PLANES>
__global__ void __launch_bounds__(256) read_mix_kernel(const double* __restrict__ in, int64_t n4, double* __restrict__ out) {
    const int64_t stride = int64_t(gridDim.x) * blockDim.x;
    for (int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n4; i += stride) {
        double v[PLANES][4];
#pragma unroll
        for (int j = 0; j < PLANES; ++j) {
            const double* p = in + (int64_t(j) * n4 + i) * 4;
            asm volatile("ld.global.nc.L1::no_allocate.v4.f64 {%0, %1, %2, %3}, [%4];"
                         : "=d"(v[j][0]), "=d"(v[j][1]), "=d"(v[j][2]), "=d"(v[j][3])
                         : "l"(p));
        }
        double a[4] = {v[0][0], v[0][1], v[0][2], v[0][3]};
#pragma unroll
        for (int j = 1; j < PLANES; ++j)
#pragma unroll
            for (int k = 0; k < 4; ++k) a[k] += v[j][k];
        asm volatile("st.global.v4.f64 [%0], {%1, %2, %3, %4};" ::"l"(out + i * 4), "d"(a[0]), "d"(a[1]), "d"(a[2]), "d"(a[3]) : "memory");
    }
}

}  // namespace eigsol_b200
