// TEST INFRASTRUCTURE: drives tests/statemachine/sim.cu (the device's control state machine of kernels.cuh compiled
// for the host) under AddressSanitizer + UndefinedBehaviorSanitizer: converging, fixed-trip, zero-iteration and
// zero-norm runs at several poll cadences.  Built and run by tests/test_sanitizers.py.
#include <cmath>
#include <complex>
#include <cstdio>
#include <vector>

extern "C" int sim_power_dense(int n, const double* a, const double* x0, int max_it, double tol, int poll, double* x_out,
                               double* lambda_out, int* iters, int* conv, int* launches_out);

int main() {
    typedef std::complex<double> cd;
    int runs = 0;
    for (int n : {1, 2, 3, 17}) {
        for (int kind = 0; kind < 3; ++kind) {
            std::vector<cd> a(size_t(n) * size_t(n), cd(0.0, 0.0));
            for (int i = 0; i < n; ++i)
                for (int j = 0; j < n; ++j) {
                    if (kind == 0) a[size_t(j) * n + i] = i == j ? cd(1.0 + i, 0.5 * i) : cd(0.01 * (i + j), 0.0);
                    else if (kind == 1) a[size_t(j) * n + i] = j == i + 1 ? cd(1.0, 0.0) : cd(0.0, 0.0);  // nilpotent: ||y|| == 0 exit
                    else a[size_t(j) * n + i] = cd(1e150 * (i == j ? 2.0 + i : 0.1), 0.0);                // extreme scale
                }
            std::vector<cd> x0(n, cd(1.0 / std::sqrt(double(n)), 0.0));
            std::vector<cd> x(n, cd(0.0, 0.0));
            for (int max_it : {0, 1, 2, 7, 300})
                for (double tol : {-1.0, 1e-10})
                    for (int poll : {1, 3, 64}) {
                        double lam[2] = {0, 0};
                        int iters = -1, conv = -1, launches = -1;
                        const int rc = sim_power_dense(n, reinterpret_cast<const double*>(a.data()), reinterpret_cast<const double*>(x0.data()),
                                                       max_it, tol, poll, reinterpret_cast<double*>(x.data()), lam, &iters, &conv, &launches);
                        if (rc != 0 || iters < 0 || iters > std::max(max_it, 0) || launches < 0) {
                            std::fprintf(stderr, "bad result rc=%d iters=%d launches=%d\n", rc, iters, launches);
                            return 2;
                        }
                        ++runs;
                    }
        }
    }
    std::printf("statemachine_sanitize: %d runs, no sanitizer report\n", runs);
    return 0;
}
