// sim.cu -- TEST INFRASTRUCTURE: drives the device's control state machine (kernels.cuh: control_wants_fold,
// control_apply, fold_launch, next_sigma -- compiled here for the HOST) through the same launch pipeline the
// engine issues, with the matvec and its three sums done in plain C++.  What it checks is the logic that never
// shows in a bandwidth number: the one-launch lag between a trip and its convergence test, the three-buffer
// ring that keeps the result vector of the stopping trip intact, power-of-two rescaling, maxIterations,
// the ||y|| == 0 exit (reference: src/power_method/power_method.hpp:68-98) and the host's poll cadence.
// Built by tests/test_state_machine.py with nvcc (no kernel is launched; runs on a CPU-only box).
#include <cmath>
#include <complex>
#include <cstring>
#include <vector>

#include "../../pcsc_eigenvalue_solver_project_b200/csrc/kernels.cuh"

using namespace eigsol_b200;
typedef std::complex<double> cd;

namespace {
struct Sim {
    int n;
    const cd* a;  // column-major n x n
    std::vector<cd> buf[kRing];
    double part[kRing][3];
    PowerState st;
    int issued = 0;

    void begin(const cd* x0, int max_it, double tol) {  // mirrors MatrixImpl::begin
        std::memset(&st, 0, sizeof st);
        double amax = 0.0;  // MatrixImpl::compute_prescale
        for (size_t k = 0; k < size_t(n) * size_t(n); ++k) {
            amax = std::fmax(amax, std::fabs(a[k].real()));
            amax = std::fmax(amax, std::fabs(a[k].imag()));
        }
        st.q[0] = 1.0;
        st.sigma[0] = prescale_from_amax(amax);
        st.tol = tol;
        st.final_q = 1.0;
        st.max_trips = max_it;
        st.stop = max_it <= 0 ? STOP_MAX_TRIPS : STOP_NONE;
        st.ctrl[0].sigma = st.sigma[0];
        st.ctrl[0].launch_no = 0;
        st.ctrl[0].active = max_it > 0 ? 1 : 0;
        for (auto& b : buf) b.assign(size_t(n), cd(0.0, 0.0));
        for (int i = 0; i < n; ++i) buf[0][size_t(i)] = x0[i];
        issued = 0;
    }
    void control(bool advance) {  // the control CTA, MODE_FUSED
        const ControlPlan p = control_wants_fold(&st);
        double s0 = 0, s1 = 0, s2 = 0;
        if (p.fold) {
            s0 = part[p.j % kRing][0];
            s1 = part[p.j % kRing][1];
            s2 = part[p.j % kRing][2];
        }
        control_apply(&st, p, true, s0, s1, s2, advance);
    }
    void launch() {  // one matvec launch: the matvec CTAs read ctrl[phase] as the PREVIOUS control CTA left it
        const int phase = issued % kRing;
        const LaunchCtrl ct = st.ctrl[phase];
        control(true);
        if (ct.active) {
            const std::vector<cd>& v = buf[phase];
            std::vector<cd>& out = buf[(phase + 1) % kRing];
            cd dot(0.0, 0.0);
            double q = 0.0;
            for (int i = 0; i < n; ++i) {
                cd acc(0.0, 0.0);
                for (int j = 0; j < n; ++j) acc += a[size_t(j) * size_t(n) + size_t(i)] * v[size_t(j)];
                const cd z = acc * ct.sigma;
                out[size_t(i)] = z;
                dot += std::conj(v[size_t(i)]) * z;
                q += std::norm(z);
            }
            part[phase][0] = dot.real();
            part[phase][1] = dot.imag();
            part[phase][2] = q;
        }
        ++issued;
    }
};
}  // namespace

// Same outputs as eigsol_b200_power_dense; complex data as interleaved (re, im) doubles for both dtypes
// (dtype 0 callers pass zero imaginary parts).  poll = trips between host polls (engine default 64).
extern "C" __attribute__((visibility("default"))) int sim_power_dense(int n, const double* a, const double* x0, int max_it,
                                                                     double tol, int poll, double* x_out, double* lambda_out,
                                                                     int* iters, int* conv, int* launches_out) {
    Sim s;
    s.n = n;
    s.a = reinterpret_cast<const cd*>(a);
    s.begin(reinterpret_cast<const cd*>(x0), max_it, tol);
    if (max_it > 0) {  // mirrors MatrixImpl::run
        long long remaining = (long long)max_it + 1;
        if (poll < 1) poll = 1;
        while (remaining > 0) {
            const int k = int(remaining < poll ? remaining : poll);
            for (int i = 0; i < k; ++i) s.launch();
            remaining -= k;
            s.control(false);
            if (s.st.stop != STOP_NONE) break;
        }
    }
    s.control(false);  // MatrixImpl::end polls once more
    const double nrm = std::sqrt(s.st.final_q);
    const std::vector<cd>& v = s.buf[s.st.final_buf];
    for (int i = 0; i < n; ++i) {
        const cd x = v[size_t(i)] / nrm;
        x_out[2 * i] = x.real();
        x_out[2 * i + 1] = x.imag();
    }
    lambda_out[0] = s.st.lambda[0];
    lambda_out[1] = s.st.lambda[1];
    *iters = s.st.trips;
    *conv = s.st.converged;
    if (launches_out) *launches_out = s.issued;
    return s.st.stop == STOP_ERROR ? 1 : 0;
}
