// spec_wiener.hpp -- host-side (double precision) Wiener interpolation look-up tables for DRS based
// channel estimation.  Follows the reference's construction so that the same pilots and (within its
// own 1e-4 de-duplication threshold) the same weights are used:
//   lib/src/phy/rx/rx_synced/channel_estimation/channel_lut.cpp:169-634 (init_lut, s0..s3),
//   .../channel_estimation/wiener.hpp:113-139 (least-norm solve; here Rpp + sigma*I is positive definite,
//   so Gaussian elimination with partial pivoting returns the same unique solution as Eigen's COD),
//   .../channel_estimation/channel_statistics.cpp:29-35, lib/src/phy/filter/bessel.cpp:28-62.
//
// One table serves one (N_step_virtual, SNR class) pair of one beta:
//   idx_pilot[t][pc][k]   first pilot of the interpolation window inside the (interlaced) pilot vector
//   idx_weight[t][pc][k]  which weight vector to use
//   weights[v][n_interp]
// t = 0 .. N_step_virtual (OFDM symbol relative to the left DRS symbol), pc = position class of the
// transmit stream (0..3, DRS subcarrier offset of the LEFT pilot symbol), k = 0 .. N_b_OCC.
#pragma once
#include <cmath>
#include <cstdint>
#include <vector>
#include "spec_re.hpp"

namespace dnrp {
namespace spec {

struct wiener_lut_t {
    uint32_t n_interp = 0, n_t = 0, Nk = 0, n_vec = 0;
    std::vector<uint16_t> idx_pilot;   // [n_t][4][Nk]
    std::vector<uint16_t> idx_weight;  // [n_t][4][Nk]
    std::vector<float> weights;        // [n_vec][n_interp]
};

inline double wiener_sinc(float n) {  // filter::sinc, rectangular.cpp
    return n == 0.0f ? 1.0f : (float)(std::sin(M_PI * n) / (M_PI * n));
}
inline double wiener_bessel_j0(float zf) {  // 8-term series, bessel.cpp:28-62, evaluated in double, returned as float
    const double z = zf;
    double ret = 0.0;
    for (uint32_t k = 0; k <= 8; ++k) {
        double fac = 1.0;
        for (uint32_t i = 1; i <= k; ++i) fac *= (double)i;
        ret += std::pow(-0.25 * z * z, (double)k) / (fac * fac);
    }
    return (float)ret;
}

// solve A x = b (n <= 14), A destroyed
inline bool solve_dense(std::vector<double>& A, std::vector<double>& b, uint32_t n) {
    for (uint32_t c = 0; c < n; ++c) {
        uint32_t piv = c;
        for (uint32_t r = c + 1; r < n; ++r)
            if (std::fabs(A[r * n + c]) > std::fabs(A[piv * n + c])) piv = r;
        if (A[piv * n + c] == 0.0) return false;
        if (piv != c) {
            for (uint32_t k = 0; k < n; ++k) std::swap(A[c * n + k], A[piv * n + k]);
            std::swap(b[c], b[piv]);
        }
        for (uint32_t r = c + 1; r < n; ++r) {
            const double f = A[r * n + c] / A[c * n + c];
            if (f == 0.0) continue;
            for (uint32_t k = c; k < n; ++k) A[r * n + k] -= f * A[c * n + k];
            b[r] -= f * b[c];
        }
    }
    for (int r = (int)n - 1; r >= 0; --r) {
        double s = b[r];
        for (uint32_t k = r + 1; k < n; ++k) s -= A[r * n + k] * b[k];
        b[r] = s / A[r * n + r];
    }
    return true;
}

inline bool build_wiener_lut(uint32_t b, uint32_t u_max, uint32_t N_step_virtual, double nu_max_hz,
                             double tau_rms_sec, double snr_db, uint32_t n_interp, wiener_lut_t& L) {
    const uint32_t N_b_OCC = 56 * b, Nk = N_b_OCC + 1, NP = N_b_OCC / 4;
    const uint32_t n_t = N_step_virtual + 1;
    const double delta_u_f = (double)u_max * SUBCARRIER_SPACING_MIN;
    const double T_u_symb = (64.0 + 8.0) / 64.0 / delta_u_f;
    const double sigma = 1.0 / std::pow(10.0, snr_db / 10.0);
    L.n_interp = n_interp; L.n_t = n_t; L.Nk = Nk;
    L.idx_pilot.assign((size_t)n_t * 4 * Nk, 0);
    L.idx_weight.assign((size_t)n_t * 4 * Nk, 0);
    L.weights.clear();
    std::vector<std::vector<float>> vecs;

    struct coord { double f, t; };
    auto corr = [&](const coord& A, const coord& B) -> double {
        // s2_calc_correlation_noisefree: float arguments like the reference's float functions
        const double df = B.f - A.f, dt = B.t - A.t;
        const double rf = wiener_sinc((float)M_PI * (float)tau_rms_sec * (float)(df * delta_u_f));
        const double rt = wiener_bessel_j0(2.0f * (float)M_PI * (float)nu_max_hz * (float)(dt * T_u_symb));
        return rf * rt;
    };

    for (uint32_t pc = 0; pc < 4; ++pc) {
        // s0: pilot coordinates in the order the receiver stores them (interlaced left/right)
        std::vector<coord> P;
        for (uint32_t i = 0; i < NP; ++i) {
            const double kl = occ2k(4 * i + pc, N_b_OCC), kr = occ2k(4 * i + (pc + 2) % 4, N_b_OCC);
            if (N_step_virtual == 0) {
                P.push_back({kl, 1.0});
            } else if (pc <= 1) {
                P.push_back({kl, 1.0});
                P.push_back({kr, 1.0 + N_step_virtual});
            } else {
                P.push_back({kr, 1.0 + N_step_virtual});
                P.push_back({kl, 1.0});
            }
        }
        if (P.size() < n_interp) return false;
        for (uint32_t t = 1; t <= n_t; ++t) {
            uint32_t prev = 0;
            std::vector<double> Rpp_lu;  // kept between subcarriers that share the pilot window
            uint32_t have_for = 0xffffffffu;
            std::vector<double> Rpp(n_interp * n_interp);
            for (uint32_t f = 0; f < Nk; ++f) {
                // s1: window of n_interp consecutive pilots with the minimum summed distance,
                // searched upwards from the previous optimum with the 1.1x early abort
                double best = 1e9;
                uint32_t opt = 0;
                for (uint32_t i = prev; i + n_interp <= P.size(); ++i) {
                    double s = 0;
                    for (uint32_t j = i; j < i + n_interp; ++j)
                        s += std::sqrt((f - P[j].f) * (f - P[j].f) + (t - P[j].t) * (t - P[j].t));
                    if (s < best) { best = s; opt = i; }
                    if (s > best * 1.1) break;
                }
                // s2: weights = Rpp^-1 rdp, normalised to sum 1
                if (f == 0 || opt != have_for) {
                    for (uint32_t r = 0; r < n_interp; ++r)
                        for (uint32_t c2 = 0; c2 < n_interp; ++c2)
                            Rpp[r * n_interp + c2] = corr(P[opt + r], P[opt + c2]) + (r == c2 ? sigma : 0.0);
                    have_for = opt;
                }
                std::vector<double> A = Rpp, w(n_interp);
                for (uint32_t r = 0; r < n_interp; ++r) w[r] = corr({(double)f, (double)t}, P[opt + r]);
                if (!solve_dense(A, w, n_interp)) return false;
                double sum = 0;
                for (double v : w) sum += v;
                std::vector<float> wf(n_interp);
                for (uint32_t r = 0; r < n_interp; ++r) wf[r] = (float)(w[r] / sum);
                // s3: reuse an already known vector within 1e-4 (first match wins)
                int known = -1;
                for (size_t v = 0; v < vecs.size() && known < 0; ++v) {
                    float md = 0.f;
                    uint32_t mi = 0;
                    for (uint32_t r = 0; r < n_interp; ++r) {
                        const float dlt = std::fabs(wf[r] - vecs[v][r]);
                        if (dlt > md) { md = dlt; mi = r; }
                    }
                    (void)mi;
                    if ((double)md < 1e-4) known = (int)v;
                }
                if (known < 0) {
                    vecs.push_back(wf);
                    known = (int)vecs.size() - 1;
                }
                const size_t o = ((size_t)(t - 1) * 4 + pc) * Nk + f;
                L.idx_pilot[o] = (uint16_t)opt;
                L.idx_weight[o] = (uint16_t)known;
                prev = opt;
            }
        }
    }
    L.n_vec = (uint32_t)vecs.size();
    for (auto& v : vecs) L.weights.insert(L.weights.end(), v.begin(), v.end());
    return L.n_vec <= 0xffff;
}

}  // namespace spec
}  // namespace dnrp
