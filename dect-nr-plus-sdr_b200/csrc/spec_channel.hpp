// spec_channel.hpp -- host-side tapped-delay-line profiles of the doubly selective simulation channel (product code).
// Reference: lib/include/dectnrp/simulation/wireless/link.hpp:84-109 (three generic power delay profiles, taken there
// from srsRAN's fading.c), lib/src/simulation/wireless/link.cpp:72-146 (scaling to a requested rms delay spread,
// delays floored to samples, equal delays merged, powers normalised) and :290-322 (rms delay spread).
#pragma once
#include <cmath>
#include <cstdint>

namespace dnrp::spec {

constexpr uint32_t TDL_PROFILES = 3, TDL_MAX_TAPS = 9;

struct tdl_profile_t {
    uint32_t n;
    double delay_ns[TDL_MAX_TAPS];
    double power_dB[TDL_MAX_TAPS];
};

inline const tdl_profile_t& tdl_profile(uint32_t idx) {
    static const tdl_profile_t P[TDL_PROFILES] = {
        {7, {0, 30, 70, 90, 110, 190, 410, 0, 0}, {0.0, -1.0, -2.0, -3.0, -8.0, (double)-17.2f, (double)-20.8f, 0, 0}},
        {9, {0, 30, 150, 310, 370, 710, 1090, 1730, 2510},
         {0.0, -1.5, (double)-1.4f, (double)-3.6f, (double)-0.6f, (double)-9.1f, -7.0, -12.0, (double)-16.9f}},
        {9, {0, 50, 120, 200, 230, 500, 1600, 2300, 5000}, {-1.0, -1.0, -1.0, 0.0, 0.0, 0.0, -3.0, -5.0, -7.0}}};
    return P[idx];
}

// rms delay spread of a generic profile (second central moment of the normalised power delay profile)
inline double tdl_tau_rms_ns(uint32_t idx) {
    const tdl_profile_t& p = tdl_profile(idx);
    double lin[TDL_MAX_TAPS], total = 0.0, mean = 0.0, m2 = 0.0;
    for (uint32_t i = 0; i < p.n; ++i) { lin[i] = std::pow(10.0, p.power_dB[i] / 10.0); total += lin[i]; }
    for (uint32_t i = 0; i < p.n; ++i) mean += p.delay_ns[i] * (lin[i] / total);
    for (uint32_t i = 0; i < p.n; ++i) m2 += (p.delay_ns[i] - mean) * (p.delay_ns[i] - mean) * (lin[i] / total);
    return std::sqrt(m2);
}

// taps at the given sample rate for the requested rms delay spread; returns the number of (merged) taps
inline uint32_t tdl_taps(double samp_rate, uint32_t idx, float tau_rms_ns, uint32_t* delay_smpl, double* power_lin) {
    const tdl_profile_t& p = tdl_profile(idx);
    const double stretch = (double)tau_rms_ns / tdl_tau_rms_ns(idx), Ts = 1.0 / samp_rate;
    uint32_t n = 0;
    for (uint32_t i = 0; i < p.n; ++i) {
        const uint32_t d = (uint32_t)std::floor(p.delay_ns[i] * 1.0e-9 * stretch / Ts);
        const double pw = std::pow(10.0, p.power_dB[i] / 10.0);
        uint32_t k = 0;
        while (k < n && delay_smpl[k] != d) ++k;
        if (k == n) { delay_smpl[n] = d; power_lin[n] = pw; ++n; } else power_lin[k] += pw;
    }
    double total = 0.0;
    for (uint32_t i = 0; i < n; ++i) total += power_lin[i];
    for (uint32_t i = 0; i < n; ++i) power_lin[i] /= total;
    return n;
}

}  // namespace dnrp::spec
