// spec_resample.hpp -- host-side design of the rational resampler's polyphase low pass (product code).
// Follows the reference's filter definition so that both sides use the same coefficients:
//   lib/src/phy/resample/resampler.cpp:55-140 (merged image/anti-alias Kaiser low pass at max(L, M), gain L, delay
//   compensation), lib/src/phy/filter/kaiser.cpp:37-118, rectangular.cpp:28-46, bessel.cpp:28-62,
//   lib/include/dectnrp/phy/resample/resampler_param.hpp:61-86 (pass / stop / attenuation per oversampling).
// The streaming state machine of the reference (history, skipped samples, subfilter order) collapses to
//   y[m] = sum_j taps[j][l] * x[i0 - j],   p = delay + m*M,  l = p % L,  i0 = p / L,   x = 0 outside the stream,
// which is what k_resample evaluates for every output sample independently.
#pragma once
#include <cmath>
#include <cstdint>
#include <vector>

namespace dnrp::spec {

struct resampler_design_t {
    uint32_t L = 1, M = 1;
    uint32_t filter_length = 1, subfilter_length = 1, delay = 0;
    std::vector<float> taps;  // [subfilter_length][L]: taps[j * L + l] = L * h[l + j * L]

    bool bypass() const { return L == M; }
    // resampler_t::get_N_samples_after_resampling (resampler.cpp:310-328): outputs of n_in samples incl. the flush
    uint32_t n_after(uint32_t n_in) const {
        if (bypass()) return n_in;
        const uint64_t last = (uint64_t)(n_in + subfilter_length - 1) * L - 1;  // largest usable up-sampled index
        return last < delay ? 0 : (uint32_t)((last - delay) / M + 1);
    }
};

inline float resample_bessel_i0(float zf) {
    const double q = 0.25 * (double)zf * (double)zf;
    double acc = 0.0, fac = 1.0;
    for (uint32_t k = 0; k <= 8; ++k) {
        if (k) fac *= (double)k;
        acc += std::pow(q, (double)k) / (fac * fac);
    }
    return (float)acc;
}

// oversampling -> (pass, stop, stop band attenuation dB); 0 if the reference defines none for this os
inline bool resampler_filter_def(uint32_t os, float& f_pass, float& f_stop, float& att_dB) {
    switch (os) {
        case 1: f_pass = 0.48f; f_stop = 0.499f; att_dB = 14.0f; return true;
        case 2: f_pass = 0.30f; f_stop = 0.499f; att_dB = 20.0f; return true;
        case 4: f_pass = 0.20f; f_stop = 0.499f; att_dB = 20.0f; return true;
        case 8: f_pass = 0.15f; f_stop = 0.499f; att_dB = 20.0f; return true;
        default: return false;
    }
}

inline bool design_resampler(uint32_t L, uint32_t M, uint32_t os, resampler_design_t& out) {
    out = resampler_design_t{};
    out.L = L; out.M = M;
    if (L == M) { out.L = out.M = 1; out.taps.assign(1, 1.0f); return true; }
    float fp, fs, att;
    if (!resampler_filter_def(os, fp, fs, att)) return false;
    const float wider = (float)(L > M ? L : M);
    const float f_pass = fp / wider, f_stop = fs / wider;
    // Kaiser parameters: attenuation from the stricter of stop band attenuation and pass band ripple (100 dB = unused)
    const float delta_stop = std::pow(10.0f, -att / 20.0f), delta_pass = std::pow(10.0f, 100.0f / 20.0f) - 1.0f;
    const float A = -20.0f * std::log10(delta_stop < delta_pass ? delta_stop : delta_pass);
    float beta = 0.0f;
    if (A > 50.0f) beta = 0.1102f * (A - 8.7f);
    else if (A >= 21.0f) beta = 0.5842f * std::pow(A - 21.0f, 0.4f) + 0.07886 * (A - 21.0f);
    const float width = f_stop - f_pass;
    const float order = (A - 7.95f) / (2.285f * 2.0f * M_PI * width);
    uint32_t N = (uint32_t)std::ceil(order + 1.0f);
    N |= 1u;  // odd: integer delay
    const float cutoff = f_pass + width / 2.0f, Nf = (float)N;
    std::vector<float> h(N);
    float sum = 0.0f;
    for (uint32_t n = 0; n < N; ++n) {
        const float nf = (float)n;
        const float win = resample_bessel_i0(beta * std::sqrt(1.0f - std::pow(2.0f * nf / (Nf - 1.0f) - 1.0f, 2.0f))) /
                          resample_bessel_i0(beta);
        const float t = 2.0f * cutoff * (nf - (Nf - 1.0f) / 2.0f);
        const float sinc = t == 0.0f ? 1.0f : std::sin(M_PI * t) / (M_PI * t);
        h[n] = win * (2.0f * cutoff * sinc);
        sum += h[n];
    }
    out.filter_length = N;
    out.delay = (N - 1) / 2;
    out.subfilter_length = (N + L - 1) / L;
    out.taps.assign((size_t)out.subfilter_length * L, 0.0f);
    for (uint32_t n = 0; n < N; ++n) out.taps[(n / L) * L + (n % L)] = (h[n] / sum) * (float)L;
    // the reference's subfilter schedule needs the first output to fall on an input sample that yields one output
    // (resampler.cpp:174-176 "known limitation"); the closed form has no such restriction
    return true;
}

}  // namespace dnrp::spec
