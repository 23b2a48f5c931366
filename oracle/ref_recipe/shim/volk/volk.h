/* <volk/volk.h> stand-in: the generic-C semantics (VOLK 3.1 *_generic kernels) of the kernels the reference calls.
 * Test infrastructure (oracle/_ref build); VOLK is not installed on this image. */
#ifndef SHIM_VOLK_H
#define SHIM_VOLK_H
#include <cmath>
#include <complex>
#include <cstdint>
#include <cstdlib>

typedef std::complex<float> lv_32fc_t;
#define lv_cmake(r, i) lv_32fc_t((r), (i))
#define lv_creal(x) ((x).real())
#define lv_cimag(x) ((x).imag())

static inline size_t volk_get_alignment() { return 64; }
static inline void* volk_malloc(size_t size, size_t alignment) {
    void* p = nullptr;
    if (posix_memalign(&p, alignment < sizeof(void*) ? sizeof(void*) : alignment, size ? size : alignment)) return nullptr;
    return p;
}
static inline void volk_free(void* p) { free(p); }

/* complex products are written out (no NaN-recovery path of operator*): real arithmetic only, as the C kernels do */
static inline lv_32fc_t shim_cmul(const lv_32fc_t a, const lv_32fc_t b) {
    return lv_32fc_t(a.real() * b.real() - a.imag() * b.imag(), a.real() * b.imag() + a.imag() * b.real());
}
static inline lv_32fc_t shim_cmulc(const lv_32fc_t a, const lv_32fc_t b) {   /* a * conj(b) */
    return lv_32fc_t(a.real() * b.real() + a.imag() * b.imag(), a.imag() * b.real() - a.real() * b.imag());
}

static inline void volk_32fc_32f_dot_prod_32fc_u(lv_32fc_t* result, const lv_32fc_t* input, const float* taps, unsigned int n) {
    float re = 0.f, im = 0.f;
    for (unsigned int i = 0; i < n; ++i) { re += input[i].real() * taps[i]; im += input[i].imag() * taps[i]; }
    *result = lv_32fc_t(re, im);
}
static inline void volk_32fc_x2_dot_prod_32fc_u(lv_32fc_t* result, const lv_32fc_t* a, const lv_32fc_t* b, unsigned int n) {
    lv_32fc_t s(0.f, 0.f);
    for (unsigned int i = 0; i < n; ++i) s += shim_cmul(a[i], b[i]);
    *result = s;
}
static inline void volk_32fc_x2_conjugate_dot_prod_32fc(lv_32fc_t* result, const lv_32fc_t* input, const lv_32fc_t* taps, unsigned int n) {
    lv_32fc_t s(0.f, 0.f);
    for (unsigned int i = 0; i < n; ++i) s += shim_cmulc(input[i], taps[i]);
    *result = s;
}
static inline void volk_32fc_x2_multiply_conjugate_32fc(lv_32fc_t* c, const lv_32fc_t* a, const lv_32fc_t* b, unsigned int n) {
    for (unsigned int i = 0; i < n; ++i) c[i] = shim_cmulc(a[i], b[i]);
}
static inline void volk_32fc_x2_divide_32fc(lv_32fc_t* c, const lv_32fc_t* a, const lv_32fc_t* b, unsigned int n) {
    for (unsigned int i = 0; i < n; ++i) {
        const float d = b[i].real() * b[i].real() + b[i].imag() * b[i].imag();
        const lv_32fc_t t = shim_cmulc(a[i], b[i]);
        c[i] = lv_32fc_t(t.real() / d, t.imag() / d);
    }
}
static inline void volk_32fc_x2_add_32fc(lv_32fc_t* c, const lv_32fc_t* a, const lv_32fc_t* b, unsigned int n) {
    for (unsigned int i = 0; i < n; ++i) c[i] = a[i] + b[i];
}
static inline void volk_32fc_32f_multiply_32fc(lv_32fc_t* c, const lv_32fc_t* a, const float* b, unsigned int n) {
    for (unsigned int i = 0; i < n; ++i) c[i] = lv_32fc_t(a[i].real() * b[i], a[i].imag() * b[i]);
}
static inline void volk_32f_s32f_multiply_32f(float* c, const float* a, const float scalar, unsigned int n) {
    for (unsigned int i = 0; i < n; ++i) c[i] = a[i] * scalar;
}
static inline void volk_32f_x2_subtract_32f(float* c, const float* a, const float* b, unsigned int n) {
    for (unsigned int i = 0; i < n; ++i) c[i] = a[i] - b[i];
}
static inline void volk_32fc_s32fc_multiply2_32fc(lv_32fc_t* c, const lv_32fc_t* a, const lv_32fc_t* scalar, unsigned int n) {
    for (unsigned int i = 0; i < n; ++i) c[i] = shim_cmul(a[i], *scalar);
}
static inline void volk_32fc_accumulator_s32fc(lv_32fc_t* result, const lv_32fc_t* in, unsigned int n) {
    lv_32fc_t s(0.f, 0.f);
    for (unsigned int i = 0; i < n; ++i) s += in[i];
    *result = s;
}
static inline void volk_32fc_accumulator_s32fc_a(lv_32fc_t* result, const lv_32fc_t* in, unsigned int n) {
    volk_32fc_accumulator_s32fc(result, in, n);
}
static inline void volk_32f_accumulator_s32f(float* result, const float* in, unsigned int n) {
    float s = 0.f;
    for (unsigned int i = 0; i < n; ++i) s += in[i];
    *result = s;
}
static inline void volk_32fc_deinterleave_real_32f(float* out, const lv_32fc_t* in, unsigned int n) {
    for (unsigned int i = 0; i < n; ++i) out[i] = in[i].real();
}
static inline void volk_32fc_index_max_32u(uint32_t* target, const lv_32fc_t* src, uint32_t n) {
    float m = -1.f;
    uint32_t idx = 0;
    for (uint32_t i = 0; i < n; ++i) {
        const float a = src[i].real() * src[i].real() + src[i].imag() * src[i].imag();
        if (a > m) { m = a; idx = i; }
    }
    *target = idx;
}
/* out[i] = in[i] * phase; phase *= phase_inc; magnitude renormalised every 512 samples and at the end (generic kernel) */
static inline void volk_32fc_s32fc_x2_rotator2_32fc(lv_32fc_t* out, const lv_32fc_t* in, const lv_32fc_t* phase_inc, lv_32fc_t* phase,
                                                    unsigned int n) {
    unsigned int i = 0, j = 0;
    for (i = 0; i < n / 512; ++i) {
        for (j = 0; j < 512; ++j) {
            *out++ = shim_cmul(*in++, *phase);
            *phase = shim_cmul(*phase, *phase_inc);
        }
        const float mag = std::sqrt(phase->real() * phase->real() + phase->imag() * phase->imag());
        *phase = lv_32fc_t(phase->real() / mag, phase->imag() / mag);
    }
    for (i = 0; i < n % 512; ++i) {
        *out++ = shim_cmul(*in++, *phase);
        *phase = shim_cmul(*phase, *phase_inc);
    }
    if (i) {
        const float mag = std::sqrt(phase->real() * phase->real() + phase->imag() * phase->imag());
        *phase = lv_32fc_t(phase->real() / mag, phase->imag() / mag);
    }
}
#endif
