/* oracle/shim/volk/volk.h -- TEST INFRASTRUCTURE ONLY (never on the product path).
 *
 * Stand-in for the external VOLK library (un-vendored dependency of the reference,
 * VOLK >= 2.1.0, /root/reference/CMakeLists.txt:58) so that the reference's own kernel
 * sources compile unmodified into oracle/_ref/.  Each function is a plain scalar loop:
 * this is what the reference's QA pins, because its tests run with VOLK_GENERIC=1
 * (/root/reference/cmake/Modules/GrTest.cmake:64).  GCC auto-vectorises the loops at -O3.
 * Only the entry points the hot path calls are provided (call sites:
 * gr-filter/lib/fir_filter.cc:129-182, gr-filter/lib/fft_filter.cc:128,259,393,
 * gr-blocks/include/gnuradio/blocks/rotator.h:52).
 */
#ifndef GRB200_ORACLE_SHIM_VOLK_H
#define GRB200_ORACLE_SHIM_VOLK_H

#include <complex>
#include <cstddef>
#include <cstdint>
#include <cstdlib>

typedef std::complex<float> lv_32fc_t;

static inline size_t volk_get_alignment(void) { return 64; }

static inline void* volk_malloc(size_t size, size_t alignment)
{
    void* p = nullptr;
    if (alignment < sizeof(void*))
        alignment = sizeof(void*);
    if (posix_memalign(&p, alignment, size ? size : alignment) != 0)
        return nullptr;
    return p;
}

static inline void volk_free(void* p) { ::free(p); }

/* real . real */
static inline void
volk_32f_x2_dot_prod_32f_a(float* result, const float* a, const float* b, unsigned int n)
{
    float acc = 0.0f;
    for (unsigned int i = 0; i < n; i++)
        acc += a[i] * b[i];
    *result = acc;
}

/* complex input . real taps */
static inline void volk_32fc_32f_dot_prod_32fc_a(lv_32fc_t* result,
                                                 const lv_32fc_t* in,
                                                 const float* taps,
                                                 unsigned int n)
{
    const float* f = reinterpret_cast<const float*>(in);
    float re = 0.0f, im = 0.0f;
    for (unsigned int i = 0; i < n; i++) {
        re += f[2 * i] * taps[i];
        im += f[2 * i + 1] * taps[i];
    }
    *result = lv_32fc_t(re, im);
}

/* complex . complex (no conjugate) */
static inline void volk_32fc_x2_dot_prod_32fc_a(lv_32fc_t* result,
                                                const lv_32fc_t* a,
                                                const lv_32fc_t* b,
                                                unsigned int n)
{
    const float* x = reinterpret_cast<const float*>(a);
    const float* y = reinterpret_cast<const float*>(b);
    float re = 0.0f, im = 0.0f;
    for (unsigned int i = 0; i < n; i++) {
        re += x[2 * i] * y[2 * i] - x[2 * i + 1] * y[2 * i + 1];
        im += x[2 * i] * y[2 * i + 1] + x[2 * i + 1] * y[2 * i];
    }
    *result = lv_32fc_t(re, im);
}

/* int16 input . complex taps */
static inline void volk_16i_32fc_dot_prod_32fc_a(lv_32fc_t* result,
                                                 const short* in,
                                                 const lv_32fc_t* taps,
                                                 unsigned int n)
{
    const float* t = reinterpret_cast<const float*>(taps);
    float re = 0.0f, im = 0.0f;
    for (unsigned int i = 0; i < n; i++) {
        re += (float)in[i] * t[2 * i];
        im += (float)in[i] * t[2 * i + 1];
    }
    *result = lv_32fc_t(re, im);
}

/* real . real, saturated to int16 */
static inline void
volk_32f_x2_dot_prod_16i_a(int16_t* result, const float* a, const float* b, unsigned int n)
{
    float acc = 0.0f;
    for (unsigned int i = 0; i < n; i++)
        acc += a[i] * b[i];
    *result = (int16_t)acc;
}

/* element-wise complex multiply */
static inline void volk_32fc_x2_multiply_32fc_a(lv_32fc_t* c,
                                                const lv_32fc_t* a,
                                                const lv_32fc_t* b,
                                                unsigned int n)
{
    const float* x = reinterpret_cast<const float*>(a);
    const float* y = reinterpret_cast<const float*>(b);
    float* z = reinterpret_cast<float*>(c);
    for (unsigned int i = 0; i < n; i++) {
        const float xr = x[2 * i], xi = x[2 * i + 1], yr = y[2 * i], yi = y[2 * i + 1];
        z[2 * i] = xr * yr - xi * yi;
        z[2 * i + 1] = xr * yi + xi * yr;
    }
}

/* complex * real (used by the fft_vcc restatement) */
static inline void volk_32fc_32f_multiply_32fc(lv_32fc_t* c,
                                               const lv_32fc_t* a,
                                               const float* b,
                                               unsigned int n)
{
    const float* x = reinterpret_cast<const float*>(a);
    float* z = reinterpret_cast<float*>(c);
    for (unsigned int i = 0; i < n; i++) {
        z[2 * i] = x[2 * i] * b[i];
        z[2 * i + 1] = x[2 * i + 1] * b[i];
    }
}

/* |z|^2 (used by the complex_to_mag_squared restatement) */
static inline void
volk_32fc_magnitude_squared_32f(float* out, const lv_32fc_t* in, unsigned int n)
{
    const float* x = reinterpret_cast<const float*>(in);
    for (unsigned int i = 0; i < n; i++)
        out[i] = x[2 * i] * x[2 * i] + x[2 * i + 1] * x[2 * i + 1];
}

/* out[i] = in[i] * phase; phase *= phase_inc; renormalised every 512 items and at the end of every call
 * (VOLK's published generic kernel volk_32fc_s32fc_x2_rotator_32fc, ROTATOR_RELOAD = 512; call site
 * gr-blocks/include/gnuradio/blocks/rotator.h:50-53) */
static inline void volk_32fc_s32fc_x2_rotator_32fc(lv_32fc_t* out,
                                                   const lv_32fc_t* in,
                                                   const lv_32fc_t phase_inc,
                                                   lv_32fc_t* phase,
                                                   unsigned int n)
{
    unsigned int i = 0;
    for (unsigned int b = 0; b < n / 512; b++) {
        for (int j = 0; j < 512; j++) {
            *out++ = *in++ * (*phase);
            (*phase) *= phase_inc;
        }
        (*phase) /= hypotf(phase->real(), phase->imag());
    }
    for (i = 0; i < n % 512; i++) {
        *out++ = *in++ * (*phase);
        (*phase) *= phase_inc;
    }
    if (i)
        (*phase) /= hypotf(phase->real(), phase->imag());
}

#endif
