// exact_math.h -- bit-reproducible single-precision libm routines for device code.
//
// The reference computes its per-sample CFO rotation with glibc's cosf/sinf
// (reference src/phy/phy.cpp:217-219, src/phy/LoRaDemod.cpp:154-156), its chirps with
// std::polar -> cosf/sinf (include/lora_phy/ChirpGenerator.hpp:37,45), the offset estimate with
// std::arg -> atan2f (src/phy/phy.cpp:118) and the detector with std::abs -> hypotf and log10f
// (include/lora_phy/LoRaDetector.hpp:60-71).  CUDA's own sinf/cosf/atan2f differ from glibc in the
// last bit for a few percent of arguments, which is enough to move a near-tie FFT argmax.  This
// header restates the *published algorithms* glibc 2.39 uses for those functions (the ARM
// optimized-routines sincosf with its double-precision polynomials, and the fdlibm float
// atanf/atan2f) with every floating-point operation spelled out un-fused, so that the CUDA build
// (IEEE double/float, no FMA contraction) and a plain C build produce identical bits.  The C build
// of this same header is compared against the host libm by scripts/check_exact_math.c (run, in its sampled mode,
// by tests/test_exact_math.py).
//
// Everything here is a pure function of its argument: no tables in memory except the 2/pi digits.
#pragma once

#include <stdint.h>

#if defined(__CUDACC__)
#define XM_HD __host__ __device__ __forceinline__
#else
#define XM_HD static inline
#endif

// ---- un-fused arithmetic -------------------------------------------------------------------------
// Device code uses the explicit round-to-nearest intrinsics, which the compiler never contracts into
// FMAs; host code relies on -ffp-contract=off (enforced by the build recipes).
//
// One step IS fused on purpose.  glibc selects its sinf/cosf at run time (ifunc): on any x86 with
// FMA it runs a build of the same source compiled with -mfma, where the one-step range reduction
// x - n*(pi/2) becomes a single fma.  Measured in the build container (glibc 2.39, FMA-capable Xeon)
// over ALL 4,278,190,080 finite floats (scripts/check_exact_math.c full):
//   XM_REDUCE_FUSED=1 : 0 sinf and 0 cosf mismatches against the host libm;
//   XM_REDUCE_FUSED=0 : 12 sinf + 22 cosf mismatches, every one with 17 < |x| < 120 (the fast
//                       reduction path, where cancellation exposes the rounding of n*(pi/2)).
// No other contraction in the polynomial changes a single result, so the fused reduction is the
// whole difference.  Set XM_REDUCE_FUSED=0 only to match a host CPU without FMA.
#ifndef XM_REDUCE_FUSED
#define XM_REDUCE_FUSED 1
#endif
#if defined(__CUDA_ARCH__)
#define XM_DFMA(a, b, c) __fma_rn((a), (b), (c))
#else
#define XM_DFMA(a, b, c) __builtin_fma((a), (b), (c))
#endif
#if defined(__CUDA_ARCH__)
#define XM_DMUL(a, b) __dmul_rn((a), (b))
#define XM_DADD(a, b) __dadd_rn((a), (b))
#define XM_DSUB(a, b) __dsub_rn((a), (b))
#define XM_FMUL(a, b) __fmul_rn((a), (b))
#define XM_FADD(a, b) __fadd_rn((a), (b))
#define XM_FSUB(a, b) __fsub_rn((a), (b))
#define XM_FDIV(a, b) __fdiv_rn((a), (b))
#else
#define XM_DMUL(a, b) ((a) * (b))
#define XM_DADD(a, b) ((a) + (b))
#define XM_DSUB(a, b) ((a) - (b))
#define XM_FMUL(a, b) ((a) * (b))
#define XM_FADD(a, b) ((a) + (b))
#define XM_FSUB(a, b) ((a) - (b))
#define XM_FDIV(a, b) ((a) / (b))
#endif

XM_HD uint32_t xm_f2u(float f)
{
#if defined(__CUDA_ARCH__)
    return __float_as_uint(f);
#else
    union { float f; uint32_t u; } v; v.f = f; return v.u;
#endif
}

XM_HD float xm_u2f(uint32_t u)
{
#if defined(__CUDA_ARCH__)
    return __uint_as_float(u);
#else
    union { float f; uint32_t u; } v; v.u = u; return v.f;
#endif
}

// ---- sinf / cosf ---------------------------------------------------------------------------------
// Algorithm: reduce to [-pi/4, pi/4] in double (one multiply-subtract for |x| < 120, a 32x96-bit
// fixed-point product with 2/pi beyond), then a degree-7 sine or degree-8 cosine polynomial in
// double, rounded once to float.

// Polynomial coefficients (double).  C[] cosine, S[] sine.
#define XM_C0 1.0
#define XM_C1 (-0x1.ffffffd0c621cp-2)
#define XM_C2 (0x1.55553e1068f19p-5)
#define XM_C3 (-0x1.6c087e89a359dp-10)
#define XM_C4 (0x1.99343027bf8c3p-16)
#define XM_S1 (-0x1.555545995a603p-3)
#define XM_S2 (0x1.1107605230bc4p-7)
#define XM_S3 (-0x1.994eb3774cf24p-13)
#define XM_HPI_INV_2P24 (0x1.45F306DC9C883p+23) /* 2/pi * 2^24 */
#define XM_HPI (0x1.921FB54442D18p0)            /* pi/2 */
#define XM_PI63 (0x1.921FB54442D18p-62)         /* pi/2 * 2^-62 */

// 2/pi as a string of bits; entry i is the 32-bit window that starts 24 bits before digit 8*i.
XM_HD uint32_t xm_inv_pio4(int i)
{
    // hex digits of 2/pi: A2F9836E 4E441529 FC2757D1 F534DDC0 DB629599 3C439041 ...
    const uint32_t w[24] = {
        0x000000a2u, 0x0000a2f9u, 0x00a2f983u, 0xa2f9836eu, 0xf9836e4eu, 0x836e4e44u,
        0x6e4e4415u, 0x4e441529u, 0x441529fcu, 0x1529fc27u, 0x29fc2757u, 0xfc2757d1u,
        0x2757d1f5u, 0x57d1f534u, 0xd1f534ddu, 0xf534ddc0u, 0x34ddc0dbu, 0xddc0db62u,
        0xc0db6295u, 0xdb629599u, 0x6295993cu, 0x95993c43u, 0x993c4390u, 0x3c439041u};
    return w[i];
}

// The three 32-bit windows the large-argument reduction multiplies by -- entries base, base+4 and base+8 of the table
// above -- are 96 consecutive bits of ONE bit string (24 zero bits, then the digits of 2/pi) starting 8*base bits in.
// Cutting them out of four 64-bit words with two funnel shifts costs ~15 instructions; indexing the table with a
// run-time subscript compiles to a 24-way select chain per entry on the GPU (~70 instructions each).
// On the device the sixteen possible triples live in constant memory instead: `base` comes from the argument's
// exponent, which the lanes of a warp almost always share, so the lookup is three uniform constant loads.
#if defined(__CUDACC__)
static __constant__ uint32_t xm_pio4_x3_table[16][4] = {
    {0x000000a2u, 0xf9836e4eu, 0x441529fcu, 0u},
    {0x0000a2f9u, 0x836e4e44u, 0x1529fc27u, 0u},
    {0x00a2f983u, 0x6e4e4415u, 0x29fc2757u, 0u},
    {0xa2f9836eu, 0x4e441529u, 0xfc2757d1u, 0u},
    {0xf9836e4eu, 0x441529fcu, 0x2757d1f5u, 0u},
    {0x836e4e44u, 0x1529fc27u, 0x57d1f534u, 0u},
    {0x6e4e4415u, 0x29fc2757u, 0xd1f534ddu, 0u},
    {0x4e441529u, 0xfc2757d1u, 0xf534ddc0u, 0u},
    {0x441529fcu, 0x2757d1f5u, 0x34ddc0dbu, 0u},
    {0x1529fc27u, 0x57d1f534u, 0xddc0db62u, 0u},
    {0x29fc2757u, 0xd1f534ddu, 0xc0db6295u, 0u},
    {0xfc2757d1u, 0xf534ddc0u, 0xdb629599u, 0u},
    {0x2757d1f5u, 0x34ddc0dbu, 0x6295993cu, 0u},
    {0x57d1f534u, 0xddc0db62u, 0x95993c43u, 0u},
    {0xd1f534ddu, 0xc0db6295u, 0x993c4390u, 0u},
    {0xf534ddc0u, 0xdb629599u, 0x3c439041u, 0u}};
#endif

XM_HD void xm_inv_pio4_x3(int base, uint32_t* w0, uint32_t* w1, uint32_t* w2)
{
#if defined(__CUDA_ARCH__)
    *w0 = xm_pio4_x3_table[base][0];
    *w1 = xm_pio4_x3_table[base][1];
    *w2 = xm_pio4_x3_table[base][2];
#else
    const uint64_t z0 = 0x000000A2F9836E4Eull, z1 = 0x441529FC2757D1F5ull, z2 = 0x34DDC0DB6295993Cull,
                   z3 = 0x4390410000000000ull;
    const int up = base >= 8;               // bit offset 8*base is 0..120: word 0 or word 1
    const uint64_t a = up ? z1 : z0, b = up ? z2 : z1, c = up ? z3 : z2;
    const int r = (base & 7) * 8;           // 0, 8, ..., 56
    const uint64_t hi = r ? ((a << r) | (b >> (64 - r))) : a;
    const uint64_t lo = r ? ((b << r) | (c >> (64 - r))) : b;
    *w0 = (uint32_t)(hi >> 32);
    *w1 = (uint32_t)hi;
    *w2 = (uint32_t)(lo >> 32);
#endif
}

// Evaluate the sine (odd==0) or cosine (odd==1) polynomial.  `neg` flips the cosine's sign (the
// reference algorithm swaps in a negated coefficient set; negating every coefficient is an exact
// operation, so flipping the result is bit-identical).
XM_HD float xm_sincos_poly(double x, double x2, int odd, int neg)
{
    if (!odd) {
        double x3 = XM_DMUL(x, x2);
        double s1 = XM_DADD(XM_S2, XM_DMUL(x2, XM_S3));
        double x7 = XM_DMUL(x3, x2);
        double s = XM_DADD(x, XM_DMUL(x3, XM_S1));
        return (float)XM_DADD(s, XM_DMUL(x7, s1));
    } else {
        double x4 = XM_DMUL(x2, x2);
        double c2 = XM_DADD(XM_C3, XM_DMUL(x2, XM_C4));
        double c1 = XM_DADD(XM_C0, XM_DMUL(x2, XM_C1));
        double x6 = XM_DMUL(x4, x2);
        double c = XM_DADD(c1, XM_DMUL(x4, XM_C2));
        double r = XM_DADD(c, XM_DMUL(x6, c2));
        return (float)(neg ? -r : r);
    }
}

// Quadrant reduction shared by sine and cosine.  Returns the reduced argument (already multiplied by
// the quadrant sign) and x*x; *quad receives the quadrant used to pick polynomial and sign.
XM_HD void xm_reduce(float y, double* xr, double* x2, int* quad_odd, int* quad_neg)
{
    const uint32_t bits = xm_f2u(y);
    const uint32_t top = (bits >> 20) & 0x7ffu;
    double x = (double)y;
    int n;
    if (top < ((0x42f00000u >> 20) & 0x7ffu)) { // |y| < 120
        double r = XM_DMUL(x, XM_HPI_INV_2P24);
        n = ((int32_t)r + 0x800000) >> 24;
#if XM_REDUCE_FUSED
        x = XM_DFMA(-(double)n, XM_HPI, x);
#else
        x = XM_DSUB(x, XM_DMUL((double)n, XM_HPI));
#endif
        const double sgn = ((n & 3) == 1 || (n & 3) == 2) ? -1.0 : 1.0;
        *quad_neg = (n & 2) != 0;
        *quad_odd = n & 1;
        *x2 = XM_DMUL(x, x);
        *xr = XM_DMUL(x, sgn);
    } else {
        const int sign = (int)(bits >> 31);
        const int base = (int)((bits >> 26) & 15u);
        const int shift = (int)((bits >> 23) & 7u);
        uint32_t xi = (bits & 0xffffffu) | 0x800000u;
        xi <<= shift;
        uint32_t w0, w1, w2;
        xm_inv_pio4_x3(base, &w0, &w1, &w2);
        uint64_t res0 = (uint64_t)(uint32_t)(xi * w0);
        const uint64_t res1 = (uint64_t)xi * w1;
        const uint64_t res2 = (uint64_t)xi * w2;
        res0 = (res2 >> 32) | (res0 << 32);
        res0 += res1;
        const uint64_t nn = (res0 + (1ull << 61)) >> 62;
        res0 -= nn << 62;
        x = XM_DMUL((double)(int64_t)res0, XM_PI63);
        n = (int)nn;
        const int q = (n + sign) & 3;
        const double sgn = (q == 1 || q == 2) ? -1.0 : 1.0;
        *quad_neg = ((n + sign) & 2) != 0;
        *quad_odd = n & 1;
        *x2 = XM_DMUL(x, x);
        *xr = XM_DMUL(x, sgn);
    }
}

XM_HD float xm_sinf(float y)
{
    const uint32_t top = (xm_f2u(y) >> 20) & 0x7ffu;
    if (top < ((0x3f490fdbu >> 20) & 0x7ffu)) { // |y| < pi/4 (by exponent/top-mantissa class)
        if (top < ((0x39800000u >> 20) & 0x7ffu)) return y; // |y| < 2^-12
        double x = (double)y;
        return xm_sincos_poly(x, XM_DMUL(x, x), 0, 0);
    }
    if (top >= 0x7f8u) return y - y; // inf / nan
    double xr, x2; int odd, neg;
    xm_reduce(y, &xr, &x2, &odd, &neg);
    return xm_sincos_poly(xr, x2, odd, neg);
}

XM_HD float xm_cosf(float y)
{
    const uint32_t top = (xm_f2u(y) >> 20) & 0x7ffu;
    if (top < ((0x3f490fdbu >> 20) & 0x7ffu)) {
        if (top < ((0x39800000u >> 20) & 0x7ffu)) return 1.0f;
        double x = (double)y;
        return xm_sincos_poly(x, XM_DMUL(x, x), 1, 0);
    }
    if (top >= 0x7f8u) return y - y;
    double xr, x2; int odd, neg;
    xm_reduce(y, &xr, &x2, &odd, &neg);
    return xm_sincos_poly(xr, x2, odd ^ 1, neg);
}

// Both at once (one reduction).  Same bits as xm_sinf / xm_cosf.
XM_HD void xm_sincosf(float y, float* s, float* c)
{
    const uint32_t top = (xm_f2u(y) >> 20) & 0x7ffu;
    if (top < ((0x3f490fdbu >> 20) & 0x7ffu)) {
        if (top < ((0x39800000u >> 20) & 0x7ffu)) { *s = y; *c = 1.0f; return; }
        double x = (double)y;
        double x2 = XM_DMUL(x, x);
        *s = xm_sincos_poly(x, x2, 0, 0);
        *c = xm_sincos_poly(x, x2, 1, 0);
        return;
    }
    if (top >= 0x7f8u) { *s = y - y; *c = y - y; return; }
    double xr, x2; int odd, neg;
    xm_reduce(y, &xr, &x2, &odd, &neg);
    *s = xm_sincos_poly(xr, x2, odd, neg);
    *c = xm_sincos_poly(xr, x2, odd ^ 1, neg);
}

// Branch-free form of xm_sincosf for code that evaluates several arguments side by side (the chirp modulator
// interleaves four): no divergent path for a scheduler to serialise, each polynomial evaluated exactly once with
// sine / cosine picking theirs.  `use_small` / `use_large` say which range reductions the caller needs: pass
// compile-time 1, 1 when nothing is known about y (both are computed, one selected), 1, 0 when the caller knows
// |y| < 120 (e.g. from a warp vote), 0, 1 when it knows |y| >= 120.  Same bits as xm_sincosf for every float
// (scripts/check_exact_math.c runs all three forms over all finite floats).  Why the separate |y| < pi/4 path of
// xm_sincosf is not needed: there the fast reduction yields n = 0 and fma(-0, pi/2, x) = x exactly, i.e. the same
// polynomial inputs.
XM_HD void xm_sincosf_paths(float y, float* s, float* c, const int use_small, const int use_large)
{
    const uint32_t bits = xm_f2u(y);
    const uint32_t top = (bits >> 20) & 0x7ffu;
    const double x = (double)y;
    double xs = 0.0, xl = 0.0;
    int ns = 0, nl = 0;
    const int sign = (int)(bits >> 31);
    if (use_small) { // |y| < 120: one multiply and one fused multiply-subtract
        const double r = XM_DMUL(x, XM_HPI_INV_2P24);
#if defined(__CUDA_ARCH__)
        const int32_t ri = __double2int_rz(r); // saturates for arguments the selection below discards
#else
        const int32_t ri = (r > -2.0e9 && r < 2.0e9) ? (int32_t)r : 0;
#endif
        ns = (ri + 0x800000) >> 24;
        xs = XM_DFMA(-(double)ns, XM_HPI, x);
    }
    if (use_large) { // |y| >= 120: 24-bit mantissa times 96 bits of 2/pi
        const int base = (int)((bits >> 26) & 15u);
        const int shift = (int)((bits >> 23) & 7u);
        uint32_t xi = (bits & 0xffffffu) | 0x800000u;
        xi <<= shift;
        uint32_t w0, w1, w2;
        xm_inv_pio4_x3(base, &w0, &w1, &w2);
        uint64_t res0 = (uint64_t)(uint32_t)(xi * w0);
        const uint64_t res1 = (uint64_t)xi * w1;
        const uint64_t res2 = (uint64_t)xi * w2;
        res0 = (res2 >> 32) | (res0 << 32);
        res0 += res1;
        const uint64_t nn = (res0 + (1ull << 61)) >> 62;
        res0 -= nn << 62;
        xl = XM_DMUL((double)(int64_t)res0, XM_PI63);
        nl = (int)nn;
    }
    const int big = use_large && (!use_small || top >= ((0x42f00000u >> 20) & 0x7ffu));
    const double xx = big ? xl : xs;
    const int n = big ? nl : ns;
    const int q = big ? ((n + sign) & 3) : (n & 3);
    const double sgn = (q == 1 || q == 2) ? -1.0 : 1.0;
    const int neg = (q & 2) != 0, odd = n & 1;
    const double x2 = XM_DMUL(xx, xx);
    const double xr = XM_DMUL(xx, sgn);
    const float ps = xm_sincos_poly(xr, x2, 0, 0);   // sine polynomial
    const float pc = xm_sincos_poly(xr, x2, 1, neg); // cosine polynomial, sign applied
    float sv = odd ? pc : ps, cv = odd ? ps : pc;
    if (top < ((0x39800000u >> 20) & 0x7ffu)) { sv = y; cv = 1.0f; } // |y| < 2^-12
    if (top >= 0x7f8u) { sv = y - y; cv = y - y; }                    // inf / nan
    *s = sv;
    *c = cv;
}

XM_HD void xm_sincosf_bf(float y, float* s, float* c) { xm_sincosf_paths(y, s, c, 1, 1); }

// |y| >= 120 by exponent class (the test glibc's sincosf makes)
XM_HD int xm_is_large_arg(float y) { return ((xm_f2u(y) >> 20) & 0x7ffu) >= ((0x42f00000u >> 20) & 0x7ffu); }

// ---- atanf / atan2f (fdlibm float algorithm as shipped in glibc 2.39) ------------------------------
XM_HD float xm_atanf(float x)
{
    const float atanhi0 = 4.6364760399e-01f, atanhi1 = 7.8539812565e-01f,
                atanhi2 = 9.8279368877e-01f, atanhi3 = 1.5707962513e+00f;
    const float atanlo0 = 5.0121582440e-09f, atanlo1 = 3.7748947079e-08f,
                atanlo2 = 3.4473217170e-08f, atanlo3 = 7.5497894159e-08f;
    const float aT0 = 3.3333334327e-01f, aT1 = -2.0000000298e-01f, aT2 = 1.4285714924e-01f,
                aT3 = -1.1111110449e-01f, aT4 = 9.0908870101e-02f, aT5 = -7.6918758452e-02f,
                aT6 = 6.6610731184e-02f, aT7 = -5.8335702866e-02f, aT8 = 4.9768779427e-02f,
                aT9 = -3.6531571299e-02f, aT10 = 1.6285819933e-02f;
    const int32_t hx = (int32_t)xm_f2u(x);
    const int32_t ix = hx & 0x7fffffff;
    int id;
    if (ix >= 0x4c000000) { // |x| >= 2^25
        if (ix > 0x7f800000) return XM_FADD(x, x); // NaN
        if (hx > 0) return XM_FADD(atanhi3, atanlo3);
        return XM_FSUB(-atanhi3, atanlo3);
    }
    if (ix < 0x3ee00000) { // |x| < 0.4375
        if (ix < 0x31000000) return x; // |x| < 2^-29
        id = -1;
    } else {
        x = xm_u2f((uint32_t)ix); // fabsf
        if (ix < 0x3f980000) {     // |x| < 1.1875
            if (ix < 0x3f300000) { // 7/16 <= |x| < 11/16
                id = 0;
                x = XM_FDIV(XM_FSUB(XM_FMUL(2.0f, x), 1.0f), XM_FADD(2.0f, x));
            } else {               // 11/16 <= |x| < 19/16
                id = 1;
                x = XM_FDIV(XM_FSUB(x, 1.0f), XM_FADD(x, 1.0f));
            }
        } else {
            if (ix < 0x401c0000) { // |x| < 2.4375
                id = 2;
                x = XM_FDIV(XM_FSUB(x, 1.5f), XM_FADD(1.0f, XM_FMUL(1.5f, x)));
            } else {               // 2.4375 <= |x| < 2^25
                id = 3;
                x = XM_FDIV(-1.0f, x);
            }
        }
    }
    const float z = XM_FMUL(x, x);
    const float w = XM_FMUL(z, z);
    // odd and even halves of the polynomial, Horner in w
    float s1 = XM_FMUL(w, aT10);
    s1 = XM_FMUL(w, XM_FADD(aT8, s1));
    s1 = XM_FMUL(w, XM_FADD(aT6, s1));
    s1 = XM_FMUL(w, XM_FADD(aT4, s1));
    s1 = XM_FMUL(w, XM_FADD(aT2, s1));
    s1 = XM_FMUL(z, XM_FADD(aT0, s1));
    float s2 = XM_FMUL(w, aT9);
    s2 = XM_FMUL(w, XM_FADD(aT7, s2));
    s2 = XM_FMUL(w, XM_FADD(aT5, s2));
    s2 = XM_FMUL(w, XM_FADD(aT3, s2));
    s2 = XM_FMUL(w, XM_FADD(aT1, s2));
    if (id < 0) return XM_FSUB(x, XM_FMUL(x, XM_FADD(s1, s2)));
    const float hi = id == 0 ? atanhi0 : id == 1 ? atanhi1 : id == 2 ? atanhi2 : atanhi3;
    const float lo = id == 0 ? atanlo0 : id == 1 ? atanlo1 : id == 2 ? atanlo2 : atanlo3;
    const float r = XM_FSUB(hi, XM_FSUB(XM_FSUB(XM_FMUL(x, XM_FADD(s1, s2)), lo), x));
    return (hx < 0) ? -r : r;
}

XM_HD float xm_atan2f(float y, float x)
{
    const float tiny = 1.0e-30f, pi_o_4 = 7.8539818525e-01f, pi_o_2 = 1.5707963705e+00f,
                pi = 3.1415927410e+00f, pi_lo = -8.7422776573e-08f;
    const int32_t hx = (int32_t)xm_f2u(x), hy = (int32_t)xm_f2u(y);
    const int32_t ix = hx & 0x7fffffff, iy = hy & 0x7fffffff;
    if (ix > 0x7f800000 || iy > 0x7f800000) return XM_FADD(x, y); // NaN
    if (hx == 0x3f800000) return xm_atanf(y);                     // x == 1
    const int m = ((hy >> 31) & 1) | ((hx >> 30) & 2);
    if (iy == 0) {
        switch (m) {
        case 0:
        case 1: return y;
        case 2: return XM_FADD(pi, tiny);
        default: return XM_FSUB(-pi, tiny);
        }
    }
    if (ix == 0) return (hy < 0) ? XM_FSUB(-pi_o_2, tiny) : XM_FADD(pi_o_2, tiny);
    if (ix == 0x7f800000) {
        if (iy == 0x7f800000) {
            switch (m) {
            case 0: return XM_FADD(pi_o_4, tiny);
            case 1: return XM_FSUB(-pi_o_4, tiny);
            case 2: return XM_FADD(XM_FMUL(3.0f, pi_o_4), tiny);
            default: return XM_FSUB(XM_FMUL(-3.0f, pi_o_4), tiny);
            }
        } else {
            switch (m) {
            case 0: return 0.0f;
            case 1: return -0.0f;
            case 2: return XM_FADD(pi, tiny);
            default: return XM_FSUB(-pi, tiny);
            }
        }
    }
    if (iy == 0x7f800000) return (hy < 0) ? XM_FSUB(-pi_o_2, tiny) : XM_FADD(pi_o_2, tiny);
    const int32_t k = (iy - ix) >> 23;
    float z;
    if (k > 60) z = XM_FADD(pi_o_2, XM_FMUL(0.5f, pi_lo));
    else if (hx < 0 && k < -60) z = 0.0f;
    else z = xm_atanf(xm_u2f(xm_f2u(XM_FDIV(y, x)) & 0x7fffffffu));
    switch (m) {
    case 0: return z;
    case 1: return xm_u2f(xm_f2u(z) ^ 0x80000000u);
    case 2: return XM_FSUB(pi, XM_FSUB(z, pi_lo));
    default: return XM_FSUB(XM_FSUB(z, pi_lo), pi);
    }
}

// ---- hypotf: sqrt of the exact double sum of squares, rounded once to float ----------------------
XM_HD float xm_hypotf(float re, float im)
{
    const uint32_t ar = xm_f2u(re) & 0x7fffffffu, ai = xm_f2u(im) & 0x7fffffffu;
    if (ar == 0x7f800000u || ai == 0x7f800000u) return xm_u2f(0x7f800000u);
    if (ar > 0x7f800000u || ai > 0x7f800000u) return XM_FADD(re, im);
    const double a = (double)re, b = (double)im;
    const double s = XM_DADD(XM_DMUL(a, a), XM_DMUL(b, b));
#if defined(__CUDA_ARCH__)
    return (float)__dsqrt_rn(s);
#else
    return (float)__builtin_sqrt(s);
#endif
}
