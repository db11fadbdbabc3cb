/*
 * mrisim_math.h -- the WALKER STREAM CONTRACT, as code that compiles identically for the
 * host (gcc, -ffp-contract=off -mfma) and for sm_100a (nvcc, -fmad=false).
 *
 * Why this exists.  The reference draws walker randomness with cuRAND (XORWOW, default
 * ordering, seeded with time(NULL): src/gpu_random.cu:18-23,48-51,69-72), bulk-generated
 * into HBM (5 uniforms + 1 normal per walker-step, src/cuda_sim.cu:929-944,976-977).  That
 * stream is closed-source, version-dependent and wall-clock seeded, so it cannot be an
 * oracle.  The contract below replaces it with the reference's OWN in-tree generator
 * (src/xorshift.h), one stream per walker, state in registers:
 *
 *   walker with GLOBAL index w owns XORShift<uint64_t>(walker_seed + w)
 *       -- splitmix64 seeding, src/xorshift.h:158-168; next_rand(), :131-141.
 *   uniform draws are XORShift::rand_pos_double(), [0,1) on a 2^-52 grid (:55-62).
 *   per step k (k = global step index of that walker, starting at 0), in this order,
 *   which is the order in which the reference's channels are consumed:
 *       ux, uy, uz      direction cube point              (cuda_sim.cu:568-570)
 *       if k is even:   u1, u2 -> Box-Muller pair (z0, z1); norm = z0, spare = z1
 *       else:           norm = spare                      (cuda_sim.cu:566)
 *       coin            membrane reflection               (cuda_sim.cu:505)
 *       kick            intracellular kick uniform        (cuda_sim.cu:795)
 *   Box-Muller: v = 1 - u1 in (0,1]; r = sqrt(-2*ln v); (s, c) = sincos(2*pi*u2);
 *       z0 = r*c; z1 = r*s.
 *
 * ln and sincos are written here from + - * / sqrt and explicit fma() only (the classic
 * fdlibm-style argument reductions and minimax coefficient sets, restated), because
 * libdevice and glibc differ in the last bits and a one-ulp difference in a step length
 * can flip a membrane decision.  With these, RNG words, normals, positions, in_cell
 * decisions and octree leaf indices are BIT-IDENTICAL between the CPU oracle and the GPU.
 *
 * Nothing in this header touches oracle/: the oracle includes the product's contract, never
 * the other way round.
 */
#ifndef MRISIM_MATH_H
#define MRISIM_MATH_H

#include <stdint.h>
#include <string.h>

#if defined(__CUDACC__)
#define MRISIM_HD __host__ __device__ __forceinline__
#else
#define MRISIM_HD static inline
#endif

/* ---- bit casts ---------------------------------------------------------------------- */
MRISIM_HD double mrisim_bits_to_double(uint64_t b) {
#if defined(__CUDA_ARCH__)
    return __longlong_as_double((long long)b);
#else
    double d; memcpy(&d, &b, sizeof d); return d;
#endif
}
MRISIM_HD uint64_t mrisim_double_to_bits(double d) {
#if defined(__CUDA_ARCH__)
    return (uint64_t)__double_as_longlong(d);
#else
    uint64_t b; memcpy(&b, &d, sizeof b); return b;
#endif
}
MRISIM_HD double mrisim_fma(double a, double b, double c) {
#if defined(__CUDA_ARCH__)
    return fma(a, b, c);
#else
    return __builtin_fma(a, b, c);
#endif
}
MRISIM_HD double mrisim_sqrt(double a) {
#if defined(__CUDA_ARCH__)
    return sqrt(a);              /* IEEE-754 correctly rounded on sm_100a */
#else
    return __builtin_sqrt(a);
#endif
}

/* ---- xoroshiro128+ (src/xorshift.h:127-141) and its splitmix64 seeding (:158-168) --- */
typedef struct mrisim_rng { uint64_t s0, s1; } mrisim_rng;

MRISIM_HD uint64_t mrisim_rotl64(uint64_t x, int k) { return (x << k) | (x >> (64 - k)); }

MRISIM_HD uint64_t mrisim_rng_next(mrisim_rng* r) {
    const uint64_t s0 = r->s0;
    uint64_t s1 = r->s1;
    const uint64_t result = s0 + s1;
    s1 ^= s0;
    r->s0 = mrisim_rotl64(s0, 55) ^ s1 ^ (s1 << 14);
    r->s1 = mrisim_rotl64(s1, 36);
    return result;
}

MRISIM_HD void mrisim_rng_seed(mrisim_rng* r, uint64_t x) {
    uint64_t z;
    z = (x += UINT64_C(0x9E3779B97F4A7C15));
    z = (z ^ (z >> 30)) * UINT64_C(0xBF58476D1CE4E5B9);
    z = (z ^ (z >> 27)) * UINT64_C(0x94D049BB133111EB);
    r->s0 = z ^ (z >> 31);
    z = (x += UINT64_C(0x9E3779B97F4A7C15));
    z = (z ^ (z >> 30)) * UINT64_C(0xBF58476D1CE4E5B9);
    z = (z ^ (z >> 27)) * UINT64_C(0x94D049BB133111EB);
    r->s1 = z ^ (z >> 31);
}

/* rand_pos_double (src/xorshift.h:55-62): [1,2) by mantissa fill, minus 1 (exact) */
MRISIM_HD double mrisim_u01_from_bits(uint64_t x) {
    return mrisim_bits_to_double((UINT64_C(0x3FF) << 52) | (x >> 12)) - 1.0;
}
MRISIM_HD double mrisim_rng_u01(mrisim_rng* r) { return mrisim_u01_from_bits(mrisim_rng_next(r)); }

/* Polynomial coefficients.  On the device they are read from the constant bank (one operand of
 * the DFMA) instead of being materialised into registers as immediates, two moves per
 * coefficient; the VALUES are the same literals either way. */
#define MRISIM_LG1 6.666666666666735130e-01
#define MRISIM_LG2 3.999999999940941908e-01
#define MRISIM_LG3 2.857142874366239149e-01
#define MRISIM_LG4 2.222219843214978396e-01
#define MRISIM_LG5 1.818357216161805012e-01
#define MRISIM_LG6 1.531383769920937332e-01
#define MRISIM_LG7 1.479819860511658591e-01
#define MRISIM_S1 -1.66666666666666324348e-01
#define MRISIM_S2 8.33333333332248946124e-03
#define MRISIM_S3 -1.98412698298579493134e-04
#define MRISIM_S4 2.75573137070700676789e-06
#define MRISIM_S5 -2.50507602534068634195e-08
#define MRISIM_S6 1.58969099521155010221e-10
#define MRISIM_C1 4.16666666666666019037e-02
#define MRISIM_C2 -1.38888888888741095749e-03
#define MRISIM_C3 2.48015872894767294178e-05
#define MRISIM_C4 -2.75573143513906633035e-07
#define MRISIM_C5 2.08757232129817482790e-09
#define MRISIM_C6 -1.13596475577881948265e-11
#if defined(__CUDACC__)
static __constant__ double mrisim_kc[19] = {
    MRISIM_LG1, MRISIM_LG2, MRISIM_LG3, MRISIM_LG4, MRISIM_LG5, MRISIM_LG6, MRISIM_LG7,
    MRISIM_S1, MRISIM_S2, MRISIM_S3, MRISIM_S4, MRISIM_S5, MRISIM_S6,
    MRISIM_C1, MRISIM_C2, MRISIM_C3, MRISIM_C4, MRISIM_C5, MRISIM_C6};
#endif
#if defined(__CUDA_ARCH__)
#define MRISIM_KC(i, lit) mrisim_kc[i]
#else
#define MRISIM_KC(i, lit) (lit)
#endif

/* ---- ln(v) for v in [2^-52, 1] ------------------------------------------------------ */
/* v = 2^k * m, m in [sqrt(1/2), sqrt(2)); f = m - 1; s = f/(2+f); ln m = f - f^2/2 + s*(f^2/2 + R(s^2)). */
MRISIM_HD double mrisim_log(double v) {
    const double ln2_hi = 6.93147180369123816490e-01;   /* 0x3fe62e42fee00000 */
    const double ln2_lo = 1.90821492927058770002e-10;   /* 0x3dea39ef35793c76 */
    const double Lg1 = MRISIM_KC(0, MRISIM_LG1), Lg2 = MRISIM_KC(1, MRISIM_LG2), Lg3 = MRISIM_KC(2, MRISIM_LG3),
                 Lg4 = MRISIM_KC(3, MRISIM_LG4), Lg5 = MRISIM_KC(4, MRISIM_LG5), Lg6 = MRISIM_KC(5, MRISIM_LG6),
                 Lg7 = MRISIM_KC(6, MRISIM_LG7);
    uint64_t b = mrisim_double_to_bits(v);
    int32_t k = (int32_t)(b >> 52) - 1023;               /* v is normal and positive */
    uint64_t mant = b & UINT64_C(0x000FFFFFFFFFFFFF);
    if (mant >= UINT64_C(0x0006A09E667F3BCD)) {          /* m >= sqrt(2): halve it */
        k += 1;
        b = mant | (UINT64_C(0x3FE) << 52);
    } else {
        b = mant | (UINT64_C(0x3FF) << 52);
    }
    const double f = mrisim_bits_to_double(b) - 1.0;     /* exact */
    const double s = f / (2.0 + f);
    const double z = s * s;
    const double w = z * z;
    const double t1 = w * mrisim_fma(w, mrisim_fma(w, Lg6, Lg4), Lg2);
    const double t2 = z * mrisim_fma(w, mrisim_fma(w, mrisim_fma(w, Lg7, Lg5), Lg3), Lg1);
    const double R = t2 + t1;
    const double hfsq = 0.5 * f * f;
    const double dk = (double)k;
    return dk * ln2_hi - ((hfsq - mrisim_fma(s, hfsq + R, dk * ln2_lo)) - f);
}

/* ---- (sin, cos)(2*pi*u) for u in [0,1) ---------------------------------------------- */
/* r = 4u (exact); n = nearest quarter turn; f = r - n in [-1/2, 1/2] (exact);
 * theta = f*pi/2 in [-pi/4, pi/4]; odd/even minimax kernels; quadrant fix-up. */
MRISIM_HD void mrisim_sincos2pi(double u, double* sn, double* cs) {
    const double pio2 = 1.57079632679489655800e+00;
    const double S1 = MRISIM_KC(7, MRISIM_S1), S2 = MRISIM_KC(8, MRISIM_S2), S3 = MRISIM_KC(9, MRISIM_S3),
                 S4 = MRISIM_KC(10, MRISIM_S4), S5 = MRISIM_KC(11, MRISIM_S5), S6 = MRISIM_KC(12, MRISIM_S6);
    const double C1 = MRISIM_KC(13, MRISIM_C1), C2 = MRISIM_KC(14, MRISIM_C2), C3 = MRISIM_KC(15, MRISIM_C3),
                 C4 = MRISIM_KC(16, MRISIM_C4), C5 = MRISIM_KC(17, MRISIM_C5), C6 = MRISIM_KC(18, MRISIM_C6);
    const double r = u * 4.0;
    const int32_t n = (int32_t)(r + 0.5);
    const double f = r - (double)n;
    const double th = f * pio2;
    const double z = th * th;
    const double ps = mrisim_fma(z, mrisim_fma(z, mrisim_fma(z, mrisim_fma(z, mrisim_fma(z, S6, S5), S4), S3), S2), S1);
    const double s = mrisim_fma(z * th, ps, th);
    const double pc = mrisim_fma(z, mrisim_fma(z, mrisim_fma(z, mrisim_fma(z, mrisim_fma(z, C6, C5), C4), C3), C2), C1);
    const double c = mrisim_fma(z * z, pc, mrisim_fma(z, -0.5, 1.0));
    /* quadrant fix-up, branch-free: n&3 = 0: (s, c)  1: (c, -s)  2: (-s, -c)  3: (-c, s) */
    const double a = (n & 1) ? c : s, b = (n & 1) ? s : c;
    *sn = (n & 2) ? -a : a;
    *cs = ((n + 1) & 2) ? -b : b;
}

/* ---- Box-Muller pair from two [0,1) uniforms ---------------------------------------- */
MRISIM_HD void mrisim_normal_pair_v(double v, double u2, double* z0, double* z1) {   /* v = 1 - u1 */
    const double r = mrisim_sqrt(-2.0 * mrisim_log(v));
    double s, c;
    mrisim_sincos2pi(u2, &s, &c);
    *z0 = r * c;
    *z1 = r * s;
}
MRISIM_HD void mrisim_normal_pair(double u1, double u2, double* z0, double* z1) {
    mrisim_normal_pair_v(1.0 - u1, u2, z0, z1);          /* exact, in (0,1] */
}

/* ---- the intracellular kick of cuda_sim.cu:632, in closed form ---------------------- */
/* With RANDOM_KICK the source expression is
 *   (in_cell) * ( phase_stdev*sqrt(1/(abs(nD-.5)*2)-1)*(((nD-.5)>0) - ((nD-.5))<0) * tau
 *                 + phase_k*1e-3*42.58*2*M_PI*7*tau )
 * whose "sign" factor parses as ((bool)-(nD-.5)) < 0 and is 0 for every nD in [0,1)
 * (SURVEY.md D4).  For nD != 0.5 the first product is finite*0 = +0 and the value is
 * in_cell * K with K the constant below; for nD == 0.5 exactly it is inf*0 = NaN, which
 * poisons the sum whatever in_cell is.  The oracle evaluates the expression verbatim; the
 * kernel uses this closed form; tests/test_kick_expression.py proves them equal. */
MRISIM_HD double mrisim_random_kick_constant(double phase_k, double tau) {
    return phase_k * 1e-3 * 42.58 * 2 * 3.14159265358979323846 * 7 * tau;
}

#endif /* MRISIM_MATH_H */
