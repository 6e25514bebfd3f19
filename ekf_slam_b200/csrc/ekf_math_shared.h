/* ekf_math_shared.h -- the one transcendental on the association / linearisation path, defined once for both sides.
 *
 * z_k(2) = wrapTo360(atan2d(delta_y, delta_x) - theta)   (EKF_SLAM.m:130, Correspondence.m:56)
 *
 * Everything else in the measurement model is +, -, *, /, sqrt: IEEE-exact on the GPU and on the host as long as
 * nothing is contracted into FMAs.  atan2 is not: CUDA's libdevice and glibc agree only to 1-2 ulp, and a 1-ulp
 * difference in a Mahalanobis cost flips an argmin on a near-tie (association mode B, Correspondence.m:74).  So the
 * library and the C oracle (the structured C oracle) evaluate THIS restatement: the classic table-free arctangent
 * (argument reduction at 7/16, 11/16, 19/16, 39/16 with atan(0.5), atan(1), atan(1.5), pi/2 split in high and low
 * parts, odd minimax polynomial of 11 terms below 7/16; < 1 ulp against a 50-digit reference, see
 * test_oracle_cpu.py), written with explicit single operations.  On the device every operation is a
 * round-to-nearest intrinsic (never contracted); the host side must be compiled with -ffp-contract=off.  Same
 * operands, same operations, same order: the two sides return the same bits for every input.
 *
 * The NumPy oracle (dense_literal.py) keeps the platform's atan2, as MATLAB does; it is compared with a
 * tolerance, not bit for bit.
 */
#ifndef EKF_MATH_SHARED_H
#define EKF_MATH_SHARED_H

#include <math.h>

#if defined(__CUDA_ARCH__)
#define EKM_FN __device__ __forceinline__
#define EKM_MUL(a, b) __dmul_rn((a), (b))
#define EKM_ADD(a, b) __dadd_rn((a), (b))
#define EKM_SUB(a, b) __dsub_rn((a), (b))
#define EKM_DIV(a, b) __ddiv_rn((a), (b))
#elif defined(__CUDACC__)
#define EKM_FN __host__ __device__ inline
#define EKM_MUL(a, b) ((a) * (b))
#define EKM_ADD(a, b) ((a) + (b))
#define EKM_SUB(a, b) ((a) - (b))
#define EKM_DIV(a, b) ((a) / (b))
#else
#define EKM_FN static inline
#define EKM_MUL(a, b) ((a) * (b))
#define EKM_ADD(a, b) ((a) + (b))
#define EKM_SUB(a, b) ((a) - (b))
#define EKM_DIV(a, b) ((a) / (b))
#endif

/* arctangent of a non-negative finite or infinite argument */
EKM_FN double ekm_atan_pos(double x) {
    const double hi0 = 4.63647609000806093515e-01, lo0 = 2.26987774529616870924e-17;   /* atan(0.5) */
    const double hi1 = 7.85398163397448278999e-01, lo1 = 3.06161699786838301793e-17;   /* atan(1.0) */
    const double hi2 = 9.82793723247329054082e-01, lo2 = 1.39033110312309984516e-17;   /* atan(1.5) */
    const double hi3 = 1.57079632679489655800e+00, lo3 = 6.12323399573676603587e-17;   /* pi/2      */
    const double a0 = 3.33333333333329318027e-01, a1 = -1.99999999998764832476e-01, a2 = 1.42857142725034663711e-01,
                 a3 = -1.11111104054623557880e-01, a4 = 9.09088713343650656196e-02, a5 = -7.69187620504482999495e-02,
                 a6 = 6.66107313738753120669e-02, a7 = -5.83357013379057348645e-02, a8 = 4.97687799461593236017e-02,
                 a9 = -3.65315727442169155270e-02, a10 = 1.62858201153657823623e-02;
    double hi, lo, t;
    int id;
    if (x >= 1.0e19) return EKM_ADD(hi3, lo3);            /* covers +inf */
    if (x < 0.4375) {                                     /* |x| < 7/16: no reduction */
        if (x < 3.7252902984619140625e-09) return x;      /* 2^-28: atan(x) == x to the last bit */
        id = -1; t = x; hi = 0.0; lo = 0.0;
    } else if (x < 1.1875) {
        if (x < 0.6875) { id = 0; t = EKM_DIV(EKM_SUB(EKM_MUL(2.0, x), 1.0), EKM_ADD(2.0, x)); hi = hi0; lo = lo0; }
        else            { id = 1; t = EKM_DIV(EKM_SUB(x, 1.0), EKM_ADD(x, 1.0)); hi = hi1; lo = lo1; }
    } else {
        if (x < 2.4375) { id = 2; t = EKM_DIV(EKM_SUB(x, 1.5), EKM_ADD(1.0, EKM_MUL(1.5, x))); hi = hi2; lo = lo2; }
        else            { id = 3; t = EKM_DIV(-1.0, x); hi = hi3; lo = lo3; }
    }
    {
        const double z = EKM_MUL(t, t);
        const double w = EKM_MUL(z, z);
        /* odd and even halves of the polynomial, Horner in w */
        double s1 = EKM_MUL(w, a10);
        double s2;
        s1 = EKM_MUL(w, EKM_ADD(a8, s1));
        s1 = EKM_MUL(w, EKM_ADD(a6, s1));
        s1 = EKM_MUL(w, EKM_ADD(a4, s1));
        s1 = EKM_MUL(w, EKM_ADD(a2, s1));
        s1 = EKM_MUL(z, EKM_ADD(a0, s1));
        s2 = EKM_MUL(w, a9);
        s2 = EKM_MUL(w, EKM_ADD(a7, s2));
        s2 = EKM_MUL(w, EKM_ADD(a5, s2));
        s2 = EKM_MUL(w, EKM_ADD(a3, s2));
        s2 = EKM_MUL(w, EKM_ADD(a1, s2));
        if (id < 0) return EKM_SUB(t, EKM_MUL(t, EKM_ADD(s1, s2)));
        return EKM_SUB(hi, EKM_SUB(EKM_SUB(EKM_MUL(t, EKM_ADD(s1, s2)), lo), t));
    }
}

/* atan2 with the IEEE / C99 conventions for zeros, infinities and NaN; result in radians, (-pi, pi]. */
EKM_FN double ekm_atan2(double y, double x) {
    const double pi = 3.1415926535897931160E+00, pi_lo = 1.2246467991473531772E-16;
    const double pio2 = 1.57079632679489655800e+00, pio4 = 7.85398163397448278999e-01;
    double z;
    if (x != x || y != y) return EKM_ADD(x, y);                               /* NaN */
    {
        const int ys = signbit(y) ? 1 : 0, xs = signbit(x) ? 1 : 0;
        if (y == 0.0) return xs ? (ys ? -pi : pi) : y;                        /* atan2(+-0, x) */
        if (x == 0.0) return ys ? -pio2 : pio2;                               /* atan2(y, +-0) */
        if (isinf(x)) {
            if (isinf(y)) return xs ? (ys ? -3.0 * pio4 : 3.0 * pio4) : (ys ? -pio4 : pio4);
            return xs ? (ys ? -pi : pi) : (ys ? -0.0 : 0.0);
        }
        if (isinf(y)) return ys ? -pio2 : pio2;
        z = ekm_atan_pos(fabs(EKM_DIV(y, x)));                                /* overflow -> pi/2, underflow -> tiny: both right */
        if (!xs) return ys ? -z : z;
        z = EKM_SUB(pi, EKM_SUB(z, pi_lo));
        return ys ? -z : z;
    }
}

#endif /* EKF_MATH_SHARED_H */
