/* TEST INFRASTRUCTURE ONLY -- structured ("5-column") CPU oracle for the EKF-SLAM hot path, plain C.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load this library.
 * The product (ekf_slam_b200) never does.
 *
 * What it is: the same filter as oracle/dense_literal.py (which restates the reference with dense n x n matrices,
 * O(n^3), usable to N ~ 2k), written with the sparsity of F and H_k taken out by hand so that one update costs O(n^2)
 * and a 10 000-landmark filter (P = 3.2 GB) can be stepped on the host: F = I + a e1 e3' + b e2 e3' touches two rows and
 * two columns of P (EKF_SLAM.m:47, :62-64); H_k has five non-zero columns c = [1 2 3 j j+1] (EKF_SLAM.m:134-138), so
 * phi_k = H P(c,c) H' + R (:141), K = P(:,c) H' phi_k^-1 (:143) and (I - K H) P = P - K (H P(c,:)) (:145).
 * It is an independent statement of the algebra (different language, different grouping of the products than both the
 * dense oracle and the CUDA kernels), compiled with -ffp-contract=off so every operation rounds once like MATLAB's
 * scalar arithmetic.  Each function cites the reference lines it follows.  atan2 comes from the restatement shared with
 * the CUDA library (ekf_slam_b200/csrc/ekf_math_shared.h, < 1.3 ulp), every other operation is IEEE-exact, so index
 * decisions that hang on a cost comparison (association mode B) can be required to match the GPU bit for bit.
 *
 * parity: pinned through oracle/dense_literal.py (which tests hold to the reference's source text as executed by
 * oracle/mini_matlab.py, tests/golden/refrun_*.npz); MathWorks' built-ins stay unpinned: the reference has no tests /
 * golden vectors and cannot be run under MATLAB/Octave here (neither is in the image).  This file itself is pinned by
 * agreement with dense_literal.py on the golden worlds
 * (tests/test_oracle_cpu.py) and by the 50-digit mpmath restatement there.
 *
 * Layout: x[n] (x, y, theta_deg, l1x, l1y, ...), P column-major with leading dimension ld >= n_cap, s[N] signatures.
 * Indices crossing this interface are 1-based like MATLAB's.
 */
#include <math.h>
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

/* The one transcendental on the association path, shared bit for bit with the CUDA library (see the header's comment):
 * with the platform's atan2 the Mahalanobis costs of mode B (Correspondence.m:74) would agree with the GPU's only to
 * 1-2 ulp and near-ties could pick different landmarks.  -DSO_PLATFORM_ATAN2 builds the oracle with libm's atan2. */
#include "../ekf_slam_b200/csrc/ekf_math_shared.h"

#define PI_ 3.14159265358979323846
#define P_(i, j) P[(size_t)(i) + (size_t)(j) * (size_t)ld]

/* ---- MATLAB runtime / Mapping Toolbox helpers (semantics as in oracle/matlab_compat.py) ---- */
static void reduce90(double x, int* m, double* r) {
    x = fmod(x, 360.0);
    double n = rint(x / 90.0);
    *r = x - 90.0 * n;
    *m = ((int)n) & 3;
}
double so_sind(double x) {                       /* EKF_SLAM.m:42,59,63,84,87-88 */
    int m; double r;
    reduce90(x, &m, &r);
    const double t = r * (PI_ / 180.0);
    return m == 0 ? sin(t) : m == 1 ? cos(t) : m == 2 ? -sin(t) : -cos(t);
}
double so_cosd(double x) {                       /* EKF_SLAM.m:42,58,64,85,87-88 */
    int m; double r;
    reduce90(x, &m, &r);
    const double t = r * (PI_ / 180.0);
    return m == 0 ? cos(t) : m == 1 ? -sin(t) : m == 2 ? -cos(t) : sin(t);
}
#ifdef SO_PLATFORM_ATAN2
double so_atan2d(double y, double x) { return atan2(y, x) * (180.0 / PI_); }   /* EKF_SLAM.m:130 */
#else
double so_atan2d(double y, double x) { return ekm_atan2(y, x) * (180.0 / PI_); }   /* EKF_SLAM.m:130 */
#endif
double so_wrapTo360(double a) {                  /* EKF_SLAM.m:50,130; Correspondence.m:56 */
    double r = fmod(a, 360.0);
    if (r != 0.0 && r < 0.0) r += 360.0;
    if (r == 0.0 && a > 0.0) r = 360.0;
    return r;
}

/* ---- host threads for the one O(n^2) loop (the downdate); plain pthreads, no OpenMP runtime needed ---- */
static int g_threads = 1;
void so_set_threads(int32_t t) { g_threads = t < 1 ? 1 : (t > 64 ? 64 : t); }
typedef struct { double* P; int64_t ld; int n, j0, j1; const double *K0, *K1, *B0, *B1; } dd_job;
static void* dd_run(void* arg) {
    const dd_job* jb = (const dd_job*)arg;
    double* P = jb->P; const int64_t ld = jb->ld;
    for (int j = jb->j0; j < jb->j1; ++j) {                                    /* (I - K H) P = P - K (H P), :145 */
        double* col = &P_(0, j);
        const double b0 = jb->B0[j], b1 = jb->B1[j];
        for (int i = 0; i < jb->n; ++i) col[i] = col[i] - (jb->K0[i] * b0 + jb->K1[i] * b1);
    }
    return 0;
}

/* ---- predict: EKF_SLAM.m:40-51 (== EKF_SLAM_UC.m:42-53), f at :56-65 ---- */
void so_predict(double* x, double* P, int64_t ld, int32_t N, const double* u, double C) {
    const int n = 3 + 2 * N;
    const double th = x[2];
    const double W[3] = {u[0] * so_cosd(th), u[0] * so_sind(th), u[1]};        /* :42 */
    const double a = -1.0 * u[0] * so_sind(th), b = u[0] * so_cosd(th);        /* F(1,3), F(2,3): :63-64, heading before the motion */
    x[0] = x[0] + u[0] * so_cosd(th + u[1]);                                   /* :58 */
    x[1] = x[1] + u[0] * so_sind(th + u[1]);                                   /* :59 */
    x[2] = th + u[1];                                                          /* :60 */
    /* F*P: rows 1,2 += a,b * row 3 */
    for (int j = 0; j < n; ++j) {
        const double p2 = P_(2, j);
        P_(0, j) = P_(0, j) + a * p2;
        P_(1, j) = P_(1, j) + b * p2;
    }
    /* (F*P)*F': columns 1,2 += a,b * column 3 */
    for (int i = 0; i < n; ++i) {
        const double p2 = P_(i, 2);
        P_(i, 0) = P_(i, 0) + p2 * a;
        P_(i, 1) = P_(i, 1) + p2 * b;
    }
    for (int i = 0; i < 3; ++i)                                                /* + Q, Q(1:3,1:3) = C*W*W' (:44) */
        for (int j = 0; j < 3; ++j) P_(i, j) = P_(i, j) + (W[i] * C) * W[j];
    x[2] = so_wrapTo360(x[2]);                                                 /* :50 */
}

/* ---- measurement model about landmark idx1: EKF_SLAM.m:125-138 == Correspondence.m:50-63 ---- */
static void meas_model(const double* x, int al, double zk[2], double H[2][5]) {
    const double d0 = x[al] - x[0], d1 = x[al + 1] - x[1];                     /* :126 */
    const double q = d0 * d0 + d1 * d1;                                        /* :127 */
    const double sq = sqrt(q), iq = 1.0 / q;
    zk[0] = sq;
    zk[1] = so_wrapTo360(so_atan2d(d1, d0) - x[2]);                            /* :130 */
    const double h[2][5] = {{-sq * d0, -sq * d1, 0.0, sq * d0, sq * d1}, {d1, -d0, -q, -d1, d0}};   /* :137-138 */
    for (int r = 0; r < 2; ++r)
        for (int k = 0; k < 5; ++k) H[r][k] = iq * h[r][k];
}

/* phi = (H P(c,c)) H' + R (:141), inverse (:143) */
static void innovation(const double* P, int64_t ld, const int c[5], double H[2][5], const double Rd[2], double Si[2][2]) {
    double HP[2][5], S[2][2];
    for (int r = 0; r < 2; ++r)
        for (int j = 0; j < 5; ++j) {
            double acc = H[r][0] * P_(c[0], c[j]);
            for (int k = 1; k < 5; ++k) acc = acc + H[r][k] * P_(c[k], c[j]);
            HP[r][j] = acc;
        }
    for (int r = 0; r < 2; ++r)
        for (int t = 0; t < 2; ++t) {
            double acc = HP[r][0] * H[t][0];
            for (int k = 1; k < 5; ++k) acc = acc + HP[r][k] * H[t][k];
            S[r][t] = r == t ? acc + Rd[r] : acc;
        }
    const double det = S[0][0] * S[1][1] - S[0][1] * S[1][0];
    const double id = 1.0 / det;
    Si[0][0] = S[1][1] * id; Si[0][1] = -S[0][1] * id;
    Si[1][0] = -S[1][0] * id; Si[1][1] = S[0][0] * id;
}

/* ---- update: EKF_SLAM.m:124-145 == EKF_SLAM_UC.m:125-146.  work: 4 n doubles. Returns 0, or -1 on a bad slot. ---- */
int32_t so_update(double* x, double* P, int64_t ld, int32_t N, const double* z, const double* Rd, int32_t idx1, double* work) {
    if (idx1 < 1 || idx1 > N) return -1;
    const int n = 3 + 2 * N, al = 3 + 2 * (idx1 - 1);
    const int c[5] = {0, 1, 2, al, al + 1};
    double zk[2], H[2][5], Si[2][2];
    meas_model(x, al, zk, H);
    innovation(P, ld, c, H, Rd, Si);
    const double nu[2] = {z[0] - zk[0], z[1] - zk[1]};                         /* :144, bearing innovation NOT wrapped (F10) */
    double* K0 = work; double* K1 = work + n; double* B0 = work + 2 * n; double* B1 = work + 3 * n;
    for (int i = 0; i < n; ++i) {
        double ph0 = 0.0, ph1 = 0.0, b0 = 0.0, b1 = 0.0;
        for (int k = 0; k < 5; ++k) {
            const double pc = P_(i, c[k]), pr = P_(c[k], i);
            ph0 = ph0 + pc * H[0][k]; ph1 = ph1 + pc * H[1][k];               /* P * H' (:143) */
            b0 = b0 + H[0][k] * pr; b1 = b1 + H[1][k] * pr;                   /* H * P  (:145) */
        }
        K0[i] = ph0 * Si[0][0] + ph1 * Si[1][0];
        K1[i] = ph0 * Si[0][1] + ph1 * Si[1][1];
        B0[i] = b0; B1[i] = b1;
    }
    for (int i = 0; i < n; ++i) x[i] = x[i] + (K0[i] * nu[0] + K1[i] * nu[1]); /* :144 */
    {
        const int T = ((size_t)n * n < 1000000u) ? 1 : g_threads;
        dd_job jobs[64];
        pthread_t th[64];
        for (int t = 0; t < T; ++t) {
            dd_job jb = {P, ld, n, (int)((int64_t)n * t / T), (int)((int64_t)n * (t + 1) / T), K0, K1, B0, B1};
            jobs[t] = jb;
            if (t > 0) pthread_create(&th[t], 0, dd_run, &jobs[t]);
        }
        dd_run(&jobs[0]);
        for (int t = 1; t < T; ++t) pthread_join(th[t], 0);
    }
    return 0;
}

/* ---- append: EKF_SLAM.m:67-98 (append.m:1-27).  Capacity is the caller's concern (ld, x, s large enough). ---- */
void so_append(double* x, double* P, int64_t ld, double* s, int32_t N, const double* u, const double* Rd, const double* pos, double sig) {
    const int n = 3 + 2 * N;
    s[N] = sig;                                                                /* :70 */
    x[n] = pos[0]; x[n + 1] = pos[1];                                          /* :79 */
    const double th = x[2];
    const double jxr[2][3] = {{1.0, 0.0, -u[0] * so_sind(th)}, {0.0, 1.0, u[0] * so_cosd(th)}};             /* :84-85 */
    const double jz[2][2] = {{so_cosd(u[1]), -u[0] * so_sind(u[1])}, {so_sind(u[1]), u[0] * so_cosd(u[1])}}; /* :87-88 */
    for (int i = 0; i < n + 2; ++i) { P_(i, n) = 0.0; P_(i, n + 1) = 0.0; P_(n, i) = 0.0; P_(n + 1, i) = 0.0; }
    double JP[2][3];
    for (int r = 0; r < 2; ++r)
        for (int j = 0; j < 3; ++j) JP[r][j] = jxr[r][0] * P_(0, j) + jxr[r][1] * P_(1, j) + jxr[r][2] * P_(2, j);
    for (int r = 0; r < 2; ++r)
        for (int cc = 0; cc < 2; ++cc)                                         /* :91 */
            P_(n + r, n + cc) = (JP[r][0] * jxr[cc][0] + JP[r][1] * jxr[cc][1] + JP[r][2] * jxr[cc][2]) +
                                ((jz[r][0] * Rd[0]) * jz[cc][0] + (jz[r][1] * Rd[1]) * jz[cc][1]);
    for (int i = 0; i < 3; ++i)
        for (int cc = 0; cc < 2; ++cc) {
            const double v = P_(i, 0) * jxr[cc][0] + P_(i, 1) * jxr[cc][1] + P_(i, 2) * jxr[cc][2];       /* :92 */
            P_(i, n + cc) = v;
            P_(n + cc, i) = v;                                                 /* :93 */
        }
    for (int t = 0; t < N; ++t) {                                              /* :94-96 */
        const int al = 3 + 2 * t;
        for (int mm = 0; mm < 2; ++mm) {
            const double p0 = P_(al + mm, 0), p1 = P_(al + mm, 1), p2 = P_(al + mm, 2);
            for (int r = 0; r < 2; ++r) {
                const double v = jxr[r][0] * p0 + jxr[r][1] * p1 + jxr[r][2] * p2;
                P_(n + r, al + mm) = v;
                P_(al + mm, n + r) = v;
            }
        }
    }
}

/* ---- association: Correspondence.m:28-88.  mode 0 = as shipped (:75, signature cost only), 1 = :74 (+ Mahalanobis).
 * Returns the 1-based index; *newLL = 1 when no landmark passes the gate (index = N + 1, :33,:40). ---- */
int32_t so_associate(const double* x, const double* P, int64_t ld, const double* s, int32_t N, const double* z, const double* Rd,
                     double s_cost, double s_thresh, int32_t mode, int32_t* newLL) {
    int index = N + 1;
    double best = INFINITY;
    *newLL = 1;
    for (int kk = 1; kk <= N; ++kk) {                                          /* :49 */
        const double d = z[2] - s[kk - 1];
        double ll = (d * (1.0 / s_cost)) * d;                                  /* :71, :75 */
        if (mode == 1) {
            const int al = 3 + 2 * (kk - 1);
            const int c[5] = {0, 1, 2, al, al + 1};
            double zk[2], H[2][5], Si[2][2];
            meas_model(x, al, zk, H);
            innovation(P, ld, c, H, Rd, Si);
            const double n0 = z[0] - zk[0], n1 = z[1] - zk[1];
            const double t0 = n0 * Si[0][0] + n1 * Si[1][0], t1 = n0 * Si[0][1] + n1 * Si[1][1];           /* :69 */
            ll = (t0 * n0 + t1 * n1) + ll;                                     /* :74 */
        }
        if (ll <= s_thresh && ll < best) { best = ll; index = kk; *newLL = 0; }                            /* :78, :81 */
    }
    return index;
}

/* ---- measure: EKF_SLAM.m:100-151 (variant 0, known correspondence incl. the idx = ii quirk, :123) and
 * EKF_SLAM_UC.m:102-152 (variant 1).  observed: k x 3 column-major [range, bearing_deg, index]; table: loc (ntab x 2
 * column-major) and index (ntab).  actions[ii]: +idx update, -signature append.  Returns 0, -1 bad slot, -2 table miss,
 * -3 capacity. ---- */
int32_t so_measure(double* x, double* P, int64_t ld, double* s, int32_t* N_io, int32_t cap, int32_t variant, int32_t mode,
                   const double* observed, int32_t k, const double* u, const double* loc, const int32_t* tidx, int32_t ntab,
                   const double* Rc, double s_cost, double s_thresh, int32_t* actions, double* work) {
    int N = *N_io;
    for (int ii = 1; ii <= k; ++ii) {                                          /* :107 */
        const double z[3] = {observed[ii - 1], observed[ii - 1 + k], observed[ii - 1 + 2 * k]};
        const double Rd[2] = {z[0] * Rc[0], z[1] * Rc[1]};                     /* :108 */
        int do_append = 0, idx = 0, lookup = 2;
        double sig = 0.0, want = 0.0;
        if (N == 0) { do_append = 1; sig = 1.0; lookup = 1; }                  /* :110-111 (F15: first non-zero index) */
        else if (variant == 0) {
            if (z[2] > (double)N) { do_append = 1; sig = z[2]; want = z[2]; }  /* :118-120 */
            else idx = ii;                                                     /* :123 (F8) */
        } else {
            int32_t newLL = 0;
            idx = so_associate(x, P, ld, s, N, z, Rd, s_cost, s_thresh, mode, &newLL);   /* EKF_SLAM_UC.m:119 */
            if (newLL) { do_append = 1; sig = (double)idx; want = (double)idx; }          /* :121-123 */
        }
        if (do_append) {
            int hit = -1;
            for (int i = 0; i < ntab && hit < 0; ++i)
                if (lookup == 1 ? (tidx[i] != 0) : ((double)tidx[i] == want)) hit = i;
            if (hit < 0) { *N_io = N; return -2; }
            if (N >= cap) { *N_io = N; return -3; }
            const double pos[2] = {loc[hit], loc[hit + ntab]};
            so_append(x, P, ld, s, N, u, Rd, pos, sig);
            if (actions) actions[ii - 1] = -(int32_t)sig;
            N += 1;
        } else {
            if (so_update(x, P, ld, N, z, Rd, idx, work) != 0) { *N_io = N; return -1; }
            if (actions) actions[ii - 1] = idx;
        }
    }
    *N_io = N;
    return 0;
}
