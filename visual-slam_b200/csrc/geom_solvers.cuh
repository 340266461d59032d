// geom_solvers.cuh -- minimal solvers used by the RANSAC kernels (double precision, one
// hypothesis per thread): null spaces by Gauss-Jordan with full pivoting, the 7-point
// fundamental-matrix solver (cubic in lambda) and the 5-point essential-matrix solver
// (Nister's elimination to a degree-10 polynomial in z).  Also a one-sided Jacobi SVD for
// the 4x4 DLT null vector and the 3x3 decomposition of E.
//
// These replace the solvers cv2.findFundamentalMat(FM_RANSAC) / cv2.findEssentialMat(RANSAC)
// run inside OpenCV (reference call sites matcher.py:192, utils.py:124-125, tracker.py:242).
#pragma once
#include <math.h>

#ifdef __CUDACC__
#define VS_HD __host__ __device__
#define VS_NOINLINE __noinline__ inline
#else
#define VS_HD
#define VS_NOINLINE inline
#endif

// Null space of an R x 9 matrix (R < 9) of full row rank: reduce to RREF with full
// pivoting, then basis vector j has a 1 at free column j and -A[r][free] at the pivot
// columns.  basis: (9 - R) x 9.
template <int R>
VS_HD inline void null_space_9(double A[R][9], double basis[9 - R][9])
{
    int colperm[9];
    for (int j = 0; j < 9; ++j) colperm[j] = j;
    for (int r = 0; r < R; ++r) {
        // full pivot search in the remaining block
        int pr = r, pc = r;
        double best = -1.0;
        for (int i = r; i < R; ++i)
            for (int j = r; j < 9; ++j) {
                double v = fabs(A[i][j]);
                if (v > best) { best = v; pr = i; pc = j; }
            }
        if (pr != r)
            for (int j = 0; j < 9; ++j) { double t = A[r][j]; A[r][j] = A[pr][j]; A[pr][j] = t; }
        if (pc != r) {
            for (int i = 0; i < R; ++i) { double t = A[i][r]; A[i][r] = A[i][pc]; A[i][pc] = t; }
            int t = colperm[r]; colperm[r] = colperm[pc]; colperm[pc] = t;
        }
        double piv = A[r][r];
        if (fabs(piv) < 1e-300) piv = 1e-300;
        double inv = 1.0 / piv;
        for (int j = 0; j < 9; ++j) A[r][j] *= inv;
        for (int i = 0; i < R; ++i) {
            if (i == r) continue;
            double f = A[i][r];
            if (f == 0.0) continue;
            for (int j = 0; j < 9; ++j) A[i][j] -= f * A[r][j];
        }
    }
    for (int k = 0; k < 9 - R; ++k) {
        double v[9];
        for (int j = 0; j < 9; ++j) v[j] = 0.0;
        v[R + k] = 1.0;
        for (int r = 0; r < R; ++r) v[r] = -A[r][R + k];
        // undo the column permutation and normalise
        double nrm = 0.0;
        for (int j = 0; j < 9; ++j) nrm += v[j] * v[j];
        nrm = 1.0 / sqrt(nrm);
        for (int j = 0; j < 9; ++j) basis[k][colperm[j]] = v[j] * nrm;
    }
    // orthonormal basis (modified Gram-Schmidt, two passes): the RREF vectors can be nearly parallel, and the
    // polynomial systems built on the basis (the 5-point elimination above all) lose digits with its conditioning;
    // OpenCV takes the basis from an SVD, i.e. orthonormal.
    for (int pass = 0; pass < 2; ++pass)
        for (int k = 0; k < 9 - R; ++k) {
            for (int p = 0; p < k; ++p) {
                double d = 0.0;
                for (int j = 0; j < 9; ++j) d += basis[k][j] * basis[p][j];
                for (int j = 0; j < 9; ++j) basis[k][j] -= d * basis[p][j];
            }
            double nrm = 0.0;
            for (int j = 0; j < 9; ++j) nrm += basis[k][j] * basis[k][j];
            nrm = nrm > 0.0 ? 1.0 / sqrt(nrm) : 0.0;
            for (int j = 0; j < 9; ++j) basis[k][j] *= nrm;
        }
}

VS_HD inline double det3(const double* m)
{
    return m[0] * (m[4] * m[8] - m[5] * m[7]) - m[1] * (m[3] * m[8] - m[5] * m[6]) +
           m[2] * (m[3] * m[7] - m[4] * m[6]);
}

// Real roots of c[3] x^3 + c[2] x^2 + c[1] x + c[0]; returns the count (0..3).
VS_HD inline int solve_cubic_real(const double c[4], double roots[3])
{
    const double scale = fabs(c[0]) + fabs(c[1]) + fabs(c[2]) + fabs(c[3]);
    if (scale == 0.0) return 0;
    if (fabs(c[3]) < 1e-14 * scale) {
        if (fabs(c[2]) < 1e-14 * scale) {
            if (fabs(c[1]) < 1e-14 * scale) return 0;
            roots[0] = -c[0] / c[1];
            return 1;
        }
        double disc = c[1] * c[1] - 4.0 * c[2] * c[0];
        if (disc < 0.0) return 0;
        double sq = sqrt(disc);
        double q = -0.5 * (c[1] + (c[1] >= 0 ? sq : -sq));
        roots[0] = q / c[2];
        roots[1] = q != 0.0 ? c[0] / q : roots[0];
        return 2;
    }
    const double a = c[2] / c[3], b = c[1] / c[3], d = c[0] / c[3];
    const double Q = (a * a - 3.0 * b) / 9.0, Rr = (2.0 * a * a * a - 9.0 * a * b + 27.0 * d) / 54.0;
    const double Q3 = Q * Q * Q;
    int n;
    if (Rr * Rr < Q3) {
        const double th = acos(Rr / sqrt(Q3)), sq = -2.0 * sqrt(Q);
        roots[0] = sq * cos(th / 3.0) - a / 3.0;
        roots[1] = sq * cos((th + 2.0 * 3.14159265358979323846) / 3.0) - a / 3.0;
        roots[2] = sq * cos((th - 2.0 * 3.14159265358979323846) / 3.0) - a / 3.0;
        n = 3;
    } else {
        double A = -cbrt(fabs(Rr) + sqrt(Rr * Rr - Q3));
        if (Rr < 0) A = -A;
        const double B = A != 0.0 ? Q / A : 0.0;
        roots[0] = (A + B) - a / 3.0;
        n = 1;
    }
    // one Newton polish per root
    for (int i = 0; i < n; ++i) {
        double x = roots[i];
        for (int it = 0; it < 2; ++it) {
            double f = ((c[3] * x + c[2]) * x + c[1]) * x + c[0];
            double fp = (3.0 * c[3] * x + 2.0 * c[2]) * x + c[1];
            if (fp != 0.0) x -= f / fp;
        }
        roots[i] = x;
    }
    return n;
}

// 7-point solver.  x1, x2: 7 correspondences (any coordinate system).  Writes up to 3
// matrices F (row-major, x2^T F x1 = 0) and returns how many.
VS_HD inline int solve_7pt(const double x1[7][2], const double x2[7][2], double F[3][9])
{
    double A[7][9], nb[2][9];
    for (int i = 0; i < 7; ++i) {
        const double u = x1[i][0], v = x1[i][1], up = x2[i][0], vp = x2[i][1];
        A[i][0] = up * u; A[i][1] = up * v; A[i][2] = up;
        A[i][3] = vp * u; A[i][4] = vp * v; A[i][5] = vp;
        A[i][6] = u; A[i][7] = v; A[i][8] = 1.0;
    }
    null_space_9<7>(A, nb);
    const double* f1 = nb[0];
    const double* f2 = nb[1];
    // det(l f1 + (1 - l) f2) is a cubic in l: interpolate it from four evaluations
    double g[4], M[9];
    const double ls[4] = {0.0, 1.0, -1.0, 2.0};
    for (int k = 0; k < 4; ++k) {
        for (int j = 0; j < 9; ++j) M[j] = ls[k] * f1[j] + (1.0 - ls[k]) * f2[j];
        g[k] = det3(M);
    }
    // p(l) = c0 + c1 l + c2 l^2 + c3 l^3 with p(0)=g0, p(1)=g1, p(-1)=g2, p(2)=g3
    double c[4];
    c[0] = g[0];
    c[2] = 0.5 * (g[1] + g[2]) - g[0];
    const double s = 0.5 * (g[1] - g[2]);              // c1 + c3
    c[3] = (g[3] - g[0] - 4.0 * c[2] - 2.0 * s) / 6.0;
    c[1] = s - c[3];
    double r[3];
    const int n = solve_cubic_real(c, r);
    for (int k = 0; k < n; ++k)
        for (int j = 0; j < 9; ++j) F[k][j] = r[k] * f1[j] + (1.0 - r[k]) * f2[j];
    return n;
}
