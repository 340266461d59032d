// geom_pose.cuh -- pieces shared by recoverPose / triangulatePoints (geom_pose.cu) and the batched tracking
// step (geom_track.cu): the per-point 4x4 DLT, cv::decomposeEssentialMat and OpenCV's cheirality rules.
#pragma once
#include <math.h>

#include "geom_5pt.cuh"
#include "geom_solvers.cuh"

namespace {

struct Proj { double m[12]; };

VS_HD inline void dlt_point(const double* P1, const double* P2, double x1, double y1, double x2, double y2,
                                 double X[4])
{
    double A[4][4], V[4][4];
    for (int c = 0; c < 4; ++c) {
        A[0][c] = x1 * P1[8 + c] - P1[c];
        A[1][c] = y1 * P1[8 + c] - P1[4 + c];
        A[2][c] = x2 * P2[8 + c] - P2[c];
        A[3][c] = y2 * P2[8 + c] - P2[4 + c];
    }
    jacobi_svd<4>(A, V);
    int best = 0;
    double bn = 1e300;
    for (int c = 0; c < 4; ++c) {
        double s = 0.0;
        for (int r = 0; r < 4; ++r) s += A[r][c] * A[r][c];
        if (s < bn) { bn = s; best = c; }
    }
    for (int r = 0; r < 4; ++r) X[r] = V[r][best];
}

VS_HD inline void mat3_mul(const double* A, const double* B, double* C)
{
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) {
            double s = 0;
            for (int k = 0; k < 3; ++k) s += A[3 * i + k] * B[3 * k + j];
            C[3 * i + j] = s;
        }
}

// cv::decomposeEssentialMat: E = U diag(s) Vt, det(U), det(Vt) forced positive,
// R1 = U W Vt, R2 = U W^T Vt, t = U[:, 2]
VS_HD inline void decompose_essential(const double* E, double* R1, double* R2, double* t)
{
    double A[3][3], V[3][3];
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) A[i][j] = E[3 * i + j];
    jacobi_svd<3>(A, V);
    double s[3];
    int order[3] = {0, 1, 2};
    for (int c = 0; c < 3; ++c) s[c] = sqrt(A[0][c] * A[0][c] + A[1][c] * A[1][c] + A[2][c] * A[2][c]);
    for (int a = 0; a < 3; ++a)
        for (int b = a + 1; b < 3; ++b)
            if (s[order[b]] > s[order[a]]) { int tt = order[a]; order[a] = order[b]; order[b] = tt; }
    double U[9], Vm[9];
    for (int c = 0; c < 3; ++c) {
        const int k = order[c];
        for (int r = 0; r < 3; ++r) {
            U[3 * r + c] = s[k] > 1e-300 ? A[r][k] / s[k] : 0.0;
            Vm[3 * r + c] = V[r][k];
        }
    }
    // third left singular vector: orthogonal complement (singular value ~ 0)
    U[2] = U[3] * U[7] - U[6] * U[4];
    U[5] = U[6] * U[1] - U[0] * U[7];
    U[8] = U[0] * U[4] - U[3] * U[1];
    double nn = sqrt(U[2] * U[2] + U[5] * U[5] + U[8] * U[8]);
    if (nn > 0) { U[2] /= nn; U[5] /= nn; U[8] /= nn; }
    if (det3(U) < 0) for (int e = 0; e < 9; ++e) U[e] = -U[e];
    double Vt[9];
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) Vt[3 * i + j] = Vm[3 * j + i];
    if (det3(Vt) < 0) for (int e = 0; e < 9; ++e) Vt[e] = -Vt[e];
    const double W[9] = {0, 1, 0, -1, 0, 0, 0, 0, 1}, Wt[9] = {0, -1, 0, 1, 0, 0, 0, 0, 1};
    double T[9];
    mat3_mul(U, W, T); mat3_mul(T, Vt, R1);
    mat3_mul(U, Wt, T); mat3_mul(T, Vt, R2);
    t[0] = U[2]; t[1] = U[5]; t[2] = U[8];
}

VS_HD inline Proj make_proj(const double* R, const double* t, double sign)
{
    Proj P;
    for (int r = 0; r < 3; ++r) {
        for (int c = 0; c < 3; ++c) P.m[4 * r + c] = R[3 * r + c];
        P.m[4 * r + 3] = sign * t[r];
    }
    return P;
}


// The same null vector for the cheirality test of recoverPose, where only signs and a coarse depth bound are
// read from it: smallest eigenvector of A^T A by inverse iteration on a Cholesky factor (a few dozen
// multiply-adds per step and 2-4 steps for an inlier correspondence, against ~7 000 instructions for the
// Jacobi SVD above).  Falls back to the SVD when the iteration has not settled (degenerate geometry).
VS_HD inline void dlt_point_fast(const double* P1, const double* P2, double x1, double y1, double x2, double y2,
                                 double X[4])
{
    double A[4][4];
    for (int c = 0; c < 4; ++c) {
        A[0][c] = x1 * P1[8 + c] - P1[c];
        A[1][c] = y1 * P1[8 + c] - P1[4 + c];
        A[2][c] = x2 * P2[8 + c] - P2[c];
        A[3][c] = y2 * P2[8 + c] - P2[4 + c];
    }
    double M[4][4];
    for (int i = 0; i < 4; ++i)
        for (int j = 0; j <= i; ++j)
            M[i][j] = A[0][i] * A[0][j] + A[1][i] * A[1][j] + A[2][i] * A[2][j] + A[3][i] * A[3][j];
    const double tr = M[0][0] + M[1][1] + M[2][2] + M[3][3];
    const double shift = 1e-14 * tr + 1e-300;
    // Cholesky of M + shift I (lower triangle in place, reciprocal diagonal kept)
    double L[4][4], inv[4];
    bool ok = true;
    for (int j = 0; j < 4; ++j) {
        double d = M[j][j] + shift;
        for (int k = 0; k < j; ++k) d -= L[j][k] * L[j][k];
        if (!(d > 0.0)) { ok = false; d = shift; }
        const double r = 1.0 / sqrt(d);
        inv[j] = r;
        L[j][j] = d * r;
        for (int i = j + 1; i < 4; ++i) {
            double v = M[i][j];
            for (int k = 0; k < j; ++k) v -= L[i][k] * L[j][k];
            L[i][j] = v * r;
        }
    }
    double x[4] = {0.5, 0.5, 0.5, 0.5};
    bool conv = false;
    for (int it = 0; it < 48 && ok; ++it) {
        double y[4];
        y[0] = x[0] * inv[0];                                           // L y' = x
        y[1] = (x[1] - L[1][0] * y[0]) * inv[1];
        y[2] = (x[2] - L[2][0] * y[0] - L[2][1] * y[1]) * inv[2];
        y[3] = (x[3] - L[3][0] * y[0] - L[3][1] * y[1] - L[3][2] * y[2]) * inv[3];
        y[3] = y[3] * inv[3];                                           // L^T y = y'
        y[2] = (y[2] - L[3][2] * y[3]) * inv[2];
        y[1] = (y[1] - L[2][1] * y[2] - L[3][1] * y[3]) * inv[1];
        y[0] = (y[0] - L[1][0] * y[1] - L[2][0] * y[2] - L[3][0] * y[3]) * inv[0];
        const double nn = y[0] * y[0] + y[1] * y[1] + y[2] * y[2] + y[3] * y[3];
        if (!(nn > 0.0) || !isfinite(nn)) { ok = false; break; }
        const double rn = 1.0 / sqrt(nn);
        double dot = 0.0;
        for (int k = 0; k < 4; ++k) { y[k] *= rn; dot += y[k] * x[k]; }
        for (int k = 0; k < 4; ++k) x[k] = y[k];
        if (1.0 - fabs(dot) < 1e-15) { conv = true; break; }
    }
    if (ok && conv) { for (int k = 0; k < 4; ++k) X[k] = x[k]; return; }
    dlt_point(P1, P2, x1, y1, x2, y2, X);
}

// cheirality bits of the four (R, t) candidates for one correspondence in normalised coordinates
// (cv::recoverPose: depth > 0 in both views and < dist)
VS_HD inline unsigned pose_bits(const Proj* cand, double x1, double y1, double x2, double y2, double dist)
{
    const double P0[12] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0};
    unsigned b = 0;
    for (int c = 0; c < 4; ++c) {
        double Q[4];
        dlt_point_fast(P0, cand[c].m, x1, y1, x2, y2, Q);
        bool ok = Q[2] * Q[3] > 0;
        const double x = Q[0] / Q[3], y = Q[1] / Q[3], z = Q[2] / Q[3];
        ok = ok && (z < dist);
        const double* P = cand[c].m;
        const double z2 = P[8] * x + P[9] * y + P[10] * z + P[11];
        ok = ok && (z2 > 0) && (z2 < dist);
        b |= (unsigned)ok << c;
    }
    return b;
}

// OpenCV's winner among the four candidates' good counts (first maximum in the order R1|t, R2|t, R1|-t, R2|-t)
VS_HD inline int pose_winner(const int good[4])
{
    if (good[0] >= good[1] && good[0] >= good[2] && good[0] >= good[3]) return 0;
    if (good[1] >= good[0] && good[1] >= good[2] && good[1] >= good[3]) return 1;
    if (good[2] >= good[0] && good[2] >= good[1] && good[2] >= good[3]) return 2;
    return 3;
}

}  // namespace
