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


// cheirality bits of the four (R, t) candidates for one correspondence in normalised coordinates
// (cv::recoverPose: depth > 0 in both views and < dist)
VS_HD inline unsigned pose_bits(const Proj* cand, double x1, double y1, double x2, double y2, double dist)
{
    const double P0[12] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0};
    unsigned b = 0;
    for (int c = 0; c < 4; ++c) {
        double Q[4];
        dlt_point(P0, cand[c].m, x1, y1, x2, y2, Q);
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
