// geom_5pt.cuh -- 5-point essential-matrix solver (Nister 2004) and small Jacobi SVDs.
#pragma once
#include "geom_solvers.cuh"
#include "geom_tables.inc"

// ---- polynomial helpers in (x, y, z, 1): linear = 4, quadratic = 10, cubic = 20 coeffs ----
VS_HD inline void poly_mul_ll(const double* a, const double* b, double* q, double s)   // q += s*a*b
{
    const signed char QI[4][4] = VS_QIDX_INIT;
    for (int i = 0; i < 4; ++i)
        for (int j = 0; j < 4; ++j) q[QI[i][j]] += s * a[i] * b[j];
}
VS_HD inline void poly_mul_ql(const double* q, const double* l, double* c, double s)   // c += s*q*l
{
    const signed char CQ[10][4] = VS_CQ_INIT;
    for (int i = 0; i < 10; ++i)
        for (int k = 0; k < 4; ++k) c[CQ[i][k]] += s * q[i] * l[k];
}

// real polynomial helpers (dense, ascending powers)
VS_HD inline void pmul(const double* a, int da, const double* b, int db, double* out)  // out = a*b
{
    for (int i = 0; i <= da + db; ++i) out[i] = 0.0;
    for (int i = 0; i <= da; ++i)
        for (int j = 0; j <= db; ++j) out[i + j] += a[i] * b[j];
}

VS_HD inline double poly_eval(const double* q, int n, double x)
{
    double f = q[n];
    for (int i = n - 1; i >= 0; --i) f = f * x + q[i];
    return f;
}

// The single root of q (degree n) in a bracket [xl, xh] over which q is monotone and changes sign; `up`: q
// increases through the root.  Newton from the midpoint, bisection whenever a step leaves the bracket.
// fl, fh: q at the bracket ends (known to every caller): the first iterate is the secant point, not the midpoint.
VS_HD inline double poly_root_in(const double* q, int n, double xl, double xh, bool up, double tol, double fl, double fh)
{
    double x = (xl * fh - xh * fl) / (fh - fl);
    if (!(x > xl && x < xh)) x = 0.5 * (xl + xh);
    for (int it = 0; it < 80; ++it) {
        double f = q[n], fp = 0.0;
        for (int i = n - 1; i >= 0; --i) { fp = fp * x + f; f = f * x + q[i]; }
        if (f == 0.0) break;
        if ((f < 0.0) == up) xl = x; else xh = x;
        double xn = x - f / fp;
        if (!(xn > xl && xn < xh)) xn = 0.5 * (xl + xh);
        const double dx = fabs(xn - x);
        x = xn;
        if (dx <= tol * (1.0 + fabs(x)) || xh - xl <= tol * (1.0 + fabs(x))) break;
    }
    return x;
}

// Real roots of a degree-10 polynomial (ascending coefficients c[0..10]).  Between two consecutive real roots of
// p' the polynomial p is monotone, so the real roots of p are bracketed by the real roots of p'; the same holds
// for p' and p'', ... down to the linear 9th derivative.  Each bracket with a sign change holds exactly one
// root; the two unbounded outer intervals are searched by doubling steps away from the outermost critical point
// (the sign at infinity says whether a root is there).  Only real arithmetic, no start values, and the cost
// follows the number of real roots (typically 2-6) -- the complex Durand-Kerner iteration this replaces spent
// ~1600 instructions per sweep on all ten roots and often ran into its sweep limit.
VS_HD inline int real_roots_deg10(const double* c, double* roots, bool* deflated)
{
    const int NMAX = 10;
    double mx = 0.0;
    for (int i = 0; i <= NMAX; ++i) mx = fmax(mx, fabs(c[i]));
    *deflated = false;
    if (!(mx > 0.0) || !isfinite(mx)) return 0;
    // effective degree: leading coefficients that vanish against the largest one put roots at infinity (solutions
    // with no component along the fourth basis vector -- pure sideways motion does that); the chain of derivatives
    // below cannot bracket through critical points at 1e12, so those roots are dropped here and the caller is told
    int N = NMAX;
    while (N > 1 && fabs(c[N]) < 1e-11 * mx) --N;
    *deflated = N < NMAX;
    if (fabs(c[N]) < 1e-11 * mx) return 0;
    // balance: x = s w with s = |c0 / cN|^(1/N) (the geometric mean of the root moduli), so that the scaled
    // coefficients are of one order of magnitude whatever the unit of the roots is
    double s = 1.0;
    if (fabs(c[0]) > 0.0) s = exp((log(fabs(c[0])) - log(fabs(c[N]))) / N);
    if (!(s > 1e-6 && s < 1e6)) s = 1.0;
    double a[NMAX + 1], sp = 1.0;
    for (int i = 0; i <= NMAX; ++i) { a[i] = i <= N ? c[i] * sp : 0.0; sp *= s; }
    { const double lead = a[N]; for (int i = 0; i <= N; ++i) a[i] /= lead; }
    double B = 0.0;
    for (int i = 0; i < N; ++i) B = fmax(B, fabs(a[i]));
    B = fmin(1.0 + B, 1e8);                         // Cauchy bound; an essential matrix from |z| > 1e8 is degenerate anyway
    double crit[NMAX], next[NMAX];
    int nc = 0;                                     // real roots of the derivative one level up, ascending
    for (int m = N - 1; m >= 0; --m) {              // level m: q = m-th derivative (up to a positive factor), degree n
        const int n = N - m;
        double q[NMAX + 1];
        for (int i = 0; i <= n; ++i) {
            double f = a[i + m];
            for (int k = 0; k < m; ++k) f *= (double)(i + m - k);   // (i+m)! / i!
            q[i] = f;
        }
        const double tol = m == 0 ? 4e-16 : 1e-11;
        int nr = 0;
        // (-inf, first anchor): q(-inf) has the sign (-1)^n (q's leading coefficient is positive)
        const double a0 = nc > 0 ? crit[0] : 0.0;
        double fprev = poly_eval(q, n, a0);
        {
            const bool neg_inf_negative = (n & 1) != 0;
            if (fprev != 0.0 && ((fprev < 0.0) != neg_inf_negative)) {
                double h = 0.25 * (1.0 + fabs(a0)), xin = a0;
                double fin = fprev;
                for (int it = 0; it < 64; ++it, h *= 2.0) {
                    const double xo = a0 - h, fo = poly_eval(q, n, xo);
                    if ((fo < 0.0) == neg_inf_negative) { next[nr++] = poly_root_in(q, n, xo, xin, neg_inf_negative, tol, fo, fin); break; }
                    xin = xo; fin = fo;
                    if (h > 2.0 * B) break;
                }
            } else if (fprev == 0.0 && nc == 0) {
                next[nr++] = a0;
            }
        }
        // between consecutive critical points
        for (int iv = 0; iv < nc; ++iv) {
            const double lo = crit[iv], hi = iv + 1 < nc ? crit[iv + 1] : 0.0;
            if (fprev == 0.0) next[nr++] = lo;       // a critical point that is itself a root
            if (iv + 1 == nc) break;
            const double fhi = poly_eval(q, n, hi);
            if (hi > lo && fprev != 0.0 && fhi != 0.0 && ((fprev < 0.0) != (fhi < 0.0)))
                next[nr++] = poly_root_in(q, n, lo, hi, fprev < 0.0, tol, fprev, fhi);
            fprev = fhi;
        }
        // (last anchor, +inf): q(+inf) > 0
        {
            const double a1 = nc > 0 ? crit[nc - 1] : 0.0;
            if (fprev < 0.0) {
                double h = 0.25 * (1.0 + fabs(a1)), xin = a1;
                double fin = fprev;
                for (int it = 0; it < 64; ++it, h *= 2.0) {
                    const double xo = a1 + h, fo = poly_eval(q, n, xo);
                    if (fo > 0.0) { next[nr++] = poly_root_in(q, n, xin, xo, true, tol, fin, fo); break; }
                    xin = xo; fin = fo;
                    if (h > 2.0 * B) break;
                }
            }
        }
        nc = nr;
        for (int i = 0; i < nr; ++i) crit[i] = next[i];
    }
    // back to the caller's unit at insertion
    int nout = 0;
    for (int k = 0; k < nc; ++k) {
        const double x = crit[k] * s;
        bool dup = false;
        for (int j = 0; j < nout; ++j) dup |= fabs(roots[j] - x) < 1e-9 * (s + fabs(x));
        if (!dup && isfinite(x)) roots[nout++] = x;
    }
    return nout;
}

// Residuals of the essential-matrix constraints at E (row-major): f[0..8] = 2 E E^T E - tr(E E^T) E, f[9] = det E.
VS_HD inline void e_constraints(const double* E, double* f, double* EEt, double* tr_out)
{
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j)
            EEt[3 * i + j] = E[3 * i] * E[3 * j] + E[3 * i + 1] * E[3 * j + 1] + E[3 * i + 2] * E[3 * j + 2];
    const double tr = EEt[0] + EEt[4] + EEt[8];
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j)
            f[3 * i + j] = 2.0 * (EEt[3 * i] * E[j] + EEt[3 * i + 1] * E[3 + j] + EEt[3 * i + 2] * E[6 + j]) - tr * E[3 * i + j];
    f[9] = det3(E);
    *tr_out = tr;
}

// Gauss-Newton polish of one solution of the five-point problem in null-space coordinates c (E = sum_k c[k] L[.][k],
// |c| = 1): the ten cubic constraints are evaluated on E itself, so the result does not inherit the conditioning of
// the elimination that produced the start value.  OpenCV does not polish, but on well-conditioned samples its roots
// are accurate to ~1e-12 and the polished solutions reproduce its inlier counts (tools/ransac_agreement.py);
// un-polished roots of an ill-conditioned elimination were off by up to 0.2 in E.  Keeps the start value unless
// the residual shrinks.
VS_HD inline double polish_residual(const double L[9][4], const double* c, double* E, double* f, double* EEt, double* tr)
{
    for (int e = 0; e < 9; ++e) E[e] = c[0] * L[e][0] + c[1] * L[e][1] + c[2] * L[e][2] + c[3] * L[e][3];
    e_constraints(E, f, EEt, tr);
    f[10] = c[0] * c[0] + c[1] * c[1] + c[2] * c[2] + c[3] * c[3] - 1.0;
    double r2 = 0.0;
    for (int i = 0; i < 11; ++i) r2 += f[i] * f[i];
    return r2;
}

// Returns the final squared residual; *start_r2 receives the residual of the start value.
// NOT inlined on the device, on purpose: with this function inlined into the loop over the roots, cicc 12.9 (-O1 and
// above; -Xcicc -O0, -G and g++ are fine, ASan/UBSan find nothing) produced code in which entries 1.. of the caller's
// root list read back as 0 after the first polish -- its local arrays were overlaid on the live list.  Repro and
// device-vs-host check of the solver: tools/ubench/solve5_device.cu.
VS_HD VS_NOINLINE double polish_5pt(const double L[9][4], double c[4], double* start_r2)
{
    double E[9], f[11], EEt[9], tr;
    double r2 = polish_residual(L, c, E, f, EEt, &tr);
    *start_r2 = r2;
    for (int it = 0; it < 12 && r2 > 1e-26 && isfinite(r2); ++it) {
#ifdef VS_POLISH_TRACE
        VS_POLISH_TRACE(it, r2, c);
#endif
        // cofactors of E (d det = sum cof_ij D_ij)
        const double cof[9] = {E[4] * E[8] - E[5] * E[7], E[5] * E[6] - E[3] * E[8], E[3] * E[7] - E[4] * E[6],
                               E[2] * E[7] - E[1] * E[8], E[0] * E[8] - E[2] * E[6], E[1] * E[6] - E[0] * E[7],
                               E[1] * E[5] - E[2] * E[4], E[2] * E[3] - E[0] * E[5], E[0] * E[4] - E[1] * E[3]};
        double J[11][4];
        for (int k = 0; k < 4; ++k) {
            double D[9], G[9];                                    // direction and D E^T
            for (int e = 0; e < 9; ++e) D[e] = L[e][k];
            for (int i = 0; i < 3; ++i)
                for (int j = 0; j < 3; ++j)
                    G[3 * i + j] = D[3 * i] * E[3 * j] + D[3 * i + 1] * E[3 * j + 1] + D[3 * i + 2] * E[3 * j + 2];
            const double dtr = 2.0 * (G[0] + G[4] + G[8]);
            for (int i = 0; i < 3; ++i)
                for (int j = 0; j < 3; ++j) {
                    double v = 0.0;
                    for (int m = 0; m < 3; ++m)
                        v += (G[3 * i + m] + G[3 * m + i]) * E[3 * m + j] + EEt[3 * i + m] * D[3 * m + j];
                    J[3 * i + j][k] = 2.0 * v - dtr * E[3 * i + j] - tr * D[3 * i + j];
                }
            double dd = 0.0;
            for (int e = 0; e < 9; ++e) dd += cof[e] * D[e];
            J[9][k] = dd;
            J[10][k] = 2.0 * c[k];
        }
        // normal equations (4 x 4): Gaussian elimination with pivoting
        double N[4][5];
        for (int a = 0; a < 4; ++a) {
            for (int b = 0; b < 4; ++b) {
                double v = 0.0;
                for (int i = 0; i < 11; ++i) v += J[i][a] * J[i][b];
                N[a][b] = v;
            }
            double g = 0.0;
            for (int i = 0; i < 11; ++i) g += J[i][a] * f[i];
            N[a][4] = -g;
        }
        bool ok = true;
        for (int a = 0; a < 4 && ok; ++a) {
            int pr = a;
            for (int r = a + 1; r < 4; ++r)
                if (fabs(N[r][a]) > fabs(N[pr][a])) pr = r;
            if (!(fabs(N[pr][a]) > 1e-300)) { ok = false; break; }
            if (pr != a)
                for (int b = 0; b < 5; ++b) { const double t = N[a][b]; N[a][b] = N[pr][b]; N[pr][b] = t; }
            const double inv = 1.0 / N[a][a];
            for (int r = 0; r < 4; ++r) {
                if (r == a) continue;
                const double m = N[r][a] * inv;
                for (int b = a; b < 5; ++b) N[r][b] -= m * N[a][b];
            }
        }
        if (!ok) break;
        double d[4], step2 = 0.0;
        for (int k = 0; k < 4; ++k) { d[k] = N[k][4] / N[k][k]; step2 += d[k] * d[k]; }
        if (!isfinite(step2)) break;
        // backtracking: halve the step until the residual does not grow
        double cn[4], rn = 0.0, lam = 1.0;
        bool moved = false;
        for (int h = 0; h < 8; ++h, lam *= 0.5) {
            for (int k = 0; k < 4; ++k) cn[k] = c[k] + lam * d[k];
            rn = polish_residual(L, cn, E, f, EEt, &tr);
            if (rn < r2) { moved = true; break; }
        }
        if (!moved) { polish_residual(L, c, E, f, EEt, &tr); break; }
        for (int k = 0; k < 4; ++k) c[k] = cn[k];
        r2 = rn;
        if (step2 * lam * lam < 1e-28) break;
    }
    return r2;
}

// One pass of Nister's elimination in the null-space basis nb (orthonormal rows): appends the polished real solutions
// that are not yet in E[0..n) and returns the new count.  *trouble is set when the pass cannot be trusted to be
// complete: a polish did not converge, two roots polished into one solution (so another one was probably missed),
// leading coefficients vanished (roots at infinity of this chart), or the number of solutions found is odd.  (A start
// value far from its solution is NOT a trigger: on a panning camera the elimination loses digits on every second
// sample, the polish recovers them, and nothing is missed.)
// The three stages of a pass, separately callable so that a kernel can give the (independent) polishes of a round's
// roots to different threads (geom_track.cu); solve_5pt_in_basis below is their serial composition.  Stages 1 and 2
// are NOT inlined on the device: cicc 12.9 overlays the local arrays of inlined callees on live arrays of the caller
// at -O1 and above (see polish_5pt) -- with these two inlined, the first solution of every sample came back wrong.
// tools/ubench/solve5_device.cu compares the device build with the host build of this very source.
//
// Stage 1: elimination in the basis nb -> L (entry e = 3i + j of E as a linear form in the basis coordinates) and one
// unit start vector per real root of the degree-10 polynomial; returns the number of starts.
VS_HD VS_NOINLINE int five_pt_starts(const double nb[4][9], double L[9][4], double starts[10][4], bool* deflated)
{
    for (int e = 0; e < 9; ++e)
        for (int k = 0; k < 4; ++k) L[e][k] = nb[k][e];

    double M[10][20];
    for (int r = 0; r < 10; ++r)
        for (int c = 0; c < 20; ++c) M[r][c] = 0.0;
    double q[10];
    // row 0: det(E)
    {
        const int idx[6][3] = {{0, 4, 8}, {1, 5, 6}, {2, 3, 7}, {2, 4, 6}, {1, 3, 8}, {0, 5, 7}};
        for (int t = 0; t < 6; ++t) {
            for (int i = 0; i < 10; ++i) q[i] = 0.0;
            poly_mul_ll(L[idx[t][0]], L[idx[t][1]], q, 1.0);
            poly_mul_ql(q, L[idx[t][2]], M[0], t < 3 ? 1.0 : -1.0);
        }
    }
    // EEt (symmetric 3x3 of quadratics) and Lambda = EEt - 0.5 trace I
    double EEt[3][3][10];
    for (int i = 0; i < 3; ++i)
        for (int j = i; j < 3; ++j) {
            for (int t = 0; t < 10; ++t) EEt[i][j][t] = 0.0;
            for (int k = 0; k < 3; ++k) poly_mul_ll(L[3 * i + k], L[3 * j + k], EEt[i][j], 1.0);
            if (j != i)
                for (int t = 0; t < 10; ++t) EEt[j][i][t] = EEt[i][j][t];
        }
    for (int t = 0; t < 10; ++t) {
        double tr = 0.5 * (EEt[0][0][t] + EEt[1][1][t] + EEt[2][2][t]);
        EEt[0][0][t] -= tr; EEt[1][1][t] -= tr; EEt[2][2][t] -= tr;
    }
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j)
            for (int k = 0; k < 3; ++k) poly_mul_ql(EEt[i][k], L[3 * k + j], M[1 + 3 * i + j], 1.0);

    // Gauss-Jordan on the first 10 columns (partial pivoting)
    for (int c = 0; c < 10; ++c) {
        int pr = c;
        double best = fabs(M[c][c]);
        for (int r = c + 1; r < 10; ++r)
            if (fabs(M[r][c]) > best) { best = fabs(M[r][c]); pr = r; }
        if (best < 1e-300) return 0;
        if (pr != c)
            for (int j = 0; j < 20; ++j) { double t = M[c][j]; M[c][j] = M[pr][j]; M[pr][j] = t; }
        double inv = 1.0 / M[c][c];
        for (int j = c; j < 20; ++j) M[c][j] *= inv;
        for (int r = 0; r < 10; ++r) {
            if (r == c) continue;
            double f = M[r][c];
            if (f == 0.0) continue;
            for (int j = c; j < 20; ++j) M[r][j] -= f * M[c][j];
        }
    }
    // B(z): rows k = e - z f, l = g - z h, m = i - z j; columns (x, y, 1) as polynomials in z
    double B[3][3][5];
    for (int r = 0; r < 3; ++r) {
        const double* e = M[4 + 2 * r];
        const double* f = M[5 + 2 * r];
        // x column (degree 3): from xz^2 (10), xz (11), x (12)
        B[r][0][0] = e[12]; B[r][0][1] = e[11] - f[12]; B[r][0][2] = e[10] - f[11]; B[r][0][3] = -f[10]; B[r][0][4] = 0.0;
        // y column (degree 3): yz^2 (13), yz (14), y (15)
        B[r][1][0] = e[15]; B[r][1][1] = e[14] - f[15]; B[r][1][2] = e[13] - f[14]; B[r][1][3] = -f[13]; B[r][1][4] = 0.0;
        // constant column (degree 4): z^3 (16), z^2 (17), z (18), 1 (19)
        B[r][2][0] = e[19]; B[r][2][1] = e[18] - f[19]; B[r][2][2] = e[17] - f[18]; B[r][2][3] = e[16] - f[17];
        B[r][2][4] = -f[16];
    }
    // det B(z): degree 10
    double det[11];
    for (int i = 0; i <= 10; ++i) det[i] = 0.0;
    {
        double t1[8], t2[8], minor[8], prod[12];
        // k_x (l_y m_1 - l_1 m_y)
        pmul(B[1][1], 3, B[2][2], 4, t1); pmul(B[1][2], 4, B[2][1], 3, t2);
        for (int i = 0; i <= 7; ++i) minor[i] = t1[i] - t2[i];
        pmul(B[0][0], 3, minor, 7, prod);
        for (int i = 0; i <= 10; ++i) det[i] += prod[i];
        // - k_y (l_x m_1 - l_1 m_x)
        pmul(B[1][0], 3, B[2][2], 4, t1); pmul(B[1][2], 4, B[2][0], 3, t2);
        for (int i = 0; i <= 7; ++i) minor[i] = t1[i] - t2[i];
        pmul(B[0][1], 3, minor, 7, prod);
        for (int i = 0; i <= 10; ++i) det[i] -= prod[i];
        // + k_1 (l_x m_y - l_y m_x)
        pmul(B[1][0], 3, B[2][1], 3, t1); pmul(B[1][1], 3, B[2][0], 3, t2);
        for (int i = 0; i <= 6; ++i) minor[i] = t1[i] - t2[i];
        minor[7] = 0.0;
        pmul(B[0][2], 4, minor, 6, prod);
        for (int i = 0; i <= 10; ++i) det[i] += prod[i];
    }
    double zs[10];
    const int nz = real_roots_deg10(det, zs, deflated);
#ifdef VS_5PT_TRACE
    VS_5PT_TRACE(det, zs, nz);
#endif
    int ns = 0;
    for (int s = 0; s < nz; ++s) {
        const double z = zs[s];
        double b[3][3];
        for (int r = 0; r < 3; ++r)
            for (int c = 0; c < 3; ++c) {
                double v = B[r][c][4];
                for (int i = 3; i >= 0; --i) v = v * z + B[r][c][i];
                b[r][c] = v;
            }
        // (x, y) from the best-conditioned pair of rows
        double bestD = 0.0, x = 0.0, y = 0.0;
        for (int p = 0; p < 3; ++p) {
            const int r0 = p, r1 = (p + 1) % 3;
            double D = b[r0][0] * b[r1][1] - b[r0][1] * b[r1][0];
            if (fabs(D) > fabs(bestD)) {
                bestD = D;
                x = (b[r0][1] * b[r1][2] - b[r0][2] * b[r1][1]) / D;
                y = (b[r0][2] * b[r1][0] - b[r0][0] * b[r1][2]) / D;
            }
        }
        if (bestD == 0.0 || !isfinite(x) || !isfinite(y)) continue;
        const double cn = 1.0 / sqrt(x * x + y * y + z * z + 1.0);
        starts[ns][0] = x * cn; starts[ns][1] = y * cn; starts[ns][2] = z * cn; starts[ns][3] = cn;
        ++ns;
    }
    return ns;
}

// Stage 2: polish one start; on success Ek = the solution as a unit-Frobenius-norm matrix (singular values
// (1/sqrt2, 1/sqrt2, 0), as cv2 returns it).  Returns 1 = solution, 0 = not a solution of the constraints
// (the pass is then "in trouble"), -1 = degenerate, skip silently.
VS_HD VS_NOINLINE int five_pt_finish(const double L[9][4], const double start[4], double Ek[9])
{
    double cc[4] = {start[0], start[1], start[2], start[3]};
    double r_start;
    const double r_end = polish_5pt(L, cc, &r_start);
    if (!(r_end < 1e-16)) return 0;
    double nrm = 0.0;
    for (int e = 0; e < 9; ++e) {
        Ek[e] = cc[0] * L[e][0] + cc[1] * L[e][1] + cc[2] * L[e][2] + cc[3] * L[e][3];
        nrm += Ek[e] * Ek[e];
    }
    if (!(nrm > 0.0) || !isfinite(nrm)) return -1;
    nrm = 1.0 / sqrt(nrm);
    for (int e = 0; e < 9; ++e) Ek[e] *= nrm;
    return 1;
}

// Stage 3: append Ek to E[0..n) unless it is already there (up to sign).  Returns the new count; *dup says it was.
VS_HD inline int five_pt_merge(double E[10][9], int n, const double Ek[9], bool* dup)
{
    *dup = false;
    for (int j = 0; j < n && !*dup; ++j) {
        double dp = 0.0, dm = 0.0;
        for (int e = 0; e < 9; ++e) {
            dp = fmax(dp, fabs(E[j][e] - Ek[e]));
            dm = fmax(dm, fabs(E[j][e] + Ek[e]));
        }
        *dup = fmin(dp, dm) < 1e-7;
    }
    if (*dup || n >= 10) return n;
    for (int e = 0; e < 9; ++e) E[n][e] = Ek[e];
    return n + 1;
}

VS_HD inline int solve_5pt_in_basis(const double nb[4][9], double E[10][9], int n, bool* trouble)
{
    const int have_before = n;
    double L[9][4], starts[10][4];
    bool deflated = false;
    const int ns = five_pt_starts(nb, L, starts, &deflated);
    if (deflated) *trouble = true;
    for (int s = 0; s < ns; ++s) {
        double Ek[9];
        const int ok = five_pt_finish(L, starts[s], Ek);
        if (ok == 0) *trouble = true;
        if (ok <= 0) continue;
        bool dup;
        n = five_pt_merge(E, n, Ek, &dup);
        if (dup && have_before <= 0) *trouble = true;
    }
    // the degree-10 polynomial has real coefficients: its real roots come in an even number (with multiplicity).
    // An odd count of distinct solutions after the first pass means one was missed or two collapsed.
    if (have_before <= 0 && (n & 1)) *trouble = true;
    return n;
}

// the fixed orthogonal mix of the Gauss-Jordan null-space basis the first pass runs in (see solve_5pt)
VS_HD inline void five_pt_rotate_basis(const double nb[4][9], double rb[4][9])
{
    for (int e = 0; e < 9; ++e) {
        const double a = nb[0][e], b = nb[1][e], c = nb[2][e], d = nb[3][e];
        rb[0][e] = 0.5 * (a + b + c + d);
        rb[1][e] = 0.5 * (a - b + c - d);
        rb[2][e] = 0.5 * (a + b - c - d);
        rb[3][e] = 0.5 * (a - b - c + d);
    }
}

VS_HD inline void five_pt_null_space(const double x1[5][2], const double x2[5][2], double nb[4][9])
{
    double A[5][9];
    for (int i = 0; i < 5; ++i) {
        const double u = x1[i][0], v = x1[i][1], up = x2[i][0], vp = x2[i][1];
        A[i][0] = up * u; A[i][1] = up * v; A[i][2] = up;
        A[i][3] = vp * u; A[i][4] = vp * v; A[i][5] = vp;
        A[i][6] = u; A[i][7] = v; A[i][8] = 1.0;
    }
    null_space_9<5>(A, nb);
}

// 5-point solver.  x1, x2: 5 correspondences in normalised camera coordinates.
// Writes up to 10 essential matrices (row-major, x2^T E x1 = 0, Frobenius norm 1); returns the count.
// The roots are those of the five-point problem itself (every solution is polished on the cubic constraints), not of
// one particular elimination: when the first pass reports trouble -- the z-coordinates of several solutions cluster in
// this basis and the degree-10 polynomial is ill-conditioned -- a second pass in a rotated basis (a fixed orthogonal
// 4 x 4 mix, so other coordinates play the role of z) contributes the solutions the first one missed.
VS_HD inline int solve_5pt(const double x1[5][2], const double x2[5][2], double E[10][9])
{
    double nb[4][9], rb[4][9];
    five_pt_null_space(x1, x2, nb);
    // First pass in a rotated basis (a fixed orthogonal 4 x 4 mix of the Gauss-Jordan basis): the un-mixed basis is
    // aligned with the coordinate structure of the constraint rows, and for pure sideways motion (a panning camera)
    // that puts solutions at infinity of the (x, y, z, 1) chart on almost every second sample; mixed, that alignment
    // is gone and the second pass -- in the un-mixed basis -- is needed for the rare clustered-root cases only.
    five_pt_rotate_basis(nb, rb);
    bool trouble = false;
    int n = solve_5pt_in_basis(rb, E, 0, &trouble);
    if (trouble) {
        bool t2 = false;
        n = solve_5pt_in_basis(nb, E, n, &t2);
    }
    return n;
}

// One-sided Jacobi SVD of an N x N matrix A (row-major, overwritten: A <- U diag(s)); V
// accumulates the right singular vectors in its columns.  Returns nothing; the singular
// values are the column norms of the result.
template <int N>
VS_HD inline void jacobi_svd(double A[N][N], double V[N][N])
{
    for (int i = 0; i < N; ++i)
        for (int j = 0; j < N; ++j) V[i][j] = i == j ? 1.0 : 0.0;
    for (int sweep = 0; sweep < 30; ++sweep) {
        double off = 0.0;
        for (int p = 0; p < N - 1; ++p)
            for (int q = p + 1; q < N; ++q) {
                double a = 0.0, b = 0.0, g = 0.0;
                for (int i = 0; i < N; ++i) { a += A[i][p] * A[i][p]; b += A[i][q] * A[i][q]; g += A[i][p] * A[i][q]; }
                if (fabs(g) <= 1e-16 * sqrt(a * b) || g == 0.0) continue;
                off = fmax(off, fabs(g) / sqrt(a * b + 1e-300));
                double zeta = (b - a) / (2.0 * g);
                double t = (zeta >= 0 ? 1.0 : -1.0) / (fabs(zeta) + sqrt(1.0 + zeta * zeta));
                double cs = 1.0 / sqrt(1.0 + t * t), sn = cs * t;
                for (int i = 0; i < N; ++i) {
                    double ap = A[i][p], aq = A[i][q];
                    A[i][p] = cs * ap - sn * aq;
                    A[i][q] = sn * ap + cs * aq;
                    double vp = V[i][p], vq = V[i][q];
                    V[i][p] = cs * vp - sn * vq;
                    V[i][q] = sn * vp + cs * vq;
                }
            }
        if (off < 1e-15) break;
    }
}
