// solvers_host.cpp -- the product's minimal solvers (csrc/geom_solvers.cuh, geom_5pt.cuh are __host__ __device__)
// compiled for the CPU so that tools/ransac_agreement.py can replay OpenCV's RANSAC loop with them in the build
// container (no GPU here).  Development tool only: not part of libvslam_b200.so, not used by tests or bench.
//   g++ -O2 -shared -fPIC -o /tmp/libvs_solvers_host.so tools/host_harness/solvers_host.cpp
#include "../../visual-slam_b200/csrc/geom_5pt.cuh"

extern "C" int h_solve_5pt(const double* x1, const double* x2, double* E)
{
    double a[5][2], b[5][2], out[10][9];
    for (int i = 0; i < 5; ++i) { a[i][0] = x1[2 * i]; a[i][1] = x1[2 * i + 1]; b[i][0] = x2[2 * i]; b[i][1] = x2[2 * i + 1]; }
    const int n = solve_5pt(a, b, out);
    for (int k = 0; k < n; ++k)
        for (int e = 0; e < 9; ++e) E[k * 9 + e] = out[k][e];
    return n;
}

extern "C" int h_solve_7pt(const double* x1, const double* x2, double* F)
{
    double a[7][2], b[7][2], out[3][9];
    for (int i = 0; i < 7; ++i) { a[i][0] = x1[2 * i]; a[i][1] = x1[2 * i + 1]; b[i][0] = x2[2 * i]; b[i][1] = x2[2 * i + 1]; }
    const int n = solve_7pt(a, b, out);
    for (int k = 0; k < n; ++k)
        for (int e = 0; e < 9; ++e) F[k * 9 + e] = out[k][e];
    return n;
}
