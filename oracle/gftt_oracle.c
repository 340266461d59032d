/*
 * oracle/gftt_oracle.c -- CPU restatement of ORBExtractor.distribute_keypoints' corner stage.
 *
 * TEST INFRASTRUCTURE ONLY (see orb_oracle.c header for the rules).
 *
 * Restates what the reference reaches through
 *   src/orbslam2/extractor.py:115-129   64 x cv2.goodFeaturesToTrack(image, n//64, 0.01, 10, mask=cell)
 * The arithmetic lives in OpenCV imgproc (cornerMinEigenVal, threshold, dilate; not vendored).
 * It follows SURVEY.md section 8(a) K8 with the float details decoded against cv2 4.13
 * in this container (DESIGN.md "GFTT arithmetic"):
 *   Dx = Sobel(1,0,3, scale): row [-1 0 1] exact, column fma(s, r[-1]+r[+1], 2s*r[0])
 *   Dy = Sobel(0,1,3, scale): row s*x[-1], fma(2s, x[0]), fma(s, x[+1]); column r[+1]-r[-1]
 *        for x < w - w%32 (cv2 vector body); the last w%32 columns multiply and add without FMA
 *   cov = (Dx*Dx, Dx*Dy, Dy*Dy) float; 3x3 unnormalised box sum accumulated in DOUBLE with
 *        OpenCV's running column sum (SUM += new row; out = (float)SUM; SUM -= old row)
 *   eig = (a + c) - sqrt((a - c)*(a - c) + b*b), a = cov_xx*0.5f, b = cov_xy, c = cov_yy*0.5f
 * Pinned by tests/test_oracle_golden.py against distribute_keypoints() outputs of the reference.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

static inline int refl(int i, int n)
{
    if (n == 1) return 0;
    while (i < 0 || i >= n) i = i < 0 ? -i : 2 * n - 2 - i;
    return i;
}

/* tail rules of cv2's separable filter (vector body uses FMA, scalar tail does not) */
static int g_row_block = 32, g_col_block = 1;
void orc_gftt_set_blocks(int row_block, int col_block) { g_row_block = row_block; g_col_block = col_block; }

void orc_min_eigen_map(const uint8_t* img, int w, int h, float* eig)
{
    volatile float sc = (float)(1.0 / (4 * 3 * 255.0));
    volatile float s1 = sc, s2 = 2.0f * sc;
    float* dx = (float*)malloc(sizeof(float) * (size_t)w * h);
    float* dy = (float*)malloc(sizeof(float) * (size_t)w * h);
    float* rowd = (float*)malloc(sizeof(float) * (size_t)w * (h + 2));   /* row pass of Dx: x[+1]-x[-1] */
    float* rows = (float*)malloc(sizeof(float) * (size_t)w * (h + 2));   /* row pass of Dy: smoothed */
    const int wrow = w - w % g_row_block, wcol = w - w % g_col_block;
    for (int y = -1; y <= h; ++y) {
        const uint8_t* p = img + (long)refl(y, h) * w;
        for (int x = 0; x < w; ++x) {
            float a = (float)p[refl(x - 1, w)], b = (float)p[x], c = (float)p[refl(x + 1, w)];
            rowd[(long)(y + 1) * w + x] = c - a;
            volatile float s = s1 * a;
            if (x < wrow) { s = fmaf(s2, b, s); s = fmaf(s1, c, s); }
            else { volatile float m = s2 * b; s = s + m; m = s1 * c; s = s + m; }
            rows[(long)(y + 1) * w + x] = s;
        }
    }
    for (int y = 0; y < h; ++y)
        for (int x = 0; x < w; ++x) {
            const float* r = rowd + (long)(y + 1) * w + x;
            volatile float pr = r[-w] + r[w];
            volatile float c0 = s2 * r[0];
            float v;
            if (x < wcol) v = fmaf(s1, pr, c0);
            else { volatile float m = s1 * pr; v = c0 + m; }
            dx[(long)y * w + x] = v;
            const float* q = rows + (long)(y + 1) * w + x;
            dy[(long)y * w + x] = q[w] - q[-w];
        }
    /* cov + 3x3 box (double running column sums) + min eigenvalue */
    double* sum = (double*)calloc((size_t)w * 3, sizeof(double));
    double* ring = (double*)malloc(sizeof(double) * (size_t)w * 3 * 3);   /* row sums of rows y-1, y, y+1 */
#define ROWSUM(dst, yy)                                                                       \
    do {                                                                                      \
        int ry = refl((yy), h);                                                               \
        for (int x = 0; x < w; ++x) {                                                         \
            double acc[3] = {0, 0, 0};                                                        \
            for (int k = -1; k <= 1; ++k) {                                                   \
                int rx = refl(x + k, w);                                                      \
                volatile float gx = dx[(long)ry * w + rx], gy = dy[(long)ry * w + rx];        \
                volatile float xx = gx * gx, xy = gx * gy, yy2 = gy * gy;                     \
                acc[0] += (double)xx; acc[1] += (double)xy; acc[2] += (double)yy2;            \
            }                                                                                 \
            (dst)[3 * x] = acc[0]; (dst)[3 * x + 1] = acc[1]; (dst)[3 * x + 2] = acc[2];      \
        }                                                                                     \
    } while (0)
    double* r0 = ring;
    double* r1 = ring + (size_t)w * 3;
    double* r2 = ring + (size_t)w * 6;
    ROWSUM(r0, -1);
    ROWSUM(r1, 0);
    for (int i = 0; i < 3 * w; ++i) sum[i] = (0.0 + r0[i]) + r1[i];
    for (int y = 0; y < h; ++y) {
        ROWSUM(r2, y + 1);
        for (int x = 0; x < w; ++x) {
            float cv[3];
            for (int k = 0; k < 3; ++k) {
                double s0 = sum[3 * x + k] + r2[3 * x + k];
                cv[k] = (float)s0;
                sum[3 * x + k] = s0 - r0[3 * x + k];
            }
            volatile float a = cv[0] * 0.5f, b = cv[1], c = cv[2] * 0.5f;
            volatile float t = a - c;
            volatile float t2 = t * t, b2 = b * b;
            volatile float u = t2 + b2;
            volatile float ac = a + c;
            eig[(long)y * w + x] = ac - sqrtf(u);
        }
        double* tmp = r0; r0 = r1; r1 = r2; r2 = tmp;
    }
    free(dx); free(dy); free(rowd); free(rows); free(sum); free(ring);
}

typedef struct { float v; int addr; } gcand_t;
static int cmp_gcand(const void* a, const void* b)
{
    const gcand_t* p = (const gcand_t*)a;
    const gcand_t* q = (const gcand_t*)b;
    if (p->v > q->v) return -1;
    if (p->v < q->v) return 1;
    return p->addr > q->addr ? -1 : p->addr < q->addr ? 1 : 0;   /* ties: larger address first */
}

/* extractor.py:85-144 corner stage: rows x cols grid of goodFeaturesToTrack calls.
 * xy[2*i], cell_count[rows*cols]; returns the number of corners. */
int orc_gftt_grid(const uint8_t* img, int h, int w, int rows, int cols, int max_per_cell, double quality,
                  double min_distance, float* xy, int32_t* cell_count)
{
    float* eig = (float*)malloc(sizeof(float) * (size_t)w * h);
    orc_min_eigen_map(img, w, h, eig);
    const int ch = h / rows, cw = w / cols;
    gcand_t* cand = (gcand_t*)malloc(sizeof(gcand_t) * (size_t)(ch * cw + 1));
    int total = 0;
    const double md2 = min_distance * min_distance;
    for (int ci = 0; ci < rows; ++ci)
        for (int cj = 0; cj < cols; ++cj) {
            const int y0 = ci * ch, y1 = (ci + 1) * ch, x0 = cj * cw, x1 = (cj + 1) * cw;
            cell_count[ci * cols + cj] = 0;
            if (max_per_cell <= 0) continue;
            float mx = -INFINITY;
            for (int y = y0; y < y1; ++y)
                for (int x = x0; x < x1; ++x) mx = fmaxf(mx, eig[(long)y * w + x]);
            const float thr = (float)((double)mx * quality);
            int n = 0;
            for (int y = (y0 > 1 ? y0 : 1); y < (y1 < h - 1 ? y1 : h - 1); ++y)
                for (int x = (x0 > 1 ? x0 : 1); x < (x1 < w - 1 ? x1 : w - 1); ++x) {
                    const float v = eig[(long)y * w + x];
                    if (!(v > thr)) continue;             /* THRESH_TOZERO then val != 0 */
                    int ismax = 1;
                    for (int dy = -1; dy <= 1 && ismax; ++dy)
                        for (int dx = -1; dx <= 1; ++dx) {
                            float nb = eig[(long)(y + dy) * w + x + dx];
                            if (nb > thr && nb > v) { ismax = 0; break; }
                        }
                    if (ismax) { cand[n].v = v; cand[n].addr = y * w + x; ++n; }
                }
            qsort(cand, n, sizeof(gcand_t), cmp_gcand);
            int base = total, acc = 0;
            for (int i = 0; i < n && acc < max_per_cell; ++i) {
                const int y = cand[i].addr / w, x = cand[i].addr % w;
                int good = 1;
                for (int j = 0; j < acc && good; ++j) {
                    double ddx = x - xy[2 * (base + j)], ddy = y - xy[2 * (base + j) + 1];
                    if (ddx * ddx + ddy * ddy < md2) good = 0;
                }
                if (!good) continue;
                xy[2 * (base + acc)] = (float)x;
                xy[2 * (base + acc) + 1] = (float)y;
                ++acc;
            }
            cell_count[ci * cols + cj] = acc;
            total += acc;
        }
    free(cand);
    free(eig);
    return total;
}
