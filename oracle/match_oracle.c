/*
 * oracle/match_oracle.c -- CPU restatement of brute-force Hamming matching.
 *
 * TEST INFRASTRUCTURE ONLY (see orb_oracle.c header for the rules).
 *
 * Restates what the reference reaches through
 *   src/orbslam2/matcher.py:29      cv2.BFMatcher(cv2.NORM_HAMMING, crossCheck=False)
 *   src/orbslam2/matcher.py:70      self.matcher.knnMatch(d1, d2, k=2)
 *   src/orbslam2/matcher.py:73-81   Lowe ratio loop (m.distance < ratio * n.distance)
 *   src/orbslam2/matcher.py:102     self.matcher.match(d1, d2, mask=mask)
 * The arithmetic lives in OpenCV (features2d BFMatcher / batchDistance); semantics
 * follow SURVEY.md section 8(a) K9 as probed against cv2 4.13: int32 distances,
 * train scanned in ascending index with strict '<' insertion (ties -> lowest
 * trainIdx, for first and second place), M = 1 -> a single neighbour.
 */
#include <stdint.h>
#include <string.h>

static inline int hamming256(const uint8_t* a, const uint8_t* b)
{
    uint64_t x[4], y[4];
    memcpy(x, a, 32);
    memcpy(y, b, 32);
    return __builtin_popcountll(x[0] ^ y[0]) + __builtin_popcountll(x[1] ^ y[1]) +
           __builtin_popcountll(x[2] ^ y[2]) + __builtin_popcountll(x[3] ^ y[3]);
}

int orc_hamming256(const uint8_t* a, const uint8_t* b) { return hamming256(a, b); }

/* knnMatch(k=2): idx[q][2], dist[q][2]; missing neighbours are (-1, -1). */
void orc_bf_knn2(const uint8_t* q, int nq, const uint8_t* t, int nt, int32_t* idx, int32_t* dist)
{
    for (int i = 0; i < nq; ++i) {
        int d1 = 1 << 30, d2 = 1 << 30, i1 = -1, i2 = -1;
        for (int j = 0; j < nt; ++j) {
            int d = hamming256(q + 32L * i, t + 32L * j);
            if (d < d1) { d2 = d1; i2 = i1; d1 = d; i1 = j; }
            else if (d < d2) { d2 = d; i2 = j; }
        }
        idx[2 * i] = i1; dist[2 * i] = i1 >= 0 ? d1 : -1;
        idx[2 * i + 1] = i2; dist[2 * i + 1] = i2 >= 0 ? d2 : -1;
    }
}

/* match(mask=): k = 1 restricted to mask[q][t] != 0; rows with no allowed train -> -1. */
void orc_bf_match_masked(const uint8_t* q, int nq, const uint8_t* t, int nt, const uint8_t* mask,
                         int32_t* idx, int32_t* dist)
{
    for (int i = 0; i < nq; ++i) {
        int d1 = 1 << 30, i1 = -1;
        for (int j = 0; j < nt; ++j) {
            if (mask && !mask[(long)i * nt + j]) continue;
            int d = hamming256(q + 32L * i, t + 32L * j);
            if (d < d1) { d1 = d; i1 = j; }
        }
        idx[i] = i1;
        dist[i] = i1 >= 0 ? d1 : -1;
    }
}

/* matcher.py:73-81: keep[q] = 1 when the row survives the ratio loop.  The reference
 * compares Python doubles of float32-valued integer distances. */
int orc_ratio_filter(const int32_t* idx, const int32_t* dist, int nq, double ratio, int ratio_test,
                     uint8_t* keep)
{
    int n = 0;
    for (int i = 0; i < nq; ++i) {
        int k = 0;
        if (idx[2 * i] >= 0 && idx[2 * i + 1] >= 0)
            k = !ratio_test || (double)(float)dist[2 * i] < ratio * (double)(float)dist[2 * i + 1];
        else if (idx[2 * i] >= 0)
            k = 1;
        keep[i] = (uint8_t)k;
        n += k;
    }
    return n;
}
