/*
 * stereo_oracle.c -- CPU ORACLE (test infrastructure, NOT product code).
 *
 * Plain-C restatement of the arithmetic that the reference
 * (ocean1100/stereo_match) hands to OpenCV on its hot path:
 *   stereo_vision/stereo_vision.py:153-166,178   cv2.StereoSGBM / cv2.StereoBM compute
 *   stereo_vision/stereo_vision.py:209, disparity_calculation.py:302   cv2.reprojectImageTo3D
 *   rectified_img_cal.py:253-254, disparity_calculation.py:207-208     cv2.remap
 *   disparity_calculation.py:286-287                                   cv2.cvtColor(BGR2GRAY)
 *
 * The arithmetic lives in a third-party dependency that is NOT under
 * /root/reference: OpenCV (calib3d / imgproc).  The reference pins no version;
 * the version in this image is opencv-python-headless 4.13.0.92.  This file
 * follows the behavioural specification in SURVEY.md Appendix A and is PINNED
 * by tests/test_oracle_vs_cv2.py (live cv2 4.13.0, bit-exact) and by the
 * committed fixtures under tests/golden/ (generated from cv2 by
 * tests/golden/make_golden.py).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load this library.  The product
 * (stereo_match_b200/) never does.
 *
 * Written for clarity, one loop nest per stage, direction-decomposed path
 * aggregation (legal because S = sat16(sum_r L_r) is order independent).
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>

#define ORC_API __attribute__((visibility("default")))

static inline int imin(int a, int b) { return a < b ? a : b; }
static inline int imax(int a, int b) { return a > b ? a : b; }
static inline int iclamp(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }
static inline int16_t wrap16(int v) { return (int16_t)(uint16_t)(v & 0xffff); }
static inline int16_t sat16(int v) { return (int16_t)iclamp(v, -32768, 32767); }

typedef struct {
    int minDisparity, numDisparities, blockSize, P1, P2, disp12MaxDiff;
    int preFilterCap, uniquenessRatio, speckleWindowSize, speckleRange, mode;
} orc_sgbm_params;

/* ------------------------------------------------------------------ */
/* per-row descriptors (clipped x-Sobel g, raw r) and   */
/* their half-pixel intervals.  Planes are H x W uint8.                */
/* ------------------------------------------------------------------ */
static void descriptors(const uint8_t* I, int H, int W, int ftzero,
                        uint8_t* g, uint8_t* r)
{
    for (int y = 0; y < H; y++) {
        int yn = y > 0 ? y - 1 : y, ys = y < H - 1 ? y + 1 : y;
        const uint8_t *c = I + (size_t)y * W, *n = I + (size_t)yn * W, *s = I + (size_t)ys * W;
        uint8_t *gy = g + (size_t)y * W, *ry = r + (size_t)y * W;
        for (int x = 1; x < W - 1; x++) {
            int v = 2 * (c[x + 1] - c[x - 1]) + (n[x + 1] - n[x - 1]) + (s[x + 1] - s[x - 1]);
            gy[x] = (uint8_t)(iclamp(v, -ftzero, ftzero) + ftzero);
            ry[x] = c[x];
        }
        gy[0] = gy[W - 1] = (uint8_t)ftzero;
        ry[0] = ry[W - 1] = (uint8_t)ftzero;
    }
}

static void halfpix(const uint8_t* p, int W, uint8_t* lo, uint8_t* hi)
{
    for (int x = 0; x < W; x++) {
        int v = p[x];
        int pl = x > 0 ? (v + p[x - 1]) / 2 : v;
        int pr = x < W - 1 ? (v + p[x + 1]) / 2 : v;
        lo[x] = (uint8_t)imin(imin(pl, pr), v);
        hi[x] = (uint8_t)imax(imax(pl, pr), v);
    }
}

static inline int bt(int u, int v, int lo1, int hi1, int lo2, int hi2)
{
    int c0 = imax(0, imax(u - hi2, lo2 - u));
    int c1 = imax(0, imax(v - hi1, lo1 - v));
    return imin(c0, c1);
}

/* effective parameters, SURVEY.md A.0 */
typedef struct {
    int minD, D, maxD, SW2, SH2, P1, P2, uniq, lrmax, ftzero, minX1, maxX1, width1, INVALID;
    int chkSW2;    /* half window of the size check */
    int rawbs;     /* blockSize as given (MODE_SGBM_3WAY's stripe overlap uses it without the fallback) */
} eff_t;

static eff_t effective(const orc_sgbm_params* p, int W)
{
    eff_t e;
    e.minD = p->minDisparity; e.D = p->numDisparities; e.maxD = e.minD + e.D;
    int bs = p->blockSize > 0 ? p->blockSize : 5;
    e.SW2 = e.SH2 = bs / 2;
    e.chkSW2 = bs / 2;                /* the "image too small" test is common code: always the 5-pixel fallback */
    e.rawbs = p->blockSize;
    /* MODE_SGBM_3WAY has its own fallback: blockSize <= 0 means a 3-pixel window there (probed against cv2 4.13.0) */
    if (p->mode == 2 && p->blockSize <= 0) e.SW2 = e.SH2 = 1;
    e.P1 = p->P1 > 0 ? p->P1 : 2;
    e.P2 = imax(p->P2 > 0 ? p->P2 : 5, e.P1 + 1);
    e.uniq = p->uniquenessRatio >= 0 ? p->uniquenessRatio : 10;
    e.lrmax = p->disp12MaxDiff > 0 ? p->disp12MaxDiff : 1;
    e.ftzero = imax(p->preFilterCap, 15) | 1;
    e.minX1 = imax(e.maxD, 0);
    e.maxX1 = W + imin(e.minD, 0);
    e.width1 = e.maxX1 - e.minX1;
    e.INVALID = (e.minD - 1) * 16;
    return e;
}

/* ------------------------------------------------------------------ */
/* SURVEY.md A.1 + A.2: C[y][x][d] = i16(P2 + block sum of BT cost).   */
/* Cvol must hold H*width1*D int16.                                    */
/* ------------------------------------------------------------------ */
/* cn = 1 (gray) or 3 (interleaved colour, any order): every colour channel contributes its own Sobel and raw
   channel, each raw one shifted by 2 on its own (probed against cv2 4.13.0 with 3-channel input) */
static int32_t* hsum_volume(const uint8_t* L, const uint8_t* R, int H, int W, int cn, const eff_t* e)
{
    const int D = e->D, w1 = e->width1;
    size_t plane = (size_t)H * W;
    uint8_t* buf = (uint8_t*)malloc(plane * 12 * cn);
    uint8_t* chan = (uint8_t*)malloc(plane * 2);
    for (int c = 0; c < cn; c++) {
        uint8_t* b = buf + plane * 12 * c;
        uint8_t *g1 = b, *r1 = b + plane, *g2 = b + 2 * plane, *r2 = b + 3 * plane;
        for (size_t i = 0; i < plane; i++) { chan[i] = L[i * cn + c]; chan[plane + i] = R[i * cn + c]; }
        descriptors(chan, H, W, e->ftzero, g1, r1);
        descriptors(chan + plane, H, W, e->ftzero, g2, r2);
        uint8_t* pl[4] = { g1, r1, g2, r2 };
        for (int k = 0; k < 4; k++)
            for (int y = 0; y < H; y++)
                halfpix(pl[k] + (size_t)y * W, W, b + (4 + 2 * k) * plane + (size_t)y * W, b + (5 + 2 * k) * plane + (size_t)y * W);
    }
    free(chan);

    /* pix[y][x][d], then horizontal sums hs[y][x][d] (clamp on cropped axis) */
    size_t rowsz = (size_t)w1 * D;
    uint32_t* pix = (uint32_t*)malloc(rowsz * sizeof(uint32_t));
    int32_t* hs = (int32_t*)malloc((size_t)H * rowsz * sizeof(int32_t));
    for (int y = 0; y < H; y++) {
        size_t o = (size_t)y * W;
        memset(pix, 0, rowsz * sizeof(uint32_t));
        for (int c = 0; c < cn; c++) {
            const uint8_t* b = buf + plane * 12 * c;
            const uint8_t *g1 = b, *r1 = b + plane, *g2 = b + 2 * plane, *r2 = b + 3 * plane;
            const uint8_t *lo[4], *hi[4];
            for (int k = 0; k < 4; k++) { lo[k] = b + (4 + 2 * k) * plane; hi[k] = b + (5 + 2 * k) * plane; }
            for (int x = 0; x < w1; x++) {
                int xi = x + e->minX1;
                for (int d = 0; d < D; d++) {
                    int xr = xi - (d + e->minD);
                    int cg = bt(g1[o + xi], g2[o + xr], lo[0][o + xi], hi[0][o + xi], lo[2][o + xr], hi[2][o + xr]);
                    int cr = bt(r1[o + xi], r2[o + xr], lo[1][o + xi], hi[1][o + xi], lo[3][o + xr], hi[3][o + xr]);
                    pix[(size_t)x * D + d] += (uint32_t)(cg + (cr >> 2));
                }
            }
        }
        int32_t* hrow = hs + (size_t)y * rowsz;
        for (int x = 0; x < w1; x++)
            for (int d = 0; d < D; d++) {
                int s = 0;
                for (int k = -e->SW2; k <= e->SW2; k++)
                    s += (int)pix[(size_t)iclamp(x + k, 0, w1 - 1) * D + d];
                hrow[(size_t)x * D + d] = s;
            }
    }
    free(pix); free(buf);
    return hs;
}

static void cost_volume(const uint8_t* L, const uint8_t* R, int H, int W, int cn,
                        const eff_t* e, int16_t* Cvol)
{
    int32_t* hs = hsum_volume(L, R, H, W, cn, e);
    size_t rowsz = (size_t)e->width1 * e->D;
    for (int y = 0; y < H; y++) {
        int16_t* crow = Cvol + (size_t)y * rowsz;
        for (size_t i = 0; i < rowsz; i++) {
            int s = e->P2;
            for (int k = -e->SH2; k <= e->SH2; k++)
                s += hs[(size_t)iclamp(y + k, 0, H - 1) * rowsz + i];
            crow[i] = wrap16(s);
        }
    }
    free(hs);
}

/* ------------------------------------------------------------------ */
/* one path direction (dx,dy) = step from predecessor   */
/* to pixel, i.e. q = p - (dx,dy).  Adds L_r into Ssum (int32).        */
/* Lout (optional) receives L_r as int16, same layout as C.            */
/* ------------------------------------------------------------------ */
/* tail0 > 0: MODE_SGBM_3WAY with D not a multiple of 8 (probed against cv2 4.13.0, tests/test_oracle_vs_cv2.py).   */
/* cv2 runs d < tail0 = 8 * (D / 8) on SIMD lanes and the rest in a scalar tail whose FIRST element takes its      */
/* d - 1 neighbour from the wrong place in the two paths it updates in place:                                      */
/*   tailmode 1 (down): from the buffer the SIMD loop has just stored to -- the NEW L_r(p, tail0 - 1) of this      */
/*                      pixel instead of the predecessor's;                                                        */
/*   tailmode 2 (<-)  : from element tail0 itself (L_r(q, tail0) + P1, which can never win: the neighbour is lost). */
/* The -> path keeps its previous pixel in a second buffer and is unaffected.  D < 8 has no SIMD part: no quirk.   */
static void aggregate_dir_ex(const int16_t* Cvol, int H, int w1, int D, int P1, int P2,
                             int dx, int dy, int tail0, int tailmode, int32_t* Ssum, int16_t* Lout)
{
    size_t rowsz = (size_t)w1 * D;
    int16_t* Lbuf = (int16_t*)malloc((size_t)H * rowsz * sizeof(int16_t));
    int* minL = (int*)malloc((size_t)H * w1 * sizeof(int));
    int y0 = dy >= 0 ? 0 : H - 1, y1 = dy >= 0 ? H : -1, ys = dy >= 0 ? 1 : -1;
    /* traversal order: rows so that predecessor row is done; inside a row,
       x in the order that makes the (dx, dy=0) predecessor available */
    int x0 = dx >= 0 ? 0 : w1 - 1, x1 = dx >= 0 ? w1 : -1, xs = dx >= 0 ? 1 : -1;
    for (int y = y0; y != y1; y += ys)
        for (int x = x0; x != x1; x += xs) {
            int qx = x - dx, qy = y - dy;
            const int16_t* Cp = Cvol + (size_t)y * rowsz + (size_t)x * D;
            int16_t* Lp = Lbuf + (size_t)y * rowsz + (size_t)x * D;
            int inside = qx >= 0 && qx < w1 && qy >= 0 && qy < H;
            const int16_t* Lq = inside ? Lbuf + (size_t)qy * rowsz + (size_t)qx * D : NULL;
            int mq = inside ? minL[(size_t)qy * w1 + qx] : 0;
            int delta = P2 + mq, mn = 32767;
            for (int d = 0; d < D; d++) {
                int a = inside ? Lq[d] : 0;
                int b = inside ? (d > 0 ? Lq[d - 1] : 32767) + P1 : 0 + P1;
                if (tail0 > 0 && d == tail0) b = (tailmode == 1 ? Lp[d - 1] : (inside ? Lq[d] : 0)) + P1;
                int c = inside ? (d < D - 1 ? Lq[d + 1] : 32767) + P1 : 0 + P1;
                /* outside: L_r(q,.)=0, minL=0 -> min(0, P1, P1, P2) = 0 */
                int v = Cp[d] + imin(imin(a, b), imin(c, delta)) - delta;
                int16_t l = wrap16(v);
                Lp[d] = l;
                if (l < mn) mn = l;
                Ssum[(size_t)y * rowsz + (size_t)x * D + d] += l;
            }
            minL[(size_t)y * w1 + x] = mn;
        }
    if (Lout) memcpy(Lout, Lbuf, (size_t)H * rowsz * sizeof(int16_t));
    free(minL); free(Lbuf);
}

static void aggregate_dir(const int16_t* Cvol, int H, int w1, int D, int P1, int P2,
                          int dx, int dy, int32_t* Ssum, int16_t* Lout)
{
    aggregate_dir_ex(Cvol, H, w1, D, P1, P2, dx, dy, 0, 0, Ssum, Lout);
}

static const int DIRS[8][2] = {
    /* pass 1 */ { 1, 0 }, { 1, 1 }, { 0, 1 }, { -1, 1 },
    /* 5th path of MODE_SGBM and first of pass 2 */ { -1, 0 },
    /* rest of pass 2 (MODE_HH) */ { 1, -1 }, { 0, -1 }, { -1, -1 }
};

/* ------------------------------------------------------------------ */
/* WTA + uniqueness + disp2 + sub-pixel + LR check.     */
/* ------------------------------------------------------------------ */
/* le != 0: MODE_SGBM_3WAY (cv2 4.13.0 works on 8 int16 SIMD lanes there; probed, tests/test_oracle_vs_cv2.py):    */
/*   - winner: lane j keeps the LAST minimum among its disparities d = j (mod 8), the lanes are then merged to  */
/*     the SMALLEST position -- differs from "first minimum" only when the minimum is tied;                    */
/*   - uniqueness: 16-bit compare S(d) < (int16)(100 * minS / (100 - ratio) + 1), the threshold wrapping       */
/*     (rejects on equality too; not at all for ratio 0).                                                      */
static void wta_rows_ex(const int16_t* S, int H, int W, const eff_t* e, int le, int16_t* disp)
{
    const int D = e->D, w1 = e->width1, INV = e->INVALID;
    int16_t* disp2 = (int16_t*)malloc(W * sizeof(int16_t));
    int* disp2cost = (int*)malloc(W * sizeof(int));
    for (int y = 0; y < H; y++) {
        int16_t* d1 = disp + (size_t)y * W;
        for (int x = 0; x < W; x++) { d1[x] = (int16_t)INV; disp2[x] = (int16_t)INV; disp2cost[x] = 32767; }
        for (int x = w1 - 1; x >= 0; x--) {
            const int16_t* Sp = S + ((size_t)y * w1 + x) * D;
            int best = 0, minS = Sp[0];
            for (int d = 1; d < D; d++) if (Sp[d] < minS) { minS = Sp[d]; best = d; }
            const int simdD = le ? ((D % 8) ? 8 * (D / 8) : D) : 0;   /* 3WAY: d < simdD on SIMD lanes, a scalar tail after */
            if (le) {
                int last[8], j;
                minS = 32767; best = 0;
                for (int d = 0; d < simdD; d++) if (Sp[d] < minS) minS = Sp[d];
                for (j = 0; j < 8; j++) last[j] = -1;
                for (int d = 0; d < simdD; d++) if (Sp[d] == minS) last[d & 7] = d;
                if (simdD > 0) {
                    best = D;
                    for (j = 0; j < 8; j++) if (last[j] >= 0 && last[j] < best) best = last[j];
                }
                for (int d = simdD; d < D; d++) if (Sp[d] < minS) { minS = Sp[d]; best = d; }   /* tail: strict, first */
            }
            /* every S saturated: cv2's search starts at (MAX_COST, bestDisp = -1) with a strict compare, finds nothing */
            /* and writes (-1 + minD) * 16 = INVALID, disp2 untouched (3WAY's non-strict lanes do find a winner)      */
            if (!le && minS >= 32767) continue;
            int d;
            const int thr16 = le && e->uniq > 0 ? wrap16(minS * 100 / (100 - e->uniq) + 1) : 0;
            for (d = 0; d < D; d++) {
                int lhs = Sp[d] * (100 - e->uniq), rhs = minS * 100;
                /* 3WAY: the scalar tail d >= simdD uses the exact 32-bit rule of the other modes */
                if ((le ? (e->uniq > 0 && (d < simdD ? Sp[d] < thr16 : lhs < rhs)) : lhs < rhs) && abs(best - d) > 1) break;
            }
            if (d < D) continue;
            int x2 = x + e->minX1 - best - e->minD;
            if (disp2cost[x2] > minS) { disp2cost[x2] = minS; disp2[x2] = (int16_t)(best + e->minD); }
            int dq;
            if (best > 0 && best < D - 1) {
                int denom2 = imax(Sp[best - 1] + Sp[best + 1] - 2 * Sp[best], 1);
                dq = best * 16 + ((Sp[best - 1] - Sp[best + 1]) * 16 + denom2) / (denom2 * 2);
            } else dq = best * 16;
            d1[x + e->minX1] = (int16_t)(dq + e->minD * 16);
        }
        for (int x = e->minX1; x < e->maxX1; x++) {
            int v = d1[x];
            if (v == INV) continue;
            int _d = v >> 4, d_ = (v + 15) >> 4, _x = x - _d, x_ = x - d_;
            if (0 <= _x && _x < W && disp2[_x] >= e->minD && abs(disp2[_x] - _d) > e->lrmax &&
                0 <= x_ && x_ < W && disp2[x_] >= e->minD && abs(disp2[x_] - d_) > e->lrmax)
                d1[x] = (int16_t)INV;
        }
    }
    free(disp2cost); free(disp2);
}

/* 3x3 median, int16, replicate border, out of place */
ORC_API void orc_median3_i16(const int16_t* src, int H, int W, int16_t* dst)
{
    for (int y = 0; y < H; y++)
        for (int x = 0; x < W; x++) {
            int16_t v[9]; int n = 0;
            for (int dy = -1; dy <= 1; dy++)
                for (int dx = -1; dx <= 1; dx++)
                    v[n++] = src[(size_t)iclamp(y + dy, 0, H - 1) * W + iclamp(x + dx, 0, W - 1)];
            for (int i = 1; i < 9; i++) {           /* insertion sort */
                int16_t k = v[i]; int j = i - 1;
                while (j >= 0 && v[j] > k) { v[j + 1] = v[j]; j--; }
                v[j + 1] = k;
            }
            dst[(size_t)y * W + x] = v[4];
        }
}

/* filterSpeckles as union-find connected components */
static int uf_find(int* p, int i) { while (p[i] != i) { p[i] = p[p[i]]; i = p[i]; } return i; }
static void uf_union(int* p, int a, int b)
{
    a = uf_find(p, a); b = uf_find(p, b);
    if (a == b) return;
    if (a < b) p[b] = a; else p[a] = b;
}

ORC_API void orc_filter_speckles(int16_t* img, int H, int W, int newVal, int maxSize, int maxDiff)
{
    size_t n = (size_t)H * W;
    int* parent = (int*)malloc(n * sizeof(int));
    int* count = (int*)calloc(n, sizeof(int));
    for (size_t i = 0; i < n; i++) parent[i] = (int)i;
    for (int y = 0; y < H; y++)
        for (int x = 0; x < W; x++) {
            int i = y * W + x, v = img[i];
            if (v == newVal) continue;
            if (x + 1 < W && img[i + 1] != newVal && abs(v - img[i + 1]) <= maxDiff) uf_union(parent, i, i + 1);
            if (y + 1 < H && img[i + W] != newVal && abs(v - img[i + W]) <= maxDiff) uf_union(parent, i, i + W);
        }
    for (size_t i = 0; i < n; i++) if (img[i] != newVal) count[uf_find(parent, (int)i)]++;
    for (size_t i = 0; i < n; i++)
        if (img[i] != newVal && count[uf_find(parent, (int)i)] <= maxSize) img[i] = (int16_t)newVal;
    free(count); free(parent);
}

static void wta_rows(const int16_t* S, int H, int W, const eff_t* e, int16_t* disp) { wta_rows_ex(S, H, W, e, 0, disp); }

/* ------------------------------------------------------------------ */
/* MODE_SGBM_3WAY of cv2 4.13.0 (probed, tests/test_oracle_vs_cv2.py): the image is cut into 4 fixed stripes   */
/* of sz = ceil(H/4) rows; stripe st starts ov = blockSize/2 + 1 + ceil(0.1*sz) rows early (clamped to the     */
/* image), restarts the vertical block sum and the downward path there, sums the paths -> , down, <- and runs */
/* the winner-take-all with the `<=` uniqueness rule.  Each stripe writes row y at buffer row (st ? 0 : ov) +  */
/* y - s0 of a buffer preset to INVALID, and output row i is buffer row ov + i % sz of stripe i / sz (identity */
/* unless the early start was clamped: then rows shift, as in cv2).                                             */
/* hs: horizontal sums [H][width1][D]; raw: H*W before the median.                                              */
/* ------------------------------------------------------------------ */
static int sgbm_3way(const int32_t* hs, int H, int W, const eff_t* e, int16_t* raw)
{
    /* the overlap takes blockSize as given, NOT the effective one (0 or negative: C division, -3 / 2 = -1) */
    const int sz = (H + 3) / 4, ov = e->rawbs / 2 + 1 + (sz + 9) / 10;   /* + ceil(0.1*sz) */
    if (ov < 0) return -1;                         /* blockSize far below zero: cv2 itself runs out of bounds */
    const size_t rowsz = (size_t)e->width1 * e->D;
    int16_t* buf = (int16_t*)malloc((size_t)4 * (sz + ov) * W * sizeof(int16_t));
    for (size_t i = 0; i < (size_t)4 * (sz + ov) * W; i++) buf[i] = (int16_t)e->INVALID;
    for (int st = 0; st < 4; st++) {
        int s0 = imax(imin(st * sz - ov, H), 0), s1 = imin((st + 1) * sz, H), hst = s1 - s0;
        if (hst <= 0) continue;
        int16_t* Cv = (int16_t*)malloc((size_t)hst * rowsz * sizeof(int16_t));
        int32_t* Ssum = (int32_t*)calloc((size_t)hst * rowsz, sizeof(int32_t));
        for (int y = s0; y < s1; y++)
            for (size_t i = 0; i < rowsz; i++) {
                int v = e->P2;
                for (int k = -e->SH2; k <= e->SH2; k++) v += hs[(size_t)iclamp(y + k, s0, H - 1) * rowsz + i];
                Cv[(size_t)(y - s0) * rowsz + i] = wrap16(v);
            }
        const int tail0 = (e->D % 8) ? 8 * (e->D / 8) : 0;
        aggregate_dir(Cv, hst, e->width1, e->D, e->P1, e->P2, 1, 0, Ssum, NULL);
        aggregate_dir_ex(Cv, hst, e->width1, e->D, e->P1, e->P2, 0, 1, tail0, 1, Ssum, NULL);
        aggregate_dir_ex(Cv, hst, e->width1, e->D, e->P1, e->P2, -1, 0, tail0, 2, Ssum, NULL);
        int16_t* S = Cv;                                                /* reuse */
        for (size_t i = 0; i < (size_t)hst * rowsz; i++) S[i] = sat16(Ssum[i]);
        int16_t* r = (int16_t*)malloc((size_t)hst * W * sizeof(int16_t));
        wta_rows_ex(S, hst, W, e, 1, r);
        for (int y = s0; y < s1; y++) {
            int idx = (st == 0 ? ov : 0) + y - s0;
            if (idx < sz + ov) memcpy(buf + ((size_t)st * (sz + ov) + idx) * W, r + (size_t)(y - s0) * W, W * sizeof(int16_t));
        }
        free(r); free(Ssum); free(Cv);
    }
    for (int i = 0; i < H; i++)
        memcpy(raw + (size_t)i * W, buf + ((size_t)(i / sz) * (sz + ov) + ov + i % sz) * W, W * sizeof(int16_t));
    free(buf);
    return 0;
}

/* ------------------------------------------------------------------ */
/* StereoSGBM.compute, cn = 1 or 3 interleaved channels.  Optional outputs (may be NULL) */
/*   Cout, Sout : H*width1*D int16 (not for MODE_SGBM_3WAY) ; raw : H*W int16 (before median)   */
/* returns 0, -1 for unsupported mode, -2 where cv2 raises cv2.error.   */
/* ------------------------------------------------------------------ */
ORC_API int orc_sgbm_compute_cn(const uint8_t* left, const uint8_t* right, int H, int W, int cn,
                                const orc_sgbm_params* p, int16_t* disp,
                                int16_t* Cout, int16_t* Sout, int16_t* raw)
{
    /* direction sets (indices into DIRS): MODE_SGBM 5 paths, MODE_HH 8, MODE_HH4 the two horizontal and the
       two vertical ones (probed against cv2 4.13.0, tests/test_oracle_vs_cv2.py) */
    unsigned dirmask;
    if (p->mode == 0) dirmask = 0x1f;
    else if (p->mode == 1) dirmask = 0xff;
    else if (p->mode == 2) dirmask = 0;
    else if (p->mode == 3) dirmask = 0x55;
    else return -1;
    if (cn != 1 && cn != 3) return -1;
    eff_t e = effective(p, W);
    size_t n = (size_t)H * W;
    /* cv2 4.13.0 (stereosgbm.cpp:511) raises cv2.error unless
       width - (minDisparity + numDisparities) > SADWindowSize/2 */
    /* MODE_HH4 is separate code in cv2 and has no such test: it computes whatever width is left (probed: bit-exact with   */
    /* this restatement down to width1 = 1, all INVALID below).  MODE_SGBM_3WAY only refuses W - maxD <= 0, but what it   */
    /* returns for 0 < W - maxD <= blockSize/2 does not follow from the algorithm: kept as an error here.                  */
    if (p->mode != 3 && !(W - e.maxD > e.chkSW2)) return -2;
    if (e.width1 <= 0) { for (size_t i = 0; i < n; i++) disp[i] = (int16_t)e.INVALID; return 0; }
    int16_t* tmp = (int16_t*)malloc(n * sizeof(int16_t));
    if (p->mode == 2) {
        if (Cout || Sout) { free(tmp); return -1; }
        int32_t* hs = hsum_volume(left, right, H, W, cn, &e);
        int rc3 = sgbm_3way(hs, H, W, &e, tmp);
        free(hs);
        if (rc3) { free(tmp); return rc3; }
    } else {
        size_t vol = (size_t)H * e.width1 * e.D;
        int16_t* Cvol = (int16_t*)malloc(vol * sizeof(int16_t));
        int32_t* Ssum = (int32_t*)calloc(vol, sizeof(int32_t));
        cost_volume(left, right, H, W, cn, &e, Cvol);
        if (p->mode == 3) {
            /* cv2 4.13.0 MODE_HH4: the rows whose vertical window reaches below the image (y > 0, y + SH2 >= H) are
               never filled, C stays at its initial value P2 (found by probing: blockSize 1 matches the plain
               4-direction sum, every larger block only with this rule) */
            for (int y = imax(1, H - e.SH2); y < H; y++)
                for (size_t i = 0; i < (size_t)e.width1 * e.D; i++) Cvol[(size_t)y * e.width1 * e.D + i] = wrap16(e.P2);
        }
        for (int k = 0; k < 8; k++)
            if (dirmask >> k & 1) aggregate_dir(Cvol, H, e.width1, e.D, e.P1, e.P2, DIRS[k][0], DIRS[k][1], Ssum, NULL);
        int16_t* S = (int16_t*)malloc(vol * sizeof(int16_t));
        for (size_t i = 0; i < vol; i++) S[i] = sat16(Ssum[i]);
        wta_rows(S, H, W, &e, tmp);
        if (Cout) memcpy(Cout, Cvol, vol * sizeof(int16_t));
        if (Sout) memcpy(Sout, S, vol * sizeof(int16_t));
        free(S); free(Ssum); free(Cvol);
    }
    if (raw) memcpy(raw, tmp, n * sizeof(int16_t));
    orc_median3_i16(tmp, H, W, disp);
    if (p->speckleWindowSize > 0)
        orc_filter_speckles(disp, H, W, e.INVALID, p->speckleWindowSize, 16 * p->speckleRange);
    free(tmp);
    return 0;
}

ORC_API int orc_sgbm_compute(const uint8_t* left, const uint8_t* right, int H, int W,
                             const orc_sgbm_params* p, int16_t* disp,
                             int16_t* Cout, int16_t* Sout, int16_t* raw)
{
    return orc_sgbm_compute_cn(left, right, H, W, 1, p, disp, Cout, Sout, raw);
}

/* stage taps for probing cv2 variants: horizontal window sums hs[H][width1][D] (int32, A.2) ... */
ORC_API int orc_sgbm_hsum(const uint8_t* left, const uint8_t* right, int H, int W, const orc_sgbm_params* p, int32_t* out)
{
    eff_t e = effective(p, W);
    if (e.width1 <= 0) return -1;
    int32_t* hs = hsum_volume(left, right, H, W, 1, &e);
    memcpy(out, hs, (size_t)H * e.width1 * e.D * sizeof(int32_t));
    free(hs);
    return 0;
}

/* ... and everything behind the aggregation: S[H][width1][D] -> final disparity (A.3 - A.5) */
ORC_API int orc_sgbm_finish(const int16_t* S, int H, int W, const orc_sgbm_params* p, int16_t* disp)
{
    eff_t e = effective(p, W);
    size_t n = (size_t)H * W;
    int16_t* tmp = (int16_t*)malloc(n * sizeof(int16_t));
    wta_rows(S, H, W, &e, tmp);
    orc_median3_i16(tmp, H, W, disp);
    if (p->speckleWindowSize > 0)
        orc_filter_speckles(disp, H, W, e.INVALID, p->speckleWindowSize, 16 * p->speckleRange);
    free(tmp);
    return 0;
}

/* ... or only A.3 (winner-take-all, uniqueness, sub-pixel, LR check), before the median */
ORC_API int orc_sgbm_wta(const int16_t* S, int H, int W, const orc_sgbm_params* p, int16_t* raw)
{
    eff_t e = effective(p, W);
    wta_rows(S, H, W, &e, raw);
    return 0;
}

/* one direction's L_r for kernel-level tests; k indexes DIRS */
ORC_API int orc_sgbm_path(const int16_t* Cvol, int H, int w1, int D, int P1, int P2, int k, int16_t* Lout)
{
    if (k < 0 || k >= 8) return -1;
    int32_t* Ssum = (int32_t*)calloc((size_t)H * w1 * D, sizeof(int32_t));
    aggregate_dir(Cvol, H, w1, D, P1, P2, DIRS[k][0], DIRS[k][1], Ssum, Lout);
    free(Ssum);
    return 0;
}

ORC_API void orc_sgbm_geometry(const orc_sgbm_params* p, int W, int* out8)
{
    eff_t e = effective(p, W);
    out8[0] = e.minX1; out8[1] = e.maxX1; out8[2] = e.width1; out8[3] = e.INVALID;
    out8[4] = e.P1; out8[5] = e.P2; out8[6] = e.ftzero; out8[7] = e.uniq;
}

/* ------------------------------------------------------------------ */
/* StereoBM.compute, XSOBEL prefilter, any minDisparity  */
/* ------------------------------------------------------------------ */
static void bm_prefilter(const uint8_t* I, int H, int W, int ftzero, uint8_t* P)
{
    for (int y = 0; y + 1 < H; y += 2) {
        int r0 = y > 0 ? y - 1 : y + 1, r1 = y, r2 = y + 1, r3 = y < H - 2 ? y + 2 : y;
        const uint8_t *a = I + (size_t)r0 * W, *b = I + (size_t)r1 * W, *c = I + (size_t)r2 * W, *d = I + (size_t)r3 * W;
        uint8_t *p0 = P + (size_t)y * W, *p1 = P + (size_t)(y + 1) * W;
        p0[0] = p0[W - 1] = p1[0] = p1[W - 1] = (uint8_t)ftzero;
        for (int x = 1; x < W - 1; x++) {
            int d0 = a[x + 1] - a[x - 1], d1 = b[x + 1] - b[x - 1];
            int d2 = c[x + 1] - c[x - 1], d3 = d[x + 1] - d[x - 1];
            p0[x] = (uint8_t)(iclamp(d0 + 2 * d1 + d2, -ftzero, ftzero) + ftzero);
            p1[x] = (uint8_t)(iclamp(d1 + 2 * d2 + d3, -ftzero, ftzero) + ftzero);
        }
    }
    if (H & 1) memset(P + (size_t)(H - 1) * W, ftzero, W);
}

typedef struct {
    int numDisparities, blockSize, preFilterCap, textureThreshold, uniquenessRatio;
    int minDisparity;     /* 0 in every direct call of the reference; -(D - 1) for the right-view matcher of
                             cv2.ximgproc.createRightMatcher (stereo_vision/stereo_vision.py:171,179) */
} orc_bm_params;

/* General minDisparity (probed against cv2 4.13.0 for negative, zero and positive values, tests/test_oracle_vs_cv2.py):
 *   lofs = max(D-1+minD, 0), rofs = -min(D-1+minD, 0), width1 = W - rofs - D + 1, FILTERED = (minD-1)*16;
 *   output columns X = x + lofs for x in [0, width1), restricted to the valid rectangle
 *   [max(0, minD+D-1) + r, W - r) x [r, H - r);
 *   window column x+k of the left image is clamped to the image, the right image is read at
 *   clamp(x+k, -rofs, W-D-rofs) + rofs + j for the internal index j = D-1-(d-minD): ONE clamp for all disparities, so
 *   near the right border (negative minD) the right window stops moving while the left one goes on (cv2 quirk);
 *   candidates are scanned from the largest disparity down with strict '<'. */
ORC_API int orc_bm_compute(const uint8_t* left, const uint8_t* right, int H, int W,
                           const orc_bm_params* p, int16_t* disp)
{
    const int D = p->numDisparities, r = p->blockSize / 2, ftzero = p->preFilterCap, minD = p->minDisparity;
    const int FILTERED = (minD - 1) * 16;
    size_t n = (size_t)H * W;
    for (size_t i = 0; i < n; i++) disp[i] = (int16_t)FILTERED;
    if (D <= 0 || (D % 16) || !(p->blockSize & 1)) return -1;
    const int lofs = imax(D - 1 + minD, 0), rofs = -imin(D - 1 + minD, 0), width1 = W - rofs - D + 1;
    if (lofs >= W || rofs >= W || width1 < 1) return 0;
    uint8_t* PL = (uint8_t*)malloc(n), *PR = (uint8_t*)malloc(n);
    bm_prefilter(left, H, W, ftzero, PL);
    bm_prefilter(right, H, W, ftzero, PR);
    const int xmin = imax(0, minD + D - 1) + r, xmax = W - r, y0 = r, y1 = H - r;
    if (xmin < xmax && y0 < y1) {
        int* sad = (int*)malloc(D * sizeof(int));
        const int ncol = width1 + 2 * r;                       /* window columns x = -r .. width1-1+r */
        int32_t* col = (int32_t*)malloc((size_t)ncol * D * sizeof(int32_t));
        int32_t* tcol = (int32_t*)malloc((size_t)ncol * sizeof(int32_t));
        for (int y = y0; y < y1; y++) {
            for (int c = 0; c < ncol; c++) {
                const int xs = c - r;
                const int lc = iclamp(xs, -lofs, W - 1 - lofs) + lofs;
                const int rc = iclamp(xs, -rofs, W - D - rofs) + rofs;
                int t = 0;
                for (int yy = y - r; yy <= y + r; yy++) t += abs(PL[(size_t)yy * W + lc] - ftzero);
                tcol[c] = t;
                for (int j = 0; j < D; j++) {
                    int s = 0;
                    for (int yy = y - r; yy <= y + r; yy++)
                        s += abs(PL[(size_t)yy * W + lc] - PR[(size_t)yy * W + rc + j]);
                    col[(size_t)c * D + j] = s;
                }
            }
            for (int x = 0; x < width1; x++) {
                const int X = x + lofs;
                if (X < xmin || X >= xmax) continue;
                int tsum = 0;
                for (int j = 0; j < D; j++) sad[j] = 0;
                for (int c = x; c <= x + 2 * r; c++) {
                    tsum += tcol[c];
                    for (int j = 0; j < D; j++) sad[j] += col[(size_t)c * D + j];
                }
                int best = 0, minsad = sad[0];                  /* internal order: first minimum = largest disparity */
                for (int j = 1; j < D; j++) if (sad[j] < minsad) { minsad = sad[j]; best = j; }
                if (tsum < p->textureThreshold) continue;
                if (p->uniquenessRatio > 0) {
                    int thresh = minsad + minsad * p->uniquenessRatio / 100, j;
                    for (j = 0; j < D; j++) if (abs(j - best) > 1 && sad[j] <= thresh) break;
                    if (j < D) continue;
                }
                int pp = best < D - 1 ? sad[best + 1] : sad[best - 1];
                int nn = best > 0 ? sad[best - 1] : sad[best + 1];
                int dd = pp + nn - 2 * sad[best] + abs(pp - nn);
                disp[(size_t)y * W + X] = (int16_t)(((D - best - 1 + minD) * 256 + (dd ? (pp - nn) * 256 / dd : 0) + 15) >> 4);
            }
        }
        free(tcol); free(col); free(sad);
    }
    free(PR); free(PL);
    return 0;
}

/* ------------------------------------------------------------------ */
/* reprojectImageTo3D.  disp already converted to f32   */
/* by the caller with NO scaling (that is what cv2 does for u8/i16/i32).*/
/* ------------------------------------------------------------------ */
ORC_API void orc_reproject(const float* disp, int H, int W, const double* Q, int handleMissing, float* xyz)
{
    float minv = INFINITY;
    if (handleMissing)
        for (size_t i = 0; i < (size_t)H * W; i++) if (disp[i] < minv) minv = disp[i];
    for (int y = 0; y < H; y++)
        for (int x = 0; x < W; x++) {
            double d = (double)disp[(size_t)y * W + x];
            double P[4];
            for (int k = 0; k < 4; k++)
                /* cv2's Matx product: s = 0; s += q * v for the four terms in order (the +0 start turns a leading -0 into +0) */
                P[k] = 0.0 + Q[4 * k + 0] * x + Q[4 * k + 1] * y + Q[4 * k + 2] * d + Q[4 * k + 3];
            double iw = 1.0 / P[3];
            float* o = xyz + ((size_t)y * W + x) * 3;
            for (int k = 0; k < 3; k++) o[k] = (float)((double)(float)P[k] * iw);
            if (handleMissing && disp[(size_t)y * W + x] == minv) o[2] = 10000.f;
        }
}

/* remap INTER_LINEAR, BORDER_CONSTANT 0, cn interleaved channels */
ORC_API void orc_remap_u8(const uint8_t* src, int sH, int sW, int cn, const float* mapx, const float* mapy,
                          int H, int W, uint8_t* dst)
{
    for (int y = 0; y < H; y++)
        for (int x = 0; x < W; x++) {
            size_t i = (size_t)y * W + x;
            int sx = (int)lrintf(mapx[i] * 32.f), sy = (int)lrintf(mapy[i] * 32.f);
            int ix = iclamp(sx >> 5, -32768, 32767), fx = sx & 31;
            int iy = iclamp(sy >> 5, -32768, 32767), fy = sy & 31;
            int w00 = imin(32767, (32 - fx) * (32 - fy) * 32), w01 = fx * (32 - fy) * 32;
            int w10 = (32 - fx) * fy * 32, w11 = fx * fy * 32;
            for (int c = 0; c < cn; c++) {
#define TAP(yy, xx) (((yy) >= 0 && (yy) < sH && (xx) >= 0 && (xx) < sW) ? src[((size_t)(yy) * sW + (xx)) * cn + c] : 0)
                int v = w00 * TAP(iy, ix) + w01 * TAP(iy, ix + 1) + w10 * TAP(iy + 1, ix) + w11 * TAP(iy + 1, ix + 1);
#undef TAP
                dst[i * cn + c] = (uint8_t)((v + 16384) >> 15);
            }
        }
}

/* cvtColor BGR2GRAY */
ORC_API void orc_bgr2gray(const uint8_t* bgr, size_t npix, uint8_t* gray)
{
    for (size_t i = 0; i < npix; i++)
        gray[i] = (uint8_t)((3735 * bgr[3 * i] + 19235 * bgr[3 * i + 1] + 9798 * bgr[3 * i + 2] + 16384) >> 15);
}
