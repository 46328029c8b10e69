/*
 * wls_oracle.c -- CPU ORACLE (test infrastructure, NOT product code) for the disparity WLS filter that the
 * reference builds at stereo_vision/stereo_vision.py:171-182 (cv2.ximgproc.createDisparityWLSFilter(left_matcher),
 * setLambda / setSigmaColor, filter(displ, gray_l, None, dispr)) and, without confidence, at
 * disparity_test.py:182-199 (createDisparityWLSFilterGeneric(False)).
 *
 * PARITY UNPINNED.  The arithmetic lives in opencv_contrib (module ximgproc: disparity_filters.cpp, fgs_filter.cpp),
 * which is absent from /root/reference, from this image and from the GPU box (cv2.ximgproc does not exist in
 * opencv-python-headless 4.13.0.92), and the reference holds no fixture for it.  This file restates the published
 * algorithm -- Min et al., "Fast Global Image Smoothing Based on Weighted Least Squares" (TIP 2014) as used by
 * DisparityWLSFilter: a left-right-consistency / depth-discontinuity confidence map, then
 *      out = FGS(confidence * disparity) / (FGS(confidence) + eps)
 * inside the valid-disparity rectangle -- with the parameter rules of createDisparityWLSFilter as recalled from the
 * opencv_contrib 4.x sources.  Every floating-point operation is written out in a fixed order (float32, no
 * contraction), so that the CUDA implementation (stereo_match_b200/csrc/wls.cu) can be compared with it bit for bit;
 * that comparison proves "CUDA == this restatement", not "== cv2.ximgproc".
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>

#define ORC_API __attribute__((visibility("default")))
#define WLS_EPS 1e-43f

typedef struct {
    int roi_x, roi_y, roi_w, roi_h;     /* valid-disparity rectangle of the left view */
    int min_disp;                       /* output outside the rectangle: 16 * (min_disp - 1) */
    int use_confidence;
    int lrc_thresh;                     /* 24 */
    int radius;                         /* depth-discontinuity radius */
    float roll_off;                     /* 0.001 */
    float lambda, sigma;
    int num_iter;                       /* 3 */
    float attenuation;                  /* 0.25 */
} orc_wls_params;

static int reflect101(int p, int n)
{
    if (n == 1) return 0;
    while (p < 0 || p >= n) p = p < 0 ? -p : 2 * n - 2 - p;
    return p;
}

/* guide weights of the smoother: -exp(-sqrt(sum of squared channel differences) / sigma) */
ORC_API void orc_wls_lut(float sigma, int n, float* lut)
{
    for (int i = 0; i < n; i++) lut[i] = -expf(-sqrtf((float)i) / sigma);
}

/* depth-discontinuity map of an int16 disparity ROI copy: max(0, 1 - roll_off * variance over a (2r+1)^2 window),
 * window clipped by BORDER_REFLECT_101 at the borders of the rectangle.  Sums in float, row-major window order. */
static void disc_map(const int16_t* disp, int stride, int w, int h, int radius, float roll_off, float* out)
{
    const int k = 2 * radius + 1;
    const float inv = 1.0f / (float)(k * k);
    for (int y = 0; y < h; y++)
        for (int x = 0; x < w; x++) {
            float s = 0.f, s2 = 0.f;
            for (int dy = -radius; dy <= radius; dy++) {
                const int16_t* row = disp + (size_t)reflect101(y + dy, h) * stride;
                for (int dx = -radius; dx <= radius; dx++) {
                    const float v = (float)row[reflect101(x + dx, w)];
                    s += v;
                    s2 += v * v;
                }
            }
            const float mean = s * inv, mean2 = s2 * inv;
            const float var = mean2 - mean * mean;
            const float c = 1.0f - roll_off * var;
            out[(size_t)y * w + x] = c > 0.f ? c : 0.f;
        }
}

/* one tridiagonal solve (I + lambda * A) u = f along a line of n elements `step` floats apart, two right-hand sides
 * sharing the weights: w[j] = guide weight between element j and j+1 (negative; w[n-1] ignored).  ct: n scratch floats */
static void solve_line(const float* w, int wstep, float lam, float* f1, float* f2, int step, int n, float* ct)
{
    float cprev = 0.f, p1 = 0.f, p2 = 0.f;
    for (int j = 0; j < n; j++) {
        const float a = j > 0 ? lam * w[(size_t)(j - 1) * wstep] : 0.f;
        const float c = j < n - 1 ? lam * w[(size_t)j * wstep] : 0.f;
        const float denom = ((1.0f - a) - c) - a * cprev;
        cprev = c / denom;
        ct[j] = cprev;
        p1 = (f1[(size_t)j * step] - a * p1) / denom;
        p2 = (f2[(size_t)j * step] - a * p2) / denom;
        f1[(size_t)j * step] = p1;
        f2[(size_t)j * step] = p2;
    }
    for (int j = n - 2; j >= 0; j--) {
        p1 = f1[(size_t)j * step] - ct[j] * p1;
        p2 = f2[(size_t)j * step] - ct[j] * p2;
        f1[(size_t)j * step] = p1;
        f2[(size_t)j * step] = p2;
    }
}

static int16_t round_sat16(float v)
{
    const float r = nearbyintf(v);        /* round half to even, like cvRound */
    if (!(r > -32768.f)) return r != r ? 0 : -32768;
    if (r > 32767.f) return 32767;
    return (int16_t)r;
}

/* dl, dr: [H][W] int16 left / right-view disparity (dr may be NULL without confidence); guide: [H][W][cn] uint8.
 * out: [H][W] int16; conf_out (optional): [H][W] float, the confidence map (0..255), zero outside the rectangle. */
ORC_API int orc_wls_filter(const int16_t* dl, const int16_t* dr, const uint8_t* guide, int cn, int H, int W,
                           const orc_wls_params* p, int16_t* out, float* conf_out)
{
    const int rx = p->roi_x, ry = p->roi_y, rw = p->roi_w, rh = p->roi_h;
    for (size_t i = 0; i < (size_t)H * W; i++) out[i] = (int16_t)(16 * (p->min_disp - 1));
    if (conf_out) memset(conf_out, 0, (size_t)H * W * sizeof(float));
    if (rw <= 0 || rh <= 0) return 0;
    if (rx < 0 || ry < 0 || rx + rw > W || ry + rh > H || (cn != 1 && cn != 3)) return 1;
    if (p->use_confidence && !dr) return 1;
    const size_t n = (size_t)rw * rh;
    float* conf = (float*)malloc(n * sizeof(float));
    float* f1 = (float*)malloc(n * sizeof(float));
    float* f2 = (float*)malloc(n * sizeof(float));
    float* chor = (float*)malloc(n * sizeof(float));
    float* cvert = (float*)malloc(n * sizeof(float));
    float* ct = (float*)malloc((size_t)(rw > rh ? rw : rh) * sizeof(float));
    const int nlut = cn * 255 * 255 + 1;
    float* lut = (float*)malloc((size_t)nlut * sizeof(float));
    orc_wls_lut(p->sigma, nlut, lut);

    /* ---- confidence ------------------------------------------------------------------------------------ */
    if (p->use_confidence) {
        /* rectangle of the right view: mirrored horizontally */
        const int qx = W - (rx + rw);
        float* dright = (float*)malloc(n * sizeof(float));
        disc_map(dl + (size_t)ry * W + rx, W, rw, rh, p->radius, p->roll_off, conf);
        disc_map(dr + (size_t)ry * W + qx, W, rw, rh, p->radius, p->roll_off, dright);
        for (int y = 0; y < rh; y++) {
            const int16_t* L = dl + (size_t)(ry + y) * W;
            const int16_t* R = dr + (size_t)(ry + y) * W;
            for (int x = rx; x < rx + rw; x++) {
                const int ridx = x - (L[x] >> 4);
                float c = conf[(size_t)y * rw + (x - rx)];
                if (ridx >= qx && ridx < qx + rw) {
                    const int s = (int)L[x] + (int)R[ridx];
                    if ((s < 0 ? -s : s) < p->lrc_thresh) {
                        const float cr = dright[(size_t)y * rw + (ridx - qx)];
                        c = c < cr ? c : cr;
                    } else
                        c = 0.f;
                }
                conf[(size_t)y * rw + (x - rx)] = 255.0f * c;
            }
        }
        free(dright);
    } else {
        for (size_t i = 0; i < n; i++) conf[i] = 1.0f;
    }

    /* ---- guide weights inside the rectangle ------------------------------------------------------------ */
    for (int y = 0; y < rh; y++)
        for (int x = 0; x < rw; x++) {
            const uint8_t* g = guide + ((size_t)(ry + y) * W + rx + x) * cn;
            int dh = 0, dv = 0;
            for (int c = 0; c < cn; c++) {
                if (x < rw - 1) { const int d = (int)g[c] - (int)g[cn + c]; dh += d * d; }
                if (y < rh - 1) { const int d = (int)g[c] - (int)g[(size_t)W * cn + c]; dv += d * d; }
            }
            chor[(size_t)y * rw + x] = x < rw - 1 ? lut[dh] : 0.f;
            cvert[(size_t)y * rw + x] = y < rh - 1 ? lut[dv] : 0.f;
        }

    /* ---- the two smoothed images ----------------------------------------------------------------------- */
    for (int y = 0; y < rh; y++)
        for (int x = 0; x < rw; x++) {
            const float d = (float)dl[(size_t)(ry + y) * W + rx + x], c = conf[(size_t)y * rw + x];
            f1[(size_t)y * rw + x] = p->use_confidence ? c * d : d;
            f2[(size_t)y * rw + x] = c;
        }
    float lam = p->lambda;
    for (int it = 0; it < p->num_iter; it++) {
        for (int y = 0; y < rh; y++)
            solve_line(chor + (size_t)y * rw, 1, lam, f1 + (size_t)y * rw, f2 + (size_t)y * rw, 1, rw, ct);
        for (int x = 0; x < rw; x++)
            solve_line(cvert + x, rw, lam, f1 + x, f2 + x, rw, rh, ct);
        lam *= p->attenuation;
    }
    for (int y = 0; y < rh; y++)
        for (int x = 0; x < rw; x++) {
            const size_t i = (size_t)y * rw + x;
            const float v = p->use_confidence ? f1[i] * (1.0f / (f2[i] + WLS_EPS)) : f1[i];
            out[(size_t)(ry + y) * W + rx + x] = round_sat16(v);
            if (conf_out) conf_out[(size_t)(ry + y) * W + rx + x] = conf[i];
        }
    free(conf); free(f1); free(f2); free(chor); free(cvert); free(ct); free(lut);
    return 0;
}
