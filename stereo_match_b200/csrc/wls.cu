// wls.cu -- the disparity WLS filter of the reference's compute_disparity (stereo_vision/stereo_vision.py:171-182:
// cv2.ximgproc.createDisparityWLSFilter(left_matcher), setLambda, setSigmaColor, filter(displ, gray_l, None, dispr);
// without confidence: disparity_test.py:182-199, createDisparityWLSFilterGeneric(False)).  SURVEY.md 8(f)-3.
//
// PARITY UNPINNED: cv2.ximgproc exists neither in this image nor on the GPU box and the reference holds no fixture.
// The kernels follow the published algorithm (fast global smoother of Min et al. 2014 with the confidence weighting
// of opencv_contrib's DisparityWLSFilter, restated in oracle/wls_oracle.c) and are bit-exact with THAT restatement:
// float32, every operation in the restatement's order, no contraction (the library is built with --fmad=false), the
// guide-weight table computed on the host with the same libm call.
//
//   confidence  c = 255 * LRC-gated min of the two depth-discontinuity maps (max(0, 1 - roll_off * local variance))
//   out         = FGS(c * d) / (FGS(c) + eps)   inside the valid-disparity rectangle, 16 * (minDisparity - 1) outside
//   FGS         = num_iter x (tridiagonal solve along every row, then along every column), lambda *= 0.25 per iteration
//
// GPU mapping: one thread per tridiagonal system, systems interleaved in memory so that every access of a warp is one
// contiguous line: the column solves run on the [rows][cols] arrays, the row solves on transposed copies (a tiled
// shared-memory transpose in between).  Both right-hand sides (c*d and c) share the weights and the elimination
// coefficients, so they go through the same sweep.
#include "common.cuh"
#include <cmath>
#include <cstdlib>
#include <new>
#include <vector>

struct b2s_wls {
    float* lut = nullptr;        // device: -exp(-sqrt(i) / sigma), i = sum of squared channel differences
    size_t lut_n = 0;
    float lut_sigma = -1.f;
    void* ws = nullptr;
    size_t ws_bytes = 0;
    cudaStream_t last_stream = nullptr;
};

__device__ __forceinline__ int wls_reflect101(int p, int n)
{
    if (n == 1) return 0;
    while (p < 0 || p >= n) p = p < 0 ? -p : 2 * n - 2 - p;
    return p;
}

// depth-discontinuity map of the rectangle (x0, y0, rw, rh) of an int16 disparity image: max(0, 1 - roll_off * var)
__global__ void __launch_bounds__(256)
k_wls_disc(const int16_t* __restrict__ disp, int H, int W, int x0, int y0, int rw, int rh, int radius, float roll_off,
           float* __restrict__ out)
{
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
    if (x >= rw) return;
    const int16_t* img = disp + (size_t)blockIdx.z * H * W + (size_t)y0 * W + x0;
    const int k = 2 * radius + 1;
    const float inv = 1.0f / (float)(k * k);
    float s = 0.f, s2 = 0.f;
    for (int dy = -radius; dy <= radius; dy++) {
        const int16_t* row = img + (size_t)wls_reflect101(y + dy, rh) * W;
        for (int dx = -radius; dx <= radius; dx++) {
            const float v = (float)row[wls_reflect101(x + dx, rw)];
            s += v;
            s2 += v * v;
        }
    }
    const float mean = s * inv, mean2 = s2 * inv;
    const float var = mean2 - mean * mean;
    const float c = 1.0f - roll_off * var;
    out[((size_t)blockIdx.z * rh + y) * rw + x] = c > 0.f ? c : 0.f;
}

// confidence (in place over the left discontinuity map), the two right-hand sides and the guide weights
template <int CN>
__global__ void __launch_bounds__(256)
k_wls_prepare(const int16_t* __restrict__ dl, const int16_t* __restrict__ dr, const uint8_t* __restrict__ guide, int H, int W,
              b2s_wls_params p, const float* __restrict__ lut, float* __restrict__ conf, const float* __restrict__ dright,
              float* __restrict__ f1, float* __restrict__ f2, float* __restrict__ chor, float* __restrict__ cvert)
{
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
    const int rw = p.roi_w, rh = p.roi_h;
    if (x >= rw) return;
    const size_t img = (size_t)blockIdx.z * H * W, i = ((size_t)blockIdx.z * rh + y) * rw + x;
    const int gx = p.roi_x + x, gy = p.roi_y + y;
    const int16_t d16 = dl[img + (size_t)gy * W + gx];
    float c = 1.0f;
    if (p.use_confidence) {
        const int qx = W - (p.roi_x + rw);                       // rectangle of the right view: mirrored
        const int ridx = gx - (d16 >> 4);
        c = conf[i];
        if (ridx >= qx && ridx < qx + rw) {
            const int s = (int)d16 + (int)dr[img + (size_t)gy * W + ridx];
            if ((s < 0 ? -s : s) < p.lrc_thresh) {
                const float cr = dright[((size_t)blockIdx.z * rh + y) * rw + (ridx - qx)];
                c = c < cr ? c : cr;
            } else
                c = 0.f;
        }
        c = 255.0f * c;
    }
    conf[i] = c;
    const float d = (float)d16;
    f1[i] = p.use_confidence ? c * d : d;
    f2[i] = c;
    const uint8_t* g = guide + (img + (size_t)gy * W + gx) * CN;
    int dh = 0, dv = 0;
#pragma unroll
    for (int ch = 0; ch < CN; ch++) {
        if (x < rw - 1) { const int t = (int)g[ch] - (int)g[CN + ch]; dh += t * t; }
        if (y < rh - 1) { const int t = (int)g[ch] - (int)g[(size_t)W * CN + ch]; dv += t * t; }
    }
    chor[i] = x < rw - 1 ? lut[dh] : 0.f;
    cvert[i] = y < rh - 1 ? lut[dv] : 0.f;
}

// [n][m] -> [m][n] per image, one or two arrays
__global__ void __launch_bounds__(256)
k_wls_transpose(const float* __restrict__ a, const float* __restrict__ b, int n, int m, float* __restrict__ at, float* __restrict__ bt)
{
    __shared__ float ta[32][33], tb[32][33];
    const size_t img = (size_t)blockIdx.z * n * m;
    const int c0 = blockIdx.x * 32, r0 = blockIdx.y * 32, tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    for (int r = ty; r < 32; r += 8)
        if (r0 + r < n && c0 + tx < m) {
            ta[r][tx] = a[img + (size_t)(r0 + r) * m + c0 + tx];
            if (b) tb[r][tx] = b[img + (size_t)(r0 + r) * m + c0 + tx];
        }
    __syncthreads();
    for (int r = ty; r < 32; r += 8)
        if (c0 + r < m && r0 + tx < n) {
            at[img + (size_t)(c0 + r) * n + r0 + tx] = ta[tx][r];
            if (b) bt[img + (size_t)(c0 + r) * n + r0 + tx] = tb[tx][r];
        }
}

// m interleaved tridiagonal systems of n unknowns: element j of system s at [j * m + s]  (oracle: solve_line)
// The recurrence is serial in j; what made it slow was one round trip to memory per step.  The operands of WLS_K steps are
// loaded together (and the next chunk's while this one is being eliminated), the arithmetic keeps the restatement's order.
#define WLS_K 8
__global__ void __launch_bounds__(128)
k_wls_solve(const float* __restrict__ w, float lam, float* __restrict__ f1, float* __restrict__ f2, float* __restrict__ ct, int n, int m)
{
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= m) return;
    const size_t img = (size_t)blockIdx.y * n * m;
    w += img + s; f1 += img + s; f2 += img + s; ct += img + s;
    float cprev = 0.f, p1 = 0.f, p2 = 0.f, wprev = 0.f;
    float wv[WLS_K], r1[WLS_K], r2[WLS_K], nw[WLS_K], n1[WLS_K], n2[WLS_K];
#pragma unroll
    for (int k = 0; k < WLS_K; k++) {
        const bool in = k < n;
        const size_t o = (size_t)(in ? k : 0) * m;
        nw[k] = in ? w[o] : 0.f; n1[k] = in ? f1[o] : 0.f; n2[k] = in ? f2[o] : 0.f;
    }
    for (int j0 = 0; j0 < n; j0 += WLS_K) {
#pragma unroll
        for (int k = 0; k < WLS_K; k++) { wv[k] = nw[k]; r1[k] = n1[k]; r2[k] = n2[k]; }
#pragma unroll
        for (int k = 0; k < WLS_K; k++) {                         // the next chunk, in flight during the elimination
            const int j = j0 + WLS_K + k;
            const bool in = j < n;
            const size_t o = (size_t)(in ? j : 0) * m;
            nw[k] = in ? w[o] : 0.f; n1[k] = in ? f1[o] : 0.f; n2[k] = in ? f2[o] : 0.f;
        }
#pragma unroll
        for (int k = 0; k < WLS_K; k++) {
            const int j = j0 + k;
            if (j < n) {
                const size_t o = (size_t)j * m;
                const float wj = wv[k];
                const float a = j > 0 ? lam * wprev : 0.f;
                const float c = j < n - 1 ? lam * wj : 0.f;
                const float denom = ((1.0f - a) - c) - a * cprev;
                cprev = c / denom;
                ct[o] = cprev;
                p1 = (r1[k] - a * p1) / denom;
                p2 = (r2[k] - a * p2) / denom;
                f1[o] = p1;
                f2[o] = p2;
                wprev = wj;
            }
        }
    }
    // back substitution from j = n - 2 down, chunks of WLS_K loaded ahead the same way
    int jt = n - 2;
#pragma unroll
    for (int k = 0; k < WLS_K; k++) {
        const int j = jt - k;
        const bool in = j >= 0;
        const size_t o = (size_t)(in ? j : 0) * m;
        nw[k] = in ? ct[o] : 0.f; n1[k] = in ? f1[o] : 0.f; n2[k] = in ? f2[o] : 0.f;
    }
    for (; jt >= 0; jt -= WLS_K) {
#pragma unroll
        for (int k = 0; k < WLS_K; k++) { wv[k] = nw[k]; r1[k] = n1[k]; r2[k] = n2[k]; }
#pragma unroll
        for (int k = 0; k < WLS_K; k++) {
            const int j = jt - WLS_K - k;
            const bool in = j >= 0;
            const size_t o = (size_t)(in ? j : 0) * m;
            nw[k] = in ? ct[o] : 0.f; n1[k] = in ? f1[o] : 0.f; n2[k] = in ? f2[o] : 0.f;
        }
#pragma unroll
        for (int k = 0; k < WLS_K; k++) {
            const int j = jt - k;
            if (j >= 0) {
                const size_t o = (size_t)j * m;
                p1 = r1[k] - wv[k] * p1;
                p2 = r2[k] - wv[k] * p2;
                f1[o] = p1;
                f2[o] = p2;
            }
        }
    }
}

__global__ void __launch_bounds__(256)
k_wls_finish(const float* __restrict__ f1, const float* __restrict__ f2, const float* __restrict__ conf, int H, int W,
             b2s_wls_params p, int16_t* __restrict__ out, float* __restrict__ conf_out)
{
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
    if (x >= W) return;
    const size_t o = ((size_t)blockIdx.z * H + y) * W + x;
    const int rx = x - p.roi_x, ry = y - p.roi_y;
    int16_t v = (int16_t)(16 * (p.min_disp - 1));
    float c = 0.f;
    if (rx >= 0 && rx < p.roi_w && ry >= 0 && ry < p.roi_h) {
        const size_t i = ((size_t)blockIdx.z * p.roi_h + ry) * p.roi_w + rx;
        const float q = p.use_confidence ? f1[i] * (1.0f / (f2[i] + 1e-43f)) : f1[i];
        const float r = rintf(q);                                  // round half to even, like cvRound
        v = !(r > -32768.f) ? (r != r ? (int16_t)0 : (int16_t)-32768) : (r > 32767.f ? (int16_t)32767 : (int16_t)r);
        c = conf[i];
    }
    out[o] = v;
    if (conf_out) conf_out[o] = c;
}

__global__ void k_wls_fill(int16_t* out, size_t n, int16_t v)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = v;
}

static size_t wls_al(size_t v) { return (v + 255) / 256 * 256; }

extern "C" int b2s_wls_create(b2s_wls_handle* out)
{
    if (!out) { b2s_set_error("wls_create: null argument"); return B2S_E_INVALID; }
    b2s_wls* h = new (std::nothrow) b2s_wls();
    if (!h) { b2s_set_error("out of host memory"); return B2S_E_NOMEM; }
    *out = h;
    return B2S_OK;
}

extern "C" void b2s_wls_destroy(b2s_wls_handle h)
{
    if (!h) return;
    if (h->lut) cudaFree(h->lut);
    if (h->ws) cudaFree(h->ws);
    delete h;
}

extern "C" int b2s_wls_filter(b2s_wls_handle h, const int16_t* disp_left, const int16_t* disp_right, const uint8_t* guide,
                              int cn, int H, int W, int batch, const b2s_wls_params* params, int16_t* out, float* conf_out,
                              void* stream)
{
    if (!h || !params) { b2s_set_error("wls_filter: bad handle"); return B2S_E_INVALID; }
    if (!disp_left || !guide || !out || H <= 0 || W <= 0 || batch <= 0) { b2s_set_error("wls_filter: null pointer or empty image"); return B2S_E_INVALID; }
    if (cn != 1 && cn != 3) { b2s_set_error("DisparityWLSFilter: left_view must have 1 or 3 channels (got %d)", cn); return B2S_E_INVALID; }
    const b2s_wls_params p = *params;
    if (p.use_confidence && !disp_right) { b2s_set_error("DisparityWLSFilter: the confidence map needs disparity_map_right"); return B2S_E_INVALID; }
    if (!(p.lambda >= 0.f) || !(p.sigma > 0.f) || p.num_iter < 1 || p.radius < 0) { b2s_set_error("DisparityWLSFilter: lambda >= 0, sigma > 0, num_iter >= 1, radius >= 0 expected"); return B2S_E_INVALID; }
    if (p.roi_w < 0 || p.roi_h < 0 || p.roi_x < 0 || p.roi_y < 0 || p.roi_x + p.roi_w > W || p.roi_y + p.roi_h > H) {
        b2s_set_error("DisparityWLSFilter: the valid-disparity rectangle (%d, %d, %d, %d) leaves the %dx%d image", p.roi_x, p.roi_y, p.roi_w, p.roi_h, W, H);
        return B2S_E_INVALID;
    }
    cudaStream_t st = (cudaStream_t)stream;
    const size_t npix = (size_t)batch * H * W;
    const int rw = p.roi_w, rh = p.roi_h;
    if (rw == 0 || rh == 0) {
        k_wls_fill<<<b2s_cdiv((long long)npix, 256), 256, 0, st>>>(out, npix, (int16_t)(16 * (p.min_disp - 1)));
        B2S_LAUNCH_CHECK();
        if (conf_out) B2S_CUDA_CHECK(cudaMemsetAsync(conf_out, 0, npix * sizeof(float), st));
        return B2S_OK;
    }
    // guide-weight table (host libm, same call as the restatement), cached per sigma
    const size_t nlut = 3 * 255 * 255 + 1;
    if (!h->lut || h->lut_sigma != p.sigma) {
        if (!h->lut) B2S_CUDA_CHECK(cudaMalloc(&h->lut, nlut * sizeof(float)));
        std::vector<float> host(nlut);
        for (size_t i = 0; i < nlut; i++) host[i] = -expf(-sqrtf((float)i) / p.sigma);
        B2S_CUDA_CHECK(cudaMemcpyAsync(h->lut, host.data(), nlut * sizeof(float), cudaMemcpyHostToDevice, st));
        B2S_CUDA_CHECK(cudaStreamSynchronize(st));               // `host` goes out of scope
        h->lut_n = nlut; h->lut_sigma = p.sigma;
    }
    // workspace: conf, dright, f1, f2, f1t, f2t, chor, chort, cvert, ct   ([batch][rh][rw] floats each)
    const size_t plane = wls_al((size_t)batch * rw * rh * sizeof(float));
    const size_t need = 10 * plane;
    if (need > h->ws_bytes) {
        if (h->ws) {
            B2S_CUDA_CHECK(cudaStreamSynchronize(st));
            if (h->last_stream && h->last_stream != st) B2S_CUDA_CHECK(cudaStreamSynchronize(h->last_stream));
            cudaFree(h->ws); h->ws = nullptr; h->ws_bytes = 0;
        }
        cudaError_t err = cudaMalloc(&h->ws, need);
        if (err != cudaSuccess) {
            cudaGetLastError();
            b2s_set_error("wls_filter: workspace allocation of %zu bytes failed: %s", need, cudaGetErrorString(err));
            return B2S_E_NOMEM;
        }
        h->ws_bytes = need;
    }
    h->last_stream = st;
    { int prc = b2s_poison(h->ws, h->ws_bytes, st); if (prc) return prc; }     // debug aid, see api.cu
    char* ws = (char*)h->ws;
    float* conf = (float*)(ws + 0 * plane); float* dright = (float*)(ws + 1 * plane);
    float* f1 = (float*)(ws + 2 * plane);   float* f2 = (float*)(ws + 3 * plane);
    float* f1t = (float*)(ws + 4 * plane);  float* f2t = (float*)(ws + 5 * plane);
    float* chor = (float*)(ws + 6 * plane); float* chort = (float*)(ws + 7 * plane);
    float* cvert = (float*)(ws + 8 * plane); float* ct = (float*)(ws + 9 * plane);

    const dim3 gpix(b2s_cdiv(rw, 256), rh, batch);
    if (p.use_confidence) {
        k_wls_disc<<<gpix, 256, 0, st>>>(disp_left, H, W, p.roi_x, p.roi_y, rw, rh, p.radius, p.roll_off, conf);
        B2S_LAUNCH_CHECK();
        k_wls_disc<<<gpix, 256, 0, st>>>(disp_right, H, W, W - (p.roi_x + rw), p.roi_y, rw, rh, p.radius, p.roll_off, dright);
        B2S_LAUNCH_CHECK();
    }
    if (cn == 1) k_wls_prepare<1><<<gpix, 256, 0, st>>>(disp_left, disp_right, guide, H, W, p, h->lut, conf, dright, f1, f2, chor, cvert);
    else k_wls_prepare<3><<<gpix, 256, 0, st>>>(disp_left, disp_right, guide, H, W, p, h->lut, conf, dright, f1, f2, chor, cvert);
    B2S_LAUNCH_CHECK();
    const dim3 gt(b2s_cdiv(rw, 32), b2s_cdiv(rh, 32), batch), gtt(b2s_cdiv(rh, 32), b2s_cdiv(rw, 32), batch);
    k_wls_transpose<<<gt, 256, 0, st>>>(chor, nullptr, rh, rw, chort, nullptr);
    B2S_LAUNCH_CHECK();
    float lam = p.lambda;
    for (int it = 0; it < p.num_iter; it++) {
        k_wls_transpose<<<gt, 256, 0, st>>>(f1, f2, rh, rw, f1t, f2t);
        B2S_LAUNCH_CHECK();
        k_wls_solve<<<dim3(b2s_cdiv(rh, 128), batch), 128, 0, st>>>(chort, lam, f1t, f2t, ct, rw, rh);      // along the rows
        B2S_LAUNCH_CHECK();
        k_wls_transpose<<<gtt, 256, 0, st>>>(f1t, f2t, rw, rh, f1, f2);
        B2S_LAUNCH_CHECK();
        k_wls_solve<<<dim3(b2s_cdiv(rw, 128), batch), 128, 0, st>>>(cvert, lam, f1, f2, ct, rh, rw);        // along the columns
        B2S_LAUNCH_CHECK();
        lam *= p.attenuation;
    }
    k_wls_finish<<<dim3(b2s_cdiv(W, 256), H, batch), 256, 0, st>>>(f1, f2, conf, H, W, p, out, conf_out);
    B2S_LAUNCH_CHECK();
    return B2S_OK;
}
