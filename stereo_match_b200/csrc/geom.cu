// geom.cu -- reprojectImageTo3D, remap(INTER_LINEAR) and BGR2GRAY.
// Reference call sites: stereo_vision/stereo_vision.py:209, disparity_calculation.py:302
// (reproject); rectified_img_cal.py:253-254, disparity_calculation.py:207-208 (remap);
// disparity_calculation.py:286-287 (gray).
#include "common.cuh"

struct QMat { double q[16]; };

template <typename T> __device__ __forceinline__ float to_f32(T v) { return (float)v; }

// A.6:  P = Q * [x, y, d, 1]^T in fp64 ; out = float( double(float(P_c)) * (1.0 / P_w) )
// fp contraction must stay off for bitwise parity with the CPU evaluation order.
template <typename T>
__global__ void __launch_bounds__(256)
k_reproject(const T* __restrict__ disp, int H, int W, QMat Q, float scale, int handle_missing,
            const float* __restrict__ minv, float* __restrict__ xyz)
{
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    const int y = blockIdx.y;
    if (x >= W) return;
    const size_t i = ((size_t)blockIdx.z * H + y) * W + x;
    float df = to_f32(disp[i]);
    if (scale != 1.0f) df = __fmul_rn(df, scale);
    const double d = (double)df;
    double P[4];
#pragma unroll
    for (int k = 0; k < 4; k++) {
        // cv2's Matx product starts its accumulator at +0.0 (s = 0; s += a*b): a leading -0 product (negative or -0 entry
        // of Q at x = 0) becomes +0, which decides the sign of a zero / infinite result
        double a = __dadd_rn(0.0, __dmul_rn(Q.q[4 * k + 0], (double)x));
        a = __dadd_rn(a, __dmul_rn(Q.q[4 * k + 1], (double)y));
        a = __dadd_rn(a, __dmul_rn(Q.q[4 * k + 2], d));
        P[k] = __dadd_rn(a, Q.q[4 * k + 3]);
    }
    const double iw = __ddiv_rn(1.0, P[3]);
    float o0 = (float)__dmul_rn((double)(float)P[0], iw);
    float o1 = (float)__dmul_rn((double)(float)P[1], iw);
    float o2 = (float)__dmul_rn((double)(float)P[2], iw);
    if (handle_missing && df == minv[blockIdx.z]) o2 = 10000.f;
    // (a transposition through shared memory for fully coalesced stores was measured slower: 301 vs 270 us per
    // 128 cfg3 pairs -- L2 already merges the three strided stores)
    float* o = xyz + i * 3;
    o[0] = o0; o[1] = o1; o[2] = o2;
}

// per-image minimum of the (scaled) disparity for handleMissingValues
template <typename T>
__global__ void __launch_bounds__(256)
k_min_f32(const T* __restrict__ disp, size_t n, float scale, float* __restrict__ minv)
{
    const T* p = disp + (size_t)blockIdx.y * n;
    float m = INFINITY;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        float v = to_f32(p[i]);
        if (scale != 1.0f) v = __fmul_rn(v, scale);
        m = fminf(m, v);
    }
    for (int o = 16; o > 0; o >>= 1) m = fminf(m, __shfl_xor_sync(0xffffffffu, m, o));
    if ((threadIdx.x & 31) == 0) {
        // float atomic min through the monotone int mapping (values may be negative)
        int bits = __float_as_int(m);
        if (bits >= 0) atomicMin((int*)&minv[blockIdx.y], bits);
        else atomicMax((unsigned int*)&minv[blockIdx.y], (unsigned int)bits);
    }
}

template <typename T>
static int reproject_t(const T* disp, int H, int W, int batch, const QMat& Q, float scale, int hm,
                       float* scratch, float* xyz, cudaStream_t st)
{
    if (hm) {
        if (!scratch) { b2s_set_error("reproject: handle_missing needs a scratch buffer of batch floats"); return B2S_E_INVALID; }
        // +inf bit pattern as the neutral element
        B2S_CUDA_CHECK(cudaMemsetAsync(scratch, 0x7f, (size_t)batch * sizeof(float), st));  // 0x7f7f7f7f = 3.39e38
        dim3 g(64, batch);
        k_min_f32<T><<<g, 256, 0, st>>>(disp, (size_t)H * W, scale, scratch);
        B2S_LAUNCH_CHECK();
    }
    dim3 grid(b2s_cdiv(W, 256), H, batch);
    k_reproject<T><<<grid, 256, 0, st>>>(disp, H, W, Q, scale, hm, scratch, xyz);
    B2S_LAUNCH_CHECK();
    return B2S_OK;
}

extern "C" int b2s_reproject(const void* disp, int dtype, int H, int W, int batch, const double* Q_host,
                             float disp_scale, int handle_missing, float* scratch, float* xyz, void* stream)
{
    if (!disp || !xyz || !Q_host || H <= 0 || W <= 0 || batch <= 0) {
        b2s_set_error("reproject: null pointer or empty image");
        return B2S_E_INVALID;
    }
    QMat Q;
    for (int i = 0; i < 16; i++) Q.q[i] = Q_host[i];
    cudaStream_t st = (cudaStream_t)stream;
    switch (dtype) {
    case B2S_DT_U8:  return reproject_t((const uint8_t*)disp, H, W, batch, Q, disp_scale, handle_missing, scratch, xyz, st);
    case B2S_DT_I16: return reproject_t((const int16_t*)disp, H, W, batch, Q, disp_scale, handle_missing, scratch, xyz, st);
    case B2S_DT_I32: return reproject_t((const int32_t*)disp, H, W, batch, Q, disp_scale, handle_missing, scratch, xyz, st);
    case B2S_DT_F32: return reproject_t((const float*)disp, H, W, batch, Q, disp_scale, handle_missing, scratch, xyz, st);
    }
    b2s_set_error("reproject: disparity type must be CV_8U, CV_16S, CV_32S or CV_32F");
    return B2S_E_INVALID;
}

// ---- map2Dto3D (reference mapTo3D.py:35-56, called at :116) ----------------------------------------------------
// Pinhole back-projection of a DEPTH image, float64 like the reference's numpy expression:
//   u = -(((col - cx) * z) / fx) ,  v = ((row - cy) * z) / fy ,  z      for the pixels with pt2D > threshold.
// (Not expressible through reprojectImageTo3D: W = 1/z would be needed, which Q * [x, y, z, 1] cannot produce.)
// The reference reshapes the masked result to the image shape, so it only works when every pixel passes; failing
// pixels are written as NaN here and counted, the host wrapper raises numpy's reshape error for them.
__device__ __forceinline__ double load_as_f64(const void* p, int dt, size_t i)
{
    switch (dt) {
    case B2S_DT_U8:  return (double)((const uint8_t*)p)[i];
    case B2S_DT_I16: return (double)((const int16_t*)p)[i];
    case B2S_DT_I32: return (double)((const int32_t*)p)[i];
    case B2S_DT_F32: return (double)((const float*)p)[i];
    case B2S_DT_U16: return (double)((const uint16_t*)p)[i];
    default:         return ((const double*)p)[i];
    }
}

__global__ void __launch_bounds__(256)
k_map2dto3d(const void* __restrict__ pt2d, int dt2d, const void* __restrict__ depth, int dtd, int H, int W,
            double cx, double cy, double fx, double fy, double threshold, double* __restrict__ xyz,
            unsigned long long* __restrict__ npass)
{
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    const int y = blockIdx.y;
    const bool in = x < W;
    bool pass = false;
    if (in) {
        const size_t i = ((size_t)blockIdx.z * H + y) * W + x;
        pass = load_as_f64(pt2d, dt2d, i) > threshold;
        double u = __longlong_as_double(0x7ff8000000000000ll), v = u, z = u;
        if (pass) {
            z = load_as_f64(depth, dtd, i);
            u = -__ddiv_rn(__dmul_rn(__dsub_rn((double)x, cx), z), fx);
            v = __ddiv_rn(__dmul_rn(__dsub_rn((double)y, cy), z), fy);
        }
        double* o = xyz + i * 3;
        o[0] = u; o[1] = v; o[2] = z;
    }
    const unsigned n = __popc(__ballot_sync(0xffffffffu, pass));
    if ((threadIdx.x & 31) == 0 && n) atomicAdd(npass, (unsigned long long)n);
}

extern "C" int b2s_map2dto3d(const void* pt2d, int pt2d_dtype, const void* depth, int depth_dtype, int H, int W, int batch,
                             double cx, double cy, double fx, double fy, double threshold, double* xyz,
                             unsigned long long* npass, void* stream)
{
    if (!pt2d || !depth || !xyz || !npass || H <= 0 || W <= 0 || batch <= 0) {
        b2s_set_error("map2Dto3D: null pointer or empty image");
        return B2S_E_INVALID;
    }
    if (pt2d_dtype < B2S_DT_U8 || pt2d_dtype > B2S_DT_F64 || depth_dtype < B2S_DT_U8 || depth_dtype > B2S_DT_F64) {
        b2s_set_error("map2Dto3D: image types must be u8, u16, i16, i32, f32 or f64");
        return B2S_E_INVALID;
    }
    cudaStream_t st = (cudaStream_t)stream;
    B2S_CUDA_CHECK(cudaMemsetAsync(npass, 0, sizeof(unsigned long long), st));
    dim3 grid(b2s_cdiv(W, 256), H, batch);
    k_map2dto3d<<<grid, 256, 0, st>>>(pt2d, pt2d_dtype, depth, depth_dtype, H, W, cx, cy, fx, fy, threshold, xyz, npass);
    B2S_LAUNCH_CHECK();
    return B2S_OK;
}

// A.7: fixed-point bilinear gather, 5 fractional bits, 15-bit weights, constant-0 border
template <int CN>
__global__ void __launch_bounds__(256)
k_remap(const uint8_t* __restrict__ src, int sH, int sW, const float* __restrict__ mapx,
        const float* __restrict__ mapy, int H, int W, uint8_t* __restrict__ dst)
{
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    const int y = blockIdx.y;
    if (x >= W) return;
    const size_t m = (size_t)y * W + x;
    const int sx = __float2int_rn(__fmul_rn(__ldg(mapx + m), 32.f));
    const int sy = __float2int_rn(__fmul_rn(__ldg(mapy + m), 32.f));
    const int ix = min(max(sx >> 5, -32768), 32767), fx = sx & 31;
    const int iy = min(max(sy >> 5, -32768), 32767), fy = sy & 31;
    const int w00 = min(32767, (32 - fx) * (32 - fy) * 32), w01 = fx * (32 - fy) * 32;
    const int w10 = (32 - fx) * fy * 32, w11 = fx * fy * 32;
    const uint8_t* s = src + (size_t)blockIdx.z * sH * sW * CN;
    const bool x0 = ix >= 0 && ix < sW, x1 = ix + 1 >= 0 && ix + 1 < sW;
    const bool y0 = iy >= 0 && iy < sH, y1 = iy + 1 >= 0 && iy + 1 < sH;
    uint8_t* o = dst + ((size_t)blockIdx.z * H * W + m) * CN;
#pragma unroll
    for (int c = 0; c < CN; c++) {
        int t00 = (y0 && x0) ? s[((size_t)iy * sW + ix) * CN + c] : 0;
        int t01 = (y0 && x1) ? s[((size_t)iy * sW + ix + 1) * CN + c] : 0;
        int t10 = (y1 && x0) ? s[((size_t)(iy + 1) * sW + ix) * CN + c] : 0;
        int t11 = (y1 && x1) ? s[((size_t)(iy + 1) * sW + ix + 1) * CN + c] : 0;
        o[c] = (uint8_t)((w00 * t00 + w01 * t01 + w10 * t10 + w11 * t11 + 16384) >> 15);
    }
}

// one bilinear sample with cv2's fixed-point weights (A.7); s = image base, constant-0 border
__device__ __forceinline__ int remap_sample_c1(const uint8_t* __restrict__ s, int sH, int sW, float mxv, float myv)
{
    const int sx = __float2int_rn(__fmul_rn(mxv, 32.f));
    const int sy = __float2int_rn(__fmul_rn(myv, 32.f));
    const int ix = min(max(sx >> 5, -32768), 32767), fx = sx & 31;
    const int iy = min(max(sy >> 5, -32768), 32767), fy = sy & 31;
    const int w00 = min(32767, (32 - fx) * (32 - fy) * 32), w01 = fx * (32 - fy) * 32;
    const int w10 = (32 - fx) * fy * 32, w11 = fx * fy * 32;
    const bool x0 = ix >= 0 && ix < sW, x1 = ix + 1 >= 0 && ix + 1 < sW;
    const bool y0 = iy >= 0 && iy < sH, y1 = iy + 1 >= 0 && iy + 1 < sH;
    const int t00 = (y0 && x0) ? s[(size_t)iy * sW + ix] : 0;
    const int t01 = (y0 && x1) ? s[(size_t)iy * sW + ix + 1] : 0;
    const int t10 = (y1 && x0) ? s[(size_t)(iy + 1) * sW + ix] : 0;
    const int t11 = (y1 && x1) ? s[(size_t)(iy + 1) * sW + ix + 1] : 0;
    return (w00 * t00 + w01 * t01 + w10 * t10 + w11 * t11 + 16384) >> 15;
}

// gray images, even W: two pixels per thread (64-bit map loads, 16-bit store)
__global__ void __launch_bounds__(256)
k_remap_c1x2(const uint8_t* __restrict__ src, int sH, int sW, const float* __restrict__ mapx,
             const float* __restrict__ mapy, int H, int W, uint8_t* __restrict__ dst)
{
    const int x = 2 * (blockIdx.x * blockDim.x + threadIdx.x);
    const int y = blockIdx.y;
    if (x >= W) return;
    const size_t m = (size_t)y * W + x;
    const float2 mxv = __ldg(reinterpret_cast<const float2*>(mapx + m));
    const float2 myv = __ldg(reinterpret_cast<const float2*>(mapy + m));
    const uint8_t* s = src + (size_t)blockIdx.z * sH * sW;
    const int a = remap_sample_c1(s, sH, sW, mxv.x, myv.x), b = remap_sample_c1(s, sH, sW, mxv.y, myv.y);
    *reinterpret_cast<uint16_t*>(dst + (size_t)blockIdx.z * H * W + m) = (uint16_t)(a | (b << 8));
}

extern "C" int b2s_remap_u8(const uint8_t* src, int sH, int sW, int cn, const float* mapx, const float* mapy,
                            int H, int W, int batch, uint8_t* dst, void* stream)
{
    if (!src || !mapx || !mapy || !dst || sH <= 0 || sW <= 0 || H <= 0 || W <= 0 || batch <= 0) {
        b2s_set_error("remap: null pointer or empty image");
        return B2S_E_INVALID;
    }
    cudaStream_t st = (cudaStream_t)stream;
    dim3 grid(b2s_cdiv(W, 256), H, batch);
    const bool vec2 = (W & 1) == 0 && ((((uintptr_t)mapx | (uintptr_t)mapy) & 7) == 0) && (((uintptr_t)dst & 1) == 0);
    if (cn == 1 && vec2) k_remap_c1x2<<<dim3(b2s_cdiv(W / 2, 256), H, batch), 256, 0, st>>>(src, sH, sW, mapx, mapy, H, W, dst);
    else if (cn == 1) k_remap<1><<<grid, 256, 0, st>>>(src, sH, sW, mapx, mapy, H, W, dst);
    else if (cn == 3) k_remap<3><<<grid, 256, 0, st>>>(src, sH, sW, mapx, mapy, H, W, dst);
    else { b2s_set_error("remap: 1 or 3 channels supported, got %d", cn); return B2S_E_UNSUPPORTED; }
    B2S_LAUNCH_CHECK();
    return B2S_OK;
}

// remap of a BGR image followed by BGR->gray, fused: the three rectified channels stay in registers
// (each rounded to u8 exactly like cv2.remap's output) and only the gray byte is written.
// Replaces cv2.remap (disparity_calculation.py:207-208) + cv2.cvtColor (:286-287) without the round trip
// of the 3-channel rectified image through HBM.
__global__ void __launch_bounds__(256)
k_remap_gray(const uint8_t* __restrict__ src, int sH, int sW, const float* __restrict__ mapx,
             const float* __restrict__ mapy, int H, int W, uint8_t* __restrict__ dst)
{
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    const int y = blockIdx.y;
    if (x >= W) return;
    const size_t m = (size_t)y * W + x;
    const int sx = __float2int_rn(__fmul_rn(__ldg(mapx + m), 32.f));
    const int sy = __float2int_rn(__fmul_rn(__ldg(mapy + m), 32.f));
    const int ix = min(max(sx >> 5, -32768), 32767), fx = sx & 31;
    const int iy = min(max(sy >> 5, -32768), 32767), fy = sy & 31;
    const int w00 = min(32767, (32 - fx) * (32 - fy) * 32), w01 = fx * (32 - fy) * 32;
    const int w10 = (32 - fx) * fy * 32, w11 = fx * fy * 32;
    const uint8_t* s = src + (size_t)blockIdx.z * sH * sW * 3;
    const bool x0 = ix >= 0 && ix < sW, x1 = ix + 1 >= 0 && ix + 1 < sW;
    const bool y0 = iy >= 0 && iy < sH, y1 = iy + 1 >= 0 && iy + 1 < sH;
    int ch[3];
#pragma unroll
    for (int c = 0; c < 3; c++) {
        int t00 = (y0 && x0) ? s[((size_t)iy * sW + ix) * 3 + c] : 0;
        int t01 = (y0 && x1) ? s[((size_t)iy * sW + ix + 1) * 3 + c] : 0;
        int t10 = (y1 && x0) ? s[((size_t)(iy + 1) * sW + ix) * 3 + c] : 0;
        int t11 = (y1 && x1) ? s[((size_t)(iy + 1) * sW + ix + 1) * 3 + c] : 0;
        ch[c] = (w00 * t00 + w01 * t01 + w10 * t10 + w11 * t11 + 16384) >> 15;
    }
    dst[(size_t)blockIdx.z * H * W + m] = (uint8_t)((3735 * ch[0] + 19235 * ch[1] + 9798 * ch[2] + 16384) >> 15);
}

extern "C" int b2s_remap_gray_u8(const uint8_t* src_bgr, int sH, int sW, const float* mapx, const float* mapy,
                                 int H, int W, int batch, uint8_t* dst_gray, void* stream)
{
    if (!src_bgr || !mapx || !mapy || !dst_gray || sH <= 0 || sW <= 0 || H <= 0 || W <= 0 || batch <= 0) {
        b2s_set_error("remap_gray: null pointer or empty image");
        return B2S_E_INVALID;
    }
    dim3 grid(b2s_cdiv(W, 256), H, batch);
    k_remap_gray<<<grid, 256, 0, (cudaStream_t)stream>>>(src_bgr, sH, sW, mapx, mapy, H, W, dst_gray);
    B2S_LAUNCH_CHECK();
    return B2S_OK;
}

// A.8
__global__ void __launch_bounds__(256)
k_bgr2gray(const uint8_t* __restrict__ bgr, size_t n, uint8_t* __restrict__ gray)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint8_t* p = bgr + 3 * i;
    gray[i] = (uint8_t)((3735 * p[0] + 19235 * p[1] + 9798 * p[2] + 16384) >> 15);
}

extern "C" int b2s_bgr2gray(const uint8_t* bgr, size_t npixels, uint8_t* gray, void* stream)
{
    if (!bgr || !gray) { b2s_set_error("bgr2gray: null pointer"); return B2S_E_INVALID; }
    if (npixels == 0) return B2S_OK;
    k_bgr2gray<<<b2s_cdiv((long long)npixels, 256), 256, 0, (cudaStream_t)stream>>>(bgr, npixels, gray);
    B2S_LAUNCH_CHECK();
    return B2S_OK;
}
