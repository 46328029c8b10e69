// cloud.cu -- point-cloud extraction behind reprojectImageTo3D.
// Reference: disparity_calculation.py:309-321 (mask = (disp < hi) & (disp > lo); points_3d[mask]; colours of the
// rectified BGR image as RGB; NaN / inf coordinates -> 0) and try_try.py:91-94 (mask = disp > disp.min()).
// The boolean-mask indexing of numpy keeps row-major order, so this is an ORDERED stream compaction:
//   k_cloud_count : kept pixels per block of 1024 pixels (blocks never straddle two images)
//   k_cloud_scan  : exclusive scan of the block counts (one CTA), per-image offsets
//   k_cloud_write : every block redoes its mask, ranks its pixels with a block-wide scan and writes its vertices
#include "common.cuh"

#define CLOUD_NT 256
#define CLOUD_PPT 4                                   // consecutive pixels per thread
#define CLOUD_PPB (CLOUD_NT * CLOUD_PPT)

template <typename T>
__device__ __forceinline__ bool cloud_keep(const T* __restrict__ disp, size_t i, double lo, double hi)
{
    const double v = (double)disp[i];
    return v > lo && v < hi;
}

template <typename T>
__global__ void __launch_bounds__(CLOUD_NT)
k_cloud_count(const T* __restrict__ disp, int npix, int nblk, double lo, double hi, int* __restrict__ blockcount)
{
    const int b = blockIdx.y;
    const int p0 = blockIdx.x * CLOUD_PPB + threadIdx.x * CLOUD_PPT;
    int n = 0;
#pragma unroll
    for (int j = 0; j < CLOUD_PPT; j++)
        if (p0 + j < npix && cloud_keep(disp, (size_t)b * npix + p0 + j, lo, hi)) n++;
    __shared__ int wsum[CLOUD_NT / 32];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) n += __shfl_xor_sync(0xffffffffu, n, o);
    if ((threadIdx.x & 31) == 0) wsum[threadIdx.x >> 5] = n;
    __syncthreads();
    if (threadIdx.x == 0) {
        int t = 0;
        for (int w = 0; w < CLOUD_NT / 32; w++) t += wsum[w];
        blockcount[(size_t)b * nblk + blockIdx.x] = t;
    }
}

// one CTA: blockoff[i] = sum of blockcount[0 .. i), offsets[b] = blockoff[b * nblk], offsets[batch] = total
__global__ void __launch_bounds__(1024)
k_cloud_scan(const int* __restrict__ blockcount, int nblk, int batch, long long* __restrict__ blockoff,
             long long* __restrict__ offsets)
{
    __shared__ long long wtot[32];
    __shared__ long long carry;
    const long long total = (long long)nblk * batch;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (long long base = 0; base < total; base += 1024) {
        const long long i = base + threadIdx.x;
        const long long v = i < total ? blockcount[i] : 0;
        long long s = v;                               // inclusive scan inside the warp
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const long long t = __shfl_up_sync(0xffffffffu, s, o);
            if ((threadIdx.x & 31) >= o) s += t;
        }
        if ((threadIdx.x & 31) == 31) wtot[threadIdx.x >> 5] = s;
        __syncthreads();
        if (threadIdx.x < 32) {
            long long w = wtot[threadIdx.x], ws = w;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const long long t = __shfl_up_sync(0xffffffffu, ws, o);
                if (threadIdx.x >= o) ws += t;
            }
            wtot[threadIdx.x] = ws - w;               // exclusive warp offsets
        }
        __syncthreads();
        const long long excl = carry + wtot[threadIdx.x >> 5] + s - v;
        if (i < total) {
            blockoff[i] = excl;
            if (i % nblk == 0) offsets[i / nblk] = excl;
        }
        __syncthreads();
        if (threadIdx.x == 1023) carry = excl + v;
        __syncthreads();
    }
    if (threadIdx.x == 0) offsets[batch] = carry;
}

__device__ __forceinline__ float cloud_finite(float v) { return (isnan(v) || isinf(v)) ? 0.0f : v; }

template <typename T>
__global__ void __launch_bounds__(CLOUD_NT)
k_cloud_write(const T* __restrict__ disp, const float* __restrict__ xyz, const uint8_t* __restrict__ bgr, int npix, int nblk,
              double lo, double hi, const long long* __restrict__ blockoff, float* __restrict__ verts, uint8_t* __restrict__ rgb)
{
    const int b = blockIdx.y;
    const int p0 = blockIdx.x * CLOUD_PPB + threadIdx.x * CLOUD_PPT;
    bool keep[CLOUD_PPT];
    int n = 0;
#pragma unroll
    for (int j = 0; j < CLOUD_PPT; j++) {
        keep[j] = p0 + j < npix && cloud_keep(disp, (size_t)b * npix + p0 + j, lo, hi);
        n += keep[j] ? 1 : 0;
    }
    // exclusive scan of the per-thread counts over the block (threads are in pixel order)
    int s = n;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, s, o);
        if ((threadIdx.x & 31) >= o) s += t;
    }
    __shared__ int wtot[CLOUD_NT / 32];
    if ((threadIdx.x & 31) == 31) wtot[threadIdx.x >> 5] = s;
    __syncthreads();
    int woff = 0;
    for (int w = 0; w < (int)(threadIdx.x >> 5); w++) woff += wtot[w];
    long long o = blockoff[(size_t)b * nblk + blockIdx.x] + woff + s - n;
#pragma unroll
    for (int j = 0; j < CLOUD_PPT; j++) {
        if (!keep[j]) continue;
        const size_t i = (size_t)b * npix + p0 + j;
        verts[3 * o] = cloud_finite(xyz[3 * i]);
        verts[3 * o + 1] = cloud_finite(xyz[3 * i + 1]);
        verts[3 * o + 2] = cloud_finite(xyz[3 * i + 2]);
        if (rgb) { rgb[3 * o] = bgr[3 * i + 2]; rgb[3 * o + 1] = bgr[3 * i + 1]; rgb[3 * o + 2] = bgr[3 * i]; }
        o++;
    }
}

static size_t cloud_al(size_t v) { return (v + 255) / 256 * 256; }

extern "C" size_t b2s_compact_points_scratch_bytes(int H, int W, int batch)
{
    if (H <= 0 || W <= 0 || batch <= 0) return 0;
    const size_t nblk = (size_t)b2s_cdiv((long long)H * W, CLOUD_PPB) * batch;
    return cloud_al(nblk * sizeof(int)) + cloud_al(nblk * sizeof(long long));
}

template <typename T>
static int cloud_run(const T* disp, const float* xyz, const uint8_t* bgr, int npix, int nblk, int batch, double lo, double hi,
                     int* blockcount, long long* blockoff, long long* offsets, float* verts, uint8_t* rgb, cudaStream_t st)
{
    dim3 grid(nblk, batch);
    k_cloud_count<T><<<grid, CLOUD_NT, 0, st>>>(disp, npix, nblk, lo, hi, blockcount);
    B2S_LAUNCH_CHECK();
    k_cloud_scan<<<1, 1024, 0, st>>>(blockcount, nblk, batch, blockoff, offsets);
    B2S_LAUNCH_CHECK();
    k_cloud_write<T><<<grid, CLOUD_NT, 0, st>>>(disp, xyz, bgr, npix, nblk, lo, hi, blockoff, verts, rgb);
    B2S_LAUNCH_CHECK();
    return B2S_OK;
}

extern "C" int b2s_compact_points(const float* xyz, const void* disp, int disp_dtype, double lo, double hi, const uint8_t* bgr,
                                  int H, int W, int batch, float* verts, uint8_t* rgb, long long* offsets, void* scratch,
                                  void* stream)
{
    if (!xyz || !disp || !verts || !offsets || !scratch || H <= 0 || W <= 0 || batch <= 0) {
        b2s_set_error("compact_points: null pointer or empty image");
        return B2S_E_INVALID;
    }
    if ((rgb != nullptr) != (bgr != nullptr)) { b2s_set_error("compact_points: colours need both the BGR image and the RGB output"); return B2S_E_INVALID; }
    if ((long long)H * W >= (1ll << 31) || batch > 65535) { b2s_set_error("compact_points: image or batch too large"); return B2S_E_UNSUPPORTED; }
    const int npix = H * W, nblk = b2s_cdiv(npix, CLOUD_PPB);
    int* blockcount = (int*)scratch;
    long long* blockoff = (long long*)((char*)scratch + cloud_al((size_t)nblk * batch * sizeof(int)));
    cudaStream_t st = (cudaStream_t)stream;
    switch (disp_dtype) {
    case B2S_DT_U8: return cloud_run((const uint8_t*)disp, xyz, bgr, npix, nblk, batch, lo, hi, blockcount, blockoff, offsets, verts, rgb, st);
    case B2S_DT_I16: return cloud_run((const int16_t*)disp, xyz, bgr, npix, nblk, batch, lo, hi, blockcount, blockoff, offsets, verts, rgb, st);
    case B2S_DT_I32: return cloud_run((const int32_t*)disp, xyz, bgr, npix, nblk, batch, lo, hi, blockcount, blockoff, offsets, verts, rgb, st);
    case B2S_DT_F32: return cloud_run((const float*)disp, xyz, bgr, npix, nblk, batch, lo, hi, blockcount, blockoff, offsets, verts, rgb, st);
    }
    b2s_set_error("compact_points: unknown disparity dtype %d", disp_dtype);
    return B2S_E_INVALID;
}
