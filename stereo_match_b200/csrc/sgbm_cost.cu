// sgbm_cost.cu -- pixel descriptors, Birchfield-Tomasi cost and block sum.
// Replaces the calcPixelCostBT / block-sum stage inside cv2.StereoSGBM.compute
// (reference call site: stereo_vision/stereo_vision.py:178).
//
// Layout in HBM
//   desc   [batch][H][W] uint2 : .x = Sobel channel word, .y = raw channel word,
//                                 word = value | lo<<8 | hi<<16 (half-pixel interval), top byte 0
//   cost   [batch][H][width1][Dp] uint16 : raw block cost (cv2's C minus P2); padded cells
//                                 d >= D hold 0 (the aggregation kernel pins their L_r at 0x7fff)
#include "common.cuh"

// ---------------------------------------------------------------------------
// descriptors
// ---------------------------------------------------------------------------
__device__ __forceinline__ int sobel_clip(const uint8_t* c, const uint8_t* n, const uint8_t* s,
                                          int x, int W, int ftzero)
{
    if (x <= 0 || x >= W - 1) return ftzero;
    int v = 2 * ((int)c[x + 1] - (int)c[x - 1]) + ((int)n[x + 1] - (int)n[x - 1]) +
            ((int)s[x + 1] - (int)s[x - 1]);
    return min(max(v, -ftzero), ftzero) + ftzero;
}

__device__ __forceinline__ uint32_t pack_desc(int pm, int p, int pp, bool has_l, bool has_r)
{
    int pl = has_l ? (p + pm) >> 1 : p;
    int pr = has_r ? (p + pp) >> 1 : p;
    int lo = min(min(pl, pr), p), hi = max(max(pl, pr), p);
    return (uint32_t)p | ((uint32_t)lo << 8) | ((uint32_t)hi << 16);
}

__global__ void __launch_bounds__(256)
k_descriptors(const uint8_t* __restrict__ img, int H, int W, int ftzero, uint2* __restrict__ desc)
{
    int x = blockIdx.x * blockDim.x + threadIdx.x;
    int y = blockIdx.y;
    size_t b = blockIdx.z;
    if (x >= W) return;
    const uint8_t* base = img + b * (size_t)H * W;
    const uint8_t* c = base + (size_t)y * W;
    const uint8_t* n = base + (size_t)(y > 0 ? y - 1 : y) * W;
    const uint8_t* s = base + (size_t)(y < H - 1 ? y + 1 : y) * W;
    int g0 = sobel_clip(c, n, s, x - 1, W, ftzero);
    int g1 = sobel_clip(c, n, s, x, W, ftzero);
    int g2 = sobel_clip(c, n, s, x + 1, W, ftzero);
    // raw channel: image value, but the two border columns are the constant ftzero
    int r0 = (x - 1 <= 0) ? ftzero : c[x - 1];
    int r1 = (x <= 0 || x >= W - 1) ? ftzero : c[x];
    int r2 = (x + 1 >= W - 1) ? ftzero : c[min(x + 1, W - 1)];
    uint2 o;
    o.x = pack_desc(g0, g1, g2, x > 0, x < W - 1);
    o.y = pack_desc(r0, r1, r2, x > 0, x < W - 1);
    desc[b * (size_t)H * W + (size_t)y * W + x] = o;
}

int b2s_launch_descriptors(const uint8_t* img, const SgbmEff& e, uint2* desc, cudaStream_t st)
{
    dim3 grid(b2s_cdiv(e.W, 256), e.H, e.batch);
    k_descriptors<<<grid, 256, 0, st>>>(img, e.H, e.W, e.ftzero, desc);
    B2S_LAUNCH_CHECK();
    return B2S_OK;
}

// ---------------------------------------------------------------------------
// BT cost of a disparity pair (d0, d0+1) in packed u16x2, biased by 0x0100 per half so
// that plain 32-bit adds/subs never borrow across halves.
//   wl : left descriptor word at x            (same for both halves)
//   wa : right descriptor word at x - d0      (low half)
//   wb : right descriptor word at x - d0 - 1  (high half)
// returns packed cost (0..255 per half)
// ---------------------------------------------------------------------------
#define B2S_BIAS 0x01000100u
__device__ __forceinline__ uint32_t bt_pair(uint32_t wl, uint32_t wa, uint32_t wb)
{
    uint32_t u = __byte_perm(wl, 0, 0x4040), lo1 = __byte_perm(wl, 0, 0x4141), hi1 = __byte_perm(wl, 0, 0x4242);
    uint32_t v = __byte_perm(wa, wb, 0x3430), lo2 = __byte_perm(wa, wb, 0x3531), hi2 = __byte_perm(wa, wb, 0x3632);
    uint32_t A = u + B2S_BIAS - hi2;       // u - hi2   (biased)
    uint32_t Bv = lo2 + B2S_BIAS - u;      // lo2 - u
    uint32_t Cv = v + B2S_BIAS - hi1;      // v - hi1
    uint32_t Dv = lo1 + B2S_BIAS - v;      // lo1 - v
    uint32_t c0 = __vimax3_u16x2(A, Bv, B2S_BIAS);
    uint32_t c1 = __vimax3_u16x2(Cv, Dv, B2S_BIAS);
    return __vminu2(c0, c1) - B2S_BIAS;
}

// One CTA: column strip [x0, x0+TW) of the cropped axis, row band [yb, yb+TR), all Dp.
// Marches down the band; per row: pixel costs for TW+2*SW2 columns -> smem, horizontal
// sum, vertical running sum through a ring of blockSize horizontal sums in smem.
#define COST_NT 512
#define COST_IPT 2   // items (x, dpair) per thread
__global__ void __launch_bounds__(COST_NT)
k_cost(const uint2* __restrict__ descL, const uint2* __restrict__ descR, SgbmEff e, int TW, int TR,
       uint16_t* __restrict__ cost, int* __restrict__ overflow_flag)
{
    extern __shared__ uint32_t smem[];
    const int DP2 = e.Dp >> 1;
    const int bs = 2 * e.SW2 + 1;
    const int PW = TW + 2 * e.SW2;                 // columns of pixel costs per row
    uint32_t* pix = smem;                          // [PW][DP2]
    uint32_t* ring = smem + (size_t)PW * DP2;      // [bs][TW][DP2]
    const int x0 = blockIdx.x * TW;
    const int yb = blockIdx.y * TR;
    const size_t b = blockIdx.z;
    const int yend = min(yb + TR, e.H);
    const uint2* dL = descL + b * (size_t)e.H * e.W;
    const uint2* dR = descR + b * (size_t)e.H * e.W;
    uint16_t* out = cost + b * (size_t)e.H * e.width1 * e.Dp;
    const int nitems = TW * DP2;

    uint32_t acc[COST_IPT];
#pragma unroll
    for (int i = 0; i < COST_IPT; i++) acc[i] = 0;
    for (int i = threadIdx.x; i < bs * nitems; i += COST_NT) ring[i] = 0;
    uint32_t vmax = 0;

    int slot = 0;
    for (int yv = yb - e.SW2; yv < yend + e.SW2; yv++) {
        const int yr = min(max(yv, 0), e.H - 1);
        __syncthreads();                           // previous row's pix fully consumed
        for (int it = threadIdx.x; it < PW * DP2; it += COST_NT) {
            int xh = it / DP2, dp = it - xh * DP2;
            int xc = min(max(x0 - e.SW2 + xh, 0), e.width1 - 1);   // clamp on the cropped axis
            int xi = xc + e.minX1;
            int d0 = 2 * dp;
            uint32_t pc = 0;
            if (d0 < e.D) {
                const uint2* rowL = dL + (size_t)yr * e.W;
                const uint2* rowR = dR + (size_t)yr * e.W;
                uint2 wl = __ldg(rowL + xi);
                int xr = xi - d0 - e.minD;
                uint2 wa = __ldg(rowR + xr);
                uint2 wb = (d0 + 1 < e.D) ? __ldg(rowR + xr - 1) : wa;
                uint32_t cg = bt_pair(wl.x, wa.x, wb.x);
                uint32_t cr = bt_pair(wl.y, wa.y, wb.y);
                pc = cg + ((cr >> 2) & 0x003f003fu);
            }
            pix[it] = pc;
        }
        __syncthreads();
#pragma unroll
        for (int i = 0; i < COST_IPT; i++) {
            int it = threadIdx.x + i * COST_NT;
            if (it < nitems) {
                int xl = it / DP2, dp = it - xl * DP2;
                uint32_t hs = 0;
                for (int k = 0; k < bs; k++) hs += pix[(xl + k) * DP2 + dp];
                uint32_t* rp = ring + (size_t)slot * nitems + it;
                acc[i] += hs - *rp;                // halves never borrow: acc >= old contribution
                *rp = hs;
                // a half can only wrap past 0xffff after first being seen >= 0x8000 (one row
                // changes it by < 0x8000), so tracking the running max catches every overflow
                vmax = __vmaxu2(vmax, acc[i]);
                int yo = yv - e.SW2;
                int x = x0 + xl;
                if (yo >= yb && x < e.width1) {
                    uint32_t v = acc[i];
                    int d0 = 2 * dp;
                    if (d0 >= e.D) v = 0u;
                    else if (d0 + 1 >= e.D) v = v & 0xffffu;
                    *reinterpret_cast<uint32_t*>(out + ((size_t)yo * e.width1 + x) * e.Dp + d0) = v;
                }
            }
        }
        slot = slot + 1 == bs ? 0 : slot + 1;
    }
    uint32_t m = max(vmax & 0xffffu, vmax >> 16);
    // Guaranteed-parity domain: cv2's SIMD lanes evaluate C + min(...) with saturating int16
    // adds before subtracting delta, its scalar tail does not.  Both agree (and agree with this
    // engine) while (raw + P2) + (minL + P2) <= 32767; minL <= raw_max + P2 bounds it.
    if (2 * (int)m + 3 * e.P2 > 32767) atomicOr(overflow_flag, 1);
}

int b2s_launch_cost(const uint2* descL, const uint2* descR, const SgbmEff& e, uint16_t* cost,
                    int* overflow_flag, cudaStream_t st)
{
    const int DP2 = e.Dp / 2, bs = 2 * e.SW2 + 1;
    int TW = (COST_NT * COST_IPT) / DP2;
    if (TW < 1) TW = 1;
    if (TW > 32) TW = 32;
    if (TW > e.width1) TW = e.width1;
    // TW*DP2 <= COST_NT*COST_IPT must hold: Dp <= 2048
    if (TW * DP2 > COST_NT * COST_IPT) {
        b2s_set_error("numDisparities too large for the cost kernel (Dp=%d)", e.Dp);
        return B2S_E_UNSUPPORTED;
    }
    size_t smem = ((size_t)(TW + 2 * e.SW2) * DP2 + (size_t)bs * TW * DP2) * sizeof(uint32_t);
    while (smem > 200 * 1024 && TW > 1) {
        TW = (TW + 1) / 2;
        smem = ((size_t)(TW + 2 * e.SW2) * DP2 + (size_t)bs * TW * DP2) * sizeof(uint32_t);
    }
    if (smem > 220 * 1024) {
        b2s_set_error("blockSize*numDisparities too large for the cost kernel (needs %zu B smem)", smem);
        return B2S_E_UNSUPPORTED;
    }
    int TR = 64;
    static size_t configured = 0;
    if (smem > 48 * 1024 && smem > configured) {
        B2S_CUDA_CHECK(cudaFuncSetAttribute(k_cost, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured = smem;
    }
    dim3 grid(b2s_cdiv(e.width1, TW), b2s_cdiv(e.H, TR), e.batch);
    k_cost<<<grid, COST_NT, smem, st>>>(descL, descR, e, TW, TR, cost, overflow_flag);
    B2S_LAUNCH_CHECK();
    return B2S_OK;
}
