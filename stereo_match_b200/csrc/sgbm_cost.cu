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

// Generic cost kernel (any blockSize, Dp <= 2048).
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
                    *reinterpret_cast<uint32_t*>(out + ((size_t)yo * e.width1 + x) * e.Dp + b2s_cell_pos(d0, e.R, e.G, e.swz)) = v;
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

// ---------------------------------------------------------------------------
// Fast cost kernel (blockSize <= 11, Dp <= 256).
// CTA = 32 columns of pixel costs (32 - (BS-1) output columns) x all Dp disparities, marching
// down a band of rows.  Warp w owns disparities [8w, 8w+8); lane l owns column l.
//   stage   : left descriptor words of the row -> smem; right descriptors -> six u16 planes
//             (value, lo, hi of the Sobel and of the raw channel) in REVERSED x order, each in
//             two copies offset by one element, so that the pair (x-d, x-d-1) is one aligned
//             32-bit word: packed u16x2 operands come straight out of LDS with no byte shuffles
//   phase A : BT cost of (column, 8 disparities) -> pix[column][d] (smem, u16x2)
//   phase B : horizontal BS-sum from smem, vertical running sum through a ring of BS rows held
//             in registers (the row loop is unrolled by BS), one 16-byte store per thread
// ---------------------------------------------------------------------------
__device__ __forceinline__ uint32_t bt2(uint32_t u, uint32_t lo1, uint32_t hi1, uint32_t v, uint32_t lo2, uint32_t hi2)
{
    uint32_t A = u + B2S_BIAS - hi2;
    uint32_t Bv = lo2 + B2S_BIAS - u;
    uint32_t Cv = v + B2S_BIAS - hi1;
    uint32_t Dv = lo1 + B2S_BIAS - v;
    return __vminu2(__vimax3_u16x2(A, Bv, B2S_BIAS), __vimax3_u16x2(Cv, Dv, B2S_BIAS));   // biased by 0x0100
}

// sum of v over lanes [l, l + BS) (packed u16x2, no overflow: BS * 317 < 65536); lanes whose window
// leaves the warp return garbage (they are the halo columns and never store)
template <int BS>
__device__ __forceinline__ uint32_t window_sum(uint32_t v)
{
    uint32_t sum = 0, p = v;              // p = sum over a window of `len` lanes starting at this lane
    int off = 0;
#pragma unroll
    for (int len = 1; len <= BS; len <<= 1) {
        if (BS & len) {                   // take this power-of-two block at lane offset `off`
            sum += off ? __shfl_down_sync(0xffffffffu, p, off) : p;
            off += len;
        }
        if (2 * len <= BS) p += __shfl_down_sync(0xffffffffu, p, len);
    }
    return sum;
}

template <int BS, int MAXW, bool PAD>
__global__ void __launch_bounds__(32 * MAXW, (MAXW == 16 && BS <= 7) ? 2 : 1)
k_cost_fast(const uint2* __restrict__ descL, const uint2* __restrict__ descR, SgbmEff e, int TR, int NELP,
            uint16_t* __restrict__ cost, int* __restrict__ overflow_flag)
{
    constexpr int SW2 = BS / 2, TW = 32 - (BS - 1);
    extern __shared__ __align__(16) unsigned char smem_c[];
    const int Dp = e.Dp;
    const int lane = threadIdx.x & 31, oct = threadIdx.x >> 5, nthreads = blockDim.x;
    // smem: [2] x { LW[32] uint2 , planes[6][2][NELP] u16 }
    const int stage_bytes = 256 + 12 * NELP * 2;
    const int x0 = blockIdx.x * TW, yb = blockIdx.y * TR;
    const size_t b = blockIdx.z;
    const int yend = min(yb + TR, e.H);
    // phase A geometry of this thread (row independent)
    const int xiA = min(max(x0 - SW2 + lane, 0), e.width1 - 1) + e.minX1;     // clamp on the cropped axis
    const int XRM = min(max(x0 - SW2 + 31, 0), e.width1 - 1) + e.minX1 - e.minD;
    const int e0 = XRM - (xiA - e.minD - 8 * oct);                            // first reversed element
    const int par = e0 & 1;
    const int roff = 256 + (par * NELP) * 2 + ((e0 - par) >> 1) * 4;          // byte offset inside a stage buffer
    const int pstride = NELP;                                                 // u32 words between planes (2 copies each)
    // phase B geometry
    const int opos = b2s_cell_pos(8 * oct, e.R, e.G, e.swz);                  // where this octet lives in the record
    const int xB = x0 + lane;
    const bool outB = lane < TW && xB < e.width1;
    uint16_t* optr = cost + ((b * e.H + yb) * (size_t)e.width1 + min(xB, e.width1 - 1)) * Dp + opos;   // running store pointer
    uint32_t pmask[4];
#pragma unroll
    for (int w = 0; w < 4; w++) {
        int d0 = 8 * oct + 2 * w;
        pmask[w] = (d0 < e.D ? 0xffffu : 0u) | (d0 + 1 < e.D ? 0xffff0000u : 0u);
    }
    // staging role of this thread: element `sel` of the reversed right row (threads beyond nel idle),
    // threads of the last warp additionally fetch the left words
    const int nel = 32 + Dp;
    const int sel = threadIdx.x;                         // nel <= blockDim.x + 32: second element for the first 32 threads
    const int xr1 = XRM - sel, xr2 = XRM - (sel + nthreads);
    const bool st1 = sel < nel && xr1 >= 0 && xr1 < e.W, st2 = sel + nthreads < nel && xr2 >= 0 && xr2 < e.W;
    const bool stL = oct == (nthreads >> 5) - 1;

    // descriptor addresses as 32-bit pixel offsets from the kernel-argument pointers (batch*H*W < 2^31 is
    // checked on the host): one IMAD.WIDE per load and nothing to keep live across the row loop
    const unsigned bimg = (unsigned)b * (unsigned)(e.H * e.W);
    auto stage_load = [&](int yv, uint2& w1, uint2& w2, uint2& wl) {
        const int yr = min(max(yv, 0), e.H - 1);
        const unsigned rb = bimg + (unsigned)(yr * e.W);
        w1 = make_uint2(0u, 0u); w2 = make_uint2(0u, 0u); wl = make_uint2(0u, 0u);
        if (st1) w1 = __ldg(descR + (rb + (unsigned)xr1));
        if (st2) w2 = __ldg(descR + (rb + (unsigned)xr2));
        if (stL) wl = __ldg(descL + (rb + (unsigned)xiA));
    };
    auto stage_put = [&](unsigned char* buf, int el, const uint2& w) {
        uint16_t* planes = reinterpret_cast<uint16_t*>(buf + 256);
        const uint32_t vals[6] = { w.x & 0xffu, (w.x >> 8) & 0xffu, (w.x >> 16) & 0xffu,
                                   w.y & 0xffu, (w.y >> 8) & 0xffu, (w.y >> 16) & 0xffu };
#pragma unroll
        for (int p = 0; p < 6; p++) {
            planes[(2 * p) * NELP + el] = (uint16_t)vals[p];
            if (el > 0) planes[(2 * p + 1) * NELP + el - 1] = (uint16_t)vals[p];
        }
    };
    auto stage_store = [&](int bufi, const uint2& w1, const uint2& w2, const uint2& wl) {
        unsigned char* buf = smem_c + bufi * stage_bytes;
        if (sel < nel) stage_put(buf, sel, w1);
        if (sel + nthreads < nel) stage_put(buf, sel + nthreads, w2);
        if (stL) reinterpret_cast<uint2*>(buf)[lane] = wl;
    };

    uint32_t acc[4] = { 0, 0, 0, 0 }, ring[BS][4], vmax = 0;
#pragma unroll
    for (int s = 0; s < BS; s++) { ring[s][0] = ring[s][1] = ring[s][2] = ring[s][3] = 0; }

    {   // first row
        uint2 w1, w2, wl;
        stage_load(yb - SW2, w1, w2, wl);
        stage_store(0, w1, w2, wl);
    }
    __syncthreads();
    int cur = 0;
    for (int yv0 = yb - SW2; yv0 < yend + SW2; yv0 += BS) {
#pragma unroll
        for (int s = 0; s < BS; s++) {
            const int yv = yv0 + s;
            if (yv >= yend + SW2) break;                   // uniform
            // next row's descriptors travel while this row is computed
            uint2 nw1, nw2, nwl;
            const bool more = yv + 1 < yend + SW2;
            if (more) stage_load(yv + 1, nw1, nw2, nwl);
            const unsigned char* const buf = smem_c + cur * stage_bytes;
            // ---- phase A: pixel costs of (column, 8 disparities) --------------------------------
            uint32_t hs[4];
            {
                const uint2 wl = reinterpret_cast<const uint2*>(buf)[lane];
                const uint32_t* const rbase = reinterpret_cast<const uint32_t*>(buf + roff);
                const uint32_t ug = __byte_perm(wl.x, 0, 0x4040), log = __byte_perm(wl.x, 0, 0x4141), hig = __byte_perm(wl.x, 0, 0x4242);
                const uint32_t ur = __byte_perm(wl.y, 0, 0x4040), lor = __byte_perm(wl.y, 0, 0x4141), hir = __byte_perm(wl.y, 0, 0x4242);
                uint32_t pc[4];
#pragma unroll
                for (int w = 0; w < 4; w++) {
                    const uint32_t cg = bt2(ug, log, hig, rbase[w], rbase[pstride + w], rbase[2 * pstride + w]);
                    const uint32_t cr = bt2(ur, lor, hir, rbase[3 * pstride + w], rbase[4 * pstride + w], rbase[5 * pstride + w]);
                    // (cr - bias) >> 2 == ((cr >> 2) & 0x7f) - 0x40 per half
                    pc[w] = cg + ((cr >> 2) & 0x007f007fu) - (B2S_BIAS + 0x00400040u);
                }
                // ---- phase B: box sums.  The BS columns are the next BS-1 lanes of this warp: window sum
                // by doubling (p2 = x + x+1, p4 = p2 + p2(x+2), ...) through shuffles, no shared memory.
#pragma unroll
                for (int w = 0; w < 4; w++) hs[w] = window_sum<BS>(pc[w]);
            }
            {
#pragma unroll
                for (int w = 0; w < 4; w++) {
                    acc[w] += hs[w] - ring[s][w];          // halves never borrow: acc >= old contribution
                    ring[s][w] = hs[w];
                }
                // a half can only wrap past 0xffff after first being seen >= 0x8000
                vmax = __vimax3_u16x2(vmax, __vmaxu2(acc[0], acc[1]), __vmaxu2(acc[2], acc[3]));
                const int yo = yv - SW2;
                if (outB && yo >= yb) {
                    uint4 o = make_uint4(acc[0], acc[1], acc[2], acc[3]);
                    if (PAD) { o.x &= pmask[0]; o.y &= pmask[1]; o.z &= pmask[2]; o.w &= pmask[3]; }
                    *reinterpret_cast<uint4*>(optr) = o;
                }
                if (yo >= yb) optr += (size_t)e.width1 * Dp;
            }
            if (more) stage_store(cur ^ 1, nw1, nw2, nwl);
            cur ^= 1;
            __syncthreads();
        }
    }
    if (PAD) vmax &= (pmask[0] | pmask[1] | pmask[2] | pmask[3]);   // coarse: padded lanes never decide
    uint32_t m = max(vmax & 0xffffu, vmax >> 16);
    if (lane >= TW) m = 0;
    if (2 * (int)m + 3 * e.P2 > 32767) atomicOr(overflow_flag, 1);
}

template <int BS, int MAXW>
static int launch_cost_fast(const uint2* descL, const uint2* descR, const SgbmEff& e, uint16_t* cost, int* flag, cudaStream_t st)
{
    constexpr int TW = 32 - (BS - 1);
    const int nel = 32 + e.Dp;
    const int NELP = (nel + 63) / 64 * 64 + 32;            // copies land 64 bytes apart modulo 128: disjoint banks
    const size_t smem = 2 * (256 + (size_t)12 * NELP * 2);
    const int TR = 75;
    dim3 grid(b2s_cdiv(e.width1, TW), b2s_cdiv(e.H, TR), e.batch);
    const int nthreads = 32 * (e.Dp / 8);
    if (e.D != e.Dp) {
        if (smem > 48 * 1024) B2S_CUDA_CHECK(cudaFuncSetAttribute(k_cost_fast<BS, MAXW, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        k_cost_fast<BS, MAXW, true><<<grid, nthreads, smem, st>>>(descL, descR, e, TR, NELP, cost, flag);
    } else {
        if (smem > 48 * 1024) B2S_CUDA_CHECK(cudaFuncSetAttribute(k_cost_fast<BS, MAXW, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        k_cost_fast<BS, MAXW, false><<<grid, nthreads, smem, st>>>(descL, descR, e, TR, NELP, cost, flag);
    }
    B2S_LAUNCH_CHECK();
    return B2S_OK;
}

static int b2s_launch_cost_generic(const uint2* descL, const uint2* descR, const SgbmEff& e, uint16_t* cost,
                                   int* overflow_flag, cudaStream_t st);

int b2s_launch_cost(const uint2* descL, const uint2* descR, const SgbmEff& e, uint16_t* cost,
                    int* overflow_flag, cudaStream_t st)
{
    const int bs = 2 * e.SW2 + 1;
    if (e.Dp <= 128) {
        switch (bs) {
        case 1: return launch_cost_fast<1, 16>(descL, descR, e, cost, overflow_flag, st);
        case 3: return launch_cost_fast<3, 16>(descL, descR, e, cost, overflow_flag, st);
        case 5: return launch_cost_fast<5, 16>(descL, descR, e, cost, overflow_flag, st);
        case 7: return launch_cost_fast<7, 16>(descL, descR, e, cost, overflow_flag, st);
        case 9: return launch_cost_fast<9, 16>(descL, descR, e, cost, overflow_flag, st);
        case 11: return launch_cost_fast<11, 16>(descL, descR, e, cost, overflow_flag, st);
        }
    } else if (e.Dp <= 256) {
        switch (bs) {
        case 1: return launch_cost_fast<1, 32>(descL, descR, e, cost, overflow_flag, st);
        case 3: return launch_cost_fast<3, 32>(descL, descR, e, cost, overflow_flag, st);
        case 5: return launch_cost_fast<5, 32>(descL, descR, e, cost, overflow_flag, st);
        case 7: return launch_cost_fast<7, 32>(descL, descR, e, cost, overflow_flag, st);
        }
    }
    return b2s_launch_cost_generic(descL, descR, e, cost, overflow_flag, st);
}

static int b2s_launch_cost_generic(const uint2* descL, const uint2* descR, const SgbmEff& e, uint16_t* cost,
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
