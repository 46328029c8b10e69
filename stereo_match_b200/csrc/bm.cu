// bm.cu -- StereoBM: x-Sobel prefilter + sliding-window SAD block matching.
// Replaces cv2.StereoBM_create(numDisparities, blockSize).compute
// (reference call site stereo_vision/stereo_vision.py:165-166,178) and, with minDisparity = -(D - 1), the right-view
// matcher cv2.ximgproc.createRightMatcher builds from it (:171,179).
//
// Any minDisparity (restated in oracle/stereo_oracle.c, probed against cv2 4.13.0): disparity index dd = d - minD in
// [0, D); the window column X' of the left image is compared with column min(X' - minD, W - 1) - dd of the right
// image -- ONE clamp for all disparities, so for negative minD the right window stops moving in the last blockSize/2
// valid columns while the left one goes on.  Valid output: X in [max(0, minD + D - 1) + r, min(W - r, W + minD)),
// y in [r, H - r); everything else FILTERED = (minD - 1) * 16.  The fast kernel covers the columns whose windows are
// not clamped, the generic kernel the (at most r) clamped ones.
#include "common.cuh"

struct BmEff {
    int H, W, batch, D, r, ftzero, tex, uniq;
    int minD, filtered;       // FILTERED = (minD - 1) * 16
    int xa, xb;               // output columns of this launch
};

// prefilter: rows are filtered with weights (1,2,1) and reflect-101 rows, except that for an
// odd H the last row is the constant ftzero (cv2 processes rows in pairs).
__global__ void __launch_bounds__(256)
k_bm_prefilter(const uint8_t* __restrict__ img, int H, int W, int ftzero, uint8_t* __restrict__ out)
{
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    const int y = blockIdx.y;
    if (x >= W) return;
    const size_t base = (size_t)blockIdx.z * H * W;
    int v = ftzero;
    const bool lastodd = (H & 1) && y == H - 1;
    if (x > 0 && x < W - 1 && !lastodd) {
        const int ya = y > 0 ? y - 1 : y + 1;
        const int yb = y + 1 < H ? y + 1 : y - 1;
        const uint8_t* a = img + base + (size_t)ya * W;
        const uint8_t* c = img + base + (size_t)y * W;
        const uint8_t* b = img + base + (size_t)yb * W;
        int s = ((int)a[x + 1] - (int)a[x - 1]) + 2 * ((int)c[x + 1] - (int)c[x - 1]) + ((int)b[x + 1] - (int)b[x - 1]);
        v = min(max(s, -ftzero), ftzero) + ftzero;
    }
    out[base + (size_t)y * W + x] = (uint8_t)v;
}

// One CTA: output columns [x0, x0+TW) of output rows [yb, yb+TR), all D disparities.
// smem: col[TW+2r][D] vertical-window column SADs (u32), sad[TW][D], tcol[TW+2r].
#define BM_NT 512
#define BM_TW 32        // widest tile; large blockSize * numDisparities use narrower ones (launch_bm_generic)
__global__ void __launch_bounds__(BM_NT)
k_bm(const uint8_t* __restrict__ PL, const uint8_t* __restrict__ PR, BmEff e, int TR, int TW, int16_t* __restrict__ disp)
{
    extern __shared__ uint32_t sm[];
    const int D = e.D, r = e.r, CW = TW + 2 * r;
    uint32_t* col = sm;                       // [CW][D]
    uint32_t* sad = col + (size_t)CW * D;     // [TW][D]
    uint32_t* tcol = sad + (size_t)TW * D;    // [CW]
    const int x0 = e.xa + blockIdx.x * TW;
    const int y0 = r + blockIdx.y * TR;
    const int y1 = min(y0 + TR, e.H - r);
    const size_t base = (size_t)blockIdx.z * e.H * e.W;
    const uint8_t* L = PL + base;
    const uint8_t* Rr = PR + base;
    const int xmax = e.xb;                    // one past the last output column of this launch
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int wlast = e.W - 1;

    for (int i = threadIdx.x; i < CW * D; i += BM_NT) col[i] = 0;
    for (int i = threadIdx.x; i < CW; i += BM_NT) tcol[i] = 0;
    __syncthreads();

    // rows enter the vertical window one at a time; the first 2r rows only warm it up
    for (int yin = y0 - r; yin < y1 + r; yin++) {
        const int yout = yin - r;             // window [yout-r, yout+r] complete after adding yin
        const int yold = yin - (2 * r + 1);   // row leaving the window
        const bool sub = yold >= y0 - r;
        for (int it = threadIdx.x; it < CW * D; it += BM_NT) {
            int xc = it / D, d = it - xc * D;
            int x = x0 - r + xc;
            if (x < e.W) {
                const int xr = min(x - e.minD, wlast) - d;        // one clamp for all disparities (see the file header)
                int a = abs((int)L[(size_t)yin * e.W + x] - (int)Rr[(size_t)yin * e.W + xr]);
                if (sub) a -= abs((int)L[(size_t)yold * e.W + x] - (int)Rr[(size_t)yold * e.W + xr]);
                col[it] += a;
            }
        }
        for (int xc = threadIdx.x; xc < CW; xc += BM_NT) {
            int x = x0 - r + xc;
            if (x < e.W) {
                int a = abs((int)L[(size_t)yin * e.W + x] - e.ftzero);
                if (sub) a -= abs((int)L[(size_t)yold * e.W + x] - e.ftzero);
                tcol[xc] += a;
            }
        }
        __syncthreads();
        if (yout >= y0) {
            for (int it = threadIdx.x; it < TW * D; it += BM_NT) {
                int xl = it / D, d = it - xl * D;
                uint32_t s = 0;
                for (int k = 0; k <= 2 * r; k++) s += col[(xl + k) * D + d];
                sad[it] = s;
            }
            __syncthreads();
            // one warp per output pixel
            for (int xl = warp; xl < TW; xl += BM_NT / 32) {
                const int x = x0 + xl;
                if (x >= xmax) break;
                const uint32_t* sp = sad + (size_t)xl * D;
                // min SAD, ties -> largest disparity
                uint32_t best_s = 0xffffffffu; int best_d = 0;
                for (int d = lane; d < D; d += 32) {
                    uint32_t s = sp[d];
                    if (s < best_s || (s == best_s && d > best_d)) { best_s = s; best_d = d; }
                }
                for (int o = 16; o > 0; o >>= 1) {
                    uint32_t os = __shfl_xor_sync(0xffffffffu, best_s, o);
                    int od = __shfl_xor_sync(0xffffffffu, best_d, o);
                    if (os < best_s || (os == best_s && od > best_d)) { best_s = os; best_d = od; }
                }
                int tsum = 0;
                for (int k = lane; k <= 2 * r; k += 32) tsum += (int)tcol[xl + k];
                for (int o = 16; o > 0; o >>= 1) tsum += __shfl_xor_sync(0xffffffffu, tsum, o);
                bool bad = tsum < e.tex;
                if (e.uniq > 0) {
                    int thresh = (int)best_s + (int)best_s * e.uniq / 100;
                    int hit = 0;
                    for (int d = lane; d < D; d += 32)
                        if (abs(d - best_d) > 1 && (int)sp[d] <= thresh) hit = 1;
                    bad = bad || __any_sync(0xffffffffu, hit);
                }
                if (lane == 0) {
                    int out = e.filtered;
                    if (!bad) {
                        int p = best_d > 0 ? (int)sp[best_d - 1] : (int)sp[best_d + 1];
                        int n = best_d < D - 1 ? (int)sp[best_d + 1] : (int)sp[best_d - 1];
                        int dd = p + n - 2 * (int)best_s + abs(p - n);
                        out = ((best_d + e.minD) * 256 + (dd ? (p - n) * 256 / dd : 0) + 15) >> 4;
                    }
                    disp[base + (size_t)yout * e.W + x] = (int16_t)out;
                }
            }
        }
        __syncthreads();
    }
}

// ---------------------------------------------------------------------------
// Fast StereoBM kernel (blockSize <= 21, D <= 512).
// CTA = 32 output columns (lane = column) x a band of rows; warp w owns disparities [16w, 16w+16).
// Every thread keeps the 16 window SADs of its column in registers and marches down the band:
//   v[d] += h(yin, d) - h(yin - BS, d),   h(y, d) = sum_{dx} |PL[y][x+dx] - PR[y][x+dx-d]|
// h is evaluated with VABSDIFF4.ACC on 4 neighbouring pixels at a time; the unaligned 4-byte windows of the
// right row come out of shared memory with funnel shifts whose amounts are compile-time constants relative
// to the thread's own base word (only the base alignment is a run-time shift).  The row that leaves the
// vertical window is re-evaluated instead of being kept (a ring of BS+2 staged rows in shared memory).
// Winner-take-all: per-thread argmin over its 16 disparities -> per-warp candidates in shared memory ->
// every thread derives the pixel's best and threshold, checks uniqueness on its own registers.
// ---------------------------------------------------------------------------
template <int BS>
__global__ void __launch_bounds__(1024)
k_bm_fast(const uint8_t* __restrict__ PL, const uint8_t* __restrict__ PR, BmEff e, int TR, int16_t* __restrict__ disp)
{
    constexpr int r = BS / 2, NWD = (BS + 3) / 4, NS = BS + 2;      // window words, staged-row ring
    constexpr uint32_t LASTMASK = (BS % 4) ? ((1u << (8 * (BS % 4))) - 1u) : 0xffffffffu;
    extern __shared__ __align__(16) unsigned char smb[];
    const int D = e.D, NW = D / 16;
    const int LBYTES = ((32 + 2 * r + 8 + 15) / 16) * 16;           // left row segment (+8 for the last funnel)
    const int RBYTES = ((D + 32 + 2 * r + 4 * NWD + 24 + 15) / 16) * 16;   // right row segment, 8 bytes of front padding
    unsigned char* const Lrows = smb;                                // [NS][LBYTES]
    unsigned char* const Rrows = Lrows + NS * LBYTES;                // [NS][RBYTES]
    uint32_t* const sad = reinterpret_cast<uint32_t*>(Rrows + NS * RBYTES);   // [D][33]
    uint32_t* const cand_s = sad + (size_t)D * 33;                   // [NW][32]
    uint32_t* const cand_d = cand_s + NW * 32;                       // [NW][32]
    uint32_t* const badf = cand_d + NW * 32;                         // [32]
    int* const tsum_s = reinterpret_cast<int*>(badf + 32);           // [32]
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nthreads = blockDim.x;
    const int xe = e.xb;
    const int x0 = e.xa + blockIdx.x * 32, x = x0 + lane;
    const int y0 = r + blockIdx.y * TR, y1 = min(y0 + TR, e.H - r);
    const size_t base = (size_t)blockIdx.z * e.H * e.W;
    const int dlo = 16 * warp;
    const int xL0 = x0 - r;                                          // image column of Lrows byte 0
    const int xR0 = x0 - r - (D - 1) - e.minD - 8;                   // image column of Rrows byte 0
    // byte offset of the right window start for d = dlo (largest start of this thread's range)
    const int smax = lane + (D - 1) - dlo + 8;
    const int rw = smax >> 2, rsh = (smax & 3) * 8;
    const int lw = lane >> 2, lsh = (lane & 3) * 8;

    auto stage = [&](int y) {                                        // row y -> ring slot (y - (y0 - r)) % NS
        const int slot = (y - (y0 - r)) % NS;
        const uint8_t* rl = PL + base + (size_t)y * e.W;
        const uint8_t* rr = PR + base + (size_t)y * e.W;
        for (int i = threadIdx.x; i < LBYTES; i += nthreads) {
            const int xx = xL0 + i;
            Lrows[slot * LBYTES + i] = (xx >= 0 && xx < e.W) ? rl[xx] : 0;
        }
        for (int i = threadIdx.x; i < RBYTES; i += nthreads) {
            const int xx = xR0 + i;
            Rrows[slot * RBYTES + i] = (xx >= 0 && xx < e.W) ? rr[xx] : 0;
        }
    };
    // window SADs of row-slot `slot` for this thread's 16 disparities: acc[j] (+)= h(d = dlo + j)
    auto hrow = [&](int slot, uint32_t (&acc)[16], int& tex) {
        const uint32_t* lwp = reinterpret_cast<const uint32_t*>(Lrows + slot * LBYTES) + lw;
        const uint32_t* rwp = reinterpret_cast<const uint32_t*>(Rrows + slot * RBYTES) + rw;
        uint32_t Lw[NWD], U[NWD + 4];
        {
            uint32_t a = lwp[0];
#pragma unroll
            for (int i = 0; i < NWD; i++) { uint32_t b = lwp[i + 1]; Lw[i] = __funnelshift_r(a, b, lsh); a = b; }
            Lw[NWD - 1] &= LASTMASK;
            a = rwp[-4];
#pragma unroll
            for (int q = -4; q < NWD; q++) { uint32_t b = rwp[q + 1]; U[q + 4] = __funnelshift_r(a, b, rsh); a = b; }
        }
        const uint32_t ft4 = (uint32_t)e.ftzero * 0x01010101u;
#pragma unroll
        for (int i = 0; i < NWD; i++) tex = __vsadu4(Lw[i], i == NWD - 1 ? (ft4 & LASTMASK) : ft4) + tex;
#pragma unroll
        for (int j = 0; j < 16; j++) {
            const int g = j >> 2, jj = j & 3;
            uint32_t a = acc[j];
#pragma unroll
            for (int i = 0; i < NWD; i++) {
                const int q = i - g;                                 // window word = bytes at smax + 4q - jj
                uint32_t w = jj ? __funnelshift_l(U[q - 1 + 4], U[q + 4], 8 * jj) : U[q + 4];
                if (i == NWD - 1) w &= LASTMASK;
                a = __vsadu4(Lw[i], w) + a;
            }
            acc[j] = a;
        }
    };

    uint32_t v[16];
#pragma unroll
    for (int j = 0; j < 16; j++) v[j] = 0;
    int tsum = 0;
    stage(y0 - r);
    __syncthreads();
    for (int yin = y0 - r; yin < y1 + r; yin++) {
        if (yin + 1 < y1 + r) stage(yin + 1);                        // lands in a slot nobody reads this step
        const int slot = (yin - (y0 - r)) % NS;
        hrow(slot, v, tsum);
        const int yold = yin - BS;
        if (yold >= y0 - r) {
            uint32_t o[16];
#pragma unroll
            for (int j = 0; j < 16; j++) o[j] = 0;
            int to = 0;
            hrow((yold - (y0 - r)) % NS, o, to);
#pragma unroll
            for (int j = 0; j < 16; j++) v[j] -= o[j];
            tsum -= to;
        }
        const int yout = yin - r;
        if (yout >= y0) {                                            // uniform
            // per-thread argmin (ties -> largest d) and the SAD row for the sub-pixel neighbours
            uint32_t bs_ = v[0]; int bd = 0;
#pragma unroll
            for (int j = 1; j < 16; j++) if (v[j] <= bs_) { bs_ = v[j]; bd = j; }
            cand_s[warp * 32 + lane] = bs_;
            cand_d[warp * 32 + lane] = (uint32_t)(dlo + bd);
#pragma unroll
            for (int j = 0; j < 16; j++) sad[(dlo + j) * 33 + lane] = v[j];
            if (warp == 0) { badf[lane] = 0; tsum_s[lane] = tsum; }
            __syncthreads();
            uint32_t best_s = cand_s[lane]; int best_d = (int)cand_d[lane];
            for (int w = 1; w < NW; w++) {
                const uint32_t cs = cand_s[w * 32 + lane];
                if (cs <= best_s) { best_s = cs; best_d = (int)cand_d[w * 32 + lane]; }
            }
            if (e.uniq > 0) {
                const uint32_t thresh = best_s + best_s * (uint32_t)e.uniq / 100u;
                bool hit = false;
#pragma unroll
                for (int j = 0; j < 16; j++) {
                    const int dd = dlo + j - best_d;
                    hit = hit || ((dd > 1 || dd < -1) && v[j] <= thresh);
                }
                if (hit) badf[lane] = 1;                             // benign race: every writer stores 1
            }
            __syncthreads();
            if (warp == 0 && x < xe) {
                int out = e.filtered;
                if (!(tsum_s[lane] < e.tex) && !badf[lane]) {
                    const int p = (int)sad[(best_d > 0 ? best_d - 1 : best_d + 1) * 33 + lane];
                    const int n = (int)sad[(best_d < D - 1 ? best_d + 1 : best_d - 1) * 33 + lane];
                    const int dd = p + n - 2 * (int)best_s + abs(p - n);
                    out = ((best_d + e.minD) * 256 + (dd ? (p - n) * 256 / dd : 0) + 15) >> 4;
                }
                disp[base + (size_t)yout * e.W + x] = (int16_t)out;
            }
        }
        __syncthreads();
    }
}

template <int BS>
static int launch_bm_fast(const uint8_t* PL, const uint8_t* PR, const BmEff& e, int16_t* disp, cudaStream_t st)
{
    constexpr int r = BS / 2, NWD = (BS + 3) / 4, NS = BS + 2;
    const int LBYTES = ((32 + 2 * r + 8 + 15) / 16) * 16;
    const int RBYTES = ((e.D + 32 + 2 * r + 4 * NWD + 24 + 15) / 16) * 16;
    const int NW = e.D / 16;
    const size_t smem = (size_t)NS * (LBYTES + RBYTES) + ((size_t)e.D * 33 + 2 * NW * 32 + 64) * 4;
    if (smem > 48 * 1024) B2S_CUDA_CHECK(cudaFuncSetAttribute(k_bm_fast<BS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int ys = e.r, ye = e.H - e.r;
    const int TR = 96;
    dim3 grid(b2s_cdiv(e.xb - e.xa, 32), b2s_cdiv(ye - ys, TR), e.batch);
    k_bm_fast<BS><<<grid, 32 * NW, smem, st>>>(PL, PR, e, TR, disp);
    B2S_LAUNCH_CHECK();
    return B2S_OK;
}

__global__ void __launch_bounds__(256)
k_fill_i16(int16_t* p, size_t n, int16_t v)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) p[i] = v;
}

static int launch_bm_generic(const uint8_t* PL, const uint8_t* PR, const BmEff& e, int16_t* disp, cudaStream_t st)
{
    // tile width: the widest whose column sums fit in shared memory ((TW + 2r) * D + TW * D + TW + 2r words)
    int TW = BM_TW;
    size_t smem = 0;
    for (; TW >= 4; TW /= 2) {
        const int CW = TW + 2 * e.r;
        smem = ((size_t)CW * e.D + (size_t)TW * e.D + CW) * sizeof(uint32_t);
        if (smem <= 220 * 1024) break;
    }
    if (TW < 4) {
        b2s_set_error("StereoBM: numDisparities * blockSize too large (%d * %d needs %zu B of shared memory at a 4-column tile)",
                      e.D, 2 * e.r + 1, smem);
        return B2S_E_UNSUPPORTED;
    }
    // the attribute belongs to the current device's context: set it before every launch (cheap), not once per process
    if (smem > 48 * 1024) B2S_CUDA_CHECK(cudaFuncSetAttribute(k_bm, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int TR = 32;
    dim3 grid(b2s_cdiv(e.xb - e.xa, TW), b2s_cdiv(e.H - 2 * e.r, TR), e.batch);
    k_bm<<<grid, BM_NT, smem, st>>>(PL, PR, e, TR, TW, disp);
    B2S_LAUNCH_CHECK();
    return B2S_OK;
}

int b2s_bm_run(const uint8_t* left, const uint8_t* right, int H, int W, int batch, const b2s_bm_params& p,
               uint8_t* PL, uint8_t* PR, int16_t* disp, cudaStream_t st)
{
    BmEff e;
    e.H = H; e.W = W; e.batch = batch; e.D = p.numDisparities; e.r = p.blockSize / 2;
    e.ftzero = p.preFilterCap; e.tex = p.textureThreshold; e.uniq = p.uniquenessRatio;
    e.minD = p.minDisparity; e.filtered = (p.minDisparity - 1) * 16;
    const size_t n = (size_t)batch * H * W;
    k_fill_i16<<<b2s_cdiv((long long)n, 256), 256, 0, st>>>(disp, n, (int16_t)e.filtered);
    B2S_LAUNCH_CHECK();
    // cv2's geometry (stereobm.cpp: lofs / rofs / width1 and getValidDisparityROI), see the file header
    const int span = e.D - 1 + e.minD;
    const int lofs = span > 0 ? span : 0, rofs = span < 0 ? -span : 0, width1 = W - rofs - e.D + 1;
    if (lofs >= W || rofs >= W || width1 < 1) return B2S_OK;
    const int xmin = lofs + e.r, xend = (W - e.r < W + e.minD) ? W - e.r : W + e.minD, ys = e.r, ye = H - e.r;
    if (xmin >= xend || ys >= ye) return B2S_OK;           // no valid pixel: everything FILTERED
    dim3 gp(b2s_cdiv(W, 256), H, batch);
    k_bm_prefilter<<<gp, 256, 0, st>>>(left, H, W, e.ftzero, PL);
    B2S_LAUNCH_CHECK();
    k_bm_prefilter<<<gp, 256, 0, st>>>(right, H, W, e.ftzero, PR);
    B2S_LAUNCH_CHECK();
    // columns whose window reaches past W - 1 + minD have a clamped right window: generic kernel
    int xfast = xend;
    if (e.minD < 0 && W + e.minD - e.r < xfast) xfast = W + e.minD - e.r;
    if (xfast < xmin) xfast = xmin;
    const bool fast = e.D <= 512 && p.blockSize >= 5 && p.blockSize <= 21;
    int rc = B2S_OK;
    if (fast && xfast > xmin) {
        e.xa = xmin; e.xb = xfast;
        switch (p.blockSize) {
        case 5: rc = launch_bm_fast<5>(PL, PR, e, disp, st); break;
        case 7: rc = launch_bm_fast<7>(PL, PR, e, disp, st); break;
        case 9: rc = launch_bm_fast<9>(PL, PR, e, disp, st); break;
        case 11: rc = launch_bm_fast<11>(PL, PR, e, disp, st); break;
        case 13: rc = launch_bm_fast<13>(PL, PR, e, disp, st); break;
        case 15: rc = launch_bm_fast<15>(PL, PR, e, disp, st); break;
        case 17: rc = launch_bm_fast<17>(PL, PR, e, disp, st); break;
        case 19: rc = launch_bm_fast<19>(PL, PR, e, disp, st); break;
        case 21: rc = launch_bm_fast<21>(PL, PR, e, disp, st); break;
        }
        if (rc) return rc;
    }
    e.xa = fast ? xfast : xmin; e.xb = xend;
    if (e.xa < e.xb) rc = launch_bm_generic(PL, PR, e, disp, st);
    return rc;
}
