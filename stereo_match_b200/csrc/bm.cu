// bm.cu -- StereoBM: x-Sobel prefilter + sliding-window SAD block matching.
// Replaces cv2.StereoBM_create(numDisparities, blockSize).compute
// (reference call site stereo_vision/stereo_vision.py:165-166,178).  minDisparity = 0 only.
#include "common.cuh"

struct BmEff {
    int H, W, batch, D, r, ftzero, tex, uniq;
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
#define BM_TW 32
__global__ void __launch_bounds__(BM_NT)
k_bm(const uint8_t* __restrict__ PL, const uint8_t* __restrict__ PR, BmEff e, int TR, int16_t* __restrict__ disp)
{
    extern __shared__ uint32_t sm[];
    const int D = e.D, r = e.r, CW = BM_TW + 2 * r;
    uint32_t* col = sm;                       // [CW][D]
    uint32_t* sad = col + (size_t)CW * D;     // [BM_TW][D]
    uint32_t* tcol = sad + (size_t)BM_TW * D; // [CW]
    const int xs = e.D - 1 + r;               // first valid output column
    const int x0 = xs + blockIdx.x * BM_TW;
    const int y0 = r + blockIdx.y * TR;
    const int y1 = min(y0 + TR, e.H - r);
    const size_t base = (size_t)blockIdx.z * e.H * e.W;
    const uint8_t* L = PL + base;
    const uint8_t* Rr = PR + base;
    const int xmax = e.W - r;                 // one past last valid output column
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

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
                int a = abs((int)L[(size_t)yin * e.W + x] - (int)Rr[(size_t)yin * e.W + x - d]);
                if (sub) a -= abs((int)L[(size_t)yold * e.W + x] - (int)Rr[(size_t)yold * e.W + x - d]);
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
            for (int it = threadIdx.x; it < BM_TW * D; it += BM_NT) {
                int xl = it / D, d = it - xl * D;
                uint32_t s = 0;
                for (int k = 0; k <= 2 * r; k++) s += col[(xl + k) * D + d];
                sad[it] = s;
            }
            __syncthreads();
            // one warp per output pixel
            for (int xl = warp; xl < BM_TW; xl += BM_NT / 32) {
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
                    int out = -16;
                    if (!bad) {
                        int p = best_d > 0 ? (int)sp[best_d - 1] : (int)sp[best_d + 1];
                        int n = best_d < D - 1 ? (int)sp[best_d + 1] : (int)sp[best_d - 1];
                        int dd = p + n - 2 * (int)best_s + abs(p - n);
                        out = (best_d * 256 + (dd ? (p - n) * 256 / dd : 0) + 15) >> 4;
                    }
                    disp[base + (size_t)yout * e.W + x] = (int16_t)out;
                }
            }
        }
        __syncthreads();
    }
}

__global__ void __launch_bounds__(256)
k_fill_i16(int16_t* p, size_t n, int16_t v)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) p[i] = v;
}

int b2s_bm_run(const uint8_t* left, const uint8_t* right, int H, int W, int batch, const b2s_bm_params& p,
               uint8_t* PL, uint8_t* PR, int16_t* disp, cudaStream_t st)
{
    BmEff e;
    e.H = H; e.W = W; e.batch = batch; e.D = p.numDisparities; e.r = p.blockSize / 2;
    e.ftzero = p.preFilterCap; e.tex = p.textureThreshold; e.uniq = p.uniquenessRatio;
    const size_t n = (size_t)batch * H * W;
    k_fill_i16<<<b2s_cdiv((long long)n, 256), 256, 0, st>>>(disp, n, (int16_t)-16);
    B2S_LAUNCH_CHECK();
    const int xs = e.D - 1 + e.r, xe = W - e.r, ys = e.r, ye = H - e.r;
    if (xs >= xe || ys >= ye) return B2S_OK;           // no valid pixel: everything FILTERED
    dim3 gp(b2s_cdiv(W, 256), H, batch);
    k_bm_prefilter<<<gp, 256, 0, st>>>(left, H, W, e.ftzero, PL);
    B2S_LAUNCH_CHECK();
    k_bm_prefilter<<<gp, 256, 0, st>>>(right, H, W, e.ftzero, PR);
    B2S_LAUNCH_CHECK();
    const int CW = BM_TW + 2 * e.r;
    size_t smem = ((size_t)CW * e.D + (size_t)BM_TW * e.D + CW) * sizeof(uint32_t);
    if (smem > 220 * 1024) {
        b2s_set_error("StereoBM: numDisparities*blockSize too large (needs %zu B shared memory)", smem);
        return B2S_E_UNSUPPORTED;
    }
    static size_t configured = 0;
    if (smem > 48 * 1024 && smem > configured) {
        B2S_CUDA_CHECK(cudaFuncSetAttribute(k_bm, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured = smem;
    }
    const int TR = 32;
    dim3 grid(b2s_cdiv(xe - xs, BM_TW), b2s_cdiv(ye - ys, TR), batch);
    k_bm<<<grid, BM_NT, smem, st>>>(PL, PR, e, TR, disp);
    B2S_LAUNCH_CHECK();
    return B2S_OK;
}
