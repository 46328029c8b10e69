// api.cu -- C ABI of libb2stereo.so (see include/b2stereo.h): handles, parameter rules,
// workspace, and the per-call kernel pipeline.  No CPU fallback anywhere.
#include "common.cuh"
#include "sgbm_wave.cuh"
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>

static thread_local char g_err[512] = "";
static thread_local uint64_t g_launches = 0;

void b2s_set_error(const char* fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
void b2s_count_launch(int n) { g_launches += (uint64_t)n; }

extern "C" const char* b2s_last_error(void) { return g_err; }
extern "C" const char* b2s_version(void) { return "b2stereo 0.1 (sm_100a)"; }
extern "C" uint64_t b2s_launch_count(void) { return g_launches; }

int b2s_bm_run(const uint8_t* left, const uint8_t* right, int H, int W, int batch, const b2s_bm_params& p,
               uint8_t* PL, uint8_t* PR, int16_t* disp, cudaStream_t st);

struct b2s_matcher {
    int kind;                 // 0 = SGBM, 1 = BM
    b2s_sgbm_params sp;
    b2s_bm_params bp;
    // workspace (device), grown on demand
    void* ws = nullptr;
    size_t ws_bytes = 0;
    int* flag = nullptr;      // overflow flag (device)
    const uint16_t* last_cost = nullptr;
    const uint16_t* last_S = nullptr;
    size_t last_elems = 0;
    cudaStream_t last_stream = nullptr;
    // opt-in stage timing: events recorded on the compute stream between stages
    bool keep_volumes = false;   // debug tap: the fused WTA pass also writes the full S volume
    bool timing = false;
    cudaEvent_t ev[B2S_NSTAGES + 1] = {};
    cudaEvent_t ev_mid = nullptr;   // between the two aggregation sweeps
    bool mid_valid = false;
    bool ev_valid = false;
};

// zeroes the last `n16` uint4 of every pair's slab (slabs `pitch16` uint4 apart): MODE_HH4's bottom cost rows
__global__ void k_zero_tail(uint4* p, size_t pitch16, size_t n16)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n16) p[(size_t)blockIdx.y * pitch16 + i] = make_uint4(0u, 0u, 0u, 0u);
}

static size_t align_up(size_t v, size_t a = 256) { return (v + a - 1) / a * a; }

// ---- SGBM parameter rules -----------------------------------
static int sgbm_effective(const b2s_sgbm_params& p, int H, int W, int batch, SgbmEff& e)
{
    if (p.numDisparities <= 0) { b2s_set_error("StereoSGBM: numDisparities must be positive"); return B2S_E_INVALID; }
    if (p.mode < B2S_MODE_SGBM || p.mode > B2S_MODE_HH4) {
        b2s_set_error("StereoSGBM: unknown mode %d", p.mode);
        return B2S_E_INVALID;
    }
    e.H = H; e.W = W; e.batch = batch;
    e.minD = p.minDisparity; e.D = p.numDisparities;
    const int bs = p.blockSize > 0 ? p.blockSize : 5;
    e.SW2 = bs / 2;
    const int chk_sw2 = bs / 2;          // the size test below is common code in cv2: always the 5-pixel fallback
    // MODE_SGBM_3WAY has its own fallback: blockSize <= 0 means a 3-pixel window there (probed against cv2 4.13.0)
    if (p.mode == B2S_MODE_SGBM_3WAY && p.blockSize <= 0) e.SW2 = 1;
    e.ov_raw = p.blockSize / 2 + 1;
    e.P1 = p.P1 > 0 ? p.P1 : 2;
    const int p2 = p.P2 > 0 ? p.P2 : 5;
    e.P2 = p2 > e.P1 + 1 ? p2 : e.P1 + 1;
    e.uniq = p.uniquenessRatio >= 0 ? p.uniquenessRatio : 10;
    e.lrmax = p.disp12MaxDiff > 0 ? p.disp12MaxDiff : 1;
    e.ftzero = (p.preFilterCap > 15 ? p.preFilterCap : 15) | 1;
    const int maxD = e.minD + e.D;
    e.minX1 = maxD > 0 ? maxD : 0;
    e.maxX1 = W + (e.minD < 0 ? e.minD : 0);
    e.width1 = e.maxX1 - e.minX1;
    e.invalid = (e.minD - 1) * 16;
    e.ndirs = p.mode == B2S_MODE_HH ? 8 : (p.mode == B2S_MODE_HH4 ? 4 : (p.mode == B2S_MODE_SGBM_3WAY ? 3 : 5));
    e.cn = 1; e.Hsrc = H; e.nst = 1; e.sz = e.ov = 0; e.uqle = 0; e.way3 = 0;
    // cv2 4.13.0 stereosgbm.cpp:511 raises cv2.error on this condition
    // MODE_HH4 is separate code in cv2 and has no such test: it computes whatever width is left (bit-exact with the
    // restatement down to one column; all INVALID when nothing is left).  MODE_SGBM_3WAY only refuses W - maxD <= 0, but
    // its output for 0 < W - maxD <= blockSize/2 does not follow from the algorithm: refused here like in the other modes.
    if (p.mode != B2S_MODE_HH4 && !(W - maxD > chk_sw2)) {
        b2s_set_error("StereoSGBM: Your input images are too small for your window size and max disparity "
                      "(expected: 'width - (params.minDisparity + params.numDisparities) > "
                      "params.calcSADWindowSize().width/2'), where %d must be greater than %d", W - maxD, chk_sw2);
        return B2S_E_INVALID;
    }
    if (e.ftzero > 127) { b2s_set_error("StereoSGBM: preFilterCap must be <= 127"); return B2S_E_UNSUPPORTED; }
    if (e.P2 > 32767) { b2s_set_error("StereoSGBM: P2 must be <= 32767 (int16 cost domain)"); return B2S_E_UNSUPPORTED; }
    // lanes per pixel / registers per lane
    // G lanes x R packed registers cover Dp = 2*R*G >= D disparities.  R <= 8: fused wavefront
    // aggregation (sgbm_wave); R in {12, 16} (D > 512): one launch per direction (sgbm_agg).
    static const int RS[] = { 2, 4, 6, 8 };
    e.G = 0;
    for (int g = 8; g <= 32 && !e.G; g *= 2)
        for (int r : RS) if (2 * r * g >= e.D) { e.G = g; e.R = r; break; }
    if (!e.G) for (int r : { 12, 16 }) if (64 * r >= e.D) { e.G = 32; e.R = r; break; }
    if (!e.G) { b2s_set_error("StereoSGBM: numDisparities > 1024 is not supported"); return B2S_E_UNSUPPORTED; }
    // (measured on B200 at D = 128: G=8/R=8 195 us per pair; G=4/R=16 287 us; G=16/R=4 237 us)
    e.simdD = e.D;
    if (p.mode == B2S_MODE_SGBM_3WAY && e.D % 8 != 0) {
        // cv2's 3WAY handles the last D % 8 disparities in a scalar tail with rules of its own: those
        // numDisparities run the per-direction kernels (sgbm_agg.cu) and k_wta, 32 lanes per pixel, natural cell order
        e.simdD = 8 * (e.D / 8);
        e.G = 32;
        for (int r : { 2, 4, 8, 12, 16 }) if (64 * r >= e.D) { e.R = r; break; }
    }
    e.Dp = 2 * e.R * e.G;
    e.swz = (e.R == 8 && e.simdD == e.D) ? 1 : 0;   // wavefront path with two 16-byte chunks per lane: permuted records
    return B2S_OK;
}

// The geometry of the cost / S volumes: the image itself, or, for MODE_SGBM_3WAY, cv2's four fixed stripes as
// independent "stripe images" (batch * 4 of them, each stripe_sz + stripe_overlap rows; SgbmEff).
static SgbmEff volume_geometry(const SgbmEff& e, int cn)
{
    SgbmEff v = e;
    v.cn = cn;
    if (e.ndirs == 3) {
        v.nst = 4;
        v.sz = (e.H + 3) / 4;
        v.ov = e.ov_raw + (v.sz + 9) / 10;                  // SADWindowSize/2 + 1 + ceil(0.1 * stripe_sz), SADWindowSize as given
        v.H = v.sz + v.ov < e.H ? v.sz + v.ov : e.H;
        v.batch = e.batch * 4;
        v.uqle = e.uniq > 0 ? 1 : 0;
        v.way3 = 1;
    }
    return v;
}

struct SgbmLayout { size_t planar, descL, descR, cost, S, Sdbg, raw, rawv, key, rec, edge, labels, counts, total; };

static bool uses_wave(const SgbmEff& e) { return e.R <= 8 && e.simdD == e.D; }

// e: the images (batch pairs of H rows); v: volume_geometry(e, cn)
static SgbmLayout sgbm_layout(const SgbmEff& e, const SgbmEff& v, bool speckle, bool keep_volumes)
{
    SgbmLayout l;
    const size_t npix = (size_t)e.batch * e.H * e.W, npixv = (size_t)v.batch * v.H * v.W;
    const size_t vol = v.width1 > 0 ? (size_t)v.batch * v.H * v.width1 * v.Dp * sizeof(uint16_t) : 0;
    SgbmEff ed = e;
    ed.batch = e.batch * v.cn;                               // one gray image per colour channel
    size_t o = 0;
    l.planar = o; o += v.cn > 1 ? 2 * align_up(npix * v.cn) : 0;
    l.descL = o; o += align_up(npix * v.cn * sizeof(uint2));
    l.descR = o; o += align_up(b2s_desc_right_bytes(ed));
    l.cost = o; o += align_up(vol);
    l.S = o; o += align_up(vol);
    l.Sdbg = o; o += (keep_volumes && uses_wave(e)) ? align_up(vol) : 0;
    l.raw = o; o += align_up(npix * sizeof(int16_t));
    l.rawv = o; o += v.nst > 1 ? align_up(npixv * sizeof(int16_t)) : 0;
    l.key = o; o += align_up(npixv * sizeof(uint32_t));
    l.rec = o; o += uses_wave(e) ? align_up(npixv * sizeof(uint2)) : 0;
    l.edge = o; o += (uses_wave(e) && v.width1 > 0) ? align_up(b2s_wave_edge_bytes(v)) : 0;
    l.labels = o; o += speckle ? align_up(npix * sizeof(int32_t)) : 0;
    l.counts = o; o += speckle ? align_up(npix * sizeof(int32_t)) : 0;
    l.total = o;
    return l;
}

// B2S_POISON=<byte value> | random (debug aid; compute-sanitizer's initcheck is not available on the GPU pool): the whole workspace
// is filled with that byte before every compute, so that a kernel that reads memory nobody wrote gives a wrong result
// every time instead of once in a thousand runs.  tests/test_gpu_poison.py runs the parity vectors under several patterns.
static int poison_byte()       // -1: off; 0..255: that byte; 256: pseudo-random bytes (B2S_POISON=random)
{
    static int v = -2;
    if (v == -2) {
        const char* s = getenv("B2S_POISON");
        v = !(s && *s) ? -1 : (!strcmp(s, "random") ? 256 : (atoi(s) & 0xff));
    }
    return v;
}

__global__ void k_poison_random(uint32_t* p, size_t n, uint32_t salt)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint32_t x = (uint32_t)i * 2654435761u + salt;          // a cheap integer hash: unrelated values in neighbouring words
    x ^= x >> 15; x *= 2246822519u; x ^= x >> 13; x *= 3266489917u; x ^= x >> 16;
    p[i] = x;
}

int b2s_poison(void* ws, size_t bytes, cudaStream_t st)
{
    const int pb = poison_byte();
    if (pb < 0 || !ws || !bytes) return B2S_OK;
    if (pb < 256) { B2S_CUDA_CHECK(cudaMemsetAsync(ws, pb, bytes, st)); return B2S_OK; }
    static uint32_t salt = 1;
    const size_t n = bytes / 4;
    k_poison_random<<<(unsigned)((n + 255) / 256), 256, 0, st>>>((uint32_t*)ws, n, salt++ * 0x9e3779b9u);
    B2S_CUDA_CHECK(cudaGetLastError());
    return B2S_OK;
}

static int ensure_ws(b2s_matcher* h, size_t bytes, cudaStream_t st)
{
    if (!h->flag) B2S_CUDA_CHECK(cudaMalloc(&h->flag, sizeof(int)));
    if (bytes > h->ws_bytes) {
        if (h->ws) {
            B2S_CUDA_CHECK(cudaStreamSynchronize(st));
            if (h->last_stream && h->last_stream != st) B2S_CUDA_CHECK(cudaStreamSynchronize(h->last_stream));
            cudaFree(h->ws);
            h->ws = nullptr; h->ws_bytes = 0;
        }
        cudaError_t err = cudaMalloc(&h->ws, bytes);
        if (err != cudaSuccess) {
            cudaGetLastError();
            b2s_set_error("workspace allocation of %zu bytes failed: %s", bytes, cudaGetErrorString(err));
            return B2S_E_NOMEM;
        }
        h->ws_bytes = bytes;
    }
    return b2s_poison(h->ws, h->ws_bytes, st);
}

static int check_images(const char* who, const void* l, const void* r, const void* d, int H, int W, int batch)
{
    if (!l || !r || !d) { b2s_set_error("%s: null image pointer", who); return B2S_E_INVALID; }
    if (H <= 0 || W <= 0 || batch <= 0) { b2s_set_error("%s: empty image (H=%d W=%d batch=%d)", who, H, W, batch); return B2S_E_INVALID; }
    if ((long long)batch * H * W >= (1ll << 31)) { b2s_set_error("%s: batch*H*W must be < 2^31", who); return B2S_E_UNSUPPORTED; }
    if (W > 32768 || H > 65535 || batch > 65535) { b2s_set_error("%s: image or batch too large", who); return B2S_E_UNSUPPORTED; }
    return B2S_OK;
}

extern "C" int b2s_sgbm_create(const b2s_sgbm_params* params, b2s_handle* out)
{
    if (!params || !out) { b2s_set_error("sgbm_create: null argument"); return B2S_E_INVALID; }
    b2s_matcher* h = new (std::nothrow) b2s_matcher();
    if (!h) { b2s_set_error("out of host memory"); return B2S_E_NOMEM; }
    h->kind = 0; h->sp = *params;
    *out = h;
    return B2S_OK;
}

extern "C" int b2s_sgbm_set_params(b2s_handle h, const b2s_sgbm_params* params)
{
    if (!h || h->kind != 0 || !params) { b2s_set_error("sgbm_set_params: bad handle"); return B2S_E_INVALID; }
    h->sp = *params;
    return B2S_OK;
}

extern "C" int b2s_sgbm_get_geometry(b2s_handle h, int W, b2s_sgbm_geometry* out)
{
    if (!h || h->kind != 0 || !out) { b2s_set_error("sgbm_get_geometry: bad handle"); return B2S_E_INVALID; }
    SgbmEff e;
    int rc = sgbm_effective(h->sp, 1, W, 1, e);
    if (rc) return rc;
    out->minX1 = e.minX1; out->maxX1 = e.maxX1; out->width1 = e.width1; out->Dp = e.Dp;
    out->invalid = e.invalid; out->P1 = e.P1; out->P2 = e.P2; out->ftzero = e.ftzero;
    out->uniquenessRatio = e.uniq; out->disp12MaxDiff = e.lrmax; out->ndirs = e.ndirs;
    out->R = e.R; out->G = e.G; out->swizzled = e.swz;
    return B2S_OK;
}

extern "C" size_t b2s_sgbm_workspace_bytes(b2s_handle h, int H, int W, int batch)
{
    if (!h || h->kind != 0) return 0;
    SgbmEff e;
    if (sgbm_effective(h->sp, H, W, batch, e)) return 0;
    return sgbm_layout(e, volume_geometry(e, 1), h->sp.speckleWindowSize > 0, h->keep_volumes).total;
}

__global__ void k_fill16(int16_t* p, size_t n, int16_t v)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) p[i] = v;
}

// interleaved colour images [n][3] -> three gray planes per image [n / (H*W)][3][H*W]
__global__ void k_split3(const uint8_t* __restrict__ src, size_t npix, size_t plane, uint8_t* __restrict__ dst)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= npix) return;
    const size_t b = i / plane, o = i - b * plane;
    const uint8_t* p = src + 3 * i;
    uint8_t* d = dst + b * 3 * plane + o;
    d[0] = p[0]; d[plane] = p[1]; d[2 * plane] = p[2];
}

extern "C" int b2s_sgbm_compute_cn(b2s_handle h, const uint8_t* left, const uint8_t* right, int H, int W, int cn, int batch,
                                   int16_t* disp, void* stream)
{
    if (!h || h->kind != 0) { b2s_set_error("sgbm_compute: bad handle"); return B2S_E_INVALID; }
    if (cn != 1 && cn != 3) { b2s_set_error("StereoSGBM: images must have 1 or 3 channels (got %d)", cn); return B2S_E_INVALID; }
    int rc = check_images("StereoSGBM", left, right, disp, H, W, batch);
    if (rc) return rc;
    SgbmEff e;
    rc = sgbm_effective(h->sp, H, W, batch, e);
    if (rc) return rc;
    const SgbmEff v = volume_geometry(e, cn);          // the cost / S volumes (stripe images for MODE_SGBM_3WAY)
    if (v.nst > 1 && v.ov < 0) {
        b2s_set_error("StereoSGBM: MODE_SGBM_3WAY with blockSize %d makes cv2's stripe overlap negative (cv2 itself runs out of "
                      "bounds there)", h->sp.blockSize);
        return B2S_E_UNSUPPORTED;
    }
    if ((long long)batch * cn * H * W >= (1ll << 31) || (long long)v.batch * v.H * W >= (1ll << 31)) {
        b2s_set_error("StereoSGBM: batch*channels*H*W must be < 2^31");
        return B2S_E_UNSUPPORTED;
    }
    cudaStream_t st = (cudaStream_t)stream;
    const size_t npix = (size_t)batch * H * W;
    h->last_cost = h->last_S = nullptr;
    if (e.width1 <= 0) {                       // nothing computable: whole map INVALID (A.0)
        k_fill16<<<b2s_cdiv((long long)npix, 256), 256, 0, st>>>(disp, npix, (int16_t)e.invalid);
        B2S_LAUNCH_CHECK();
        return B2S_OK;
    }
    const bool speckle = h->sp.speckleWindowSize > 0;
    const SgbmLayout l = sgbm_layout(e, v, speckle, h->keep_volumes);
    rc = ensure_ws(h, l.total, st);
    if (rc) return rc;
    char* ws = (char*)h->ws;
    uint2* descL = (uint2*)(ws + l.descL);
    void* descR = (void*)(ws + l.descR);         // uint2 per pixel or reversed u16 planes (fast cost kernel)
    uint16_t* cost = (uint16_t*)(ws + l.cost);
    uint16_t* S = (uint16_t*)(ws + l.S);
    int16_t* raw = (int16_t*)(ws + l.raw);
    int16_t* rawv = v.nst > 1 ? (int16_t*)(ws + l.rawv) : raw;     // per-stripe rows before they are put together
    uint32_t* key = (uint32_t*)(ws + l.key);

    B2S_CUDA_CHECK(cudaMemsetAsync(h->flag, 0, sizeof(int), st));
    h->ev_valid = false; h->mid_valid = false;
#define B2S_MARK(i) do { if (h->timing) B2S_CUDA_CHECK(cudaEventRecord(h->ev[i], st)); } while (0)
    B2S_MARK(0);
    SgbmEff ed = e;                              // descriptors: every colour channel is a gray image of its own
    ed.batch = batch * cn;
    if (cn == 3) {
        uint8_t* pl = (uint8_t*)(ws + l.planar);
        uint8_t* pr = pl + align_up(npix * 3);
        k_split3<<<b2s_cdiv((long long)npix, 256), 256, 0, st>>>(left, npix, (size_t)H * W, pl);
        B2S_LAUNCH_CHECK();
        k_split3<<<b2s_cdiv((long long)npix, 256), 256, 0, st>>>(right, npix, (size_t)H * W, pr);
        B2S_LAUNCH_CHECK();
        left = pl; right = pr;
    }
    if ((rc = b2s_launch_descriptors(left, ed, descL, st))) return rc;
    if ((rc = b2s_launch_descriptors_right(right, ed, descR, st))) return rc;
    B2S_MARK(1);
    if ((rc = b2s_launch_cost(descL, descR, v, cost, h->flag, st))) return rc;
    if (e.ndirs == 4 && e.SW2 > 0 && H > 1) {
        // cv2 4.13.0's MODE_HH4 never fills the block cost of the rows whose window reaches below the image (its
        // copy of the row loop has no branch for them): C stays at P2 there, i.e. raw cost 0 (probed against cv2,
        // tests/test_oracle_vs_cv2.py).  Rows y >= max(1, H - SW2) of every pair:
        const int y0 = H - e.SW2 > 1 ? H - e.SW2 : 1;
        const size_t rowb = (size_t)e.width1 * e.Dp * sizeof(uint16_t);
        const size_t n16 = (size_t)(H - y0) * rowb / 16;           // Dp is a multiple of 32 cells: whole uint4s
        dim3 grid((unsigned)b2s_cdiv((long long)n16, 256), (unsigned)batch);
        k_zero_tail<<<grid, 256, 0, st>>>(reinterpret_cast<uint4*>(cost + (size_t)y0 * e.width1 * e.Dp), (size_t)H * rowb / 16, n16);
        B2S_LAUNCH_CHECK();
    }
    B2S_MARK(2);
    uint16_t* Sfull = S;
    if (uses_wave(e)) {
        // two fused sweeps; the second one ends in the winner-take-all (stage 2 = both sweeps)
        Sfull = h->keep_volumes ? (uint16_t*)(ws + l.Sdbg) : nullptr;
        if ((rc = b2s_launch_wave(v, cost, S, Sfull, ws + l.edge, (uint2*)(ws + l.rec), key, st, h->timing ? h->ev_mid : nullptr))) return rc;
        h->mid_valid = h->timing;
        B2S_MARK(3);
        if ((rc = b2s_launch_finish_rec(v, (const uint2*)(ws + l.rec), key, rawv, st))) return rc;
    } else {
        // one launch per direction (table of b2s_dir): MODE_SGBM 0-4, MODE_HH 0-7, MODE_HH4 -> , down, <- , up,
        // MODE_SGBM_3WAY -> , down, <-
        const int last = e.ndirs == 4 ? 6 : (e.ndirs == 3 ? 4 : e.ndirs - 1), stepk = (e.ndirs == 4 || e.ndirs == 3) ? 2 : 1;
        for (int k = 0; k <= last; k += stepk)
            if ((rc = b2s_launch_aggregate(cost, S, v, k, k == 0, st))) return rc;
        B2S_MARK(3);
        if ((rc = b2s_launch_wta(S, v, rawv, key, st))) return rc;
    }
    if (v.nst > 1 && (rc = b2s_launch_assemble_stripes(v, rawv, raw, st))) return rc;
    B2S_MARK(4);
    if ((rc = b2s_launch_median3(raw, H, W, batch, disp, st))) return rc;
    B2S_MARK(5);
    if (speckle) {
        rc = b2s_launch_speckles(disp, H, W, batch, e.invalid, h->sp.speckleWindowSize, 16 * h->sp.speckleRange,
                                 (int32_t*)(ws + l.labels), (int32_t*)(ws + l.counts), st);
        if (rc) return rc;
    }
    B2S_MARK(6);
#undef B2S_MARK
    h->ev_valid = h->timing;
    h->last_cost = cost; h->last_S = Sfull; h->last_stream = st;
    h->last_elems = (size_t)v.batch * v.H * v.width1 * v.Dp;
    return B2S_OK;
}

extern "C" int b2s_sgbm_compute(b2s_handle h, const uint8_t* left, const uint8_t* right, int H, int W, int batch,
                                int16_t* disp, void* stream)
{
    return b2s_sgbm_compute_cn(h, left, right, H, W, 1, batch, disp, stream);
}

extern "C" int b2s_sgbm_debug_volumes(b2s_handle h, uint16_t* cost_out, uint16_t* S_out, size_t elems, void* stream)
{
    if (!h || h->kind != 0 || !h->last_cost) { b2s_set_error("debug_volumes: no completed compute on this handle"); return B2S_E_INVALID; }
    if (S_out && !h->last_S) { b2s_set_error("debug_volumes: the S volume is only kept after b2s_sgbm_keep_volumes(h, 1)"); return B2S_E_INVALID; }
    if (elems > h->last_elems) { b2s_set_error("debug_volumes: %zu elements requested, volume has %zu", elems, h->last_elems); return B2S_E_INVALID; }
    cudaStream_t st = (cudaStream_t)stream;
    if (cost_out) B2S_CUDA_CHECK(cudaMemcpyAsync(cost_out, h->last_cost, elems * sizeof(uint16_t), cudaMemcpyDeviceToDevice, st));
    if (S_out) B2S_CUDA_CHECK(cudaMemcpyAsync(S_out, h->last_S, elems * sizeof(uint16_t), cudaMemcpyDeviceToDevice, st));
    return B2S_OK;
}

extern "C" int b2s_sgbm_keep_volumes(b2s_handle h, int on)
{
    if (!h || h->kind != 0) { b2s_set_error("keep_volumes: bad handle"); return B2S_E_INVALID; }
    h->keep_volumes = on != 0;
    return B2S_OK;
}

extern "C" int b2s_sgbm_overflow_flag(b2s_handle h, int* flag_host)
{
    if (!h || h->kind != 0 || !flag_host) { b2s_set_error("overflow_flag: bad argument"); return B2S_E_INVALID; }
    *flag_host = 0;
    if (!h->flag) return B2S_OK;
    B2S_CUDA_CHECK(cudaStreamSynchronize(h->last_stream));
    B2S_CUDA_CHECK(cudaMemcpy(flag_host, h->flag, sizeof(int), cudaMemcpyDeviceToHost));
    return B2S_OK;
}

// ---- BM -------------------------------------------------------------------
static int bm_validate(const b2s_bm_params& p)
{
    if (p.numDisparities <= 0 || p.numDisparities % 16 != 0) {
        b2s_set_error("StereoBM: numDisparities must be positive and divisible by 16");
        return B2S_E_INVALID;
    }
    if (p.blockSize % 2 == 0 || p.blockSize < 5 || p.blockSize > 255) {
        b2s_set_error("StereoBM: SADWindowSize must be odd, be within 5..255 and be not larger than image width or height");
        return B2S_E_INVALID;
    }
    if (p.preFilterCap < 1 || p.preFilterCap > 63) { b2s_set_error("StereoBM: preFilterCap must be within 1..63"); return B2S_E_INVALID; }
    if (p.textureThreshold < 0) { b2s_set_error("StereoBM: texture threshold must be non-negative"); return B2S_E_INVALID; }
    if (p.uniquenessRatio < 0) { b2s_set_error("StereoBM: uniqueness ratio must be non-negative"); return B2S_E_INVALID; }
    if (p.disp12MaxDiff >= 0 && p.disp12MaxDiff < 4096) {
        b2s_set_error("StereoBM: disp12MaxDiff in 0..4095 (cv2's validateDisparity) is not supported; use -1 (off, the cv2 "
                      "default) or a value that never rejects (cv2.ximgproc's factories set 1000000)");
        return B2S_E_UNSUPPORTED;
    }
    if (p.minDisparity < -32000 || p.minDisparity > 32000) { b2s_set_error("StereoBM: minDisparity out of range"); return B2S_E_INVALID; }
    return B2S_OK;
}

extern "C" int b2s_bm_create(const b2s_bm_params* params, b2s_handle* out)
{
    if (!params || !out) { b2s_set_error("bm_create: null argument"); return B2S_E_INVALID; }
    b2s_matcher* h = new (std::nothrow) b2s_matcher();
    if (!h) { b2s_set_error("out of host memory"); return B2S_E_NOMEM; }
    h->kind = 1; h->bp = *params;
    *out = h;
    return B2S_OK;
}

extern "C" int b2s_bm_compute(b2s_handle h, const uint8_t* left, const uint8_t* right, int H, int W, int batch,
                              int16_t* disp, void* stream)
{
    if (!h || h->kind != 1) { b2s_set_error("bm_compute: bad handle"); return B2S_E_INVALID; }
    int rc = check_images("StereoBM", left, right, disp, H, W, batch);
    if (rc) return rc;
    if ((rc = bm_validate(h->bp))) return rc;
    if (h->bp.blockSize >= W || h->bp.blockSize >= H) {          // cv2 4.13.0 refuses SADWindowSize >= min(width, height) (probed)
        b2s_set_error("StereoBM: SADWindowSize must be odd, be within 5..255 and be not larger than image width or height");
        return B2S_E_INVALID;
    }
    cudaStream_t st = (cudaStream_t)stream;
    const size_t npix = (size_t)batch * H * W;
    // cv2 4.13.0 StereoBM: filterSpeckles(disp, FILTERED, speckleWindowSize, speckleRange) -- the range is NOT scaled by
    // 16 here (only StereoSGBM scales it) and the filter is skipped for a negative range (probed against cv2)
    const bool speckle = h->bp.speckleWindowSize > 0 && h->bp.speckleRange >= 0;
    size_t need = 2 * align_up(npix) + (speckle ? 2 * align_up(npix * sizeof(int32_t)) : 0);
    if ((rc = ensure_ws(h, need, st))) return rc;
    uint8_t* PL = (uint8_t*)h->ws;
    uint8_t* PR = PL + align_up(npix);
    if ((rc = b2s_bm_run(left, right, H, W, batch, h->bp, PL, PR, disp, st))) return rc;
    if (speckle) {
        int32_t* labels = (int32_t*)(PR + align_up(npix));
        int32_t* counts = (int32_t*)((char*)labels + align_up(npix * sizeof(int32_t)));
        rc = b2s_launch_speckles(disp, H, W, batch, (h->bp.minDisparity - 1) * 16, h->bp.speckleWindowSize, h->bp.speckleRange, labels, counts, st);
        if (rc) return rc;
    }
    h->last_stream = st;
    return B2S_OK;
}

extern "C" int b2s_sgbm_enable_timing(b2s_handle h, int on)
{
    if (!h || h->kind != 0) { b2s_set_error("enable_timing: bad handle"); return B2S_E_INVALID; }
    if (on && !h->ev[0])
    {
        for (int i = 0; i <= B2S_NSTAGES; i++) B2S_CUDA_CHECK(cudaEventCreate(&h->ev[i]));
        B2S_CUDA_CHECK(cudaEventCreate(&h->ev_mid));
    }
    h->timing = on != 0;
    h->ev_valid = false;
    return B2S_OK;
}

extern "C" int b2s_sgbm_stage_ms(b2s_handle h, float* ms_host)
{
    if (!h || h->kind != 0 || !ms_host) { b2s_set_error("stage_ms: bad argument"); return B2S_E_INVALID; }
    if (!h->ev_valid) { b2s_set_error("stage_ms: no timed compute on this handle (call b2s_sgbm_enable_timing first)"); return B2S_E_INVALID; }
    B2S_CUDA_CHECK(cudaEventSynchronize(h->ev[B2S_NSTAGES]));
    for (int i = 0; i < B2S_NSTAGES; i++) B2S_CUDA_CHECK(cudaEventElapsedTime(&ms_host[i], h->ev[i], h->ev[i + 1]));
    return B2S_OK;
}

extern "C" int b2s_sgbm_sweep_ms(b2s_handle h, float* ms2_host)
{
    if (!h || h->kind != 0 || !ms2_host) { b2s_set_error("sweep_ms: bad argument"); return B2S_E_INVALID; }
    if (!h->ev_valid) { b2s_set_error("sweep_ms: no timed compute on this handle (call b2s_sgbm_enable_timing first)"); return B2S_E_INVALID; }
    ms2_host[0] = ms2_host[1] = 0.f;
    if (!h->mid_valid) return B2S_OK;
    B2S_CUDA_CHECK(cudaEventSynchronize(h->ev[3]));
    B2S_CUDA_CHECK(cudaEventElapsedTime(&ms2_host[0], h->ev[2], h->ev_mid));
    B2S_CUDA_CHECK(cudaEventElapsedTime(&ms2_host[1], h->ev_mid, h->ev[3]));
    return B2S_OK;
}

// The reference's per-pair chain as one call (disparity_calculation.py:207-208 remap, :286-287 gray,
// :289 compute_disparity, :302 reprojectImageTo3D): every stage is enqueued on `stream`, nothing synchronises.
extern "C" int b2s_pipeline(b2s_handle sgbm, const uint8_t* left_raw, const uint8_t* right_raw, int sH, int sW, int cn,
                            const float* mapx_l, const float* mapy_l, const float* mapx_r, const float* mapy_r,
                            int H, int W, int batch, uint8_t* rect_l, uint8_t* rect_r, int16_t* disp,
                            const double* Q_host, float disp_scale, int handle_missing, float* scratch, float* xyz,
                            void* stream)
{
    if (!sgbm || sgbm->kind != 0) { b2s_set_error("pipeline: bad StereoSGBM handle"); return B2S_E_INVALID; }
    if (!rect_l || !rect_r) { b2s_set_error("pipeline: rect_l / rect_r scratch buffers are required"); return B2S_E_INVALID; }
    if (cn != 1 && cn != 3) { b2s_set_error("pipeline: 1 or 3 channels supported, got %d", cn); return B2S_E_UNSUPPORTED; }
    int rc;
    if (cn == 1) {
        if ((rc = b2s_remap_u8(left_raw, sH, sW, 1, mapx_l, mapy_l, H, W, batch, rect_l, stream))) return rc;
        if ((rc = b2s_remap_u8(right_raw, sH, sW, 1, mapx_r, mapy_r, H, W, batch, rect_r, stream))) return rc;
    } else {
        if ((rc = b2s_remap_gray_u8(left_raw, sH, sW, mapx_l, mapy_l, H, W, batch, rect_l, stream))) return rc;
        if ((rc = b2s_remap_gray_u8(right_raw, sH, sW, mapx_r, mapy_r, H, W, batch, rect_r, stream))) return rc;
    }
    if ((rc = b2s_sgbm_compute(sgbm, rect_l, rect_r, H, W, batch, disp, stream))) return rc;
    if (xyz) {
        if (!Q_host) { b2s_set_error("pipeline: Q is required when xyz is requested"); return B2S_E_INVALID; }
        if ((rc = b2s_reproject(disp, B2S_DT_I16, H, W, batch, Q_host, disp_scale, handle_missing, scratch, xyz, stream))) return rc;
    }
    return B2S_OK;
}

extern "C" void b2s_destroy(b2s_handle h)
{
    if (!h) return;
    for (int i = 0; i <= B2S_NSTAGES; i++) if (h->ev[i]) cudaEventDestroy(h->ev[i]);
    if (h->ev_mid) cudaEventDestroy(h->ev_mid);
    if (h->ws) cudaFree(h->ws);
    if (h->flag) cudaFree(h->flag);
    delete h;
}

// ---- stand-alone stages -----------------------------------------------------
extern "C" int b2s_median3_i16(const int16_t* src, int H, int W, int batch, int16_t* dst, void* stream)
{
    if (!src || !dst || H <= 0 || W <= 0 || batch <= 0) { b2s_set_error("median3: bad argument"); return B2S_E_INVALID; }
    if (src == dst) { b2s_set_error("median3: in-place operation is not supported"); return B2S_E_INVALID; }
    return b2s_launch_median3(src, H, W, batch, dst, (cudaStream_t)stream);
}

extern "C" int b2s_filter_speckles(int16_t* img, int H, int W, int batch, int newVal, int maxSpeckleSize, int maxDiff,
                                   int32_t* labels, int32_t* counts, void* stream)
{
    if (!img || !labels || !counts || H <= 0 || W <= 0 || batch <= 0) { b2s_set_error("filterSpeckles: bad argument"); return B2S_E_INVALID; }
    return b2s_launch_speckles(img, H, W, batch, newVal, maxSpeckleSize, maxDiff, labels, counts, (cudaStream_t)stream);
}
