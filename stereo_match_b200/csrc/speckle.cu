// speckle.cu -- cv2.filterSpeckles as parallel connected-component labelling.
// Runs inside StereoSGBM.compute / StereoBM.compute when speckleWindowSize > 0 (reference call
// site stereo_vision/stereo_vision.py:178) and stand-alone through b2s_filter_speckles.
//
// cv2's sequential flood fill is order independent, so any exact 4-connected labelling under
// the predicate |a-b| <= maxDiff removes the same pixels.  Four kernels:
//   runs   : every pixel gets the index of the first pixel of its horizontal run (block-wide
//            max-scan along the row) -> horizontal connectivity costs no atomics at all
//   merge  : vertical edges are unioned (atomicMin union-find on run starts), skipping an edge
//            whose left neighbour edge already joins the same two runs
//   count  : the last pixel of every run adds the run length to its root and compresses
//   apply  : components of size <= maxSpeckleSize become newVal
// labels/counts: int32 per pixel (labels: -1 = invalid pixel, else a pixel index <= own index).
#include "common.cuh"

__device__ __forceinline__ int uf_find(int32_t* L, int i)
{
    // path halving; every value ever stored in L[i] is an ancestor-or-self with a smaller index
    int p = __ldcg(L + i);
    while (p != i) {
        int gp = __ldcg(L + p);
        if (gp != p) L[i] = gp;
        i = p; p = gp;
    }
    return i;
}

__device__ __forceinline__ int uf_find_ro(const int32_t* L, int i)
{
    int p = __ldcg(L + i);
    while (p != i) { i = p; p = __ldcg(L + i); }
    return i;
}

__device__ __forceinline__ void uf_union(int32_t* L, int a, int b)
{
    bool done;
    do {
        a = uf_find(L, a);
        b = uf_find(L, b);
        if (a < b) { int old = atomicMin(&L[b], a); done = (old == b); b = old; }
        else if (b < a) { int old = atomicMin(&L[a], b); done = (old == a); a = old; }
        else done = true;
    } while (!done);
}

#define SPK_NT 256
__global__ void __launch_bounds__(SPK_NT)
k_speckle_runs(const int16_t* __restrict__ img, int H, int W, int newVal, int maxDiff,
               int32_t* __restrict__ labels, int32_t* __restrict__ counts)
{
    __shared__ int warp_max[SPK_NT / 32];
    __shared__ int carry_s;
    const size_t rowbase = ((size_t)blockIdx.y * H + blockIdx.x) * W;
    const int16_t* row = img + rowbase;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    int carry = -1;
    for (int x0 = 0; x0 < W; x0 += SPK_NT) {
        const int x = x0 + threadIdx.x;
        int v = newVal, left = newVal;
        if (x < W) { v = row[x]; if (x > 0) left = row[x - 1]; }
        const bool valid = x < W && v != newVal;
        const bool connL = valid && left != newVal && abs(v - left) <= maxDiff;
        int s = (valid && !connL) ? x : -1;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            int t = __shfl_up_sync(0xffffffffu, s, o);
            if (lane >= o) s = max(s, t);
        }
        if (lane == 31) warp_max[wid] = s;
        __syncthreads();
        int pre = carry;
        for (int w = 0; w < wid; w++) pre = max(pre, warp_max[w]);
        s = max(s, pre);
        if (threadIdx.x == SPK_NT - 1) carry_s = s;
        if (x < W) {
            labels[rowbase + x] = valid ? (int32_t)(rowbase + s) : -1;
            if (valid && !connL) counts[rowbase + x] = 0;     // only run starts can become roots
        }
        __syncthreads();
        carry = carry_s;
    }
}

__global__ void __launch_bounds__(256)
k_speckle_merge(const int16_t* __restrict__ img, int H, int W, int newVal, int maxDiff, int32_t* labels)
{
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    const int y = blockIdx.y;
    if (x >= W || y + 1 >= H) return;
    const size_t i = ((size_t)blockIdx.z * H + y) * W + x;
    const int v = img[i], w = img[i + W];
    if (v == newVal || w == newVal || abs(v - w) > maxDiff) return;
    bool startA = true, startB = true;
    if (x > 0) {
        const int vl = img[i - 1], wl = img[i + W - 1];
        startA = !(vl != newVal && abs(v - vl) <= maxDiff);
        startB = !(wl != newVal && abs(w - wl) <= maxDiff);
        if (!startA && !startB && abs(vl - wl) <= maxDiff)
            return;                                   // the edge to the left joins the same two runs
    }
    // union-find runs on run starts only, so the label of a non-start pixel stays its run start
    const int sa = startA ? (int)i : __ldcg(labels + i);
    const int sb = startB ? (int)(i + W) : __ldcg(labels + i + W);
    uf_union(labels, sa, sb);
}

__global__ void __launch_bounds__(256)
k_speckle_count(const int16_t* __restrict__ img, int H, int W, int newVal, int maxDiff, int32_t* labels,
                int32_t* counts)
{
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    if (x >= W) return;
    const size_t i = ((size_t)blockIdx.z * H + blockIdx.y) * W + x;
    const int v = img[i];
    if (v == newVal) return;
    if (x + 1 < W) {                                  // only the last pixel of a run works
        const int r = img[i + 1];
        if (r != newVal && abs(v - r) <= maxDiff) return;
    }
    bool start = true;
    if (x > 0) { const int l = img[i - 1]; start = !(l != newVal && abs(v - l) <= maxDiff); }
    const int s = start ? (int)i : __ldcg(labels + i);
    const int root = uf_find_ro(labels, s);
    atomicAdd(&counts[root], (int)i - s + 1);
    if (root != s) labels[s] = root;                  // every run start now points at its root
}

// Only small components change, and a small component only has short runs: the last pixel of every run
// looks up its component size once and, if it is a speckle, clears its own run.
__global__ void __launch_bounds__(256)
k_speckle_apply(int16_t* img, int H, int W, int newVal, int maxDiff, int maxSize, const int32_t* __restrict__ labels,
                const int32_t* __restrict__ counts)
{
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    if (x >= W) return;
    const size_t i = ((size_t)blockIdx.z * H + blockIdx.y) * W + x;
    const int v = img[i];
    if (v == newVal) return;
    if (x + 1 < W) {                                  // only the last pixel of a run works
        const int r = img[i + 1];
        if (r != newVal && abs(v - r) <= maxDiff) return;
    }
    const int l = __ldg(labels + i);                  // run start (or, for a start, its root)
    const int root = __ldg(labels + l);               // root of the run start (a root points at itself)
    if (__ldg(counts + root) > maxSize) return;
    // walk back to the run start: l for a non-start pixel; a start pixel has no connected left neighbour
    int j = x;
    int cur = v;
    while (true) {
        const size_t k = i - (size_t)(x - j);
        int left = newVal;
        if (j > 0) left = img[k - 1];
        img[k] = (int16_t)newVal;
        if (j == 0 || left == newVal || abs(cur - left) > maxDiff) break;
        cur = left; j--;
    }
}

int b2s_launch_speckles(int16_t* img, int H, int W, int batch, int newVal, int maxSize, int maxDiff,
                        int32_t* labels, int32_t* counts, cudaStream_t st)
{
    const size_t n = (size_t)batch * H * W;
    if (n >= (size_t)1 << 31) { b2s_set_error("filterSpeckles: batch*H*W must be < 2^31"); return B2S_E_INVALID; }
    dim3 grid(b2s_cdiv(W, 256), H, batch);
    k_speckle_runs<<<dim3(H, batch), SPK_NT, 0, st>>>(img, H, W, newVal, maxDiff, labels, counts);
    B2S_LAUNCH_CHECK();
    k_speckle_merge<<<grid, 256, 0, st>>>(img, H, W, newVal, maxDiff, labels);
    B2S_LAUNCH_CHECK();
    k_speckle_count<<<grid, 256, 0, st>>>(img, H, W, newVal, maxDiff, labels, counts);
    B2S_LAUNCH_CHECK();
    k_speckle_apply<<<grid, 256, 0, st>>>(img, H, W, newVal, maxDiff, maxSize, labels, counts);
    B2S_LAUNCH_CHECK();
    return B2S_OK;
}

// ---- b2s_filter_speckles_noalloc: the same four kernels with a built-in scratch ------------------------------------
// No counterpart in the reference: added for callers that cannot call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  labels + counts live in a file-scope
// device array of a fixed, modest size: 2 x B2S_SPECKLE_NOALLOC_MAX_PIXELS int32 = 8 MiB, which every GPU that loads
// the library holds until the process exits.  Nothing here allocates: no cudaMalloc / cudaFree, no allocating library
// routine.  In host code `g_speckle_scratch` is only a host-side stand-in for the symbol, NOT a device address: the
// address is resolved once per device, with that device current, through cudaGetSymbolAddress and cached.
//
// Every call on a device shares the one array, and the array keeps the last call's contents.  Nothing has to be cleared
// between calls: k_speckle_runs rewrites labels[] for every pixel of the call and zeroes counts[] at every run start,
// and run starts are the only cells of counts[] the later kernels read or add to.  Calls that overlap on the GPU
// (different streams without an ordering between them, concurrent host threads) RACE on the scratch: serialise them.
static __device__ __align__(256) int32_t g_speckle_scratch[2 * (size_t)B2S_SPECKLE_NOALLOC_MAX_PIXELS];

#include <mutex>
static int speckle_scratch_for_current_device(int32_t** out)
{
    constexpr int MAXDEV = 64;
    static int32_t* cached[MAXDEV] = {};
    static std::mutex mu;
    int dev = -1;
    B2S_CUDA_CHECK(cudaGetDevice(&dev));
    if (dev < 0 || dev >= MAXDEV) { b2s_set_error("filterSpeckles (no-alloc): device ordinal %d out of range", dev); return B2S_E_UNSUPPORTED; }
    std::lock_guard<std::mutex> lock(mu);
    if (!cached[dev]) {
        int32_t* p = nullptr;
        B2S_CUDA_CHECK(cudaGetSymbolAddress((void**)&p, g_speckle_scratch));      // the DEVICE address on the current device
        if (!p) { b2s_set_error("filterSpeckles (no-alloc): cudaGetSymbolAddress returned a null address"); return B2S_E_CUDA; }
        cached[dev] = p;
    }
    *out = cached[dev];
    return B2S_OK;
}

extern "C" int b2s_filter_speckles_noalloc(int16_t* img, int H, int W, int batch, int newVal, int maxSpeckleSize, int maxDiff,
                                           void* stream)
{
    if (!img || H <= 0 || W <= 0 || batch <= 0) { b2s_set_error("filterSpeckles (no-alloc): bad argument"); return B2S_E_INVALID; }
    const long long n = (long long)batch * H * W;
    if (n > (long long)B2S_SPECKLE_NOALLOC_MAX_PIXELS) {
        b2s_set_error("filterSpeckles (no-alloc): batch*H*W = %lld pixels exceeds the built-in scratch of %d pixels "
                      "(B2S_SPECKLE_NOALLOC_MAX_PIXELS); use b2s_filter_speckles with caller-provided scratch", n,
                      (int)B2S_SPECKLE_NOALLOC_MAX_PIXELS);
        return B2S_E_UNSUPPORTED;
    }
    int32_t* scratch = nullptr;
    const int rc = speckle_scratch_for_current_device(&scratch);
    if (rc) return rc;
    return b2s_launch_speckles(img, H, W, batch, newVal, maxSpeckleSize, maxDiff, scratch, scratch + B2S_SPECKLE_NOALLOC_MAX_PIXELS,
                               (cudaStream_t)stream);
}
