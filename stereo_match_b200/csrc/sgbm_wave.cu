// sgbm_wave.cu -- host side of the fused wavefront aggregation (see sgbm_wave.cuh).
#include "sgbm_wave.cuh"

// L_r <= C + P2 and C <= blockSize^2 * (2*ftzero + 63) (Sobel channel BT cost <= 2*ftzero, raw channel <= 255 >> 2):
// if S1 (<= 32767) plus four such terms cannot reach 65536, the packed sums never carry between the halves
static int wave_fastacc(const SgbmEff& e)
{
    const long long bs = 2 * e.SW2 + 1;
    const long long lmax = bs * bs * (2LL * e.ftzero + 63) * e.cn + e.P2;     // cn colour channels add up
    return (4 * lmax <= 32768) ? 1 : 0;
}
static int wave_diag(const SgbmEff& e) { return (e.ndirs == 4 || e.ndirs == 3) ? 0 : 1; }

// does the lean kernel (sgbm_wave8.cu) run the sweeps with vertical directions?
static bool wave_lean(const SgbmEff& e)
{
    WaveArgs a;
    a.diag = wave_diag(e); a.fastacc = wave_fastacc(e);
    return b2s_wave8_ok(e, a, 1);
}

int b2s_wave_threads(const SgbmEff& e)
{
    if (wave_lean(e)) return b2s_wave8_rows_per_band(e) * e.G;
    if (e.G != 8) return 256;
    // the 8-path sweeps with the widest lanes run a little faster in 16-row bands (see sgbm_wave.cuh); otherwise take
    // 16-row bands only where 32-row bands would compute at least 5 % more rows (short images: the 3WAY stripes)
    if (e.ndirs == 8 && e.R == 8) return 128;
    const int r32 = (e.H + 31) / 32 * 32, r16 = (e.H + 15) / 16 * 16;
    return (20 * r32 >= 21 * r16) ? 128 : 256;
}

int b2s_wave_rows_per_band(const SgbmEff& e) { return b2s_wave_threads(e) / e.G; }

static size_t al256(size_t v) { return (v + 255) / 256 * 256; }

struct EdgeLayout { size_t edge, edge_min, progress, total; int nbands; };
static EdgeLayout edge_layout(const SgbmEff& e)
{
    EdgeLayout l;
    l.nbands = (e.H + b2s_wave_rows_per_band(e) - 1) / b2s_wave_rows_per_band(e);
    const size_t items = (size_t)e.batch * l.nbands;
    size_t o = 0;
    l.edge = o; o += al256(items * e.width1 * 3 * e.Dp * sizeof(uint16_t));
    l.edge_min = o; o += al256(items * e.width1 * 4 * sizeof(uint32_t));
    l.progress = o; o += al256((items + 1) * sizeof(int));      // progress[items] then the ticket
    l.total = o;
    return l;
}

size_t b2s_wave_edge_bytes(const SgbmEff& e) { return edge_layout(e).total; }

static int launch_pass(const SgbmEff& e, const WaveArgs& a, int vert, int final, cudaStream_t st)
{
    if (b2s_wave8_ok(e, a, vert)) return b2s_launch_wave8(e, a, final, st);
    switch (e.G) {
    case 8: return b2s_wave_threads(e) == 128 ? b2s_launch_wave_g8s(e, a, vert, final, st) : b2s_launch_wave_g8(e, a, vert, final, st);
    case 16: return b2s_launch_wave_g16(e, a, vert, final, st);
    case 32: return b2s_launch_wave_g32(e, a, vert, final, st);
    }
    b2s_set_error("wave: no kernel for G=%d", e.G);
    return B2S_E_UNSUPPORTED;
}

// cost -> (S1) -> WTA records + disp2 keys.  S holds S1 after the call; S_dbg (optional) the full S.
int b2s_launch_wave(const SgbmEff& e, const uint16_t* cost, uint16_t* S, uint16_t* S_dbg, void* edge_ws,
                    uint2* rec, uint32_t* disp2key, cudaStream_t st, cudaEvent_t between_sweeps)
{
    const EdgeLayout l = edge_layout(e);
    char* ws = (char*)edge_ws;
    const size_t items = (size_t)e.batch * l.nbands;
    WaveArgs a;
    a.cost = cost; a.S = S; a.S_out = nullptr;
    a.edge = (uint16_t*)(ws + l.edge);
    a.edge_min = (uint32_t*)(ws + l.edge_min);
    a.progress = (int*)(ws + l.progress);
    a.ticket = (unsigned int*)(a.progress + items);
    a.rec = rec; a.disp2key = disp2key;
    const int mul = 100 - e.uniq;
    a.magic = mul > 0 ? ((1ull << 40) + (unsigned long long)mul - 1) / (unsigned long long)mul : 0ull;
    a.nbands = l.nbands;
    a.diag = wave_diag(e);
    a.zero = 0u;
    // A band trails the one above by the publication interval plus the row skew.  When all work items are resident at
    // once (small batches: the latency of the band chain is what the caller sees) publish every 8 pixels; otherwise
    // every 32 -- each publication costs the publishing warp a gpu-scope release (+1.5 % per 128 pairs at 8)
    a.pubw = (items < 2 * 148) ? 8 : 32;
    a.fastacc = wave_fastacc(e);
    int rc;
    B2S_CUDA_CHECK(cudaMemsetAsync(disp2key, 0xff, (size_t)e.batch * e.H * e.W * sizeof(uint32_t), st));
    // pass 1: -> , down-right, down, down-left (MODE_HH4, MODE_SGBM_3WAY: -> , down)
    B2S_CUDA_CHECK(cudaMemsetAsync(a.progress, 0, (items + 1) * sizeof(int), st));
    a.flip = 0;
    if ((rc = launch_pass(e, a, 1, 0, st))) return rc;
    if (between_sweeps) B2S_CUDA_CHECK(cudaEventRecord(between_sweeps, st));     // opt-in stage timing (b2s_sgbm_sweep_ms)
    a.flip = 1;
    a.S_out = S_dbg;
    if (e.ndirs == 8 || e.ndirs == 4) {   // MODE_HH: the mirrored sweep adds <- , up-left, up, up-right (MODE_HH4: <- , up)
        B2S_CUDA_CHECK(cudaMemsetAsync(a.progress, 0, (items + 1) * sizeof(int), st));
        rc = launch_pass(e, a, 1, 1, st);
    } else {                     // MODE_SGBM, MODE_SGBM_3WAY: only <-
        rc = launch_pass(e, a, 0, 1, st);
    }
    return rc;
}
