// sgbm_wave_g8s.cu -- the G = 8 wavefront kernels once more with 128 threads (16 rows) per CTA.
#include "sgbm_wave_impl.cuh"
int b2s_launch_wave_g8s(const SgbmEff& e, const WaveArgs& a, int vert, int final, cudaStream_t st)
{
    return wave::launch_g<8, 128>(e, a, vert, final, st);
}
