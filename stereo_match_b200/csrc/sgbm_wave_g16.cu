// sgbm_wave_g16.cu -- instantiations of the wavefront aggregation kernel for G = 16 lanes per pixel.
#include "sgbm_wave_impl.cuh"
int b2s_launch_wave_g16(const SgbmEff& e, const WaveArgs& a, int vert, int final, cudaStream_t st)
{
    return wave::launch_g<16, 256>(e, a, vert, final, st);
}
