// pda_kernels_ref.cu — reference-draws instantiation of the evaluation kernel (sm_100a,
// compiled with -fmad=false like the replay unit).
//
// Same kernel, same explicit-rounding arithmetic as replay mode, but the uniform words come from
// where the reference takes them: one cuRAND scrambled Sobol64 sequence per thread (dimension
// tidx % 20000, offset 1234, /root/reference/src/mc.cu:366-455, gen_rand.cuh:7-25), drawn in the
// reference's order and only when the reference draws, and an xor128 stream per thread seeded
// once from the first Sobol word; both persist from one evaluation to the next (quirk Q8).
// With identical draws the per-sample efficiencies, histogram and chi-square equal those of the
// reference's own class mc on the same GPU bit for bit (tests/test_gpu_reference_mode.py).
#include "pda_eval.cuh"

namespace pda {

__global__ void pda_setup_ref_kernel(const unsigned long long* __restrict__ dv,
                                     const unsigned long long* __restrict__ scramble, long long NN,
                                     ulonglong2* sob_state, uint4* xs_state)
{
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= NN) return;
    const unsigned dim = (unsigned)(t % 20000);
    const unsigned long long* d = dv + 64u * dim;
    // curand_init(direction_vectors, scramble_c, offset = 1234, state) + skipahead
    unsigned long long i = 1234ull, x = scramble[dim];
    const unsigned long long gray = i ^ (i >> 1);
    for (int k = 0; k < 64; ++k)
        if (gray & (1ull << k)) x ^= d[k];
    // llseed = curand(state); rk_seed(llseed, rk_state)
    const unsigned long long llseed = x;
    x ^= d[__ffsll((long long)~i) - 1];
    i++;
    Xor128 g;
    g.seed(llseed);
    sob_state[t] = make_ulonglong2(i, x);
    xs_state[t] = make_uint4(g.s0, g.s1, g.s2, g.s3);
}

LaunchCfg eval_launch_cfg_ref(int device, int layout, bool with_bg, int nb)
{
    return eval_launch_cfg_t<kModeReference>(device, layout, with_bg, nb);
}

cudaError_t launch_eval_ref(const EvalArgs& a, int layout, bool with_bg, LaunchCfg cfg, cudaStream_t s, bool dense)
{
    return launch_eval_t<kModeReference>(a, layout, with_bg, cfg, s, dense);
}

cudaError_t launch_setup_ref(const unsigned long long* dv, const unsigned long long* scramble, long long NN,
                             ulonglong2* sob_state, uint4* xs_state, cudaStream_t s)
{
    if (NN <= 0) return cudaSuccess;
    pda_setup_ref_kernel<<<(unsigned)((NN + 255) / 256), 256, 0, s>>>(dv, scramble, NN, sob_state, xs_state);
    return cudaGetLastError();
}

}  // namespace pda
