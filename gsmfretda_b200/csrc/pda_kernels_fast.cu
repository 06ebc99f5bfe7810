// pda_kernels_fast.cu — normal-mode instantiation of the evaluation kernel (sm_100a).
//
// Compiled WITHOUT -fmad=false: FP32 arithmetic with fused multiply-adds and the hardware
// lg2/ex2/rcp units (`__logf`, `__expf`, `__fdividef`), binomial by FP32 inversion on a Philox
// word.  Same algorithm and same quirks as the replay instantiation in pda_kernels.cu (role of
// /root/reference/src/mc.cu:83-197, gen_rand.cuh, binom.cuh), statistically equivalent results
// (tests/test_gpu_normal_mode.py: KS p > 0.01, chi-square within 1 %), not the same bits.
#include "pda_eval.cuh"

namespace pda {

LaunchCfg eval_launch_cfg_fast(int device, int layout, bool with_bg, int nb)
{
    return eval_launch_cfg_t<kModeNormal>(device, layout, with_bg, nb);
}

cudaError_t launch_eval_fast(const EvalArgs& a, int layout, bool with_bg, LaunchCfg cfg, cudaStream_t s, bool dense)
{
    return launch_eval_t<kModeNormal>(a, layout, with_bg, cfg, s, dense);
}

}  // namespace pda
