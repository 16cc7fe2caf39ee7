// fs_fast_f32.cu — fp32 instantiations of the IDM-ring hot kernel (own
// translation unit so the variants compile in parallel with the rest).
#include "fs_ring_fast.cuh"
#include "fs_ring_hot.cuh"
#include "fs_launch.h"

cudaError_t fs_launch_fast22_f32(int shape, int variant, int n_envs, cudaStream_t st,
                                 const fs::FastP<float>& c, const fs::RingBufs<float>& b,
                                 const fs::StepArgs<float>& a) {
  return fs::launch_fast22<float>(shape, variant, n_envs, st, c, b, a);
}

cudaError_t fs_launch_hot22_f32(int shape, bool noise, int n_envs, cudaStream_t st,
                                const fs::FastP<float>& c, const fs::RingBufs<float>& b,
                                const fs::StepArgs<float>& a) {
  return fs::launch_hot22<float>(shape, noise, n_envs, st, c, b, a);
}
