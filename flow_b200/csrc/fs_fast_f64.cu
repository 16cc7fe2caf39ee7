// fs_fast_f64.cu — fp64 (parity mode) instantiations of the IDM-ring hot kernel.
#include "fs_ring_fast.cuh"
#include "fs_ring_hot.cuh"
#include "fs_launch.h"

cudaError_t fs_launch_fast22_f64(int shape, int variant, int n_envs, cudaStream_t st,
                                 const fs::FastP<double>& c, const fs::RingBufs<double>& b,
                                 const fs::StepArgs<double>& a) {
  return fs::launch_fast22<double>(shape, variant, n_envs, st, c, b, a);
}

cudaError_t fs_launch_hot22_f64(int shape, bool noise, int n_envs, cudaStream_t st,
                                const fs::FastP<double>& c, const fs::RingBufs<double>& b,
                                const fs::StepArgs<double>& a) {
  return fs::launch_hot22<double>(shape, noise, n_envs, st, c, b, a);
}
