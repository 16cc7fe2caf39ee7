// fs_mg_f32.cu — float instantiations of the kernels of fs_merge.cuh (own translation unit so the
// kernel families compile in parallel).
#include "fs_launch.h"

template <>
cudaError_t fs_launch_mg<float>(int cap, int grid, int block, size_t smem, cudaStream_t st,
                         const fs::Cfg<float>& c, const fs::MgBufs<float>& b, const fs::StepArgs<float>& a) {
  return fs::launch_mg_impl<float>(cap, grid, block, smem, st, c, b, a);
}

template <>
cudaError_t fs_launch_mg_reset<float>(int E, cudaStream_t st, const fs::Cfg<float>& c,
                               const fs::MgBufs<float>& b, const uint8_t* mask, int n_initial,
                               float* length) {
  return fs::launch_mg_reset_impl<float>(E, st, c, b, mask, n_initial, length);
}
