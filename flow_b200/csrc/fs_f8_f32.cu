// fs_f8_f32.cu — float instantiations of the kernels of fs_fig8.cuh (own translation unit so the
// kernel families compile in parallel).
#include "fs_launch.h"

template <>
cudaError_t fs_launch_f8<float>(int V, int grid, int block, size_t smem, cudaStream_t st,
                         const fs::Cfg<float>& c, const fs::F8Bufs<float>& b, const fs::StepArgs<float>& a) {
  return fs::launch_f8_impl<float>(V, grid, block, smem, st, c, b, a);
}

template <>
cudaError_t fs_launch_f8_reset<float>(int E, cudaStream_t st, const fs::Cfg<float>& c,
                               const fs::F8Bufs<float>& b, const fs::F8Geo& g, const uint8_t* mask,
                               float* length) {
  return fs::launch_f8_reset_impl<float>(E, st, c, b, g, mask, length);
}
