// fs_f8_f64.cu — double instantiations of the kernels of fs_fig8.cuh (own translation unit so the
// kernel families compile in parallel).
#include "fs_launch.h"

template <>
cudaError_t fs_launch_f8<double>(int V, int grid, int block, size_t smem, cudaStream_t st,
                         const fs::Cfg<double>& c, const fs::F8Bufs<double>& b, const fs::StepArgs<double>& a) {
  return fs::launch_f8_impl<double>(V, grid, block, smem, st, c, b, a);
}

template <>
cudaError_t fs_launch_f8_reset<double>(int E, cudaStream_t st, const fs::Cfg<double>& c,
                               const fs::F8Bufs<double>& b, const fs::F8Geo& g, const uint8_t* mask,
                               double* length) {
  return fs::launch_f8_reset_impl<double>(E, st, c, b, g, mask, length);
}
