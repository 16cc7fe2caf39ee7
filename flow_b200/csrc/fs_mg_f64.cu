// fs_mg_f64.cu — double instantiations of the kernels of fs_merge.cuh (own translation unit so the
// kernel families compile in parallel).
#include "fs_launch.h"

template <>
cudaError_t fs_launch_mg<double>(int cap, int grid, int block, size_t smem, cudaStream_t st,
                         const fs::Cfg<double>& c, const fs::MgBufs<double>& b, const fs::StepArgs<double>& a) {
  return fs::launch_mg_impl<double>(cap, grid, block, smem, st, c, b, a);
}

template <>
cudaError_t fs_launch_mg_reset<double>(int E, cudaStream_t st, const fs::Cfg<double>& c,
                               const fs::MgBufs<double>& b, const uint8_t* mask, int n_initial,
                               double* length) {
  return fs::launch_mg_reset_impl<double>(E, st, c, b, mask, n_initial, length);
}
