// fs_ml_f64.cu — double instantiations of the kernels of fs_multilane.cuh (own translation unit so the
// kernel families compile in parallel).
#include "fs_launch.h"

template <>
cudaError_t fs_launch_ml<double>(int V, bool uniform, int grid, int block, size_t smem, cudaStream_t st, const fs::Cfg<double>& c, const fs::MlBufs<double>& b, const fs::StepArgs<double>& a) {
  return fs::launch_ml_impl<double>(V, uniform, grid, block, smem, st, c, b, a);
}
