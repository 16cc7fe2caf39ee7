// fs_ml_f32.cu — float instantiations of the kernels of fs_multilane.cuh (own translation unit so the
// kernel families compile in parallel).
#include "fs_launch.h"

template <>
cudaError_t fs_launch_ml<float>(int V, bool uniform, int grid, int block, size_t smem, cudaStream_t st, const fs::Cfg<float>& c, const fs::MlBufs<float>& b, const fs::StepArgs<float>& a) {
  return fs::launch_ml_impl<float>(V, uniform, grid, block, smem, st, c, b, a);
}
