// fs_hw_f32.cu — float instantiations of the kernels of fs_highway.cuh (own translation unit so the
// kernel families compile in parallel).
#include "fs_launch.h"

template <>
cudaError_t fs_launch_hw<float>(int cap, bool multi_agent, bool single_periodic, bool generic, int grid, int block, size_t smem, cudaStream_t st, const fs::Cfg<float>& c, const fs::HwBufs<float>& b, const fs::StepArgs<float>& a) {
  return fs::launch_hw_impl<float>(cap, multi_agent, single_periodic, generic, grid, block, smem, st, c, b, a);
}
