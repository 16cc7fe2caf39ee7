// fs_hw_f64.cu — double instantiations of the kernels of fs_highway.cuh (own translation unit so the
// kernel families compile in parallel).
#include "fs_launch.h"

template <>
cudaError_t fs_launch_hw<double>(int cap, bool multi_agent, bool single_periodic, bool generic, int grid, int block, size_t smem, cudaStream_t st, const fs::Cfg<double>& c, const fs::HwBufs<double>& b, const fs::StepArgs<double>& a) {
  return fs::launch_hw_impl<double>(cap, multi_agent, single_periodic, generic, grid, block, smem, st, c, b, a);
}
