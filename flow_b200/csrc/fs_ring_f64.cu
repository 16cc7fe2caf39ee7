// fs_ring_f64.cu — double instantiations of the kernels of fs_ring.cuh (own translation unit so the
// kernel families compile in parallel).
#include "fs_launch.h"

template <>
cudaError_t fs_launch_ring1<double>(int V, int E, cudaStream_t st, const fs::Cfg<double>& c, const fs::RingBufs<double>& b, const fs::StepArgs<double>& a) {
  return fs::launch_ring1_impl<double>(V, E, st, c, b, a);
}
