// fs_ring_f32.cu — float instantiations of the kernels of fs_ring.cuh (own translation unit so the
// kernel families compile in parallel).
#include "fs_launch.h"

template <>
cudaError_t fs_launch_ring1<float>(int V, int E, cudaStream_t st, const fs::Cfg<float>& c, const fs::RingBufs<float>& b, const fs::StepArgs<float>& a) {
  return fs::launch_ring1_impl<float>(V, E, st, c, b, a);
}
