// fs_launch.h — launch entry points of the per-dtype kernel translation units.
#pragma once
#include <cuda_runtime.h>
#include "fs_common.cuh"
#include "fs_ring.cuh"
#include "fs_ring_fast.cuh"
#include "fs_ring_hot.cuh"

cudaError_t fs_launch_fast22_f32(int shape, int variant, int n_envs, cudaStream_t st,
                                 const fs::FastP<float>& c, const fs::RingBufs<float>& b,
                                 const fs::StepArgs<float>& a);
cudaError_t fs_launch_fast22_f64(int shape, int variant, int n_envs, cudaStream_t st,
                                 const fs::FastP<double>& c, const fs::RingBufs<double>& b,
                                 const fs::StepArgs<double>& a);
cudaError_t fs_launch_hot22_f32(int shape, bool noise, int n_envs, cudaStream_t st,
                                const fs::FastP<float>& c, const fs::RingBufs<float>& b,
                                const fs::StepArgs<float>& a);
cudaError_t fs_launch_hot22_f64(int shape, bool noise, int n_envs, cudaStream_t st,
                                const fs::FastP<double>& c, const fs::RingBufs<double>& b,
                                const fs::StepArgs<double>& a);
