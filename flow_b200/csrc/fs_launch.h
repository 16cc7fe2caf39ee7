// fs_launch.h — launch entry points of the per-kernel-family, per-dtype translation units
// (fs_fast_*.cu, fs_ring_*.cu, fs_ml_*.cu, fs_hw_*.cu, fs_mg_*.cu, fs_f8_*.cu): the kernel templates are instantiated
// there so that the families compile in parallel; flowstep.cu only sees these declarations.
#pragma once
#include <cuda_runtime.h>
#include "fs_common.cuh"
#include "fs_ring.cuh"
#include "fs_ring_fast.cuh"
#include "fs_multilane.cuh"
#include "fs_highway.cuh"
#include "fs_merge.cuh"
#include "fs_fig8.cuh"

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

// generic ring kernel (fs_ring.cuh), multi-lane ring (fs_multilane.cuh), open highway (fs_highway.cuh)
template <typename T>
cudaError_t fs_launch_ring1(int V, int E, cudaStream_t st, const fs::Cfg<T>& c,
                            const fs::RingBufs<T>& b, const fs::StepArgs<T>& a);
template <typename T>
cudaError_t fs_launch_ml(int V, bool uniform, int grid, int block, size_t smem, cudaStream_t st,
                         const fs::Cfg<T>& c, const fs::MlBufs<T>& b, const fs::StepArgs<T>& a);
template <typename T>
cudaError_t fs_launch_hw(int cap, bool multi_agent, bool single_periodic, bool generic, int grid,
                         int block, size_t smem, cudaStream_t st, const fs::Cfg<T>& c,
                         const fs::HwBufs<T>& b, const fs::StepArgs<T>& a);
// merge network (fs_merge.cuh)
template <typename T>
cudaError_t fs_launch_mg(int cap, int grid, int block, size_t smem, cudaStream_t st,
                         const fs::Cfg<T>& c, const fs::MgBufs<T>& b, const fs::StepArgs<T>& a);
template <typename T>
cudaError_t fs_launch_mg_reset(int E, cudaStream_t st, const fs::Cfg<T>& c, const fs::MgBufs<T>& b,
                               const uint8_t* mask, int n_initial, T* length);
// figure-eight network (fs_fig8.cuh)
template <typename T>
cudaError_t fs_launch_f8(int V, int grid, int block, size_t smem, cudaStream_t st,
                         const fs::Cfg<T>& c, const fs::F8Bufs<T>& b, const fs::StepArgs<T>& a);
template <typename T>
cudaError_t fs_launch_f8_reset(int E, cudaStream_t st, const fs::Cfg<T>& c, const fs::F8Bufs<T>& b,
                               const fs::F8Geo& g, const uint8_t* mask, T* length);
