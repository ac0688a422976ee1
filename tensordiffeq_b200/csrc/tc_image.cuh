// Layout of the tensor-core operand images of the hidden weight matrices (read by tc_kernel.cuh, written by the
// pack kernels in optim_kernels.cuh).
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

namespace tdq {

constexpr int kTcPlaneF = 128 * 32;              // floats in one [128][32] plane (16 KB)
constexpr uint32_t kTcPlaneB = kTcPlaneF * 4;
constexpr uint32_t kTcStageB = 2 * kTcPlaneB;    // one weight K-block: hi + lo
constexpr int kTcDwCol = 384;
constexpr int kTcMaxL = 5;

// floats of weight images per hidden matrix: [direction 2][K-block 4][128][32] (the fp32 words as they are: the
// tensor core reads their tf32 part, the kernel derives the lo parts on chip)
constexpr size_t kTcImgFloats = (size_t)2 * 4 * kTcPlaneF;
__host__ __device__ inline size_t tc_img_off(int m, int dir, int kb) {
    return (((size_t)m * 2 + dir) * 4 + kb) * kTcPlaneF;
}
// float offset of (row, k) in a [128][32] K-major SWIZZLE_128B plane
__host__ __device__ inline int tc_plane_off(int row, int k) { return row * 32 + ((((k >> 2) ^ (row & 7))) << 2) + (k & 3); }

// Order of one hidden matrix' gradient block in the tensor-core kernel's slab: 4-column groups outermost, so that
// the 32 lanes of a warp (32 consecutive rows k, the same 4 columns j) write 512 contiguous bytes.
__host__ __device__ inline int tc_slab_off(int k, int j, int HP) { return ((j >> 2) * HP + k) * 4 + (j & 3); }

// scatter one hidden-matrix element W_m[k][j] into both images (called by the pack kernels)
__device__ __forceinline__ void tc_img_store(float* img, int m, int k, int j, float val) {
    img[tc_img_off(m, 0, k >> 5) + tc_plane_off(j, k & 31)] = val;      // direction 0 (forward GEMM): A[M = j][K = k]
    img[tc_img_off(m, 1, j >> 5) + tc_plane_off(k, j & 31)] = val;      // direction 1 (data gradient): A[M = k][K = j]
}

}  // namespace tdq
