// K17: column sums of a bf16 [rows, cols] matrix in fp32 - the bias gradient of a Linear layer (db = sum over
// tokens of dY).  ATen's generic reduction reaches ~0.5 TB/s on these shapes (8 - 32 MB, 96 calls per GPT-2 step);
// here every thread owns 8 adjacent columns (one 16-byte load per row), a CTA of 32 x 8 threads covers 256 columns
// and a slice of the rows, and the per-slice partial sums land in partial[slices][cols] (deterministic; the caller
// folds the few slices).  Reference counterpart: the framework's bias gradient under model.fit
// (keras_model_ops.py:156).
#include <algorithm>

#include "common.cuh"

namespace mfl {
namespace colsum {

constexpr int kTX = 32, kTY = 8, kUnroll = 4;
using V = VecTraits<MFL_BFLOAT16>;

__global__ void __launch_bounds__(kTX * kTY) colsum_partial_kernel(const __nv_bfloat16* __restrict__ x, int64_t ld,
                                                                   float* __restrict__ partial, int rows, int cols,
                                                                   int rows_per_slice) {
  __shared__ float red[kTY][kTX * 8 + 8];
  const int c0 = (blockIdx.x * kTX + threadIdx.x) * 8;
  const int r0 = blockIdx.y * rows_per_slice, r1 = min(rows, r0 + rows_per_slice);
  float acc[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) acc[k] = 0.f;
  if (c0 < cols) {
    for (int r = r0 + threadIdx.y; r < r1; r += kTY * kUnroll) {
      uint4 v[kUnroll];
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        const int rr = r + u * kTY;
        v[u] = rr < r1 ? ld_stream(x + (int64_t)rr * ld + c0) : make_uint4(0, 0, 0, 0);
      }
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        float f[8];
        V::unpack(v[u], f);
#pragma unroll
        for (int k = 0; k < 8; ++k) acc[k] += f[k];
      }
    }
  }
#pragma unroll
  for (int k = 0; k < 8; ++k) red[threadIdx.y][threadIdx.x * 8 + k] = acc[k];
  __syncthreads();
  const int t = threadIdx.y * kTX + threadIdx.x;  // 256 threads fold the 8 row lanes of 256 columns
  const int c = blockIdx.x * kTX * 8 + t;
  if (c < cols) {
    float s = 0.f;
#pragma unroll
    for (int y = 0; y < kTY; ++y) s += red[y][t];
    partial[(size_t)blockIdx.y * cols + c] = s;
  }
}

}  // namespace colsum
}  // namespace mfl

using namespace mfl;

extern "C" int mfl_colsum_slices(int rows) {
  return std::max(1, std::min(64, rows / 64));
}

// partial: [mfl_colsum_slices(rows), cols] fp32, fully overwritten.  cols % 8 == 0, ld % 8 == 0, 16-byte aligned base.
extern "C" int mfl_colsum_bf16(float* partial, const void* x, int64_t ld, int rows, int cols, void* stream) {
  MFL_ARG(partial && x && rows > 0 && cols > 0, "colsum: null tensor or empty shape");
  MFL_ARG(cols % 8 == 0 && ld % 8 == 0 && ((uintptr_t)x & 15) == 0, "colsum: columns and row pitch must be multiples of 8");
  const int slices = mfl_colsum_slices(rows);
  const int rps = (rows + slices - 1) / slices;
  dim3 grid((cols + colsum::kTX * 8 - 1) / (colsum::kTX * 8), slices), block(colsum::kTX, colsum::kTY);
  colsum::colsum_partial_kernel<<<grid, block, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(x), ld, partial, rows, cols, rps);
  return check(cudaGetLastError(), "colsum_partial_kernel launch");
}
