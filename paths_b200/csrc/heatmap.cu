// Importance heat-map compositing (heatmap_visualise.py:142-178): every patch of every magnification level fills
// its rectangle of an [L, H, W] canvas (in base-level pixel space) with `importance + 1e-4`, then the levels are folded
// coarse-to-fine (`coarse += 0.5 * fine` wherever the finer level is non-zero), zeros of the result are replaced by the
// smallest positive value and an alpha mask (0.5 where covered) is produced.  Integer rectangle arithmetic and IEEE
// float/double adds only, so the result is bit-identical to the numpy restatement in oracle/heatmap_oracle.py.
//
// HBM-bound: the canvas is written once per level (owner pass) and read once in the fold: ~ (4 + 8) * L * H * W bytes.
#include "../../include/paths_b200.h"
#include "common.cuh"

namespace paths {
namespace {

__device__ __forceinline__ int shr_floor(int v, int d) { return v >> d; }       // python `//` by 2^d (floor, also for v < 0)

// Pass 1: owner[d, y, x] = 1 + index of the LAST patch of level d whose rectangle covers the pixel (numpy's sequential slice
// assignment lets later patches overwrite earlier ones, heatmap_visualise.py:152-165).  One block per patch.
__global__ void __launch_bounds__(256) heat_owner_kernel(const int* __restrict__ locs_yx, const int* __restrict__ level_off, int L, int P,
                                                         int H, int W, int* __restrict__ owner) {
  pdl_grid_sync();
  const int i = blockIdx.x;
  int d = 0;
  while (d + 1 < L && i >= level_off[d + 1]) ++d;
  const int size = shr_floor(P, d);
  const int y = shr_floor(locs_yx[2 * i], d), x = shr_floor(locs_yx[2 * i + 1], d);
  // overlap test exactly as written in the reference (closed comparison): max(y,0) <= min(y+size, H)
  const int y1 = max(y, 0), y2 = min(y + size, H), x1 = max(x, 0), x2 = min(x + size, W);
  if (!(y1 <= y2 && x1 <= x2)) return;
  const int w = x2 - x1, h = y2 - y1;
  if (w <= 0 || h <= 0) return;
  const int local = i - level_off[d] + 1;
  int* base = owner + (size_t)d * H * W;
  for (int e = threadIdx.x; e < w * h; e += blockDim.x) {
    const int yy = y1 + e / w, xx = x1 + e % w;
    atomicMax(base + (size_t)yy * W + xx, local);
  }
}

// Pass 2: per pixel, rebuild the L canvas values from the owners, fold fine -> coarse, emit level 0 and track the smallest
// positive value (positive doubles order like their bit patterns).
__global__ void __launch_bounds__(256) heat_fold_kernel(const int* __restrict__ owner, const float* __restrict__ imp,
                                                        const int* __restrict__ level_off, int L, int H, int W, double* __restrict__ out,
                                                        unsigned long long* __restrict__ min_bits) {
  pdl_grid_sync();
  const size_t n = (size_t)H * W;
  unsigned long long local_min = ~0ull;
  for (size_t p = (size_t)blockIdx.x * blockDim.x + threadIdx.x; p < n; p += (size_t)gridDim.x * blockDim.x) {
    double acc = 0.0;                                    // value of level d after folding levels d+1.. into it
    for (int d = L - 1; d >= 0; --d) {
      const int o = owner[(size_t)d * n + p];
      // numpy 2: float32 importance + python float 1e-4 is a float32 add, stored into the float64 canvas
      const double v = o ? (double)__fadd_rn(imp[level_off[d] + o - 1], 1e-4f) : 0.0;
      acc = (d == L - 1) ? v : ((acc != 0.0) ? __dadd_rn(v, __dmul_rn(acc, 0.5)) : v);
    }
    out[p] = acc;
    if (acc > 0.0) { const unsigned long long b = (unsigned long long)__double_as_longlong(acc); if (b < local_min) local_min = b; }
  }
  for (int o = 16; o > 0; o >>= 1) { const unsigned long long t = __shfl_xor_sync(0xffffffffu, local_min, o); if (t < local_min) local_min = t; }
  if ((threadIdx.x & 31) == 0 && local_min != ~0ull) atomicMin(min_bits, local_min);
}

// Pass 3: alpha = 0.5 where covered (> 0), zeros replaced by the smallest positive value (heatmap_visualise.py:175-176).
__global__ void __launch_bounds__(256) heat_fill_kernel(double* __restrict__ out, float* __restrict__ alpha, size_t n,
                                                        const unsigned long long* __restrict__ min_bits) {
  pdl_grid_sync();
  const unsigned long long mb = *min_bits;
  const double mn = mb == ~0ull ? 0.0 : __longlong_as_double((long long)mb);
  for (size_t p = (size_t)blockIdx.x * blockDim.x + threadIdx.x; p < n; p += (size_t)gridDim.x * blockDim.x) {
    const double v = out[p];
    alpha[p] = v > 0.0 ? 0.5f : 0.f;
    if (v == 0.0) out[p] = mn;
  }
}

}  // namespace
}  // namespace paths

using namespace paths;

extern "C" {

size_t paths_heatmap_workspace_bytes(int L, int H, int W) { return (size_t)L * H * W * sizeof(int) + 256; }

int paths_heatmap_composite(const int32_t* locs_yx, const float* importance, const int32_t* level_off, int32_t n_patches, int L,
                            int patch_size, int H, int W, void* workspace, size_t workspace_bytes, double* out, float* alpha,
                            void* stream) {
  PATHS_REQUIRE(level_off && workspace && out && alpha && (n_patches == 0 || (locs_yx && importance)), "heatmap: null argument");
  PATHS_REQUIRE(L >= 1 && L <= 30 && H > 0 && W > 0 && patch_size > 0, "heatmap: bad shape (L=%d H=%d W=%d P=%d)", L, H, W, patch_size);
  PATHS_REQUIRE(workspace_bytes >= paths_heatmap_workspace_bytes(L, H, W), "heatmap: workspace too small");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  char* ws = static_cast<char*>(workspace);
  unsigned long long* min_bits = reinterpret_cast<unsigned long long*>(ws);
  int* owner = reinterpret_cast<int*>(ws + 256);
  PATHS_CK(cudaMemsetAsync(owner, 0, (size_t)L * H * W * sizeof(int), st));
  PATHS_CK(cudaMemsetAsync(min_bits, 0xff, sizeof(unsigned long long), st));
  if (n_patches > 0) {
    pdl(heat_owner_kernel, n_patches, 256, 0, st)(locs_yx, level_off, L, patch_size, H, W, owner);
    PATHS_LAUNCH_CK();
  }
  const size_t n = (size_t)H * W;
  int blocks = (int)((n + 255) / 256);
  if (blocks > 8 * kNumSMs) blocks = 8 * kNumSMs;
  pdl(heat_fold_kernel, blocks, 256, 0, st)(owner, importance, level_off, L, H, W, out, min_bits);
  PATHS_LAUNCH_CK();
  pdl(heat_fill_kernel, blocks, 256, 0, st)(out, alpha, n, min_bits);
  PATHS_LAUNCH_CK();
  return 0;
}

}  // extern "C"
