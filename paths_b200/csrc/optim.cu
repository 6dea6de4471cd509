// Fused AdamW over the flat parameter buffer (train.py:49,66: torch.optim.AdamW(model.parameters(), lr, weight_decay)).
// One HBM pass: 16 B read + 12 B written per live parameter.  Semantics follow torch.optim.AdamW (non-amsgrad):
//   p *= 1 - lr*wd;  m = b1 m + (1-b1) g;  v = b2 v + (1-b2) g^2;  p -= (lr / (1-b1^t)) * m / (sqrt(v)/sqrt(1-b2^t) + eps)
// Parameters whose gradient the reference leaves at None (classifiers of levels 0..L-2) are skipped entirely, as torch
// skips them; the dead encoder / cross-attention parameters (exact-zero gradients) sit in the tail [n_live, n_total) and
// only receive the decoupled weight decay (their moments stay zero), which is bit-identical to torch's update for g = 0.
// Modules the forward pass never evaluates (procs.0.hctx_mlp of the RNN variant, every hctx_mlp without hierarchical_ctx)
// also sit in the tail but keep .grad = None in the reference: they arrive in the skip list and are left untouched.
#include "../../include/paths_b200.h"
#include "common.cuh"

namespace paths {
namespace {

struct SkipList { long long a[16], b[16]; int n; };

__global__ void __launch_bounds__(256) adamw_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m,
                                                    float* __restrict__ v, long long n_live, long long n_total, SkipList skip,
                                                    float lr, float b1, float b2, float eps, float wd, float inv_bc1, float inv_sqrt_bc2) {
  pdl_grid_sync();
  const long long n4 = n_total / 4;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) {
    const long long e0 = i * 4;
    float4 pv = reinterpret_cast<float4*>(p)[i];
    float* pe = reinterpret_cast<float*>(&pv);
    if (e0 + 3 < n_live) {
      bool any_skip = false;
      for (int s = 0; s < skip.n; ++s) any_skip |= (e0 + 3 >= skip.a[s] && e0 < skip.b[s]);
      const float4 gv = reinterpret_cast<const float4*>(g)[i];
      float4 mv = reinterpret_cast<float4*>(m)[i], vv = reinterpret_cast<float4*>(v)[i];
      const float* ge = reinterpret_cast<const float*>(&gv);
      float* me = reinterpret_cast<float*>(&mv);
      float* ve = reinterpret_cast<float*>(&vv);
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        bool sk = false;
        if (any_skip)
          for (int s = 0; s < skip.n; ++s) sk |= (e0 + k >= skip.a[s] && e0 + k < skip.b[s]);
        if (sk) continue;
        float x = pe[k] * (1.f - lr * wd);
        me[k] = b1 * me[k] + (1.f - b1) * ge[k];
        ve[k] = b2 * ve[k] + (1.f - b2) * ge[k] * ge[k];
        const float denom = sqrtf(ve[k]) * inv_sqrt_bc2 + eps;
        pe[k] = x - (lr * inv_bc1) * (me[k] / denom);
      }
      reinterpret_cast<float4*>(m)[i] = mv;
      reinterpret_cast<float4*>(v)[i] = vv;
    } else {
      // (n_live is a multiple of 64 floats by construction, so a float4 never straddles the live / dead boundary)
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        bool sk = e0 + k < n_live;
        for (int s = 0; s < skip.n; ++s) sk |= (e0 + k >= skip.a[s] && e0 + k < skip.b[s]);   // never-evaluated modules: grad None
        if (!sk) pe[k] *= (1.f - lr * wd);
      }
    }
    reinterpret_cast<float4*>(p)[i] = pv;
  }
}

}  // namespace
}  // namespace paths

extern "C" int paths_adamw_step(float* params, const float* grads, float* exp_avg, float* exp_avg_sq, int64_t n_live, int64_t n_total,
                                const int64_t* h_skip, int n_skip, float lr, float beta1, float beta2, float eps, float weight_decay,
                                int step, void* stream) {
  using namespace paths;
  PATHS_REQUIRE(params && grads && exp_avg && exp_avg_sq, "adamw: null buffer");
  PATHS_REQUIRE(n_live >= 0 && n_total >= n_live && n_live % 4 == 0 && n_total % 4 == 0, "adamw: sizes must be multiples of 4 floats");
  PATHS_REQUIRE(n_skip >= 0 && n_skip <= 16 && step >= 1, "adamw: at most 16 skipped ranges; step counts from 1");
  SkipList sk; sk.n = n_skip;
  for (int i = 0; i < n_skip; ++i) { sk.a[i] = h_skip[2 * i]; sk.b[i] = h_skip[2 * i + 1]; }
  const double bc1 = 1.0 - pow((double)beta1, step), bc2 = 1.0 - pow((double)beta2, step);
  long long blocks = (n_total / 4 + 255) / 256;
  if (blocks > 8 * kNumSMs) blocks = 8 * kNumSMs;
  if (blocks < 1) return 0;
  pdl(adamw_kernel, (unsigned)blocks, 256, 0, static_cast<cudaStream_t>(stream))(params, grads, exp_avg, exp_avg_sq, n_live, n_total, sk, lr,
                                                                               beta1, beta2, eps, weight_decay, (float)(1.0 / bc1),
                                                                               (float)(1.0 / sqrt(bc2)));
  PATHS_LAUNCH_CK();
  return 0;
}
