// Varlen softmax self-attention over packed token rows (the self-attention of nn.TransformerDecoderLayer as used by
// model/aggregator.py:70-72, key-padding mask replaced by per-slide segments).  head_dim is 32 for the shipped configs:
// 128 MMA-FLOP per exp, i.e. the op is SFU/latency-bound, not tensor-bound, and at T <= 4K+1 tokens it is <1 % of the
// level's FLOPs, so it stays on the SIMT pipes: one warp per query, lanes over keys (scores) / head dims (outputs),
// K/V tiles of 32 keys staged in shared memory, online softmax.  Long bags (20 k tokens) tile over keys the same way.
#include "common.cuh"
#include "rowops.cuh"

namespace paths {

namespace {

constexpr int QW = 4;            // queries per warp
constexpr int AW = 4;            // warps per block
constexpr int QB = QW * AW;      // queries per block
constexpr int KT = 32;           // keys per tile
constexpr int MAXDH = 64;

struct Seg { int t0, T; };
__device__ __forceinline__ Seg seg_of(const int* seg, int s) { Seg r; r.t0 = seg[s] + s; r.T = seg[s + 1] - seg[s] + 1; return r; }

// out[q, h*dh + c] = sum_j softmax_j(scale * q.k_j) v_j[c];  lse[q, h] = logsumexp_j(scale * q.k_j)
__global__ void __launch_bounds__(AW * 32) attn_fwd_kernel(const float* __restrict__ qkv, const int* __restrict__ seg,
                                                           float* __restrict__ out, float* __restrict__ lse, int d, int heads,
                                                           int nq_mode) {
  const int s = blockIdx.z, h = blockIdx.y, dh = d / heads;
  const Seg sg = seg_of(seg, s);
  const int nq = nq_mode ? 1 : sg.T;
  const int q0 = blockIdx.x * QB;
  if (q0 >= nq) return;
  __shared__ float Ks[KT][MAXDH + 1], Vs[KT][MAXDH + 1];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const float scale = rsqrtf((float)dh);
  const int ld = 3 * d;

  float qa[QW], qb[QW], m[QW], l[QW], oa[QW], ob[QW];
#pragma unroll
  for (int i = 0; i < QW; ++i) {
    const int q = q0 + w * QW + i;
    qa[i] = qb[i] = 0.f;
    if (q < nq) {
      const float* qp = qkv + (size_t)(sg.t0 + q) * ld + h * dh;
      if (lane < dh) qa[i] = qp[lane] * scale;
      if (lane + 32 < dh) qb[i] = qp[lane + 32] * scale;
    }
    m[i] = -INFINITY; l[i] = 0.f; oa[i] = ob[i] = 0.f;
  }
  for (int k0 = 0; k0 < sg.T; k0 += KT) {
    for (int e = threadIdx.x; e < KT * dh; e += AW * 32) {
      const int j = e / dh, c = e % dh;
      float kv = 0.f, vv = 0.f;
      if (k0 + j < sg.T) {
        const float* rp = qkv + (size_t)(sg.t0 + k0 + j) * ld + h * dh + c;
        kv = rp[d]; vv = rp[2 * d];
      }
      Ks[j][c] = kv; Vs[j][c] = vv;
    }
    __syncthreads();
    const bool valid = k0 + lane < sg.T;
#pragma unroll
    for (int i = 0; i < QW; ++i) {
      if (q0 + w * QW + i >= nq) continue;            // warp-uniform
      float sc = 0.f;
      for (int c = 0; c < dh; ++c) {
        const float qc = __shfl_sync(0xffffffffu, c < 32 ? qa[i] : qb[i], c & 31);
        sc = fmaf(qc, Ks[lane][c], sc);
      }
      sc = valid ? sc : -INFINITY;
      const float mn = fmaxf(m[i], warp_max(sc));
      const float p = valid ? expf(sc - mn) : 0.f;
      const float corr = expf(m[i] - mn);
      l[i] = l[i] * corr + warp_sum(p);
      oa[i] *= corr; ob[i] *= corr;
      const int jmax = min(KT, sg.T - k0);
      for (int j = 0; j < jmax; ++j) {
        const float pj = __shfl_sync(0xffffffffu, p, j);
        oa[i] = fmaf(pj, Vs[j][lane], oa[i]);
        if (dh > 32) ob[i] = fmaf(pj, Vs[j][lane + 32], ob[i]);
      }
      m[i] = mn;
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < QW; ++i) {
    const int q = q0 + w * QW + i;
    if (q >= nq) continue;
    const size_t orow = nq_mode ? (size_t)s : (size_t)(sg.t0 + q);
    const float inv = 1.f / l[i];
    if (lane < dh) out[orow * d + h * dh + lane] = oa[i] * inv;
    if (lane + 32 < dh) out[orow * d + h * dh + lane + 32] = ob[i] * inv;
    if (lane == 0) lse[orow * heads + h] = m[i] + logf(l[i]);
  }
}

// dq (and Dvec[q, h] = dout . out) for every query: dS_j = p_j (dout.v_j - D), dq = scale * sum_j dS_j k_j
__global__ void __launch_bounds__(AW * 32) attn_bwd_dq_kernel(const float* __restrict__ qkv, const int* __restrict__ seg,
                                                              const float* __restrict__ out, const float* __restrict__ lse,
                                                              const float* __restrict__ dout, float* __restrict__ dqkv,
                                                              float* __restrict__ Dvec, int d, int heads, int nq_mode) {
  const int s = blockIdx.z, h = blockIdx.y, dh = d / heads;
  const Seg sg = seg_of(seg, s);
  const int nq = nq_mode ? 1 : sg.T;
  const int q0 = blockIdx.x * QB;
  if (q0 >= nq) return;
  __shared__ float Ks[KT][MAXDH + 1], Vs[KT][MAXDH + 1];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const float scale = rsqrtf((float)dh);
  const int ld = 3 * d;

  float qa[QW], qb[QW], ga[QW], gb[QW], Dq[QW], L[QW], da[QW], db[QW];
#pragma unroll
  for (int i = 0; i < QW; ++i) {
    const int q = q0 + w * QW + i;
    qa[i] = qb[i] = ga[i] = gb[i] = 0.f; Dq[i] = 0.f; L[i] = 0.f; da[i] = db[i] = 0.f;
    if (q < nq) {
      const float* qp = qkv + (size_t)(sg.t0 + q) * ld + h * dh;
      const size_t orow = nq_mode ? (size_t)s : (size_t)(sg.t0 + q);
      float dd = 0.f;
      if (lane < dh) { qa[i] = qp[lane] * scale; ga[i] = dout[orow * d + h * dh + lane]; dd += ga[i] * out[orow * d + h * dh + lane]; }
      if (lane + 32 < dh) { qb[i] = qp[lane + 32] * scale; gb[i] = dout[orow * d + h * dh + lane + 32]; dd += gb[i] * out[orow * d + h * dh + lane + 32]; }
      Dq[i] = warp_sum(dd);
      L[i] = lse[orow * heads + h];
      if (lane == 0) Dvec[orow * heads + h] = Dq[i];
    }
  }
  for (int k0 = 0; k0 < sg.T; k0 += KT) {
    for (int e = threadIdx.x; e < KT * dh; e += AW * 32) {
      const int j = e / dh, c = e % dh;
      float kv = 0.f, vv = 0.f;
      if (k0 + j < sg.T) {
        const float* rp = qkv + (size_t)(sg.t0 + k0 + j) * ld + h * dh + c;
        kv = rp[d]; vv = rp[2 * d];
      }
      Ks[j][c] = kv; Vs[j][c] = vv;
    }
    __syncthreads();
    const bool valid = k0 + lane < sg.T;
#pragma unroll
    for (int i = 0; i < QW; ++i) {
      if (q0 + w * QW + i >= nq) continue;
      float sc = 0.f, dp = 0.f;
      for (int c = 0; c < dh; ++c) {
        const float qc = __shfl_sync(0xffffffffu, c < 32 ? qa[i] : qb[i], c & 31);
        const float gc = __shfl_sync(0xffffffffu, c < 32 ? ga[i] : gb[i], c & 31);
        sc = fmaf(qc, Ks[lane][c], sc);
        dp = fmaf(gc, Vs[lane][c], dp);
      }
      const float p = valid ? expf(sc - L[i]) : 0.f;
      const float ds = p * (dp - Dq[i]);
      const int jmax = min(KT, sg.T - k0);
      for (int j = 0; j < jmax; ++j) {
        const float dsj = __shfl_sync(0xffffffffu, ds, j);
        da[i] = fmaf(dsj, Ks[j][lane], da[i]);
        if (dh > 32) db[i] = fmaf(dsj, Ks[j][lane + 32], db[i]);
      }
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < QW; ++i) {
    const int q = q0 + w * QW + i;
    if (q >= nq) continue;
    float* dq = dqkv + (size_t)(sg.t0 + q) * ld + h * dh;
    if (lane < dh) dq[lane] = da[i] * scale;
    if (lane + 32 < dh) dq[lane + 32] = db[i] * scale;
  }
}

// dk_j = sum_i dS_ij (scale q_i), dv_j = sum_i p_ij dout_i : one warp per key, lanes over queries, query tiles in smem.
__global__ void __launch_bounds__(AW * 32) attn_bwd_dkv_kernel(const float* __restrict__ qkv, const int* __restrict__ seg,
                                                               const float* __restrict__ lse, const float* __restrict__ dout,
                                                               const float* __restrict__ Dvec, float* __restrict__ dqkv, int d,
                                                               int heads, int nq_mode) {
  const int s = blockIdx.z, h = blockIdx.y, dh = d / heads;
  const Seg sg = seg_of(seg, s);
  const int nq = nq_mode ? 1 : sg.T;
  const int j0 = blockIdx.x * QB;
  if (j0 >= sg.T) return;
  __shared__ float Qs[KT][MAXDH + 1], Gs[KT][MAXDH + 1];
  __shared__ float Ls[KT], Ds[KT];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const float scale = rsqrtf((float)dh);
  const int ld = 3 * d;

  float ka[QW], kb[QW], va[QW], vb[QW], dka[QW], dkb[QW], dva[QW], dvb[QW];
#pragma unroll
  for (int i = 0; i < QW; ++i) {
    const int j = j0 + w * QW + i;
    ka[i] = kb[i] = va[i] = vb[i] = 0.f; dka[i] = dkb[i] = dva[i] = dvb[i] = 0.f;
    if (j < sg.T) {
      const float* rp = qkv + (size_t)(sg.t0 + j) * ld + h * dh;
      if (lane < dh) { ka[i] = rp[d + lane]; va[i] = rp[2 * d + lane]; }
      if (lane + 32 < dh) { kb[i] = rp[d + lane + 32]; vb[i] = rp[2 * d + lane + 32]; }
    }
  }
  for (int i0 = 0; i0 < nq; i0 += KT) {
    for (int e = threadIdx.x; e < KT * dh; e += AW * 32) {
      const int i = e / dh, c = e % dh;
      float qv = 0.f, gv = 0.f;
      if (i0 + i < nq) {
        const size_t orow = nq_mode ? (size_t)s : (size_t)(sg.t0 + i0 + i);
        qv = qkv[(size_t)(sg.t0 + i0 + i) * ld + h * dh + c] * scale;
        gv = dout[orow * d + h * dh + c];
      }
      Qs[i][c] = qv; Gs[i][c] = gv;
    }
    if (threadIdx.x < KT) {
      const int i = threadIdx.x;
      float lv = 0.f, dv = 0.f;
      if (i0 + i < nq) {
        const size_t orow = nq_mode ? (size_t)s : (size_t)(sg.t0 + i0 + i);
        lv = lse[orow * heads + h]; dv = Dvec[orow * heads + h];
      }
      Ls[i] = lv; Ds[i] = dv;
    }
    __syncthreads();
    const bool valid = i0 + lane < nq;
#pragma unroll
    for (int i = 0; i < QW; ++i) {
      if (j0 + w * QW + i >= sg.T) continue;
      float sc = 0.f, dp = 0.f;
      for (int c = 0; c < dh; ++c) {
        const float kc = __shfl_sync(0xffffffffu, c < 32 ? ka[i] : kb[i], c & 31);
        const float vc = __shfl_sync(0xffffffffu, c < 32 ? va[i] : vb[i], c & 31);
        sc = fmaf(kc, Qs[lane][c], sc);
        dp = fmaf(vc, Gs[lane][c], dp);
      }
      const float p = valid ? expf(sc - Ls[lane]) : 0.f;
      const float ds = p * (dp - Ds[lane]);
      const int imax = min(KT, nq - i0);
      for (int ii = 0; ii < imax; ++ii) {
        const float pi = __shfl_sync(0xffffffffu, p, ii);
        const float dsi = __shfl_sync(0xffffffffu, ds, ii);
        dva[i] = fmaf(pi, Gs[ii][lane], dva[i]);
        dka[i] = fmaf(dsi, Qs[ii][lane], dka[i]);
        if (dh > 32) { dvb[i] = fmaf(pi, Gs[ii][lane + 32], dvb[i]); dkb[i] = fmaf(dsi, Qs[ii][lane + 32], dkb[i]); }
      }
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < QW; ++i) {
    const int j = j0 + w * QW + i;
    if (j >= sg.T) continue;
    float* rp = dqkv + (size_t)(sg.t0 + j) * ld + h * dh;
    if (lane < dh) { rp[d + lane] = dka[i]; rp[2 * d + lane] = dva[i]; }
    if (lane + 32 < dh) { rp[d + lane + 32] = dkb[i]; rp[2 * d + lane + 32] = dvb[i]; }
  }
}

}  // namespace

int attn_fwd(const float* qkv, const int* seg, float* out, float* lse, int d, int heads, int B, int max_T, int nq_mode, cudaStream_t st) {
  PATHS_REQUIRE(d % heads == 0 && d / heads <= MAXDH, "attention: head_dim %d unsupported (max %d)", d / heads, MAXDH);
  if (B <= 0) return 0;
  dim3 grid(nq_mode ? 1 : (max_T + QB - 1) / QB, heads, B);
  attn_fwd_kernel<<<grid, AW * 32, 0, st>>>(qkv, seg, out, lse, d, heads, nq_mode);
  PATHS_LAUNCH_CK();
  return 0;
}

// dqkv must be zero-initialised by the caller when nq_mode == 1 (dq rows of non-query tokens are not written).
int attn_bwd(const float* qkv, const int* seg, const float* out, const float* lse, const float* dout, float* dqkv, float* Dvec, int d,
             int heads, int B, int max_T, int nq_mode, cudaStream_t st) {
  if (B <= 0) return 0;
  dim3 gq(nq_mode ? 1 : (max_T + QB - 1) / QB, heads, B);
  attn_bwd_dq_kernel<<<gq, AW * 32, 0, st>>>(qkv, seg, out, lse, dout, dqkv, Dvec, d, heads, nq_mode);
  PATHS_LAUNCH_CK();
  dim3 gk((max_T + QB - 1) / QB, heads, B);
  attn_bwd_dkv_kernel<<<gk, AW * 32, 0, st>>>(qkv, seg, lse, dout, Dvec, dqkv, d, heads, nq_mode);
  PATHS_LAUNCH_CK();
  return 0;
}

}  // namespace paths
