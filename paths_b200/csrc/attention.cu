// Varlen softmax self-attention over packed token rows (the self-attention of nn.TransformerDecoderLayer as used by
// model/aggregator.py:70-72, key-padding mask replaced by per-slide segments).  head_dim is 32 for the shipped configs:
// 128 MMA-FLOP per exp, i.e. the op is SFU/latency-bound, not tensor-bound, and at T <= 4K+1 tokens it is <1 % of the
// level's FLOPs, so it stays on the SIMT pipes: one warp per query, lanes over keys (scores) / head dims (outputs),
// K/V tiles of 32 keys staged in shared memory, online softmax.  Long bags (20 k tokens) tile over keys the same way.
#include <stdlib.h>

#include "common.cuh"
#include "rowops.cuh"

namespace paths {

namespace {

constexpr int QW = 4;            // queries per warp
constexpr int AW = 4;            // warps per block
constexpr int QB = QW * AW;      // queries per block
constexpr int KT = 32;           // keys per tile
constexpr int MAXDH = 64;


// out[q, h*dh + c] = sum_j softmax_j(scale * q.k_j) v_j[c];  lse[q, h] = logsumexp_j(scale * q.k_j)
__global__ void __launch_bounds__(AW * 32) attn_fwd_kernel(const float* __restrict__ qkv, const SegGrp seg,
                                                           float* __restrict__ out, float* __restrict__ lse, int d, int heads,
                                                           int nq_mode, int skip_T, const Drop dr) {
  pdl_grid_sync();
  const int s = blockIdx.z, h = blockIdx.y, dh = d / heads;
  const Seg sg = seg_of(seg, s);
  if (sg.T <= skip_T) return;                        // short bags belong to the *_small kernels
  const uint64_t dkey = pick(dr.key, sg.g);          // dropout of the attention probabilities (dr.thresh != 0): the normaliser
                                                     // sums every probability, the value sum only the kept ones / (1 - p)
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
      float pd = p;
      if (dr.thresh)
        pd = drop_keep(dkey, (uint32_t)(sg.t0 - sg.r0 + q0 + w * QW + i), ((uint32_t)h << 24) | (uint32_t)(k0 + lane), dr.thresh) ? p * dr.inv_keep : 0.f;
      const int jmax = min(KT, sg.T - k0);
      for (int j = 0; j < jmax; ++j) {
        const float pj = __shfl_sync(0xffffffffu, pd, j);
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
    const size_t orow = nq_mode ? (size_t)sg.os : (size_t)(sg.t0 + q);
    const float inv = 1.f / l[i];
    if (lane < dh) out[orow * d + h * dh + lane] = oa[i] * inv;
    if (lane + 32 < dh) out[orow * d + h * dh + lane + 32] = ob[i] * inv;
    if (lane == 0) lse[orow * heads + h] = m[i] + logf(l[i]);
  }
}

// dq (and Dvec[q, h] = dout . out) for every query: dS_j = p_j (dout.v_j - D), dq = scale * sum_j dS_j k_j
__global__ void __launch_bounds__(AW * 32) attn_bwd_dq_kernel(const float* __restrict__ qkv, const SegGrp seg,
                                                              const float* __restrict__ out, const float* __restrict__ lse,
                                                              const float* __restrict__ dout, float* __restrict__ dqkv,
                                                              float* __restrict__ Dvec, int d, int heads, int nq_mode, int skip_T,
                                                              const Drop dr) {
  pdl_grid_sync();
  const int s = blockIdx.z, h = blockIdx.y, dh = d / heads;
  const Seg sg = seg_of(seg, s);
  if (sg.T <= skip_T) return;
  const uint64_t dkey = pick(dr.key, sg.g);
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
      const size_t orow = nq_mode ? (size_t)sg.os : (size_t)(sg.t0 + q);
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
      if (dr.thresh)                                       // d(kept probability) = dout.v / (1 - p), dropped ones get none
        dp = drop_keep(dkey, (uint32_t)(sg.t0 - sg.r0 + q0 + w * QW + i), ((uint32_t)h << 24) | (uint32_t)(k0 + lane), dr.thresh) ? dp * dr.inv_keep : 0.f;
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
__global__ void __launch_bounds__(AW * 32) attn_bwd_dkv_kernel(const float* __restrict__ qkv, const SegGrp seg,
                                                               const float* __restrict__ lse, const float* __restrict__ dout,
                                                               const float* __restrict__ Dvec, float* __restrict__ dqkv, int d,
                                                               int heads, int nq_mode, int skip_T, const Drop dr) {
  pdl_grid_sync();
  const int s = blockIdx.z, h = blockIdx.y, dh = d / heads;
  const Seg sg = seg_of(seg, s);
  if (sg.T <= skip_T) return;
  const uint64_t dkey = pick(dr.key, sg.g);
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
        const size_t orow = nq_mode ? (size_t)sg.os : (size_t)(sg.t0 + i0 + i);
        qv = qkv[(size_t)(sg.t0 + i0 + i) * ld + h * dh + c] * scale;
        gv = dout[orow * d + h * dh + c];
      }
      Qs[i][c] = qv; Gs[i][c] = gv;
    }
    if (threadIdx.x < KT) {
      const int i = threadIdx.x;
      float lv = 0.f, dv = 0.f;
      if (i0 + i < nq) {
        const size_t orow = nq_mode ? (size_t)sg.os : (size_t)(sg.t0 + i0 + i);
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
      float pm = p;
      if (dr.thresh) {
        const bool keep = drop_keep(dkey, (uint32_t)(sg.t0 - sg.r0 + i0 + lane), ((uint32_t)h << 24) | (uint32_t)(j0 + w * QW + i), dr.thresh);
        pm = keep ? p * dr.inv_keep : 0.f;
        dp = keep ? dp * dr.inv_keep : 0.f;
      }
      const float ds = p * (dp - Ds[lane]);
      const int imax = min(KT, nq - i0);
      for (int ii = 0; ii < imax; ++ii) {
        const float pi = __shfl_sync(0xffffffffu, pm, ii);
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

// ====================================================================================================================
// Short-bag kernels (head_dim 32, T <= 32*NKC <= 128 tokens per slide: every level of the BRCA-shaped workload).  One CTA
// per (head, slide) keeps the head's Q (pre-scaled), K, V (and dO in the backward) in shared memory as [T][33] rows --
// stride 33 makes the row-major ([j][lane]), the transposed ([lane + 32m][c]) and the broadcast access all conflict-free --
// and each warp works on blocks of SQ queries (or keys) with the whole score row in registers: SQ x NKC accumulators per
// lane instead of one shuffle per multiply.  Slides longer than 32*NKC tokens are left to the tiled kernels above.
constexpr int SQ = 8;            // queries (keys in the key-major phase) per warp pass
constexpr int SW = 4;            // warps per CTA
constexpr int SP = 33;           // padded row stride (floats)

// DROP: dropout of the probabilities (common.cuh: Drop) -- the normaliser sums every probability, the value sum the kept ones.
template <int NKC, bool DROP>
__device__ __forceinline__ void attn_fwd_small_body(const float* __restrict__ qkv, const SegGrp& seg, float* __restrict__ out,
                                                    float* __restrict__ lse, int d, int heads, int nq_mode, const Drop& dr) {
  constexpr int TP = 32 * NKC;
  const int h = blockIdx.x, s = blockIdx.y;
  const Seg sg = seg_of(seg, s);
  const uint64_t dkey = DROP ? pick(dr.key, sg.g) : 0;
  const uint32_t hrow0 = (uint32_t)(sg.t0 - sg.r0), hcol = (uint32_t)h << 24;
  const int T = sg.T;
  if (T > TP) return;
  const int nq = nq_mode ? 1 : T;
  extern __shared__ float sm[];
  float* Qs = sm;                       // [TP][SP]
  float* Ks = Qs + TP * SP;
  float* Vs = Ks + TP * SP;
  float* Pw = Vs + TP * SP;             // [SW][SQ][TP]
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const float scale = rsqrtf(32.f);
  const int ld = 3 * d;
  {
    // every load of the head's Q/K/V slice is issued before the first shared-memory store: (row, 16-byte quad) items, 8 per row
    constexpr int NIT = TP * 8 / (SW * 32);
    float4 q4[NIT], k4[NIT], v4[NIT];
#pragma unroll
    for (int k = 0; k < NIT; ++k) {
      const int item = threadIdx.x + k * SW * 32, j = item >> 3, c4 = (item & 7) * 4;
      q4[k] = k4[k] = v4[k] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (j < T) {
        const float* rp = qkv + (size_t)(sg.t0 + j) * ld + h * 32 + c4;
        if (j < nq) q4[k] = __ldg(reinterpret_cast<const float4*>(rp));
        k4[k] = __ldg(reinterpret_cast<const float4*>(rp + d));
        v4[k] = __ldg(reinterpret_cast<const float4*>(rp + 2 * d));
      }
    }
#pragma unroll
    for (int k = 0; k < NIT; ++k) {
      const int item = threadIdx.x + k * SW * 32, j = item >> 3, c4 = (item & 7) * 4;
      float* qd = Qs + j * SP + c4; float* kd = Ks + j * SP + c4; float* vd = Vs + j * SP + c4;
      qd[0] = q4[k].x * scale; qd[1] = q4[k].y * scale; qd[2] = q4[k].z * scale; qd[3] = q4[k].w * scale;
      kd[0] = k4[k].x; kd[1] = k4[k].y; kd[2] = k4[k].z; kd[3] = k4[k].w;
      vd[0] = v4[k].x; vd[1] = v4[k].y; vd[2] = v4[k].z; vd[3] = v4[k].w;
    }
  }
  __syncthreads();
  float* P = Pw + w * SQ * TP;
  for (int i0 = w * SQ; i0 < nq; i0 += SW * SQ) {
    float sc[SQ][NKC];
#pragma unroll
    for (int i = 0; i < SQ; ++i)
#pragma unroll
      for (int m = 0; m < NKC; ++m) sc[i][m] = 0.f;
#pragma unroll 4
    for (int c = 0; c < 32; ++c) {
      float kv[NKC];
#pragma unroll
      for (int m = 0; m < NKC; ++m) kv[m] = Ks[(lane + 32 * m) * SP + c];
#pragma unroll
      for (int i = 0; i < SQ; ++i) {
        const float qv = Qs[(i0 + i) * SP + c];        // rows past nq are zero (inside TP because i0 + SQ <= TP)
#pragma unroll
        for (int m = 0; m < NKC; ++m) sc[i][m] = fmaf(qv, kv[m], sc[i][m]);
      }
    }
    float linv[SQ], lsev[SQ];
#pragma unroll
    for (int i = 0; i < SQ; ++i) {
      float mx = -INFINITY;
#pragma unroll
      for (int m = 0; m < NKC; ++m) {
        if (lane + 32 * m >= T) sc[i][m] = -INFINITY;
        mx = fmaxf(mx, sc[i][m]);
      }
      mx = warp_max(mx);
      float sum = 0.f;
#pragma unroll
      for (int m = 0; m < NKC; ++m) {
        const float p = expf(sc[i][m] - mx);            // exp(-inf) = 0 at masked keys
        if (DROP) P[i * TP + lane + 32 * m] = drop_keep(dkey, hrow0 + i0 + i, hcol | (uint32_t)(lane + 32 * m), dr.thresh) ? p * dr.inv_keep : 0.f;
        else P[i * TP + lane + 32 * m] = p;
        sum += p;
      }
      sum = warp_sum(sum);
      linv[i] = 1.f / sum; lsev[i] = mx + logf(sum);
    }
    __syncwarp();
    float o[SQ];
#pragma unroll
    for (int i = 0; i < SQ; ++i) o[i] = 0.f;
    const int T4 = (T + 3) & ~3;
    for (int j = 0; j < T4; j += 4) {
      const float v0 = Vs[j * SP + lane], v1 = Vs[(j + 1) * SP + lane], v2 = Vs[(j + 2) * SP + lane], v3 = Vs[(j + 3) * SP + lane];
#pragma unroll
      for (int i = 0; i < SQ; ++i) {
        const float4 p = *reinterpret_cast<const float4*>(P + i * TP + j);
        o[i] = fmaf(p.x, v0, fmaf(p.y, v1, fmaf(p.z, v2, fmaf(p.w, v3, o[i]))));
      }
    }
#pragma unroll
    for (int i = 0; i < SQ; ++i) {
      const int q = i0 + i;
      if (q < nq) {
        const size_t orow = nq_mode ? (size_t)sg.os : (size_t)(sg.t0 + q);
        out[orow * d + h * 32 + lane] = o[i] * linv[i];
        if (lane == 0) lse[orow * heads + h] = lsev[i];
      }
    }
    __syncwarp();
  }
}

// The launch is compiled for the longest bag (32 * NKC tokens); a CTA whose own bag is shorter runs the body compiled for its
// length class: the score rows of a 57-token bag then span 64 key slots instead of 96 (a third less FMA work in the score phases).
template <int NKC, bool DROP>
__global__ void __launch_bounds__(SW * 32) attn_fwd_small_kernel(const float* __restrict__ qkv, const SegGrp seg,
                                                                 float* __restrict__ out, float* __restrict__ lse, int d, int heads,
                                                                 int nq_mode, const Drop dr) {
  pdl_grid_sync();
  const int T = seg_of(seg, blockIdx.y).T;
  if (NKC >= 2 && T <= 32) attn_fwd_small_body<1, DROP>(qkv, seg, out, lse, d, heads, nq_mode, dr);
  else if (NKC >= 3 && T <= 64) attn_fwd_small_body<2, DROP>(qkv, seg, out, lse, d, heads, nq_mode, dr);
  else if (NKC >= 4 && T <= 96) attn_fwd_small_body<3, DROP>(qkv, seg, out, lse, d, heads, nq_mode, dr);
  else attn_fwd_small_body<NKC, DROP>(qkv, seg, out, lse, d, heads, nq_mode, dr);
}

// Backward of the short-bag attention: phase A is query-major (dq), phase B key-major (dk, dv); both recompute the
// probabilities from the saved log-sum-exp, so nothing is accumulated across warps and no atomics are needed.
template <int NKC, bool DROP>
__device__ __forceinline__ void attn_bwd_small_body(const float* __restrict__ qkv, const SegGrp& seg, const float* __restrict__ out,
                                                    const float* __restrict__ lse, const float* __restrict__ dout,
                                                    float* __restrict__ dqkv, int d, int heads, int nq_mode, const Drop& dr) {
  constexpr int TP = 32 * NKC;
  const int h = blockIdx.x, s = blockIdx.y;
  const Seg sg = seg_of(seg, s);
  const uint64_t dkey = DROP ? pick(dr.key, sg.g) : 0;
  const uint32_t hrow0 = (uint32_t)(sg.t0 - sg.r0), hcol = (uint32_t)h << 24;
  const int T = sg.T;
  if (T > TP) return;
  const int nq = nq_mode ? 1 : T;
  extern __shared__ float sm[];
  float* Qs = sm;                       // [TP][SP], pre-scaled
  float* Ks = Qs + TP * SP;
  float* Vs = Ks + TP * SP;
  float* Gs = Vs + TP * SP;             // dO
  float* Ls = Gs + TP * SP;             // [TP] log-sum-exp per query
  float* Ds = Ls + TP;                  // [TP] dO . O per query
  float* Pw = Ds + TP;                  // [SW][2][SQ][TP]
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const float scale = rsqrtf(32.f);
  const int ld = 3 * d;
  {
    constexpr int NIT = TP * 8 / (SW * 32);
    float4 a4[NIT], b4[NIT], c4r[NIT];
#pragma unroll
    for (int k = 0; k < NIT; ++k) {                     // K, V
      const int item = threadIdx.x + k * SW * 32, j = item >> 3, c4 = (item & 7) * 4;
      a4[k] = b4[k] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (j < T) {
        const float* rp = qkv + (size_t)(sg.t0 + j) * ld + h * 32 + c4;
        a4[k] = __ldg(reinterpret_cast<const float4*>(rp + d));
        b4[k] = __ldg(reinterpret_cast<const float4*>(rp + 2 * d));
      }
    }
#pragma unroll
    for (int k = 0; k < NIT; ++k) {
      const int item = threadIdx.x + k * SW * 32, j = item >> 3, c4 = (item & 7) * 4;
      float* kd = Ks + j * SP + c4; float* vd = Vs + j * SP + c4;
      kd[0] = a4[k].x; kd[1] = a4[k].y; kd[2] = a4[k].z; kd[3] = a4[k].w;
      vd[0] = b4[k].x; vd[1] = b4[k].y; vd[2] = b4[k].z; vd[3] = b4[k].w;
    }
#pragma unroll
    for (int k = 0; k < NIT; ++k) {                     // Q, dO, O (-> D), lse
      const int item = threadIdx.x + k * SW * 32, j = item >> 3, c4 = (item & 7) * 4;
      a4[k] = b4[k] = c4r[k] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (j < nq) {
        const size_t orow = nq_mode ? (size_t)sg.os : (size_t)(sg.t0 + j);
        a4[k] = __ldg(reinterpret_cast<const float4*>(qkv + (size_t)(sg.t0 + j) * ld + h * 32 + c4));
        b4[k] = __ldg(reinterpret_cast<const float4*>(dout + orow * d + h * 32 + c4));
        c4r[k] = __ldg(reinterpret_cast<const float4*>(out + orow * d + h * 32 + c4));
      }
    }
#pragma unroll
    for (int k = 0; k < NIT; ++k) {
      const int item = threadIdx.x + k * SW * 32, j = item >> 3, c4 = (item & 7) * 4;
      float* qd = Qs + j * SP + c4; float* gd = Gs + j * SP + c4;
      qd[0] = a4[k].x * scale; qd[1] = a4[k].y * scale; qd[2] = a4[k].z * scale; qd[3] = a4[k].w * scale;
      gd[0] = b4[k].x; gd[1] = b4[k].y; gd[2] = b4[k].z; gd[3] = b4[k].w;
      float dd = b4[k].x * c4r[k].x + b4[k].y * c4r[k].y + b4[k].z * c4r[k].z + b4[k].w * c4r[k].w;
      dd += __shfl_xor_sync(0xffffffffu, dd, 1);
      dd += __shfl_xor_sync(0xffffffffu, dd, 2);
      dd += __shfl_xor_sync(0xffffffffu, dd, 4);
      if ((item & 7) == 0) {
        Ds[j] = dd;
        Ls[j] = (j < nq) ? lse[(nq_mode ? (size_t)sg.os : (size_t)(sg.t0 + j)) * heads + h] : 0.f;
      }
    }
  }
  __syncthreads();
  float* P = Pw + w * 2 * SQ * TP;      // probabilities (phase B)
  float* S = P + SQ * TP;               // dS
  // ---- phase A: dq_i = scale * sum_j dS_ij k_j,  dS_ij = p_ij (dO_i . v_j - D_i)
  for (int i0 = w * SQ; i0 < nq; i0 += SW * SQ) {
    float sc[SQ][NKC], dp[SQ][NKC];
#pragma unroll
    for (int i = 0; i < SQ; ++i)
#pragma unroll
      for (int m = 0; m < NKC; ++m) { sc[i][m] = 0.f; dp[i][m] = 0.f; }
#pragma unroll 2
    for (int c = 0; c < 32; ++c) {
      float kv[NKC], vv[NKC];
#pragma unroll
      for (int m = 0; m < NKC; ++m) { kv[m] = Ks[(lane + 32 * m) * SP + c]; vv[m] = Vs[(lane + 32 * m) * SP + c]; }
#pragma unroll
      for (int i = 0; i < SQ; ++i) {
        const float qv = Qs[(i0 + i) * SP + c], gv = Gs[(i0 + i) * SP + c];
#pragma unroll
        for (int m = 0; m < NKC; ++m) { sc[i][m] = fmaf(qv, kv[m], sc[i][m]); dp[i][m] = fmaf(gv, vv[m], dp[i][m]); }
      }
    }
#pragma unroll
    for (int i = 0; i < SQ; ++i) {
      const float L = Ls[i0 + i], Dq = Ds[i0 + i];
#pragma unroll
      for (int m = 0; m < NKC; ++m) {
        const float p = (lane + 32 * m < T) ? expf(sc[i][m] - L) : 0.f;
        float dpm = dp[i][m];                             // d(kept probability) = dO.v / (1 - p); dropped ones get none
        if (DROP) dpm = drop_keep(dkey, hrow0 + i0 + i, hcol | (uint32_t)(lane + 32 * m), dr.thresh) ? dpm * dr.inv_keep : 0.f;
        S[i * TP + lane + 32 * m] = p * (dpm - Dq);
      }
    }
    __syncwarp();
    float a[SQ];
#pragma unroll
    for (int i = 0; i < SQ; ++i) a[i] = 0.f;
    const int T4 = (T + 3) & ~3;
    for (int j = 0; j < T4; j += 4) {
      const float k0 = Ks[j * SP + lane], k1 = Ks[(j + 1) * SP + lane], k2 = Ks[(j + 2) * SP + lane], k3 = Ks[(j + 3) * SP + lane];
#pragma unroll
      for (int i = 0; i < SQ; ++i) {
        const float4 ds = *reinterpret_cast<const float4*>(S + i * TP + j);
        a[i] = fmaf(ds.x, k0, fmaf(ds.y, k1, fmaf(ds.z, k2, fmaf(ds.w, k3, a[i]))));
      }
    }
#pragma unroll
    for (int i = 0; i < SQ; ++i)
      if (i0 + i < nq) dqkv[(size_t)(sg.t0 + i0 + i) * ld + h * 32 + lane] = a[i] * scale;
    __syncwarp();
  }
  // ---- phase B: dv_j = sum_i p_ij dO_i,  dk_j = sum_i dS_ij (scale q_i); lanes run over the queries
  const int nqc = (nq + 31) >> 5;
  for (int j0 = w * SQ; j0 < T; j0 += SW * SQ) {
    float sc[SQ][NKC], dp[SQ][NKC];
#pragma unroll
    for (int i = 0; i < SQ; ++i)
#pragma unroll
      for (int m = 0; m < NKC; ++m) { sc[i][m] = 0.f; dp[i][m] = 0.f; }
#pragma unroll 2
    for (int c = 0; c < 32; ++c) {
      float qv[NKC], gv[NKC];
#pragma unroll
      for (int m = 0; m < NKC; ++m) {
        qv[m] = 0.f; gv[m] = 0.f;
        if (m < nqc) { qv[m] = Qs[(lane + 32 * m) * SP + c]; gv[m] = Gs[(lane + 32 * m) * SP + c]; }
      }
#pragma unroll
      for (int i = 0; i < SQ; ++i) {
        const float kv = Ks[(j0 + i) * SP + c], vv = Vs[(j0 + i) * SP + c];
#pragma unroll
        for (int m = 0; m < NKC; ++m) { sc[i][m] = fmaf(kv, qv[m], sc[i][m]); dp[i][m] = fmaf(vv, gv[m], dp[i][m]); }
      }
    }
#pragma unroll
    for (int m = 0; m < NKC; ++m) {
      const int qi = lane + 32 * m;
      const float L = Ls[qi], Dq = Ds[qi];
#pragma unroll
      for (int i = 0; i < SQ; ++i) {
        const float p = (qi < nq) ? expf(sc[i][m] - L) : 0.f;
        float pm = p, dpm = dp[i][m];
        if (DROP) {
          const bool keep = drop_keep(dkey, hrow0 + qi, hcol | (uint32_t)(j0 + i), dr.thresh);
          pm = keep ? p * dr.inv_keep : 0.f; dpm = keep ? dpm * dr.inv_keep : 0.f;
        }
        P[i * TP + qi] = pm;
        S[i * TP + qi] = p * (dpm - Dq);
      }
    }
    __syncwarp();
    float dk[SQ], dv[SQ];
#pragma unroll
    for (int i = 0; i < SQ; ++i) { dk[i] = 0.f; dv[i] = 0.f; }
    const int n4 = (nq + 3) & ~3;
    for (int q = 0; q < n4; q += 4) {
      const float g0 = Gs[q * SP + lane], g1 = Gs[(q + 1) * SP + lane], g2 = Gs[(q + 2) * SP + lane], g3 = Gs[(q + 3) * SP + lane];
      const float q0 = Qs[q * SP + lane], q1 = Qs[(q + 1) * SP + lane], q2 = Qs[(q + 2) * SP + lane], q3 = Qs[(q + 3) * SP + lane];
#pragma unroll
      for (int i = 0; i < SQ; ++i) {
        const float4 p = *reinterpret_cast<const float4*>(P + i * TP + q);
        const float4 ds = *reinterpret_cast<const float4*>(S + i * TP + q);
        dv[i] = fmaf(p.x, g0, fmaf(p.y, g1, fmaf(p.z, g2, fmaf(p.w, g3, dv[i]))));
        dk[i] = fmaf(ds.x, q0, fmaf(ds.y, q1, fmaf(ds.z, q2, fmaf(ds.w, q3, dk[i]))));
      }
    }
#pragma unroll
    for (int i = 0; i < SQ; ++i) {
      if (j0 + i < T) {
        float* rp = dqkv + (size_t)(sg.t0 + j0 + i) * ld + h * 32 + lane;
        rp[d] = dk[i]; rp[2 * d] = dv[i];
      }
    }
    __syncwarp();
  }
}

template <int NKC, bool DROP>
__global__ void __launch_bounds__(SW * 32) attn_bwd_small_kernel(const float* __restrict__ qkv, const SegGrp seg,
                                                                 const float* __restrict__ out, const float* __restrict__ lse,
                                                                 const float* __restrict__ dout, float* __restrict__ dqkv, int d,
                                                                 int heads, int nq_mode, const Drop dr) {
  pdl_grid_sync();
  const int T = seg_of(seg, blockIdx.y).T;
  if (NKC >= 2 && T <= 32) attn_bwd_small_body<1, DROP>(qkv, seg, out, lse, dout, dqkv, d, heads, nq_mode, dr);
  else if (NKC >= 3 && T <= 64) attn_bwd_small_body<2, DROP>(qkv, seg, out, lse, dout, dqkv, d, heads, nq_mode, dr);
  else if (NKC >= 4 && T <= 96) attn_bwd_small_body<3, DROP>(qkv, seg, out, lse, dout, dqkv, d, heads, nq_mode, dr);
  else attn_bwd_small_body<NKC, DROP>(qkv, seg, out, lse, dout, dqkv, d, heads, nq_mode, dr);
}

}  // namespace


namespace {
// ====================================================================================================================
// Row-per-thread-pair kernels (head_dim 32, bags of at most NT tokens: every level of the BRCA-shaped workload).  One CTA per
// (head, slide) stages the head's K and V rows in shared memory as float4 rows; the two lanes of a pair own query i and one
// 16-wide half of the head each: pre-scaled query half, running max / sum and 16 output accumulators live in registers, the
// pair's partial dot products meet through one shuffle.  Every K / V access is a 16-byte load of a row all pairs of the warp
// read at the same time (two distinct addresses per warp: broadcast, no bank conflicts), and the score row never touches
// memory.  Keys are consumed four at a time so the online-softmax rescale is paid once per four keys.
__device__ __forceinline__ float dot4(const float4& a, const float4& b, float acc) {
  return fmaf(a.x, b.x, fmaf(a.y, b.y, fmaf(a.z, b.z, fmaf(a.w, b.w, acc))));
}
__device__ __forceinline__ void axpy4(float s, const float4& x, float4& y) {
  y.x = fmaf(s, x.x, y.x); y.y = fmaf(s, x.y, y.y); y.z = fmaf(s, x.z, y.z); y.w = fmaf(s, x.w, y.w);
}
__device__ __forceinline__ float pair_sum(float v) { return v + __shfl_xor_sync(0xffffffffu, v, 1); }

template <int NT>
__global__ void __launch_bounds__(2 * NT) attn_fwd_tq_kernel(const float* __restrict__ qkv, const SegGrp seg, float* __restrict__ out,
                                                             float* __restrict__ lse, int d, int heads) {
  pdl_grid_sync();
  const int h = blockIdx.x;
  const Seg sg = seg_of(seg, blockIdx.y);
  const int T = sg.T;
  if (T > NT) return;                                    // longer bags belong to the tiled kernels
  extern __shared__ float4 sm4[];
  float4* K4 = sm4;                                      // [NT + 3][8]  (3 zero rows of slack for the 4-key blocks)
  float4* V4 = sm4 + (NT + 3) * 8;
  const int ld = 3 * d;
  const float* base = qkv + (size_t)sg.t0 * ld + h * 32;
  for (int idx = threadIdx.x; idx < (NT + 3) * 8; idx += 2 * NT) {
    const int j = idx >> 3, c = idx & 7;
    float4 kv = make_float4(0.f, 0.f, 0.f, 0.f), vv = kv;
    if (j < T) {
      kv = __ldg(reinterpret_cast<const float4*>(base + (size_t)j * ld + d) + c);
      vv = __ldg(reinterpret_cast<const float4*>(base + (size_t)j * ld + 2 * d) + c);
    }
    K4[idx] = kv; V4[idx] = vv;
  }
  const int i = threadIdx.x >> 1, hf = (threadIdx.x & 1) * 4;    // query row, first float4 of this lane's half
  const bool live = i < T;                                        // both lanes of a pair agree; dead pairs still shuffle
  float4 q[4];
  {
    const float scale = rsqrtf(32.f);
    const float4* qp = reinterpret_cast<const float4*>(base + (size_t)(live ? i : 0) * ld) + hf;
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      q[c] = __ldg(qp + c);
      q[c].x *= scale; q[c].y *= scale; q[c].z *= scale; q[c].w *= scale;
    }
  }
  __syncthreads();
  float4 o[4];
#pragma unroll
  for (int c = 0; c < 4; ++c) o[c] = make_float4(0.f, 0.f, 0.f, 0.f);
  float m = -INFINITY, l = 0.f;
  for (int j0 = 0; j0 < T; j0 += 4) {
    float sc[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const float4* kr = K4 + (j0 + u) * 8 + hf;
      const float a = dot4(q[0], kr[0], 0.f) + dot4(q[1], kr[1], 0.f), b = dot4(q[2], kr[2], 0.f) + dot4(q[3], kr[3], 0.f);
      sc[u] = pair_sum(a + b);
      if (j0 + u >= T) sc[u] = -INFINITY;
    }
    const float mn = fmaxf(fmaxf(m, fmaxf(sc[0], sc[1])), fmaxf(sc[2], sc[3]));
    const float corr = expf(m - mn);                     // exp(-inf) = 0 on the first block
    float p[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) p[u] = expf(sc[u] - mn);
    l = l * corr + ((p[0] + p[1]) + (p[2] + p[3]));
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      o[c].x *= corr; o[c].y *= corr; o[c].z *= corr; o[c].w *= corr;
#pragma unroll
      for (int u = 0; u < 4; ++u) axpy4(p[u], V4[(j0 + u) * 8 + hf + c], o[c]);
    }
    m = mn;
  }
  if (!live) return;
  const float inv = 1.f / l;
  float4* op = reinterpret_cast<float4*>(out + (size_t)(sg.t0 + i) * d + h * 32) + hf;
#pragma unroll
  for (int c = 0; c < 4; ++c) op[c] = make_float4(o[c].x * inv, o[c].y * inv, o[c].z * inv, o[c].w * inv);
  if (hf == 0) lse[(size_t)(sg.t0 + i) * heads + h] = m + logf(l);
}

// Backward, same layout.  Phase A (pair = query i): dS_ij = p_ij (dO_i . v_j - D_i), dq_i = scale * sum_j dS_ij k_j.
// Phase B (pair = key j): dv_j = sum_i p_ij dO_i, dk_j = sum_i dS_ij (scale q_i).  Probabilities are recomputed from the saved
// log-sum-exp in both phases; nothing is accumulated across pairs, so no atomics and a fixed summation order.
template <int NT>
__global__ void __launch_bounds__(2 * NT, NT == 96 ? 3 : NT == 128 ? 2 : NT == 64 ? 4 : 8) attn_bwd_tq_kernel(const float* __restrict__ qkv, const SegGrp seg, const float* __restrict__ out,
                                                             const float* __restrict__ lse, const float* __restrict__ dout,
                                                             float* __restrict__ dqkv, int d, int heads) {
  pdl_grid_sync();
  const int h = blockIdx.x;
  const Seg sg = seg_of(seg, blockIdx.y);
  const int T = sg.T;
  if (T > NT) return;
  extern __shared__ float4 sm4[];
  float4* Q4 = sm4;                                      // [NT][8] pre-scaled queries
  float4* K4 = Q4 + NT * 8;
  float4* V4 = K4 + NT * 8;
  float4* G4 = V4 + NT * 8;                              // dO
  float* Ls = reinterpret_cast<float*>(G4 + NT * 8);     // [NT] log-sum-exp per query
  float* Ds = Ls + NT;                                   // [NT] dO . O per query
  const int ld = 3 * d;
  const float scale = rsqrtf(32.f);
  const float* base = qkv + (size_t)sg.t0 * ld + h * 32;
  for (int idx = threadIdx.x; idx < NT * 8; idx += 2 * NT) {
    const int j = idx >> 3, c = idx & 7;
    float4 qv = make_float4(0.f, 0.f, 0.f, 0.f), kv = qv, vv = qv, gv = qv;
    if (j < T) {
      const float4* rp = reinterpret_cast<const float4*>(base + (size_t)j * ld) + c;
      qv = __ldg(rp); kv = __ldg(rp + d / 4); vv = __ldg(rp + d / 2);
      gv = __ldg(reinterpret_cast<const float4*>(dout + (size_t)(sg.t0 + j) * d + h * 32) + c);
      qv.x *= scale; qv.y *= scale; qv.z *= scale; qv.w *= scale;
    }
    Q4[idx] = qv; K4[idx] = kv; V4[idx] = vv; G4[idx] = gv;
  }
  const int i = threadIdx.x >> 1, hf = (threadIdx.x & 1) * 4;
  const bool live = i < T;
  float L = 0.f, Dq = 0.f;
  {
    float part = 0.f;
    if (live) {
      const float4* op = reinterpret_cast<const float4*>(out + (size_t)(sg.t0 + i) * d + h * 32) + hf;
      const float4* gp = reinterpret_cast<const float4*>(dout + (size_t)(sg.t0 + i) * d + h * 32) + hf;
#pragma unroll
      for (int c = 0; c < 4; ++c) part = dot4(__ldg(op + c), __ldg(gp + c), part);
      L = lse[(size_t)(sg.t0 + i) * heads + h];
    }
    Dq = pair_sum(part);
    if (hf == 0) { Ls[i] = L; Ds[i] = Dq; }
  }
  __syncthreads();
  {
    // ---- phase A: this pair's query row
    float4 q[4], g[4], dq[4];
#pragma unroll
    for (int c = 0; c < 4; ++c) { q[c] = Q4[(live ? i : 0) * 8 + hf + c]; g[c] = G4[(live ? i : 0) * 8 + hf + c]; dq[c] = make_float4(0.f, 0.f, 0.f, 0.f); }
    for (int j = 0; j < T; ++j) {
      float4 k[4];
      float sa = 0.f, sb = 0.f, da = 0.f, db = 0.f;
#pragma unroll
      for (int c = 0; c < 4; ++c) k[c] = K4[j * 8 + hf + c];
      sa = dot4(q[0], k[0], dot4(q[1], k[1], 0.f)); sb = dot4(q[2], k[2], dot4(q[3], k[3], 0.f));
      da = dot4(g[0], V4[j * 8 + hf], dot4(g[1], V4[j * 8 + hf + 1], 0.f)); db = dot4(g[2], V4[j * 8 + hf + 2], dot4(g[3], V4[j * 8 + hf + 3], 0.f));
      const float sc = pair_sum(sa + sb), dp = pair_sum(da + db);
      const float ds = expf(sc - L) * (dp - Dq);
#pragma unroll
      for (int c = 0; c < 4; ++c) axpy4(ds, k[c], dq[c]);
    }
    if (live) {
      float4* dqp = reinterpret_cast<float4*>(dqkv + (size_t)(sg.t0 + i) * ld + h * 32) + hf;
#pragma unroll
      for (int c = 0; c < 4; ++c) dqp[c] = make_float4(dq[c].x * scale, dq[c].y * scale, dq[c].z * scale, dq[c].w * scale);
    }
  }
  {
    // ---- phase B: this pair's key row
    float4 k[4], v[4], dk[4], dv[4];
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      k[c] = K4[(live ? i : 0) * 8 + hf + c]; v[c] = V4[(live ? i : 0) * 8 + hf + c];
      dk[c] = make_float4(0.f, 0.f, 0.f, 0.f); dv[c] = dk[c];
    }
    for (int qi = 0; qi < T; ++qi) {
      float4 a[4], b[4];
#pragma unroll
      for (int c = 0; c < 4; ++c) { a[c] = Q4[qi * 8 + hf + c]; b[c] = G4[qi * 8 + hf + c]; }
      const float sa = dot4(k[0], a[0], dot4(k[1], a[1], 0.f)), sb = dot4(k[2], a[2], dot4(k[3], a[3], 0.f));
      const float da = dot4(v[0], b[0], dot4(v[1], b[1], 0.f)), db = dot4(v[2], b[2], dot4(v[3], b[3], 0.f));
      const float sc = pair_sum(sa + sb), dp = pair_sum(da + db);
      const float p = expf(sc - Ls[qi]);
      const float ds = p * (dp - Ds[qi]);
#pragma unroll
      for (int c = 0; c < 4; ++c) { axpy4(ds, a[c], dk[c]); axpy4(p, b[c], dv[c]); }
    }
    if (live) {
      float4* rp = reinterpret_cast<float4*>(dqkv + (size_t)(sg.t0 + i) * ld + h * 32) + hf;
#pragma unroll
      for (int c = 0; c < 4; ++c) { rp[d / 4 + c] = dk[c]; rp[d / 2 + c] = dv[c]; }
    }
  }
}

template <int NT>
int launch_fwd_tq(const float* qkv, const SegGrp& seg, float* out, float* lse, int d, int heads, int B, cudaStream_t st) {
  static bool cfg = false;
  const int bytes = 2 * (NT + 3) * 8 * (int)sizeof(float4);
  if (!cfg) { PATHS_CK(cudaFuncSetAttribute(attn_fwd_tq_kernel<NT>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes)); cfg = true; }
  pdl(attn_fwd_tq_kernel<NT>, dim3(heads, B), 2 * NT, bytes, st)(qkv, seg, out, lse, d, heads);
  PATHS_LAUNCH_CK();
  return 0;
}
template <int NT>
int launch_bwd_tq(const float* qkv, const SegGrp& seg, const float* out, const float* lse, const float* dout, float* dqkv, int d, int heads,
                  int B, cudaStream_t st) {
  static bool cfg = false;
  const int bytes = 4 * NT * 8 * (int)sizeof(float4) + 2 * NT * (int)sizeof(float);
  if (!cfg) { PATHS_CK(cudaFuncSetAttribute(attn_bwd_tq_kernel<NT>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes)); cfg = true; }
  pdl(attn_bwd_tq_kernel<NT>, dim3(heads, B), 2 * NT, bytes, st)(qkv, seg, out, lse, dout, dqkv, d, heads);
  PATHS_LAUNCH_CK();
  return 0;
}
}  // namespace

namespace {
// ====================================================================================================================
// Single-query attention (nq_mode 1: the last decoder layer is read at the special token only, aggregator.py:75), head_dim 32.
// One warp per (slide, head), any bag length: lanes run over the keys of a 32-key chunk for the scores (each lane reads its
// key row), then over the head dims for the value sum (coalesced rows, probabilities by shuffle); online softmax over chunks.
// keys j0 = first, first + stride, ... in chunks of 32: running max m, normaliser l, value sum o (this lane's head dim)
__device__ __forceinline__ void nq1_fwd_chunks(const float* __restrict__ base, int ld, int d, int T, int first, int stride, float qv, int lane,
                                               float& m, float& l, float& o) {
  for (int j0 = first; j0 < T; j0 += stride) {
    const int j = j0 + lane;
    float4 k4[8];
    if (j < T) {
      const float4* kr = reinterpret_cast<const float4*>(base + (size_t)j * ld + d);
#pragma unroll
      for (int c = 0; c < 8; ++c) k4[c] = __ldg(kr + c);
    } else {
#pragma unroll
      for (int c = 0; c < 8; ++c) k4[c] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    float sc = 0.f;
#pragma unroll
    for (int c = 0; c < 8; ++c) {                        // shuffles are executed by every lane
      sc = fmaf(__shfl_sync(0xffffffffu, qv, 4 * c), k4[c].x, sc);
      sc = fmaf(__shfl_sync(0xffffffffu, qv, 4 * c + 1), k4[c].y, sc);
      sc = fmaf(__shfl_sync(0xffffffffu, qv, 4 * c + 2), k4[c].z, sc);
      sc = fmaf(__shfl_sync(0xffffffffu, qv, 4 * c + 3), k4[c].w, sc);
    }
    if (j >= T) sc = -INFINITY;
    const float mn = fmaxf(m, warp_max(sc));
    const float p = (j < T) ? expf(sc - mn) : 0.f;
    const float corr = expf(m - mn);
    l = l * corr + warp_sum(p);
    o *= corr;
    const int jmax = min(32, T - j0);
#pragma unroll 8
    for (int jj = 0; jj < 32; ++jj) {                    // fixed trip count: the value-row loads of a chunk are all in flight
      const float vv = (jj < jmax) ? __ldg(base + (size_t)(j0 + jj) * ld + 2 * d + lane) : 0.f;
      o = fmaf(__shfl_sync(0xffffffffu, p, jj), vv, o);
    }
    m = mn;
  }
}

constexpr int NQ1_WIDE_T = 1024;     // bags above this many tokens get a whole CTA per (slide, head) instead of one warp
constexpr int NQ1_WIDE_WARPS = 32;

__global__ void __launch_bounds__(128) attn_fwd_nq1_kernel(const float* __restrict__ qkv, const SegGrp seg,
                                                           float* __restrict__ out, float* __restrict__ lse, int d, int heads, int B,
                                                           int skip_above) {
  pdl_grid_sync();
  const int wid = blockIdx.x * 4 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (wid >= B * heads) return;
  const int s = wid / heads, h = wid % heads;
  const Seg sg = seg_of(seg, s);
  if (sg.T > skip_above) return;                         // the wide kernel's bag
  const int ld = 3 * d;
  const float* base = qkv + (size_t)sg.t0 * ld + h * 32;
  const float qv = base[lane] * rsqrtf(32.f);
  float m = -INFINITY, l = 0.f, o = 0.f;
  nq1_fwd_chunks(base, ld, d, sg.T, 0, 32, qv, lane, m, l, o);
  out[(size_t)sg.os * d + h * 32 + lane] = o / l;
  if (lane == 0) lse[(size_t)sg.os * heads + h] = m + logf(l);
}

// Long bags (the 20 k-token bag of cfg 4): one CTA of 32 warps per (slide, head), the warps take interleaved 32-key chunks with
// their own running (m, l, o) and are merged through shared memory.
__global__ void __launch_bounds__(NQ1_WIDE_WARPS * 32) attn_fwd_nq1_wide_kernel(const float* __restrict__ qkv, const SegGrp seg,
                                                                                float* __restrict__ out, float* __restrict__ lse, int d,
                                                                                int heads) {
  pdl_grid_sync();
  const int s = blockIdx.x / heads, h = blockIdx.x % heads;
  const Seg sg = seg_of(seg, s);
  if (sg.T <= NQ1_WIDE_T) return;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int ld = 3 * d;
  const float* base = qkv + (size_t)sg.t0 * ld + h * 32;
  const float qv = base[lane] * rsqrtf(32.f);
  float m = -INFINITY, l = 0.f, o = 0.f;
  nq1_fwd_chunks(base, ld, d, sg.T, w * 32, NQ1_WIDE_WARPS * 32, qv, lane, m, l, o);
  __shared__ float sm[NQ1_WIDE_WARPS], sl[NQ1_WIDE_WARPS], so[NQ1_WIDE_WARPS][32];
  if (lane == 0) { sm[w] = m; sl[w] = l; }
  so[w][lane] = o;
  __syncthreads();
  if (w == 0) {
    float M = -INFINITY;
    for (int k = 0; k < NQ1_WIDE_WARPS; ++k) M = fmaxf(M, sm[k]);
    float Ls = 0.f, O = 0.f;
    for (int k = 0; k < NQ1_WIDE_WARPS; ++k) {           // fixed order: deterministic
      const float c = expf(sm[k] - M);                   // warps that saw no key: exp(-inf) = 0
      Ls = fmaf(sl[k], c, Ls); O = fmaf(so[k][lane], c, O);
    }
    out[(size_t)sg.os * d + h * 32 + lane] = O / Ls;
    if (lane == 0) lse[(size_t)sg.os * heads + h] = M + logf(Ls);
  }
}

// Backward of the single-query attention; writes the q, k and v parts of every dqkv row of the slide (q part is zero
// except at the special token), so the caller needs no zero-fill.
// keys j0 = first, first + stride, ...: dk, dv of every key row (and zeros in the q part of the non-query rows); returns this
// lane's share of dq (unscaled)
__device__ __forceinline__ float nq1_bwd_chunks(const float* __restrict__ base, float* __restrict__ dbase, int ld, int d, int T, int first,
                                                int stride, float qv, float g, float Dq, float L, int lane) {
  float dq = 0.f;
  for (int j0 = first; j0 < T; j0 += stride) {
    const int j = j0 + lane;
    float sc = 0.f, dp = 0.f;
    float4 k4[8], v4[8];
    if (j < T) {
      const float4* kr = reinterpret_cast<const float4*>(base + (size_t)j * ld + d);
      const float4* vr = reinterpret_cast<const float4*>(base + (size_t)j * ld + 2 * d);
#pragma unroll
      for (int c = 0; c < 8; ++c) { k4[c] = __ldg(kr + c); v4[c] = __ldg(vr + c); }
    } else {
#pragma unroll
      for (int c = 0; c < 8; ++c) { k4[c] = make_float4(0.f, 0.f, 0.f, 0.f); v4[c] = k4[c]; }
    }
#pragma unroll
    for (int c = 0; c < 8; ++c) {
      sc = fmaf(__shfl_sync(0xffffffffu, qv, 4 * c), k4[c].x, sc);     dp = fmaf(__shfl_sync(0xffffffffu, g, 4 * c), v4[c].x, dp);
      sc = fmaf(__shfl_sync(0xffffffffu, qv, 4 * c + 1), k4[c].y, sc); dp = fmaf(__shfl_sync(0xffffffffu, g, 4 * c + 1), v4[c].y, dp);
      sc = fmaf(__shfl_sync(0xffffffffu, qv, 4 * c + 2), k4[c].z, sc); dp = fmaf(__shfl_sync(0xffffffffu, g, 4 * c + 2), v4[c].z, dp);
      sc = fmaf(__shfl_sync(0xffffffffu, qv, 4 * c + 3), k4[c].w, sc); dp = fmaf(__shfl_sync(0xffffffffu, g, 4 * c + 3), v4[c].w, dp);
    }
    const float p = (j < T) ? expf(sc - L) : 0.f;
    const float ds = p * (dp - Dq);
    const int jmax = min(32, T - j0);
#pragma unroll 8
    for (int jj = 0; jj < 32; ++jj) {
      const float pj = __shfl_sync(0xffffffffu, p, jj), dsj = __shfl_sync(0xffffffffu, ds, jj);
      if (jj < jmax) {
        const size_t r = (size_t)(j0 + jj) * ld;
        dq = fmaf(dsj, __ldg(base + r + d + lane), dq);
        if (j0 + jj > 0) dbase[r + lane] = 0.f;
        dbase[r + d + lane] = dsj * qv;
        dbase[r + 2 * d + lane] = pj * g;
      }
    }
  }
  return dq;
}

__global__ void __launch_bounds__(128) attn_bwd_nq1_kernel(const float* __restrict__ qkv, const SegGrp seg,
                                                           const float* __restrict__ out, const float* __restrict__ lse,
                                                           const float* __restrict__ dout, float* __restrict__ dqkv, int d, int heads,
                                                           int B, int skip_above) {
  pdl_grid_sync();
  const int wid = blockIdx.x * 4 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (wid >= B * heads) return;
  const int s = wid / heads, h = wid % heads;
  const Seg sg = seg_of(seg, s);
  if (sg.T > skip_above) return;
  const int ld = 3 * d;
  const float scale = rsqrtf(32.f);
  const float* base = qkv + (size_t)sg.t0 * ld + h * 32;
  float* dbase = dqkv + (size_t)sg.t0 * ld + h * 32;
  const float qv = base[lane] * scale;
  const float g = dout[(size_t)sg.os * d + h * 32 + lane];
  const float Dq = warp_sum(g * out[(size_t)sg.os * d + h * 32 + lane]);
  const float L = lse[(size_t)sg.os * heads + h];
  const float dq = nq1_bwd_chunks(base, dbase, ld, d, sg.T, 0, 32, qv, g, Dq, L, lane);
  dbase[lane] = dq * scale;
}

// ... and for the long bags: a CTA of 32 warps per (slide, head), warps over interleaved chunks, dq merged in a fixed order
__global__ void __launch_bounds__(NQ1_WIDE_WARPS * 32) attn_bwd_nq1_wide_kernel(const float* __restrict__ qkv, const SegGrp seg,
                                                                                const float* __restrict__ out, const float* __restrict__ lse,
                                                                                const float* __restrict__ dout, float* __restrict__ dqkv,
                                                                                int d, int heads) {
  pdl_grid_sync();
  const int s = blockIdx.x / heads, h = blockIdx.x % heads;
  const Seg sg = seg_of(seg, s);
  if (sg.T <= NQ1_WIDE_T) return;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int ld = 3 * d;
  const float scale = rsqrtf(32.f);
  const float* base = qkv + (size_t)sg.t0 * ld + h * 32;
  float* dbase = dqkv + (size_t)sg.t0 * ld + h * 32;
  const float qv = base[lane] * scale;
  const float g = dout[(size_t)sg.os * d + h * 32 + lane];
  const float Dq = warp_sum(g * out[(size_t)sg.os * d + h * 32 + lane]);
  const float L = lse[(size_t)sg.os * heads + h];
  const float dq = nq1_bwd_chunks(base, dbase, ld, d, sg.T, w * 32, NQ1_WIDE_WARPS * 32, qv, g, Dq, L, lane);
  __shared__ float sd[NQ1_WIDE_WARPS][32];
  sd[w][lane] = dq;
  __syncthreads();                                        // (also orders every warp's zero-fill of row 0's q part before the write below)
  if (w == 0) {
    float t = 0.f;
    for (int k = 0; k < NQ1_WIDE_WARPS; ++k) t += sd[k][lane];
    dbase[lane] = t * scale;
  }
}
}  // namespace

namespace {
constexpr int SMALL_MAX_T = 128;
inline size_t small_fwd_smem(int nkc) { return (size_t)(3 * 32 * nkc * SP + SW * SQ * 32 * nkc) * sizeof(float); }
inline size_t small_bwd_smem(int nkc) { return (size_t)(4 * 32 * nkc * SP + 2 * 32 * nkc + SW * 2 * SQ * 32 * nkc) * sizeof(float); }

template <int NKC, bool DROP>
int launch_fwd_small_d(const float* qkv, const SegGrp& seg, float* out, float* lse, int d, int heads, int B, int nq_mode, const Drop& dr,
                       cudaStream_t st) {
  static bool cfg = false;
  const int bytes = (int)small_fwd_smem(NKC);
  if (!cfg) { PATHS_CK(cudaFuncSetAttribute(attn_fwd_small_kernel<NKC, DROP>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes)); cfg = true; }
  pdl(attn_fwd_small_kernel<NKC, DROP>, dim3(heads, B), SW * 32, bytes, st)(qkv, seg, out, lse, d, heads, nq_mode, dr);
  PATHS_LAUNCH_CK();
  return 0;
}
template <int NKC>
int launch_fwd_small(const float* qkv, const SegGrp& seg, float* out, float* lse, int d, int heads, int B, int nq_mode, cudaStream_t st,
                     const Drop& dr) {
  return dr.thresh ? launch_fwd_small_d<NKC, true>(qkv, seg, out, lse, d, heads, B, nq_mode, dr, st)
                   : launch_fwd_small_d<NKC, false>(qkv, seg, out, lse, d, heads, B, nq_mode, dr, st);
}
template <int NKC, bool DROP>
int launch_bwd_small_d(const float* qkv, const SegGrp& seg, const float* out, const float* lse, const float* dout, float* dqkv, int d, int heads,
                       int B, int nq_mode, const Drop& dr, cudaStream_t st) {
  static bool cfg = false;
  const int bytes = (int)small_bwd_smem(NKC);
  if (!cfg) { PATHS_CK(cudaFuncSetAttribute(attn_bwd_small_kernel<NKC, DROP>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes)); cfg = true; }
  pdl(attn_bwd_small_kernel<NKC, DROP>, dim3(heads, B), SW * 32, bytes, st)(qkv, seg, out, lse, dout, dqkv, d, heads, nq_mode, dr);
  PATHS_LAUNCH_CK();
  return 0;
}
template <int NKC>
int launch_bwd_small(const float* qkv, const SegGrp& seg, const float* out, const float* lse, const float* dout, float* dqkv, int d, int heads,
                     int B, int nq_mode, cudaStream_t st, const Drop& dr) {
  return dr.thresh ? launch_bwd_small_d<NKC, true>(qkv, seg, out, lse, dout, dqkv, d, heads, B, nq_mode, dr, st)
                   : launch_bwd_small_d<NKC, false>(qkv, seg, out, lse, dout, dqkv, d, heads, B, nq_mode, dr, st);
}
}  // namespace

int attn_fwd(const float* qkv, const int* seg, float* out, float* lse, int d, int heads, int B, int max_T, int nq_mode, cudaStream_t st) {
  return attn_fwd_grp(qkv, seg_single(seg, B), out, lse, d, heads, max_T, nq_mode, st);
}
int attn_fwd_grp(const float* qkv, const SegGrp& seg, float* out, float* lse, int d, int heads, int max_T, int nq_mode, cudaStream_t st,
                 const Drop* drop, int tc_passes, int* cls) {
  PATHS_REQUIRE(d % heads == 0 && d / heads <= MAXDH, "attention: head_dim %d unsupported (max %d)", d / heads, MAXDH);
  const int B = seg.n * seg.B;                       // virtual slides
  if (B <= 0) return 0;
  int skip_T = 0;
  const Drop dr = drop ? *drop : no_drop();
  PATHS_REQUIRE(!dr.thresh || (max_T < (1 << 24) && heads < 256), "attention dropout: bag of %d tokens / %d heads outside the hash's index range",
                max_T, heads);
  if (d / heads == 32 && nq_mode && !dr.thresh) {    // (with dropout the short-bag / tiled kernels serve the single-query layer too)
    pdl(attn_fwd_nq1_kernel, (B * heads + 3) / 4, 128, 0, st)(qkv, seg, out, lse, d, heads, B, max_T > NQ1_WIDE_T ? NQ1_WIDE_T : 0x7fffffff);
    PATHS_LAUNCH_CK();
    if (max_T > NQ1_WIDE_T) {
      pdl(attn_fwd_nq1_wide_kernel, B * heads, NQ1_WIDE_WARPS * 32, 0, st)(qkv, seg, out, lse, d, heads);
      PATHS_LAUNCH_CK();
    }
    return 0;
  }
  static const bool no_tc = getenv("PATHS_B200_ATTN_SIMT") != nullptr;  // keep the SIMT short-bag kernels in the tensor-core modes
  static const bool use_tq = getenv("PATHS_B200_ATTN_TQ") != nullptr;   // experimental pair-per-row kernels (not faster: off)
  if (tc_passes && !no_tc && !nq_mode && attn_tc_supported(d, heads)) {  // bags of at most 128 tokens on the tensor cores (attention_tc.cu)
    return attn_tc_fwd(tc_passes, qkv, seg, out, lse, d, heads, max_T, st, dr, cls);      // every bag length (short classes + tiled)
  } else if (use_tq && d / heads == 32 && d % 4 == 0 && !nq_mode && !dr.thresh) {
    const int nkc = max_T <= 32 ? 1 : max_T <= 64 ? 2 : max_T <= 96 ? 3 : 4;
    int r = 0;
    switch (nkc) {
      case 1: r = launch_fwd_tq<32>(qkv, seg, out, lse, d, heads, B, st); break;
      case 2: r = launch_fwd_tq<64>(qkv, seg, out, lse, d, heads, B, st); break;
      case 3: r = launch_fwd_tq<96>(qkv, seg, out, lse, d, heads, B, st); break;
      default: r = launch_fwd_tq<128>(qkv, seg, out, lse, d, heads, B, st); break;
    }
    if (r) return r;
    if (max_T <= SMALL_MAX_T) return 0;
    skip_T = SMALL_MAX_T;
  } else if (d / heads == 32) {                // whole-head-in-shared-memory kernels
    const int nkc = max_T <= 32 ? 1 : max_T <= 64 ? 2 : max_T <= 96 ? 3 : 4;
    int r = 0;
    switch (nkc) {
      case 1: r = launch_fwd_small<1>(qkv, seg, out, lse, d, heads, B, nq_mode, st, dr); break;
      case 2: r = launch_fwd_small<2>(qkv, seg, out, lse, d, heads, B, nq_mode, st, dr); break;
      case 3: r = launch_fwd_small<3>(qkv, seg, out, lse, d, heads, B, nq_mode, st, dr); break;
      default: r = launch_fwd_small<4>(qkv, seg, out, lse, d, heads, B, nq_mode, st, dr); break;
    }
    if (r) return r;
    if (max_T <= SMALL_MAX_T) return 0;
    skip_T = SMALL_MAX_T;
  }
  dim3 grid(nq_mode ? 1 : (max_T + QB - 1) / QB, heads, B);
  pdl(attn_fwd_kernel, grid, AW * 32, 0, st)(qkv, seg, out, lse, d, heads, nq_mode, skip_T, dr);
  PATHS_LAUNCH_CK();
  return 0;
}

// dqkv must be zero-initialised by the caller when nq_mode == 1 and attn_bwd_writes_all(d, heads) is false (the dq rows of
// non-query tokens are then not written).
bool attn_bwd_writes_all(int d, int heads) { return d / heads == 32; }
int attn_bwd(const float* qkv, const int* seg, const float* out, const float* lse, const float* dout, float* dqkv, float* Dvec, int d,
             int heads, int B, int max_T, int nq_mode, cudaStream_t st) {
  return attn_bwd_grp(qkv, seg_single(seg, B), out, lse, dout, dqkv, Dvec, d, heads, max_T, nq_mode, st);
}
int attn_bwd_grp(const float* qkv, const SegGrp& seg, const float* out, const float* lse, const float* dout, float* dqkv, float* Dvec, int d,
                 int heads, int max_T, int nq_mode, cudaStream_t st, const Drop* drop, int tc_passes, int* cls) {
  const int B = seg.n * seg.B;
  if (B <= 0) return 0;
  int skip_T = 0;
  const Drop dr = drop ? *drop : no_drop();
  if (d / heads == 32 && nq_mode && !dr.thresh) {    // (with dropout and nq_mode the caller zero-fills dqkv: only query rows get a dq)
    pdl(attn_bwd_nq1_kernel, (B * heads + 3) / 4, 128, 0, st)(qkv, seg, out, lse, dout, dqkv, d, heads, B, max_T > NQ1_WIDE_T ? NQ1_WIDE_T : 0x7fffffff);
    PATHS_LAUNCH_CK();
    if (max_T > NQ1_WIDE_T) {
      pdl(attn_bwd_nq1_wide_kernel, B * heads, NQ1_WIDE_WARPS * 32, 0, st)(qkv, seg, out, lse, dout, dqkv, d, heads);
      PATHS_LAUNCH_CK();
    }
    return 0;
  }
  static const bool no_tc = getenv("PATHS_B200_ATTN_SIMT") != nullptr;
  static const bool use_tq = getenv("PATHS_B200_ATTN_TQ") != nullptr;
  if (tc_passes && !no_tc && !nq_mode && attn_tc_supported(d, heads)) {
    return attn_tc_bwd(tc_passes, qkv, seg, out, lse, dout, dqkv, Dvec, d, heads, max_T, st, dr, cls);
  } else if (use_tq && d / heads == 32 && d % 4 == 0 && !nq_mode && !dr.thresh) {
    const int nkc = max_T <= 32 ? 1 : max_T <= 64 ? 2 : max_T <= 96 ? 3 : 4;
    int r = 0;
    switch (nkc) {
      case 1: r = launch_bwd_tq<32>(qkv, seg, out, lse, dout, dqkv, d, heads, B, st); break;
      case 2: r = launch_bwd_tq<64>(qkv, seg, out, lse, dout, dqkv, d, heads, B, st); break;
      case 3: r = launch_bwd_tq<96>(qkv, seg, out, lse, dout, dqkv, d, heads, B, st); break;
      default: r = launch_bwd_tq<128>(qkv, seg, out, lse, dout, dqkv, d, heads, B, st); break;
    }
    if (r) return r;
    if (max_T <= SMALL_MAX_T) return 0;
    skip_T = SMALL_MAX_T;
  } else if (d / heads == 32) {
    const int nkc = max_T <= 32 ? 1 : max_T <= 64 ? 2 : max_T <= 96 ? 3 : 4;
    int r = 0;
    switch (nkc) {
      case 1: r = launch_bwd_small<1>(qkv, seg, out, lse, dout, dqkv, d, heads, B, nq_mode, st, dr); break;
      case 2: r = launch_bwd_small<2>(qkv, seg, out, lse, dout, dqkv, d, heads, B, nq_mode, st, dr); break;
      case 3: r = launch_bwd_small<3>(qkv, seg, out, lse, dout, dqkv, d, heads, B, nq_mode, st, dr); break;
      default: r = launch_bwd_small<4>(qkv, seg, out, lse, dout, dqkv, d, heads, B, nq_mode, st, dr); break;
    }
    if (r) return r;
    if (max_T <= SMALL_MAX_T) return 0;
    skip_T = SMALL_MAX_T;
  }
  dim3 gq(nq_mode ? 1 : (max_T + QB - 1) / QB, heads, B);
  pdl(attn_bwd_dq_kernel, gq, AW * 32, 0, st)(qkv, seg, out, lse, dout, dqkv, Dvec, d, heads, nq_mode, skip_T, dr);
  PATHS_LAUNCH_CK();
  dim3 gk((max_T + QB - 1) / QB, heads, B);
  pdl(attn_bwd_dkv_kernel, gk, AW * 32, 0, st)(qkv, seg, lse, dout, Dvec, dqkv, d, heads, nq_mode, skip_T, dr);
  PATHS_LAUNCH_CK();
  return 0;
}

}  // namespace paths
