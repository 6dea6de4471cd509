// Short-bag varlen self-attention on the 5th-generation tensor cores (tcgen05 + TMEM), sm_100a.
//
// The self-attention of nn.TransformerDecoderLayer as model/aggregator.py:70-72 uses it (head_dim 32, key-padding mask replaced by
// per-slide segments), for bags of at most 128 tokens -- every level of the BRCA-shaped workload.  One CTA per (head, slide):
//
//   forward    S = (scale Q) K^T          tcgen05.mma, M = 128 query rows, N = padded keys, K = 32
//              P = exp(S - rowmax)        threads read their accumulator row from TMEM (one TMEM lane per query), write P as bf16 tiles
//              O = P V / rowsum           tcgen05.mma, N = 32, K = padded keys
//   backward   S, dP = dO V^T             two products into neighbouring TMEM columns
//              P = exp(S - lse), dS = P o (dP - D)      (D = dO . O per query, computed while the rows are staged)
//              dQ = scale dS K,  dK = dS^T (scale Q),  dV = P^T dO      three products; dS and P are read as K-major (dQ) and,
//              from the same shared-memory tiles, as M-major operands (the transposed products)
//
// Precision: every operand is split x = hi + lo (bf16 each); NP = 2 issues hi.hi + hi.lo + lo.hi into the same fp32 TMEM
// accumulator (~2^-17 per product: the fp32-parity mode), NP = 1 issues hi.hi only (plain bf16: the fast mode).
//
// Shared-memory operand tiles (all bf16, 128-byte rows, 8-row / 1024-byte SWIZZLE_128B atoms, written by the CTA's own threads --
// the rows come from fp32 activations, so there is no TMA on this path):
//   natural tile  [token rows][hi 32 | lo 32]    K-major over head_dim: Q, K (S), dO, V (dP).  The lo half of a row is the
//                                                second 64 bytes of the same swizzled row: a descriptor start + 64 B selects it.
//   transposed    [32 head_dim rows][tokens]     K-major over tokens, one 4 KB block per 64 tokens and plane: V (O), K (dQ),
//                                                Q (dK), dO (dV) as the N = 32 operand of a product that contracts over tokens.
//   score tile    [query rows][keys]             one block per 64 keys and plane: P, dS.  K-major operand of O / dQ; read as an
//                                                M-major operand (M = keys, K = queries) by dK / dV.
// Bags above 128 tokens are tiled (the *_long kernels further down).  Two size classes of short bags share the code: NKB = 1 (bags up to 64 tokens, 64 threads, 64 TMEM columns forward / 128 backward, several CTAs
// per SM) and NKB = 2 (65..128 tokens, 128 threads).  A launch covers every (head, slide); CTAs of the other class exit at once.
#include "common.cuh"
#include "rowops.cuh"
#include "tc_ptx.cuh"

namespace paths {
namespace {
using namespace tc;

constexpr int DH = 32;
// V as the N = 32 operand of O = P V straight from its natural tile ([token rows][hi 32 | lo 32], read M-major: reduction index =
// the rows, output index along the row, the lo half at descriptor start + 64 B) instead of a transposed copy written with 2-byte
// stores -- the staging of the short-bag kernels is L1 / shared-memory bound (profiles/r2_attn_tc_full_672.ncu_summary.txt).
constexpr bool V_NATURAL = true;

__device__ __forceinline__ uint32_t pack_bf16(float a, float b) {
  const __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<const uint32_t*>(&h);
}
// (hi, lo) split of 8 consecutive values into two 16-byte chunks
__device__ __forceinline__ void split8(const float* x, uint4& hi, uint4& lo) {
  uint32_t h[4], l[4];
#pragma unroll
  for (int e = 0; e < 4; ++e) {
    const __nv_bfloat162 hh = __floats2bfloat162_rn(x[2 * e], x[2 * e + 1]);
    const float2 hf = __bfloat1622float2(hh);
    h[e] = *reinterpret_cast<const uint32_t*>(&hh);
    l[e] = pack_bf16(x[2 * e] - hf.x, x[2 * e + 1] - hf.y);
  }
  hi = make_uint4(h[0], h[1], h[2], h[3]); lo = make_uint4(l[0], l[1], l[2], l[3]);
}
// natural tile: row r = [hi(x[0..32)) | lo(x[0..32))], 16-byte chunk c of the row stored at chunk c ^ (r & 7)
__device__ __forceinline__ void store_natural(uint8_t* tile, int r, const float* x) {
  uint8_t* row = tile + (r >> 3) * 1024 + (r & 7) * 128;
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    uint4 hi, lo;
    split8(x + 8 * c, hi, lo);
    *reinterpret_cast<uint4*>(row + ((c ^ (r & 7)) << 4)) = hi;
    *reinterpret_cast<uint4*>(row + (((c + 4) ^ (r & 7)) << 4)) = lo;
  }
}
__device__ __forceinline__ void zero_natural(uint8_t* tile, int r) {
  uint8_t* row = tile + (r >> 3) * 1024 + (r & 7) * 128;
#pragma unroll
  for (int c = 0; c < 8; ++c) *reinterpret_cast<uint4*>(row + (c << 4)) = make_uint4(0, 0, 0, 0);
}
// cooperative staging of an NT-row natural tile by NT threads from rows src + r * ld (32 floats each): item i = (row i / 4, 8-float
// group i % 4), so a warp reads 8 whole 128-byte rows per pass (four passes) instead of 32 rows 16 bytes at a time
template <int NT>
__device__ __forceinline__ void stage_natural(uint8_t* tile, const float* __restrict__ src, int ld, int T, int t, float mul) {
  float4 a[4], b[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int i = t + k * NT, r = i >> 2, g = i & 3;
    if (r < T) {
      const float4* p = reinterpret_cast<const float4*>(src + (size_t)r * ld + g * 8);
      a[k] = __ldg(p); b[k] = __ldg(p + 1);
    } else {
      a[k] = make_float4(0.f, 0.f, 0.f, 0.f); b[k] = a[k];
    }
  }
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int i = t + k * NT, r = i >> 2, g = i & 3;
    const float x[8] = {a[k].x * mul, a[k].y * mul, a[k].z * mul, a[k].w * mul, b[k].x * mul, b[k].y * mul, b[k].z * mul, b[k].w * mul};
    uint4 hi, lo;
    split8(x, hi, lo);
    uint8_t* row = tile + (r >> 3) * 1024 + (r & 7) * 128;
    *reinterpret_cast<uint4*>(row + ((g ^ (r & 7)) << 4)) = hi;
    *reinterpret_cast<uint4*>(row + (((g + 4) ^ (r & 7)) << 4)) = lo;
  }
}
// the same for dO, with rowdot[r] = sum_c dO[r][c] * O[r][c] left in shared memory (the four lanes of a row are adjacent)
template <int NT>
__device__ __forceinline__ void stage_natural_dot(uint8_t* tile, const float* __restrict__ src, const float* __restrict__ oth, int ld,
                                                  int T, int t, float* rowdot) {
  float4 a[4], b[4], oa[4], ob[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int i = t + k * NT, r = i >> 2, g = i & 3;
    if (r < T) {
      const float4* p = reinterpret_cast<const float4*>(src + (size_t)r * ld + g * 8);
      const float4* q = reinterpret_cast<const float4*>(oth + (size_t)r * ld + g * 8);
      a[k] = __ldg(p); b[k] = __ldg(p + 1); oa[k] = __ldg(q); ob[k] = __ldg(q + 1);
    } else {
      a[k] = make_float4(0.f, 0.f, 0.f, 0.f); b[k] = a[k]; oa[k] = a[k]; ob[k] = a[k];
    }
  }
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int i = t + k * NT, r = i >> 2, g = i & 3;
    const float x[8] = {a[k].x, a[k].y, a[k].z, a[k].w, b[k].x, b[k].y, b[k].z, b[k].w};
    float dot = a[k].x * oa[k].x;
    dot = fmaf(a[k].y, oa[k].y, dot); dot = fmaf(a[k].z, oa[k].z, dot); dot = fmaf(a[k].w, oa[k].w, dot);
    dot = fmaf(b[k].x, ob[k].x, dot); dot = fmaf(b[k].y, ob[k].y, dot); dot = fmaf(b[k].z, ob[k].z, dot); dot = fmaf(b[k].w, ob[k].w, dot);
    dot += __shfl_xor_sync(0xffffffffu, dot, 1);
    dot += __shfl_xor_sync(0xffffffffu, dot, 2);
    if (g == 0) rowdot[r] = dot;
    uint4 hi, lo;
    split8(x, hi, lo);
    uint8_t* row = tile + (r >> 3) * 1024 + (r & 7) * 128;
    *reinterpret_cast<uint4*>(row + ((g ^ (r & 7)) << 4)) = hi;
    *reinterpret_cast<uint4*>(row + (((g + 4) ^ (r & 7)) << 4)) = lo;
  }
}
// Block staging of the tiled (long-bag) kernels: a block of 64 token rows x 32 floats is read by the CTA's 128 threads, two items
// (row, 8-float group) each -- a warp covers 8 whole 128-byte rows per pass -- into registers (Blk), and written to its natural tile
// later: the loads of block j + 1 are issued before block j's products are waited for, so the global round trip is off the
// per-block dependency chain.  rows_live <= 0 gives a block of zeros.
struct Blk { float4 a[2], b[2]; };
__device__ __forceinline__ void blk_load(Blk& g, const float* __restrict__ src, int ld, int rows_live, int t) {
#pragma unroll
  for (int k = 0; k < 2; ++k) {
    const int i = t + 128 * k, r = i >> 2, grp = i & 3;
    if (r < rows_live) {
      const float4* p = reinterpret_cast<const float4*>(src + (size_t)r * ld + grp * 8);
      g.a[k] = __ldg(p); g.b[k] = __ldg(p + 1);
    } else {
      g.a[k] = make_float4(0.f, 0.f, 0.f, 0.f); g.b[k] = g.a[k];
    }
  }
}
__device__ __forceinline__ void blk_store(const Blk& g, uint8_t* tile, int t, float mul) {
#pragma unroll
  for (int k = 0; k < 2; ++k) {
    const int i = t + 128 * k, r = i >> 2, grp = i & 3;
    const float x[8] = {g.a[k].x * mul, g.a[k].y * mul, g.a[k].z * mul, g.a[k].w * mul, g.b[k].x * mul, g.b[k].y * mul, g.b[k].z * mul, g.b[k].w * mul};
    uint4 hi, lo;
    split8(x, hi, lo);
    uint8_t* row = tile + (r >> 3) * 1024 + (r & 7) * 128;
    *reinterpret_cast<uint4*>(row + ((grp ^ (r & 7)) << 4)) = hi;
    *reinterpret_cast<uint4*>(row + (((grp + 4) ^ (r & 7)) << 4)) = lo;
  }
}
// transposed planes: element (head_dim c, token j) of plane p at  p * plane_bytes + (j / 64) * 4096 + swizzled (row c, column j % 64)
template <int NP>
__device__ __forceinline__ void store_transposed(uint8_t* base, int plane_bytes, int j, const float* x) {
  uint8_t* blk = base + (j >> 6) * 4096 + ((j & 7) << 1);
  const int jc = (j >> 3) & 7;
#pragma unroll
  for (int c = 0; c < DH; ++c) {
    const __nv_bfloat16 h = __float2bfloat16_rn(x[c]);
    const uint32_t off = (c >> 3) * 1024 + (c & 7) * 128 + ((jc ^ (c & 7)) << 4);
    *reinterpret_cast<__nv_bfloat16*>(blk + off) = h;
    if (NP == 2) *reinterpret_cast<__nv_bfloat16*>(blk + plane_bytes + off) = __float2bfloat16_rn(x[c] - __bfloat162float(h));
  }
}
// score tile: 8 consecutive keys (chunk kc of key block kb) of query row q; block = ROWS x 128 bytes
template <int NP, int ROWS>
__device__ __forceinline__ void store_score8(uint8_t* base, int plane_bytes, int q, int kb, int kc, const float* x) {
  uint8_t* p = base + kb * (ROWS * 128) + (q >> 3) * 1024 + (q & 7) * 128 + ((kc ^ (q & 7)) << 4);
  uint4 hi, lo;
  split8(x, hi, lo);
  *reinterpret_cast<uint4*>(p) = hi;
  if (NP == 2) *reinterpret_cast<uint4*>(p + plane_bytes) = lo;
}

// K-major operand over 16 reduction elements starting at byte `koff` of the 128-byte rows (8-row groups 1024 bytes apart)
__device__ __forceinline__ uint64_t desc_k(uint32_t tile, uint32_t koff) { return umma_desc(tile + koff, 16, 1024); }
// M-major operand: 16 reduction rows starting at row 16 * ks, output index along the 128-byte rows, 64-wide blocks `lbo` bytes apart
__device__ __forceinline__ uint64_t desc_m(uint32_t tile, int ks, uint32_t lbo) { return umma_desc(tile + ks * 2048, lbo, 1024); }

// hi.hi (+ hi.lo + lo.hi) of one 16-deep reduction step; *acc tracks whether the accumulator already holds a term
template <int NP>
__device__ __forceinline__ void mma_terms(uint32_t d_tmem, uint64_t a_hi, uint64_t a_lo, uint64_t b_hi, uint64_t b_lo, uint32_t idesc,
                                          uint32_t* acc) {
  if (NP == 2) {
    umma_bf16(d_tmem, a_lo, b_hi, idesc, *acc);
    umma_bf16(d_tmem, a_hi, b_lo, idesc, 1u);
    umma_bf16(d_tmem, a_hi, b_hi, idesc, 1u);
  } else {
    umma_bf16(d_tmem, a_hi, b_hi, idesc, *acc);
  }
  *acc = 1u;
}

__device__ __forceinline__ void load_row32(const float* p, float* x) {
#pragma unroll
  for (int c = 0; c < 8; ++c) {
    const float4 v = __ldg(reinterpret_cast<const float4*>(p) + c);
    x[4 * c] = v.x; x[4 * c + 1] = v.y; x[4 * c + 2] = v.z; x[4 * c + 3] = v.w;
  }
}
__device__ __forceinline__ void store_row32(float* p, const uint32_t* r, float mul) {
#pragma unroll
  for (int c = 0; c < 8; ++c)
    reinterpret_cast<float4*>(p)[c] = make_float4(__uint_as_float(r[4 * c]) * mul, __uint_as_float(r[4 * c + 1]) * mul,
                                                  __uint_as_float(r[4 * c + 2]) * mul, __uint_as_float(r[4 * c + 3]) * mul);
}

template <int NKB, int NP>
struct FwdCfg {
  static constexpr int NT = 64 * NKB;                       // threads = token rows of the class
  static constexpr int NAT = NT * 128;                      // natural tile bytes
  static constexpr int VT_PLANE = NKB * 4096;               // transposed V, one plane (V_NATURAL == false)
  static constexpr int P_PLANE = NKB * NT * 128;            // score tile, one plane
  static constexpr int OFF_Q = 0, OFF_K = NAT, OFF_VT = 2 * NAT;
  // 64-row class: P (hi, lo: 2 x 8 KB) takes the place of Q | K once S is done, and the 64 rows an M = 128 operand reads past a
  // block land in the next plane / the V tile -- 25 KB per CTA, eight CTAs per SM.  128-row class: P has its own 64 KB.
  static constexpr int V_BYTES = V_NATURAL ? NAT : NP * VT_PLANE;
  static constexpr int OFF_P = NKB == 1 ? 0 : OFF_VT + V_BYTES;
  static constexpr int SMEM = (NKB == 1 ? OFF_VT + V_BYTES : OFF_P + NP * P_PLANE) + 1024;
  static constexpr int TMEM_COLS = 64 * NKB;                // S; O reuses its columns
};

// ------------------------------------------------------------------------------------------------------------------ forward
template <int NKB, int NP, bool DROP>
__global__ void __launch_bounds__(64 * NKB) attn_tc_fwd_kernel(const float* __restrict__ qkv, const SegGrp seg, float* __restrict__ out,
                                                               float* __restrict__ lse, int d, int heads, const Drop dr,
                                                               const int* __restrict__ list) {
  pdl_grid_sync();
  using cfg = FwdCfg<NKB, NP>;
  constexpr int NT = cfg::NT;
  const int h = blockIdx.x;
  // list (attn_tc_classify_kernel): the bags of this class, served by a grid of resident CTAs; without one, CTA y = bag y
  int n_items = gridDim.y;
  if (list) {
    n_items = list[0];
    if ((int)blockIdx.y >= n_items) return;
  } else {
    const int T0 = seg_of(seg, blockIdx.y).T;
    if (T0 > NT || (NKB == 2 && T0 <= 64)) return;          // the other class's launch (or the tiled kernels) serves this bag
  }
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);   // offset from the __shared__ array: stays STS / LDS
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_slot;
  const int t = threadIdx.x, warp = t >> 5;
  if (t == 0) { mbar_init(&bar, 1); fence_barrier_init(); }
  if (warp == 0) tmem_alloc(&tmem_slot, cfg::TMEM_COLS);

  const float scale = rsqrtf((float)DH);
  const int ld = 3 * d;
#pragma unroll 1
  for (int it = blockIdx.y; it < n_items; it += gridDim.y) {
  const Seg sg = seg_of(seg, list ? list[2 + it] : it);
  const int T = sg.T;
  {
    const float* rp = qkv + (size_t)sg.t0 * ld + h * DH;
    stage_natural<NT>(smem + cfg::OFF_Q, rp, ld, T, t, scale);
    stage_natural<NT>(smem + cfg::OFF_K, rp + d, ld, T, t, 1.f);
    stage_natural<NT>(smem + cfg::OFF_VT, rp + 2 * d, ld, T, t, 1.f);   // rows past the bag: zeros (0 * garbage must stay 0)
  }
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_slot;
  const uint32_t sQ = smem_u32(smem + cfg::OFF_Q), sK = smem_u32(smem + cfg::OFF_K), sVt = smem_u32(smem + cfg::OFF_VT),
                 sP = smem_u32(smem + cfg::OFF_P);
  if (t == 0) {                                              // S = (scale Q) K^T : M = 128, N = NT keys, K = 32
    constexpr uint32_t idesc = umma_idesc(128, NT, 0, 0);
    uint32_t acc = 0;
#pragma unroll
    for (int ks = 0; ks < 2; ++ks)
      mma_terms<NP>(tmem_base, desc_k(sQ, ks * 32), desc_k(sQ, 64 + ks * 32), desc_k(sK, ks * 32), desc_k(sK, 64 + ks * 32), idesc, &acc);
    umma_commit(&bar);
  }
  mbar_wait(&bar, 0);
  tc_fence_after();

  const uint32_t lane_addr = tmem_base + ((uint32_t)(warp * 32) << 16);
  const uint64_t dkey = DROP ? pick(dr.key, sg.g) : 0;
  const uint32_t hrow = (uint32_t)(sg.t0 - sg.r0 + t), hcol = (uint32_t)h << 24;
  float mx = -INFINITY;
#pragma unroll 1
  for (int c0 = 0; c0 < NT; c0 += 32) {
    uint32_t r[32];
    tmem_ld32(lane_addr + c0, r);
#pragma unroll
    for (int i = 0; i < 32; ++i)
      if (c0 + i < T) mx = fmaxf(mx, __uint_as_float(r[i]));
  }
  float sum = 0.f;
#pragma unroll 1
  for (int c0 = 0; c0 < NT; c0 += 32) {
    uint32_t r[32];
    tmem_ld32(lane_addr + c0, r);
#pragma unroll
    for (int q = 0; q < 4; ++q) {                            // 8 keys at a time; groups past the bag (uniform per CTA) store zeros
      float p[8];
      if (c0 + 8 * q >= T) {
#pragma unroll
        for (int i = 0; i < 8; ++i) p[i] = 0.f;
      } else {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const int j = c0 + 8 * q + i;
          const float e = (j < T) ? __expf(__uint_as_float(r[8 * q + i]) - mx) : 0.f;
          sum += e;                                          // the normaliser sums every probability, the value sum the kept ones
          p[i] = e;
          if (DROP) p[i] = drop_keep(dkey, hrow, hcol | (uint32_t)j, dr.thresh) ? e * dr.inv_keep : 0.f;
        }
      }
      store_score8<NP, NT>(smem + cfg::OFF_P, cfg::P_PLANE, t, c0 >> 6, ((c0 & 63) >> 3) + q, p);
    }
  }
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (t == 0) {                                              // O = P V : M = 128, N = 32, K = NT keys (S's columns are free now)
    constexpr uint32_t idesc = umma_idesc(128, DH, 0, 0), idesc_bm = umma_idesc(128, DH, 0, 1);
    uint32_t acc = 0;
#pragma unroll
    for (int kb = 0; kb < NKB; ++kb)
#pragma unroll
      for (int ks = 0; ks < 4; ++ks) {
        const uint32_t a = sP + kb * (NT * 128);
        if (V_NATURAL) {
          const uint32_t b = sVt + (kb * 64 + ks * 16) * 128;              // 16 token rows of the natural tile
          mma_terms<NP>(tmem_base, desc_k(a, ks * 32), desc_k(a + cfg::P_PLANE, ks * 32), desc_m(b, 0, 8192), desc_m(b + 64, 0, 8192), idesc_bm, &acc);
        } else {
          const uint32_t b = sVt + kb * 4096;
          mma_terms<NP>(tmem_base, desc_k(a, ks * 32), desc_k(a + cfg::P_PLANE, ks * 32), desc_k(b, ks * 32), desc_k(b + cfg::VT_PLANE, ks * 32),
                        idesc, &acc);
        }
      }
    umma_commit(&bar);
  }
  mbar_wait(&bar, 1);
  tc_fence_after();
  {
    uint32_t r[32];
    tmem_ld32(lane_addr, r);
    if (t < T) {
      store_row32(out + (size_t)(sg.t0 + t) * d + h * DH, r, 1.f / sum);
      lse[(size_t)(sg.t0 + t) * heads + h] = mx + logf(sum);
    }
  }
  tc_fence_before();
  __syncthreads();
  }                                                          // next bag of the list: same barrier phases (two commits per bag)
  if (warp == 0) tmem_dealloc(tmem_slot, cfg::TMEM_COLS);
}

// ----------------------------------------------------------------------------------------------------------------- backward
template <int NKB, int NP>
struct BwdCfg {
  static constexpr int NT = 64 * NKB;
  static constexpr int NAT = NT * 128;
  static constexpr int S_PLANE = NKB * NT * 128;            // a score tile (P or dS), one plane
  // natural Q, K, V, dO: K-major operands of S and dP, and (K, Q, dO) the M-major N = 32 operands of dQ, dK, dV -- no transposed
  // copies; then the score tiles P, dS (+ the 64 rows an M = 128 operand of the 64-row class reads past the last block)
  static constexpr int OFF_Q = 0, OFF_K = NAT, OFF_V = 2 * NAT, OFF_DO = 3 * NAT;
  static constexpr int OFF_P = 4 * NAT, OFF_DS = OFF_P + NP * S_PLANE;
  static constexpr int SMEM = OFF_DS + NP * S_PLANE + (NKB == 1 ? 8192 : 0) + 1024;
  static constexpr int TMEM_COLS = 128 * NKB;               // S | dP; dQ, dK, dV reuse the columns
};

template <int NKB, int NP, bool DROP>
__global__ void __launch_bounds__(64 * NKB) attn_tc_bwd_kernel(const float* __restrict__ qkv, const SegGrp seg, const float* __restrict__ out,
                                                               const float* __restrict__ lse, const float* __restrict__ dout,
                                                               float* __restrict__ dqkv, int d, int heads, const Drop dr,
                                                               const int* __restrict__ list) {
  pdl_grid_sync();
  using cfg = BwdCfg<NKB, NP>;
  constexpr int NT = cfg::NT;
  const int h = blockIdx.x;
  int n_items = gridDim.y;
  if (list) {
    n_items = list[0];
    if ((int)blockIdx.y >= n_items) return;
  } else {
    const int T0 = seg_of(seg, blockIdx.y).T;
    if (T0 > NT || (NKB == 2 && T0 <= 64)) return;
  }
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);   // offset from the __shared__ array: stays STS / LDS
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_slot;
  const int t = threadIdx.x, warp = t >> 5;
  if (t == 0) { mbar_init(&bar, 1); fence_barrier_init(); }
  if (warp == 0) tmem_alloc(&tmem_slot, cfg::TMEM_COLS);

  const float scale = rsqrtf((float)DH);
  const int ld = 3 * d;
  __shared__ float s_rowdot[NT];
#pragma unroll 1
  for (int it = blockIdx.y; it < n_items; it += gridDim.y) {
  const Seg sg = seg_of(seg, list ? list[2 + it] : it);
  const int T = sg.T;
  {
    const float* rp = qkv + (size_t)sg.t0 * ld + h * DH;
    const size_t o0 = (size_t)sg.t0 * d + h * DH;
    stage_natural<NT>(smem + cfg::OFF_Q, rp, ld, T, t, scale);
    stage_natural<NT>(smem + cfg::OFF_K, rp + d, ld, T, t, 1.f);
    stage_natural<NT>(smem + cfg::OFF_V, rp + 2 * d, ld, T, t, 1.f);
    stage_natural_dot<NT>(smem + cfg::OFF_DO, dout + o0, out + o0, d, T, t, s_rowdot);
  }
  const float L = t < T ? lse[(size_t)(sg.t0 + t) * heads + h] : 0.f;
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const float Dq = s_rowdot[t];
  const uint32_t tmem_base = tmem_slot;
  const uint32_t s0 = smem_u32(smem);
  if (t == 0) {                                              // S = (scale Q) K^T -> columns [0, NT);  dP = dO V^T -> [NT, 2 NT)
    constexpr uint32_t idesc = umma_idesc(128, NT, 0, 0);
    uint32_t acc = 0;
#pragma unroll
    for (int ks = 0; ks < 2; ++ks)
      mma_terms<NP>(tmem_base, desc_k(s0 + cfg::OFF_Q, ks * 32), desc_k(s0 + cfg::OFF_Q, 64 + ks * 32), desc_k(s0 + cfg::OFF_K, ks * 32),
                    desc_k(s0 + cfg::OFF_K, 64 + ks * 32), idesc, &acc);
    acc = 0;
#pragma unroll
    for (int ks = 0; ks < 2; ++ks)
      mma_terms<NP>(tmem_base + NT, desc_k(s0 + cfg::OFF_DO, ks * 32), desc_k(s0 + cfg::OFF_DO, 64 + ks * 32), desc_k(s0 + cfg::OFF_V, ks * 32),
                    desc_k(s0 + cfg::OFF_V, 64 + ks * 32), idesc, &acc);
    umma_commit(&bar);
  }
  mbar_wait(&bar, 0);                                        // both products done: the natural tiles may be overwritten
  tc_fence_after();

  const uint32_t lane_addr = tmem_base + ((uint32_t)(warp * 32) << 16);
  const uint64_t dkey = DROP ? pick(dr.key, sg.g) : 0;
  const uint32_t hrow = (uint32_t)(sg.t0 - sg.r0 + t), hcol = (uint32_t)h << 24;
#pragma unroll 1
  for (int c0 = 0; c0 < NT; c0 += 32) {
    uint32_t rs[32], rp[32];
    tmem_ld32(lane_addr + c0, rs);
    tmem_ld32(lane_addr + NT + c0, rp);
#pragma unroll
    for (int q = 0; q < 4; ++q) {                            // 8 keys at a time; groups past the bag (uniform per CTA) store zeros
      float pv[8], ds[8];
      if (c0 + 8 * q >= T) {
#pragma unroll
        for (int i = 0; i < 8; ++i) pv[i] = ds[i] = 0.f;
      } else {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const int j = c0 + 8 * q + i;
          const float p = (t < T && j < T) ? __expf(__uint_as_float(rs[8 * q + i]) - L) : 0.f;   // rows / columns past the bag: exact zeros
          float pm = p, dpm = __uint_as_float(rp[8 * q + i]);  // (they are contracted over by dK / dV, resp. dQ)
          if (DROP) {
            const bool keep = drop_keep(dkey, hrow, hcol | (uint32_t)j, dr.thresh);
            pm = keep ? p * dr.inv_keep : 0.f; dpm = keep ? dpm * dr.inv_keep : 0.f;
          }
          pv[i] = pm;
          ds[i] = p * (dpm - Dq);
        }
      }
      store_score8<NP, NT>(smem + cfg::OFF_P, cfg::S_PLANE, t, c0 >> 6, ((c0 & 63) >> 3) + q, pv);
      store_score8<NP, NT>(smem + cfg::OFF_DS, cfg::S_PLANE, t, c0 >> 6, ((c0 & 63) >> 3) + q, ds);
    }
  }
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (t == 0) {
    constexpr uint32_t idk = umma_idesc(128, DH, 0, 1);      // A K-major, B M-major
    constexpr uint32_t idm = umma_idesc(128, DH, 1, 1);      // A M-major (the transposed products), B M-major
    const uint32_t sP = s0 + cfg::OFF_P, sDS = s0 + cfg::OFF_DS, sK = s0 + cfg::OFF_K, sQ = s0 + cfg::OFF_Q, sDO = s0 + cfg::OFF_DO;
    uint32_t acc = 0;                                        // dQ[q, c] = sum_j dS[q, j] K[j, c]        -> columns [0, 32)
#pragma unroll
    for (int kb = 0; kb < NKB; ++kb)
#pragma unroll
      for (int ks = 0; ks < 4; ++ks) {
        const uint32_t a = sDS + kb * (NT * 128), b = sK + (kb * 64 + ks * 16) * 128;      // 16 key rows of the natural tile
        mma_terms<NP>(tmem_base, desc_k(a, ks * 32), desc_k(a + cfg::S_PLANE, ks * 32), desc_m(b, 0, 8192), desc_m(b + 64, 0, 8192), idk, &acc);
      }
    acc = 0;                                                 // dK[j, c] = sum_q dS[q, j] (scale Q)[q, c]   -> columns [32, 64)
#pragma unroll
    for (int ks = 0; ks < NT / 16; ++ks) {
      const uint32_t b = sQ + ks * 16 * 128;
      mma_terms<NP>(tmem_base + 32, desc_m(sDS, ks, NT * 128), desc_m(sDS + cfg::S_PLANE, ks, NT * 128), desc_m(b, 0, 8192), desc_m(b + 64, 0, 8192),
                    idm, &acc);
    }
    acc = 0;                                                 // dV[j, c] = sum_q P[q, j] dO[q, c]           -> columns [64, 96)
#pragma unroll
    for (int ks = 0; ks < NT / 16; ++ks) {
      const uint32_t b = sDO + ks * 16 * 128;
      mma_terms<NP>(tmem_base + 64, desc_m(sP, ks, NT * 128), desc_m(sP + cfg::S_PLANE, ks, NT * 128), desc_m(b, 0, 8192), desc_m(b + 64, 0, 8192),
                    idm, &acc);
    }
    umma_commit(&bar);
  }
  mbar_wait(&bar, 1);
  tc_fence_after();
  {
    float* rp = dqkv + (size_t)(sg.t0 + t) * ld + h * DH;
    uint32_t r[32];
    tmem_ld32(lane_addr, r);
    if (t < T) store_row32(rp, r, scale);
    tmem_ld32(lane_addr + 32, r);
    if (t < T) store_row32(rp + d, r, 1.f);
    tmem_ld32(lane_addr + 64, r);
    if (t < T) store_row32(rp + 2 * d, r, 1.f);
  }
  tc_fence_before();
  __syncthreads();
  }
  if (warp == 0) tmem_dealloc(tmem_slot, cfg::TMEM_COLS);
}

// =====================================================================================================================
// Long bags (more than 128 tokens: the 20 k-token bag of cfg 4, the 257-token bags K = 64 produces): the same products, tiled.
// A CTA owns 128 queries (forward, dQ) or 128 keys (dK / dV) and walks the other dimension in blocks of 64; accumulators that need
// no rescaling (dQ, dK, dV: the log-sum-exp is known) stay in TMEM across blocks, the forward pass keeps its running output row
// in registers (one query per thread) and rescales it there (online softmax).
constexpr int LB = 64;                                      // block of the walked dimension

template <int NP>
struct LongFwdCfg {
  static constexpr int OFF_Q = 0, OFF_K = 16384, OFF_V = OFF_K + LB * 128, OFF_P = OFF_V + LB * 128;    // K, V: natural 64-row tiles
  static constexpr int P_PLANE = 128 * 128;
  static constexpr int SMEM = OFF_P + NP * P_PLANE + 1024;
  static constexpr int TMEM_COLS = 128;                      // S [0, 64), P V [64, 96)
};

template <int NP, bool DROP>
__global__ void __launch_bounds__(128) attn_tc_fwd_long_kernel(const float* __restrict__ qkv, const SegGrp seg, float* __restrict__ out,
                                                               float* __restrict__ lse, int d, int heads, const Drop dr) {
  pdl_grid_sync();
  using cfg = LongFwdCfg<NP>;
  const int h = blockIdx.x;
  const Seg sg = seg_of(seg, blockIdx.z);
  const int T = sg.T, q0 = blockIdx.y * 128;
  if (T <= 128 || q0 >= T) return;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);   // offset from the __shared__ array: stays STS / LDS
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_slot;
  const int t = threadIdx.x, warp = t >> 5;
  if (t == 0) { mbar_init(&bar, 1); fence_barrier_init(); }
  if (warp == 0) tmem_alloc(&tmem_slot, cfg::TMEM_COLS);
  const float scale = rsqrtf((float)DH);
  const int ld = 3 * d;
  const bool live = q0 + t < T;
  {
    float x[DH];
    if (live) {
      load_row32(qkv + (size_t)(sg.t0 + q0 + t) * ld + h * DH, x);
#pragma unroll
      for (int c = 0; c < DH; ++c) x[c] *= scale;
      store_natural(smem + cfg::OFF_Q, t, x);
    } else {
      zero_natural(smem + cfg::OFF_Q, t);
    }
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_slot;
  const uint32_t lane_addr = tmem_base + ((uint32_t)(warp * 32) << 16);
  const uint32_t sQ = smem_u32(smem + cfg::OFF_Q), sK = smem_u32(smem + cfg::OFF_K), sV = smem_u32(smem + cfg::OFF_V),
                 sP = smem_u32(smem + cfg::OFF_P);
  const uint64_t dkey = DROP ? pick(dr.key, sg.g) : 0;
  const uint32_t hrow = (uint32_t)(sg.t0 - sg.r0 + q0 + t), hcol = (uint32_t)h << 24;
  float o[DH];
#pragma unroll
  for (int c = 0; c < DH; ++c) o[c] = 0.f;
  float mx = -INFINITY, sum = 0.f;
  uint32_t phase = 0;
  const float* kbase = qkv + (size_t)sg.t0 * ld + h * DH + d;  // K rows of the bag; V rows are d floats further
  Blk gK, gV;
  blk_load(gK, kbase, ld, T, t);
  blk_load(gV, kbase + d, ld, T, t);
#pragma unroll 1
  for (int k0 = 0; k0 < T; k0 += LB) {
    blk_store(gK, smem + cfg::OFF_K, t, 1.f);                // the block's keys and values (natural tiles: V is read M-major)
    blk_store(gV, smem + cfg::OFF_V, t, 1.f);
    if (k0 + LB < T) {                                       // next block's rows: in flight while this block's products run
      blk_load(gK, kbase + (size_t)(k0 + LB) * ld, ld, T - k0 - LB, t);
      blk_load(gV, kbase + d + (size_t)(k0 + LB) * ld, ld, T - k0 - LB, t);
    }
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    if (t == 0) {                                            // S = (scale Q) K_blk^T : N = 64
      constexpr uint32_t idesc = umma_idesc(128, LB, 0, 0);
      uint32_t acc = 0;
#pragma unroll
      for (int ks = 0; ks < 2; ++ks)
        mma_terms<NP>(tmem_base, desc_k(sQ, ks * 32), desc_k(sQ, 64 + ks * 32), desc_k(sK, ks * 32), desc_k(sK, 64 + ks * 32), idesc, &acc);
      umma_commit(&bar);
    }
    mbar_wait(&bar, phase); phase ^= 1;
    tc_fence_after();
    float bmx = mx;
#pragma unroll 1
    for (int c0 = 0; c0 < LB; c0 += 32) {
      uint32_t r[32];
      tmem_ld32(lane_addr + c0, r);
#pragma unroll
      for (int i = 0; i < 32; ++i)
        if (k0 + c0 + i < T) bmx = fmaxf(bmx, __uint_as_float(r[i]));
    }
    const float corr = expf(mx - bmx);                       // exp(-inf) = 0 on the first block
    mx = bmx;
    sum *= corr;
#pragma unroll 1
    for (int c0 = 0; c0 < LB; c0 += 32) {
      uint32_t r[32];
      tmem_ld32(lane_addr + c0, r);
      float p[32];
#pragma unroll
      for (int i = 0; i < 32; ++i) {
        const float e = (k0 + c0 + i < T) ? __expf(__uint_as_float(r[i]) - mx) : 0.f;
        sum += e;
        p[i] = e;
        if (DROP) p[i] = drop_keep(dkey, hrow, hcol | (uint32_t)(k0 + c0 + i), dr.thresh) ? e * dr.inv_keep : 0.f;
      }
#pragma unroll
      for (int q = 0; q < 4; ++q) store_score8<NP, 128>(smem + cfg::OFF_P, cfg::P_PLANE, t, 0, (c0 >> 3) + q, p + 8 * q);
    }
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    if (t == 0) {                                            // block's P V -> columns [64, 96), fresh
      constexpr uint32_t idesc = umma_idesc(128, DH, 0, 1);  // B = 16 value rows of the natural tile, M-major
      uint32_t acc = 0;
#pragma unroll
      for (int ks = 0; ks < 4; ++ks) {
        const uint32_t b = sV + ks * 16 * 128;
        mma_terms<NP>(tmem_base + 64, desc_k(sP, ks * 32), desc_k(sP + cfg::P_PLANE, ks * 32), desc_m(b, 0, 8192), desc_m(b + 64, 0, 8192), idesc,
                      &acc);
      }
      umma_commit(&bar);
    }
    mbar_wait(&bar, phase); phase ^= 1;
    tc_fence_after();
    {
      uint32_t r[32];
      tmem_ld32(lane_addr + 64, r);
#pragma unroll
      for (int c = 0; c < DH; ++c) o[c] = fmaf(o[c], corr, __uint_as_float(r[c]));
    }
    tc_fence_before();                                       // the next block's MMAs overwrite S / P V and the staged tiles
    __syncthreads();
    tc_fence_after();
  }
  if (live) {
    const float inv = 1.f / sum;
    float* op = out + (size_t)(sg.t0 + q0 + t) * d + h * DH;
#pragma unroll
    for (int c = 0; c < 8; ++c) reinterpret_cast<float4*>(op)[c] = make_float4(o[4 * c] * inv, o[4 * c + 1] * inv, o[4 * c + 2] * inv, o[4 * c + 3] * inv);
    lse[(size_t)(sg.t0 + q0 + t) * heads + h] = mx + logf(sum);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem_base, cfg::TMEM_COLS);
}

template <int NP>
struct LongDqCfg {
  static constexpr int OFF_Q = 0, OFF_DO = 16384, OFF_K = 32768, OFF_V = OFF_K + LB * 128, OFF_DS = OFF_V + LB * 128;
  static constexpr int S_PLANE = 128 * 128;
  static constexpr int SMEM = OFF_DS + NP * S_PLANE + 1024;
  static constexpr int TMEM_COLS = 256;                      // S [0, 64), dP [64, 128), dQ [128, 160)
};

// dQ of 128 queries: walks the keys in blocks of 64, dQ accumulates in TMEM.  Also writes D = dO . O of its queries (dvec) for the
// dK / dV kernel.
template <int NP, bool DROP>
__global__ void __launch_bounds__(128) attn_tc_bwd_dq_long_kernel(const float* __restrict__ qkv, const SegGrp seg, const float* __restrict__ out,
                                                                  const float* __restrict__ lse, const float* __restrict__ dout,
                                                                  float* __restrict__ dqkv, float* __restrict__ dvec, int d, int heads,
                                                                  const Drop dr) {
  pdl_grid_sync();
  using cfg = LongDqCfg<NP>;
  const int h = blockIdx.x;
  const Seg sg = seg_of(seg, blockIdx.z);
  const int T = sg.T, q0 = blockIdx.y * 128;
  if (T <= 128 || q0 >= T) return;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);   // offset from the __shared__ array: stays STS / LDS
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_slot;
  const int t = threadIdx.x, warp = t >> 5;
  if (t == 0) { mbar_init(&bar, 1); fence_barrier_init(); }
  if (warp == 0) tmem_alloc(&tmem_slot, cfg::TMEM_COLS);
  const float scale = rsqrtf((float)DH);
  const int ld = 3 * d;
  const bool live = q0 + t < T;
  float Dq = 0.f, L = 0.f;
  {
    float x[DH], y[DH];
    if (live) {
      load_row32(qkv + (size_t)(sg.t0 + q0 + t) * ld + h * DH, x);
#pragma unroll
      for (int c = 0; c < DH; ++c) x[c] *= scale;
      store_natural(smem + cfg::OFF_Q, t, x);
      const size_t orow = (size_t)(sg.t0 + q0 + t) * d + h * DH;
      load_row32(dout + orow, x);
      load_row32(out + orow, y);
#pragma unroll
      for (int c = 0; c < DH; ++c) Dq = fmaf(x[c], y[c], Dq);
      store_natural(smem + cfg::OFF_DO, t, x);
      L = lse[(size_t)(sg.t0 + q0 + t) * heads + h];
      dvec[(size_t)(sg.t0 + q0 + t) * heads + h] = Dq;
    } else {
      zero_natural(smem + cfg::OFF_Q, t);
      zero_natural(smem + cfg::OFF_DO, t);
    }
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_slot;
  const uint32_t lane_addr = tmem_base + ((uint32_t)(warp * 32) << 16);
  const uint32_t s0 = smem_u32(smem);
  const uint64_t dkey = DROP ? pick(dr.key, sg.g) : 0;
  const uint32_t hrow = (uint32_t)(sg.t0 - sg.r0 + q0 + t), hcol = (uint32_t)h << 24;
  uint32_t phase = 0, acc_dq = 0;
  const float* kbase = qkv + (size_t)sg.t0 * ld + h * DH + d;
  Blk gK, gV;
  blk_load(gK, kbase, ld, T, t);
  blk_load(gV, kbase + d, ld, T, t);
#pragma unroll 1
  for (int k0 = 0; k0 < T; k0 += LB) {
    blk_store(gK, smem + cfg::OFF_K, t, 1.f);                // natural tiles: K is the K-major operand of S and the M-major one of dQ
    blk_store(gV, smem + cfg::OFF_V, t, 1.f);
    if (k0 + LB < T) {
      blk_load(gK, kbase + (size_t)(k0 + LB) * ld, ld, T - k0 - LB, t);
      blk_load(gV, kbase + d + (size_t)(k0 + LB) * ld, ld, T - k0 - LB, t);
    }
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    if (t == 0) {
      constexpr uint32_t idesc = umma_idesc(128, LB, 0, 0);
      uint32_t acc = 0;
#pragma unroll
      for (int ks = 0; ks < 2; ++ks)
        mma_terms<NP>(tmem_base, desc_k(s0 + cfg::OFF_Q, ks * 32), desc_k(s0 + cfg::OFF_Q, 64 + ks * 32), desc_k(s0 + cfg::OFF_K, ks * 32),
                      desc_k(s0 + cfg::OFF_K, 64 + ks * 32), idesc, &acc);
      acc = 0;
#pragma unroll
      for (int ks = 0; ks < 2; ++ks)
        mma_terms<NP>(tmem_base + 64, desc_k(s0 + cfg::OFF_DO, ks * 32), desc_k(s0 + cfg::OFF_DO, 64 + ks * 32), desc_k(s0 + cfg::OFF_V, ks * 32),
                      desc_k(s0 + cfg::OFF_V, 64 + ks * 32), idesc, &acc);
      umma_commit(&bar);
    }
    mbar_wait(&bar, phase); phase ^= 1;
    tc_fence_after();
#pragma unroll 1
    for (int c0 = 0; c0 < LB; c0 += 32) {
      uint32_t rs[32], rp[32];
      tmem_ld32(lane_addr + c0, rs);
      tmem_ld32(lane_addr + 64 + c0, rp);
      float ds[32];
#pragma unroll
      for (int i = 0; i < 32; ++i) {
        const float p = (live && k0 + c0 + i < T) ? __expf(__uint_as_float(rs[i]) - L) : 0.f;
        float dpm = __uint_as_float(rp[i]);
        if (DROP) dpm = drop_keep(dkey, hrow, hcol | (uint32_t)(k0 + c0 + i), dr.thresh) ? dpm * dr.inv_keep : 0.f;
        ds[i] = p * (dpm - Dq);
      }
#pragma unroll
      for (int q = 0; q < 4; ++q) store_score8<NP, 128>(smem + cfg::OFF_DS, cfg::S_PLANE, t, 0, (c0 >> 3) + q, ds + 8 * q);
    }
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    if (t == 0) {                                            // dQ += dS_blk K_blk : columns [128, 160), accumulated over the blocks
      constexpr uint32_t idesc = umma_idesc(128, DH, 0, 1);
      const uint32_t a = s0 + cfg::OFF_DS;
#pragma unroll
      for (int ks = 0; ks < 4; ++ks) {
        const uint32_t b = s0 + cfg::OFF_K + ks * 16 * 128;  // 16 key rows of the natural tile
        mma_terms<NP>(tmem_base + 128, desc_k(a, ks * 32), desc_k(a + cfg::S_PLANE, ks * 32), desc_m(b, 0, 8192), desc_m(b + 64, 0, 8192), idesc,
                      &acc_dq);
      }
      umma_commit(&bar);
    }
    mbar_wait(&bar, phase); phase ^= 1;                      // the staged tiles are free again
    tc_fence_after();
  }
  {
    uint32_t r[32];
    tmem_ld32(lane_addr + 128, r);
    if (live) store_row32(dqkv + (size_t)(sg.t0 + q0 + t) * ld + h * DH, r, scale);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem_base, cfg::TMEM_COLS);
}

template <int NP>
struct LongDkvCfg {
  static constexpr int OFF_K = 0, OFF_V = 16384, OFF_Q = 32768, OFF_DO = OFF_Q + LB * 128, OFF_P = OFF_DO + LB * 128;
  static constexpr int S_PLANE = 128 * 128;                  // [128 key rows][64 queries]
  static constexpr int OFF_DS = OFF_P + NP * S_PLANE;
  static constexpr int SMEM = OFF_DS + NP * S_PLANE + 1024;
  static constexpr int TMEM_COLS = 256;                      // S^T [0, 64), dP^T [64, 128), dK [128, 160), dV [160, 192)
};

// dK, dV of 128 keys: walks the queries in blocks of 64.  The products are formed transposed -- S^T = K (scale Q)^T, dP^T = V dO^T,
// one TMEM lane per KEY -- so the score tiles P^T, dS^T are written row-wise by the key's thread and are K-major operands of
// dV += P^T dO and dK += dS^T (scale Q); the per-query log-sum-exp and D = dO . O come from shared memory.
template <int NP, bool DROP>
__global__ void __launch_bounds__(128) attn_tc_bwd_dkv_long_kernel(const float* __restrict__ qkv, const SegGrp seg, const float* __restrict__ lse,
                                                                   const float* __restrict__ dout, const float* __restrict__ dvec,
                                                                   float* __restrict__ dqkv, int d, int heads, const Drop dr) {
  pdl_grid_sync();
  using cfg = LongDkvCfg<NP>;
  const int h = blockIdx.x;
  const Seg sg = seg_of(seg, blockIdx.z);
  const int T = sg.T, j0 = blockIdx.y * 128;
  if (T <= 128 || j0 >= T) return;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);   // offset from the __shared__ array: stays STS / LDS
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_slot;
  __shared__ float sL[LB], sD[LB];
  const int t = threadIdx.x, warp = t >> 5;
  if (t == 0) { mbar_init(&bar, 1); fence_barrier_init(); }
  if (warp == 0) tmem_alloc(&tmem_slot, cfg::TMEM_COLS);
  const float scale = rsqrtf((float)DH);
  const int ld = 3 * d;
  const bool live = j0 + t < T;                              // this thread's key
  {
    float x[DH];
    if (live) {
      const float* rp = qkv + (size_t)(sg.t0 + j0 + t) * ld + h * DH;
      load_row32(rp + d, x);
      store_natural(smem + cfg::OFF_K, t, x);
      load_row32(rp + 2 * d, x);
      store_natural(smem + cfg::OFF_V, t, x);
    } else {
      zero_natural(smem + cfg::OFF_K, t);
      zero_natural(smem + cfg::OFF_V, t);
    }
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_slot;
  const uint32_t lane_addr = tmem_base + ((uint32_t)(warp * 32) << 16);
  const uint32_t s0 = smem_u32(smem);
  const uint64_t dkey = DROP ? pick(dr.key, sg.g) : 0;
  const uint32_t hrow0 = (uint32_t)(sg.t0 - sg.r0), hcol = ((uint32_t)h << 24) | (uint32_t)(j0 + t);
  uint32_t phase = 0, acc_dk = 0, acc_dv = 0;
  const float* qbase = qkv + (size_t)sg.t0 * ld + h * DH;
  const float* obase = dout + (size_t)sg.t0 * d + h * DH;
  const float* lbase = lse + (size_t)sg.t0 * heads + h;
  const float* dbase = dvec + (size_t)sg.t0 * heads + h;
  Blk gQ, gO;
  float gL = 0.f, gD = 0.f;                                  // (threads t < 64: the block's per-query log-sum-exp and D)
  blk_load(gQ, qbase, ld, T, t);
  blk_load(gO, obase, d, T, t);
  if (t < LB && t < T) { gL = __ldg(lbase + (size_t)t * heads); gD = __ldg(dbase + (size_t)t * heads); }
#pragma unroll 1
  for (int q0 = 0; q0 < T; q0 += LB) {
    // the block's queries: scale Q and dO as natural tiles (K-major operands of S^T / dP^T, M-major ones of dK / dV); lse, D
    blk_store(gQ, smem + cfg::OFF_Q, t, scale);
    blk_store(gO, smem + cfg::OFF_DO, t, 1.f);
    if (t < LB) { sL[t] = gL; sD[t] = gD; }
    if (q0 + LB < T) {
      const int nq = q0 + LB;
      blk_load(gQ, qbase + (size_t)nq * ld, ld, T - nq, t);
      blk_load(gO, obase + (size_t)nq * d, d, T - nq, t);
      gL = 0.f; gD = 0.f;
      if (t < LB && nq + t < T) { gL = __ldg(lbase + (size_t)(nq + t) * heads); gD = __ldg(dbase + (size_t)(nq + t) * heads); }
    }
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    if (t == 0) {                                            // S^T = K (scale Q_blk)^T, dP^T = V dO_blk^T : M = 128 keys, N = 64 queries
      constexpr uint32_t idesc = umma_idesc(128, LB, 0, 0);
      uint32_t acc = 0;
#pragma unroll
      for (int ks = 0; ks < 2; ++ks)
        mma_terms<NP>(tmem_base, desc_k(s0 + cfg::OFF_K, ks * 32), desc_k(s0 + cfg::OFF_K, 64 + ks * 32), desc_k(s0 + cfg::OFF_Q, ks * 32),
                      desc_k(s0 + cfg::OFF_Q, 64 + ks * 32), idesc, &acc);
      acc = 0;
#pragma unroll
      for (int ks = 0; ks < 2; ++ks)
        mma_terms<NP>(tmem_base + 64, desc_k(s0 + cfg::OFF_V, ks * 32), desc_k(s0 + cfg::OFF_V, 64 + ks * 32), desc_k(s0 + cfg::OFF_DO, ks * 32),
                      desc_k(s0 + cfg::OFF_DO, 64 + ks * 32), idesc, &acc);
      umma_commit(&bar);
    }
    mbar_wait(&bar, phase); phase ^= 1;
    tc_fence_after();
#pragma unroll 1
    for (int c0 = 0; c0 < LB; c0 += 32) {
      uint32_t rs[32], rp[32];
      tmem_ld32(lane_addr + c0, rs);
      tmem_ld32(lane_addr + 64 + c0, rp);
      float pv[32], ds[32];
#pragma unroll
      for (int i = 0; i < 32; ++i) {
        const int q = q0 + c0 + i;
        const float p = (live && q < T) ? __expf(__uint_as_float(rs[i]) - sL[c0 + i]) : 0.f;
        float pm = p, dpm = __uint_as_float(rp[i]);
        if (DROP) {
          const bool keep = drop_keep(dkey, hrow0 + (uint32_t)q, hcol, dr.thresh);
          pm = keep ? p * dr.inv_keep : 0.f; dpm = keep ? dpm * dr.inv_keep : 0.f;
        }
        pv[i] = pm;
        ds[i] = p * (dpm - sD[c0 + i]);
      }
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        store_score8<NP, 128>(smem + cfg::OFF_P, cfg::S_PLANE, t, 0, (c0 >> 3) + q, pv + 8 * q);
        store_score8<NP, 128>(smem + cfg::OFF_DS, cfg::S_PLANE, t, 0, (c0 >> 3) + q, ds + 8 * q);
      }
    }
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    if (t == 0) {                                            // dK += dS^T (scale Q_blk), dV += P^T dO_blk : contract over the 64 queries
      constexpr uint32_t idesc = umma_idesc(128, DH, 0, 1);
      const uint32_t aS = s0 + cfg::OFF_DS, aP = s0 + cfg::OFF_P;
#pragma unroll
      for (int ks = 0; ks < 4; ++ks) {
        const uint32_t b = s0 + cfg::OFF_Q + ks * 16 * 128;  // 16 query rows of the natural tile
        mma_terms<NP>(tmem_base + 128, desc_k(aS, ks * 32), desc_k(aS + cfg::S_PLANE, ks * 32), desc_m(b, 0, 8192), desc_m(b + 64, 0, 8192), idesc,
                      &acc_dk);
      }
#pragma unroll
      for (int ks = 0; ks < 4; ++ks) {
        const uint32_t b = s0 + cfg::OFF_DO + ks * 16 * 128;
        mma_terms<NP>(tmem_base + 160, desc_k(aP, ks * 32), desc_k(aP + cfg::S_PLANE, ks * 32), desc_m(b, 0, 8192), desc_m(b + 64, 0, 8192), idesc,
                      &acc_dv);
      }
      umma_commit(&bar);
    }
    mbar_wait(&bar, phase); phase ^= 1;
    tc_fence_after();
  }
  {
    float* rp = dqkv + (size_t)(sg.t0 + j0 + t) * ld + h * DH;
    uint32_t r[32];
    tmem_ld32(lane_addr + 128, r);
    if (live) store_row32(rp + d, r, 1.f);
    tmem_ld32(lane_addr + 160, r);
    if (live) store_row32(rp + 2 * d, r, 1.f);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem_base, cfg::TMEM_COLS);
}

template <typename K>
int set_smem_once(K kern, int bytes, bool* done) {
  if (*done) return 0;
  PATHS_CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
  *done = true;
  return 0;
}

// CTAs of `kern` the device holds at once (occupancy x SMs), cached in *slots
template <typename K>
int resident_slots(K kern, int threads, int smem, int* slots) {
  if (*slots) return 0;
  int dev = 0, sms = 0, per_sm = 0;
  PATHS_CK(cudaGetDevice(&dev));
  PATHS_CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  PATHS_CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, threads, smem));
  *slots = sms * (per_sm > 0 ? per_sm : 1);
  return 0;
}
// grid rows (CTAs per head): one per bag without a class list; with one, as many as stay resident -- they walk the list
inline int list_rows(const int* list, int bags, int resident, int heads) {
  if (!list) return bags;
  const int rows = resident / heads > 0 ? resident / heads : 1;
  return rows < bags ? rows : bags;
}

// Bags of the 65..128-token class (list[0] of them, list[2 ...]) and of the tiled class (list[1], list[2 + bags ...]) of a grouped
// launch: the bag lengths live on the device, and a grid over every bag spends more time retiring the CTAs of the other classes
// (each holds its shared memory until its segment offsets arrive) than computing.  One block; order is irrelevant to the results.
__global__ void __launch_bounds__(1024) attn_tc_classify_kernel(const SegGrp seg, int* __restrict__ list) {
  pdl_grid_sync();
  __shared__ int n_mid, n_long;
  if (threadIdx.x == 0) { n_mid = 0; n_long = 0; }
  __syncthreads();
  const int bags = seg.n * seg.B;
  for (int sv = threadIdx.x; sv < bags; sv += blockDim.x) {
    const int T = seg_of(seg, sv).T;
    if (T > 128) list[2 + bags + atomicAdd(&n_long, 1)] = sv;
    else if (T > 64) list[2 + atomicAdd(&n_mid, 1)] = sv;
  }
  __syncthreads();
  if (threadIdx.x == 0) { list[0] = n_mid; list[1] = n_long; }
}
int classify(const SegGrp& seg, int* list, cudaStream_t st) {
  pdl(attn_tc_classify_kernel, 1, 1024, 0, st)(seg, list);
  PATHS_LAUNCH_CK();
  return 0;
}

template <int NKB, int NP, bool DROP>
int launch_fwd(const float* qkv, const SegGrp& seg, float* out, float* lse, int d, int heads, cudaStream_t st, const Drop& dr, const int* list) {
  static bool done = false;
  static int resident = 0;
  auto kern = attn_tc_fwd_kernel<NKB, NP, DROP>;
  if (int r = set_smem_once(kern, FwdCfg<NKB, NP>::SMEM, &done)) return r;
  if (int r = resident_slots(kern, 64 * NKB, FwdCfg<NKB, NP>::SMEM, &resident)) return r;
  pdl(kern, dim3(heads, list_rows(list, seg.n * seg.B, resident, heads)), 64 * NKB, FwdCfg<NKB, NP>::SMEM, st)(qkv, seg, out, lse, d, heads, dr,
                                                                                                             list);
  PATHS_LAUNCH_CK();
  return 0;
}
template <int NKB, int NP, bool DROP>
int launch_bwd(const float* qkv, const SegGrp& seg, const float* out, const float* lse, const float* dout, float* dqkv, int d, int heads,
               cudaStream_t st, const Drop& dr, const int* list) {
  static bool done = false;
  static int resident = 0;
  auto kern = attn_tc_bwd_kernel<NKB, NP, DROP>;
  if (int r = set_smem_once(kern, BwdCfg<NKB, NP>::SMEM, &done)) return r;
  if (int r = resident_slots(kern, 64 * NKB, BwdCfg<NKB, NP>::SMEM, &resident)) return r;
  pdl(kern, dim3(heads, list_rows(list, seg.n * seg.B, resident, heads)), 64 * NKB, BwdCfg<NKB, NP>::SMEM, st)(qkv, seg, out, lse, dout, dqkv, d,
                                                                                                             heads, dr, list);
  PATHS_LAUNCH_CK();
  return 0;
}

template <int NP, bool DROP>
int launch_fwd_long(const float* qkv, const SegGrp& seg, float* out, float* lse, int d, int heads, int max_T, cudaStream_t st, const Drop& dr) {
  static bool done = false;
  auto kern = attn_tc_fwd_long_kernel<NP, DROP>;
  if (int r = set_smem_once(kern, LongFwdCfg<NP>::SMEM, &done)) return r;
  pdl(kern, dim3(heads, (max_T + 127) / 128, seg.n * seg.B), 128, LongFwdCfg<NP>::SMEM, st)(qkv, seg, out, lse, d, heads, dr);
  PATHS_LAUNCH_CK();
  return 0;
}
template <int NP, bool DROP>
int launch_bwd_long(const float* qkv, const SegGrp& seg, const float* out, const float* lse, const float* dout, float* dqkv, float* dvec, int d,
                    int heads, int max_T, cudaStream_t st, const Drop& dr) {
  static bool done = false;
  auto kq = attn_tc_bwd_dq_long_kernel<NP, DROP>;
  auto kk = attn_tc_bwd_dkv_long_kernel<NP, DROP>;
  if (!done) {
    PATHS_CK(cudaFuncSetAttribute(kq, cudaFuncAttributeMaxDynamicSharedMemorySize, LongDqCfg<NP>::SMEM));
    PATHS_CK(cudaFuncSetAttribute(kk, cudaFuncAttributeMaxDynamicSharedMemorySize, LongDkvCfg<NP>::SMEM));
    done = true;
  }
  const dim3 grid(heads, (max_T + 127) / 128, seg.n * seg.B);
  pdl(kq, grid, 128, LongDqCfg<NP>::SMEM, st)(qkv, seg, out, lse, dout, dqkv, dvec, d, heads, dr);
  PATHS_LAUNCH_CK();
  pdl(kk, grid, 128, LongDkvCfg<NP>::SMEM, st)(qkv, seg, lse, dout, dvec, dqkv, d, heads, dr);
  PATHS_LAUNCH_CK();
  return 0;
}

}  // namespace

bool attn_tc_supported(int d, int heads) { return heads > 0 && d % heads == 0 && d / heads == DH && d % 4 == 0; }

#define ATTN_TC_DISPATCH(FN, NKB_, ...)                                                \
  do {                                                                                 \
    int r_;                                                                            \
    if (passes == 3) r_ = dr.thresh ? FN<NKB_, 2, true>(__VA_ARGS__) : FN<NKB_, 2, false>(__VA_ARGS__); \
    else r_ = dr.thresh ? FN<NKB_, 1, true>(__VA_ARGS__) : FN<NKB_, 1, false>(__VA_ARGS__);             \
    if (r_) return r_;                                                                 \
  } while (0)

int attn_tc_fwd(int passes, const float* qkv, const SegGrp& seg, float* out, float* lse, int d, int heads, int max_T, cudaStream_t st,
                const Drop& dr, int* cls) {
  PATHS_REQUIRE(attn_tc_supported(d, heads) && (passes == 1 || passes == 3), "attn_tc_fwd: head_dim must be 32");
  if (seg.n * seg.B <= 0) return 0;
  if (cls && max_T > 64)
    if (int r = classify(seg, cls, st)) return r;
  const int* none = nullptr;
  ATTN_TC_DISPATCH(launch_fwd, 1, qkv, seg, out, lse, d, heads, st, dr, none);
  if (max_T > 64) ATTN_TC_DISPATCH(launch_fwd, 2, qkv, seg, out, lse, d, heads, st, dr, cls);
  if (max_T > 128) {                                         // long bags: tiled over keys (attn_tc_fwd_long_kernel)
    int r;
    if (passes == 3) r = dr.thresh ? launch_fwd_long<2, true>(qkv, seg, out, lse, d, heads, max_T, st, dr) : launch_fwd_long<2, false>(qkv, seg, out, lse, d, heads, max_T, st, dr);
    else r = dr.thresh ? launch_fwd_long<1, true>(qkv, seg, out, lse, d, heads, max_T, st, dr) : launch_fwd_long<1, false>(qkv, seg, out, lse, d, heads, max_T, st, dr);
    if (r) return r;
  }
  return 0;
}
int attn_tc_bwd(int passes, const float* qkv, const SegGrp& seg, const float* out, const float* lse, const float* dout, float* dqkv, float* dvec,
                int d, int heads, int max_T, cudaStream_t st, const Drop& dr, int* cls) {
  PATHS_REQUIRE(attn_tc_supported(d, heads) && (passes == 1 || passes == 3), "attn_tc_bwd: head_dim must be 32");
  if (seg.n * seg.B <= 0) return 0;
  if (cls && max_T > 64)
    if (int r = classify(seg, cls, st)) return r;
  const int* none = nullptr;
  ATTN_TC_DISPATCH(launch_bwd, 1, qkv, seg, out, lse, dout, dqkv, d, heads, st, dr, none);
  if (max_T > 64) ATTN_TC_DISPATCH(launch_bwd, 2, qkv, seg, out, lse, dout, dqkv, d, heads, st, dr, cls);
  if (max_T > 128) {                                         // long bags: dQ per query tile, then dK / dV per key tile
    int r;
    if (passes == 3) r = dr.thresh ? launch_bwd_long<2, true>(qkv, seg, out, lse, dout, dqkv, dvec, d, heads, max_T, st, dr)
                                   : launch_bwd_long<2, false>(qkv, seg, out, lse, dout, dqkv, dvec, d, heads, max_T, st, dr);
    else r = dr.thresh ? launch_bwd_long<1, true>(qkv, seg, out, lse, dout, dqkv, dvec, d, heads, max_T, st, dr)
                       : launch_bwd_long<1, false>(qkv, seg, out, lse, dout, dqkv, dvec, d, heads, max_T, st, dr);
    if (r) return r;
  }
  return 0;
}
#undef ATTN_TC_DISPATCH

}  // namespace paths
