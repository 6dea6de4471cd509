// tcgen05 / TMEM / TMA contractions for sm_100a.
//
//   gemm_nt_tc : C[M,N] (+)= act(A . B^T + bias)     A = fp32 activations (RowSrc: gathered / concatenated rows),
//                                                     B = weights, pre-split into bf16 (hi, lo) planes and fed by TMA
//   gemm_tn_tc : C[NI,NJ] += P^T . Q                  both operands fp32 activations (weight gradients)
//
// Precision: operands are split as x = hi + lo with hi = bf16(x), lo = bf16(x - hi).  `passes == 3` issues
// hi*hi + hi*lo + lo*hi into the same fp32 TMEM accumulator (error ~2^-17 per product: the fp32-parity mode);
// `passes == 1` issues hi*hi only (plain bf16: the fast mode).  Activations stay fp32 in HBM (the canonical storage
// the parity oracle is compared against); a coalesced split kernel turns the rows a contraction needs -- gathered and
// concatenated through RowSrc ([x || h_parent], token gathers) -- into (hi, lo) planes right before the contraction.
//
// A coalesced split kernel writes the planes (performing the gather / concat); the contraction kernel is then a pure
// TMA -> tcgen05 pipeline: CTA = 10 warps: 1 TMA warp, 1 MMA warp (one elected thread issues tcgen05.mma), 8 epilogue warps.
// One 128 x BN fp32 accumulator tile in TMEM per CTA; smem ring of `NS` stages, 64-deep (128-byte, SWIZZLE_128B) K-blocks.
#include <cuda.h>
#include <cuda_bf16.h>

#include <mutex>
#include <vector>

#include "common.cuh"
#include "tc_ptx.cuh"

namespace paths {

namespace {

constexpr int BM = 128;
constexpr int BK = 64;                    // bf16 elements per k-block = one 128-byte swizzle row
constexpr int EPI_WARPS = 8, TMA_WARP = 8, MMA_WARP = 9;
constexpr int LOADER_WARP0 = 10, LOADER_WARPS = 8;           // only present when an operand is converted in-kernel
constexpr int NTHREADS_TMA = 10 * 32, NTHREADS_CVT = (10 + LOADER_WARPS) * 32;

using namespace tc;

// Tile layouts in shared memory (bf16, 128-byte rows, 8-row / 1024-byte swizzle atoms):
//  K-major  tile [R rows x 64 k]   : byte(r, k)  = (r/8)*1024 + (r%8)*128 + (((k/8) ^ (r%8)) * 16) + (k%8)*2
//  MN-major tile [64 k x R cols]   : byte(k, mn) = (mn/64)*8192 + (k/8)*1024 + (k%8)*128 + ((((mn%64)/8) ^ (k%8)) * 16) + (mn%8)*2
// Both are what a SWIZZLE_128B TMA box {64 elements, rows} produces, so TMA-written and producer-written tiles agree.
__device__ __forceinline__ uint32_t sw128_off(int row, int elem /* 0..63 */) {
  return (uint32_t)((row >> 3) * 1024 + (row & 7) * 128 + ((((elem >> 3) ^ (row & 7))) << 4) + (elem & 7) * 2);
}

// Gate non-linearities of the epilogue (tensor-core modes only; the fp32 mode's row kernels keep expf / tanhf): ex2.approx and
// rcp.approx, ~2e-7 absolute -- two orders below the split-bf16 products they are applied to -- at a quarter of the instructions
// and without tanhf's range branches.
__device__ __forceinline__ float sigmoid_fast(float x) { return __fdividef(1.f, 1.f + __expf(-x)); }
__device__ __forceinline__ float tanh_fast(float x) { return 1.f - __fdividef(2.f, 1.f + __expf(2.f * x)); }
// the plain-bf16 mode (stated tolerance 1e-2): one MUFU.TANH per element (tanh.approx.f32, ~5e-4 absolute), sigmoid through it
__device__ __forceinline__ float tanh_mufu(float x) { float y; asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ float sigmoid_mufu(float x) { return fmaf(0.5f, tanh_mufu(0.5f * x), 0.5f); }

// CG = 2: CTA pair (cta_group::2) -- the pair's tile is 256 x BN, each CTA stages its 128 rows of A and HALF of the B tile
template <int BN, int NP, int CG = 1>
struct Cfg {
  static constexpr int A_PLANE = BM * BK * 2;
  static constexpr int B_PLANE = (BN / CG) * BK * 2;
  static constexpr int STAGE = NP * (A_PLANE + B_PLANE);
  static constexpr int NS = (200 * 1024) / STAGE >= 4 ? 4 : (200 * 1024) / STAGE;
  static constexpr int EPI_STAGE = 8 * 4096;                 // per-epilogue-warp 32 x 32 fp32 transpose buffers
  static constexpr int SMEM = NS * STAGE + EPI_STAGE + 1024;
  static_assert(SMEM <= 227 * 1024 + 1024 - 2048, "shared memory budget");
  static_assert(NS >= 2, "need at least two stages");
};

// Epilogue shared by both kernels: thread (warp w, lane) owns accumulator row (w%4)*32+lane and the column half w/4.
template <int BN, typename F>
__device__ __forceinline__ void epilogue_rows(uint32_t tmem_base, int warp, int lane, F&& emit) {
  const int q = warp & 3, half = warp >> 2;
  constexpr int COLS = BN / 2;
#pragma unroll 1
  for (int c0 = 0; c0 < COLS; c0 += 32) {
    uint32_t r[32];
    const int col = half * COLS + c0;
    tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)col, r);
    emit(q * 32 + lane, col, r);
  }
}

// =================================================================================================== split kernel
// fp32 rows (gathered / concatenated through RowSrc) -> bf16 planes stacked as [2 * plane_rows, ldp] (hi rows first).
// Rows [M, round_up(M, 64)) are zero-filled so a reduction over rows (TN) may run whole 64-row blocks.
__global__ void __launch_bounds__(256) split_planes_kernel(RowSrc A, int M_host, const int* __restrict__ M_dev, int K, Planes dst, int col0,
                                                           int want_lo) {
  pdl_grid_sync();
  const int M = M_dev ? *M_dev : M_host;
  const int Mpad = planes_rows_to_write(dst, M);
  const int lane = threadIdx.x & 31;
  for (int m = blockIdx.x * 8 + (threadIdx.x >> 5); m < Mpad; m += gridDim.x * 8) {
    for (int k = lane * 4; k < K; k += 128) {
      float4 x = make_float4(0.f, 0.f, 0.f, 0.f);
      if (m < M) {
        const float* p = rowsrc_ptr(A, m, k);
        if (p) x = __ldg(reinterpret_cast<const float4*>(p));
      }
      planes_store4(dst, m, col0 + k, x, want_lo != 0);
    }
  }
}

// Grouped form (common.cuh: Grp): rows are global rows of a concatenated buffer, blockIdx.y is the row group.  Every group's
// region is zero-filled from its live count up to round_up(max live count over the launch's groups, 64): the weight-gradient
// contraction gives every output tile the k-block count of the longest group, so the shorter groups must read zeros there
// (the engine spaces the regions by at least that much).
// With `csum` the pass also accumulates the column sums of the live rows into csum.p[group] (K <= 512): the bias gradient
// that would otherwise need its own pass over the same tensor.
// J = float4 column groups per lane in use (ceil(min(K, 512) / 128)), R = rows per warp iteration: a 128-column row is ONE
// 16-byte load per lane, so several rows are loaded before the first is stored to keep enough bytes in flight per SM.
template <int J, int R>
__global__ void __launch_bounds__(256) split_planes_grp_kernel(const float* __restrict__ A, int lda, int K, Planes dst, Grp grp, int want_lo,
                                                               MutL csum) {
  pdl_grid_sync();
  const int gi = blockIdx.y;
  const int M = grp_count(grp, gi);
  int mx = 0;
#pragma unroll
  for (int g = 0; g < kMaxGroups; ++g) {
    const int c = g < grp.n ? (grp.cnt[g] ? *grp.cnt[g] : grp.hcnt[g]) : 0;
    mx = c > mx ? c : mx;
  }
  const int Mfill = (mx + 63) & ~63;
  const int r0 = pick(grp.roff, gi);
  const int lane = threadIdx.x & 31;
  float* cs = pick(csum.p, gi);
  const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
  float4 acc[J];
#pragma unroll
  for (int j = 0; j < J; ++j) acc[j] = z4;
  const int stride = gridDim.x * 8;
  for (int m0 = blockIdx.x * 8 + (threadIdx.x >> 5); m0 < Mfill; m0 += R * stride) {
    float4 x[R][J];
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const int m = m0 + r * stride;
      const float* row = A + (size_t)(r0 + m) * lda;
#pragma unroll
      for (int j = 0; j < J; ++j) {
        const int k = lane * 4 + 128 * j;
        x[r][j] = z4;
        if (m < M && k < K) x[r][j] = __ldg(reinterpret_cast<const float4*>(row + k));
      }
    }
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const int m = m0 + r * stride;
      if (m >= Mfill) continue;                              // warp-uniform
#pragma unroll
      for (int j = 0; j < J; ++j) {
        const int k = lane * 4 + 128 * j;
        if (k < K) {
          planes_store4(dst, r0 + m, k, x[r][j], want_lo != 0);
          acc[j].x += x[r][j].x; acc[j].y += x[r][j].y; acc[j].z += x[r][j].z; acc[j].w += x[r][j].w;
        }
      }
      if (J == 4) {
        const float* row = A + (size_t)(r0 + m) * lda;
        for (int k = lane * 4 + 512; k < K; k += 128) {      // wider rows (no column sums there)
          float4 xw = z4;
          if (m < M) xw = __ldg(reinterpret_cast<const float4*>(row + k));
          planes_store4(dst, r0 + m, k, xw, want_lo != 0);
        }
      }
    }
  }
  if (cs == nullptr) return;                               // block-uniform
  __shared__ float4 part[8][128];
  const int w = threadIdx.x >> 5;
#pragma unroll
  for (int j = 0; j < 4; ++j) part[w][lane + 32 * j] = j < J ? acc[j < J ? j : 0] : z4;
  __syncthreads();
  if (threadIdx.x < 128) {
    const int j = threadIdx.x >> 5, ln = threadIdx.x & 31;
    const int k = ln * 4 + 128 * j;                          // column of part[*][threadIdx.x]
    if (k < K) {
      float4 t = part[0][threadIdx.x];
#pragma unroll
      for (int ww = 1; ww < 8; ++ww) { const float4 u = part[ww][threadIdx.x]; t.x += u.x; t.y += u.y; t.z += u.z; t.w += u.w; }
      atomicAdd(cs + k, t.x); atomicAdd(cs + k + 1, t.y); atomicAdd(cs + k + 2, t.z); atomicAdd(cs + k + 3, t.w);
    }
  }
}

// =================================================================================================== TMA-fed UMMA kernel
// D[128 x BN] (+)= sum over reduction blocks of A_tile . B_tile^T, both operands arriving by TMA as 64x64 bf16 boxes of
// (hi, lo) planes.  Operand orientation:
//   X_MN == false (K-major):  plane is [rows = output index, cols = reduction index]   (activations in NT, weights [N,K])
//   X_MN == true  (MN-major): plane is [rows = reduction index, cols = output index]   (weights [K,N] in dX, both in TN)
// The reduction length may live on the device (TN: number of packed rows); so may the output row count (NT).
struct Operand { int row0, col0, plane_rows; };

// In-kernel operand conversion: the loader warps read an fp32 operand tile straight from its RowSrc (gather / concat
// included), split it into bf16 (hi, lo) and store it in the SWIZZLE_128B layout a TMA box would have produced -- no separate
// split pass, no bf16 copy in HBM.  K-major tile: [W rows (output index) x 64 k]; MN-major tile: [64 k x W cols (output index)].
// Work item = one 16-byte bf16 chunk (8 elements = two float4 loads); W * 8 items per tile, 4 per thread per batch.
// Stream-K partition shared by the contraction kernel and tn_reduce_kernel.  k1 >= 0 (more CTAs than tiles): CTA t < tiles
// owns k-blocks [0, k1) of tile t; the other G - tiles CTAs split the tails [k1, nkb) of all tiles (tile-major, q k-blocks
// each; tl = nkb - k1, total = tiles * tl).  k1 < 0: plain tile-major stream-K over all G CTAs.
struct SkPlan { int k1, tl; long long q, total; };
// One work unit of the contraction kernel: output tile origin, k-block range, row group, A-row limit, reduction row window.
struct Unit { int i0, j0, kb_begin, kb_end, g; };
__host__ __device__ __forceinline__ SkPlan sk_plan(int tiles, int nkb, int G) {
  SkPlan p;
  if (G > tiles) {
    p.k1 = (int)(((long long)nkb * tiles) / G);
    p.tl = nkb - p.k1;
    p.total = (long long)tiles * p.tl;
    p.q = (p.total + (G - tiles) - 1) / (G - tiles);
    if (p.q < 1) p.q = 1;
  } else {
    p.k1 = -1; p.tl = nkb; p.total = (long long)tiles * nkb; p.q = (p.total + G - 1) / G;
  }
  return p;
}

constexpr int LB = 4;                                       // items per thread per batch
struct CvtRegs { float4 x[LB][2]; uint32_t off[LB]; };

// issue the global loads of batch `bi` (items bi*LB .. bi*LB+LB-1 of this thread) of a W-wide tile
template <int W, bool MN>
__device__ __forceinline__ void cvt_issue(const RowSrc& S, int bi, int o0, int o_lim, int r0, int r_lim, int ltid, CvtRegs& g) {
  constexpr int NLT = LOADER_WARPS * 32;
#pragma unroll
  for (int n = 0; n < LB; ++n) {
    const int item = ltid + (bi * LB + n) * NLT;
    int o, k;
    if (!MN) {
      const int r = item >> 3, ch = item & 7;
      o = o0 + r; k = r0 + ch * 8;
      g.off[n] = (uint32_t)((r >> 3) * 1024 + (r & 7) * 128 + ((ch ^ (r & 7)) << 4));
    } else {
      const int kr = item / (W / 8), ch = item % (W / 8);
      o = o0 + ch * 8; k = r0 + kr;
      g.off[n] = (uint32_t)((ch >> 3) * 8192 + (kr >> 3) * 1024 + (kr & 7) * 128 + (((ch & 7) ^ (kr & 7)) << 4));
    }
    const float* p = (o < o_lim && k < r_lim) ? (MN ? rowsrc_ptr(S, k, o) : rowsrc_ptr(S, o, k)) : nullptr;
    g.x[n][0] = g.x[n][1] = make_float4(0.f, 0.f, 0.f, 0.f);
    if (p) { g.x[n][0] = __ldg(reinterpret_cast<const float4*>(p)); g.x[n][1] = __ldg(reinterpret_cast<const float4*>(p) + 1); }
  }
}
// split the loaded batch into (hi, lo) and store it into the swizzled operand tile
template <int NP>
__device__ __forceinline__ void cvt_store(const CvtRegs& g, uint8_t* dst, int plane_bytes) {
#pragma unroll
  for (int n = 0; n < LB; ++n) {
    const float v[8] = {g.x[n][0].x, g.x[n][0].y, g.x[n][0].z, g.x[n][0].w, g.x[n][1].x, g.x[n][1].y, g.x[n][1].z, g.x[n][1].w};
    uint32_t hi[4], lo[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const __nv_bfloat162 h = __floats2bfloat162_rn(v[2 * e], v[2 * e + 1]);
      const float2 hf = __bfloat1622float2(h);
      const __nv_bfloat162 l = __floats2bfloat162_rn(v[2 * e] - hf.x, v[2 * e + 1] - hf.y);
      hi[e] = *reinterpret_cast<const uint32_t*>(&h); lo[e] = *reinterpret_cast<const uint32_t*>(&l);
    }
    *reinterpret_cast<uint4*>(dst + g.off[n]) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
    if (NP == 2) *reinterpret_cast<uint4*>(dst + plane_bytes + g.off[n]) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
  }
}

// SK (stream-K, used by the weight-gradient contractions whose reduction runs over the packed rows): the linearised
// (output tile, k-block) space is cut into gridDim.x equal contiguous ranges, one per CTA, so every SM gets the same number of
// k-blocks whatever the tile count; each (CTA, tile) segment leaves a partial 128 x BN tile in a workspace slot
// [blockIdx.x * segmax + ordinal] (row-major) and tn_reduce_kernel adds the slots of a tile in
// CTA order into C: deterministic, no atomics.
// A_CVT / B_CVT: the operand comes as fp32 rows (RowSrc sa / sb) and is converted by the loader warps instead of arriving as
// TMA boxes of pre-split planes.
// CG = 2 (NT form with both operands TMA-fed): the CTAs of a 2-cluster work as a pair on a 256 x BN tile with ONE
// tcgen05.mma.cta_group::2 per step, issued by rank 0.  What bounds the single-CTA form is the L2 -> shared-memory fill rate
// (A 128 x 64 + B BN x 64 per 64-deep k-block and SM: 48 B/clk/SM at the 76 % tensor activity the gate contraction reaches, the
// chip's L2 -> SM ceiling); the pair halves the B bytes each SM has to pull.  Protocol: both TMA warps load into their own
// shared memory and count bytes on rank 0's full barrier; rank 0's commit multicasts "stage free" / "accumulator full" to both
// CTAs; the epilogue warps of both CTAs release the accumulator on rank 0's barrier.
template <int BN, int NP, bool A_MN, bool B_MN, bool SK, bool A_CVT, bool B_CVT, int CG = 1>
__global__ void __launch_bounds__((A_CVT || B_CVT) ? NTHREADS_CVT : NTHREADS_TMA, 1)
gemm_tma_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, Operand oa, Operand ob,
                const RowSrc sa, const RowSrc sb, float* __restrict__ C, int ldc, const GemmGrp gg, int b_grp_rows, int OI_host, int OJ,
                int R_host, int flags, int segmax) {
  pdl_grid_sync();
  // Persistent: one CTA per SM walks work units u = blockIdx.x, += gridDim.x; unit = (row tile, column tile, reduction split)
  // with column tiles fastest so CTAs that run together share A rows in L2.  Two TMEM accumulators: the epilogue of unit n
  // drains one while the MMA warp fills the other.
  static_assert(CG == 1 || (CG == 2 && !A_CVT && !B_CVT && BN == 256), "CTA pairs: TMA-fed operands, BN = 256");
  using cfg = Cfg<BN, NP, CG>;
  constexpr int NS = cfg::NS;
  constexpr int TM = BM * CG;                                // rows of a work unit's tile
  constexpr uint32_t TMA_BYTES = (A_CVT ? 0 : NP * cfg::A_PLANE) + (B_CVT ? 0 : NP * cfg::B_PLANE);
  const uint32_t rank = CG == 2 ? cluster_ctarank() : 0u;    // rank 0 leads the pair
  constexpr uint32_t FULL_COUNT = (TMA_BYTES ? 1 : 0) + ((A_CVT || B_CVT) ? LOADER_WARPS : 0);
  // Row groups (common.cuh): the NT form walks the live 128-row tiles of every group of the concatenated activation buffer
  // (weights / bias of the tile's group); the TN form (SK) produces one [OI_host x OJ] output per group, reducing over the
  // group's rows -- every output tile gets the k-block count of the longest group, shorter groups read zeros past their end.
  __shared__ int g_cnt[kMaxGroups], g_cum[kMaxGroups + 1], g_max, g_roff[kMaxGroups], g_wlev[kMaxGroups];
  __shared__ const float* g_bias[kMaxGroups];
  if (threadIdx.x == 0) {
    int cum = 0, mx = 0;
    g_cum[0] = 0;
#pragma unroll
    for (int g = 0; g < kMaxGroups; ++g) {                 // compile-time indices into the parameter struct (see pick())
      const int c = g < gg.rows.n ? (gg.rows.cnt[g] ? *gg.rows.cnt[g] : gg.rows.hcnt[g]) : 0;
      g_cnt[g] = c; g_roff[g] = gg.rows.roff[g]; g_wlev[g] = gg.wlev[g]; g_bias[g] = gg.bias[g];
      cum += (c + TM - 1) / TM;
      g_cum[g + 1] = cum;
      mx = c > mx ? c : mx;
    }
    g_max = mx;
  }

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);   // offset from the __shared__ array: stays STS / LDS
  __shared__ uint64_t full_bar[NS], empty_bar[NS], acc_full[2], acc_empty[2];
  __shared__ uint32_t tmem_slot;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    for (int s = 0; s < NS; ++s) { mbar_init(&full_bar[s], FULL_COUNT); mbar_init(&empty_bar[s], 1); }
    for (int b = 0; b < 2; ++b) { mbar_init(&acc_full[b], 1); mbar_init(&acc_empty[b], EPI_WARPS * CG); }
    fence_barrier_init();
  }
  if (warp == MMA_WARP) { if (CG == 2) tmem_alloc_pair(&tmem_slot, 2 * BN); else tmem_alloc(&tmem_slot, 2 * BN); }
  tc_fence_before();
  __syncthreads();
  if (CG == 2) cluster_sync_all();                           // the peer's barriers exist before anything is signalled across
  tc_fence_after();
  const uint32_t tmem_base = tmem_slot;
  // (stream-K over pairs: the plan below is the single-CTA one with the pair as the "CTA" and the 256-row tile as the tile)
  const int vcta = (int)(blockIdx.x / CG), vG = (int)(gridDim.x / CG);
  const int tiles_i = SK ? (OI_host + TM - 1) / TM : g_cum[kMaxGroups], tiles_j = (OJ + BN - 1) / BN;
  const int nkb_total = ((SK ? g_max : R_host) + BK - 1) / BK;
  const int units = SK ? gg.rows.n * tiles_i * tiles_j : tiles_i * tiles_j;
  const long long sk_total = (long long)units * nkb_total;
  const long long sk_q = (sk_total + vG - 1) / vG;
  const long long sk_lo = (long long)vcta * sk_q;
  const long long sk_hi = sk_lo + sk_q < sk_total ? sk_lo + sk_q : sk_total;

  // n-th work unit of this CTA: (output tile origin, k-block range); false when the CTA has no more work
  auto get_unit = [&](int n, Unit& u) -> bool {
    int t;
    int& kb_begin = u.kb_begin; int& kb_end = u.kb_end;
    if (SK) {
      if (nkb_total == 0) return false;
      const SkPlan pl = sk_plan(units, nkb_total, vG);
      if (pl.k1 >= 0) {
        // more CTAs than tiles: CTA t < tiles sweeps k-blocks [0, k1) of tile t in lock step with the others (all of them read
        // the same reduction rows at the same time: operands stay L2-resident); the remaining CTAs stream over the tails
        if (vcta < units) {
          if (n > 0 || pl.k1 == 0) return false;
          t = vcta; kb_begin = 0; kb_end = pl.k1;
        } else {
          const long long lo = (long long)(vcta - units) * pl.q, hi0 = lo + pl.q;
          const long long hi = hi0 < pl.total ? hi0 : pl.total;
          t = (int)(lo / pl.tl) + n;
          const long long t_lo = (long long)t * pl.tl;
          const long long b = lo > t_lo ? lo : t_lo;
          const long long e = hi < t_lo + pl.tl ? hi : t_lo + pl.tl;
          if (b >= e) return false;
          kb_begin = pl.k1 + (int)(b - t_lo); kb_end = pl.k1 + (int)(e - t_lo);
        }
      } else {
        t = (int)(sk_lo / nkb_total) + n;
        const long long t_lo = (long long)t * nkb_total;
        const long long b = sk_lo > t_lo ? sk_lo : t_lo;
        const long long e = sk_hi < t_lo + nkb_total ? sk_hi : t_lo + nkb_total;
        if (b >= e) return false;
        kb_begin = (int)(b - t_lo); kb_end = (int)(e - t_lo);
      }
    } else {
      t = vcta + n * vG;
      if (t >= units || nkb_total == 0) return false;
      kb_begin = 0; kb_end = nkb_total;
    }
    if (SK) {
      const int tpg = tiles_i * tiles_j, tt = t % tpg;
      u.g = t / tpg;
      u.j0 = (tt % tiles_j) * BN; u.i0 = (tt / tiles_j) * TM + (int)rank * BM;
    } else {
      const int r = t / tiles_j;
      int g = 0;
      for (int k = 1; k < gg.rows.n; ++k) g = r >= g_cum[k] ? k : g;
      u.g = g;
      u.j0 = (t % tiles_j) * BN; u.i0 = g_roff[g] + (r - g_cum[g]) * TM + (int)rank * BM;
    }
    return true;
  };
  // A-row limit of a unit's group / reduction row window (NT: the K extent; TN: the group's rows)
  auto a_olim = [&](const Unit& u) -> int { return SK ? OI_host : g_roff[u.g] + g_cnt[u.g]; };
  auto r_base = [&](const Unit& u) -> int { return SK ? g_roff[u.g] : 0; };
  auto r_lim = [&](const Unit& u) -> int { return SK ? g_roff[u.g] + g_cnt[u.g] : R_host; };

  if (warp >= LOADER_WARP0) {
    if (A_CVT || B_CVT) {
      // Software-pipelined by one batch: the loads of batch b+1 are in flight while batch b is converted and stored, so the
      // loader never sits idle behind a DRAM round trip.  A batch = LB 16-byte chunks per thread of one operand of one k-block.
      constexpr int NLT = LOADER_WARPS * 32;
      constexpr int NA = A_CVT ? (BM * 8 / NLT) / LB : 0, NB = B_CVT ? (BN * 8 / NLT) / LB : 0, NSUB = NA + NB;
      static_assert((BM * 8 / NLT) % LB == 0 && (BN * 8 / NLT) % LB == 0, "loader batch size");
      const int ltid = threadIdx.x - LOADER_WARP0 * 32;
      struct Pos { int n, kb, sub; Unit u; uint32_t it; bool ok; };
      Pos cur; cur.n = 0; cur.sub = 0; cur.it = 0;
      cur.ok = get_unit(0, cur.u); cur.kb = cur.u.kb_begin;
      auto advance = [&](Pos p) -> Pos {
        if (++p.sub < NSUB) return p;
        p.sub = 0; ++p.it;
        if (++p.kb < p.u.kb_end) return p;
        ++p.n;
        p.ok = get_unit(p.n, p.u); p.kb = p.u.kb_begin;
        return p;
      };
      auto issue = [&](const Pos& p, CvtRegs& g) {
        if (A_CVT && p.sub < NA) cvt_issue<BM, A_MN>(sa, p.sub, p.u.i0, a_olim(p.u), r_base(p.u) + p.kb * BK, r_lim(p.u), ltid, g);
        else if (B_CVT) cvt_issue<BN, B_MN>(sb, p.sub - NA, p.u.j0, OJ, r_base(p.u) + p.kb * BK, r_lim(p.u), ltid, g);
      };
      auto retire = [&](const Pos& p, const CvtRegs& g) {
        const int s = p.it % NS;
        if (p.sub == 0) mbar_wait(&empty_bar[s], ((p.it / NS) & 1) ^ 1);
        uint8_t* a_base = smem + (size_t)s * cfg::STAGE;
        if (A_CVT && p.sub < NA) cvt_store<NP>(g, a_base, cfg::A_PLANE);
        else cvt_store<NP>(g, a_base + NP * cfg::A_PLANE, cfg::B_PLANE);
        if (p.sub == NSUB - 1) {
          fence_proxy_async();
          __syncwarp();
          if (lane == 0) mbar_arrive(&full_bar[s]);
        }
      };
      CvtRegs g0, g1;
      if (cur.ok) issue(cur, g0);
      while (cur.ok) {
        Pos nxt = advance(cur);
        if (nxt.ok) issue(nxt, g1);
        retire(cur, g0);
        if (!nxt.ok) break;
        cur = advance(nxt);
        if (cur.ok) issue(cur, g0);
        retire(nxt, g1);
      }
    }
  } else if (warp == TMA_WARP) {
    if (lane == 0 && TMA_BYTES != 0) {
      uint32_t it = 0;
      Unit u;
      for (int n = 0; get_unit(n, u); ++n) {
        const int i0 = u.i0, j0 = u.j0;
        const int b_row0 = ob.row0 + (SK ? 0 : g_wlev[u.g] * b_grp_rows);   // the group's member of a stacked packed weight
        const int rb = r_base(u);
        for (int kb = u.kb_begin; kb < u.kb_end; ++kb, ++it) {
          const int s = it % NS;
          const uint32_t ph = (it / NS) & 1;
          mbar_wait(&empty_bar[s], ph ^ 1);
          uint8_t* a_base = smem + (size_t)s * cfg::STAGE;
          uint8_t* b_base = a_base + NP * cfg::A_PLANE;
          if (CG == 2) {
            // both CTAs load their A rows and their half of the B tile; every byte is counted on the leader's barrier
            if (rank == 0) mbar_arrive_expect_tx(&full_bar[s], 2 * TMA_BYTES);
            const uint32_t fb = mapa_rank(smem_u32(&full_bar[s]), 0);
            const int jh = j0 + (int)rank * (BN / 2);
#pragma unroll
            for (int pl = 0; pl < NP; ++pl) {
#pragma unroll
              for (int bx = 0; bx < BM / 64; ++bx) {
                uint8_t* dst = a_base + pl * cfg::A_PLANE + bx * 8192;
                if (!A_MN) tma_load_2d_pair(dst, &tmA, oa.col0 + kb * BK, pl * oa.plane_rows + oa.row0 + i0 + bx * 64, fb);
                else       tma_load_2d_pair(dst, &tmA, oa.col0 + i0 + bx * 64, pl * oa.plane_rows + oa.row0 + rb + kb * BK, fb);
              }
#pragma unroll
              for (int bx = 0; bx < BN / 2 / 64; ++bx) {
                uint8_t* dst = b_base + pl * cfg::B_PLANE + bx * 8192;
                if (!B_MN) tma_load_2d_pair(dst, &tmB, ob.col0 + kb * BK, pl * ob.plane_rows + b_row0 + jh + bx * 64, fb);
                else       tma_load_2d_pair(dst, &tmB, ob.col0 + jh + bx * 64, pl * ob.plane_rows + b_row0 + rb + kb * BK, fb);
              }
            }
            continue;
          }
          mbar_arrive_expect_tx(&full_bar[s], TMA_BYTES);
#pragma unroll
          for (int pl = 0; pl < NP; ++pl) {
#pragma unroll
            for (int bx = 0; bx < (A_CVT ? 0 : BM / 64); ++bx) {
              uint8_t* dst = a_base + pl * cfg::A_PLANE + bx * 8192;
              if (!A_MN) tma_load_2d(dst, &tmA, oa.col0 + kb * BK, pl * oa.plane_rows + oa.row0 + i0 + bx * 64, &full_bar[s]);
              else       tma_load_2d(dst, &tmA, oa.col0 + i0 + bx * 64, pl * oa.plane_rows + oa.row0 + rb + kb * BK, &full_bar[s]);
            }
#pragma unroll
            for (int bx = 0; bx < (B_CVT ? 0 : BN / 64); ++bx) {
              uint8_t* dst = b_base + pl * cfg::B_PLANE + bx * 8192;
              if (!B_MN) tma_load_2d(dst, &tmB, ob.col0 + kb * BK, pl * ob.plane_rows + b_row0 + j0 + bx * 64, &full_bar[s]);
              else       tma_load_2d(dst, &tmB, ob.col0 + j0 + bx * 64, pl * ob.plane_rows + b_row0 + rb + kb * BK, &full_bar[s]);
            }
          }
        }
      }
    }
  } else if (warp == MMA_WARP) {
    if (lane == 0 && rank == 0) {
      constexpr uint32_t idesc = umma_idesc(TM, BN, A_MN ? 1 : 0, B_MN ? 1 : 0);
      uint32_t it = 0, tcount = 0;
      Unit u;
      for (int n = 0; get_unit(n, u); ++n) {
        const int kb_begin = u.kb_begin, kb_end = u.kb_end;
        const uint32_t buf = tcount & 1, aph = (tcount >> 1) & 1;
        mbar_wait(&acc_empty[buf], aph ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + buf * BN;
        for (int kb = kb_begin; kb < kb_end; ++kb, ++it) {
          const int s = it % NS;
          const uint32_t ph = (it / NS) & 1;
          mbar_wait(&full_bar[s], ph);
          tc_fence_after();
          const uint32_t a_addr = smem_u32(smem + (size_t)s * cfg::STAGE);
          const uint32_t b_addr = a_addr + NP * cfg::A_PLANE;
#pragma unroll
          for (int ks = 0; ks < BK / 16; ++ks) {
            const uint64_t a_hi = A_MN ? umma_desc(a_addr + ks * 2048, 8192, 1024) : umma_desc(a_addr + ks * 32, 16, 1024);
            const uint64_t b_hi = B_MN ? umma_desc(b_addr + ks * 2048, 8192, 1024) : umma_desc(b_addr + ks * 32, 16, 1024);
            const uint32_t acc0 = (kb > kb_begin || ks) ? 1u : 0u;
            if (NP == 2) {
              const uint64_t a_lo = A_MN ? umma_desc(a_addr + cfg::A_PLANE + ks * 2048, 8192, 1024)
                                         : umma_desc(a_addr + cfg::A_PLANE + ks * 32, 16, 1024);
              const uint64_t b_lo = B_MN ? umma_desc(b_addr + cfg::B_PLANE + ks * 2048, 8192, 1024)
                                         : umma_desc(b_addr + cfg::B_PLANE + ks * 32, 16, 1024);
              if (CG == 2) {
                umma_bf16_pair(d_tmem, a_lo, b_hi, idesc, acc0);
                umma_bf16_pair(d_tmem, a_hi, b_lo, idesc, 1u);
                umma_bf16_pair(d_tmem, a_hi, b_hi, idesc, 1u);
              } else {
                umma_bf16(d_tmem, a_lo, b_hi, idesc, acc0);
                umma_bf16(d_tmem, a_hi, b_lo, idesc, 1u);
                umma_bf16(d_tmem, a_hi, b_hi, idesc, 1u);
              }
            } else {
              if (CG == 2) umma_bf16_pair(d_tmem, a_hi, b_hi, idesc, acc0);
              else umma_bf16(d_tmem, a_hi, b_hi, idesc, acc0);
            }
          }
          if (CG == 2) umma_commit_pair(&empty_bar[s], 3); else umma_commit(&empty_bar[s]);
        }
        if (CG == 2) umma_commit_pair(&acc_full[buf], 3); else umma_commit(&acc_full[buf]);
        ++tcount;
      }
    }
  } else {
    // ------------------------------------------------ epilogue warps: TMEM -> registers -> shared (transpose) -> global
    // tcgen05.ld hands lane r the 32 columns of accumulator row r; stored as they come, one instruction would touch 32
    // different 128-byte lines.  Each warp therefore transposes its 32 x 32 chunk through a private 4 KB XOR-swizzled buffer
    // (16-byte slot q of row r lives at slot q ^ (r & 7): conflict-free both ways) and writes rows out 8 lanes x 16 B = one
    // full line per row, 4 rows per instruction.
    float* stg = reinterpret_cast<float*>(smem + (size_t)NS * cfg::STAGE) + warp * 1024;
    const int q = warp & 3, half = warp >> 2;
    const int orow = lane >> 3, oq = lane & 7;               // output mapping: row within a group of 4, 16-byte slot
    uint32_t tcount = 0;
    Unit u;
    for (int n = 0; get_unit(n, u); ++n) {
      const int i0 = u.i0, j0 = u.j0;
      const float* bias = SK ? nullptr : g_bias[u.g];
      const uint32_t buf = tcount & 1, aph = (tcount >> 1) & 1;
      mbar_wait(&acc_full[buf], aph);
      tc_fence_after();
      float* Cb; int ldo, oi_lim, oj_lim, ci0, cj0;
      if (SK) {                                              // partial tile -> workspace slot, row-major [BM][BN]
        Cb = C + (((size_t)vcta * segmax + n) * CG + rank) * (BM * BN); ldo = BN; oi_lim = BM; oj_lim = BN; ci0 = 0; cj0 = 0;
      } else {
        Cb = C; ldo = ldc; oi_lim = a_olim(u); oj_lim = OJ; ci0 = i0; cj0 = j0;
      }
      const bool vec = (ldo % 4 == 0) && ((reinterpret_cast<uintptr_t>(Cb) & 15) == 0);
      // RowAdd: this lane's 8 output rows are the same for every column chunk of the tile -- their slots are read once, and each
      // chunk's addends are requested before the accumulator chunk is fetched and transposed (the gather latency hides behind it)
      const bool has_add = !SK && gg.add.p != nullptr;
      int ai[8];
      if (has_add) {
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const int i = ci0 + q * 32 + 4 * k + orow;
          ai[k] = i < oi_lim ? __ldg(gg.add.idx + i) : -1;
        }
      }
#pragma unroll 1
      for (int c0 = 0; c0 < BN / 2; c0 += 32) {
        uint32_t r[32];
        const int col = half * (BN / 2) + c0;
        // interior chunk (warp-uniform): all 32 rows and 32 columns inside the output, 16-byte stores possible -- the lean path
        // below.  The general path costs ~1300 issued instructions per 32 x 32 chunk (bounds, tails and every epilogue variant
        // predicated per store step), which made the epilogue the pace of every contraction with a short reduction.
        const bool interior = vec && (ci0 + q * 32 + 32 <= oi_lim) && (cj0 + col + 32 <= oj_lim);
        float4 a4[8];
        if (has_add && interior) {
          const float* ap0 = gg.add.p + (cj0 + col + oq * 4);
#pragma unroll
          for (int k = 0; k < 8; ++k)
            a4[k] = ai[k] >= 0 ? __ldg(reinterpret_cast<const float4*>(ap0 + (size_t)ai[k] * gg.add.ld)) : make_float4(0.f, 0.f, 0.f, 0.f);
        } else if (has_add) {
          const int nn = cj0 + col + oq * 4;
#pragma unroll
          for (int k = 0; k < 8; ++k) {
            a4[k] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (ai[k] >= 0 && nn < oj_lim) {
              const float* ap = gg.add.p + (size_t)ai[k] * gg.add.ld + nn;
              if (nn + 3 < oj_lim) a4[k] = __ldg(reinterpret_cast<const float4*>(ap));
              else { a4[k].x = ap[0]; if (nn + 1 < oj_lim) a4[k].y = ap[1]; if (nn + 2 < oj_lim) a4[k].z = ap[2]; }
            }
          }
        }
        tmem_ld32(tmem_base + buf * BN + ((uint32_t)(q * 32) << 16) + (uint32_t)col, r);
#pragma unroll
        for (int j = 0; j < 8; ++j)
          *reinterpret_cast<uint4*>(stg + lane * 32 + ((j ^ (lane & 7)) << 2)) = make_uint4(r[4 * j], r[4 * j + 1], r[4 * j + 2], r[4 * j + 3]);
        __syncwarp();
        const int nn = cj0 + col + oq * 4;                   // first of this lane's 4 output columns
        if (interior) {
          float4 bv = make_float4(0.f, 0.f, 0.f, 0.f);
          if (!SK && bias) bv = make_float4(bias[nn], bias[nn + 1], bias[nn + 2], bias[nn + 3]);
          const bool relu = !SK && (flags & GEMM_RELU), accum = !SK && (flags & GEMM_ACCUM), gates = !SK && (flags & GEMM_GATES);
          const int gc = flags >> 8;
          const bool as_tanh = nn >= 2 * gc && nn < 3 * gc;
          float* cp = Cb + (size_t)(ci0 + q * 32 + orow) * ldo + nn;
          const size_t step = (size_t)4 * ldo;
#pragma unroll
          for (int k = 0; k < 8; ++k, cp += step) {
            const int row = 4 * k + orow;
            float4 w = *reinterpret_cast<const float4*>(stg + row * 32 + ((oq ^ (row & 7)) << 2));
            if (!SK) {
              if (has_add) { w.x += a4[k].x; w.y += a4[k].y; w.z += a4[k].z; w.w += a4[k].w; }
              w.x += bv.x; w.y += bv.y; w.z += bv.z; w.w += bv.w;
              if (relu) { w.x = fmaxf(w.x, 0.f); w.y = fmaxf(w.y, 0.f); w.z = fmaxf(w.z, 0.f); w.w = fmaxf(w.w, 0.f); }
              if (gates) {
                if (NP == 1) {
                  if (as_tanh) { w.x = tanh_mufu(w.x); w.y = tanh_mufu(w.y); w.z = tanh_mufu(w.z); w.w = tanh_mufu(w.w); }
                  else { w.x = sigmoid_mufu(w.x); w.y = sigmoid_mufu(w.y); w.z = sigmoid_mufu(w.z); w.w = sigmoid_mufu(w.w); }
                } else if (as_tanh) { w.x = tanh_fast(w.x); w.y = tanh_fast(w.y); w.z = tanh_fast(w.z); w.w = tanh_fast(w.w); }
                else { w.x = sigmoid_fast(w.x); w.y = sigmoid_fast(w.y); w.z = sigmoid_fast(w.z); w.w = sigmoid_fast(w.w); }
              }
              if (accum) { const float4 c = *reinterpret_cast<const float4*>(cp); w.x += c.x; w.y += c.y; w.z += c.z; w.w += c.w; }
            }
            *reinterpret_cast<float4*>(cp) = w;
          }
          __syncwarp();
          continue;
        }
        float4 bv = make_float4(0.f, 0.f, 0.f, 0.f);
        if (!SK && bias) {
          if (nn + 3 < oj_lim) bv = make_float4(bias[nn], bias[nn + 1], bias[nn + 2], bias[nn + 3]);
          else { if (nn < oj_lim) bv.x = bias[nn]; if (nn + 1 < oj_lim) bv.y = bias[nn + 1]; if (nn + 2 < oj_lim) bv.z = bias[nn + 2]; }
        }
#pragma unroll
        for (int r4 = 0; r4 < 32; r4 += 4) {
          const int row = r4 + orow;
          float4 w = *reinterpret_cast<const float4*>(stg + row * 32 + ((oq ^ (row & 7)) << 2));
          const int i = ci0 + q * 32 + row;
          if (i >= oi_lim || nn >= oj_lim) continue;
          float* cp = Cb + (size_t)i * ldo + nn;
          if (!SK) {
            if (has_add) { const float4 a = a4[r4 >> 2]; w.x += a.x; w.y += a.y; w.z += a.z; w.w += a.w; }   // shared-parent term (RowAdd)
            w.x += bv.x; w.y += bv.y; w.z += bv.z; w.w += bv.w;
            if (flags & GEMM_RELU) { w.x = fmaxf(w.x, 0.f); w.y = fmaxf(w.y, 0.f); w.z = fmaxf(w.z, 0.f); w.w = fmaxf(w.w, 0.f); }
            if (flags & GEMM_GATES) {                        // column ranges are multiples of 4: one branch per float4
              const int gc = flags >> 8;
              if (NP == 1) {
                if (nn >= 2 * gc && nn < 3 * gc) { w.x = tanh_mufu(w.x); w.y = tanh_mufu(w.y); w.z = tanh_mufu(w.z); w.w = tanh_mufu(w.w); }
                else { w.x = sigmoid_mufu(w.x); w.y = sigmoid_mufu(w.y); w.z = sigmoid_mufu(w.z); w.w = sigmoid_mufu(w.w); }
              } else if (nn >= 2 * gc && nn < 3 * gc) { w.x = tanh_fast(w.x); w.y = tanh_fast(w.y); w.z = tanh_fast(w.z); w.w = tanh_fast(w.w); }
              else { w.x = sigmoid_fast(w.x); w.y = sigmoid_fast(w.y); w.z = sigmoid_fast(w.z); w.w = sigmoid_fast(w.w); }
            }
          }
          if (vec && nn + 3 < oj_lim) {
            if (!SK && (flags & GEMM_ACCUM)) { const float4 c = *reinterpret_cast<const float4*>(cp); w.x += c.x; w.y += c.y; w.z += c.z; w.w += c.w; }
            *reinterpret_cast<float4*>(cp) = w;
          } else {
            const float o[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
            for (int e = 0; e < 4; ++e)
              if (nn + e < oj_lim) cp[e] = (!SK && (flags & GEMM_ACCUM)) ? cp[e] + o[e] : o[e];
          }
        }
        __syncwarp();
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if (CG == 2) mbar_arrive_cluster(mapa_rank(smem_u32(&acc_empty[buf]), 0));   // the leader's MMA warp waits for both CTAs
        else mbar_arrive(&acc_empty[buf]);
      }
      ++tcount;
    }
  }
  tc_fence_before();
  __syncthreads();
  if (CG == 2) cluster_sync_all();                           // no signal may still be under way to a CTA that has exited
  if (warp == MMA_WARP) { if (CG == 2) tmem_dealloc_pair(tmem_base, 2 * BN); else tmem_dealloc(tmem_base, 2 * BN); }
}

// ---------------------------------------------------------------------------------------------- packed weights
// fp32 [R, Cc] (leading dim ld) -> bf16 planes stacked as [2R, ldp]: rows [0,R) = hi, [R,2R) = lo.
struct PackJob { const float* src; __nv_bfloat16* dst; int R, Cc, ld, ldp; };

__global__ void __launch_bounds__(256) pack_weights_kernel(const PackJob* __restrict__ jobs) {
  pdl_grid_sync();
  const PackJob jb = jobs[blockIdx.y];
  const size_t total = (size_t)jb.R * jb.ldp;
  for (size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
    const int r = (int)(e / jb.ldp), c = (int)(e % jb.ldp);
    const float x = c < jb.Cc ? jb.src[(size_t)r * jb.ld + c] : 0.f;
    const __nv_bfloat16 h = __float2bfloat16_rn(x);
    jb.dst[e] = h;
    jb.dst[total + e] = __float2bfloat16_rn(x - __bfloat162float(h));
  }
}

struct Packed {
  const float* src; int R, Cc, ld, ldp;
  int nstack;                          // > 1: head of a stack of same-shaped matrices packed back to back (2R rows apart); the
  __nv_bfloat16* dst;                  //      map then spans the whole stack so a grouped contraction can address member g
  CUtensorMap map;
};

std::mutex g_mu;
std::vector<Packed> g_packed;

typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                             const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                             CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeFn g_encode = nullptr;

int get_encode() {
  if (g_encode) return 0;
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult qres;
  PATHS_CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres));
  PATHS_REQUIRE(fn != nullptr && qres == cudaDriverEntryPointSuccess, "cuTensorMapEncodeTiled is not available from this driver");
  g_encode = reinterpret_cast<EncodeFn>(fn);
  return 0;
}

int make_map(Packed& p) {
  if (int r = get_encode()) return r;
  cuuint64_t gdim[2] = {(cuuint64_t)p.ldp, (cuuint64_t)(2 * p.R) * (cuuint64_t)(p.nstack > 1 ? p.nstack : 1)};
  cuuint64_t gstr[1] = {(cuuint64_t)p.ldp * 2};
  cuuint32_t box[2] = {64, 64};
  cuuint32_t estr[2] = {1, 1};
  CUresult rc = g_encode(&p.map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, p.dst, gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                         CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  PATHS_REQUIRE(rc == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed (%d) for a [%d x %d] weight", (int)rc, 2 * p.R, p.ldp);
  return 0;
}

const Packed* find_packed(const float* B, int ldb, int* row0, int* col0) {
  for (const Packed& p : g_packed) {
    if (p.ld != ldb) continue;
    const ptrdiff_t off = B - p.src;
    if (off < 0 || off >= (ptrdiff_t)p.R * p.ld) continue;
    *row0 = (int)(off / p.ld); *col0 = (int)(off % p.ld);
    return &p;
  }
  return nullptr;
}

template <typename Kern>
int set_smem(Kern kern, int bytes) {
  PATHS_CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
  return 0;
}

}  // namespace
int split_rows(const RowSrc& A, int M_host, const int* M_dev, int K, const Planes& dst, int col0, bool want_lo, cudaStream_t st);
namespace {
// activation planes: scratch provided by the engine (or by the stand-alone entry points)
struct Scratch { void* p[2]; size_t bytes[2]; } g_scratch = {{nullptr, nullptr}, {0, 0}};

struct MapKey { const void* ptr; int rows, cols, ldp; };
struct MapEnt { MapKey k; CUtensorMap map; };
std::vector<MapEnt> g_maps;          // tiny cache: scratch pointers and shapes repeat every step

int encode_map(CUtensorMap* out, const void* ptr, int rows, int cols, int ldp) {
  if (int r = get_encode()) return r;
  cuuint64_t gdim[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t gstr[1] = {(cuuint64_t)ldp * 2};
  cuuint32_t box[2] = {64, 64};
  cuuint32_t estr[2] = {1, 1};
  CUresult rc = g_encode(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), gdim, gstr, box, estr,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  PATHS_REQUIRE(rc == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed (%d) for a [%d x %d] plane (ld %d)", (int)rc, rows, cols, ldp);
  return 0;
}

int cached_map(const CUtensorMap** out, const void* ptr, int rows, int cols, int ldp) {
  for (MapEnt& e : g_maps)
    if (e.k.ptr == ptr && e.k.rows == rows && e.k.cols == cols && e.k.ldp == ldp) { *out = &e.map; return 0; }
  if (g_maps.size() > 4096) g_maps.clear();
  g_maps.reserve(4200);
  MapEnt e;
  e.k = MapKey{ptr, rows, cols, ldp};
  if (int r = encode_map(&e.map, ptr, rows, cols, ldp)) return r;
  g_maps.push_back(e);
  *out = &g_maps.back().map;
  return 0;
}

// splits A (fp32, RowSrc) into scratch `which`; returns the planes record
int split_to_scratch(int which, const RowSrc& A, int M_host, const int* M_dev, int K, bool want_lo, Planes* out, cudaStream_t st) {
  const size_t need = planes_bytes(M_host, K);
  PATHS_REQUIRE(g_scratch.p[which] && need <= g_scratch.bytes[which], "tensor-core scratch %d too small: need %zu bytes, have %zu", which,
                need, g_scratch.bytes[which]);
  *out = planes_at(g_scratch.p[which], M_host, K);
  return split_rows(A, M_host, M_dev, K, *out, 0, want_lo, st);
}

template <int BN, int NP, bool A_MN, bool B_MN, bool SK, bool A_CVT, bool B_CVT>
int launch_tma(const CUtensorMap* ma, const CUtensorMap* mb, Operand oa, Operand ob, const RowSrc& sa, const RowSrc& sb, float* C, int ldc,
               const GemmGrp& gg, int b_grp_rows, int OI_host, int OJ, int R_host, int flags, int grid_sk, int segmax, cudaStream_t st) {
  auto kern = gemm_tma_kernel<BN, NP, A_MN, B_MN, SK, A_CVT, B_CVT>;
  static bool configured = false;
  if (!configured) { if (int r = set_smem(kern, Cfg<BN, NP>::SMEM)) return r; configured = true; }
  static const CUtensorMap dummy = {};                       // stands in for the map of an operand converted in-kernel
  long row_tiles = 0;                                         // by capacity: the live tile count is only known on the device
  for (int g = 0; g < gg.rows.n; ++g) row_tiles += (gg.rows.hcnt[g] + BM - 1) / BM;
  long units = (long)((OJ + BN - 1) / BN) * row_tiles;
  if (units < 1) units = 1;
  const int grid = SK ? grid_sk : (units < kNumSMs ? (int)units : kNumSMs);
  pdl(kern, grid, (A_CVT || B_CVT) ? NTHREADS_CVT : NTHREADS_TMA, Cfg<BN, NP>::SMEM, st)(
      A_CVT ? dummy : *ma, B_CVT ? dummy : *mb, oa, ob, sa, sb, C, ldc, gg, b_grp_rows, OI_host, OJ, R_host, flags, segmax);
  PATHS_LAUNCH_CK();
  return 0;
}

// CTA-pair launch (gemm_tma_kernel<..., CG = 2>): clusters of 2, one pair per TPC.  Off with PATHS_B200_NO_PAIR=1 (A/B runs).
bool pair_enabled() {
  static const bool on = getenv("PATHS_B200_NO_PAIR") == nullptr;
  return on;
}
// fewest 256 x 256 tiles worth a pair launch: half a wave; PATHS_B200_FORCE_PAIR=1 (tests) sends every eligible contraction there
long pair_min_units() {
  static const long n = getenv("PATHS_B200_FORCE_PAIR") ? 1 : kNumSMs / 2;
  return n;
}
// maximum number of co-resident pairs of a kernel instance (clusters of 2 at one CTA per SM)
template <typename K>
int pair_capacity(K kern, int smem, int* out) {
  cudaLaunchConfig_t cfg = {};
  cfg.blockDim = dim3(NTHREADS_TMA); cfg.dynamicSmemBytes = smem; cfg.gridDim = dim3(kNumSMs);
  cudaLaunchAttribute at;
  at.id = cudaLaunchAttributeClusterDimension;
  at.val.clusterDim.x = 2; at.val.clusterDim.y = 1; at.val.clusterDim.z = 1;
  cfg.attrs = &at; cfg.numAttrs = 1;
  if (int r = set_smem(kern, smem)) return r;
  int nc = 0;
  PATHS_CK(cudaOccupancyMaxActiveClusters(&nc, kern, &cfg));
  PATHS_REQUIRE(nc >= 1, "gemm_tc: the device cannot co-schedule a CTA pair of the contraction kernel");
  *out = nc < kNumSMs / 2 ? nc : kNumSMs / 2;
  return 0;
}
// SK = true: pairs_sk pairs share the stream-K plan (tn_plan with cg = 2)
template <int NP, bool A_MN, bool B_MN, bool SK>
int launch_pair(const CUtensorMap* ma, const CUtensorMap* mb, Operand oa, Operand ob, float* C, int ldc, const GemmGrp& gg, int b_grp_rows,
                int OI_host, int OJ, int R_host, int flags, int pairs_sk, int segmax, cudaStream_t st) {
  constexpr int BN = 256;
  auto kern = gemm_tma_kernel<BN, NP, A_MN, B_MN, SK, false, false, 2>;
  static int max_pairs = 0;
  cudaLaunchConfig_t cfg = {};
  cfg.blockDim = dim3(NTHREADS_TMA); cfg.dynamicSmemBytes = Cfg<BN, NP, 2>::SMEM; cfg.stream = st;
  cudaLaunchAttribute at[2];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
  at[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  if (!max_pairs)
    if (int r = pair_capacity(kern, Cfg<BN, NP, 2>::SMEM, &max_pairs)) return r;
  int pairs;
  if (SK) {
    PATHS_REQUIRE(pairs_sk >= 1, "gemm_tn_tc: empty stream-K plan");
    pairs = pairs_sk;                                        // fixed by the host's plan (slot layout); segments are independent
  } else {
    long row_tiles = 0;
    for (int g = 0; g < gg.rows.n; ++g) row_tiles += (gg.rows.hcnt[g] + 2 * BM - 1) / (2 * BM);
    long units = (long)((OJ + BN - 1) / BN) * row_tiles;
    if (units < 1) units = 1;
    pairs = units < max_pairs ? (int)units : max_pairs;
  }
  cfg.gridDim = dim3(2 * pairs); cfg.numAttrs = pdl_enabled() ? 2 : 1;
  static const RowSrc none = row_src(nullptr, 0, 0);
  cudaLaunchKernelEx(&cfg, kern, *ma, *mb, oa, ob, none, none, C, ldc, gg, b_grp_rows, OI_host, OJ, R_host, flags, segmax);
  PATHS_LAUNCH_CK();
  return 0;
}

// C[tile] += sum over the stream-K segments of the tile, in a fixed order (deterministic).  A block owns 256 / PG consecutive
// float4 outputs of one tile; PG thread groups each sum every PG-th segment and are combined through shared memory.
// CG = 2: the partial tiles were produced by CTA pairs -- plan and slots are those of the 256-row pair tile T that holds this
// 128-row tile, whose partials sit at [slot of T][half = i-tile & 1]; G counts pairs.
template <int BN, int PG, int CG = 1>
__global__ void __launch_bounds__(256) tn_reduce_kernel(const float4* __restrict__ ws, const GemmGrp gg, int ldc, int NI, int NJ, int G,
                                                        int segmax) {
  pdl_grid_sync();
  constexpr int OUT = 256 / PG;
  int R = 0;
#pragma unroll
  for (int g = 0; g < kMaxGroups; ++g) {
    const int c = g < gg.rows.n ? (gg.rows.cnt[g] ? *gg.rows.cnt[g] : gg.rows.hcnt[g]) : 0;
    R = c > R ? c : R;
  }
  const int nkb = (R + BK - 1) / BK;
  if (nkb == 0) return;
  const int tiles_j = (NJ + BN - 1) / BN;
  const int tpg = ((NI + BM - 1) / BM) * tiles_j;            // output tiles per group
  float* __restrict__ C = pick(gg.cout, blockIdx.x / tpg);
  const int tt = blockIdx.x % tpg;
  // t, units: the tile this block sums and the tile count IN THE PRODUCER'S PLAN (pair tiles when CG = 2)
  const int tpg_p = ((NI + BM * CG - 1) / (BM * CG)) * tiles_j;
  const int t = CG == 1 ? (int)blockIdx.x : (int)(blockIdx.x / tpg) * tpg_p + (tt / tiles_j / 2) * tiles_j + tt % tiles_j;
  const int half = CG == 1 ? 0 : (tt / tiles_j) & 1;
  const int units = gg.rows.n * tpg_p;
  const SkPlan pl = sk_plan(units, nkb, G);
  const int o = threadIdx.x % OUT, g = threadIdx.x / OUT;
  const int elem = blockIdx.y * OUT + o;                 // float4 index inside the tile, row-major
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
  // contributions in a fixed order: (lock-step CTA t, if any), then the streaming CTAs that touched the tile
  const int c_base = pl.k1 >= 0 ? units : 0;
  const long long t_lo = (long long)t * pl.tl;
  const int c_first = pl.tl > 0 ? (int)(t_lo / pl.q) : 0, c_last = pl.tl > 0 ? (int)((t_lo + pl.tl - 1) / pl.q) : -1;
  const int extra = pl.k1 > 0 ? 1 : 0;
  for (int k = g; k < extra + (c_last - c_first + 1); k += PG) {
    size_t slot;
    if (k < extra) slot = (size_t)t * segmax;
    else {
      const int c = c_first + (k - extra);
      slot = (size_t)(c_base + c) * segmax + (t - (int)(((long long)c * pl.q) / pl.tl));
    }
    const float4 v = __ldcs(ws + (slot * CG + half) * (BM * BN / 4) + elem);
    acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
  }
  if (PG > 1) {
    __shared__ float4 red[PG][OUT];
    red[g][o] = acc;
    __syncthreads();
    if (g != 0) return;
#pragma unroll
    for (int k = 1; k < PG; ++k) { const float4 v = red[k][o]; acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w; }
  }
  const int row = elem / (BN / 4), c4 = elem % (BN / 4);
  const int i = (tt / tiles_j) * BM + row, j = (tt % tiles_j) * BN + c4 * 4;
  if (i >= NI || j >= NJ) return;
  float* cp = C + (size_t)i * ldc + j;
  const bool vec = (ldc % 4 == 0) && ((reinterpret_cast<uintptr_t>(C) & 15) == 0);
  if (vec && j + 3 < NJ) {
    float4 w = *reinterpret_cast<float4*>(cp);
    w.x += acc.x; w.y += acc.y; w.z += acc.z; w.w += acc.w;
    *reinterpret_cast<float4*>(cp) = w;
  } else {
    const float av[4] = {acc.x, acc.y, acc.z, acc.w};
    for (int k = 0; k < 4 && j + k < NJ; ++k) cp[k] += av[k];
  }
}

struct TnWs { void* p; size_t bytes; } g_tn_ws = {nullptr, 0};

}  // namespace

// ---- registry API (engine.cu registers every weight matrix it contracts against) -----------------------------------
size_t packed_bytes(int R, int Cc) { return (size_t)2 * R * ((Cc + 7) / 8 * 8) * sizeof(__nv_bfloat16) + 256; }

int register_packed(const float* src, int R, int Cc, int ld, void* dst, int nstack) {
  std::lock_guard<std::mutex> lk(g_mu);
  for (auto it = g_packed.begin(); it != g_packed.end();)
    it = (it->src == src) ? g_packed.erase(it) : it + 1;
  Packed p;
  p.src = src; p.R = R; p.Cc = Cc; p.ld = ld; p.ldp = (Cc + 7) / 8 * 8; p.nstack = nstack;
  p.dst = reinterpret_cast<__nv_bfloat16*>((reinterpret_cast<uintptr_t>(dst) + 255) & ~uintptr_t(255));
  if (int r = make_map(p)) return r;
  g_packed.push_back(p);
  return 0;
}

// where the registry currently keeps the packed copy of `src` (nullptr: not registered)
const void* packed_dst_of(const float* src) {
  std::lock_guard<std::mutex> lk(g_mu);
  for (const Packed& p : g_packed)
    if (p.src == src) return p.dst;
  return nullptr;
}

void unregister_packed(const float* src) {
  std::lock_guard<std::mutex> lk(g_mu);
  for (auto it = g_packed.begin(); it != g_packed.end();)
    it = (it->src == src) ? g_packed.erase(it) : it + 1;
}

// Builds the device job table for the given registered weights (once, at bind time) ...
int build_pack_jobs(const float* const* srcs, int n, void* jobs_dev, int* n_jobs, cudaStream_t st) {
  std::vector<PackJob> jobs;
  {
    std::lock_guard<std::mutex> lk(g_mu);
    for (int i = 0; i < n; ++i)
      for (const Packed& p : g_packed)
        if (p.src == srcs[i]) { jobs.push_back(PackJob{p.src, p.dst, p.R, p.Cc, p.ld, p.ldp}); break; }
  }
  *n_jobs = (int)jobs.size();
  if (jobs.empty()) return 0;
  PATHS_CK(cudaMemcpyAsync(jobs_dev, jobs.data(), jobs.size() * sizeof(PackJob), cudaMemcpyHostToDevice, st));
  PATHS_CK(cudaStreamSynchronize(st));
  return 0;
}
// ... and re-splits them into bf16 (hi, lo) planes: one launch per step (the weights change with every optimizer step).
int run_pack_jobs(const void* jobs_dev, int n_jobs, cudaStream_t st) {
  if (n_jobs <= 0) return 0;
  pdl(pack_weights_kernel, dim3(296, (unsigned)n_jobs), 256, 0, st)(reinterpret_cast<const PackJob*>(jobs_dev));
  PATHS_LAUNCH_CK();
  return 0;
}
size_t pack_job_bytes() { return sizeof(PackJob); }

bool gemm_tc_supported_nt(const RowSrc& A, const float* B, int ldb, int b_kn, int ldc, int N, int K) {
  if (K % 8 != 0 || A.k0 % 4 != 0 || A.ld0 % 4 != 0 || (A.p1 && A.ld1 % 4 != 0)) return false;
  if ((reinterpret_cast<uintptr_t>(A.p0) & 15) || (A.p1 && (reinterpret_cast<uintptr_t>(A.p1) & 15))) return false;
  std::lock_guard<std::mutex> lk(g_mu);
  int r0, c0;
  return find_packed(B, ldb, &r0, &c0) != nullptr;
}
bool gemm_tc_supported_tn(int ldp, const RowSrc& Q, int ldc, int NI, int NJ) {
  if (ldp % 4 != 0 || NI % 8 != 0 || NJ % 8 != 0) return false;
  if (Q.k0 % 4 != 0 || Q.ld0 % 4 != 0 || (Q.p1 && Q.ld1 % 4 != 0)) return false;
  if ((reinterpret_cast<uintptr_t>(Q.p0) & 15) || (Q.p1 && (reinterpret_cast<uintptr_t>(Q.p1) & 15))) return false;
  return NI >= 8 && NJ >= 8;
}

void set_tc_scratch(void* a, size_t a_bytes, void* b, size_t b_bytes) {
  g_scratch.p[0] = a; g_scratch.bytes[0] = a_bytes; g_scratch.p[1] = b; g_scratch.bytes[1] = b_bytes;
}
size_t tc_scratch_bytes(int rows, int cols) { return planes_bytes(rows, cols) + 256; }

int split_rows(const RowSrc& A, int M_host, const int* M_dev, int K, const Planes& dst, int col0, bool want_lo, cudaStream_t st) {
  PATHS_REQUIRE(dst.p && K % 4 == 0 && col0 % 4 == 0 && col0 + K <= dst.ldp, "split_rows: bad destination window (col0 %d, K %d, ldp %d)", col0,
                K, dst.ldp);
  if (M_host <= 0) return 0;
  int g = (planes_pad_rows(M_host) + 7) / 8; if (g > 16 * kNumSMs) g = 16 * kNumSMs;
  pdl(split_planes_kernel, g, 256, 0, st)(A, M_host, M_dev, K, dst, col0, want_lo ? 1 : 0);
  PATHS_LAUNCH_CK();
  return 0;
}

// C[M,N] (+)= act(A . B^T + bias); A either as planes (columns [a_col0, a_col0 + K)) or as fp32 rows converted in-kernel
static int nt_impl(int passes, const Planes* A, int a_col0, const RowSrc* Asrc, const float* B, int ldb, int b_kn, float* C, int ldc,
                   const float* bias, int M_host, const int* M_dev, int N, int K, int flags, cudaStream_t st, const GemmGrp* grp = nullptr,
                   const RowAdd* add = nullptr) {
  if (M_host <= 0 || N <= 0) return 0;
  GemmGrp gg;
  if (grp) gg = *grp;
  else {
    gg.rows = grp_single(M_host, M_dev);
    for (int i = 0; i < kMaxGroups; ++i) { gg.wlev[i] = 0; gg.bias[i] = nullptr; gg.cout[i] = nullptr; }
    gg.bias[0] = bias;
    gg.add = add ? *add : no_row_add();
    PATHS_REQUIRE(!gg.add.p || (gg.add.idx && gg.add.ld % 4 == 0 && !(reinterpret_cast<uintptr_t>(gg.add.p) & 15)), "gemm_nt_tc: bad row addend");
  }
  const Packed* pk; int r0, c0;
  {
    std::lock_guard<std::mutex> lk(g_mu);
    pk = find_packed(B, ldb, &r0, &c0);
  }
  PATHS_REQUIRE(pk, "gemm_nt_tc: weight matrix was not registered for the tensor-core path");
  const CUtensorMap* ma = nullptr;
  Operand oa; oa.row0 = 0; oa.col0 = a_col0; oa.plane_rows = 0;
  RowSrc none = row_src(nullptr, 0, 0);
  if (A) {
    PATHS_REQUIRE(A->p && a_col0 % 8 == 0 && M_host <= A->rows, "gemm_nt_tc: bad activation planes");
    if (int r = cached_map(&ma, A->p, 2 * A->rows, a_col0 + K, A->ldp)) return r;
    oa.plane_rows = A->rows;
  }
  Operand ob; ob.row0 = r0; ob.col0 = c0; ob.plane_rows = pk->R;
  const int bgr = 2 * pk->R;                                   // rows between the members of a stacked packed weight
  if (grp)
    for (int g = 0; g < gg.rows.n; ++g)
      PATHS_REQUIRE(gg.wlev[g] >= 0 && gg.wlev[g] < (pk->nstack > 1 ? pk->nstack : 1), "gemm_nt_tc: group %d addresses member %d of a %d-stack",
                    g, gg.wlev[g], pk->nstack);
  const bool big = N > 128;
  // CTA pairs for the wide outputs with TMA-fed operands, when there are enough 256-row tiles to keep the pairs busy
  long pair_tiles = 0;
  for (int g = 0; g < gg.rows.n; ++g) pair_tiles += (gg.rows.hcnt[g] + 2 * BM - 1) / (2 * BM);
  const bool pair = A && big && pair_enabled() && pair_tiles * ((N + 255) / 256) >= pair_min_units();
  if (pair) {
    if (passes == 3) return b_kn ? launch_pair<2, false, true, false>(ma, &pk->map, oa, ob, C, ldc, gg, bgr, M_host, N, K, flags, 0, 0, st)
                                 : launch_pair<2, false, false, false>(ma, &pk->map, oa, ob, C, ldc, gg, bgr, M_host, N, K, flags, 0, 0, st);
    return b_kn ? launch_pair<1, false, true, false>(ma, &pk->map, oa, ob, C, ldc, gg, bgr, M_host, N, K, flags, 0, 0, st)
                : launch_pair<1, false, false, false>(ma, &pk->map, oa, ob, C, ldc, gg, bgr, M_host, N, K, flags, 0, 0, st);
  }
#define NT_CASE(BN_, NP_, MN_)                                                                                                         \
  do {                                                                                                                                 \
    if (A) return launch_tma<BN_, NP_, false, MN_, false, false, false>(ma, &pk->map, oa, ob, none, none, C, ldc, gg, bgr, M_host,     \
                                                                        N, K, flags, 0, 0, st);                                      \
    return launch_tma<BN_, NP_, false, MN_, false, true, false>(nullptr, &pk->map, oa, ob, *Asrc, none, C, ldc, gg, bgr, M_host, N,    \
                                                                K, flags, 0, 0, st);                                                 \
  } while (0)
  if (passes == 3) {
    if (big) { if (b_kn) NT_CASE(256, 2, true); else NT_CASE(256, 2, false); }
    else     { if (b_kn) NT_CASE(128, 2, true); else NT_CASE(128, 2, false); }
  } else {
    if (big) { if (b_kn) NT_CASE(256, 1, true); else NT_CASE(256, 1, false); }
    else     { if (b_kn) NT_CASE(128, 1, true); else NT_CASE(128, 1, false); }
  }
#undef NT_CASE
  return 0;
}

int gemm_nt_tc_p(int passes, const Planes& A, int a_col0, const float* B, int ldb, int b_kn, float* C, int ldc, const float* bias, int M_host,
                 const int* M_dev, int N, int K, int flags, cudaStream_t st, const RowAdd* add) {
  return nt_impl(passes, &A, a_col0, nullptr, B, ldb, b_kn, C, ldc, bias, M_host, M_dev, N, K, flags, st, nullptr, add);
}

// fp32 rows qualify for the in-kernel conversion when every 8-element chunk lies inside one 16-byte-aligned source row
static bool cvt_ok(const RowSrc& A) {
  if (A.k0 % 8 != 0 || A.ld0 % 4 != 0 || (A.p1 && A.ld1 % 4 != 0)) return false;
  return !(reinterpret_cast<uintptr_t>(A.p0) & 15) && !(A.p1 && (reinterpret_cast<uintptr_t>(A.p1) & 15));
}

int gemm_nt_tc(int passes, const RowSrc& A, const float* B, int ldb, int b_kn, float* C, int ldc, const float* bias, int M_host,
               const int* M_dev, int N, int K, int flags, cudaStream_t st) {
  if (M_host <= 0 || N <= 0) return 0;
  if (cvt_ok(A)) return nt_impl(passes, nullptr, 0, &A, B, ldb, b_kn, C, ldc, bias, M_host, M_dev, N, K, flags, st);
  Planes pa;
  if (int r = split_to_scratch(0, A, M_host, M_dev, K, passes == 3, &pa, st)) return r;
  return gemm_nt_tc_p(passes, pa, 0, B, ldb, b_kn, C, ldc, bias, M_host, M_dev, N, K, flags, st, nullptr);
}

// Stream-K plan of a weight-gradient contraction: CTAs and workspace slots per CTA.
// cg = 2: the plan of CTA pairs (tiles of 256 rows, G pairs)
static void tn_plan(int NI, int NJ, int M_host, int ngroups, int* G, int* segmax, int cg = 1) {
  const int BN = NJ > 128 ? 256 : 128;
  const int tiles = ngroups * ((NI + BM * cg - 1) / (BM * cg)) * ((NJ + BN - 1) / BN);
  const long total = (long)tiles * ((M_host + BK - 1) / BK);
  long g = total / 4;                            // at least ~4 k-blocks per CTA
  if (g > kNumSMs / cg) g = kNumSMs / cg;
  if (g < 1) g = 1;
  *G = (int)g;
  *segmax = (tiles + *G - 1) / *G + 2;
  if (*G > tiles) {                              // tail CTAs of the lock-step plan touch ceil(tiles / (G - tiles)) + 1 tiles at most
    const int s2 = (tiles + (*G - tiles) - 1) / (*G - tiles) + 2;
    if (s2 > *segmax) *segmax = s2;
  }
}
static bool tn_pair_ok(int NI, int NJ) { return NJ > 128 && NI % (2 * BM) == 0 && pair_enabled(); }
size_t tn_workspace_bytes(int NI, int NJ, int M_host, int ngroups) {
  int G, segmax;
  tn_plan(NI, NJ, M_host, ngroups, &G, &segmax);
  size_t need = (size_t)G * segmax * BM * (NJ > 128 ? 256 : 128) * sizeof(float);
  if (tn_pair_ok(NI, NJ)) {                                  // the pair plan: two half-tiles per slot
    tn_plan(NI, NJ, M_host, ngroups, &G, &segmax, 2);
    const size_t n2 = (size_t)G * segmax * 2 * BM * 256 * sizeof(float);
    if (n2 > need) need = n2;
  }
  return need;
}
void set_tn_workspace(void* p, size_t bytes) { g_tn_ws.p = p; g_tn_ws.bytes = bytes; }

// C[NI,NJ] += P^T . Q; operands as planes (column windows p_col0 / q_col0) or as fp32 rows converted in-kernel
static int tn_impl(int passes, const Planes* P, int p_col0, const Planes* Q, int q_col0, const RowSrc* Ps, const RowSrc* Qs, float* C, int ldc,
                   int M_host, const int* M_dev, int NI, int NJ, cudaStream_t st, const GemmGrp* grp = nullptr) {
  if (M_host <= 0 || NI <= 0 || NJ <= 0) return 0;
  GemmGrp gg;
  if (grp) gg = *grp;
  else {
    gg.rows = grp_single(M_host, M_dev);
    for (int i = 0; i < kMaxGroups; ++i) { gg.wlev[i] = 0; gg.bias[i] = nullptr; gg.cout[i] = nullptr; }
    gg.cout[0] = C;
  }
  const int ng = gg.rows.n;
  const CUtensorMap *ma = nullptr, *mb = nullptr;
  Operand oa, ob;
  oa.row0 = 0; oa.col0 = p_col0; oa.plane_rows = 0;
  ob.row0 = 0; ob.col0 = q_col0; ob.plane_rows = 0;
  RowSrc none = row_src(nullptr, 0, 0);
  if (P) {
    PATHS_REQUIRE(P->p && Q && Q->p && p_col0 % 8 == 0 && q_col0 % 8 == 0 && M_host <= P->rows && M_host <= Q->rows,
                  "gemm_tn_tc: bad activation planes");
    if (int r = cached_map(&ma, P->p, 2 * P->rows, p_col0 + NI, P->ldp)) return r;
    if (int r = cached_map(&mb, Q->p, 2 * Q->rows, q_col0 + NJ, Q->ldp)) return r;
    oa.plane_rows = P->rows; ob.plane_rows = Q->rows;
  }
  const bool big = NJ > 128;
  const int BN = big ? 256 : 128;
  const int tiles = ng * ((NI + BM - 1) / BM) * ((NJ + BN - 1) / BN);
  int G, segmax;
  // CTA pairs when both operands are TMA-fed planes and the output rows tile by 256 (the LSTM block's weight gradients)
  const bool pair = P && tn_pair_ok(NI, NJ) && (long)tiles * ((M_host + BK - 1) / BK) / 8 >= pair_min_units();
  tn_plan(NI, NJ, M_host, ng, &G, &segmax, pair ? 2 : 1);
  const size_t need = (size_t)G * segmax * (pair ? 2 : 1) * BM * BN * sizeof(float);
  PATHS_REQUIRE(g_tn_ws.p && need <= g_tn_ws.bytes, "gemm_tn_tc: stream-K workspace too small: need %zu bytes, have %zu", need, g_tn_ws.bytes);
  float* ws = static_cast<float*>(g_tn_ws.p);
  if (pair) {
    int r = passes == 3 ? launch_pair<2, true, true, true>(ma, mb, oa, ob, ws, 0, gg, 0, NI, NJ, M_host, 0, G, segmax, st)
                        : launch_pair<1, true, true, true>(ma, mb, oa, ob, ws, 0, gg, 0, NI, NJ, M_host, 0, G, segmax, st);
    if (r) return r;
    if (4 * tiles >= 2 * G)
      pdl(tn_reduce_kernel<256, 1, 2>, dim3(tiles, BM * 256 / 4 / 256), 256, 0, st)(reinterpret_cast<const float4*>(ws), gg, ldc, NI, NJ, G, segmax);
    else
      pdl(tn_reduce_kernel<256, 8, 2>, dim3(tiles, BM * 256 / 4 / 32), 256, 0, st)(reinterpret_cast<const float4*>(ws), gg, ldc, NI, NJ, G, segmax);
    PATHS_LAUNCH_CK();
    return 0;
  }
#define TN_CASE(BN_, NP_)                                                                                                              \
  do {                                                                                                                                 \
    int r;                                                                                                                             \
    if (P) r = launch_tma<BN_, NP_, true, true, true, false, false>(ma, mb, oa, ob, none, none, ws, 0, gg, 0, NI, NJ, M_host, 0, G,    \
                                                                    segmax, st);                                                     \
    else r = launch_tma<BN_, NP_, true, true, true, true, true>(nullptr, nullptr, oa, ob, *Ps, *Qs, ws, 0, gg, 0, NI, NJ, M_host, 0,  \
                                                                G, segmax, st);                                                      \
    if (r) return r;                                                                                                                   \
    if (4 * tiles >= G)                                                                                                                \
      pdl(tn_reduce_kernel<BN_, 1>, dim3(tiles, BM * BN_ / 4 / 256), 256, 0, st)(reinterpret_cast<const float4*>(ws), gg, ldc, NI, NJ, \
                                                                               G, segmax);                                            \
    else                                                                                                                               \
      pdl(tn_reduce_kernel<BN_, 8>, dim3(tiles, BM * BN_ / 4 / 32), 256, 0, st)(reinterpret_cast<const float4*>(ws), gg, ldc, NI, NJ,  \
                                                                              G, segmax);                                             \
    PATHS_LAUNCH_CK();                                                                                                                 \
    return 0;                                                                                                                          \
  } while (0)
  if (passes == 3) { if (big) TN_CASE(256, 2); else TN_CASE(128, 2); }
  else             { if (big) TN_CASE(256, 1); else TN_CASE(128, 1); }
#undef TN_CASE
  return 0;
}

int gemm_tn_tc_p(int passes, const Planes& P, int p_col0, const Planes& Q, int q_col0, float* C, int ldc, int M_host, const int* M_dev, int NI,
                 int NJ, cudaStream_t st) {
  return tn_impl(passes, &P, p_col0, &Q, q_col0, nullptr, nullptr, C, ldc, M_host, M_dev, NI, NJ, st);
}

int split_rows_grp(const float* A, int lda, int K, const Planes& dst, const Grp& grp, bool want_lo, const MutL* colsum, cudaStream_t st) {
  PATHS_REQUIRE(colsum == nullptr || K <= 512, "split_rows_grp: column sums need K <= 512 (K %d)", K);
  PATHS_REQUIRE(dst.p && K % 4 == 0 && K <= dst.ldp && lda % 4 == 0 && (reinterpret_cast<uintptr_t>(A) & 15) == 0,
                "split_rows_grp: bad operands (K %d, ldp %d, lda %d)", K, dst.ldp, lda);
  const int cap = grp_max_rows(grp);
  if (cap <= 0) return 0;
  int g = (cap + 64 + 7) / 8;
  const int lim = grp.n > 1 ? 4 * kNumSMs : 16 * kNumSMs;
  if (g > lim) g = lim;
  if (colsum && g > kNumSMs) g = kNumSMs;                    // fewer blocks: fewer same-address atomics at the end
  const MutL cs = colsum ? *colsum : mutl1(nullptr);
  const dim3 grid(g, grp.n);
  if (K <= 128) pdl(split_planes_grp_kernel<1, 4>, grid, 256, 0, st)(A, lda, K, dst, grp, want_lo ? 1 : 0, cs);
  else if (K <= 256) pdl(split_planes_grp_kernel<2, 2>, grid, 256, 0, st)(A, lda, K, dst, grp, want_lo ? 1 : 0, cs);
  else if (K <= 384) pdl(split_planes_grp_kernel<3, 2>, grid, 256, 0, st)(A, lda, K, dst, grp, want_lo ? 1 : 0, cs);
  else pdl(split_planes_grp_kernel<4, 1>, grid, 256, 0, st)(A, lda, K, dst, grp, want_lo ? 1 : 0, cs);
  PATHS_LAUNCH_CK();
  return 0;
}
// grouped contractions on pre-split operand planes (rows of the planes = global rows of the concatenated buffers)
int gemm_nt_tc_grp_p(int passes, const Planes& A, const float* B, int ldb, int b_kn, float* C, int ldc, const GemmGrp& gg, int N, int K, int flags,
                     cudaStream_t st) {
  const int cap = grp_max_rows(gg.rows);
  if (cap <= 0 || N <= 0) return 0;
  return nt_impl(passes, &A, 0, nullptr, B, ldb, b_kn, C, ldc, nullptr, cap, nullptr, N, K, flags, st, &gg);
}
int gemm_tn_tc_grp_p(int passes, const Planes& P, const Planes& Q, int ldc, const GemmGrp& gg, int NI, int NJ, cudaStream_t st) {
  const int cap = grp_max_rows(gg.rows);
  if (cap <= 0 || NI <= 0 || NJ <= 0) return 0;
  return tn_impl(passes, &P, 0, &Q, 0, nullptr, nullptr, nullptr, ldc, cap, nullptr, NI, NJ, st, &gg);
}

// ---- grouped forms: one launch for the same contraction of several row groups (levels) of concatenated activation buffers.
// A / P / Q are addressed with GLOBAL rows of the concatenated buffer; B is the first member of a stacked packed weight.
bool gemm_tc_supported_nt_grp(const RowSrc& A, const float* B, int ldb, int N, int K) {
  if (!cvt_ok(A) || K % 8 != 0) return false;
  std::lock_guard<std::mutex> lk(g_mu);
  int r0, c0;
  const Packed* pk = find_packed(B, ldb, &r0, &c0);
  return pk != nullptr;
}
int gemm_nt_tc_grp(int passes, const RowSrc& A, const float* B, int ldb, int b_kn, float* C, int ldc, const GemmGrp& gg, int N, int K, int flags,
                   cudaStream_t st) {
  const int cap = grp_max_rows(gg.rows);
  if (cap <= 0 || N <= 0) return 0;
  PATHS_REQUIRE(cvt_ok(A), "gemm_nt_tc_grp: activation rows must be 16-byte aligned");
  return nt_impl(passes, nullptr, 0, &A, B, ldb, b_kn, C, ldc, nullptr, cap, nullptr, N, K, flags, st, &gg);
}
int gemm_tn_tc_grp(int passes, const float* P, int ldp, const RowSrc& Q, int ldc, const GemmGrp& gg, int NI, int NJ, cudaStream_t st) {
  const int cap = grp_max_rows(gg.rows);
  if (cap <= 0 || NI <= 0 || NJ <= 0) return 0;
  const RowSrc Ps = row_src(P, ldp, NI);
  PATHS_REQUIRE(cvt_ok(Ps) && cvt_ok(Q), "gemm_tn_tc_grp: activation rows must be 16-byte aligned");
  return tn_impl(passes, nullptr, 0, nullptr, 0, &Ps, &Q, nullptr, ldc, cap, nullptr, NI, NJ, st, &gg);
}

int gemm_tn_tc(int passes, const float* P, int ldp, const RowSrc& Q, float* C, int ldc, int M_host, const int* M_dev, int NI, int NJ,
               cudaStream_t st) {
  if (M_host <= 0 || NI <= 0 || NJ <= 0) return 0;
  const RowSrc Ps = row_src(P, ldp, NI);
  if (cvt_ok(Ps) && cvt_ok(Q)) return tn_impl(passes, nullptr, 0, nullptr, 0, &Ps, &Q, C, ldc, M_host, M_dev, NI, NJ, st);
  Planes pp, pq;
  if (int r = split_to_scratch(0, Ps, M_host, M_dev, NI, passes == 3, &pp, st)) return r;
  if (int r = split_to_scratch(1, Q, M_host, M_dev, NJ, passes == 3, &pq, st)) return r;
  return gemm_tn_tc_p(passes, pp, 0, pq, 0, C, ldc, M_host, M_dev, NI, NJ, st);
}

}  // namespace paths
