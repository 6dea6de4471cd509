// Shared device/host helpers for the PATHS B200 kernels (sm_100a only).
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

namespace paths {

void set_error(const char* fmt, ...);          // records a message for paths_last_error()

#define PATHS_CK(expr)                                                                           \
  do {                                                                                           \
    cudaError_t _e = (expr);                                                                     \
    if (_e != cudaSuccess) {                                                                     \
      ::paths::set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(_e));  \
      return -1;                                                                                 \
    }                                                                                            \
  } while (0)

#define PATHS_LAUNCH_CK() PATHS_CK(cudaGetLastError())

#define PATHS_REQUIRE(cond, ...)                 \
  do {                                           \
    if (!(cond)) {                               \
      ::paths::set_error(__VA_ARGS__);           \
      return -2;                                 \
    }                                            \
  } while (0)

constexpr int kNumSMs = 148;                     // B200

// ---- programmatic dependent launch ---------------------------------------------------------------------------------------
// A step is ~270 dependent launches, most of them a few tens of microseconds long: the drain of one grid and the fill of the
// next are a measurable share of it.  Every kernel of the library is launched with the programmatic-stream-serialization
// attribute and starts with pdl_grid_sync(): `launch_dependents` lets the NEXT kernel of the stream become resident as soon as
// every CTA of this one has started (its CTAs fill the SMs while ours drain), `wait` blocks until the PREVIOUS kernel has
// completed and flushed its memory -- so no kernel touches global memory before its predecessor is done, and completion stays
// transitive because every CTA of every kernel executes the wait before anything else.  Opt-in: PATHS_B200_PDL=1 (without the
// launch attribute griddepcontrol is a no-op) -- on the multi-stream step it measured slower, see DESIGN.md §3.7b.
#ifdef __CUDACC__
__device__ __forceinline__ void pdl_grid_sync() {
#ifdef PATHS_PDL_EARLY                                    // early trigger: dependents become resident while this grid still runs
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
#endif                                                    // default: implicit trigger as each CTA exits (dependents fill the tail only)
  asm volatile("griddepcontrol.wait;" ::: "memory");
}
#endif
bool pdl_enabled();
template <typename... KA>
struct PdlLaunch {
  void (*kern)(KA...); dim3 grid, block; size_t smem; cudaStream_t st;
  template <typename... A>
  void operator()(A&&... a) const {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute at;
    at.id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at.val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = &at; cfg.numAttrs = pdl_enabled() ? 1 : 0;
    cudaLaunchKernelEx(&cfg, kern, KA(static_cast<A&&>(a))...);        // errors surface through PATHS_LAUNCH_CK()
  }
};
template <typename... KA>
inline PdlLaunch<KA...> pdl(void (*kern)(KA...), dim3 grid, dim3 block, size_t smem, cudaStream_t st) {
  return PdlLaunch<KA...>{kern, grid, block, smem, st};
}

// A-operand row source: logical row m of a [M, K] matrix whose columns [0,k0) come from table p0 and
// columns [k0,K) from table p1, each optionally through a row-index array (idx < 0 => zero row).
// This is how [x || h_parent] (model/interface.py:49) and packed-token gathers reach a GEMM without
// a materialised concat/gather.
struct RowSrc {
  const float* p0; int ld0; const int* idx0; int k0;
  const float* p1; int ld1; const int* idx1;
};

inline RowSrc row_src(const float* p, int ld, int K, const int* idx = nullptr) {
  RowSrc r; r.p0 = p; r.ld0 = ld; r.idx0 = idx; r.k0 = K; r.p1 = nullptr; r.ld1 = 0; r.idx1 = nullptr; return r;
}
inline RowSrc row_src2(const float* p0, int ld0, const int* idx0, int k0, const float* p1, int ld1, const int* idx1) {
  RowSrc r; r.p0 = p0; r.ld0 = ld0; r.idx0 = idx0; r.k0 = k0; r.p1 = p1; r.ld1 = ld1; r.idx1 = idx1; return r;
}

// Returns pointer to 4 consecutive floats (m, k..k+3) or nullptr for a zero row.  k must not straddle k0.
__device__ __forceinline__ const float* rowsrc_ptr(const RowSrc& s, int m, int k) {
  if (k < s.k0) {
    int r = s.idx0 ? s.idx0[m] : m;
    return r < 0 ? nullptr : s.p0 + (size_t)r * s.ld0 + k;
  }
  int r = s.idx1 ? s.idx1[m] : m;
  return r < 0 ? nullptr : s.p1 + (size_t)r * s.ld1 + (k - s.k0);
}

// Tensor-core operand storage of an activation: bf16 planes stacked as [2 * rows, ldp] -- rows [0, rows) hold hi = bf16(x),
// rows [rows, 2 rows) hold lo = bf16(x - hi).  Written by the kernel that produces the activation (or by split_rows) and read
// by TMA as 64x64 SWIZZLE_128B boxes.  `rows` is the padded capacity: live rows are followed by zero rows up to the next
// multiple of 64 so a reduction over rows (weight gradients) may run whole 64-row blocks.
struct Planes { __nv_bfloat16* p; int rows; int ldp; };
inline int planes_pad_rows(int M) { return ((M + 63) / 64) * 64 + 64; }
inline size_t planes_bytes(int M_cap, int K) { return (size_t)2 * planes_pad_rows(M_cap) * ((K + 7) / 8 * 8) * sizeof(__nv_bfloat16) + 256; }
inline Planes planes_at(void* mem, int M_cap, int K) {
  Planes pl;
  pl.p = reinterpret_cast<__nv_bfloat16*>((reinterpret_cast<uintptr_t>(mem) + 255) & ~uintptr_t(255));
  pl.rows = planes_pad_rows(M_cap); pl.ldp = (K + 7) / 8 * 8;
  return pl;
}
inline Planes no_planes() { Planes pl; pl.p = nullptr; pl.rows = 0; pl.ldp = 0; return pl; }

__device__ __forceinline__ void planes_store4(const Planes& pl, int m, int k, float4 v, bool want_lo) {
  __nv_bfloat162 h0 = __floats2bfloat162_rn(v.x, v.y), h1 = __floats2bfloat162_rn(v.z, v.w);
  uint2 hv;
  hv.x = *reinterpret_cast<uint32_t*>(&h0); hv.y = *reinterpret_cast<uint32_t*>(&h1);
  *reinterpret_cast<uint2*>(pl.p + (size_t)m * pl.ldp + k) = hv;
  if (want_lo) {
    const float2 f0 = __bfloat1622float2(h0), f1 = __bfloat1622float2(h1);
    __nv_bfloat162 l0 = __floats2bfloat162_rn(v.x - f0.x, v.y - f0.y), l1 = __floats2bfloat162_rn(v.z - f1.x, v.w - f1.y);
    uint2 lv;
    lv.x = *reinterpret_cast<uint32_t*>(&l0); lv.y = *reinterpret_cast<uint32_t*>(&l1);
    *reinterpret_cast<uint2*>(pl.p + ((size_t)pl.rows + m) * pl.ldp + k) = lv;
  }
}
// number of rows a producer must cover: live rows plus the zero rows up to the next multiple of 64
__device__ __forceinline__ int planes_rows_to_write(const Planes& pl, int M) {
  int Mpad = (M + 63) & ~63;
  return Mpad > pl.rows ? pl.rows : Mpad;
}

__device__ __forceinline__ float sigmoidf_(float x) { return 1.0f / (1.0f + expf(-x)); }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// ---- row groups --------------------------------------------------------------------------------------------------------
// The aggregators (decoders) of the L magnification levels are independent problems with the same shapes but their own weights
// and their own row counts.  Their token-shaped buffers are laid out back to back in one allocation (level l owns rows
// [roff[l], roff[l] + capacity_l), roff a multiple of 128) so that ONE launch can serve every level: a kernel receives the
// buffer's base pointer plus a Grp describing which row regions are live, and per-group parameter pointers (PtrL / MutL).
constexpr int kMaxGroups = 8;        // == PATHS_MAX_LEVELS
struct Grp {
  int n;                             // number of groups
  int roff[kMaxGroups];              // first row of each group in the concatenated buffer
  int hcnt[kMaxGroups];              // host-side row count (capacity when cnt[g] != nullptr)
  const int* cnt[kMaxGroups];        // live row count on the device, or nullptr => hcnt
};
struct PtrL { const float* p[kMaxGroups]; };
struct MutL { float* p[kMaxGroups]; };
struct IdxL { const int* p[kMaxGroups]; };
// Select entry g of a per-group table held in kernel parameters with compile-time indices only: a dynamically indexed
// parameter array would be copied to local memory by every thread at kernel entry.
template <typename T>
__host__ __device__ __forceinline__ T pick(const T (&a)[kMaxGroups], int g) {
  T r = a[0];
#pragma unroll
  for (int k = 1; k < kMaxGroups; ++k) r = (g == k) ? a[k] : r;
  return r;
}
__host__ __device__ __forceinline__ int grp_count(const Grp& g, int i) {
#ifdef __CUDA_ARCH__
  const int* c = pick(g.cnt, i);
  return c ? *c : pick(g.hcnt, i);
#else
  return g.hcnt[i];
#endif
}
inline Grp grp_single(int M_host, const int* M_dev, int roff = 0) {
  Grp g; g.n = 1;
  for (int i = 0; i < kMaxGroups; ++i) { g.roff[i] = 0; g.hcnt[i] = 0; g.cnt[i] = nullptr; }
  g.roff[0] = roff; g.hcnt[0] = M_host; g.cnt[0] = M_dev;
  return g;
}
inline int grp_max_rows(const Grp& g) { int m = 0; for (int i = 0; i < g.n; ++i) if (g.hcnt[i] > m) m = g.hcnt[i]; return m; }
inline PtrL ptrl1(const float* p) { PtrL l; for (int i = 0; i < kMaxGroups; ++i) l.p[i] = nullptr; l.p[0] = p; return l; }
inline MutL mutl1(float* p) { MutL l; for (int i = 0; i < kMaxGroups; ++i) l.p[i] = nullptr; l.p[0] = p; return l; }
inline IdxL idxl1(const int* p) { IdxL l; for (int i = 0; i < kMaxGroups; ++i) l.p[i] = nullptr; l.p[0] = p; return l; }

// Grouped contraction: group g reads/writes rows of the concatenated activation buffers per `rows`, contracts against the
// g-th member of a stacked packed weight (wlev[g] = index inside the stack), adds bias[g]; the weight-gradient form (TN)
// reduces over the group's rows into cout[g].
// RowAdd: rows gathered into the NT epilogue before bias / activation: C[i, :] = act(acc + add[idx[i], :] + bias), idx[i] < 0 = none.
// What the gate contraction of a child level uses for the h_parent half of [x || h_parent] . W^T: the four children of a patch
// share that product, so it is computed once per selected parent and added here (interface.py:49).
struct RowAdd { const float* p; const int* idx; int ld; };
inline RowAdd no_row_add() { RowAdd a; a.p = nullptr; a.idx = nullptr; a.ld = 0; return a; }
struct GemmGrp {
  Grp rows;
  int wlev[kMaxGroups];
  const float* bias[kMaxGroups];
  float* cout[kMaxGroups];
  RowAdd add{nullptr, nullptr, 0};
};

// ---- dropout (nn.Transformer(dropout=p) in training mode, model/aggregator.py:25-33) ------------------------------------
// torch's generator stream cannot be reproduced, so keep/drop decisions come from a counter-based hash: an element is kept
// iff fmix64(site_key ^ (row << 32 | col)) >> 32 >= p * 2^32, where site_key mixes the step's seed with (level, layer, site),
// `row` is the token's packed row inside its level and `col` the feature column (attention probabilities: head << 24 | key).
// Forward and backward recompute the same decision; oracle/paths_oracle.py restates the hash to build explicit masks.
enum DropSite { DROP_ATTN = 0, DROP_SA_OUT = 1, DROP_CROSS_OUT = 2, DROP_FFN_INNER = 3, DROP_FFN_OUT = 4 };
__host__ __device__ __forceinline__ uint64_t fmix64(uint64_t x) {
  x ^= x >> 33; x *= 0xff51afd7ed558ccdULL; x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ULL; x ^= x >> 33;
  return x;
}
inline uint64_t drop_site_key(uint64_t seed, int level, int layer, int site) {
  return fmix64(seed ^ (0x9E3779B97F4A7C15ULL * (uint64_t)((level * 8 + layer) * 8 + site + 1)));
}
__device__ __forceinline__ bool drop_keep(uint64_t key, uint32_t row, uint32_t col, uint32_t thresh) {
  return (uint32_t)(fmix64(key ^ (((uint64_t)row << 32) | col)) >> 32) >= thresh;
}
struct Drop {
  uint32_t thresh;                   // 0: dropout off
  float inv_keep;                    // 1 / (1 - p)
  uint64_t key[kMaxGroups];          // site key of each row group (level)
};
inline Drop no_drop() { Drop d; d.thresh = 0; d.inv_keep = 1.f; for (int i = 0; i < kMaxGroups; ++i) d.key[i] = 0; return d; }
inline uint32_t drop_thresh(float p) {
  const double t = (double)p * 4294967296.0;
  return t <= 0.0 ? 0u : t >= 4294967295.0 ? 4294967295u : (uint32_t)t;
}

// GEMM_GATES: the output is the LSTM gate pre-activation block [f | r | m | o] (widths C, C, C, D; C = flags >> 8): the
// epilogue applies sigmoid / sigmoid / tanh / sigmoid by column range, so the gates reach HBM already activated.
enum GemmFlags { GEMM_RELU = 1, GEMM_ACCUM = 2, GEMM_GATES = 4 };

// Precision modes of the dense contractions.
//   0: fp32 SIMT FFMA (exact fp32 semantics, the parity anchor)
//   1: tcgen05 bf16x3 split (hi*hi + hi*lo + lo*hi, fp32 accumulate in TMEM; ~2^-17 relative)
//   2: tcgen05 bf16 single pass (fast mode, stated 1e-2)
enum Precision { PREC_FP32_SIMT = 0, PREC_BF16X3_TC = 1, PREC_BF16_TC = 2 };

// ---- dense contraction entry points (gemm_simt.cu / gemm_tc.cu) -------------------------------------------
// C[m, n] (+)= act( sum_k A(m,k) * B(n,k) + bias[n] ),  m < M (M read from *M_dev when M_dev != nullptr, else M_host)
// b_kn == 0: B is [N, K] row-major with leading dim ldb (weight as stored by nn.Linear)
// b_kn == 1: B is [K, N] row-major with leading dim ldb (used for dX = dY . W)
int gemm_nt(int prec, const RowSrc& A, const float* B, int ldb, int b_kn, float* C, int ldc, const float* bias,
            int M_host, const int* M_dev, int N, int K, int flags, cudaStream_t st);

// C[i, j] += sum_m P[m, i] * Q(m, j)   (dW = dY^T . X).  P is [M, NI] with leading dim ldp, Q a RowSrc with NJ columns.
// Accumulates with fp32 atomics (split over m); C must be initialised (zeroed grads).
int gemm_tn(int prec, const float* P, int ldp, const RowSrc& Q, float* C, int ldc, int M_host, const int* M_dev,
            int NI, int NJ, cudaStream_t st);

// An activation as the contraction wrappers see it: fp32 rows (f, ld) for the SIMT mode and/or bf16 planes for the tensor-core
// modes; col0 selects a column window of either.  A null `pl.p` in a tensor-core mode means "split f on the fly".
struct Act { const float* f; int ld; Planes pl; int col0; };
inline Act act(const float* f, int ld, const Planes& pl, int col0 = 0) { Act a; a.f = f; a.ld = ld; a.pl = pl; a.col0 = col0; return a; }
inline Act act_f(const float* f, int ld) { return act(f, ld, no_planes(), 0); }

// out[n] += sum_m X[m, n]
int colsum_accum(const float* X, int ld, float* out, int M_host, const int* M_dev, int N, cudaStream_t st);

}  // namespace paths
