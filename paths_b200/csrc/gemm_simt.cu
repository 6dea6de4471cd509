// fp32 SIMT (FFMA) contractions.  This is the exact-fp32 engine: it anchors parity (precision mode 0) and is
// the on-device reference the tcgen05 path (gemm_tc.cu) is checked against.  128x128x16 tiles, 256 threads,
// 8x8 outputs per thread, register-prefetch double buffering.
#include "common.cuh"

namespace paths {

namespace {

constexpr int BM = 128, BN = 128, BK = 16, PAD = 4, NT = 256;

__device__ __forceinline__ void mma_tile(const float (*As)[BM + PAD], const float (*Bs)[BN + PAD], int ty, int tx,
                                         float (&acc)[8][8]) {
#pragma unroll
  for (int k = 0; k < BK; ++k) {
    float4 a0 = *reinterpret_cast<const float4*>(&As[k][ty * 4]);
    float4 a1 = *reinterpret_cast<const float4*>(&As[k][64 + ty * 4]);
    float4 b0 = *reinterpret_cast<const float4*>(&Bs[k][tx * 4]);
    float4 b1 = *reinterpret_cast<const float4*>(&Bs[k][64 + tx * 4]);
    float a[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
    float b[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
  }
}

// ---- C = A . B^T ---------------------------------------------------------------------------------------
template <bool B_KN>
__global__ void __launch_bounds__(NT) gemm_nt_kernel(RowSrc A, const float* __restrict__ B, int ldb,
                                                     float* __restrict__ C, int ldc, const float* __restrict__ bias,
                                                     int M_host, const int* __restrict__ M_dev, int N, int K, int flags) {
  pdl_grid_sync();
  __shared__ __align__(16) float As[2][BK][BM + PAD];
  __shared__ __align__(16) float Bs[2][BK][BN + PAD];
  const int M = M_dev ? *M_dev : M_host;
  const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;
  if (m0 >= M) return;
  const int t = threadIdx.x, ty = t / 16, tx = t % 16;

  float4 ra[2], rb[2];
  auto load_tiles = [&](int kb) {
    const int k = kb * BK + (t % 4) * 4;
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      const int row = t / 4 + 64 * i;
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (m0 + row < M && k < K) {
        const float* p = rowsrc_ptr(A, m0 + row, k);
        if (p) v = *reinterpret_cast<const float4*>(p);
      }
      ra[i] = v;
    }
    if (!B_KN) {
#pragma unroll
      for (int i = 0; i < 2; ++i) {
        const int row = t / 4 + 64 * i;
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (n0 + row < N && k < K) v = *reinterpret_cast<const float4*>(B + (size_t)(n0 + row) * ldb + k);
        rb[i] = v;
      }
    } else {
#pragma unroll
      for (int i = 0; i < 2; ++i) {
        const int kr = kb * BK + t / 32 + 8 * i, n = n0 + (t % 32) * 4;
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (kr < K && n < N) {
          const float* p = B + (size_t)kr * ldb + n;
          if (n + 3 < N) v = *reinterpret_cast<const float4*>(p);
          else { v.x = p[0]; if (n + 1 < N) v.y = p[1]; if (n + 2 < N) v.z = p[2]; }
        }
        rb[i] = v;
      }
    }
  };
  auto store_tiles = [&](int buf) {
    const int kq = (t % 4) * 4;
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      const int row = t / 4 + 64 * i;
      As[buf][kq + 0][row] = ra[i].x; As[buf][kq + 1][row] = ra[i].y;
      As[buf][kq + 2][row] = ra[i].z; As[buf][kq + 3][row] = ra[i].w;
    }
    if (!B_KN) {
#pragma unroll
      for (int i = 0; i < 2; ++i) {
        const int row = t / 4 + 64 * i;
        Bs[buf][kq + 0][row] = rb[i].x; Bs[buf][kq + 1][row] = rb[i].y;
        Bs[buf][kq + 2][row] = rb[i].z; Bs[buf][kq + 3][row] = rb[i].w;
      }
    } else {
#pragma unroll
      for (int i = 0; i < 2; ++i)
        *reinterpret_cast<float4*>(&Bs[buf][t / 32 + 8 * i][(t % 32) * 4]) = rb[i];
    }
  };

  float acc[8][8];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;

  const int nkb = (K + BK - 1) / BK;
  load_tiles(0);
  store_tiles(0);
  __syncthreads();
  for (int kb = 0; kb < nkb; ++kb) {
    const int buf = kb & 1;
    if (kb + 1 < nkb) load_tiles(kb + 1);
    mma_tile(As[buf], Bs[buf], ty, tx, acc);
    if (kb + 1 < nkb) store_tiles(buf ^ 1);
    __syncthreads();
  }

  const bool vec = (N % 4 == 0) && (ldc % 4 == 0);
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int m = m0 + (i < 4 ? ty * 4 + i : 64 + ty * 4 + (i - 4));
    if (m >= M) continue;
#pragma unroll
    for (int jq = 0; jq < 2; ++jq) {
      const int n = n0 + (jq == 0 ? tx * 4 : 64 + tx * 4);
      if (n >= N) continue;
      float v[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        float x = acc[i][jq * 4 + j];
        if (bias && n + j < N) x += bias[n + j];
        if (flags & GEMM_RELU) x = fmaxf(x, 0.f);
        v[j] = x;
      }
      float* cp = C + (size_t)m * ldc + n;
      if (vec && n + 3 < N) {
        float4 o = make_float4(v[0], v[1], v[2], v[3]);
        if (flags & GEMM_ACCUM) { float4 c = *reinterpret_cast<float4*>(cp); o.x += c.x; o.y += c.y; o.z += c.z; o.w += c.w; }
        *reinterpret_cast<float4*>(cp) = o;
      } else {
#pragma unroll
        for (int j = 0; j < 4; ++j)
          if (n + j < N) cp[j] = (flags & GEMM_ACCUM) ? cp[j] + v[j] : v[j];
      }
    }
  }
}

// ---- C += P^T . Q  (reduction over rows) -------------------------------------------------------------------
__global__ void __launch_bounds__(NT) gemm_tn_kernel(const float* __restrict__ P, int ldp, RowSrc Q, float* __restrict__ C,
                                                     int ldc, int M_host, const int* __restrict__ M_dev, int NI, int NJ, int pvec) {
  pdl_grid_sync();
  __shared__ __align__(16) float Ps[2][BK][BM + PAD];
  __shared__ __align__(16) float Qs[2][BK][BN + PAD];
  const int M = M_dev ? *M_dev : M_host;
  const int i0 = blockIdx.y * BM, j0 = blockIdx.x * BN;
  const int nsteps_total = (M + BK - 1) / BK;
  const int per = (nsteps_total + gridDim.z - 1) / gridDim.z;
  const int s_begin = blockIdx.z * per, s_end = min(nsteps_total, s_begin + per);
  if (s_begin >= s_end) return;
  const int t = threadIdx.x, ty = t / 16, tx = t % 16;

  float4 rp[2], rq[2];
  auto load_tiles = [&](int s) {
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      const int m = s * BK + t / 32 + 8 * r;
      const int i = i0 + (t % 32) * 4, j = j0 + (t % 32) * 4;
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f), w = v;
      if (m < M) {
        if (i < NI) {
          const float* p = P + (size_t)m * ldp + i;
          if (pvec && i + 3 < NI) v = *reinterpret_cast<const float4*>(p);
          else { v.x = p[0]; if (i + 1 < NI) v.y = p[1]; if (i + 2 < NI) v.z = p[2]; if (i + 3 < NI) v.w = p[3]; }
        }
        if (j < NJ) {
          const float* q = rowsrc_ptr(Q, m, j);
          if (q) {
            if (j + 3 < NJ) w = *reinterpret_cast<const float4*>(q);
            else { w.x = q[0]; if (j + 1 < NJ) w.y = q[1]; if (j + 2 < NJ) w.z = q[2]; }
          }
        }
      }
      rp[r] = v; rq[r] = w;
    }
  };
  auto store_tiles = [&](int buf) {
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      *reinterpret_cast<float4*>(&Ps[buf][t / 32 + 8 * r][(t % 32) * 4]) = rp[r];
      *reinterpret_cast<float4*>(&Qs[buf][t / 32 + 8 * r][(t % 32) * 4]) = rq[r];
    }
  };

  float acc[8][8];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;

  load_tiles(s_begin);
  store_tiles(0);
  __syncthreads();
  for (int s = s_begin; s < s_end; ++s) {
    const int buf = (s - s_begin) & 1;
    if (s + 1 < s_end) load_tiles(s + 1);
    mma_tile(Ps[buf], Qs[buf], ty, tx, acc);
    if (s + 1 < s_end) store_tiles(buf ^ 1);
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int ii = i0 + (i < 4 ? ty * 4 + i : 64 + ty * 4 + (i - 4));
    if (ii >= NI) continue;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int jj = j0 + (j < 4 ? tx * 4 + j : 64 + tx * 4 + (j - 4));
      if (jj < NJ) atomicAdd(C + (size_t)ii * ldc + jj, acc[i][j]);
    }
  }
}

// out[n] += sum_m X[m, n].  Vector path (N % 4 == 0, ld % 4 == 0, N <= 1024): a thread keeps one float4 column group and
// walks rows (independent loads, unrolled), one shared-memory combine and one atomic per column per block.
__global__ void __launch_bounds__(256) colsum_kernel(const float* __restrict__ X, int ld, MutL outl, Grp grp, int N) {
  pdl_grid_sync();
  const int M = grp_count(grp, blockIdx.y);                // blockIdx.y = row group of the concatenated buffer
  X += (size_t)pick(grp.roff, blockIdx.y) * ld;
  float* out = pick(outl.p, blockIdx.y);
  const int ngrp = N >> 2;
  const int rpb = 256 / ngrp > 0 ? 256 / ngrp : 1;
  const int g = threadIdx.x % ngrp, rsub = threadIdx.x / ngrp;
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
  if (rsub < rpb) {
#pragma unroll 8
    for (int m = blockIdx.x * rpb + rsub; m < M; m += gridDim.x * rpb) {
      const float4 v = __ldg(reinterpret_cast<const float4*>(X + (size_t)m * ld) + g);
      acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
    }
  }
  __shared__ float4 part[256];
  part[threadIdx.x] = acc;
  __syncthreads();
  if (threadIdx.x < ngrp) {
    float4 t = part[threadIdx.x];
    for (int k = 1; k < rpb; ++k) { const float4 u = part[k * ngrp + threadIdx.x]; t.x += u.x; t.y += u.y; t.z += u.z; t.w += u.w; }
    float* o = out + 4 * threadIdx.x;
    atomicAdd(o, t.x); atomicAdd(o + 1, t.y); atomicAdd(o + 2, t.z); atomicAdd(o + 3, t.w);
  }
}

// scalar fallback for odd widths (the 4-wide classifier bias)
__global__ void __launch_bounds__(256) colsum_scalar_kernel(const float* __restrict__ X, int ld, float* __restrict__ out,
                                                            int M_host, const int* __restrict__ M_dev, int N) {
  pdl_grid_sync();
  __shared__ float red[8][33];
  const int M = M_dev ? *M_dev : M_host;
  const int n = blockIdx.x * 32 + threadIdx.x % 32, y = threadIdx.x / 32;
  float s = 0.f;
  if (n < N)
    for (int m = blockIdx.y * 8 + y; m < M; m += gridDim.y * 8) s += X[(size_t)m * ld + n];
  red[y][threadIdx.x % 32] = s;
  __syncthreads();
  if (y == 0 && n < N) {
    float tot = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) tot += red[i][threadIdx.x];
    if (tot != 0.f) atomicAdd(out + n, tot);
  }
}

}  // namespace

int gemm_nt_simt(const RowSrc& A, const float* B, int ldb, int b_kn, float* C, int ldc, const float* bias, int M_host,
                 const int* M_dev, int N, int K, int flags, cudaStream_t st) {
  PATHS_REQUIRE(K % 4 == 0 && A.k0 % 4 == 0 && A.ld0 % 4 == 0 && (A.p1 == nullptr || A.ld1 % 4 == 0),
                "gemm_nt: K (%d), k0 (%d) and leading dims must be multiples of 4", K, A.k0);
  PATHS_REQUIRE(b_kn || ldb % 4 == 0, "gemm_nt: ldb must be a multiple of 4");
  if (M_host <= 0 || N <= 0) return 0;
  dim3 grid((N + BN - 1) / BN, (M_host + BM - 1) / BM);
  if (b_kn) pdl(gemm_nt_kernel<true>, grid, NT, 0, st)(A, B, ldb, C, ldc, bias, M_host, M_dev, N, K, flags);
  else pdl(gemm_nt_kernel<false>, grid, NT, 0, st)(A, B, ldb, C, ldc, bias, M_host, M_dev, N, K, flags);
  PATHS_LAUNCH_CK();
  return 0;
}

int gemm_tn_simt(const float* P, int ldp, const RowSrc& Q, float* C, int ldc, int M_host, const int* M_dev, int NI, int NJ,
                 cudaStream_t st) {
  PATHS_REQUIRE(Q.k0 % 4 == 0 && Q.ld0 % 4 == 0 && (Q.p1 == nullptr || Q.ld1 % 4 == 0),
                "gemm_tn: k0 and leading dims of Q must be multiples of 4");
  if (M_host <= 0 || NI <= 0 || NJ <= 0) return 0;
  const int tiles = ((NI + BM - 1) / BM) * ((NJ + BN - 1) / BN);
  int splits = (2 * kNumSMs + tiles - 1) / tiles;
  const int max_splits = (M_host + 4 * BK - 1) / (4 * BK);
  if (splits > max_splits) splits = max_splits;
  if (splits < 1) splits = 1;
  dim3 grid((NJ + BN - 1) / BN, (NI + BM - 1) / BM, splits);
  pdl(gemm_tn_kernel, grid, NT, 0, st)(P, ldp, Q, C, ldc, M_host, M_dev, NI, NJ, ldp % 4 == 0 ? 1 : 0);
  PATHS_LAUNCH_CK();
  return 0;
}

static bool colsum_vec_ok(const float* X, int ld, int N) {
  return N % 4 == 0 && ld % 4 == 0 && N <= 1024 && N >= 8 && (reinterpret_cast<uintptr_t>(X) & 15) == 0;
}
int colsum_accum_grp(const float* X, int ld, const MutL& out, const Grp& grp, int N, cudaStream_t st) {
  const int cap = grp_max_rows(grp);
  if (cap <= 0 || N <= 0) return 0;
  PATHS_REQUIRE(colsum_vec_ok(X, ld, N), "colsum_accum_grp: width %d / leading dim %d unsupported", N, ld);
  const int rpb = 256 / (N / 4) > 0 ? 256 / (N / 4) : 1;
  int g = (cap + 8 * rpb - 1) / (8 * rpb);                // few blocks: the final same-address atomics serialise in L2
  if (g > 96) g = 96;
  if (g < 1) g = 1;
  pdl(colsum_kernel, dim3(g, grp.n), 256, 0, st)(X, ld, out, grp, N);
  PATHS_LAUNCH_CK();
  return 0;
}
int colsum_accum(const float* X, int ld, float* out, int M_host, const int* M_dev, int N, cudaStream_t st) {
  if (M_host <= 0 || N <= 0) return 0;
  if (colsum_vec_ok(X, ld, N)) {
    return colsum_accum_grp(X, ld, mutl1(out), grp_single(M_host, M_dev), N, st);
  } else {
    int splits = (M_host + 255) / 256;
    if (splits > 64) splits = 64;
    dim3 grid((N + 31) / 32, splits);
    pdl(colsum_scalar_kernel, grid, 256, 0, st)(X, ld, out, M_host, M_dev, N);
  }
  PATHS_LAUNCH_CK();
  return 0;
}

}  // namespace paths
