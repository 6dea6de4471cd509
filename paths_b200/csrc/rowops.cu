// Row-wise (HBM-bound) kernels of the PATHS level: feature gather, LSTM gate math, importance + positional
// encoding, residual LayerNorms, the classifier head and the survival loss -- forward and backward.
// One warp per packed row, lanes striding the feature dimension (coalesced 128 B per access).
#include "common.cuh"
#include "rowops.cuh"

namespace paths {

namespace {

constexpr int WPB = 8;                 // warps per block
constexpr int TPB = WPB * 32;

__device__ __forceinline__ int row_count(int M_host, const int* M_dev) { return M_dev ? *M_dev : M_host; }

#define ROW_LOOP(m, M)                                                                  \
  const int lane = threadIdx.x & 31;                                                    \
  for (int m = blockIdx.x * WPB + (threadIdx.x >> 5); m < (M); m += gridDim.x * WPB)

// ---------------------------------------------------------------------------------------------------------
// gather: X[m,:] = grid(slide, level)[x*H + y, :]         (data_utils/slide.py:331 `fts[new_locs[:,0], new_locs[:,1]]`)
// BF16: the store holds the grids as bfloat16 (PyramidStore(dtype=torch.bfloat16): half the HBM footprint and gather traffic;
// the features are rounded once, at load) -- rows are widened to fp32 on the way into X.
template <bool BF16>
__global__ void __launch_bounds__(TPB) gather_rows_kernel(const float* const* __restrict__ grid_ptr, const int* __restrict__ gh,
                                                          int level, int L, const int* __restrict__ slide_ids,
                                                          const int* __restrict__ row_slide, const int* __restrict__ row_xy,
                                                          float* __restrict__ X, int D, int M_host, const int* M_dev, Planes XP,
                                                          int want_lo) {
  pdl_grid_sync();
  const int M = row_count(M_host, M_dev);
  // XP (optional): the rows also leave as the bf16 (hi, lo) operand planes of the gate contraction, with the zero rows up to the
  // next multiple of 64 the weight-gradient form reads -- the split pass over X then never runs
  const int Mw = XP.p ? planes_rows_to_write(XP, M) : M;
  ROW_LOOP(m, Mw) {
    if (m >= M) {
      for (int c = lane; c < D / 4; c += 32) planes_store4(XP, m, 4 * c, make_float4(0.f, 0.f, 0.f, 0.f), want_lo != 0);
      continue;
    }
    const int s = row_slide[m];
    const int sid = slide_ids ? slide_ids[s] : s;
    const size_t cell = (size_t)row_xy[2 * m] * gh[(size_t)sid * L + level] + row_xy[2 * m + 1];
    float4* dst = reinterpret_cast<float4*>(X + (size_t)m * D);
    if (BF16) {
      const uint2* s2 = reinterpret_cast<const uint2*>(reinterpret_cast<const __nv_bfloat16*>(grid_ptr[(size_t)sid * L + level]) + cell * D);
      for (int c = lane; c < D / 4; c += 32) {
        const uint2 v = __ldg(s2 + c);
        const float2 a = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&v.x));
        const float2 b = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&v.y));
        const float4 x = make_float4(a.x, a.y, b.x, b.y);
        dst[c] = x;
        if (XP.p) planes_store4(XP, m, 4 * c, x, want_lo != 0);
      }
    } else {
      const float4* s4 = reinterpret_cast<const float4*>(grid_ptr[(size_t)sid * L + level] + cell * D);
      for (int c = lane; c < D / 4; c += 32) {
        const float4 x = __ldg(s4 + c);
        dst[c] = x;
        if (XP.p) planes_store4(XP, m, 4 * c, x, want_lo != 0);
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------------------
// LSTM cell, model/interface.py:49-56.  G holds the pre-activations [f | r | m | o] (+bias) of one row.
// Every kernel below can emit its outputs as fp32 rows (SIMT mode / pointwise consumers) and/or as bf16 (hi, lo) planes
// (the operand format of the tensor-core contractions), so no separate split pass over the activations is needed.
__device__ __forceinline__ float4 ld4(const float* p) { return *reinterpret_cast<const float4*>(p); }
__device__ __forceinline__ void st4(float* p, float4 v) { *reinterpret_cast<float4*>(p) = v; }

__global__ void __launch_bounds__(TPB) lstm_point1_fwd_kernel(float* __restrict__ G, const float* __restrict__ Cprev,
                                                              const int* __restrict__ parent, float* __restrict__ Cn, Planes CnP, int want_lo,
                                                              int activated, int D, int C, int M_host, const int* M_dev) {
  pdl_grid_sync();
  const int M = row_count(M_host, M_dev);
  const int ldg = 3 * C + D;
  const int Mw = CnP.p ? planes_rows_to_write(CnP, M) : M;
  ROW_LOOP(m, Mw) {
    if (m >= M) {                                            // zero rows behind the live ones (planes only)
      for (int c4 = lane; c4 < C / 4; c4 += 32) planes_store4(CnP, m, 4 * c4, make_float4(0.f, 0.f, 0.f, 0.f), want_lo);
      continue;
    }
    float* g = G + (size_t)m * ldg;
    const int p = parent ? parent[m] : -1;
    const float* cp = (Cprev && p >= 0) ? Cprev + (size_t)p * C : nullptr;
    for (int c4 = lane; c4 < C / 4; c4 += 32) {
      const int c = 4 * c4;
      float4 f = ld4(g + c), r = ld4(g + C + c), mm = ld4(g + 2 * C + c);
      const float4 cprev = cp ? ld4(cp + c) : make_float4(0.f, 0.f, 0.f, 0.f);
      if (!activated) {                                      // (the tensor-core gate contraction activates in its epilogue)
        f.x = sigmoidf_(f.x); f.y = sigmoidf_(f.y); f.z = sigmoidf_(f.z); f.w = sigmoidf_(f.w);
        r.x = sigmoidf_(r.x); r.y = sigmoidf_(r.y); r.z = sigmoidf_(r.z); r.w = sigmoidf_(r.w);
        mm.x = tanhf(mm.x); mm.y = tanhf(mm.y); mm.z = tanhf(mm.z); mm.w = tanhf(mm.w);
        st4(g + c, f); st4(g + C + c, r); st4(g + 2 * C + c, mm);
      }
      const float4 cn = make_float4(cprev.x * f.x + r.x * mm.x, cprev.y * f.y + r.y * mm.y, cprev.z * f.z + r.z * mm.z,
                                    cprev.w * f.w + r.w * mm.w);
      st4(Cn + (size_t)m * C + c, cn);
      if (CnP.p) planes_store4(CnP, m, c, cn, want_lo);
    }
    if (!activated)
      for (int j4 = lane; j4 < D / 4; j4 += 32) {
        float4 o = ld4(g + 3 * C + 4 * j4);
        o.x = sigmoidf_(o.x); o.y = sigmoidf_(o.y); o.z = sigmoidf_(o.z); o.w = sigmoidf_(o.w);
        st4(g + 3 * C + 4 * j4, o);
      }
  }
}

// t = tanh(P), h' = o * t, Y = X + h'                       (interface.py:56, paths.py:89)
__global__ void __launch_bounds__(TPB) lstm_point2_fwd_kernel(float* __restrict__ Pt, const float* __restrict__ G,
                                                              const float* __restrict__ X, float* __restrict__ H,
                                                              float* __restrict__ Y, Planes YP, int want_lo, int D, int C, int M_host,
                                                              const int* M_dev) {
  pdl_grid_sync();
  const int M = row_count(M_host, M_dev);
  const int ldg = 3 * C + D;
  const int Mw = YP.p ? planes_rows_to_write(YP, M) : M;
  ROW_LOOP(m, Mw) {
    if (m >= M) {                                            // zero rows behind the live ones (planes only)
      for (int j4 = lane; j4 < D / 4; j4 += 32) planes_store4(YP, m, 4 * j4, make_float4(0.f, 0.f, 0.f, 0.f), want_lo);
      continue;
    }
    // four column groups per pass, all twelve 16-byte loads issued before the first store (P is read and written through the
    // same pointer, so the compiler will not move the loads of a later group above the stores of an earlier one by itself)
    for (int j0 = lane; j0 < D / 4; j0 += 4 * 32) {
      float4 t[4], o[4], x[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int j = 4 * (j0 + 32 * u);
        t[u] = o[u] = x[u] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (j < D) { t[u] = ld4(Pt + (size_t)m * D + j); o[u] = ld4(G + (size_t)m * ldg + 3 * C + j); x[u] = ld4(X + (size_t)m * D + j); }
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int j = 4 * (j0 + 32 * u);
        if (j >= D) continue;
        float4 tt = t[u];
        tt.x = tanhf(tt.x); tt.y = tanhf(tt.y); tt.z = tanhf(tt.z); tt.w = tanhf(tt.w);
        st4(Pt + (size_t)m * D + j, tt);                     // (recomputing tanh in the backward kernel costs more than these 4 KB/row)
        const float4 h = make_float4(o[u].x * tt.x, o[u].y * tt.y, o[u].z * tt.z, o[u].w * tt.w);
        st4(H + (size_t)m * D + j, h);
        const float4 y = make_float4(x[u].x + h.x, x[u].y + h.y, x[u].z + h.z, x[u].w + h.w);
        if (Y) st4(Y + (size_t)m * D + j, y);
        if (YP.p) planes_store4(YP, m, j, y, want_lo);
      }
    }
  }
}

// dH = dY + dHc;  dP = dH*o*(1-t^2), dGo = dH*t*o*(1-o);  column sums of both (the bias gradients of mem_to_out and of the
// output gate) are accumulated on the way.  Threads own float4 column groups, blocks stride over rows.
constexpr int CG = 4;                                     // float4 column groups per thread: D <= 4 * 4 * 256
//   child_of != nullptr: dHc is the CHILD level's per-row gradient table and row m adds the rows child_of[4m .. 4m+3] (>= 0) in
//   that order -- the child -> parent sum of parent_sum_kernel without the [M, D] round trip through HBM.
// Two rows per block iteration, every load of both rows issued before the first use: the child rows hang off an index load, and
// one row at a time leaves too few bytes in flight per SM to cover the two dependent latencies.
constexpr int P2R = 2;
template <int CGT>                                        // float4 column groups per thread actually in use: ceil(D / 1024)
__global__ void __launch_bounds__(256) lstm_point2_bwd_kernel(const float* __restrict__ dY, const float* __restrict__ dHc,
                                                              const int* __restrict__ child_of,
                                                              const float* __restrict__ Pt, const float* __restrict__ G,
                                                              float* __restrict__ dP, Planes dPP, float* __restrict__ dG, Planes dGP,
                                                              float* __restrict__ db_mo, float* __restrict__ db_o, int want_lo, int D, int C,
                                                              int M_host, const int* M_dev) {
  pdl_grid_sync();
  const int M = row_count(M_host, M_dev);
  const int ldg = 3 * C + D;
  const int Mw = dPP.p ? planes_rows_to_write(dPP, M) : M;
  const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
  float4 aP[CGT], aG[CGT];
#pragma unroll
  for (int q = 0; q < CGT; ++q) aP[q] = aG[q] = z4;
  for (int m0 = blockIdx.x; m0 < Mw; m0 += P2R * gridDim.x) {
    int4 ch[P2R];
#pragma unroll
    for (int r = 0; r < P2R; ++r) {
      const int m = m0 + r * gridDim.x;
      ch[r] = make_int4(-1, -1, -1, -1);
      if (child_of && m < M) ch[r] = __ldg(reinterpret_cast<const int4*>(child_of) + m);
    }
#pragma unroll
    for (int q = 0; q < CGT; ++q) {
      const int j = 4 * (threadIdx.x + 256 * q);
      if (j >= D) break;
      float4 dh[P2R], t[P2R], o[P2R], e[P2R][4];
#pragma unroll
      for (int r = 0; r < P2R; ++r) {
        const int m = m0 + r * gridDim.x;
        dh[r] = t[r] = o[r] = z4;
        e[r][0] = e[r][1] = e[r][2] = e[r][3] = z4;
        if (m < M) {
          if (dY) dh[r] = ld4(dY + (size_t)m * D + j);
          t[r] = ld4(Pt + (size_t)m * D + j);
          o[r] = ld4(G + (size_t)m * ldg + 3 * C + j);
          if (dHc && child_of) {
            if (ch[r].x >= 0) e[r][0] = ld4(dHc + (size_t)ch[r].x * D + j);
            if (ch[r].y >= 0) e[r][1] = ld4(dHc + (size_t)ch[r].y * D + j);
            if (ch[r].z >= 0) e[r][2] = ld4(dHc + (size_t)ch[r].z * D + j);
            if (ch[r].w >= 0) e[r][3] = ld4(dHc + (size_t)ch[r].w * D + j);
          } else if (dHc) {
            e[r][0] = ld4(dHc + (size_t)m * D + j);
          }
        }
      }
#pragma unroll
      for (int r = 0; r < P2R; ++r) {
        const int m = m0 + r * gridDim.x;
        if (m >= Mw) continue;
        float4 vp = z4, vg = z4;
        if (m < M) {
          float4 s4 = e[r][0];                                 // ((c0 + c1) + c2) + c3: the order of parent_sum_kernel
          s4.x += e[r][1].x; s4.y += e[r][1].y; s4.z += e[r][1].z; s4.w += e[r][1].w;
          s4.x += e[r][2].x; s4.y += e[r][2].y; s4.z += e[r][2].z; s4.w += e[r][2].w;
          s4.x += e[r][3].x; s4.y += e[r][3].y; s4.z += e[r][3].z; s4.w += e[r][3].w;
          const float4 d4 = make_float4(dh[r].x + s4.x, dh[r].y + s4.y, dh[r].z + s4.z, dh[r].w + s4.w);
          const float4 tt = t[r], oo = o[r];
          vp = make_float4(d4.x * oo.x * (1.f - tt.x * tt.x), d4.y * oo.y * (1.f - tt.y * tt.y), d4.z * oo.z * (1.f - tt.z * tt.z),
                           d4.w * oo.w * (1.f - tt.w * tt.w));
          vg = make_float4(d4.x * tt.x * oo.x * (1.f - oo.x), d4.y * tt.y * oo.y * (1.f - oo.y), d4.z * tt.z * oo.z * (1.f - oo.z),
                           d4.w * tt.w * oo.w * (1.f - oo.w));
          aP[q].x += vp.x; aP[q].y += vp.y; aP[q].z += vp.z; aP[q].w += vp.w;
          aG[q].x += vg.x; aG[q].y += vg.y; aG[q].z += vg.z; aG[q].w += vg.w;
          if (dP) st4(dP + (size_t)m * D + j, vp);
          if (dG) st4(dG + (size_t)m * ldg + 3 * C + j, vg);
        }
        if (dPP.p) planes_store4(dPP, m, j, vp, want_lo);
        if (dGP.p) planes_store4(dGP, m, 3 * C + j, vg, want_lo);
      }
    }
  }
#pragma unroll
  for (int q = 0; q < CGT; ++q) {
    const int j = 4 * (threadIdx.x + 256 * q);
    if (j >= D) break;
    atomicAdd(db_mo + j, aP[q].x); atomicAdd(db_mo + j + 1, aP[q].y); atomicAdd(db_mo + j + 2, aP[q].z); atomicAdd(db_mo + j + 3, aP[q].w);
    atomicAdd(db_o + j, aG[q].x); atomicAdd(db_o + j + 1, aG[q].y); atomicAdd(db_o + j + 2, aG[q].z); atomicAdd(db_o + j + 3, aG[q].w);
  }
}

// dc' = dCn + dCc;  df, dr, dm -> dG[:, :3C] (+ their column sums -> gate bias gradients);  dCrow = dc' * f (gradient w.r.t. the
// inherited c of this row).  A block covers 256 / (C/4) rows at a time; thread = (row slot, float4 column group).
__global__ void __launch_bounds__(256) lstm_point1_bwd_kernel(const float* __restrict__ dCn, const float* __restrict__ dCc,
                                                              const int* __restrict__ child_of,
                                                              const float* __restrict__ G, const float* __restrict__ Cprev,
                                                              const int* __restrict__ parent, float* __restrict__ dG, Planes dGP,
                                                              float* __restrict__ dCrow, float* __restrict__ db_g, int want_lo, int D, int C,
                                                              int M_host, const int* M_dev) {
  pdl_grid_sync();
  const int M = row_count(M_host, M_dev);
  const int ldg = 3 * C + D;
  const int G4 = C / 4, R = 256 / G4;
  const int c4 = threadIdx.x % G4, sub = threadIdx.x / G4;
  const int c = 4 * c4;
  const int Mw = dGP.p ? planes_rows_to_write(dGP, M) : M;
  float4 af = make_float4(0.f, 0.f, 0.f, 0.f), ar = af, am = af;
  if (sub < R) {
    for (int m = blockIdx.x * R + sub; m < Mw; m += gridDim.x * R) {
      float4 df = make_float4(0.f, 0.f, 0.f, 0.f), dr = df, dm = df;
      if (m < M) {
        const float* g = G + (size_t)m * ldg;
        const int p = parent ? parent[m] : -1;
        float4 dc = ld4(dCn + (size_t)m * C + c);
        if (dCc && child_of) {                                  // children's rows of the child level's dCrow, summed in index order
          const int4 ch = __ldg(reinterpret_cast<const int4*>(child_of) + m);
          float4 e = make_float4(0.f, 0.f, 0.f, 0.f);
          if (ch.x >= 0) { const float4 u = ld4(dCc + (size_t)ch.x * C + c); e.x += u.x; e.y += u.y; e.z += u.z; e.w += u.w; }
          if (ch.y >= 0) { const float4 u = ld4(dCc + (size_t)ch.y * C + c); e.x += u.x; e.y += u.y; e.z += u.z; e.w += u.w; }
          if (ch.z >= 0) { const float4 u = ld4(dCc + (size_t)ch.z * C + c); e.x += u.x; e.y += u.y; e.z += u.z; e.w += u.w; }
          if (ch.w >= 0) { const float4 u = ld4(dCc + (size_t)ch.w * C + c); e.x += u.x; e.y += u.y; e.z += u.z; e.w += u.w; }
          dc.x += e.x; dc.y += e.y; dc.z += e.z; dc.w += e.w;
        } else if (dCc) { const float4 e = ld4(dCc + (size_t)m * C + c); dc.x += e.x; dc.y += e.y; dc.z += e.z; dc.w += e.w; }
        const float4 f = ld4(g + c), r = ld4(g + C + c), mm = ld4(g + 2 * C + c);
        const float4 cprev = (Cprev && p >= 0) ? ld4(Cprev + (size_t)p * C + c) : make_float4(0.f, 0.f, 0.f, 0.f);
        df = make_float4(dc.x * cprev.x * f.x * (1.f - f.x), dc.y * cprev.y * f.y * (1.f - f.y), dc.z * cprev.z * f.z * (1.f - f.z),
                         dc.w * cprev.w * f.w * (1.f - f.w));
        dr = make_float4(dc.x * mm.x * r.x * (1.f - r.x), dc.y * mm.y * r.y * (1.f - r.y), dc.z * mm.z * r.z * (1.f - r.z),
                         dc.w * mm.w * r.w * (1.f - r.w));
        dm = make_float4(dc.x * r.x * (1.f - mm.x * mm.x), dc.y * r.y * (1.f - mm.y * mm.y), dc.z * r.z * (1.f - mm.z * mm.z),
                         dc.w * r.w * (1.f - mm.w * mm.w));
        af.x += df.x; af.y += df.y; af.z += df.z; af.w += df.w;
        ar.x += dr.x; ar.y += dr.y; ar.z += dr.z; ar.w += dr.w;
        am.x += dm.x; am.y += dm.y; am.z += dm.z; am.w += dm.w;
        if (dG) { float* dg = dG + (size_t)m * ldg; st4(dg + c, df); st4(dg + C + c, dr); st4(dg + 2 * C + c, dm); }
        if (dCrow) st4(dCrow + (size_t)m * C + c, make_float4(dc.x * f.x, dc.y * f.y, dc.z * f.z, dc.w * f.w));
      }
      if (dGP.p) { planes_store4(dGP, m, c, df, want_lo); planes_store4(dGP, m, C + c, dr, want_lo); planes_store4(dGP, m, 2 * C + c, dm, want_lo); }
    }
  }
  __shared__ float4 red[3][256];
  red[0][threadIdx.x] = af; red[1][threadIdx.x] = ar; red[2][threadIdx.x] = am;
  __syncthreads();
  if (sub == 0) {
#pragma unroll
    for (int gsel = 0; gsel < 3; ++gsel) {
      float4 s4 = make_float4(0.f, 0.f, 0.f, 0.f);
      for (int r = 0; r < R; ++r) { const float4 v = red[gsel][r * G4 + c4]; s4.x += v.x; s4.y += v.y; s4.z += v.z; s4.w += v.w; }
      float* o = db_g + gsel * C + c;
      atomicAdd(o, s4.x); atomicAdd(o + 1, s4.y); atomicAdd(o + 2, s4.z); atomicAdd(o + 3, s4.w);
    }
  }
}

// Deterministic child -> parent sum of state gradients (autograd of `torch.cat((ctx_patch,)*4)[filter]`, slide.py:318-329).
__global__ void __launch_bounds__(TPB) parent_sum_kernel(const int* __restrict__ child_of, const float* __restrict__ dHrow,
                                                         const float* __restrict__ dCrow, float* __restrict__ dHc,
                                                         float* __restrict__ dCc, int D, int C, int M_host, const int* M_dev) {
  pdl_grid_sync();
  const int M = row_count(M_host, M_dev);
  ROW_LOOP(m, M) {
    int ch[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) ch[j] = child_of[4 * m + j];
    for (int c = lane; c < D; c += 32) {
      float s = 0.f;
#pragma unroll
      for (int j = 0; j < 4; ++j) if (ch[j] >= 0) s += dHrow[(size_t)ch[j] * D + c];
      dHc[(size_t)m * D + c] = s;
    }
    for (int c = lane; c < C; c += 32) {
      float s = 0.f;
#pragma unroll
      for (int j = 0; j < 4; ++j) if (ch[j] >= 0) s += dCrow[(size_t)ch[j] * C + c];
      dCc[(size_t)m * C + c] = s;
    }
  }
}

// ---------------------------------------------------------------------------------------------------------
// importance + token: u = relu(U+b1); alpha = sigmoid(u.w2+b2); tok = alpha*V + b_in + PE      (paths.py:95-98,120-124;
// aggregator.py:43-56; utils.py:47-67).  UV = Y.[W1;Win]^T is produced by the preceding GEMM (row scaling commutes).
__device__ __forceinline__ float pos_enc(int pe_mode, const float* __restrict__ div, int d, int c, int x, int y, int i) {
  if (pe_mode == 2) {                       // 2-D: first half from x, second half from y; d/4 frequencies
    const int half = d / 2;
    const int cc = c < half ? c : c - half;
    const float pos = (float)(c < half ? x : y);
    const float a = pos * div[cc >> 1];
    return (cc & 1) ? cosf(a) : sinf(a);
  }
  const float a = (float)i * div[c >> 1];    // 1-D over the row index inside the bag (utils.py:16-23)
  return (c & 1) ? cosf(a) : sinf(a);
}

__global__ void __launch_bounds__(TPB) imp_tok_fwd_kernel(float* __restrict__ UV, const float* __restrict__ b1,
                                                          const float* __restrict__ w2, const float* __restrict__ b2,
                                                          const float* __restrict__ bin, const float* __restrict__ div,
                                                          const int* __restrict__ row_slide, const int* __restrict__ row_xy,
                                                          const int* __restrict__ seg, float* __restrict__ alpha,
                                                          float* __restrict__ tok, int I, int d, int imp_mul, int pe_mode,
                                                          int mode, int M_host, const int* M_dev) {
  pdl_grid_sync();
  // mode 0: importance + token (LSTM path); 1: importance only; 2: token only, unscaled (RNN variant: V = Z.W_in^T)
  const int M = row_count(M_host, M_dev);
  const int ld = I + d;
  ROW_LOOP(m, M) {
    float* uv = UV + (size_t)m * ld;
    float a = 1.f;
    if (mode != 2) {
      float s = 0.f;
      for (int j = lane; j < I; j += 32) {
        const float u = fmaxf(uv[j] + b1[j], 0.f);
        uv[j] = u;
        s += u * w2[j];
      }
      s = warp_sum(s) + b2[0];
      a = sigmoidf_(s);
      if (lane == 0) alpha[m] = a;
      if (mode == 1) continue;
    }
    const int sl = row_slide[m];
    const int x = row_xy[2 * m], y = row_xy[2 * m + 1], i = m - seg[sl];
    float* t = tok + (size_t)(m + sl + 1) * d;
    const float sc = (imp_mul && mode == 0) ? a : 1.f;
    for (int c = lane; c < d; c += 32) t[c] = sc * uv[I + c] + bin[c] + pos_enc(pe_mode, div, d, c, x, y, i);
  }
}

// mode 0 with 16-byte accesses (I, d multiples of 4, at most 256): a lane owns the float4 column groups lane + 32 q, a warp works on
// two rows at a time with their loads issued before the first use (same scheme as imp_tok_bwd4_kernel below).
template <int Q>
__global__ void __launch_bounds__(TPB) imp_tok_fwd4_kernel(float* __restrict__ UV, const float* __restrict__ b1,
                                                           const float* __restrict__ w2, const float* __restrict__ b2,
                                                           const float* __restrict__ bin, const float* __restrict__ div,
                                                           const int* __restrict__ row_slide, const int* __restrict__ row_xy,
                                                           const int* __restrict__ seg, float* __restrict__ alpha,
                                                           float* __restrict__ tok, int I, int d, int imp_mul, int pe_mode,
                                                           int M_host, const int* M_dev) {
  pdl_grid_sync();
  constexpr int R = 2;
  const int M = row_count(M_host, M_dev);
  const int ld = I + d;
  const int lane = threadIdx.x & 31;
  const int wid = blockIdx.x * WPB + (threadIdx.x >> 5), nw = gridDim.x * WPB;
  const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
  float4 b14[Q], w24[Q], bin4[Q];
#pragma unroll
  for (int q = 0; q < Q; ++q) {
    const int c = 4 * lane + 128 * q;
    b14[q] = c < I ? ld4(b1 + c) : z4; w24[q] = c < I ? ld4(w2 + c) : z4; bin4[q] = c < d ? ld4(bin + c) : z4;
  }
  const float b2v = b2[0];
  for (int m0 = wid; m0 < M; m0 += R * nw) {
    float4 u4[R][Q], v4[R][Q];
    int sl[R], px[R], py[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const int m = m0 + r * nw;
      sl[r] = px[r] = py[r] = 0;
#pragma unroll
      for (int q = 0; q < Q; ++q) u4[r][q] = v4[r][q] = z4;
      if (m < M) {
        const float* uv = UV + (size_t)m * ld;
        sl[r] = row_slide[m]; px[r] = row_xy[2 * m]; py[r] = row_xy[2 * m + 1];
#pragma unroll
        for (int q = 0; q < Q; ++q) {
          const int c = 4 * lane + 128 * q;
          if (c < I) u4[r][q] = ld4(uv + c);
          if (c < d) v4[r][q] = ld4(uv + I + c);
        }
      }
    }
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const int m = m0 + r * nw;
      if (m >= M) continue;                                  // warp-uniform
      float* uv = UV + (size_t)m * ld;
      float s = 0.f;
#pragma unroll
      for (int q = 0; q < Q; ++q) {
        const int j = 4 * lane + 128 * q;
        if (j < I) {
          const float4 u = make_float4(fmaxf(u4[r][q].x + b14[q].x, 0.f), fmaxf(u4[r][q].y + b14[q].y, 0.f), fmaxf(u4[r][q].z + b14[q].z, 0.f),
                                       fmaxf(u4[r][q].w + b14[q].w, 0.f));
          st4(uv + j, u);
          s += (u.x * w24[q].x + u.y * w24[q].y) + (u.z * w24[q].z + u.w * w24[q].w);
        }
      }
      s = warp_sum(s) + b2v;
      const float a = sigmoidf_(s);
      if (lane == 0) alpha[m] = a;
      const int i = m - seg[sl[r]];
      float* t = tok + (size_t)(m + sl[r] + 1) * d;
      const float sc = imp_mul ? a : 1.f;
#pragma unroll
      for (int q = 0; q < Q; ++q) {
        const int c = 4 * lane + 128 * q;
        if (c < d) {
          const float4 v = v4[r][q];
          st4(t + c, make_float4(sc * v.x + bin4[q].x + pos_enc(pe_mode, div, d, c, px[r], py[r], i),
                                 sc * v.y + bin4[q].y + pos_enc(pe_mode, div, d, c + 1, px[r], py[r], i),
                                 sc * v.z + bin4[q].z + pos_enc(pe_mode, div, d, c + 2, px[r], py[r], i),
                                 sc * v.w + bin4[q].w + pos_enc(pe_mode, div, d, c + 3, px[r], py[r], i)));
        }
      }
    }
  }
}

__global__ void __launch_bounds__(TPB) special_fill_kernel(const float* __restrict__ special, const int* __restrict__ seg,
                                                           float* __restrict__ tok, int d, int B) {
  pdl_grid_sync();
  ROW_LOOP(s, B) {
    float* t = tok + (size_t)(seg[s] + s) * d;
    for (int c = lane; c < d; c += 32) t[c] = special[c];
  }
}

// backward of imp_tok: dUV = [du | dV]; accumulates db1, dw2, db2, dbin.  Columns are lane-strided so every lane
// keeps private partial sums for its columns across the rows its warp visits (d, I <= 256).
__global__ void __launch_bounds__(TPB) imp_tok_bwd_kernel(const float* __restrict__ dtok, const float* __restrict__ UV,
                                                          const float* __restrict__ alpha, const float* __restrict__ w2,
                                                          const int* __restrict__ row_slide, float* __restrict__ dUV, Planes dUVP,
                                                          int want_lo, float* __restrict__ db1, float* __restrict__ dw2,
                                                          float* __restrict__ db2, float* __restrict__ dbin, int I, int d, int imp_mul,
                                                          int mode, int M_host, const int* M_dev) {
  pdl_grid_sync();
  // mode 0: full backward of imp_tok_fwd; 2: token part only (dV = dtok, dbin), the importance head is handled elsewhere
  const int M = row_count(M_host, M_dev);
  const int ld = I + d;
  if (mode == 2) imp_mul = 0;
  float a_b1[8], a_w2[8], a_bin[8], a_b2 = 0.f;
#pragma unroll
  for (int q = 0; q < 8; ++q) a_b1[q] = a_w2[q] = a_bin[q] = 0.f;
  const int Mw = dUVP.p ? planes_rows_to_write(dUVP, M) : M;
  auto put = [&](int m, int col, float v) {
    if (dUV) dUV[(size_t)m * ld + col] = v;
    if (dUVP.p) {
      const __nv_bfloat16 h = __float2bfloat16_rn(v);
      dUVP.p[(size_t)m * dUVP.ldp + col] = h;
      if (want_lo) dUVP.p[((size_t)dUVP.rows + m) * dUVP.ldp + col] = __float2bfloat16_rn(v - __bfloat162float(h));
    }
  };
  ROW_LOOP(m, Mw) {
    if (m >= M) {
      for (int col = lane; col < ld; col += 32) put(m, col, 0.f);
      continue;
    }
    const float* uv = UV + (size_t)m * ld;
    const float* dt = dtok + (size_t)(m + row_slide[m] + 1) * d;
    const float a = alpha[m];
    float da = 0.f;
#pragma unroll
    for (int q = 0; q < 8; ++q) {
      const int c = lane + 32 * q;
      if (c < d) {
        const float g = dt[c];
        a_bin[q] += g;
        da += g * uv[I + c];
        put(m, I + c, imp_mul ? a * g : g);
      }
    }
    float ds = 0.f;
    if (imp_mul) { da = warp_sum(da); ds = da * a * (1.f - a); }
    a_b2 += ds;
    if (mode == 2) continue;
#pragma unroll
    for (int q = 0; q < 8; ++q) {
      const int j = lane + 32 * q;
      if (j < I) {
        const float u = uv[j];
        a_w2[q] += ds * u;
        const float du = (u > 0.f) ? ds * w2[j] : 0.f;
        a_b1[q] += du;
        put(m, j, du);
      }
    }
  }
  // block reduction through shared memory, then one atomic per column per block
  __shared__ float red[WPB][3][256];
  const int w = threadIdx.x >> 5;
#pragma unroll
  for (int q = 0; q < 8; ++q) {
    red[w][0][lane + 32 * q] = a_b1[q]; red[w][1][lane + 32 * q] = a_w2[q]; red[w][2][lane + 32 * q] = a_bin[q];
  }
  __shared__ float red_b2[WPB];
  if (lane == 0) red_b2[w] = a_b2;
  __syncthreads();
  for (int c = threadIdx.x; c < 256; c += TPB) {
    float s0 = 0.f, s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int ww = 0; ww < WPB; ++ww) { s0 += red[ww][0][c]; s1 += red[ww][1][c]; s2 += red[ww][2][c]; }
    if (c < I && mode != 2) { if (s0 != 0.f) atomicAdd(db1 + c, s0); if (s1 != 0.f) atomicAdd(dw2 + c, s1); }
    if (c < d && s2 != 0.f) atomicAdd(dbin + c, s2);
  }
  if (threadIdx.x == 0 && mode != 2) {
    float s = 0.f;
#pragma unroll
    for (int ww = 0; ww < WPB; ++ww) s += red_b2[ww];
    if (s != 0.f) atomicAdd(db2, s);
  }
}

// The same backward with 16-byte accesses (I, d multiples of 4, at most 256): a lane owns the float4 column groups lane + 32 q,
// a warp works on two rows at a time with every load issued before the first use, and the grid is several blocks per SM -- the
// scalar kernel above is bound by the latency of its per-row chain (slide index -> token row -> loads -> warp sum -> stores).
constexpr int ITR = 2;                                    // rows per warp iteration
template <int Q>                                           // float4 column groups per lane: ceil(max(I, d) / 128)
__global__ void __launch_bounds__(TPB) imp_tok_bwd4_kernel(const float* __restrict__ dtok, const float* __restrict__ UV,
                                                           const float* __restrict__ alpha, const float* __restrict__ w2,
                                                           const int* __restrict__ row_slide, float* __restrict__ dUV, Planes dUVP,
                                                           int want_lo, float* __restrict__ db1, float* __restrict__ dw2,
                                                           float* __restrict__ db2, float* __restrict__ dbin, int I, int d, int imp_mul,
                                                           int M_host, const int* M_dev) {
  pdl_grid_sync();
  const int M = row_count(M_host, M_dev);
  const int ld = I + d;
  const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
  float4 a_b1[Q], a_w2[Q], a_bin[Q];
  float a_b2 = 0.f;
#pragma unroll
  for (int q = 0; q < Q; ++q) a_b1[q] = a_w2[q] = a_bin[q] = z4;
  const int Mw = dUVP.p ? planes_rows_to_write(dUVP, M) : M;
  const int lane = threadIdx.x & 31;
  const int wid = blockIdx.x * WPB + (threadIdx.x >> 5), nw = gridDim.x * WPB;
  float4 w4[Q];
#pragma unroll
  for (int q = 0; q < Q; ++q) { const int j = 4 * lane + 128 * q; w4[q] = j < I ? ld4(w2 + j) : z4; }
  auto put4 = [&](int m, int col, const float4& v) {
    if (dUV) st4(dUV + (size_t)m * ld + col, v);
    if (dUVP.p) planes_store4(dUVP, m, col, v, want_lo != 0);
  };
  for (int m0 = wid; m0 < Mw; m0 += ITR * nw) {
    float4 g4[ITR][Q], v4[ITR][Q], u4[ITR][Q];
    float a[ITR];
#pragma unroll
    for (int r = 0; r < ITR; ++r) {
      const int m = m0 + r * nw;
      a[r] = 0.f;
#pragma unroll
      for (int q = 0; q < Q; ++q) g4[r][q] = v4[r][q] = u4[r][q] = z4;
      if (m < M) {
        const float* uv = UV + (size_t)m * ld;
        const float* dt = dtok + (size_t)(m + row_slide[m] + 1) * d;
        a[r] = alpha[m];
#pragma unroll
        for (int q = 0; q < Q; ++q) {
          const int c = 4 * lane + 128 * q;
          if (c < d) { g4[r][q] = ld4(dt + c); v4[r][q] = ld4(uv + I + c); }
          if (c < I) u4[r][q] = ld4(uv + c);
        }
      }
    }
#pragma unroll
    for (int r = 0; r < ITR; ++r) {
      const int m = m0 + r * nw;
      if (m >= Mw) continue;                                 // warp-uniform
      if (m >= M) {
#pragma unroll
        for (int q = 0; q < Q; ++q) { const int c = 4 * lane + 128 * q; if (c < I) put4(m, c, z4); if (c < d) put4(m, I + c, z4); }
        continue;
      }
      float da = 0.f;
#pragma unroll
      for (int q = 0; q < Q; ++q) {
        const int c = 4 * lane + 128 * q;
        if (c < d) {
          const float4 g = g4[r][q], v = v4[r][q];
          a_bin[q].x += g.x; a_bin[q].y += g.y; a_bin[q].z += g.z; a_bin[q].w += g.w;
          da += g.x * v.x + g.y * v.y + g.z * v.z + g.w * v.w;
          put4(m, I + c, imp_mul ? make_float4(a[r] * g.x, a[r] * g.y, a[r] * g.z, a[r] * g.w) : g);
        }
      }
      float ds = 0.f;
      if (imp_mul) { da = warp_sum(da); ds = da * a[r] * (1.f - a[r]); }
      a_b2 += ds;
#pragma unroll
      for (int q = 0; q < Q; ++q) {
        const int j = 4 * lane + 128 * q;
        if (j < I) {
          const float4 u = u4[r][q];
          a_w2[q].x += ds * u.x; a_w2[q].y += ds * u.y; a_w2[q].z += ds * u.z; a_w2[q].w += ds * u.w;
          const float4 du = make_float4(u.x > 0.f ? ds * w4[q].x : 0.f, u.y > 0.f ? ds * w4[q].y : 0.f, u.z > 0.f ? ds * w4[q].z : 0.f,
                                        u.w > 0.f ? ds * w4[q].w : 0.f);
          a_b1[q].x += du.x; a_b1[q].y += du.y; a_b1[q].z += du.z; a_b1[q].w += du.w;
          put4(m, j, du);
        }
      }
    }
  }
  __shared__ float4 red[WPB][3][32 * Q];
  const int w = threadIdx.x >> 5;
#pragma unroll
  for (int q = 0; q < Q; ++q) { red[w][0][lane + 32 * q] = a_b1[q]; red[w][1][lane + 32 * q] = a_w2[q]; red[w][2][lane + 32 * q] = a_bin[q]; }
  __shared__ float red_b2[WPB];
  if (lane == 0) red_b2[w] = a_b2;
  __syncthreads();
  for (int e = threadIdx.x; e < 3 * 32 * Q; e += TPB) {
    const int k = e / (32 * Q), c4 = e % (32 * Q), c = 4 * c4;
    float4 t = red[0][k][c4];
#pragma unroll
    for (int ww = 1; ww < WPB; ++ww) { const float4 u = red[ww][k][c4]; t.x += u.x; t.y += u.y; t.z += u.z; t.w += u.w; }
    float* o = k == 0 ? db1 : k == 1 ? dw2 : dbin;
    if (c < (k == 2 ? d : I)) { atomicAdd(o + c, t.x); atomicAdd(o + c + 1, t.y); atomicAdd(o + c + 2, t.z); atomicAdd(o + c + 3, t.w); }
  }
  if (threadIdx.x == 0) {
    float sb = 0.f;
#pragma unroll
    for (int ww = 0; ww < WPB; ++ww) sb += red_b2[ww];
    if (sb != 0.f) atomicAdd(db2, sb);
  }
}

// ---- RNN variant (lstm = false, paths.py:96-109): Z = alpha * X + hctx ------------------------------------------------
__global__ void __launch_bounds__(TPB) rnn_z_fwd_kernel(const float* __restrict__ X, const float* __restrict__ alpha,
                                                        const float* __restrict__ hctx, float* __restrict__ Z, int D, int imp_mul,
                                                        int M_host, const int* M_dev) {
  pdl_grid_sync();
  const int M = row_count(M_host, M_dev);
  ROW_LOOP(m, M) {
    const float a = imp_mul ? alpha[m] : 1.f;
    for (int c = lane; c < D; c += 32) {
      float z = a * X[(size_t)m * D + c];
      if (hctx) z += hctx[(size_t)m * D + c];
      Z[(size_t)m * D + c] = z;
    }
  }
}

// dZ (+= dZc from the children); d alpha = <dZ, X>; backward of the importance head -> dUV[:, :I], db1, dw2, db2
__global__ void __launch_bounds__(TPB) rnn_alpha_bwd_kernel(float* __restrict__ dZ, const float* __restrict__ dZc,
                                                            const float* __restrict__ X, const float* __restrict__ UV,
                                                            const float* __restrict__ alpha, const float* __restrict__ w2,
                                                            float* __restrict__ dUV, float* __restrict__ db1, float* __restrict__ dw2,
                                                            float* __restrict__ db2, int D, int I, int d, int imp_mul, int M_host,
                                                            const int* M_dev) {
  pdl_grid_sync();
  const int M = row_count(M_host, M_dev);
  const int ld = I + d;
  float a_b1[8], a_w2[8], a_b2 = 0.f;
#pragma unroll
  for (int q = 0; q < 8; ++q) a_b1[q] = a_w2[q] = 0.f;
  ROW_LOOP(m, M) {
    float da = 0.f;
    for (int c = lane; c < D; c += 32) {
      float g = dZ[(size_t)m * D + c];
      if (dZc) { g += dZc[(size_t)m * D + c]; dZ[(size_t)m * D + c] = g; }
      da += g * X[(size_t)m * D + c];
    }
    const float a = alpha[m];
    float ds = 0.f;
    if (imp_mul) { da = warp_sum(da); ds = da * a * (1.f - a); }
    a_b2 += ds;
    const float* uv = UV + (size_t)m * ld;
    float* duv = dUV + (size_t)m * ld;
#pragma unroll
    for (int q = 0; q < 8; ++q) {
      const int j = lane + 32 * q;
      if (j < I) {
        const float u = uv[j];
        a_w2[q] += ds * u;
        const float du = (u > 0.f) ? ds * w2[j] : 0.f;
        a_b1[q] += du;
        duv[j] = du;
      }
    }
  }
  __shared__ float red[WPB][2][256];
  const int w = threadIdx.x >> 5;
#pragma unroll
  for (int q = 0; q < 8; ++q) { red[w][0][lane + 32 * q] = a_b1[q]; red[w][1][lane + 32 * q] = a_w2[q]; }
  __shared__ float red_b2[WPB];
  if (lane == 0) red_b2[w] = a_b2;
  __syncthreads();
  for (int c = threadIdx.x; c < I; c += TPB) {
    float s0 = 0.f, s1 = 0.f;
#pragma unroll
    for (int ww = 0; ww < WPB; ++ww) { s0 += red[ww][0][c]; s1 += red[ww][1][c]; }
    if (s0 != 0.f) atomicAdd(db1 + c, s0);
    if (s1 != 0.f) atomicAdd(dw2 + c, s1);
  }
  if (threadIdx.x == 0) {
    float s = 0.f;
#pragma unroll
    for (int ww = 0; ww < WPB; ++ww) s += red_b2[ww];
    if (s != 0.f) atomicAdd(db2, s);
  }
}

// ---- slide_ctx_mode "concat" (paths.py:134-137): logits = Wc . [F_0 | ... | F_l] + bc ------------------------------------
__global__ void __launch_bounds__(TPB) head_concat_fwd_kernel(FList fl, const float* __restrict__ Wc, const float* __restrict__ bc,
                                                              float* __restrict__ logits, int d, int nl, int B) {
  pdl_grid_sync();
  const int ldw = fl.n * d;
  ROW_LOOP(s, B) {
    for (int k = 0; k < nl; ++k) {
      float acc = 0.f;
      for (int j = 0; j < fl.n; ++j)
        for (int c = lane; c < d; c += 32) acc += fl.f[j][(size_t)s * d + c] * Wc[(size_t)k * ldw + j * d + c];
      acc = warp_sum(acc);
      if (lane == 0) logits[(size_t)s * nl + k] = acc + bc[k];
    }
  }
}
// dF_j = dlogits . Wc[:, j*d:(j+1)*d] for every level j <= l
__global__ void __launch_bounds__(TPB) head_concat_bwd_kernel(const float* __restrict__ dlogits, const float* __restrict__ Wc, GList gl,
                                                              int d, int nl, int B) {
  pdl_grid_sync();
  const int ldw = gl.n * d;
  ROW_LOOP(s, B) {
    for (int j = 0; j < gl.n; ++j)
      for (int c = lane; c < d; c += 32) {
        float g = 0.f;
        for (int k = 0; k < nl; ++k) g += dlogits[(size_t)s * nl + k] * Wc[(size_t)k * ldw + j * d + c];
        gl.g[j][(size_t)s * d + c] = g;
      }
  }
}

// dspecial[c] += sum_b dtok[special token of slide b, c]: a block sums 8 slides (one per warp) through shared memory and adds
// its 8-row partial with one atomic per column.
__global__ void __launch_bounds__(256) special_grad_kernel(const float* __restrict__ dtok, GIdx sg, MutL dsp, int d, int B) {
  pdl_grid_sync();
  __shared__ float part[8][256];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int b = blockIdx.x * 8 + w;
  const int* seg = pick(sg.idx.p, blockIdx.y);
  float* dspecial = pick(dsp.p, blockIdx.y);
  const float* r = b < B ? dtok + (size_t)(pick(sg.roff, blockIdx.y) + seg[b] + b) * d : nullptr;
#pragma unroll
  for (int q = 0; q < 8; ++q) {
    const int c = lane + 32 * q;
    part[w][c] = (r && c < d) ? r[c] : 0.f;
  }
  __syncthreads();
  const int c = threadIdx.x;
  if (c < d) {
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) s += part[k][c];
    atomicAdd(dspecial + c, s);
  }
}

// ---------------------------------------------------------------------------------------------------------
// y = LayerNorm(res[ridx] + add + bias) (post-norm residual block of nn.TransformerDecoderLayer), saving xhat and rstd.
// Grouped (common.cuh): blockIdx.y is the row group; buffers are concatenated (rows grp.roff[g] + m), parameters per group.
// With a row index list, res rows are ri.roff[g] + ridx[m] of the (token-shaped) res buffer.
__global__ void __launch_bounds__(TPB) res_ln_fwd_kernel(const float* __restrict__ res, GIdx ri, const float* __restrict__ add, PtrL biasl,
                                                         PtrL gammal, PtrL betal, float* __restrict__ y, float* __restrict__ xhat,
                                                         float* __restrict__ rstd, int d, float eps, Grp grp) {
  pdl_grid_sync();
  const int gi = blockIdx.y;
  const int M = grp_count(grp, gi);
  const size_t r0 = pick(grp.roff, gi);
  const int* ridx = pick(ri.idx.p, gi);
  if (res) res += (size_t)(ridx ? pick(ri.roff, gi) : r0) * d;
  if (add) add += r0 * d;
  y += r0 * d; xhat += r0 * d; rstd += r0;
  const float* __restrict__ bias = pick(biasl.p, gi); const float* __restrict__ gamma = pick(gammal.p, gi);
  const float* __restrict__ beta = pick(betal.p, gi);
  ROW_LOOP(m, M) {
    const float* r = res ? res + (size_t)(ridx ? ridx[m] : m) * d : nullptr;
    float v[8];
    float s = 0.f;
#pragma unroll
    for (int q = 0; q < 8; ++q) {
      const int c = lane + 32 * q;
      float x = 0.f;
      if (c < d) {
        if (r) x = r[c];
        if (add) x += add[(size_t)m * d + c];
        if (bias) x += bias[c];
      }
      v[q] = x; s += x;
    }
    const float mean = warp_sum(s) / d;
    float var = 0.f;
#pragma unroll
    for (int q = 0; q < 8; ++q) { const int c = lane + 32 * q; if (c < d) { const float t = v[q] - mean; var += t * t; } }
    var = warp_sum(var) / d;
    const float rs = rsqrtf(var + eps);
    if (lane == 0) rstd[m] = rs;
#pragma unroll
    for (int q = 0; q < 8; ++q) {
      const int c = lane + 32 * q;
      if (c < d) {
        const float xh = (v[q] - mean) * rs;
        xhat[(size_t)m * d + c] = xh;
        y[(size_t)m * d + c] = xh * gamma[c] + beta[c];
      }
    }
  }
}

// dx = rstd * (g - mean(g) - xhat*mean(g*xhat)), g = dy*gamma.  dx is written (or accumulated) to dx[didx[m]];
// dgamma/dbeta and (optionally) colsum(dx) -> dbias are accumulated.
__global__ void __launch_bounds__(TPB) ln_bwd_kernel(const float* __restrict__ dy, const float* __restrict__ dy2,
                                                     const float* __restrict__ xhat, const float* __restrict__ rstd, PtrL gammal,
                                                     float* __restrict__ dx, GIdx di, int accum, MutL dgammal, MutL dbetal, MutL dbiasl,
                                                     int d, Grp grp) {
  pdl_grid_sync();
  const int gi = blockIdx.y;
  const int M = grp_count(grp, gi);
  const size_t r0 = pick(grp.roff, gi);
  const int* didx = pick(di.idx.p, gi);
  dy += r0 * d; if (dy2) dy2 += r0 * d;
  xhat += r0 * d; rstd += r0;
  dx += (size_t)(didx ? pick(di.roff, gi) : r0) * d;
  const float* __restrict__ gamma = pick(gammal.p, gi);
  float* dgamma = pick(dgammal.p, gi); float* dbeta = pick(dbetal.p, gi); float* dbias = pick(dbiasl.p, gi);
  float a_g[8], a_b[8], a_x[8];
#pragma unroll
  for (int q = 0; q < 8; ++q) a_g[q] = a_b[q] = a_x[q] = 0.f;
  ROW_LOOP(m, M) {
    float g[8], xh[8];
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int q = 0; q < 8; ++q) {
      const int c = lane + 32 * q;
      g[q] = 0.f; xh[q] = 0.f;
      if (c < d) {
        float dyv = dy[(size_t)m * d + c];
        if (dy2) dyv += dy2[(size_t)m * d + c];
        xh[q] = xhat[(size_t)m * d + c];
        a_g[q] += dyv * xh[q]; a_b[q] += dyv;
        g[q] = dyv * gamma[c];
        s1 += g[q]; s2 += g[q] * xh[q];
      }
    }
    s1 = warp_sum(s1) / d; s2 = warp_sum(s2) / d;
    const float rs = rstd[m];
    float* o = dx + (size_t)(didx ? didx[m] : m) * d;
#pragma unroll
    for (int q = 0; q < 8; ++q) {
      const int c = lane + 32 * q;
      if (c < d) {
        const float v = rs * (g[q] - s1 - xh[q] * s2);
        a_x[q] += v;
        o[c] = accum ? o[c] + v : v;
      }
    }
  }
  __shared__ float red[WPB][3][256];
  const int w = threadIdx.x >> 5;
#pragma unroll
  for (int q = 0; q < 8; ++q) {
    red[w][0][lane + 32 * q] = a_g[q]; red[w][1][lane + 32 * q] = a_b[q]; red[w][2][lane + 32 * q] = a_x[q];
  }
  __syncthreads();
  for (int c = threadIdx.x; c < d; c += TPB) {
    float s0 = 0.f, s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int ww = 0; ww < WPB; ++ww) { s0 += red[ww][0][c]; s1 += red[ww][1][c]; s2 += red[ww][2][c]; }
    if (s0 != 0.f) atomicAdd(dgamma + c, s0);
    if (s1 != 0.f) atomicAdd(dbeta + c, s1);
    if (dbias && s2 != 0.f) atomicAdd(dbias + c, s2);
  }
}

// ---- 16-byte forms of the two LayerNorm kernels (d % 4 == 0): a lane owns the float4 column groups lane + 32 q (Q = ceil(d / 128)),
// a warp works on LNR rows at a time with every load issued before the first use.  The scalar forms above issue four times the
// instructions and keep one 4-byte load per lane in flight (2.5 TB/s on the grouped launches).
constexpr int LNR = 2;
template <int Q>
__global__ void __launch_bounds__(TPB) res_ln_fwd4_kernel(const float* __restrict__ res, GIdx ri, const float* __restrict__ add, PtrL biasl,
                                                          PtrL gammal, PtrL betal, float* __restrict__ y, float* __restrict__ xhat,
                                                          float* __restrict__ rstd, int d, float eps, Grp grp) {
  pdl_grid_sync();
  const int gi = blockIdx.y;
  const int M = grp_count(grp, gi);
  const size_t r0 = pick(grp.roff, gi);
  const int* ridx = pick(ri.idx.p, gi);
  if (res) res += (size_t)(ridx ? pick(ri.roff, gi) : r0) * d;
  if (add) add += r0 * d;
  y += r0 * d; xhat += r0 * d; rstd += r0;
  const float* __restrict__ bias = pick(biasl.p, gi); const float* __restrict__ gamma = pick(gammal.p, gi);
  const float* __restrict__ beta = pick(betal.p, gi);
  const int lane = threadIdx.x & 31;
  const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
  float4 b4[Q], g4[Q], e4[Q];
#pragma unroll
  for (int q = 0; q < Q; ++q) {
    const int c = 4 * lane + 128 * q;
    b4[q] = (bias && c < d) ? ld4(bias + c) : z4;
    g4[q] = c < d ? ld4(gamma + c) : z4;
    e4[q] = c < d ? ld4(beta + c) : z4;
  }
  const int stride = gridDim.x * WPB;
  const float inv_d = 1.f / d;
  for (int m0 = blockIdx.x * WPB + (threadIdx.x >> 5); m0 < M; m0 += LNR * stride) {
    float4 v[LNR][Q];
#pragma unroll
    for (int r = 0; r < LNR; ++r) {
      const int m = m0 + r * stride;
#pragma unroll
      for (int q = 0; q < Q; ++q) {
        const int c = 4 * lane + 128 * q;
        float4 x = z4;
        if (m < M && c < d) {
          if (res) x = ld4(res + (size_t)(ridx ? ridx[m] : m) * d + c);
          if (add) { const float4 a = ld4(add + (size_t)m * d + c); x.x += a.x; x.y += a.y; x.z += a.z; x.w += a.w; }
          x.x += b4[q].x; x.y += b4[q].y; x.z += b4[q].z; x.w += b4[q].w;
        }
        v[r][q] = x;
      }
    }
#pragma unroll
    for (int r = 0; r < LNR; ++r) {
      const int m = m0 + r * stride;
      if (m >= M) continue;                                  // warp-uniform
      float s = 0.f;
#pragma unroll
      for (int q = 0; q < Q; ++q) s += (v[r][q].x + v[r][q].y) + (v[r][q].z + v[r][q].w);
      const float mean = warp_sum(s) * inv_d;
      float var = 0.f;
#pragma unroll
      for (int q = 0; q < Q; ++q) {
        if (4 * lane + 128 * q < d) {
          const float tx = v[r][q].x - mean, ty = v[r][q].y - mean, tz = v[r][q].z - mean, tw = v[r][q].w - mean;
          var += (tx * tx + ty * ty) + (tz * tz + tw * tw);
        }
      }
      var = warp_sum(var) * inv_d;
      const float rs = rsqrtf(var + eps);
      if (lane == 0) rstd[m] = rs;
#pragma unroll
      for (int q = 0; q < Q; ++q) {
        const int c = 4 * lane + 128 * q;
        if (c < d) {
          const float4 xh = make_float4((v[r][q].x - mean) * rs, (v[r][q].y - mean) * rs, (v[r][q].z - mean) * rs, (v[r][q].w - mean) * rs);
          st4(xhat + (size_t)m * d + c, xh);
          st4(y + (size_t)m * d + c, make_float4(xh.x * g4[q].x + e4[q].x, xh.y * g4[q].y + e4[q].y, xh.z * g4[q].z + e4[q].z,
                                                  xh.w * g4[q].w + e4[q].w));
        }
      }
    }
  }
}

template <int Q>
__global__ void __launch_bounds__(TPB) ln_bwd4_kernel(const float* __restrict__ dy, const float* __restrict__ dy2,
                                                      const float* __restrict__ xhat, const float* __restrict__ rstd, PtrL gammal,
                                                      float* __restrict__ dx, GIdx di, int accum, MutL dgammal, MutL dbetal, MutL dbiasl,
                                                      int d, Grp grp) {
  pdl_grid_sync();
  const int gi = blockIdx.y;
  const int M = grp_count(grp, gi);
  const size_t r0 = pick(grp.roff, gi);
  const int* didx = pick(di.idx.p, gi);
  dy += r0 * d; if (dy2) dy2 += r0 * d;
  xhat += r0 * d; rstd += r0;
  dx += (size_t)(didx ? pick(di.roff, gi) : r0) * d;
  const float* __restrict__ gamma = pick(gammal.p, gi);
  float* dgamma = pick(dgammal.p, gi); float* dbeta = pick(dbetal.p, gi); float* dbias = pick(dbiasl.p, gi);
  const int lane = threadIdx.x & 31;
  const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
  float4 a_g[Q], a_b[Q], a_x[Q], g4[Q];
#pragma unroll
  for (int q = 0; q < Q; ++q) { a_g[q] = a_b[q] = a_x[q] = z4; const int c = 4 * lane + 128 * q; g4[q] = c < d ? ld4(gamma + c) : z4; }
  const int stride = gridDim.x * WPB;
  const float inv_d = 1.f / d;
  for (int m0 = blockIdx.x * WPB + (threadIdx.x >> 5); m0 < M; m0 += LNR * stride) {
    float4 dv[LNR][Q], xh[LNR][Q], ov[LNR][Q];
    float rs[LNR];
#pragma unroll
    for (int r = 0; r < LNR; ++r) {
      const int m = m0 + r * stride;
      rs[r] = 0.f;
#pragma unroll
      for (int q = 0; q < Q; ++q) {
        const int c = 4 * lane + 128 * q;
        dv[r][q] = xh[r][q] = ov[r][q] = z4;
        if (m < M && c < d) {
          dv[r][q] = ld4(dy + (size_t)m * d + c);
          if (dy2) { const float4 a = ld4(dy2 + (size_t)m * d + c); dv[r][q].x += a.x; dv[r][q].y += a.y; dv[r][q].z += a.z; dv[r][q].w += a.w; }
          xh[r][q] = ld4(xhat + (size_t)m * d + c);
          if (accum) ov[r][q] = ld4(dx + (size_t)(didx ? didx[m] : m) * d + c);
        }
      }
      if (m < M) rs[r] = rstd[m];
    }
#pragma unroll
    for (int r = 0; r < LNR; ++r) {
      const int m = m0 + r * stride;
      if (m >= M) continue;                                  // warp-uniform
      float4 g[Q];
      float s1 = 0.f, s2 = 0.f;
#pragma unroll
      for (int q = 0; q < Q; ++q) {
        const float4 a = dv[r][q], x = xh[r][q];
        a_g[q].x += a.x * x.x; a_g[q].y += a.y * x.y; a_g[q].z += a.z * x.z; a_g[q].w += a.w * x.w;
        a_b[q].x += a.x; a_b[q].y += a.y; a_b[q].z += a.z; a_b[q].w += a.w;
        g[q] = make_float4(a.x * g4[q].x, a.y * g4[q].y, a.z * g4[q].z, a.w * g4[q].w);
        s1 += (g[q].x + g[q].y) + (g[q].z + g[q].w);
        s2 += (g[q].x * x.x + g[q].y * x.y) + (g[q].z * x.z + g[q].w * x.w);
      }
      s1 = warp_sum(s1) * inv_d; s2 = warp_sum(s2) * inv_d;
      float* o = dx + (size_t)(didx ? didx[m] : m) * d;
#pragma unroll
      for (int q = 0; q < Q; ++q) {
        const int c = 4 * lane + 128 * q;
        if (c < d) {
          const float4 x = xh[r][q];
          const float4 vv = make_float4(rs[r] * (g[q].x - s1 - x.x * s2), rs[r] * (g[q].y - s1 - x.y * s2), rs[r] * (g[q].z - s1 - x.z * s2),
                                        rs[r] * (g[q].w - s1 - x.w * s2));
          a_x[q].x += vv.x; a_x[q].y += vv.y; a_x[q].z += vv.z; a_x[q].w += vv.w;
          st4(o + c, make_float4(ov[r][q].x + vv.x, ov[r][q].y + vv.y, ov[r][q].z + vv.z, ov[r][q].w + vv.w));
        }
      }
    }
  }
  __shared__ float4 red[WPB][3][32 * Q];
  const int w = threadIdx.x >> 5;
#pragma unroll
  for (int q = 0; q < Q; ++q) { red[w][0][lane + 32 * q] = a_g[q]; red[w][1][lane + 32 * q] = a_b[q]; red[w][2][lane + 32 * q] = a_x[q]; }
  __syncthreads();
  for (int e = threadIdx.x; e < 3 * 32 * Q; e += TPB) {
    const int k = e / (32 * Q), c4 = e % (32 * Q), c = 4 * c4;
    float* o = k == 0 ? dgamma : k == 1 ? dbeta : dbias;
    if (c >= d || o == nullptr) continue;
    float4 t = red[0][k][c4];
#pragma unroll
    for (int ww = 1; ww < WPB; ++ww) { const float4 u = red[ww][k][c4]; t.x += u.x; t.y += u.y; t.z += u.z; t.w += u.w; }
    atomicAdd(o + c, t.x); atomicAdd(o + c + 1, t.y); atomicAdd(o + c + 2, t.z); atomicAdd(o + c + 3, t.w);
  }
}

// relu backward in place: dH *= (H > 0)
// dH *= (H > 0); optionally dbias[c] += sum_m dH[m, c] in the same pass (the bias gradient of the layer that fed the ReLU).
// ld % 4 == 0; a thread keeps one float4 column group and walks rows, so the column sums need one shared-memory combine
// and one atomic per column per block.
// With `outp` the masked gradient is written as tensor-core operand planes (global rows of the concatenated buffer, zero rows
// up to round_up(longest group, 64) like split_planes_grp_kernel) instead of back to dH: its only consumers are contractions.
__global__ void __launch_bounds__(256) relu_bwd_kernel(float* __restrict__ dH, const float* __restrict__ H, int ld, Grp grp, MutL dbiasl,
                                                       Planes outp, int want_lo, float scale) {
  pdl_grid_sync();
  const int M = grp_count(grp, blockIdx.y);
  const int r0 = pick(grp.roff, blockIdx.y);
  dH += (size_t)r0 * ld; H += (size_t)r0 * ld;
  float* dbias = pick(dbiasl.p, blockIdx.y);
  int Mfill = M;
  if (outp.p) {
    int mx = 0;
#pragma unroll
    for (int g = 0; g < kMaxGroups; ++g) {
      const int c = g < grp.n ? (grp.cnt[g] ? *grp.cnt[g] : grp.hcnt[g]) : 0;
      mx = c > mx ? c : mx;
    }
    Mfill = (mx + 63) & ~63;
  }
  const int ngrp = ld >> 2;                               // float4 groups per row
  const int rpb = 256 / ngrp > 0 ? 256 / ngrp : 1;        // rows per block iteration (ld <= 1024)
  const int g = threadIdx.x % ngrp, rsub = threadIdx.x / ngrp;
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
  if (rsub < rpb) {
#pragma unroll 4
    for (int m = blockIdx.x * rpb + rsub; m < Mfill; m += gridDim.x * rpb) {
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (m < M) {
        float4* dp = reinterpret_cast<float4*>(dH + (size_t)m * ld) + g;
        const float4 h = __ldg(reinterpret_cast<const float4*>(H + (size_t)m * ld) + g);
        v = *dp;
        // (scale != 1: H went through dropout after the relu -- its zeros cover both masks, the survivors carry 1 / (1 - p))
        v.x = h.x > 0.f ? v.x * scale : 0.f; v.y = h.y > 0.f ? v.y * scale : 0.f; v.z = h.z > 0.f ? v.z * scale : 0.f;
        v.w = h.w > 0.f ? v.w * scale : 0.f;
        if (!outp.p) *dp = v;
        acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
      }
      if (outp.p) planes_store4(outp, r0 + m, 4 * g, v, want_lo != 0);
    }
  }
  if (!dbias) return;
  __shared__ float4 part[256];
  part[threadIdx.x] = acc;
  __syncthreads();
  if (threadIdx.x < ngrp) {
    float4 t = part[threadIdx.x];
    for (int k = 1; k < rpb; ++k) { const float4 u = part[k * ngrp + threadIdx.x]; t.x += u.x; t.y += u.y; t.z += u.z; t.w += u.w; }
    float* o = dbias + 4 * threadIdx.x;
    atomicAdd(o, t.x); atomicAdd(o + 1, t.y); atomicAdd(o + 2, t.z); atomicAdd(o + 3, t.w);
  }
}

__global__ void __launch_bounds__(256) dropout_fwd_kernel(const float* __restrict__ in, float* __restrict__ out, int cols, PtrL biasl,
                                                          Grp grp, GIdx hrow, Drop dr) {
  pdl_grid_sync();
  const int gi = blockIdx.y;
  const int M = grp_count(grp, gi);
  const size_t r0 = (size_t)pick(grp.roff, gi);
  const float* bias = pick(biasl.p, gi);
  const int* hr = pick(hrow.idx.p, gi);
  const uint64_t key = pick(dr.key, gi);
  const int n4 = cols >> 2;
  const size_t total = (size_t)M * n4;
  for (size_t e = (size_t)blockIdx.x * 256 + threadIdx.x; e < total; e += (size_t)gridDim.x * 256) {
    const int m = (int)(e / n4), c = 4 * (int)(e % n4);
    const uint32_t row = hr ? (uint32_t)hr[m] : (uint32_t)m;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (in) v = ld4(in + (r0 + m) * cols + c);
    if (bias) { const float4 b = ld4(bias + c); v.x += b.x; v.y += b.y; v.z += b.z; v.w += b.w; }
    v.x = drop_keep(key, row, c, dr.thresh) ? v.x * dr.inv_keep : 0.f;
    v.y = drop_keep(key, row, c + 1, dr.thresh) ? v.y * dr.inv_keep : 0.f;
    v.z = drop_keep(key, row, c + 2, dr.thresh) ? v.z * dr.inv_keep : 0.f;
    v.w = drop_keep(key, row, c + 3, dr.thresh) ? v.w * dr.inv_keep : 0.f;
    st4(out + (r0 + m) * cols + c, v);
  }
}

__global__ void __launch_bounds__(256) dropout_bwd_kernel(const float* __restrict__ g, float* __restrict__ out, int cols, MutL dbiasl,
                                                          Grp grp, GIdx hrow, Drop dr) {
  pdl_grid_sync();
  const int gi = blockIdx.y;
  const int M = grp_count(grp, gi);
  const size_t r0 = (size_t)pick(grp.roff, gi);
  float* dbias = pick(dbiasl.p, gi);
  const int* hr = pick(hrow.idx.p, gi);
  const uint64_t key = pick(dr.key, gi);
  const int ngrp = cols >> 2;
  const int rpb = 256 / ngrp > 0 ? 256 / ngrp : 1;
  const int cg = threadIdx.x % ngrp, rsub = threadIdx.x / ngrp;
  const int c = 4 * cg;
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
  if (rsub < rpb) {
    for (int m = blockIdx.x * rpb + rsub; m < M; m += gridDim.x * rpb) {
      const uint32_t row = hr ? (uint32_t)hr[m] : (uint32_t)m;
      float4 v = ld4(g + (r0 + m) * cols + c);
      v.x = drop_keep(key, row, c, dr.thresh) ? v.x * dr.inv_keep : 0.f;
      v.y = drop_keep(key, row, c + 1, dr.thresh) ? v.y * dr.inv_keep : 0.f;
      v.z = drop_keep(key, row, c + 2, dr.thresh) ? v.z * dr.inv_keep : 0.f;
      v.w = drop_keep(key, row, c + 3, dr.thresh) ? v.w * dr.inv_keep : 0.f;
      if (out) st4(out + (r0 + m) * cols + c, v);
      acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
    }
  }
  if (!dbias) return;
  __shared__ float4 part[256];
  part[threadIdx.x] = acc;
  __syncthreads();
  if (threadIdx.x < ngrp) {
    float4 t = part[threadIdx.x];
    for (int k = 1; k < rpb; ++k) { const float4 u = part[k * ngrp + threadIdx.x]; t.x += u.x; t.y += u.y; t.z += u.z; t.w += u.w; }
    float* o = dbias + 4 * threadIdx.x;
    atomicAdd(o, t.x); atomicAdd(o + 1, t.y); atomicAdd(o + 2, t.z); atomicAdd(o + 3, t.w);
  }
}

__global__ void zero_rows_kernel(float* __restrict__ X, size_t ld, int M_host, const int* M_dev, int M_extra) {
  pdl_grid_sync();
  const size_t n = (size_t)(row_count(M_host, M_dev) + M_extra) * ld;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) X[i] = 0.f;
}

// dst[idx[m], :] += src[m, :]   (m < B rows; idx unique)
__global__ void __launch_bounds__(TPB) scatter_add_rows_kernel(const float* __restrict__ src, GIdx di, float* __restrict__ dst, int d, Grp grp) {
  pdl_grid_sync();
  const int B = grp_count(grp, blockIdx.y);
  src += (size_t)pick(grp.roff, blockIdx.y) * d;
  dst += (size_t)pick(di.roff, blockIdx.y) * d;
  const int* idx = pick(di.idx.p, blockIdx.y);
  ROW_LOOP(m, B) {
    float* o = dst + (size_t)idx[m] * d;
    for (int c = lane; c < d; c += 32) o[c] += src[(size_t)m * d + c];
  }
}

// ---------------------------------------------------------------------------------------------------------
// head: F = z (+ Fprev);  logits = F.Wc^T + bc                         (paths.py:130-139)
__global__ void __launch_bounds__(TPB) head_fwd_kernel(const float* __restrict__ z, const float* __restrict__ Fprev,
                                                       const float* __restrict__ Wc, const float* __restrict__ bc,
                                                       float* __restrict__ F, float* __restrict__ logits, int d, int nl, int B) {
  pdl_grid_sync();
  ROW_LOOP(s, B) {
    float v[8];
#pragma unroll
    for (int q = 0; q < 8; ++q) {
      const int c = lane + 32 * q;
      v[q] = 0.f;
      if (c < d) { v[q] = z[(size_t)s * d + c] + (Fprev ? Fprev[(size_t)s * d + c] : 0.f); F[(size_t)s * d + c] = v[q]; }
    }
    for (int k = 0; k < nl; ++k) {
      float acc = 0.f;
#pragma unroll
      for (int q = 0; q < 8; ++q) { const int c = lane + 32 * q; if (c < d) acc += v[q] * Wc[(size_t)k * d + c]; }
      acc = warp_sum(acc);
      if (lane == 0) logits[(size_t)s * nl + k] = acc + bc[k];
    }
  }
}

// dF = dF_in + dlogits.Wc      (dWc, dbc come from gemm_tn / colsum_accum in the engine)
__global__ void __launch_bounds__(TPB) head_bwd_kernel(const float* __restrict__ dlogits, const float* __restrict__ dF_in,
                                                       const float* __restrict__ Wc, float* __restrict__ dF, int d, int nl, int B) {
  pdl_grid_sync();
  ROW_LOOP(s, B) {
#pragma unroll
    for (int q = 0; q < 8; ++q) {
      const int c = lane + 32 * q;
      if (c < d) {
        float g = dF_in ? dF_in[(size_t)s * d + c] : 0.f;
        if (dlogits)
          for (int k = 0; k < nl; ++k) g += dlogits[(size_t)s * nl + k] * Wc[(size_t)k * d + c];
        dF[(size_t)s * d + c] = g;
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------------------
// survival NLL (utils.py:283-305, alpha=0.4, eps=1e-7, mean over the batch) on hazards = sigmoid(logits) (utils.py:269),
// and its gradient w.r.t. the logits.  task 1: cross-entropy (utils.py:277).
__global__ void loss_kernel(const float* __restrict__ logits, const int* __restrict__ ybin, const int* __restrict__ cens,
                            float* __restrict__ hazards, float* __restrict__ loss_per, float* __restrict__ risk,
                            float* __restrict__ dlogits, int nl, int B, int task, float alpha, float eps,
                            float scale /* 1 / global batch */) {
  pdl_grid_sync();
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= B) return;
  const float* lg = logits + (size_t)s * nl;
  const int y = ybin[s];
  if (task == 0) {
    const float c = (float)cens[s];
    float surv_y = 1.f, surv_y1 = 1.f, hy = 0.f;
    float surv = 1.f, rk = 0.f;                                         // eval.py:60-61: risk = -sum(cumprod(1 - hazards))
    for (int k = 0; k < nl; ++k) {
      const float h = sigmoidf_(lg[k]);
      hazards[(size_t)s * nl + k] = h;
      surv *= (1.f - h); rk -= surv;
      if (k < y) surv_y *= (1.f - h);
      if (k <= y) surv_y1 *= (1.f - h);
      if (k == y) hy = h;
    }
    const float unc = -(1.f - c) * (logf(fmaxf(surv_y, eps)) + logf(fmaxf(hy, eps)));
    const float cen = -c * logf(fmaxf(surv_y1, eps));
    loss_per[s] = (1.f - alpha) * (cen + unc) + alpha * unc;
    if (risk) risk[s] = rk;
    // d loss / d unc = 1 ; d loss / d cen = (1 - alpha)
    if (dlogits) {
      const float w_sy = (surv_y > eps) ? -(1.f - c) : 0.f;              // coefficient of log S_pad[y]
      const float w_hy = (hy > eps) ? -(1.f - c) : 0.f;                  // coefficient of log h[y]
      const float w_sy1 = (surv_y1 > eps) ? -(1.f - alpha) * c : 0.f;    // coefficient of log S_pad[y+1]
      for (int k = 0; k < nl; ++k) {
        const float h = hazards[(size_t)s * nl + k];
        float dh = 0.f;                                                   // d/dh_k
        if (k < y) dh += w_sy * (-1.f / (1.f - h));
        if (k <= y) dh += w_sy1 * (-1.f / (1.f - h));
        if (k == y) dh += w_hy / h;
        dlogits[(size_t)s * nl + k] = scale * dh * h * (1.f - h);
      }
    }
  } else {
    float mx = -INFINITY;
    for (int k = 0; k < nl; ++k) mx = fmaxf(mx, lg[k]);
    float se = 0.f;
    for (int k = 0; k < nl; ++k) se += expf(lg[k] - mx);
    const float lse = mx + logf(se);
    loss_per[s] = lse - lg[y];
    if (risk) risk[s] = 0.f;                                            // no risk score for the classification task
    for (int k = 0; k < nl; ++k) {
      hazards[(size_t)s * nl + k] = lg[k];
      if (dlogits) dlogits[(size_t)s * nl + k] = scale * (expf(lg[k] - lse) - (k == y ? 1.f : 0.f));
    }
  }
}

__global__ void loss_reduce_kernel(const float* __restrict__ loss_per, float* __restrict__ loss, int B, float scale) {
  pdl_grid_sync();
  __shared__ float red[256];
  float s = 0.f;
  for (int i = threadIdx.x; i < B; i += 256) s += loss_per[i];
  red[threadIdx.x] = s;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) { if (threadIdx.x < o) red[threadIdx.x] += red[threadIdx.x + o]; __syncthreads(); }
  if (threadIdx.x == 0) loss[0] = red[0] * scale;
}

inline int row_grid(int M) { int g = (M + WPB - 1) / WPB; if (g > 8 * kNumSMs) g = 8 * kNumSMs; return g < 1 ? 1 : g; }

}  // namespace

// ---- shared-parent form of the gate contraction (interface.py:49: gates = W [x || h_parent]) -------------------------------------
// The (up to) four children of a selected patch see the same h_parent, so W_h h_parent is one product per SELECTED PARENT ("slot"
// s * kcap + position in the slide's selected list), not per child.  Index tables of one level pair, one launch:
//   slot4[m]   (rows of the parent level)  = (slot of row m, -1, -1, -1) if any child of m survived, else all -1 -- the layout
//                                            lstm_point2_bwd's child gather reads, so it picks up ONE row of the slot-shaped dH table
//   sel_rows[i] (slots)                    = the parent row of slot i, -1 for the holes of slides with fewer selected rows
//   add_idx[r] (rows of the child level)   = slot of row r's parent, -1 for the rows of a fallback slide (no parent, slide.py:336-352)
__global__ void __launch_bounds__(256) parent_slots_kernel(const int* __restrict__ child_of, const int* __restrict__ count, int cap,
                                                           const int* __restrict__ sel, const int* __restrict__ sel_count, int B, int kcap,
                                                           const int* __restrict__ n_row_slide, const int* __restrict__ n_row_parent,
                                                           const int* __restrict__ n_parent_local, const int* __restrict__ n_count, int n_cap,
                                                           int* __restrict__ slot4, int* __restrict__ sel_rows, int* __restrict__ add_idx) {
  pdl_grid_sync();
  const int M = min(count[0], cap), Mn = min(n_count[0], n_cap), slots = B * kcap;
  const int total = max(max(M, Mn), slots);
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    if (i < M) {
      const int4 ch = __ldg(reinterpret_cast<const int4*>(child_of) + i);
      const int kid = ch.x >= 0 ? ch.x : ch.y >= 0 ? ch.y : ch.z >= 0 ? ch.z : ch.w;
      const int slot = (kid >= 0 && kid < Mn) ? n_row_slide[kid] * kcap + n_parent_local[kid] : -1;
      reinterpret_cast<int4*>(slot4)[i] = make_int4(slot, -1, -1, -1);
    }
    if (i < slots) {
      const int sl = i / kcap, j = i - sl * kcap;
      sel_rows[i] = j < sel_count[sl] ? sel[i] : -1;
    }
    if (i < Mn) add_idx[i] = n_row_parent[i] >= 0 ? n_row_slide[i] * kcap + n_parent_local[i] : -1;
  }
}
// dGsum[slot] = sum of dG over the surviving children of the slot's parent row, planes in, planes out (hi + lo summed in fp32 and
// split again); every slot row is written (holes and the padding rows up to a multiple of 64 as zeros).
__global__ void __launch_bounds__(256) child_sum_planes_kernel(const Planes src, const int* __restrict__ child_of, const int* __restrict__ sel_rows,
                                                               int slots, int cols, Planes dst, int want_lo) {
  pdl_grid_sync();
  const int rows_w = planes_rows_to_write(dst, slots);
  for (int i = blockIdx.x; i < rows_w; i += gridDim.x) {
    int4 ch = make_int4(-1, -1, -1, -1);
    if (i < slots) {
      const int pr = __ldg(sel_rows + i);
      if (pr >= 0) ch = __ldg(reinterpret_cast<const int4*>(child_of) + pr);
    }
    const int kid[4] = {ch.x, ch.y, ch.z, ch.w};
    for (int k = 4 * threadIdx.x; k < cols; k += 4 * blockDim.x) {
      uint2 hv[4], lv[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        hv[q] = lv[q] = make_uint2(0u, 0u);
        if (kid[q] >= 0) {
          hv[q] = __ldg(reinterpret_cast<const uint2*>(src.p + (size_t)kid[q] * src.ldp + k));
          if (want_lo) lv[q] = __ldg(reinterpret_cast<const uint2*>(src.p + ((size_t)src.rows + kid[q]) * src.ldp + k));
        }
      }
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        // bf16 -> fp32 is a 16-bit shift; hi + lo is exact in fp32 (lo has at most 8 significant bits below hi's last)
        acc.x += __uint_as_float(hv[q].x << 16) + __uint_as_float(lv[q].x << 16);
        acc.y += __uint_as_float(hv[q].x & 0xffff0000u) + __uint_as_float(lv[q].x & 0xffff0000u);
        acc.z += __uint_as_float(hv[q].y << 16) + __uint_as_float(lv[q].y << 16);
        acc.w += __uint_as_float(hv[q].y & 0xffff0000u) + __uint_as_float(lv[q].y & 0xffff0000u);
      }
      planes_store4(dst, i, k, acc, want_lo != 0);
    }
  }
}

// ---- host launchers -----------------------------------------------------------------------------------------
int gather_rows(const float* const* grid_ptr, const int* gh, int level, int L, const int* slide_ids, const int* row_slide,
                const int* row_xy, float* X, int D, int M_host, const int* M_dev, cudaStream_t st, int bf16, const Planes* XP, int want_lo) {
  PATHS_REQUIRE(D % 4 == 0, "gather_rows: D must be a multiple of 4");
  if (M_host <= 0) return 0;
  const Planes xp = XP ? *XP : no_planes();
  PATHS_REQUIRE(!xp.p || (D <= xp.ldp && M_host <= xp.rows), "gather_rows: operand planes too small");
  const int g = row_grid(M_host + (xp.p ? 64 : 0));
  if (bf16) pdl(gather_rows_kernel<true>, g, TPB, 0, st)(grid_ptr, gh, level, L, slide_ids, row_slide, row_xy, X, D, M_host, M_dev, xp, want_lo);
  else pdl(gather_rows_kernel<false>, g, TPB, 0, st)(grid_ptr, gh, level, L, slide_ids, row_slide, row_xy, X, D, M_host, M_dev, xp, want_lo);
  PATHS_LAUNCH_CK();
  return 0;
}
int lstm_point1_fwd(float* G, const float* Cprev, const int* parent, float* Cn, const Planes& CnP, int want_lo, int activated, int D, int C,
                    int M_host,
                    const int* M_dev, cudaStream_t st) {
  PATHS_REQUIRE(D % 4 == 0 && C % 4 == 0, "lstm: D and C must be multiples of 4");
  if (M_host <= 0) return 0;
  pdl(lstm_point1_fwd_kernel, row_grid(M_host + 64), TPB, 0, st)(G, Cprev, parent, Cn, CnP, want_lo, activated, D, C, M_host, M_dev);
  PATHS_LAUNCH_CK();
  return 0;
}
int lstm_point2_fwd(float* Pt, const float* G, const float* X, float* H, float* Y, const Planes& YP, int want_lo, int D, int C, int M_host,
                    const int* M_dev, cudaStream_t st) {
  if (M_host <= 0) return 0;
  pdl(lstm_point2_fwd_kernel, row_grid(M_host + 64), TPB, 0, st)(Pt, G, X, H, Y, YP, want_lo, D, C, M_host, M_dev);
  PATHS_LAUNCH_CK();
  return 0;
}
int lstm_point2_bwd(const float* dY, const float* dHc, const int* child_of, const float* Pt, const float* G, float* dP, const Planes& dPP,
                    float* dG, const Planes& dGP, float* db_mo, float* db_o, int want_lo, int D, int C, int M_host, const int* M_dev, cudaStream_t st) {
  PATHS_REQUIRE(D <= 4 * CG * 256, "lstm backward: patch_embed_dim above %d unsupported", 4 * CG * 256);
  if (M_host <= 0) return 0;
  // one resident wave: the blocks stride over the rows and end with 2 D atomics each
  static int occ[2] = {0, 0};
  const int wide = D > 1024 ? 1 : 0;
  if (!occ[wide]) {
    if (wide) PATHS_CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ[1], lstm_point2_bwd_kernel<CG>, 256, 0));
    else PATHS_CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ[0], lstm_point2_bwd_kernel<1>, 256, 0));
    if (occ[wide] < 1) occ[wide] = 1;
    if (occ[wide] > 4) occ[wide] = 4;
  }
  int g = (M_host + 64 + P2R - 1) / P2R; if (g > occ[wide] * kNumSMs) g = occ[wide] * kNumSMs;
  if (wide) pdl(lstm_point2_bwd_kernel<CG>, g, 256, 0, st)(dY, dHc, child_of, Pt, G, dP, dPP, dG, dGP, db_mo, db_o, want_lo, D, C, M_host, M_dev);
  else pdl(lstm_point2_bwd_kernel<1>, g, 256, 0, st)(dY, dHc, child_of, Pt, G, dP, dPP, dG, dGP, db_mo, db_o, want_lo, D, C, M_host, M_dev);
  PATHS_LAUNCH_CK();
  return 0;
}
int lstm_point1_bwd(const float* dCn, const float* dCc, const int* child_of, const float* G, const float* Cprev, const int* parent, float* dG,
                    const Planes& dGP, float* dCrow, float* db_g, int want_lo, int D, int C, int M_host, const int* M_dev, cudaStream_t st) {
  PATHS_REQUIRE(C / 4 <= 256, "lstm backward: hierarchical_ctx_mlp_hidden_dim above 1024 unsupported");
  if (M_host <= 0) return 0;
  const int R = 256 / (C / 4);
  int g = (M_host + 64 + R - 1) / R; if (g > 3 * kNumSMs) g = 3 * kNumSMs;
  pdl(lstm_point1_bwd_kernel, g, 256, 0, st)(dCn, dCc, child_of, G, Cprev, parent, dG, dGP, dCrow, db_g, want_lo, D, C, M_host, M_dev);
  PATHS_LAUNCH_CK();
  return 0;
}
int parent_slots(const int* child_of, const int* count, int cap, const int* sel, const int* sel_count, int B, int kcap, const int* n_row_slide,
                 const int* n_row_parent, const int* n_parent_local, const int* n_count, int n_cap, int* slot4, int* sel_rows, int* add_idx,
                 cudaStream_t st) {
  int total = cap > n_cap ? cap : n_cap;
  if (B * kcap > total) total = B * kcap;
  int g = (total + 255) / 256; if (g > 4 * kNumSMs) g = 4 * kNumSMs;
  pdl(parent_slots_kernel, g, 256, 0, st)(child_of, count, cap, sel, sel_count, B, kcap, n_row_slide, n_row_parent, n_parent_local, n_count, n_cap,
                                          slot4, sel_rows, add_idx);
  PATHS_LAUNCH_CK();
  return 0;
}
int child_sum_planes(const Planes& src, const int* child_of, const int* sel_rows, int slots, int cols, const Planes& dst, int want_lo,
                     cudaStream_t st) {
  PATHS_REQUIRE(src.p && dst.p && cols % 4 == 0 && cols <= src.ldp && cols <= dst.ldp && slots <= dst.rows, "child_sum_planes: bad planes");
  if (slots <= 0) return 0;
  int g = slots + 64; if (g > 8 * kNumSMs) g = 8 * kNumSMs;
  pdl(child_sum_planes_kernel, g, 256, 0, st)(src, child_of, sel_rows, slots, cols, dst, want_lo);
  PATHS_LAUNCH_CK();
  return 0;
}
int parent_sum(const int* child_of, const float* dHrow, const float* dCrow, float* dHc, float* dCc, int D, int C, int M_host,
               const int* M_dev, cudaStream_t st) {
  if (M_host <= 0) return 0;
  pdl(parent_sum_kernel, row_grid(M_host), TPB, 0, st)(child_of, dHrow, dCrow, dHc, dCc, D, C, M_host, M_dev);
  PATHS_LAUNCH_CK();
  return 0;
}
int imp_tok_fwd(float* UV, const float* b1, const float* w2, const float* b2, const float* bin, const float* div, const int* row_slide,
                const int* row_xy, const int* seg, float* alpha, float* tok, int I, int d, int imp_mul, int pe_mode, int mode, int M_host,
                const int* M_dev, cudaStream_t st) {
  PATHS_REQUIRE(I <= 256 && d <= 256, "imp_tok: hidden dims above 256 are not supported (I=%d, d=%d)", I, d);
  if (M_host <= 0) return 0;
  const bool vec = mode == 0 && I % 4 == 0 && d % 4 == 0 &&
                   ((reinterpret_cast<uintptr_t>(UV) | reinterpret_cast<uintptr_t>(b1) | reinterpret_cast<uintptr_t>(w2) |
                     reinterpret_cast<uintptr_t>(bin) | reinterpret_cast<uintptr_t>(tok)) & 15) == 0;
  if (vec) {
    int g = (M_host + 2 * WPB - 1) / (2 * WPB);
    if (g > 8 * kNumSMs) g = 8 * kNumSMs;
    if (I <= 128 && d <= 128) pdl(imp_tok_fwd4_kernel<1>, g, TPB, 0, st)(UV, b1, w2, b2, bin, div, row_slide, row_xy, seg, alpha, tok, I, d, imp_mul, pe_mode, M_host, M_dev);
    else pdl(imp_tok_fwd4_kernel<2>, g, TPB, 0, st)(UV, b1, w2, b2, bin, div, row_slide, row_xy, seg, alpha, tok, I, d, imp_mul, pe_mode, M_host, M_dev);
  } else {
    pdl(imp_tok_fwd_kernel, row_grid(M_host), TPB, 0, st)(UV, b1, w2, b2, bin, div, row_slide, row_xy, seg, alpha, tok, I, d, imp_mul,
                                                         pe_mode, mode, M_host, M_dev);
  }
  PATHS_LAUNCH_CK();
  return 0;
}
int special_fill(const float* special, const int* seg, float* tok, int d, int B, cudaStream_t st) {
  pdl(special_fill_kernel, row_grid(B), TPB, 0, st)(special, seg, tok, d, B);
  PATHS_LAUNCH_CK();
  return 0;
}
int imp_tok_bwd(const float* dtok, const float* UV, const float* alpha, const float* w2, const int* row_slide, float* dUV, const Planes& dUVP,
                int want_lo, float* db1, float* dw2, float* db2, float* dbin, int I, int d, int imp_mul, int mode, int M_host, const int* M_dev,
                cudaStream_t st) {
  if (M_host <= 0) return 0;
  int g = row_grid(M_host + 64);
  const bool vec = mode == 0 && I % 4 == 0 && d % 4 == 0 && I <= 256 && d <= 256 && ((reinterpret_cast<uintptr_t>(dtok) | reinterpret_cast<uintptr_t>(UV) |
                   reinterpret_cast<uintptr_t>(w2) | reinterpret_cast<uintptr_t>(dUV)) & 15) == 0 && (!dUVP.p || dUVP.ldp % 4 == 0);
  if (vec) {
    const int wide = (I > 128 || d > 128) ? 1 : 0;
    static int occ[2] = {0, 0};
    if (!occ[wide]) {
      if (wide) PATHS_CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ[1], imp_tok_bwd4_kernel<2>, TPB, 0));
      else PATHS_CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ[0], imp_tok_bwd4_kernel<1>, TPB, 0));
      if (occ[wide] < 1) occ[wide] = 1;
    }
    g = (M_host + 64 + WPB * ITR - 1) / (WPB * ITR);
    if (g > occ[wide] * kNumSMs) g = occ[wide] * kNumSMs;     // one resident wave: each block ends with up to 3 * 256 atomics
    if (wide) pdl(imp_tok_bwd4_kernel<2>, g, TPB, 0, st)(dtok, UV, alpha, w2, row_slide, dUV, dUVP, want_lo, db1, dw2, db2, dbin, I, d, imp_mul, M_host, M_dev);
    else pdl(imp_tok_bwd4_kernel<1>, g, TPB, 0, st)(dtok, UV, alpha, w2, row_slide, dUV, dUVP, want_lo, db1, dw2, db2, dbin, I, d, imp_mul, M_host, M_dev);
    PATHS_LAUNCH_CK();
    return 0;
  }
  if (g > 2 * kNumSMs) g = 2 * kNumSMs;
  pdl(imp_tok_bwd_kernel, g, TPB, 0, st)(dtok, UV, alpha, w2, row_slide, dUV, dUVP, want_lo, db1, dw2, db2, dbin, I, d, imp_mul, mode, M_host,
                                        M_dev);
  PATHS_LAUNCH_CK();
  return 0;
}
int rnn_z_fwd(const float* X, const float* alpha, const float* hctx, float* Z, int D, int imp_mul, int M_host, const int* M_dev,
              cudaStream_t st) {
  if (M_host <= 0) return 0;
  pdl(rnn_z_fwd_kernel, row_grid(M_host), TPB, 0, st)(X, alpha, hctx, Z, D, imp_mul, M_host, M_dev);
  PATHS_LAUNCH_CK();
  return 0;
}
int rnn_alpha_bwd(float* dZ, const float* dZc, const float* X, const float* UV, const float* alpha, const float* w2, float* dUV, float* db1,
                  float* dw2, float* db2, int D, int I, int d, int imp_mul, int M_host, const int* M_dev, cudaStream_t st) {
  if (M_host <= 0) return 0;
  int g = row_grid(M_host); if (g > 2 * kNumSMs) g = 2 * kNumSMs;
  pdl(rnn_alpha_bwd_kernel, g, TPB, 0, st)(dZ, dZc, X, UV, alpha, w2, dUV, db1, dw2, db2, D, I, d, imp_mul, M_host, M_dev);
  PATHS_LAUNCH_CK();
  return 0;
}
int head_concat_fwd(const FList& fl, const float* Wc, const float* bc, float* logits, int d, int nl, int B, cudaStream_t st) {
  pdl(head_concat_fwd_kernel, row_grid(B), TPB, 0, st)(fl, Wc, bc, logits, d, nl, B);
  PATHS_LAUNCH_CK();
  return 0;
}
int head_concat_bwd(const float* dlogits, const float* Wc, const GList& gl, int d, int nl, int B, cudaStream_t st) {
  pdl(head_concat_bwd_kernel, row_grid(B), TPB, 0, st)(dlogits, Wc, gl, d, nl, B);
  PATHS_LAUNCH_CK();
  return 0;
}
int special_grad_grp(const float* dtok, const GIdx& seg, const MutL& dspecial, int d, int B, int ngroups, cudaStream_t st) {
  PATHS_REQUIRE(d <= 256, "special_grad: trans_dim above 256 is not supported (d=%d)", d);
  pdl(special_grad_kernel, dim3((B + 7) / 8, ngroups), 256, 0, st)(dtok, seg, dspecial, d, B);
  PATHS_LAUNCH_CK();
  return 0;
}
int special_grad(const float* dtok, const int* seg, float* dspecial, int d, int B, cudaStream_t st) {
  return special_grad_grp(dtok, gidx1(seg), mutl1(dspecial), d, B, 1, st);
}
int res_ln_fwd_grp(const float* res, const GIdx& ridx, const float* add, const PtrL& bias, const PtrL& gamma, const PtrL& beta, float* y,
                   float* xhat, float* rstd, int d, float eps, const Grp& grp, cudaStream_t st) {
  PATHS_REQUIRE(d <= 256, "layer norm width above 256 is not supported (d=%d)", d);
  const int cap = grp_max_rows(grp);
  if (cap <= 0) return 0;
  int g = row_grid(cap);
  if (grp.n > 1 && g > 2 * kNumSMs) g = 2 * kNumSMs;
  const bool al16 = ((reinterpret_cast<uintptr_t>(res) | reinterpret_cast<uintptr_t>(add) | reinterpret_cast<uintptr_t>(y) |
                      reinterpret_cast<uintptr_t>(xhat)) & 15) == 0;
  bool vec = d % 4 == 0 && al16;
  for (int i = 0; i < grp.n; ++i)
    vec = vec && ((reinterpret_cast<uintptr_t>(bias.p[i]) | reinterpret_cast<uintptr_t>(gamma.p[i]) | reinterpret_cast<uintptr_t>(beta.p[i])) & 15) == 0;
  if (vec) {
    g = (cap + WPB * LNR - 1) / (WPB * LNR);
    if (g > 4 * kNumSMs) g = 4 * kNumSMs;
    if (d <= 128) pdl(res_ln_fwd4_kernel<1>, dim3(g, grp.n), TPB, 0, st)(res, ridx, add, bias, gamma, beta, y, xhat, rstd, d, eps, grp);
    else pdl(res_ln_fwd4_kernel<2>, dim3(g, grp.n), TPB, 0, st)(res, ridx, add, bias, gamma, beta, y, xhat, rstd, d, eps, grp);
  } else {
    pdl(res_ln_fwd_kernel, dim3(g, grp.n), TPB, 0, st)(res, ridx, add, bias, gamma, beta, y, xhat, rstd, d, eps, grp);
  }
  PATHS_LAUNCH_CK();
  return 0;
}
int res_ln_fwd(const float* res, const int* ridx, const float* add, const float* bias, const float* gamma, const float* beta, float* y,
               float* xhat, float* rstd, int d, float eps, int M_host, const int* M_dev, int M_extra, cudaStream_t st) {
  PATHS_REQUIRE(M_extra == 0, "res_ln_fwd: M_extra is no longer supported");
  return res_ln_fwd_grp(res, gidx1(ridx), add, ptrl1(bias), ptrl1(gamma), ptrl1(beta), y, xhat, rstd, d, eps, grp_single(M_host, M_dev), st);
}
int ln_bwd_grp(const float* dy, const float* dy2, const float* xhat, const float* rstd, const PtrL& gamma, float* dx, const GIdx& didx, int accum,
               const MutL& dgamma, const MutL& dbeta, const MutL& dbias, int d, const Grp& grp, cudaStream_t st) {
  const int cap = grp_max_rows(grp);
  if (cap <= 0) return 0;
  int g = row_grid(cap);
  const int lim = grp.n > 1 ? kNumSMs : 2 * kNumSMs;
  if (g > lim) g = lim;
  const bool al16 = ((reinterpret_cast<uintptr_t>(dy) | reinterpret_cast<uintptr_t>(dy2) | reinterpret_cast<uintptr_t>(dx) |
                      reinterpret_cast<uintptr_t>(xhat)) & 15) == 0;
  bool vec = d % 4 == 0 && d <= 256 && al16;
  for (int i = 0; i < grp.n; ++i) vec = vec && (reinterpret_cast<uintptr_t>(gamma.p[i]) & 15) == 0;
  if (vec) {
    if (d <= 128) pdl(ln_bwd4_kernel<1>, dim3(g, grp.n), TPB, 0, st)(dy, dy2, xhat, rstd, gamma, dx, didx, accum, dgamma, dbeta, dbias, d, grp);
    else pdl(ln_bwd4_kernel<2>, dim3(g, grp.n), TPB, 0, st)(dy, dy2, xhat, rstd, gamma, dx, didx, accum, dgamma, dbeta, dbias, d, grp);
  } else {
    pdl(ln_bwd_kernel, dim3(g, grp.n), TPB, 0, st)(dy, dy2, xhat, rstd, gamma, dx, didx, accum, dgamma, dbeta, dbias, d, grp);
  }
  PATHS_LAUNCH_CK();
  return 0;
}
int ln_bwd(const float* dy, const float* dy2, const float* xhat, const float* rstd, const float* gamma, float* dx, const int* didx, int accum,
           float* dgamma, float* dbeta, float* dbias, int d, int M_host, const int* M_dev, int M_extra, cudaStream_t st) {
  PATHS_REQUIRE(M_extra == 0, "ln_bwd: M_extra is no longer supported");
  return ln_bwd_grp(dy, dy2, xhat, rstd, ptrl1(gamma), dx, gidx1(didx), accum, mutl1(dgamma), mutl1(dbeta), mutl1(dbias), d,
                    grp_single(M_host, M_dev), st);
}
int relu_bwd_grp(float* dH, const float* H, int ld, const Grp& grp, const MutL& dbias, cudaStream_t st, const Planes* outp, int want_lo,
                 float scale) {
  const int cap = grp_max_rows(grp);
  if (cap <= 0) return 0;
  PATHS_REQUIRE(ld % 4 == 0 && ld <= 1024, "relu_bwd: row width %d unsupported (multiple of 4, <= 1024)", ld);
  const int rpb = 256 / (ld / 4) > 0 ? 256 / (ld / 4) : 1;
  int g = (cap + rpb - 1) / rpb;
  const int lim = grp.n > 1 ? kNumSMs : 2 * kNumSMs;
  if (g > lim) g = lim;
  PATHS_REQUIRE(outp == nullptr || (outp->p && ld <= outp->ldp), "relu_bwd: output planes narrower than the rows");
  pdl(relu_bwd_kernel, dim3(g, grp.n), 256, 0, st)(dH, H, ld, grp, dbias, outp ? *outp : no_planes(), want_lo, scale);
  PATHS_LAUNCH_CK();
  return 0;
}
int dropout_fwd_grp(const float* in, float* out, int cols, const PtrL& bias, const Grp& grp, const GIdx& hrow, const Drop& dr, cudaStream_t st) {
  PATHS_REQUIRE(cols % 4 == 0 && cols > 0, "dropout: row width %d is not a multiple of 4", cols);
  const int cap = grp_max_rows(grp);
  if (cap <= 0) return 0;
  size_t g = ((size_t)cap * (cols >> 2) + 255) / 256;
  if (g > (size_t)8 * kNumSMs) g = 8 * kNumSMs;
  pdl(dropout_fwd_kernel, dim3((unsigned)g, grp.n), 256, 0, st)(in, out, cols, bias, grp, hrow, dr);
  PATHS_LAUNCH_CK();
  return 0;
}
int dropout_bwd_grp(const float* g, float* out, int cols, const MutL& dbias, const Grp& grp, const GIdx& hrow, const Drop& dr, cudaStream_t st) {
  PATHS_REQUIRE(cols % 4 == 0 && cols > 0 && cols <= 1024, "dropout backward: row width %d unsupported", cols);
  const int cap = grp_max_rows(grp);
  if (cap <= 0) return 0;
  const int rpb = 256 / (cols >> 2) > 0 ? 256 / (cols >> 2) : 1;
  int nb = (cap + rpb - 1) / rpb;
  if (nb > 2 * kNumSMs) nb = 2 * kNumSMs;
  pdl(dropout_bwd_kernel, dim3(nb, grp.n), 256, 0, st)(g, out, cols, dbias, grp, hrow, dr);
  PATHS_LAUNCH_CK();
  return 0;
}
int relu_bwd(float* dH, const float* H, int ld, int M_host, const int* M_dev, int M_extra, float* dbias, cudaStream_t st) {
  PATHS_REQUIRE(M_extra == 0, "relu_bwd: M_extra is no longer supported");
  return relu_bwd_grp(dH, H, ld, grp_single(M_host, M_dev), mutl1(dbias), st, nullptr, 0, 1.f);
}
int zero_rows(float* X, int ld, int M_host, const int* M_dev, int M_extra, cudaStream_t st) {
  if (M_host + M_extra <= 0) return 0;
  pdl(zero_rows_kernel, 4 * kNumSMs, 256, 0, st)(X, (size_t)ld, M_host, M_dev, M_extra);
  PATHS_LAUNCH_CK();
  return 0;
}
int scatter_add_rows_grp(const float* src, const GIdx& idx, float* dst, int d, const Grp& grp, cudaStream_t st) {
  const int cap = grp_max_rows(grp);
  if (cap <= 0) return 0;
  pdl(scatter_add_rows_kernel, dim3(row_grid(cap), grp.n), TPB, 0, st)(src, idx, dst, d, grp);
  PATHS_LAUNCH_CK();
  return 0;
}
int scatter_add_rows(const float* src, const int* idx, float* dst, int d, int B, cudaStream_t st) {
  return scatter_add_rows_grp(src, gidx1(idx), dst, d, grp_single(B, nullptr), st);
}
int head_fwd(const float* z, const float* Fprev, const float* Wc, const float* bc, float* F, float* logits, int d, int nl, int B, cudaStream_t st) {
  PATHS_REQUIRE(d <= 256 && nl <= 32, "head: d <= 256 and at most 32 logits supported");
  pdl(head_fwd_kernel, row_grid(B), TPB, 0, st)(z, Fprev, Wc, bc, F, logits, d, nl, B);
  PATHS_LAUNCH_CK();
  return 0;
}
int head_bwd(const float* dlogits, const float* dF_in, const float* Wc, float* dF, int d, int nl, int B, cudaStream_t st) {
  pdl(head_bwd_kernel, row_grid(B), TPB, 0, st)(dlogits, dF_in, Wc, dF, d, nl, B);
  PATHS_LAUNCH_CK();
  return 0;
}
int loss_fwd_bwd(const float* logits, const int* ybin, const int* cens, float* hazards, float* loss_per, float* loss, float* risk, float* dlogits,
                 int nl, int B, int task, float scale, cudaStream_t st) {
  pdl(loss_kernel, (B + 127) / 128, 128, 0, st)(logits, ybin, cens, hazards, loss_per, risk, dlogits, nl, B, task, 0.4f, 1e-7f, scale);
  PATHS_LAUNCH_CK();
  pdl(loss_reduce_kernel, 1, 256, 0, st)(loss_per, loss, B, scale);
  PATHS_LAUNCH_CK();
  return 0;
}

}  // namespace paths

// ---- level-0 features from the collated, zero-padded batch["fts"] [B, N, D] (dataset.py:217) into packed rows --------
namespace paths {
namespace {
__global__ void __launch_bounds__(256) pack_padded_kernel(const float* __restrict__ fts, int N, const int* __restrict__ seg,
                                                          const int* __restrict__ row_slide, float* __restrict__ X, int D,
                                                          int M_host, const int* M_dev) {
  pdl_grid_sync();
  const int M = M_dev ? *M_dev : M_host;
  const int lane = threadIdx.x & 31;
  for (int m = blockIdx.x * 8 + (threadIdx.x >> 5); m < M; m += gridDim.x * 8) {
    const int s = row_slide[m], i = m - seg[s];
    const float4* src = reinterpret_cast<const float4*>(fts + ((size_t)s * N + i) * D);
    float4* dst = reinterpret_cast<float4*>(X + (size_t)m * D);
    for (int c = lane; c < D / 4; c += 32) dst[c] = __ldg(src + c);
  }
}
}  // namespace
int pack_padded_rows(const float* fts, int N, const int* seg, const int* row_slide, float* X, int D, int M_host, const int* M_dev,
                     cudaStream_t st) {
  if (M_host <= 0) return 0;
  int g = (M_host + 7) / 8; if (g > 8 * kNumSMs) g = 8 * kNumSMs;
  pdl(pack_padded_kernel, g, 256, 0, st)(fts, N, seg, row_slide, X, D, M_host, M_dev);
  PATHS_LAUNCH_CK();
  return 0;
}
}  // namespace paths

// ---- occupancy bitmap of a [cells, D] grid (replaces the per-step `fts[x, y].sum() != 0` probe, slide.py:324) -----
namespace paths {
namespace {
__global__ void __launch_bounds__(256) occupancy_kernel(const float* __restrict__ grid, int cells, int D, unsigned int* __restrict__ bits,
                                                        int* __restrict__ count) {
  pdl_grid_sync();
  const int lane = threadIdx.x & 31;
  int local = 0;
  for (int c = blockIdx.x * 8 + (threadIdx.x >> 5); c < cells; c += gridDim.x * 8) {
    const float4* r = reinterpret_cast<const float4*>(grid + (size_t)c * D);
    bool nz = false;
    for (int i = lane; i < D / 4; i += 32) { const float4 v = __ldg(r + i); nz |= (v.x != 0.f) | (v.y != 0.f) | (v.z != 0.f) | (v.w != 0.f); }
    const bool any = __any_sync(0xffffffffu, nz);
    if (lane == 0 && any) { atomicOr(bits + (c >> 5), 1u << (c & 31)); ++local; }
  }
  if (lane == 0 && local) atomicAdd(count, local);
}
}  // namespace
}  // namespace paths

extern "C" int paths_build_occupancy(const float* grid, int cells, int D, uint32_t* bits, int32_t* count, void* stream) {
  using namespace paths;
  PATHS_REQUIRE(D % 4 == 0, "occupancy: D must be a multiple of 4");
  if (cells <= 0) return 0;
  int g = (cells + 7) / 8; if (g > 8 * kNumSMs) g = 8 * kNumSMs;
  pdl(occupancy_kernel, g, 256, 0, static_cast<cudaStream_t>(stream))(grid, cells, D, bits, count);
  PATHS_LAUNCH_CK();
  return 0;
}
