// Per-slide top-K by importance, 4-child expansion, bounds + occupancy filter and order-preserving compaction into the
// packed row tables of the next magnification level.  Restates data_utils/slide.py:294-352 (PreprocessedSlide.iter) on
// the device: no importance.cpu() sync (utils.py:250), no per-slide Python loop (utils.py:252-258), and the
// `fts[...].sum(dim=1) != 0` background probe (slide.py:324; 4 KB read per candidate) becomes one bit of a per-level
// occupancy bitmap built when the pyramid is loaded.  Integer outputs only: bit-exact against the reference.
#include "common.cuh"
#include "rowops.cuh"

namespace paths {

namespace {

constexpr int ST = 256;                    // threads per block
constexpr int MAX_N_BITS = 65536;          // largest bag one slide may have at a level with K != -1

__device__ __forceinline__ bool occ_bit(const unsigned int* occ, unsigned long long off, int cell) {
  return (occ[off + (cell >> 5)] >> (cell & 31)) & 1u;
}

struct Cand { int parent_row, pi, blk, cx, cy; bool ok; };

// candidate j of a slide with k kept rows: block-major (slide.py:312-315): blk 0:(2x,2y) 1:(2x,2y+1) 2:(2x+1,2y) 3:(2x+1,2y+1)
__device__ __forceinline__ Cand candidate(const SelectArgs& a, int s, int sid, int k, int j) {
  Cand c;
  c.blk = j / k; c.pi = j - c.blk * k;
  c.parent_row = a.sel[(size_t)s * a.kcap + c.pi];
  const int x = a.row_xy[2 * c.parent_row], y = a.row_xy[2 * c.parent_row + 1];
  c.cx = 2 * x + (c.blk >> 1); c.cy = 2 * y + (c.blk & 1);
  const size_t g = (size_t)sid * a.L + a.level + 1;
  const int W = a.gw[g], H = a.gh[g];
  c.ok = c.cx < W && c.cy < H && occ_bit(a.occ, a.occ_off[g], c.cx * H + c.cy);
  return c;
}

__global__ void __launch_bounds__(ST) select_topk_kernel(SelectArgs a) {
  pdl_grid_sync();
  const int s = blockIdx.x, t = threadIdx.x;
  const int sid = a.slide_ids ? a.slide_ids[s] : s;
  // A bag above the level's per-slide bound (a fallback bag that fitted the batch-wide slack, slide.py:336-352) has already
  // raised the overflow flag and the host re-runs the step with worst-case buffers: clamp so this pass stays inside sel[] and
  // the `taken` bitmap whatever the bag holds.
  const int base = a.seg[s], n = min(a.seg[s + 1] - base, a.keep < 0 ? min(a.maxseg, a.kcap) : min(a.maxseg, MAX_N_BITS));
  const int k = a.keep < 0 ? n : min(n, a.keep);
  __shared__ unsigned int taken[MAX_N_BITS / 32];
  __shared__ float rv[ST / 32];
  __shared__ int ri[ST / 32];
  __shared__ int cnt_red[ST / 32];

  for (int i = t; i < (a.seg[s + 1] - base) * 4; i += ST) a.child_of[(size_t)base * 4 + i] = -1;

  if (a.keep < 0) {
    for (int i = t; i < n; i += ST) a.sel[(size_t)s * a.kcap + i] = base + i;           // keep all, original order (slide.py:294)
  } else {
    for (int i = t; i < (n + 31) / 32; i += ST) taken[i] = 0u;
    __syncthreads();
    for (int r = 0; r < k; ++r) {
      // argmax over the not-yet-taken rows; ties resolve to the lowest row index (torch.topk leaves tie order unspecified)
      float bv = -INFINITY; int bi = 0x7fffffff;
      for (int i = t; i < n; i += ST) {
        if ((taken[i >> 5] >> (i & 31)) & 1u) continue;
        const float v = a.alpha[base + i];
        if (v > bv || (v == bv && i < bi) || bi == 0x7fffffff) { bv = v; bi = i; }
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
        const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
        if (oi != 0x7fffffff && (bi == 0x7fffffff || ov > bv || (ov == bv && oi < bi))) { bv = ov; bi = oi; }
      }
      if ((t & 31) == 0) { rv[t >> 5] = bv; ri[t >> 5] = bi; }
      __syncthreads();
      if (t == 0) {
        for (int w = 1; w < ST / 32; ++w)
          if (ri[w] != 0x7fffffff && (bi == 0x7fffffff || rv[w] > bv || (rv[w] == bv && ri[w] < bi))) { bv = rv[w]; bi = ri[w]; }
        taken[bi >> 5] |= 1u << (bi & 31);
        a.sel[(size_t)s * a.kcap + r] = base + bi;
      }
      __syncthreads();
    }
  }
  for (int i = k + t; i < a.kcap; i += ST) a.sel[(size_t)s * a.kcap + i] = -1;
  __syncthreads();                                  // sel[] of this slide is visible to the whole block

  int c = 0;
  for (int j = t; j < 4 * k; j += ST) c += candidate(a, s, sid, k, j).ok ? 1 : 0;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
  if ((t & 31) == 0) cnt_red[t >> 5] = c;
  __syncthreads();
  if (t == 0) {
    int tot = 0;
    for (int w = 0; w < ST / 32; ++w) tot += cnt_red[w];
    int fb = 0;
    if (tot == 0) {                                  // slide.py:336-352: no child survives -> every non-background cell
      const size_t g = (size_t)sid * a.L + a.level + 1;
      const int nb = a.nonbg[g];
      if (nb > 0) { tot = nb; fb = 1; } else { tot = a.gw[g] * a.gh[g]; fb = 2; }
    }
    a.sel_count[s] = k;
    a.slide_cnt[s] = tot;
    a.slide_fb[s] = fb;
  }
}

__global__ void __launch_bounds__(1024) select_scan_kernel(SelectArgs a) {
  pdl_grid_sync();
  __shared__ int part[1024];
  const int t = threadIdx.x;
  const int per = (a.B + 1023) / 1024;
  int loc = 0;
  int big = 0;                                     // largest bag of the next level
  for (int i = t * per; i < min(a.B, (t + 1) * per); ++i) { loc += a.slide_cnt[i]; big = max(big, a.slide_cnt[i]); }
  part[t] = loc;
  // one slide's bag above the next level's per-slide bound (the no-child fallback can do that while the batch total still
  // fits): the attention launches and the next select are sized from that bound, so the step must be re-run
  if (big > a.n_maxseg) { a.overflow[0] = 1; atomicMax(a.overflow + 1, a.n_cap + big); }
  __syncthreads();
  for (int o = 1; o < 1024; o <<= 1) {             // Hillis-Steele inclusive scan
    const int v = t >= o ? part[t - o] : 0;
    __syncthreads();
    part[t] += v;
    __syncthreads();
  }
  int run = part[t] - loc;
  for (int i = t * per; i < min(a.B, (t + 1) * per); ++i) { a.n_seg[i] = min(run, a.n_cap); run += a.slide_cnt[i]; }
  if (t == 1023) {
    const int total = part[1023];
    a.n_seg[a.B] = min(total, a.n_cap);
    if (total > a.n_cap) { a.overflow[0] = 1; atomicMax(a.overflow + 1, total); }
    const int live = min(total, a.n_cap);
    a.n_count[0] = live;
    a.n_ntok[0] = live + a.B;
  }
}

__global__ void __launch_bounds__(ST) select_write_kernel(SelectArgs a) {
  pdl_grid_sync();
  const int s = blockIdx.x, t = threadIdx.x, lane = t & 31, w = t >> 5;
  const int sid = a.slide_ids ? a.slide_ids[s] : s;
  const int k = a.sel_count[s], fb = a.slide_fb[s];
  const int out0 = a.n_seg[s];
  const size_t g = (size_t)sid * a.L + a.level + 1;
  const int W = a.gw[g], H = a.gh[g];
  const int ncand = fb ? W * H : 4 * k;
  __shared__ int wtot[ST / 32];
  int run = 0;
  for (int j0 = 0; j0 < ncand; j0 += ST) {
    const int j = j0 + t;
    bool ok = false;
    Cand c;
    if (j < ncand) {
      if (!fb) { c = candidate(a, s, sid, k, j); ok = c.ok; }
      else {
        c.parent_row = -1; c.pi = j; c.blk = 0; c.cx = j / H; c.cy = j - c.cx * H;
        ok = (fb == 2) || occ_bit(a.occ, a.occ_off[g], j);
      }
    }
    const unsigned int bal = __ballot_sync(0xffffffffu, ok);
    if (lane == 0) wtot[w] = __popc(bal);
    __syncthreads();
    int off = run;
    for (int ww = 0; ww < w; ++ww) off += wtot[ww];
    int tot = 0;
    for (int ww = 0; ww < ST / 32; ++ww) tot += wtot[ww];
    if (ok) {
      const int dst = out0 + off + __popc(bal & ((1u << lane) - 1u));
      if (dst < a.n_cap) {
        a.n_row_slide[dst] = s;
        a.n_row_xy[2 * dst] = c.cx; a.n_row_xy[2 * dst + 1] = c.cy;
        a.n_row_parent[dst] = c.parent_row;
        a.n_parent_local[dst] = c.pi;
        if (!fb) a.child_of[(size_t)c.parent_row * 4 + c.blk] = dst;
      }
    }
    run += tot;
    __syncthreads();
  }
}

__global__ void __launch_bounds__(1024) level0_scan_kernel(const int* __restrict__ gw, const int* __restrict__ gh, int L,
                                                           const int* __restrict__ slide_ids, int B, int cap, int* count, int* ntok,
                                                           int* seg, int* overflow) {
  pdl_grid_sync();
  __shared__ int part[1024];
  const int t = threadIdx.x;
  const int per = (B + 1023) / 1024;
  auto cells = [&](int i) { const size_t g = (size_t)(slide_ids ? slide_ids[i] : i) * L; return gw[g] * gh[g]; };
  int loc = 0;
  for (int i = t * per; i < min(B, (t + 1) * per); ++i) loc += cells(i);
  part[t] = loc;
  __syncthreads();
  for (int o = 1; o < 1024; o <<= 1) {
    const int v = t >= o ? part[t - o] : 0;
    __syncthreads();
    part[t] += v;
    __syncthreads();
  }
  int run = part[t] - loc;
  for (int i = t * per; i < min(B, (t + 1) * per); ++i) { seg[i] = min(run, cap); run += cells(i); }
  if (t == 1023) {
    const int total = part[1023];
    seg[B] = min(total, cap);
    if (total > cap) { overflow[0] = 1; atomicMax(overflow + 1, total); }
    count[0] = min(total, cap);
    ntok[0] = min(total, cap) + B;
  }
}

// level 0 = every grid cell, background included, x-major (slide.py:258-269)
__global__ void __launch_bounds__(ST) level0_fill_kernel(const int* __restrict__ gh, int L, const int* __restrict__ slide_ids,
                                                         const int* __restrict__ seg, int cap, int* row_slide, int* row_xy,
                                                         int* row_parent, int* parent_local) {
  pdl_grid_sync();
  const int s = blockIdx.x;
  const int H = gh[(size_t)(slide_ids ? slide_ids[s] : s) * L];
  const int base = seg[s], n = seg[s + 1] - base;
  for (int i = threadIdx.x; i < n; i += ST) {
    const int r = base + i;
    if (r >= cap) break;
    row_slide[r] = s;
    row_xy[2 * r] = i / H; row_xy[2 * r + 1] = i % H;
    row_parent[r] = -1;
    parent_local[r] = i;
  }
}

}  // namespace

int select_expand(const SelectArgs& a, cudaStream_t st) {
  if (a.B <= 0) return 0;
  pdl(select_topk_kernel, a.B, ST, 0, st)(a);
  PATHS_LAUNCH_CK();
  pdl(select_scan_kernel, 1, 1024, 0, st)(a);
  PATHS_LAUNCH_CK();
  pdl(select_write_kernel, a.B, ST, 0, st)(a);
  PATHS_LAUNCH_CK();
  return 0;
}

int level0_rows(const int* gw, const int* gh, int L, const int* slide_ids, int B, int cap, int* count, int* ntok, int* seg, int* row_slide,
                int* row_xy, int* row_parent, int* parent_local, int* overflow, cudaStream_t st) {
  if (B <= 0) return 0;
  pdl(level0_scan_kernel, 1, 1024, 0, st)(gw, gh, L, slide_ids, B, cap, count, ntok, seg, overflow);
  PATHS_LAUNCH_CK();
  pdl(level0_fill_kernel, B, ST, 0, st)(gh, L, slide_ids, seg, cap, row_slide, row_xy, row_parent, parent_local);
  PATHS_LAUNCH_CK();
  return 0;
}

}  // namespace paths
