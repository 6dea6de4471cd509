// Host launchers of the row-wise kernels (rowops.cu), the varlen attention (attention.cu) and the
// select/expand kernels (select.cu).  All pointers are device pointers; M_dev, when non-null, overrides M_host
// as the live row count (M_host is then the capacity used to size the grid).
#pragma once
#include "common.cuh"

namespace paths {

// Per-group row index lists into a concatenated buffer: group g's entry idx.p[g][m] is relative to row roff[g] of that buffer.
struct GIdx { IdxL idx; int roff[kMaxGroups]; };
inline GIdx gidx1(const int* idx, int roff = 0) {
  GIdx g; g.idx = idxl1(idx);
  for (int i = 0; i < kMaxGroups; ++i) g.roff[i] = 0;
  g.roff[0] = roff;
  return g;
}

int gather_rows(const float* const* grid_ptr, const int* gh, int level, int L, const int* slide_ids, const int* row_slide,
                const int* row_xy, float* X, int D, int M_host, const int* M_dev, cudaStream_t st, int bf16 = 0, const Planes* XP = nullptr,
                int want_lo = 0);
int pack_padded_rows(const float* fts, int N, const int* seg, const int* row_slide, float* X, int D, int M_host, const int* M_dev,
                     cudaStream_t st);
// Outputs come as fp32 rows (pointer may be null) and/or bf16 (hi, lo) planes (Planes.p may be null); want_lo = 0 skips the
// lo plane (single-pass bf16 mode).  The backward kernels also accumulate the bias gradients (column sums) they pass over.
int lstm_point1_fwd(float* G, const float* Cprev, const int* parent, float* Cn, const Planes& CnP, int want_lo, int activated, int D, int C,
                    int M_host,
                    const int* M_dev, cudaStream_t st);
int lstm_point2_fwd(float* Pt, const float* G, const float* X, float* H, float* Y, const Planes& YP, int want_lo, int D, int C, int M_host,
                    const int* M_dev, cudaStream_t st);
// child_of != nullptr: dHc / dCc are the CHILD level's per-row gradient tables and each row sums its children's rows itself
int lstm_point2_bwd(const float* dY, const float* dHc, const int* child_of, const float* Pt, const float* G, float* dP, const Planes& dPP,
                    float* dG, const Planes& dGP, float* db_mo, float* db_o, int want_lo, int D, int C, int M_host, const int* M_dev, cudaStream_t st);
int lstm_point1_bwd(const float* dCn, const float* dCc, const int* child_of, const float* G, const float* Cprev, const int* parent, float* dG,
                    const Planes& dGP, float* dCrow, float* db_g, int want_lo, int D, int C, int M_host, const int* M_dev, cudaStream_t st);
// shared-parent gate contraction (rowops.cu: parent_slots_kernel, child_sum_planes_kernel)
int parent_slots(const int* child_of, const int* count, int cap, const int* sel, const int* sel_count, int B, int kcap, const int* n_row_slide,
                 const int* n_row_parent, const int* n_parent_local, const int* n_count, int n_cap, int* slot4, int* sel_rows, int* add_idx,
                 cudaStream_t st);
int child_sum_planes(const Planes& src, const int* child_of, const int* sel_rows, int slots, int cols, const Planes& dst, int want_lo,
                     cudaStream_t st);
int parent_sum(const int* child_of, const float* dHrow, const float* dCrow, float* dHc, float* dCc, int D, int C, int M_host,
               const int* M_dev, cudaStream_t st);
int imp_tok_fwd(float* UV, const float* b1, const float* w2, const float* b2, const float* bin, const float* div, const int* row_slide,
                const int* row_xy, const int* seg, float* alpha, float* tok, int I, int d, int imp_mul, int pe_mode, int mode, int M_host,
                const int* M_dev, cudaStream_t st);
int special_fill(const float* special, const int* seg, float* tok, int d, int B, cudaStream_t st);
int imp_tok_bwd(const float* dtok, const float* UV, const float* alpha, const float* w2, const int* row_slide, float* dUV, const Planes& dUVP,
                int want_lo, float* db1, float* dw2, float* db2, float* dbin, int I, int d, int imp_mul, int mode, int M_host, const int* M_dev,
                cudaStream_t st);
// RNN variant (lstm = false, paths.py:96-109) and slide_ctx_mode "concat" (paths.py:134-137)
struct FList { const float* f[8]; int n; };
struct GList { float* g[8]; int n; };
int rnn_z_fwd(const float* X, const float* alpha, const float* hctx, float* Z, int D, int imp_mul, int M_host, const int* M_dev,
              cudaStream_t st);
int rnn_alpha_bwd(float* dZ, const float* dZc, const float* X, const float* UV, const float* alpha, const float* w2, float* dUV, float* db1,
                  float* dw2, float* db2, int D, int I, int d, int imp_mul, int M_host, const int* M_dev, cudaStream_t st);
int head_concat_fwd(const FList& fl, const float* Wc, const float* bc, float* logits, int d, int nl, int B, cudaStream_t st);
int head_concat_bwd(const float* dlogits, const float* Wc, const GList& gl, int d, int nl, int B, cudaStream_t st);
int special_grad(const float* dtok, const int* seg, float* dspecial, int d, int B, cudaStream_t st);
// grouped forms (common.cuh: Grp): buffers are concatenated over the groups, parameters come per group
int res_ln_fwd_grp(const float* res, const GIdx& ridx, const float* add, const PtrL& bias, const PtrL& gamma, const PtrL& beta, float* y,
                   float* xhat, float* rstd, int d, float eps, const Grp& grp, cudaStream_t st);
int ln_bwd_grp(const float* dy, const float* dy2, const float* xhat, const float* rstd, const PtrL& gamma, float* dx, const GIdx& didx, int accum,
               const MutL& dgamma, const MutL& dbeta, const MutL& dbias, int d, const Grp& grp, cudaStream_t st);
int relu_bwd_grp(float* dH, const float* H, int ld, const Grp& grp, const MutL& dbias, cudaStream_t st, const Planes* outp = nullptr,
                 int want_lo = 0, float scale = 1.f);
int scatter_add_rows_grp(const float* src, const GIdx& idx, float* dst, int d, const Grp& grp, cudaStream_t st);
int special_grad_grp(const float* dtok, const GIdx& seg, const MutL& dspecial, int d, int B, int ngroups, cudaStream_t st);
int colsum_accum_grp(const float* X, int ld, const MutL& out, const Grp& grp, int N, cudaStream_t st);
// Dropout of a sublayer output (common.cuh: Drop).  Rows are hashed by their level-local token row: the row itself, or
// hrow.idx[g][m] for slide-shaped buffers (the special token's row).
//   fwd: out[m, :] = keep ? ((in ? in[m, :] : 0) + bias_g) / (1 - p) : 0          (in == out allowed; cols % 4 == 0)
//   bwd: out[m, :] = keep ? g[m, :] / (1 - p) : 0  (out may be null);  dbias_g += column sums of it   (cols <= 1024)
int dropout_fwd_grp(const float* in, float* out, int cols, const PtrL& bias, const Grp& grp, const GIdx& hrow, const Drop& dr, cudaStream_t st);
int dropout_bwd_grp(const float* g, float* out, int cols, const MutL& dbias, const Grp& grp, const GIdx& hrow, const Drop& dr, cudaStream_t st);
int res_ln_fwd(const float* res, const int* ridx, const float* add, const float* bias, const float* gamma, const float* beta, float* y,
               float* xhat, float* rstd, int d, float eps, int M_host, const int* M_dev, int M_extra, cudaStream_t st);
int ln_bwd(const float* dy, const float* dy2, const float* xhat, const float* rstd, const float* gamma, float* dx, const int* didx, int accum,
           float* dgamma, float* dbeta, float* dbias, int d, int M_host, const int* M_dev, int M_extra, cudaStream_t st);
int relu_bwd(float* dH, const float* H, int ld, int M_host, const int* M_dev, int M_extra, float* dbias, cudaStream_t st);
int zero_rows(float* X, int ld, int M_host, const int* M_dev, int M_extra, cudaStream_t st);
int scatter_add_rows(const float* src, const int* idx, float* dst, int d, int B, cudaStream_t st);
int head_fwd(const float* z, const float* Fprev, const float* Wc, const float* bc, float* F, float* logits, int d, int nl, int B, cudaStream_t st);
int head_bwd(const float* dlogits, const float* dF_in, const float* Wc, float* dF, int d, int nl, int B, cudaStream_t st);
int loss_fwd_bwd(const float* logits, const int* ybin, const int* cens, float* hazards, float* loss_per, float* loss, float* risk, float* dlogits,
                 int nl, int B, int task, float scale, cudaStream_t st);

// attention.cu -- varlen softmax self-attention over packed token rows.  qkv is [ntok, 3d] (q | k | v), heads of width dh.
// Slide s owns token rows [seg[s]+s, seg[s+1]+s+1).  nq_mode 0: every token is a query; 1: only token 0 of each slide
// (the last decoder layer is read at the special token only, aggregator.py:75); outputs are then compact [B, d].
struct SegGrp { int n, B; const int* seg[kMaxGroups]; int roff[kMaxGroups]; int oroff[kMaxGroups]; };
inline SegGrp seg_single(const int* seg, int B) {
  SegGrp g; g.n = 1; g.B = B;
  for (int i = 0; i < kMaxGroups; ++i) { g.seg[i] = nullptr; g.roff[i] = 0; g.oroff[i] = 0; }
  g.seg[0] = seg;
  return g;
}
#ifdef __CUDACC__
// Virtual slide sv of a grouped launch = (group g = sv / B, slide s = sv % B): its tokens are rows roff[g] + seg[g][s] + s ...
// of the concatenated qkv buffer, its compact (single-query) output row is oroff[g] + s.
struct Seg { int t0, T, os, g, r0; };            // (+ row group and its first row: dropout hashes level-local token rows)
__device__ __forceinline__ Seg seg_of(const SegGrp& sg, int sv) {
  const int g = sv / sg.B, s = sv - g * sg.B;
  const int* seg = pick(sg.seg, g);
  Seg r; r.r0 = pick(sg.roff, g); r.g = g; r.t0 = r.r0 + seg[s] + s; r.T = seg[s + 1] - seg[s] + 1; r.os = pick(sg.oroff, g) + s;
  return r;
}
#endif
// Tensor-core (tcgen05) attention, attention_tc.cu: head_dim 32, every token a query; bags of at most 128 tokens as one tile per
// (head, slide), longer ones tiled.  passes = 3: Q K^T, P V and the four backward products as three-term bf16 splits
// (fp32-parity); 1: plain bf16.  dvec: [tokens, heads] scratch (D = dO . O of the long bags).
bool attn_tc_supported(int d, int heads);
// cls (optional): attn_cls_ints(bags) ints of scratch -- the bags of 65..128 tokens are listed there first and served by a resident grid
inline size_t attn_cls_ints(int bags) { return 2 + 2 * (size_t)bags; }
int attn_tc_fwd(int passes, const float* qkv, const SegGrp& seg, float* out, float* lse, int d, int heads, int max_T, cudaStream_t st,
                const Drop& dr, int* cls = nullptr);
int attn_tc_bwd(int passes, const float* qkv, const SegGrp& seg, const float* out, const float* lse, const float* dout, float* dqkv, float* dvec,
                int d, int heads, int max_T, cudaStream_t st, const Drop& dr, int* cls = nullptr);
// drop (optional): dropout of the attention probabilities, hashed on (level-local query row, head << 24 | key index)
// tc_passes: 0 = SIMT fp32 kernels only; 3 / 1 = short bags on the tensor cores (attn_tc_*) in split / plain bf16
int attn_fwd_grp(const float* qkv, const SegGrp& seg, float* out, float* lse, int d, int heads, int max_T, int nq_mode, cudaStream_t st,
                 const Drop* drop = nullptr, int tc_passes = 0, int* cls = nullptr);
int attn_bwd_grp(const float* qkv, const SegGrp& seg, const float* out, const float* lse, const float* dout, float* dqkv, float* Dvec, int d,
                 int heads, int max_T, int nq_mode, cudaStream_t st, const Drop* drop = nullptr, int tc_passes = 0, int* cls = nullptr);
int attn_fwd(const float* qkv, const int* seg, float* out, float* lse, int d, int heads, int B, int max_T, int nq_mode, cudaStream_t st);
bool attn_bwd_writes_all(int d, int heads);
int attn_bwd(const float* qkv, const int* seg, const float* out, const float* lse, const float* dout, float* dqkv, float* Dvec, int d,
             int heads, int B, int max_T, int nq_mode, cudaStream_t st);

// select.cu -- per-slide top-K by importance, 4-child expansion, bounds + occupancy filter, order-preserving compaction
// (data_utils/slide.py:294-352).
struct SelectArgs {
  // level l (input)
  const float* alpha; const int* seg; const int* row_xy; const int* count; int B; int cap;
  int keep;                       // K (-1 = keep all)
  int kcap;                       // per-slide bound on kept rows (host): K, or the per-slide row bound when K = -1
  int maxseg;                     // per-slide bound on the rows of level l the buffers were sized for
  int n_maxseg;                   // ... and of level l+1
  // pyramid store, level l+1
  const int* gw; const int* gh; const unsigned long long* occ_off; const unsigned int* occ; const int* nonbg; int level; int L;
  const int* slide_ids;
  // level l outputs
  int* sel;                       // [B * kcap] kept rows (global), importance-descending; -1 padded
  int* sel_count;                 // [B]
  int* child_of;                  // [cap * 4]
  // level l+1 outputs
  int* n_count; int* n_seg; int* n_row_slide; int* n_row_xy; int* n_row_parent; int* n_parent_local; int n_cap;
  int* n_ntok;
  int* overflow;                  // [2]: flag, rows needed
  // scratch
  int* slide_cnt;                 // [B] children per slide (after filter / fallback)
  int* slide_fb;                  // [B] fallback flag
};
int select_expand(const SelectArgs& a, cudaStream_t st);
int level0_rows(const int* gw, const int* gh, int L, const int* slide_ids, int B, int cap, int* count, int* ntok, int* seg, int* row_slide,
                int* row_xy, int* row_parent, int* parent_local, int* overflow, cudaStream_t st);

}  // namespace paths
