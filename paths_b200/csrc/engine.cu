// The level engine: owns the packed row tables and activation buffers of every magnification level inside one
// caller-provided workspace and sequences the kernels of the hierarchical forward/backward pass
// (utils.py:228-279 inference_end2end + loss.backward()).  All row counts live on the device: nothing here
// synchronises with the host, so a whole step is one uninterrupted stream of launches.
#include <stdarg.h>
#include <stdlib.h>
#include <string.h>

#include <mutex>
#include <new>
#include <vector>

#include <nvtx3/nvToolsExt.h>       // header-only NVTX 3: a no-op unless a profiler injects its library

#include "../../include/paths_b200.h"
#include "common.cuh"
#include "rowops.cuh"

namespace paths {

bool pdl_enabled() {
  static const bool on = [] { const char* v = getenv("PATHS_B200_PDL"); return !(v && v[0] == '0'); }();   // on unless PATHS_B200_PDL=0
  return on;
}

static thread_local char g_err[1024] = "";
void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int gemm_nt_simt(const RowSrc& A, const float* B, int ldb, int b_kn, float* C, int ldc, const float* bias, int M_host,
                 const int* M_dev, int N, int K, int flags, cudaStream_t st);
int gemm_tn_simt(const float* P, int ldp, const RowSrc& Q, float* C, int ldc, int M_host, const int* M_dev, int NI, int NJ,
                 cudaStream_t st);
int gemm_nt_tc(int passes, const RowSrc& A, const float* B, int ldb, int b_kn, float* C, int ldc, const float* bias, int M_host,
               const int* M_dev, int N, int K, int flags, cudaStream_t st);
int gemm_tn_tc(int passes, const float* P, int ldp, const RowSrc& Q, float* C, int ldc, int M_host, const int* M_dev, int NI, int NJ,
               cudaStream_t st);
bool gemm_tc_supported_nt(const RowSrc& A, const float* B, int ldb, int b_kn, int ldc, int N, int K);
bool gemm_tc_supported_tn(int ldp, const RowSrc& Q, int ldc, int NI, int NJ);
int gemm_nt_tc_p(int passes, const Planes& A, int a_col0, const float* B, int ldb, int b_kn, float* C, int ldc, const float* bias, int M_host,
                 const int* M_dev, int N, int K, int flags, cudaStream_t st, const RowAdd* add = nullptr);
int gemm_tn_tc_p(int passes, const Planes& P, int p_col0, const Planes& Q, int q_col0, float* C, int ldc, int M_host, const int* M_dev, int NI,
                 int NJ, cudaStream_t st);
int split_rows(const RowSrc& A, int M_host, const int* M_dev, int K, const Planes& dst, int col0, bool want_lo, cudaStream_t st);
bool gemm_tc_supported_nt_grp(const RowSrc& A, const float* B, int ldb, int N, int K);
int gemm_nt_tc_grp(int passes, const RowSrc& A, const float* B, int ldb, int b_kn, float* C, int ldc, const GemmGrp& gg, int N, int K, int flags,
                   cudaStream_t st);
int gemm_tn_tc_grp(int passes, const float* P, int ldp, const RowSrc& Q, int ldc, const GemmGrp& gg, int NI, int NJ, cudaStream_t st);
int split_rows_grp(const float* A, int lda, int K, const Planes& dst, const Grp& grp, bool want_lo, const MutL* colsum, cudaStream_t st);
int gemm_nt_tc_grp_p(int passes, const Planes& A, const float* B, int ldb, int b_kn, float* C, int ldc, const GemmGrp& gg, int N, int K, int flags,
                     cudaStream_t st);
int gemm_tn_tc_grp_p(int passes, const Planes& P, const Planes& Q, int ldc, const GemmGrp& gg, int NI, int NJ, cudaStream_t st);
size_t packed_bytes(int R, int Cc);
int register_packed(const float* src, int R, int Cc, int ld, void* dst, int nstack);
void unregister_packed(const float* src);
const void* packed_dst_of(const float* src);
int build_pack_jobs(const float* const* srcs, int n, void* jobs_dev, int* n_jobs, cudaStream_t st);
int run_pack_jobs(const void* jobs_dev, int n_jobs, cudaStream_t st);
size_t pack_job_bytes();
void set_tc_scratch(void* a, size_t a_bytes, void* b, size_t b_bytes);
size_t tc_scratch_bytes(int rows, int cols);
size_t tn_workspace_bytes(int NI, int NJ, int M_host, int ngroups);
void set_tn_workspace(void* p, size_t bytes);

int gemm_nt(int prec, const RowSrc& A, const float* B, int ldb, int b_kn, float* C, int ldc, const float* bias, int M_host,
            const int* M_dev, int N, int K, int flags, cudaStream_t st) {
  if (prec != PREC_FP32_SIMT && gemm_tc_supported_nt(A, B, ldb, b_kn, ldc, N, K))
    return gemm_nt_tc(prec == PREC_BF16X3_TC ? 3 : 1, A, B, ldb, b_kn, C, ldc, bias, M_host, M_dev, N, K, flags, st);
  return gemm_nt_simt(A, B, ldb, b_kn, C, ldc, bias, M_host, M_dev, N, K, flags, st);
}
int gemm_tn(int prec, const float* P, int ldp, const RowSrc& Q, float* C, int ldc, int M_host, const int* M_dev, int NI, int NJ,
            cudaStream_t st) {
  if (prec != PREC_FP32_SIMT && gemm_tc_supported_tn(ldp, Q, ldc, NI, NJ))
    return gemm_tn_tc(prec == PREC_BF16X3_TC ? 3 : 1, P, ldp, Q, C, ldc, M_host, M_dev, NI, NJ, st);
  return gemm_tn_simt(P, ldp, Q, C, ldc, M_host, M_dev, NI, NJ, st);
}

// Contractions on activations that may already exist as tensor-core planes (written by their producer kernel).
int gemm_nt_act(int prec, const Act& A, const float* B, int ldb, int b_kn, float* C, int ldc, const float* bias, int M_host, const int* M_dev,
                int N, int K, int flags, cudaStream_t st, const RowAdd* add = nullptr) {
  if (prec != PREC_FP32_SIMT && A.pl.p)
    return gemm_nt_tc_p(prec == PREC_BF16X3_TC ? 3 : 1, A.pl, A.col0, B, ldb, b_kn, C, ldc, bias, M_host, M_dev, N, K, flags, st, add);
  PATHS_REQUIRE(A.f && !add, "gemm_nt: activation has neither fp32 rows nor planes (row addends need the planes form)");
  return gemm_nt(prec, row_src(A.f + A.col0, A.ld, K), B, ldb, b_kn, C, ldc, bias, M_host, M_dev, N, K, flags, st);
}
int gemm_tn_act(int prec, const Act& P, const Act& Q, const RowSrc* Qsrc, float* C, int ldc, int M_host, const int* M_dev, int NI, int NJ,
                cudaStream_t st) {
  if (prec != PREC_FP32_SIMT && P.pl.p && Q.pl.p)
    return gemm_tn_tc_p(prec == PREC_BF16X3_TC ? 3 : 1, P.pl, P.col0, Q.pl, Q.col0, C, ldc, M_host, M_dev, NI, NJ, st);
  PATHS_REQUIRE(P.f && (Q.f || Qsrc), "gemm_tn: activation has neither fp32 rows nor planes");
  return gemm_tn(prec, P.f + P.col0, P.ld, Qsrc ? *Qsrc : row_src(Q.f + Q.col0, Q.ld, NJ), C, ldc, M_host, M_dev, NI, NJ, st);
}

namespace {

struct Bump {                       // two-pass workspace carver: base == nullptr only measures
  char* base; size_t off;
  template <typename T> T* take(size_t n) {
    off = (off + 255) & ~size_t(255);
    T* p = base ? reinterpret_cast<T*>(base + off) : nullptr;
    off += n * sizeof(T);
    return p;
  }
};

struct LayerBuf { float *qkv, *O, *lse, *y1, *xh1, *rs1, *y2, *xh2, *rs2, *Hf, *y3, *xh3, *rs3; };

struct Level {
  int cap, maxseg, kcap, ncap_tok;
  int *count, *ntok, *seg, *tseg, *row_slide, *row_xy, *row_parent, *parent_local, *sel, *sel_count, *child_of;
  float *X, *G, *Cn, *Pt, *H, *Y, *UV, *alpha, *tok;
  LayerBuf layer[PATHS_MAX_LAYERS];
  float *xhf, *rsf, *z, *F, *logits, *dlogits, *dHrow, *dCrow;
  float *dFc;                        // slide_ctx_mode "concat": d loss / d F_l written by the last level's head backward
  Planes AxhP, CnP, YP;              // tensor-core operand planes written by the producers: [x || h_parent], c', Y
  // shared-parent gate contraction (rowops.cu: parent_slots_kernel): W_h h_parent once per selected parent of level l - 1 ("slot")
  bool shared;                       // this level (l >= 1) runs it: tensor-core modes, fewer parent slots than child rows
  bool tables_ok;                    // the slot tables describe the level's current rows (built by do_select; dropped when a caller takes
                                     // the level's ROW_PARENT buffer to fill the tables by hand, as the level API of runtime.py does)
  bool ran_pl, ran_shared;           // what the last lstm_forward of the level used: its backward (and its parent's) must match
  bool x_planes;                     // the gather of this pass already wrote x into AxhP (lstm_forward skips its split pass over X)
  int *slot4, *sel_rows;             // as the parent level: [cap, 4] slot of a row with surviving children; [B * kcap] row of a slot
  int* add_idx;                      // as the child level: [cap] slot of the row's parent
  Planes HselP;                      // as the child level: h of the selected parents, [B * kcap(l-1), D] planes (kept for dW)
  // decoder buffers.  Every token-shaped ([tokens, *]) and slide-shaped ([B, *]) decoder buffer of the L levels is ONE allocation
  // with the levels back to back (level l at rows tok_off[l] resp. l * B): lv[0]'s pointer is the base of the concatenated
  // buffer, lv[l]'s the level's slice, so the same kernels serve one level or, in a single grouped launch, any set of levels.
  int tok_off;                                           // first token row of this level in the concatenated buffers
  float *A1, *Fo;                                        // forward temporaries [tokens, d]
  float *A1b, *Fob;                                      // ... and their [B, d] counterparts (last layer: special-token rows only)
  float *ga, *gb, *gO, *gt, *dqkv, *dHf, *Dvec;          // backward, token-shaped
  float* Dvecb;                                          // dO.O of the single-query (last) layer, slide-shaped
  int* attn_cls;                                         // bag-class list of the set this level leads (rowops.cuh: attn_cls_ints)
  float *gab, *gbb, *gOb, *dHfb;                         // backward, slide-shaped (last layer)
  float *dFl, *dy3s;                                     // [B, d]: d loss / d F_l, gradient entering the last decoder layer
};

struct LayerP { const float *Wqkv, *bqkv, *Wo, *bo, *bcross, *W1, *b1, *W2, *b2, *n1w, *n1b, *n2w, *n2b, *n3w, *n3b; };
struct LevelP { const float *Wuv, *b1, *w2, *b2, *bin, *special, *nfw, *nfb, *Wc, *bc, *Wh0, *bh0, *Wh2, *bh2; LayerP layer[PATHS_MAX_LAYERS]; };
struct Params { const float *Wg, *bg, *Wmo, *bmo; LevelP level[PATHS_MAX_LEVELS]; };

__global__ void tseg_kernel(const int* __restrict__ seg, int* __restrict__ tseg, int B) {
  pdl_grid_sync();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < B) tseg[i] = seg[i] + i;
}

}  // namespace
}  // namespace paths

using namespace paths;

struct paths_engine {
  paths_dims_t dm;
  int B;
  Level lv[PATHS_MAX_LEVELS];
  // shared scratch
  int maxcap, maxtok;
  float *dY, *dP, *dG, *dCn, *dHc, *dCc, *dUV;          // backward, row-shaped (LSTM block: main stream only)
  // the aggregator (decoder) of a level depends only on that level's tokens, and nothing downstream but the slide vector
  // depends on it: each level's decoder runs on its own stream, overlapping the LSTM / select chain of the deeper levels
  cudaStream_t side[PATHS_MAX_LEVELS];
  cudaStream_t mainp;                                    // engine-owned high-priority stream of the LSTM / select chain
  cudaEvent_t ev_in, ev_out;                             // fork from / join into the caller's stream
  cudaEvent_t ev_tok[PATHS_MAX_LEVELS], ev_z[PATHS_MAX_LEVELS], ev_head[PATHS_MAX_LEVELS], ev_dtok[PATHS_MAX_LEVELS];
  bool have_streams;
  Planes dGP, dPP, dUVP;                                 // gradient planes written by the pointwise backward kernels
  float* Gh;                                             // shared-parent gate term W_h h_parent, [slots, 3C + D] (one level at a time)
  Planes dGsumP;                                         // ... and the child-summed gate gradient of the slots
  float *hazards, *loss, *loss_per, *risk;
  int *overflow, *slide_cnt, *slide_fb;
  // host-visible copy of the overflow flag: copied to pinned memory right after the LAST select of a forward pass (the rest of
  // the pass is then still queued), so paths_engine_overflow_poll waits for a point the GPU passes well before the step ends
  int* h_over; cudaEvent_t ev_sel; bool sel_recorded;
  // ev_grad[l]: recorded by the backward pass once every gradient of procs[l] is final (after level 0: the shared LSTM's too),
  // so a data-parallel caller can start the level's all-reduce while the shallower levels are still running
  cudaEvent_t ev_grad[PATHS_MAX_LEVELS];
  const int* slide_ids;
  Params P, Gr;                                          // parameters and their gradients (same layout)
  float* grads_base; size_t grads_floats;
  const float* div;
  paths_store_t store;
  bool have_params, have_store;
  int64_t launches;
  size_t ws_bytes;
  char* ws_base;
  // bf16 (hi, lo) planes of every weight matrix the tensor-core path contracts against
  char* tc_scratch[2]; size_t tc_scratch_each;
  char* tn_ws; size_t tn_ws_bytes;                       // stream-K partial tiles of the weight-gradient contractions
  char* dec_tn_ws[PATHS_MAX_LEVELS]; size_t dec_tn_ws_bytes;   // ... of the decoder sets (one per concurrently running set)
  const float* dtok_base;                                // gradient w.r.t. the tokens, concatenated (one of lv[0].ga / lv[0].gt)
  bool stacked;                                          // decoder weights are packed as per-kind stacks over the levels
  bool dec_planes;                                       // aggregator contractions on pre-split planes (PATHS_B200_DEC_NO_PLANES unsets)
  int S, Bs;                                             // row stride between the levels' regions: token-shaped / slide-shaped buffers
  Planes dplD[2], dplF[2], dpl3;                         // decoder operand planes (bf16 hi, lo) of width d, F, 3d over all L * S rows
  Planes dplDb[2], dplFb[2];                             // ... and for the slide-shaped operands of the last layer (L * Bs rows): the two
                                                         // row spaces overlap, and level sets run concurrently on different streams
  // operand planes of the forward pass's saved activations (layer input, attention output, FFN input, FFN hidden), kept for the
  // backward pass: its weight-gradient contractions read them as they are instead of splitting the fp32 tensors again
  Planes svT[PATHS_MAX_LAYERS], svO[PATHS_MAX_LAYERS], svY2[PATHS_MAX_LAYERS], svHf[PATHS_MAX_LAYERS];
  int sv_prec[PATHS_MAX_LEVELS];                         // precision the level's saved planes were written in (-1: none)
  unsigned sv_set[PATHS_MAX_LEVELS];                     // levels of the forward set that wrote them (its zero-fill covers their longest)
  // dropout of the aggregators (paths_engine_set_dropout): probability and seed of the NEXT forward pass; each level's forward
  // pass records what it used so that its backward pass repeats the same keep/drop decisions
  float drop_p; uint64_t drop_seed;
  float lv_drop_p[PATHS_MAX_LEVELS]; uint64_t lv_drop_seed[PATHS_MAX_LEVELS];
  // level sets of the whole-step entry points: each set's aggregators run as one grouped pass on the set's own stream
  int n_fsets, n_bsets;
  int fset_n[PATHS_MAX_LEVELS], fset_lv[PATHS_MAX_LEVELS][PATHS_MAX_LEVELS];
  int bset_n[PATHS_MAX_LEVELS], bset_lv[PATHS_MAX_LEVELS][PATHS_MAX_LEVELS];
  char* packed_arena; size_t packed_arena_bytes;
  void* pack_jobs; int n_pack_jobs;
  std::vector<const float*> packed_srcs;
  // optional per-kernel timing (CUDA events on the launch stream) for bench.py's roofline line
  bool prof_on;
  std::vector<cudaEvent_t> prof_ev[PATHS_PROF__COUNT];     // start/stop pairs
  struct Site { int line; const char* what; cudaEvent_t a, b; };
  bool site_on; std::vector<Site> sites; int site_skip;
};

namespace {

size_t carve(paths_engine* e, char* base) {
  const paths_dims_t& dm = e->dm;
  const int B = e->B, D = dm.D, C = dm.C, d = dm.d, I = dm.I, F = dm.F;
  Bump b{base, 0};
  int maxcap = 0, max_slots = 0;
  // uniform region stride (a multiple of 64 rows): a level's region can be zero-filled up to the longest level's row count
  // without touching its neighbour (what the grouped weight-gradient contraction on TMA-fed planes needs)
  int S = 64;
  for (int l = 0; l < dm.L; ++l) {
    Level& v = e->lv[l];
    v.ncap_tok = v.cap + B;
    if (((v.ncap_tok + 63) & ~63) > S) S = (v.ncap_tok + 63) & ~63;
    if (v.cap > maxcap) maxcap = v.cap;
  }
  const int Bs = (B + 63) & ~63;
  e->S = S; e->Bs = Bs;
  for (int l = 0; l < dm.L; ++l) e->lv[l].tok_off = l * S;
  const size_t Ttot = (size_t)dm.L * S;
  // concatenated decoder buffers: one allocation of [Ttot, cols] (token-shaped) or [L * B, cols] (slide-shaped), sliced per level
  auto cat_tok = [&](size_t cols, float* Level::*m) {
    float* base_p = b.take<float>(Ttot * cols);
    for (int l = 0; l < dm.L; ++l) e->lv[l].*m = base_p ? base_p + (size_t)e->lv[l].tok_off * cols : nullptr;
  };
  auto cat_b = [&](size_t cols, float* Level::*m) {
    float* base_p = b.take<float>((size_t)dm.L * Bs * cols);
    for (int l = 0; l < dm.L; ++l) e->lv[l].*m = base_p ? base_p + (size_t)l * Bs * cols : nullptr;
  };
  auto cat_layer = [&](int j, bool tokens, size_t cols, float* LayerBuf::*m) {
    const size_t rows = tokens ? Ttot : (size_t)dm.L * Bs;
    float* base_p = b.take<float>(rows * cols);
    for (int l = 0; l < dm.L; ++l)
      e->lv[l].layer[j].*m = base_p ? base_p + (tokens ? (size_t)e->lv[l].tok_off : (size_t)l * Bs) * cols : nullptr;
  };
  cat_tok(d, &Level::tok);
  for (int j = 0; j < dm.layers; ++j) {
    const bool tk = j != dm.layers - 1;                 // the last layer is evaluated at the special tokens only: [B, *] rows
    cat_layer(j, true, 3 * d, &LayerBuf::qkv);
    cat_layer(j, tk, d, &LayerBuf::O); cat_layer(j, tk, dm.heads, &LayerBuf::lse);
    cat_layer(j, tk, d, &LayerBuf::y1); cat_layer(j, tk, d, &LayerBuf::xh1); cat_layer(j, tk, 1, &LayerBuf::rs1);
    cat_layer(j, tk, d, &LayerBuf::y2); cat_layer(j, tk, d, &LayerBuf::xh2); cat_layer(j, tk, 1, &LayerBuf::rs2);
    cat_layer(j, tk, F, &LayerBuf::Hf);
    cat_layer(j, tk, d, &LayerBuf::y3); cat_layer(j, tk, d, &LayerBuf::xh3); cat_layer(j, tk, 1, &LayerBuf::rs3);
  }
  cat_b(d, &Level::xhf); cat_b(1, &Level::rsf); cat_b(d, &Level::z);
  cat_tok(d, &Level::A1); cat_tok(d, &Level::Fo); cat_b(d, &Level::A1b); cat_b(d, &Level::Fob);
  cat_tok(d, &Level::ga); cat_tok(d, &Level::gb); cat_tok(d, &Level::gO); cat_tok(d, &Level::gt);
  cat_tok(3 * d, &Level::dqkv); cat_tok(F, &Level::dHf); cat_tok(dm.heads, &Level::Dvec);
  cat_b(dm.heads, &Level::Dvecb);
  cat_b(d, &Level::gab); cat_b(d, &Level::gbb); cat_b(d, &Level::gOb); cat_b(F, &Level::dHfb);
  cat_b(d, &Level::dFl); cat_b(d, &Level::dy3s);
  {
    auto mk = [&](int cols) {
      Planes pl = planes_at(b.take<char>(planes_bytes((int)Ttot, cols)), (int)Ttot, cols);
      if (!base) pl.p = nullptr;
      return pl;
    };
    const bool want = d % 8 == 0 && F % 8 == 0;
    e->dplD[0] = want ? mk(d) : no_planes(); e->dplD[1] = want ? mk(d) : no_planes();
    e->dplF[0] = want ? mk(F) : no_planes(); e->dplF[1] = want ? mk(F) : no_planes();
    e->dpl3 = want ? mk(3 * d) : no_planes();
    auto mkb = [&](int cols) {
      Planes pl = planes_at(b.take<char>(planes_bytes(dm.L * Bs, cols)), dm.L * Bs, cols);
      if (!base) pl.p = nullptr;
      return pl;
    };
    e->dplDb[0] = want ? mkb(d) : no_planes(); e->dplDb[1] = want ? mkb(d) : no_planes();
    e->dplFb[0] = want ? mkb(F) : no_planes(); e->dplFb[1] = want ? mkb(F) : no_planes();
    for (int j = 0; j < PATHS_MAX_LAYERS; ++j) {
      const bool on = want && j < dm.layers, tk = j != dm.layers - 1;
      e->svT[j] = on ? mk(d) : no_planes();
      e->svO[j] = on ? (tk ? mk(d) : mkb(d)) : no_planes();
      e->svY2[j] = on ? (tk ? mk(d) : mkb(d)) : no_planes();
      e->svHf[j] = on ? (tk ? mk(F) : mkb(F)) : no_planes();
    }
  }
  {
    const int dec_shapes[][2] = {{3 * d, d}, {d, d}, {F, d}, {d, F}};
    size_t tnb = 0;
    int tokmax = 1;
    for (int l = 0; l < dm.L; ++l) if (e->lv[l].ncap_tok > tokmax) tokmax = e->lv[l].ncap_tok;
    for (const auto& sh : dec_shapes)
      for (int ng = 1; ng <= dm.L; ++ng) {
        const size_t need = tn_workspace_bytes(sh[0], sh[1], tokmax, ng);
        if (need > tnb) tnb = need;
      }
    e->dec_tn_ws_bytes = tnb;
    for (int l = 0; l < dm.L; ++l) e->dec_tn_ws[l] = b.take<char>(tnb);
  }
  for (int l = 0; l < dm.L; ++l) {
    Level& v = e->lv[l];
    const size_t cap = v.cap;
    v.count = b.take<int>(1); v.ntok = b.take<int>(1);
    v.seg = b.take<int>(B + 1); v.tseg = b.take<int>(B);
    v.row_slide = b.take<int>(cap); v.row_xy = b.take<int>(2 * cap); v.row_parent = b.take<int>(cap);
    v.parent_local = b.take<int>(cap);
    v.sel = b.take<int>((size_t)B * (v.kcap > 0 ? v.kcap : 1)); v.sel_count = b.take<int>(B);
    v.child_of = b.take<int>(4 * cap);
    v.attn_cls = b.take<int>(attn_cls_ints(dm.L * B));
    v.slot4 = b.take<int>(4 * cap); v.sel_rows = b.take<int>((size_t)B * (v.kcap > 0 ? v.kcap : 1)); v.add_idx = b.take<int>(cap);
    v.X = b.take<float>(cap * D); v.G = b.take<float>(dm.lstm ? cap * (3 * C + D) : 4); v.Cn = b.take<float>(cap * C);
    v.Pt = b.take<float>(cap * D); v.H = b.take<float>(cap * D); v.Y = b.take<float>(cap * D);
    v.UV = b.take<float>(cap * (I + d)); v.alpha = b.take<float>(cap);
    v.F = b.take<float>((size_t)B * d); v.logits = b.take<float>((size_t)B * dm.nl); v.dlogits = b.take<float>((size_t)B * dm.nl);
    v.dHrow = b.take<float>(cap * D); v.dCrow = b.take<float>(cap * C);
    v.dFc = b.take<float>((size_t)B * d);
    if (dm.lstm) {
      // shared-parent levels keep x only (the h half lives once per slot in HselP)
      const int xcols = (v.shared || l == 0) ? D : 2 * D;
      v.AxhP = planes_at(b.take<char>(planes_bytes(v.cap, xcols)), v.cap, xcols);
      v.CnP = planes_at(b.take<char>(planes_bytes(v.cap, C)), v.cap, C);
      v.YP = planes_at(b.take<char>(planes_bytes(v.cap, D)), v.cap, D);
      const int slots = v.shared ? B * e->lv[l - 1].kcap : 0;
      v.HselP = slots ? planes_at(b.take<char>(planes_bytes(slots, D)), slots, D) : no_planes();
      if (slots > max_slots) max_slots = slots;
      if (!base) { v.AxhP.p = nullptr; v.CnP.p = nullptr; v.YP.p = nullptr; v.HselP.p = nullptr; }
    } else {
      v.AxhP = v.CnP = v.YP = v.HselP = no_planes();
    }
  }
  e->maxcap = maxcap; e->maxtok = maxcap + B;
  const size_t mc = maxcap, mt = (size_t)maxcap + B;
  e->dY = b.take<float>(mc * D); e->dP = b.take<float>(mc * D); e->dG = b.take<float>(dm.lstm ? mc * (3 * C + D) : 4);
  e->dCn = b.take<float>(mc * C); e->dHc = b.take<float>(mc * D); e->dCc = b.take<float>(mc * C);
  e->dUV = b.take<float>(mc * (I + d));
  if (dm.lstm) {
    e->dGP = planes_at(b.take<char>(planes_bytes(maxcap, 3 * C + D)), maxcap, 3 * C + D);
    e->Gh = max_slots ? b.take<float>((size_t)max_slots * (3 * C + D)) : nullptr;
    e->dGsumP = max_slots ? planes_at(b.take<char>(planes_bytes(max_slots, 3 * C + D)), max_slots, 3 * C + D) : no_planes();
    if (!base) e->dGsumP.p = nullptr;
    e->dPP = planes_at(b.take<char>(planes_bytes(maxcap, D)), maxcap, D);
    e->dUVP = planes_at(b.take<char>(planes_bytes(maxcap, I + d)), maxcap, I + d);
  } else {
    e->dGP = e->dPP = e->dUVP = e->dGsumP = no_planes();
    e->Gh = nullptr;
  }
  e->hazards = b.take<float>((size_t)B * dm.nl); e->loss = b.take<float>(1); e->loss_per = b.take<float>(B);
  e->risk = b.take<float>(B);
  e->overflow = b.take<int>(2); e->slide_cnt = b.take<int>(B); e->slide_fb = b.take<int>(B);
  {
    size_t pb = dm.lstm ? packed_bytes(3 * C + D, 2 * D) + packed_bytes(D, C) : 0;
    const size_t per_level = packed_bytes(I + d, D) + (dm.lstm ? 0 : packed_bytes(C, D) + packed_bytes(D, C)) +
                             (size_t)dm.layers * (packed_bytes(3 * d, d) + packed_bytes(d, d) + packed_bytes(F, d) + packed_bytes(d, F));
    pb += per_level * dm.L;
    e->packed_arena_bytes = pb;
    e->packed_arena = b.take<char>(pb);
    e->pack_jobs = b.take<char>((size_t)(2 + dm.L * (3 + 4 * dm.layers)) * 64);
    int cols = 2 * D;
    if (3 * C + D > cols) cols = 3 * C + D;
    if (F > cols) cols = F;
    if (3 * d > cols) cols = 3 * d;
    if (I + d > cols) cols = I + d;
    e->tc_scratch_each = tc_scratch_bytes(e->maxtok, cols);
    e->tc_scratch[0] = b.take<char>(e->tc_scratch_each);
    e->tc_scratch[1] = b.take<char>(e->tc_scratch_each);
    const int tn_shapes[][2] = {{3 * C + D, 2 * D}, {3 * C + D, D}, {D, C}, {C, D}, {I + d, D}, {I, D}, {d, D}, {3 * d, d}, {d, d}, {F, d}, {d, F}};
    size_t tnb = 0;
    for (const auto& sh : tn_shapes) {
      const size_t need = tn_workspace_bytes(sh[0], sh[1], e->maxtok, 1);
      if (need > tnb) tnb = need;
    }
    e->tn_ws_bytes = tnb;
    e->tn_ws = b.take<char>(tnb);
  }
  return (b.off + 255) & ~size_t(255);
}

int check_dims(const paths_dims_t* dm, int B, const int32_t* caps, const int32_t* maxseg) {
  PATHS_REQUIRE(dm && caps && maxseg, "null argument");
  PATHS_REQUIRE(dm->L >= 1 && dm->L <= PATHS_MAX_LEVELS, "num_levels %d outside [1, %d]", dm->L, PATHS_MAX_LEVELS);
  PATHS_REQUIRE(dm->layers >= 1 && dm->layers <= PATHS_MAX_LAYERS, "trans_layers %d outside [1, %d]", dm->layers, PATHS_MAX_LAYERS);
  PATHS_REQUIRE(dm->D % 4 == 0 && dm->C % 4 == 0 && dm->d % 4 == 0 && dm->I % 4 == 0 && dm->F % 4 == 0,
                "feature dims must be multiples of 4 (D=%d C=%d d=%d I=%d F=%d)", dm->D, dm->C, dm->d, dm->I, dm->F);
  PATHS_REQUIRE(dm->d % dm->heads == 0 && dm->d / dm->heads <= 64, "head_dim must be <= 64");
  PATHS_REQUIRE(dm->d <= 256 && dm->I <= 256 && dm->nl <= 32, "trans_dim / importance hidden dim above 256 unsupported");
  PATHS_REQUIRE(dm->pe_mode == 1 || dm->pe_mode == 2, "pos_encoding_mode must be 1d or 2d");
  PATHS_REQUIRE(dm->residual >= 0 && dm->residual <= 2, "slide_ctx_mode must be none (0), residual (1) or concat (2)");
  PATHS_REQUIRE(dm->lstm == 1 || dm->lstm == 0, "lstm must be 0 or 1");
  PATHS_REQUIRE(dm->lstm == 0 || dm->hier_ctx == 1, "the LSTM requires hierarchical_ctx (config.py:106)");
  PATHS_REQUIRE(B >= 1, "empty batch");
  for (int l = 0; l < dm->L; ++l) {
    PATHS_REQUIRE(caps[l] >= 1 && maxseg[l] >= 1, "level %d: capacity must be positive", l);
    PATHS_REQUIRE(l == dm->L - 1 || dm->K[l] < 0 || maxseg[l] <= 65536, "level %d: bags above 65536 rows need top_k = -1", l);
  }
  return 0;
}

void fill_level_meta(paths_engine* e, const int32_t* caps, const int32_t* maxseg) {
  for (int l = 0; l < e->dm.L; ++l) {
    Level& v = e->lv[l];
    v.cap = caps[l]; v.maxseg = maxseg[l];
    const int K = e->dm.K[l];
    v.kcap = (l == e->dm.L - 1) ? 1 : (K < 0 ? maxseg[l] : (K < maxseg[l] ? K : maxseg[l]));
  }
  // shared-parent gate contraction (decided on the dimensions alone: the precision mode can change during the engine's life):
  // it pays when a level has fewer parent slots than rows, which is what top-K selection with four children per patch gives
  static const bool off = getenv("PATHS_B200_NO_SHARED_PARENT") != nullptr;
  const paths_dims_t& dm = e->dm;
  const bool dims_ok = dm.lstm && dm.D % 8 == 0 && dm.C % 8 == 0 && (dm.I + dm.d) % 8 == 0;
  for (int l = 0; l < dm.L; ++l)
    e->lv[l].shared = !off && dims_ok && l >= 1 && (long)e->B * e->lv[l - 1].kcap <= (long)e->lv[l].cap;
}

void bind(Params& P, const float* base, const int64_t* off, const paths_dims_t& dm) {
  auto at = [&](int64_t o) { return o < 0 ? nullptr : base + o; };
  P.Wg = at(off[PATHS_P_WG]); P.bg = at(off[PATHS_P_BG]); P.Wmo = at(off[PATHS_P_WMO]); P.bmo = at(off[PATHS_P_BMO]);
  for (int l = 0; l < dm.L; ++l) {
    const int64_t* o = off + PATHS_P_LSTM_COUNT + (size_t)l * PATHS_LEVEL_STRIDE;
    LevelP& q = P.level[l];
    q.Wuv = at(o[PATHS_P_WUV]); q.b1 = at(o[PATHS_P_B1]); q.w2 = at(o[PATHS_P_W2]); q.b2 = at(o[PATHS_P_B2]);
    q.bin = at(o[PATHS_P_BIN]); q.special = at(o[PATHS_P_SPECIAL]); q.nfw = at(o[PATHS_P_NFW]); q.nfb = at(o[PATHS_P_NFB]);
    q.Wc = at(o[PATHS_P_WC]); q.bc = at(o[PATHS_P_BC]);
    q.Wh0 = at(o[PATHS_P_WH0]); q.bh0 = at(o[PATHS_P_BH0]); q.Wh2 = at(o[PATHS_P_WH2]); q.bh2 = at(o[PATHS_P_BH2]);
    for (int j = 0; j < dm.layers; ++j) {
      const int64_t* p = o + PATHS_P_LEVEL_HEAD_COUNT + (size_t)j * PATHS_P_LAYER_COUNT;
      LayerP& y = q.layer[j];
      y.Wqkv = at(p[PATHS_P_WQKV]); y.bqkv = at(p[PATHS_P_BQKV]); y.Wo = at(p[PATHS_P_WO]); y.bo = at(p[PATHS_P_BO]);
      y.bcross = at(p[PATHS_P_BCROSS]); y.W1 = at(p[PATHS_P_FW1]); y.b1 = at(p[PATHS_P_FB1]); y.W2 = at(p[PATHS_P_FW2]);
      y.b2 = at(p[PATHS_P_FB2]); y.n1w = at(p[PATHS_P_N1W]); y.n1b = at(p[PATHS_P_N1B]); y.n2w = at(p[PATHS_P_N2W]);
      y.n2b = at(p[PATHS_P_N2B]); y.n3w = at(p[PATHS_P_N3W]); y.n3b = at(p[PATHS_P_N3B]);
    }
  }
}

// Producer-written operand planes need every contraction extent of the LSTM block to satisfy the TMA/UMMA alignment rules.
inline bool planes_mode(const paths_dims_t& dm) {
  return dm.precision != PREC_FP32_SIMT && dm.lstm && dm.D % 8 == 0 && dm.C % 8 == 0 && (dm.I + dm.d) % 8 == 0;
}

// One tagged region of a step: an NVTX range on the host thread that enqueues it (the reference has no tracing at
// all; these are the ranges a timeline of the step is read by) and, while paths_engine_profile is on, a CUDA-event pair on the
// region's stream.
static const char* const kProfNames[PATHS_PROF__COUNT] = {"paths/gates_fwd", "paths/gates_dW", "paths/gates_dH", "paths/lstm_fwd",
                                                          "paths/decoder_fwd", "paths/select_gather", "paths/decoder_bwd", "paths/lstm_bwd"};
// tensor-core passes of the short-bag attention: 0 in the fp32 SIMT mode, else the contraction modes' split (3) / plain bf16 (1)
inline int attn_tc_passes(const paths_dims_t& dm) { return dm.precision == PREC_BF16X3_TC ? 3 : dm.precision == PREC_BF16_TC ? 1 : 0; }

struct ProfScope {
  paths_engine* e; int tag; cudaStream_t st; bool on; bool open;
  ProfScope(paths_engine* e_, int tag_, cudaStream_t st_) : e(e_), tag(tag_), st(st_), on(e_->prof_on), open(true) {
    nvtxRangePushA(kProfNames[tag]);
    if (on) { cudaEvent_t a; cudaEventCreate(&a); cudaEventRecord(a, st); e->prof_ev[tag].push_back(a); }
  }
  void end() {
    if (on) { cudaEvent_t b; cudaEventCreate(&b); cudaEventRecord(b, st); e->prof_ev[tag].push_back(b); on = false; }
    if (open) { nvtxRangePop(); open = false; }
  }
  ~ProfScope() { end(); }
};

// Developer aid: with PATHS_B200_PROF_DUMP=<file> every RUN() call site is bracketed by CUDA events on its stream and the
// per-call-site totals (warm, in-situ) are written to <file> when the engine is destroyed.  Use with PATHS_B200_NO_STREAMS=1.
struct SiteScope {
  paths_engine* e; int line; const char* what; cudaStream_t st; cudaEvent_t a;
  SiteScope(paths_engine* e_, int line_, const char* what_, cudaStream_t st_) : e(e_), line(line_), what(what_), st(st_), a(nullptr) {
    if (e->site_on) { cudaEventCreate(&a); cudaEventRecord(a, st); }
  }
  ~SiteScope() {
    if (a) { cudaEvent_t b; cudaEventCreate(&b); cudaEventRecord(b, st); e->sites.push_back(paths_engine::Site{line, what, a, b}); }
  }
};

#define RUN(call)                          \
  do {                                     \
    SiteScope _ss(e, __LINE__, #call, st); \
    int _r = (call);                       \
    if (_r) return _r;                     \
    e->launches++;                         \
  } while (0)

inline float* W_(const float* p) { return const_cast<float*>(p); }   // gradient views share the Params struct

// ---- forward of one level, part 1: patch contextualiser (LSTM / RNN step), importance, tokens
int lstm_forward(paths_engine* e, int l, cudaStream_t st) {
  const paths_dims_t& dm = e->dm;
  set_tc_scratch(e->tc_scratch[0], e->tc_scratch_each, e->tc_scratch[1], e->tc_scratch_each);
  ProfScope ps_blk(e, PATHS_PROF_LSTM_FWD, st);
  const int B = e->B, D = dm.D, C = dm.C, d = dm.d, I = dm.I, F = dm.F, prec = dm.precision;
  Level& v = e->lv[l];
  const LevelP& p = e->P.level[l];
  const int M = v.cap;
  const int* Md = v.count;
  const int ldg = 3 * C + D;

  const bool has_parent = l > 0;
  if (dm.lstm) {
    // ---- LSTM cell step for every patch (interface.py:49-56).  In the tensor-core modes every contraction operand of this
    // block lives as bf16 planes written by its producer: [x || h_parent] by one split pass that also performs the parent
    // gather (kept for the weight gradient), c' and Y by the gate kernels.
    // (a shared-parent level whose tables were filled by hand has no slot tables and only x-wide planes: fp32-row operands)
    const bool pl = planes_mode(dm) && !(has_parent && v.shared && !v.tables_ok);
    const int lo = prec == PREC_BF16X3_TC ? 1 : 0;
    const int Kx = has_parent ? 2 * D : D;
    RowSrc Axh = has_parent ? row_src2(v.X, D, nullptr, D, e->lv[l - 1].H, D, v.row_parent) : row_src(v.X, D, D);
    const bool shp = pl && has_parent && v.shared;       // W_h h_parent once per selected parent, added by the gate epilogue
    v.ran_pl = pl; v.ran_shared = shp;
    const bool x_ready = v.x_planes && pl;               // consumed by this pass: a later pass without a gather splits again
    v.x_planes = false;
    if (shp) {
      Level& pv = e->lv[l - 1];
      const int slots = B * pv.kcap;
      RUN(split_rows(row_src(pv.H, D, D, pv.sel_rows), slots, nullptr, D, v.HselP, 0, lo != 0, st));
      if (!x_ready) RUN(split_rows(row_src(v.X, D, D), M, Md, D, v.AxhP, 0, lo != 0, st));
      ProfScope ps(e, PATHS_PROF_GATES_FWD, st);
      RUN(gemm_nt_act(prec, act(nullptr, 0, v.HselP), e->P.Wg + D, 2 * D, 0, e->Gh, ldg, nullptr, slots, nullptr, ldg, D, 0, st));
      RowAdd ra; ra.p = e->Gh; ra.idx = v.add_idx; ra.ld = ldg;
      RUN(gemm_nt_act(prec, act(nullptr, 0, v.AxhP), e->P.Wg, 2 * D, 0, v.G, ldg, e->P.bg, M, Md, ldg, D, GEMM_GATES | (C << 8), st, &ra));
    } else {
      if (pl && !(x_ready && !has_parent)) RUN(split_rows(Axh, M, Md, Kx, v.AxhP, 0, lo != 0, st));
      ProfScope ps(e, PATHS_PROF_GATES_FWD, st);
      // tensor-core path: the epilogue applies the gate non-linearities, G reaches HBM activated (written once)
      if (pl) RUN(gemm_nt_act(prec, act(nullptr, 0, v.AxhP), e->P.Wg, 2 * D, 0, v.G, ldg, e->P.bg, M, Md, ldg, Kx, GEMM_GATES | (C << 8), st));
      else RUN(gemm_nt(prec, Axh, e->P.Wg, 2 * D, 0, v.G, ldg, e->P.bg, M, Md, ldg, Kx, 0, st));
    }
    RUN(lstm_point1_fwd(v.G, has_parent ? e->lv[l - 1].Cn : nullptr, has_parent ? v.row_parent : nullptr, v.Cn, pl ? v.CnP : no_planes(), lo,
                        pl ? 1 : 0, D, C, M, Md, st));
    RUN(gemm_nt_act(prec, act(v.Cn, C, pl ? v.CnP : no_planes()), e->P.Wmo, C, 0, v.Pt, D, e->P.bmo, M, Md, D, C, 0, st));
    RUN(lstm_point2_fwd(v.Pt, v.G, v.X, v.H, pl ? nullptr : v.Y, pl ? v.YP : no_planes(), lo, D, C, M, Md, st));

    // ---- importance + token projection share one contraction over Y (paths.py:95-98, aggregator.py:48)
    RUN(gemm_nt_act(prec, act(v.Y, D, pl ? v.YP : no_planes()), p.Wuv, D, 0, v.UV, I + d, nullptr, M, Md, I + d, D, 0, st));
    RUN(imp_tok_fwd(v.UV, p.b1, p.w2, p.b2, p.bin, e->div, v.row_slide, v.row_xy, v.seg, v.alpha, v.tok, I, d, dm.imp_mul,
                    dm.pe_mode, 0, M, Md, st));
  } else {
    // ---- RNN variant (paths.py:95-109): alpha from X; Z = alpha X + MLP_h(Z_parent); tokens from Z.  Z lives in v.H.
    RUN(gemm_nt(prec, row_src(v.X, D, D), p.Wuv, D, 0, v.UV, I + d, nullptr, M, Md, I, D, 0, st));
    RUN(imp_tok_fwd(v.UV, p.b1, p.w2, p.b2, p.bin, e->div, v.row_slide, v.row_xy, v.seg, v.alpha, v.tok, I, d, dm.imp_mul,
                    dm.pe_mode, 1, M, Md, st));
    const bool hctx = has_parent && dm.hier_ctx;
    if (hctx) {
      RUN(gemm_nt(prec, row_src(e->lv[l - 1].H, D, D, v.row_parent), p.Wh0, D, 0, v.Cn, C, p.bh0, M, Md, C, D, GEMM_RELU, st));
      RUN(gemm_nt(prec, row_src(v.Cn, C, C), p.Wh2, C, 0, v.Pt, D, p.bh2, M, Md, D, C, 0, st));
    }
    RUN(rnn_z_fwd(v.X, v.alpha, hctx ? v.Pt : nullptr, v.H, D, dm.imp_mul, M, Md, st));
    RUN(gemm_nt(prec, row_src(v.H, D, D), p.Wuv + (size_t)I * D, D, 0, v.UV + I, I + d, nullptr, M, Md, d, D, 0, st));
    RUN(imp_tok_fwd(v.UV, p.b1, p.w2, p.b2, p.bin, e->div, v.row_slide, v.row_xy, v.seg, v.alpha, v.tok, I, d, dm.imp_mul,
                    dm.pe_mode, 2, M, Md, st));
  }
  RUN(special_fill(p.special, v.seg, v.tok, d, B, st));
  return 0;
}

// ---- level sets ---------------------------------------------------------------------------------------------------------
// The aggregators of different levels are independent problems of identical shapes: a DecSet names the levels one pass of
// decoder_forward / decoder_backward serves, every kernel of the pass being ONE grouped launch over the levels' row regions of
// the concatenated buffers (common.cuh: Grp).  A set of one level reproduces the per-level behaviour (level API, fallbacks).
struct DecSet {
  int n; int lv[PATHS_MAX_LEVELS];
  Grp gt, gb;                          // token-shaped / slide-shaped row groups
  SegGrp sg;                           // slide segments for the attention kernels
  GIdx tsegI, segI;                    // special-token rows (tseg) and segment offsets, relative to the level's token region
  int max_T;
};

DecSet make_set(const paths_engine* e, const int* lvls, int n) {
  DecSet S;
  S.n = n; S.max_T = 1;
  S.gt.n = S.gb.n = n; S.sg.n = n; S.sg.B = e->B;
  for (int g = 0; g < kMaxGroups; ++g) {
    S.gt.roff[g] = S.gb.roff[g] = 0; S.gt.hcnt[g] = S.gb.hcnt[g] = 0; S.gt.cnt[g] = S.gb.cnt[g] = nullptr;
    S.sg.seg[g] = nullptr; S.sg.roff[g] = S.sg.oroff[g] = 0;
    S.tsegI.idx.p[g] = S.segI.idx.p[g] = nullptr; S.tsegI.roff[g] = S.segI.roff[g] = 0;
    S.lv[g] = 0;
  }
  for (int g = 0; g < n; ++g) {
    const int l = lvls[g];
    const Level& v = e->lv[l];
    S.lv[g] = l;
    S.gt.roff[g] = v.tok_off; S.gt.hcnt[g] = v.ncap_tok; S.gt.cnt[g] = v.ntok;
    S.gb.roff[g] = l * e->Bs; S.gb.hcnt[g] = e->B; S.gb.cnt[g] = nullptr;
    S.sg.seg[g] = v.seg; S.sg.roff[g] = v.tok_off; S.sg.oroff[g] = l * e->Bs;
    S.tsegI.idx.p[g] = v.tseg; S.tsegI.roff[g] = v.tok_off;
    S.segI.idx.p[g] = v.seg; S.segI.roff[g] = v.tok_off;
    if (v.maxseg + 1 > S.max_T) S.max_T = v.maxseg + 1;
  }
  return S;
}

template <typename Fn> PtrL PL(const DecSet& S, Fn f) { PtrL r; for (int g = 0; g < kMaxGroups; ++g) r.p[g] = g < S.n ? f(S.lv[g]) : nullptr; return r; }
template <typename Fn> MutL ML(const DecSet& S, Fn f) { MutL r; for (int g = 0; g < kMaxGroups; ++g) r.p[g] = g < S.n ? const_cast<float*>(f(S.lv[g])) : nullptr; return r; }
inline PtrL PL0() { return ptrl1(nullptr); }
inline MutL ML0() { return mutl1(nullptr); }
inline GIdx noidx() { return gidx1(nullptr); }

// C rows of every group (+)= act(A rows . W_level^T + bias_level): grouped tensor-core launch when the weights are stacked,
// per-level contractions otherwise (fp32 SIMT mode, odd dimensions).  A and C are bases of concatenated buffers.
// An operand of the grouped contractions as bf16 (hi, lo) planes in one of the engine's scratch plane buffers: split once
// (one coalesced pass, which also zero-fills the rows the weight-gradient form reads past a short group), then fed by TMA to
// every contraction that uses it -- the small-K aggregator contractions are latency-bound when their loader warps convert
// fp32 rows themselves (one 32 KB batch in flight per SM), and TMA keeps the whole smem ring in flight.
struct OpPl { Planes pl; bool ready; };
inline OpPl oppl(const Planes& pl) { OpPl o; o.pl = pl; o.ready = false; return o; }
inline bool dec_planes_on(const paths_engine* e) {
  return e->dec_planes && e->dm.precision != PREC_FP32_SIMT && e->stacked && e->dplD[0].p != nullptr;
}
int ensure_split(paths_engine* e, const Grp& rows, const float* A, int lda, int K, OpPl* op, cudaStream_t st, const MutL* colsum = nullptr) {
  if (op->ready) return 0;
  if (int r = split_rows_grp(A, lda, K, op->pl, rows, e->dm.precision == PREC_BF16X3_TC, colsum, st)) return r;
  op->ready = true;
  return 0;
}

template <typename WF>
int gemm_nt_set(paths_engine* e, const DecSet& S, const Grp& rows, const float* A, int lda, WF W, int ldb, int b_kn, float* C, int ldc,
                const PtrL& bias, int N, int K, int flags, cudaStream_t st, OpPl* ap = nullptr) {
  const int prec = e->dm.precision;
  if (ap && dec_planes_on(e) && K % 8 == 0 && lda == K) {
    GemmGrp gg;
    gg.rows = rows;
    for (int g = 0; g < kMaxGroups; ++g) { gg.wlev[g] = g < S.n ? S.lv[g] : 0; gg.bias[g] = bias.p[g]; gg.cout[g] = nullptr; }
    if (!ap->ready) e->launches++;
    if (int r = ensure_split(e, rows, A, lda, K, ap, st)) return r;
    return gemm_nt_tc_grp_p(prec == PREC_BF16X3_TC ? 3 : 1, ap->pl, W(0), ldb, b_kn, C, ldc, gg, N, K, flags, st);
  }
  if (prec != PREC_FP32_SIMT && e->stacked && gemm_tc_supported_nt_grp(row_src(A, lda, K), W(0), ldb, N, K)) {
    GemmGrp gg;
    gg.rows = rows;
    for (int g = 0; g < kMaxGroups; ++g) { gg.wlev[g] = g < S.n ? S.lv[g] : 0; gg.bias[g] = bias.p[g]; gg.cout[g] = nullptr; }
    return gemm_nt_tc_grp(prec == PREC_BF16X3_TC ? 3 : 1, row_src(A, lda, K), W(0), ldb, b_kn, C, ldc, gg, N, K, flags, st);
  }
  for (int g = 0; g < S.n; ++g) {
    const size_t r0 = rows.roff[g];
    if (int r = gemm_nt(prec, row_src(A + r0 * lda, lda, K), W(S.lv[g]), ldb, b_kn, C + r0 * ldc, ldc, bias.p[g], rows.hcnt[g], rows.cnt[g], N, K,
                        flags, st))
      return r;
  }
  return 0;
}
// dW_level[NI, NJ] += P rows^T . Q rows over every group's rows
template <typename GF>
int gemm_tn_set(paths_engine* e, const DecSet& S, const Grp& rows, const float* P, int ldp, const float* Q, int ldq, GF dW, int ldc, int NI, int NJ,
                cudaStream_t st, OpPl* pp = nullptr, OpPl* qp = nullptr) {
  const int prec = e->dm.precision;
  if (pp && qp && dec_planes_on(e) && NI % 8 == 0 && NJ % 8 == 0 && ldp == NI && ldq == NJ) {
    GemmGrp gg;
    gg.rows = rows;
    for (int g = 0; g < kMaxGroups; ++g) { gg.wlev[g] = 0; gg.bias[g] = nullptr; gg.cout[g] = g < S.n ? const_cast<float*>(dW(S.lv[g])) : nullptr; }
    e->launches += (pp->ready ? 0 : 1) + (qp->ready ? 0 : 1);
    if (int r = ensure_split(e, rows, P, ldp, NI, pp, st)) return r;
    if (int r = ensure_split(e, rows, Q, ldq, NJ, qp, st)) return r;
    return gemm_tn_tc_grp_p(prec == PREC_BF16X3_TC ? 3 : 1, pp->pl, qp->pl, ldc, gg, NI, NJ, st);
  }
  if (prec != PREC_FP32_SIMT && gemm_tc_supported_tn(ldp, row_src(Q, ldq, NJ), ldc, NI, NJ) && ldp % 8 == 0 && ldq % 8 == 0) {
    GemmGrp gg;
    gg.rows = rows;
    for (int g = 0; g < kMaxGroups; ++g) { gg.wlev[g] = 0; gg.bias[g] = nullptr; gg.cout[g] = g < S.n ? const_cast<float*>(dW(S.lv[g])) : nullptr; }
    return gemm_tn_tc_grp(prec == PREC_BF16X3_TC ? 3 : 1, P, ldp, row_src(Q, ldq, NJ), ldc, gg, NI, NJ, st);
  }
  for (int g = 0; g < S.n; ++g) {
    const size_t r0 = rows.roff[g];
    if (int r = gemm_tn(prec, P + r0 * ldp, ldp, row_src(Q + r0 * ldq, ldq, NJ), const_cast<float*>(dW(S.lv[g])), ldc, rows.hcnt[g], rows.cnt[g], NI, NJ,
                        st))
      return r;
  }
  return 0;
}

// keep/drop parameters of one dropout site of layer j for every level of the set (seeds as recorded by the levels' forward)
Drop set_drop(const paths_engine* e, const DecSet& S, int j, int site) {
  Drop dr = no_drop();
  const float p = e->lv_drop_p[S.lv[0]];
  dr.thresh = drop_thresh(p);
  dr.inv_keep = 1.f / (1.f - p);
  for (int g = 0; g < S.n; ++g) dr.key[g] = drop_site_key(e->lv_drop_seed[S.lv[g]], S.lv[g], j, site);
  return dr;
}

// ---- part 2: transformer aggregator -- decoder layers over packed tokens, last layer evaluated at the special token only.
// Reads only the set's tokens and writes only the set's rows of the decoder buffers (safe on the set's own stream).
int decoder_forward(paths_engine* e, const DecSet& S, cudaStream_t st) {
  const paths_dims_t& dm = e->dm;
  const int d = dm.d, F = dm.F;
  Level& v0 = e->lv[0];                                  // base pointers of the concatenated buffers
  const LevelP* P = e->P.level;
  ProfScope ps_dec(e, PATHS_PROF_DEC_FWD, st);
  unsigned set_mask = 0;
  for (int g = 0; g < S.n; ++g) set_mask |= 1u << S.lv[g];
  for (int g = 0; g < S.n; ++g) { e->sv_prec[S.lv[g]] = dec_planes_on(e) ? dm.precision : -1; e->sv_set[S.lv[g]] = set_mask; }
  for (int g = 0; g < S.n; ++g) { e->lv_drop_p[S.lv[g]] = e->drop_p; e->lv_drop_seed[S.lv[g]] = e->drop_seed; }
  const bool drop = e->drop_p > 0.f;                     // training-mode dropout at the five sites of nn.TransformerDecoderLayer
  const float* t_in = v0.tok;
  for (int j = 0; j < dm.layers; ++j) {
    LayerBuf& y = v0.layer[j];
    const bool last = j == dm.layers - 1;
    const Grp& gr = last ? S.gb : S.gt;
    float* A1 = last ? v0.A1b : v0.A1;
    float* Fo = last ? v0.Fob : v0.Fo;
    // (the last layer's slide-shaped operands have their own plane buffers: their rows alias token rows of other levels)
    OpPl pT = oppl(e->svT[j]), pO = oppl(e->svO[j]), pY2 = oppl(e->svY2[j]), pHf = oppl(e->svHf[j]);
    RUN(gemm_nt_set(e, S, S.gt, t_in, d, [&](int l) { return P[l].layer[j].Wqkv; }, d, 0, y.qkv, 3 * d,
                    PL(S, [&](int l) { return P[l].layer[j].bqkv; }), 3 * d, d, 0, st, &pT));
    const GIdx hrow = last ? S.tsegI : noidx();          // slide-shaped rows hash as their special token's row
    const PtrL bo = PL(S, [&](int l) { return P[l].layer[j].bo; }), bcross = PL(S, [&](int l) { return P[l].layer[j].bcross; }),
               b2 = PL(S, [&](int l) { return P[l].layer[j].b2; });
    Drop dA;
    if (drop) dA = set_drop(e, S, j, DROP_ATTN);
    RUN(attn_fwd_grp(y.qkv, S.sg, y.O, y.lse, d, dm.heads, S.max_T, last ? 1 : 0, st, drop ? &dA : nullptr, attn_tc_passes(dm), e->lv[S.lv[0]].attn_cls));
    RUN(gemm_nt_set(e, S, gr, y.O, d, [&](int l) { return P[l].layer[j].Wo; }, d, 0, A1, d, PL0(), d, d, 0, st, &pO));
    // dropout1(W_o a + b_o), dropout2(b_cross) and dropout3(W_2 h + b_2) are applied to the branch INCLUDING its bias before
    // the residual sum; the cross-attention branch (a bias only) is materialised in Fo, free until the FFN writes it
    if (drop) RUN(dropout_fwd_grp(A1, A1, d, bo, gr, hrow, set_drop(e, S, j, DROP_SA_OUT), st));
    RUN(res_ln_fwd_grp(t_in, last ? S.tsegI : noidx(), A1, drop ? PL0() : bo,
                       PL(S, [&](int l) { return P[l].layer[j].n1w; }), PL(S, [&](int l) { return P[l].layer[j].n1b; }), y.y1, y.xh1, y.rs1, d,
                       1e-5f, gr, st));
    if (drop) RUN(dropout_fwd_grp(nullptr, Fo, d, bcross, gr, hrow, set_drop(e, S, j, DROP_CROSS_OUT), st));
    RUN(res_ln_fwd_grp(y.y1, noidx(), drop ? Fo : nullptr, drop ? PL0() : bcross,
                       PL(S, [&](int l) { return P[l].layer[j].n2w; }), PL(S, [&](int l) { return P[l].layer[j].n2b; }), y.y2, y.xh2, y.rs2, d,
                       1e-5f, gr, st));
    RUN(gemm_nt_set(e, S, gr, y.y2, d, [&](int l) { return P[l].layer[j].W1; }, d, 0, y.Hf, F, PL(S, [&](int l) { return P[l].layer[j].b1; }), F,
                    d, GEMM_RELU, st, &pY2));
    if (drop) RUN(dropout_fwd_grp(y.Hf, y.Hf, F, PL0(), gr, hrow, set_drop(e, S, j, DROP_FFN_INNER), st));
    RUN(gemm_nt_set(e, S, gr, y.Hf, F, [&](int l) { return P[l].layer[j].W2; }, F, 0, Fo, d, PL0(), d, F, 0, st, &pHf));
    if (drop) RUN(dropout_fwd_grp(Fo, Fo, d, b2, gr, hrow, set_drop(e, S, j, DROP_FFN_OUT), st));
    RUN(res_ln_fwd_grp(y.y2, noidx(), Fo, drop ? PL0() : b2, PL(S, [&](int l) { return P[l].layer[j].n3w; }),
                       PL(S, [&](int l) { return P[l].layer[j].n3b; }), y.y3, y.xh3, y.rs3, d, 1e-5f, gr, st));
    t_in = y.y3;
  }
  RUN(res_ln_fwd_grp(t_in, noidx(), nullptr, PL0(), PL(S, [&](int l) { return P[l].nfw; }), PL(S, [&](int l) { return P[l].nfb; }), v0.z, v0.xhf,
                     v0.rsf, d, 1e-5f, S.gb, st));
  return 0;
}

// ---- part 3: residual slide context + classifier (paths.py:130-139); needs F of the previous level
int head_forward(paths_engine* e, int l, cudaStream_t st) {
  const paths_dims_t& dm = e->dm;
  const int B = e->B, d = dm.d;
  Level& v = e->lv[l];
  const LevelP& p = e->P.level[l];
  if (dm.residual == 2) {
    RUN(head_fwd(v.z, nullptr, nullptr, nullptr, v.F, nullptr, d, 0, B, st));          // F_l = z_l
    FList fl; fl.n = l + 1;
    for (int j = 0; j <= l; ++j) fl.f[j] = e->lv[j].F;
    RUN(head_concat_fwd(fl, p.Wc, p.bc, v.logits, d, dm.nl, B, st));
  } else {
    RUN(head_fwd(v.z, (dm.residual == 1 && l > 0) ? e->lv[l - 1].F : nullptr, p.Wc, p.bc, v.F, v.logits, d, dm.nl, B, st));
  }
  return 0;
}

int level_forward(paths_engine* e, int l, cudaStream_t st) {
  if (int r = lstm_forward(e, l, st)) return r;
  if (int r = decoder_forward(e, make_set(e, &l, 1), st)) return r;
  return head_forward(e, l, st);
}

// ---- backward of one level, part 1: classifier + residual slide context -> dFl (needs dFl of level l+1)
int head_backward(paths_engine* e, int l, int use_dlogits, cudaStream_t st) {
  const paths_dims_t& dm = e->dm;
  const int B = e->B, d = dm.d, prec = dm.precision;
  Level& v = e->lv[l];
  const LevelP& p = e->P.level[l];
  const LevelP& g = e->Gr.level[l];
  const bool last_level = l == dm.L - 1;
  const float* dlog = use_dlogits ? v.dlogits : nullptr;
  if (dm.residual == 2) {
    // concat: the classifier of the level whose logits reach the loss sees every F_j, j <= l (paths.py:134-137)
    if (dlog) {
      GList gl; gl.n = l + 1;
      for (int j = 0; j <= l; ++j) gl.g[j] = e->lv[j].dFc;
      RUN(head_concat_bwd(dlog, p.Wc, gl, d, dm.nl, B, st));
      for (int j = 0; j <= l; ++j)
        RUN(gemm_tn(prec, dlog, dm.nl, row_src(e->lv[j].F, d, d), W_(g.Wc) + (size_t)j * d, (l + 1) * d, B, nullptr, dm.nl, d, st));
      RUN(colsum_accum(dlog, dm.nl, W_(g.bc), B, nullptr, dm.nl, st));
    } else if (last_level) {
      RUN(zero_rows(v.dFc, d, B, nullptr, 0, st));
    }
    RUN(head_bwd(nullptr, v.dFc, p.Wc, v.dFl, d, dm.nl, B, st));
  } else {
    const float* dF_in = (dm.residual == 1 && !last_level) ? e->lv[l + 1].dFl : nullptr;
    RUN(head_bwd(dlog, dF_in, p.Wc, v.dFl, d, dm.nl, B, st));
    if (dlog) {
      RUN(gemm_tn(prec, dlog, dm.nl, row_src(v.F, d, d), W_(g.Wc), d, B, nullptr, dm.nl, d, st));
      RUN(colsum_accum(dlog, dm.nl, W_(g.bc), B, nullptr, dm.nl, st));
    }
  }
  return 0;
}

// ---- part 2: aggregator backward of a level set, dFl -> gradient of the tokens (e->dtok_base) and of the decoder parameters.
// Touches only the set's rows of the decoder buffers, the set's workspace and parameter gradients (safe on the set's own stream).
int decoder_backward(paths_engine* e, const DecSet& S, int ws_slot, cudaStream_t st) {
  const paths_dims_t& dm = e->dm;
  set_tn_workspace(e->dec_tn_ws[ws_slot], e->dec_tn_ws_bytes);
  const int B = e->B, d = dm.d, F = dm.F;
  Level& v0 = e->lv[0];
  const LevelP* P = e->P.level;
  const LevelP* G = e->Gr.level;
  ProfScope ps_dec(e, PATHS_PROF_DEC_BWD, st);
  // the forward pass left its operand planes, zero-filled to the longest level of ITS set: usable when this set is inside it
  bool saved = dec_planes_on(e);
  unsigned set_mask = 0;
  for (int g = 0; g < S.n; ++g) set_mask |= 1u << S.lv[g];
  for (int g = 0; g < S.n; ++g) saved = saved && e->sv_prec[S.lv[g]] == dm.precision && (set_mask & ~e->sv_set[S.lv[g]]) == 0;
  // ---- final LayerNorm at the special tokens
  RUN(ln_bwd_grp(v0.dFl, nullptr, v0.xhf, v0.rsf, PL(S, [&](int l) { return P[l].nfw; }), v0.dy3s, noidx(), 0,
                 ML(S, [&](int l) { return G[l].nfw; }), ML(S, [&](int l) { return G[l].nfb; }), ML0(), d, S.gb, st));

  const float* dy3 = v0.dy3s;
  float* ga = v0.ga; float* gtk = v0.gt;                 // token-shaped ping-pong pair
  for (int j = dm.layers - 1; j >= 0; --j) {
    LayerBuf& y = v0.layer[j];
    const bool last = j == dm.layers - 1;
    const Grp& gr = last ? S.gb : S.gt;
    float* gA = last ? v0.gab : ga;
    float* gB = last ? v0.gbb : v0.gb;
    float* gO = last ? v0.gOb : v0.gO;
    float* dHf = last ? v0.dHfb : v0.dHf;
    const float* t_in = j == 0 ? v0.tok : v0.layer[j - 1].y3;
    auto Wl = [&](int l) -> const LayerP& { return P[l].layer[j]; };
    auto Gl = [&](int l) -> const LayerP& { return G[l].layer[j]; };
    // With dropout the gradient of a branch is the masked, rescaled copy of the residual-sum gradient (and its column sums the
    // gradient of the branch's bias): it is written to gB, which is free at both points where a copy is needed.
    const bool drop = e->lv_drop_p[S.lv[0]] > 0.f;
    for (int g = 1; g < S.n; ++g)
      PATHS_REQUIRE((e->lv_drop_p[S.lv[g]] > 0.f) == drop && e->lv_drop_p[S.lv[g]] == e->lv_drop_p[S.lv[0]],
                    "decoder backward: the levels of a set ran their forward passes with different dropout");
    const GIdx hrow = last ? S.tsegI : noidx();
    // LN3 -> gA = d(y2 + ffn); db2
    RUN(ln_bwd_grp(dy3, nullptr, y.xh3, y.rs3, PL(S, [&](int l) { return Wl(l).n3w; }), gA, noidx(), 0, ML(S, [&](int l) { return Gl(l).n3w; }),
                   ML(S, [&](int l) { return Gl(l).n3b; }), drop ? ML0() : ML(S, [&](int l) { return Gl(l).b2; }), d, gr, st));
    const float* gF = gA;                                  // gradient of the FFN branch output
    if (drop) {
      RUN(dropout_bwd_grp(gA, gB, d, ML(S, [&](int l) { return Gl(l).b2; }), gr, hrow, set_drop(e, S, j, DROP_FFN_OUT), st));
      gF = gB;
    }
    // operand planes: each tensor is split once and feeds both the data-gradient and the weight-gradient contraction
    OpPl pA = oppl(last ? e->dplDb[0] : e->dplD[0]), pHf = oppl(saved ? e->svHf[j] : last ? e->dplFb[0] : e->dplF[0]),
         pdHf = oppl(last ? e->dplFb[1] : e->dplF[1]), pY2 = oppl(saved ? e->svY2[j] : last ? e->dplDb[1] : e->dplD[1]);
    pHf.ready = pY2.ready = saved;
    RUN(gemm_nt_set(e, S, gr, gF, d, [&](int l) { return Wl(l).W2; }, F, 1, dHf, F, PL0(), F, d, 0, st, &pA));
    RUN(gemm_tn_set(e, S, gr, gF, d, y.Hf, F, [&](int l) { return Gl(l).W2; }, F, d, F, st, &pA, &pHf));
    const float inner_scale = drop ? 1.f / (1.f - e->lv_drop_p[S.lv[0]]) : 1.f;    // Hf was saved after the inner dropout
    if (dec_planes_on(e) && F % 8 == 0) {                  // masked gradient straight to operand planes (no fp32 write, no split pass)
      RUN(relu_bwd_grp(dHf, y.Hf, F, gr, ML(S, [&](int l) { return Gl(l).b1; }), st, &pdHf.pl, e->dm.precision == PREC_BF16X3_TC, inner_scale));
      pdHf.ready = true;
    } else {
      RUN(relu_bwd_grp(dHf, y.Hf, F, gr, ML(S, [&](int l) { return Gl(l).b1; }), st, nullptr, 0, inner_scale));
    }
    RUN(gemm_tn_set(e, S, gr, dHf, F, y.y2, d, [&](int l) { return Gl(l).W1; }, d, F, d, st, &pdHf, &pY2));
    RUN(gemm_nt_set(e, S, gr, dHf, F, [&](int l) { return Wl(l).W1; }, d, 1, gA, d, PL0(), d, F, GEMM_ACCUM, st, &pdHf));     // gA = dy2
    // LN2 (cross-attention over an empty memory contributes only its output bias), then LN1
    RUN(ln_bwd_grp(gA, nullptr, y.xh2, y.rs2, PL(S, [&](int l) { return Wl(l).n2w; }), gB, noidx(), 0, ML(S, [&](int l) { return Gl(l).n2w; }),
                   ML(S, [&](int l) { return Gl(l).n2b; }), drop ? ML0() : ML(S, [&](int l) { return Gl(l).bcross; }), d, gr, st));
    if (drop)                                              // the cross-attention branch is its bias alone: only d b_cross
      RUN(dropout_bwd_grp(gB, nullptr, d, ML(S, [&](int l) { return Gl(l).bcross; }), gr, hrow, set_drop(e, S, j, DROP_CROSS_OUT), st));
    RUN(ln_bwd_grp(gB, nullptr, y.xh1, y.rs1, PL(S, [&](int l) { return Wl(l).n1w; }), gA, noidx(), 0, ML(S, [&](int l) { return Gl(l).n1w; }),
                   ML(S, [&](int l) { return Gl(l).n1b; }), drop ? ML0() : ML(S, [&](int l) { return Gl(l).bo; }), d, gr, st));      // gA = da
    const float* gS = gA;                                  // gradient of the self-attention branch output
    if (drop) {
      RUN(dropout_bwd_grp(gA, gB, d, ML(S, [&](int l) { return Gl(l).bo; }), gr, hrow, set_drop(e, S, j, DROP_SA_OUT), st));
      gS = gB;
    }
    OpPl pA2 = oppl(last ? e->dplDb[0] : e->dplD[0]), pO = oppl(saved ? e->svO[j] : last ? e->dplDb[1] : e->dplD[1]);
    pO.ready = saved;
    RUN(gemm_nt_set(e, S, gr, gS, d, [&](int l) { return Wl(l).Wo; }, d, 1, gO, d, PL0(), d, d, 0, st, &pA2));
    RUN(gemm_tn_set(e, S, gr, gS, d, y.O, d, [&](int l) { return Gl(l).Wo; }, d, d, d, st, &pA2, &pO));
    if (last && (!attn_bwd_writes_all(d, dm.heads) || drop))
      for (int g = 0; g < S.n; ++g)
        RUN(zero_rows(v0.dqkv + (size_t)S.gt.roff[g] * 3 * d, 3 * d, S.gt.hcnt[g], S.gt.cnt[g], 0, st));
    Drop dA;
    if (drop) dA = set_drop(e, S, j, DROP_ATTN);
    RUN(attn_bwd_grp(y.qkv, S.sg, y.O, y.lse, gO, v0.dqkv, last ? v0.Dvecb : v0.Dvec, d, dm.heads, S.max_T, last ? 1 : 0, st, drop ? &dA : nullptr,
                     attn_tc_passes(dm), e->lv[S.lv[0]].attn_cls));
    OpPl p3 = oppl(e->dpl3), pT = oppl(saved ? e->svT[j] : e->dplD[0]);
    pT.ready = saved;
    const MutL dbqkv = ML(S, [&](int l) { return Gl(l).bqkv; });
    const bool fused_bias = dec_planes_on(e) && 3 * d <= 512 && d % 8 == 0;
    if (fused_bias) RUN(ensure_split(e, S.gt, v0.dqkv, 3 * d, 3 * d, &p3, st, &dbqkv));      // split pass also sums the columns: d b_qkv
    RUN(gemm_tn_set(e, S, S.gt, v0.dqkv, 3 * d, t_in, d, [&](int l) { return Gl(l).Wqkv; }, d, 3 * d, d, st, &p3, &pT));
    if (!fused_bias) RUN(colsum_accum_grp(v0.dqkv, 3 * d, dbqkv, S.gt, 3 * d, st));
    if (last) {
      RUN(gemm_nt_set(e, S, S.gt, v0.dqkv, 3 * d, [&](int l) { return Wl(l).Wqkv; }, d, 1, gtk, d, PL0(), d, 3 * d, 0, st, &p3));
      RUN(scatter_add_rows_grp(gA, S.tsegI, gtk, d, S.gb, st));
      dy3 = gtk;
    } else {
      // ga holds da (the residual branch); add the attention branch, then hand the buffer over as the next layer's dy3
      RUN(gemm_nt_set(e, S, S.gt, v0.dqkv, 3 * d, [&](int l) { return Wl(l).Wqkv; }, d, 1, ga, d, PL0(), d, 3 * d, GEMM_ACCUM, st, &p3));
      float* tmp = ga; ga = gtk; gtk = tmp;
      dy3 = gtk;
    }
  }
  e->dtok_base = dy3;
  RUN(special_grad_grp(dy3, S.segI, ML(S, [&](int l) { return G[l].special; }), d, B, S.n, st));
  return 0;
}

// ---- part 3: importance / projection, LSTM (or RNN) step, gradients handed to the parent rows
int lstm_backward(paths_engine* e, int l, cudaStream_t st) {
  const paths_dims_t& dm = e->dm;
  set_tc_scratch(e->tc_scratch[0], e->tc_scratch_each, e->tc_scratch[1], e->tc_scratch_each);
  set_tn_workspace(e->tn_ws, e->tn_ws_bytes);
  const int D = dm.D, C = dm.C, d = dm.d, I = dm.I, prec = dm.precision;
  Level& v = e->lv[l];
  const LevelP& p = e->P.level[l];
  const LevelP& g = e->Gr.level[l];
  const int M = v.cap;
  const int* Md = v.count;
  const int ldg = 3 * C + D;
  const bool last_level = l == dm.L - 1;
  ProfScope ps_blk(e, PATHS_PROF_LSTM_BWD, st);
  const float* dtok = e->dtok_base + (size_t)v.tok_off * d;
  const bool has_children = !last_level;
  const bool has_parent = l > 0;
  if (!dm.lstm) {
    // ---- RNN variant: tok = Z.W_in^T + b_in + PE;  Z = alpha X + hctx;  hctx = Wh2 relu(Wh0 Z_parent + bh0) + bh2
    const bool hctx = has_parent && dm.hier_ctx;
    RUN(imp_tok_bwd(dtok, v.UV, v.alpha, p.w2, v.row_slide, e->dUV, no_planes(), 0, W_(g.b1), W_(g.w2), W_(g.b2), W_(g.bin), I, d, dm.imp_mul, 2,
                    M, Md, st));
    RUN(gemm_nt(prec, row_src(e->dUV + I, I + d, d), p.Wuv + (size_t)I * D, D, 1, e->dY, D, nullptr, M, Md, D, d, 0, st));   // dZ
    RUN(gemm_tn(prec, e->dUV + I, I + d, row_src(v.H, D, D), W_(g.Wuv) + (size_t)I * D, D, M, Md, d, D, st));
    if (has_children && dm.hier_ctx)
      RUN(parent_sum(v.child_of, e->lv[l + 1].dHrow, nullptr, e->dHc, nullptr, D, 0, M, Md, st));
    RUN(rnn_alpha_bwd(e->dY, (has_children && dm.hier_ctx) ? e->dHc : nullptr, v.X, v.UV, v.alpha, p.w2, e->dUV, W_(g.b1), W_(g.w2),
                      W_(g.b2), D, I, d, dm.imp_mul, M, Md, st));
    RUN(gemm_tn(prec, e->dUV, I + d, row_src(v.X, D, D), W_(g.Wuv), D, M, Md, I, D, st));
    if (hctx) {
      RUN(gemm_tn(prec, e->dY, D, row_src(v.Cn, C, C), W_(g.Wh2), C, M, Md, D, C, st));
      RUN(colsum_accum(e->dY, D, W_(g.bh2), M, Md, D, st));
      RUN(gemm_nt(prec, row_src(e->dY, D, D), p.Wh2, C, 1, e->dCn, C, nullptr, M, Md, C, D, 0, st));
      RUN(relu_bwd(e->dCn, v.Cn, C, M, Md, 0, nullptr, st));
      RUN(gemm_tn(prec, e->dCn, C, row_src(e->lv[l - 1].H, D, D, v.row_parent), W_(g.Wh0), D, M, Md, C, D, st));
      RUN(colsum_accum(e->dCn, C, W_(g.bh0), M, Md, C, st));
      RUN(gemm_nt(prec, row_src(e->dCn, C, C), p.Wh0, D, 1, v.dHrow, D, nullptr, M, Md, D, C, 0, st));
    }
    return 0;
  }
  const bool pl = v.ran_pl;
  const int lo = prec == PREC_BF16X3_TC ? 1 : 0;
  const Planes none = no_planes();
  RUN(imp_tok_bwd(dtok, v.UV, v.alpha, p.w2, v.row_slide, pl ? nullptr : e->dUV, pl ? e->dUVP : none, lo, W_(g.b1), W_(g.w2), W_(g.b2),
                  W_(g.bin), I, d, dm.imp_mul, 0, M, Md, st));
  const Act aUV = act(e->dUV, I + d, pl ? e->dUVP : none), aY = act(v.Y, D, pl ? v.YP : none);
  RUN(gemm_nt_act(prec, aUV, p.Wuv, D, 1, e->dY, D, nullptr, M, Md, D, I + d, 0, st));
  RUN(gemm_tn_act(prec, aUV, aY, nullptr, W_(g.Wuv), D, M, Md, I + d, D, st));
  // ---- state gradients arriving from the children of this level's selected rows
  // ---- LSTM cell (the pointwise kernels sum the state gradients of a row's children (slide.py:318-329) as they read them, and
  //      accumulate the bias gradients of mem_to_out and of the four gates)
  const int* kids = has_children ? v.child_of : nullptr;
  // (a shared-parent child level leaves ONE dH row per slot: slot4 is the child table that reads it)
  const int* kids_h = has_children ? (e->lv[l + 1].ran_shared ? v.slot4 : v.child_of) : nullptr;
  RUN(lstm_point2_bwd(e->dY, has_children ? e->lv[l + 1].dHrow : nullptr, kids_h, v.Pt, v.G, pl ? nullptr : e->dP, pl ? e->dPP : none, pl ? nullptr : e->dG,
                      pl ? e->dGP : none, W_(e->Gr.bmo), W_(e->Gr.bg) + 3 * C, lo, D, C, M, Md, st));
  const Act aP = act(e->dP, D, pl ? e->dPP : none), aCn = act(v.Cn, C, pl ? v.CnP : none);
  RUN(gemm_nt_act(prec, aP, e->P.Wmo, C, 1, e->dCn, C, nullptr, M, Md, C, D, 0, st));
  RUN(gemm_tn_act(prec, aP, aCn, nullptr, W_(e->Gr.Wmo), C, M, Md, D, C, st));
  RUN(lstm_point1_bwd(e->dCn, has_children ? e->lv[l + 1].dCrow : nullptr, kids, v.G, has_parent ? e->lv[l - 1].Cn : nullptr,
                      has_parent ? v.row_parent : nullptr, pl ? nullptr : e->dG, pl ? e->dGP : none, has_parent ? v.dCrow : nullptr,
                      W_(e->Gr.bg), lo, D, C, M, Md, st));
  const Act aG = act(e->dG, ldg, pl ? e->dGP : none);
  if (v.ran_shared) {
    // dH_parent = (sum of dG over the parent's children) W_h and dW_h = (that sum)^T h_parent: slot-shaped, not row-shaped
    Level& pv = e->lv[l - 1];
    const int slots = e->B * pv.kcap;
    RUN(child_sum_planes(e->dGP, pv.child_of, pv.sel_rows, slots, ldg, e->dGsumP, lo, st));
    const Act aGs = act(nullptr, 0, e->dGsumP);
    {
      ProfScope ps(e, PATHS_PROF_GATES_DH, st);
      RUN(gemm_nt_act(prec, aGs, e->P.Wg + D, 2 * D, 1, v.dHrow, D, nullptr, slots, nullptr, D, ldg, 0, st));
    }
    ProfScope ps(e, PATHS_PROF_GATES_DW, st);
    RUN(gemm_tn_act(prec, aG, act(nullptr, 0, v.AxhP), nullptr, W_(e->Gr.Wg), 2 * D, M, Md, ldg, D, st));
    RUN(gemm_tn_act(prec, aGs, act(nullptr, 0, v.HselP), nullptr, W_(e->Gr.Wg) + D, 2 * D, slots, nullptr, ldg, D, st));
    return 0;
  }
  if (has_parent) {
    ProfScope ps(e, PATHS_PROF_GATES_DH, st);
    RUN(gemm_nt_act(prec, aG, e->P.Wg + D, 2 * D, 1, v.dHrow, D, nullptr, M, Md, D, ldg, 0, st));
  }
  RowSrc Axh = has_parent ? row_src2(v.X, D, nullptr, D, e->lv[l - 1].H, D, v.row_parent) : row_src(v.X, D, D);
  {
    ProfScope ps(e, PATHS_PROF_GATES_DW, st);
    RUN(gemm_tn_act(prec, aG, act(nullptr, 0, pl ? v.AxhP : none), &Axh, W_(e->Gr.Wg), 2 * D, M, Md, ldg, has_parent ? 2 * D : D, st));
  }
  return 0;
}

int level_backward(paths_engine* e, int l, int use_dlogits, cudaStream_t st) {
  if (int r = head_backward(e, l, use_dlogits, st)) return r;
  if (int r = decoder_backward(e, make_set(e, &l, 1), l, st)) return r;
  return lstm_backward(e, l, st);
}

int do_select(paths_engine* e, int l, cudaStream_t st) {
  const paths_dims_t& dm = e->dm;
  PATHS_REQUIRE(l >= 0 && l < dm.L - 1, "select: level %d has no successor", l);
  PATHS_REQUIRE(e->have_store, "select: no pyramid store bound");
  Level& v = e->lv[l];
  Level& n = e->lv[l + 1];
  ProfScope ps_sel(e, PATHS_PROF_SELECT, st);
  SelectArgs a;
  a.alpha = v.alpha; a.seg = v.seg; a.row_xy = v.row_xy; a.count = v.count; a.B = e->B; a.cap = v.cap;
  a.keep = dm.K[l]; a.kcap = v.kcap; a.maxseg = v.maxseg; a.n_maxseg = n.maxseg;
  a.gw = e->store.gw; a.gh = e->store.gh; a.occ_off = reinterpret_cast<const unsigned long long*>(e->store.occ_off);
  a.occ = e->store.occ; a.nonbg = e->store.nonbg; a.level = l; a.L = dm.L; a.slide_ids = e->slide_ids;
  a.sel = v.sel; a.sel_count = v.sel_count; a.child_of = v.child_of;
  a.n_count = n.count; a.n_seg = n.seg; a.n_row_slide = n.row_slide; a.n_row_xy = n.row_xy; a.n_row_parent = n.row_parent;
  a.n_parent_local = n.parent_local; a.n_cap = n.cap; a.n_ntok = n.ntok;
  a.overflow = e->overflow; a.slide_cnt = e->slide_cnt; a.slide_fb = e->slide_fb;
  RUN(select_expand(a, st)); e->launches += 2;
  pdl(tseg_kernel, (e->B + 255) / 256, 256, 0, st)(n.seg, n.tseg, e->B);
  PATHS_LAUNCH_CK(); e->launches++;
  // a shared-parent level's operand planes hold x alone: the gather writes them as it goes
  n.x_planes = n.shared && planes_mode(dm);
  RUN(gather_rows(e->store.grid_ptr, e->store.gh, l + 1, dm.L, e->slide_ids, n.row_slide, n.row_xy, n.X, dm.D, n.cap, n.count, st, e->store.bf16,
                  n.x_planes ? &n.AxhP : nullptr, dm.precision == PREC_BF16X3_TC));
  n.tables_ok = n.shared;
  if (n.shared)
    RUN(parent_slots(v.child_of, v.count, v.cap, v.sel, v.sel_count, e->B, v.kcap, n.row_slide, n.row_parent, n.parent_local, n.count, n.cap,
                     v.slot4, v.sel_rows, n.add_idx, st));
  return 0;
}

// Publishes the overflow flag as it stands after the selects enqueued so far (see paths_engine::h_over).
int publish_overflow(paths_engine* e, cudaStream_t st) {
  PATHS_CK(cudaMemcpyAsync(e->h_over, e->overflow, 2 * sizeof(int), cudaMemcpyDeviceToHost, st));
  // inside a stream capture (CUDA-graph replay of the step) a plain record would only become a graph edge: the host waits on
  // this event, so it must be a real event-record node of the graph
  cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
  PATHS_CK(cudaStreamIsCapturing(st, &cap));
  if (cap == cudaStreamCaptureStatusActive) PATHS_CK(cudaEventRecordWithFlags(e->ev_sel, st, cudaEventRecordExternal));
  else PATHS_CK(cudaEventRecord(e->ev_sel, st));
  e->sel_recorded = true;
  return 0;
}

// N > 0: fts is the collated zero-padded batch [B, N, D];  N < 0: fts holds the level-0 rows already packed ([-N, D], slide-major,
// cells x-major: exactly the order of the level-0 row table);  fts == nullptr: rows are gathered from the pyramid store.
int do_begin(paths_engine* e, const int32_t* slide_ids, const float* fts_padded, int N, cudaStream_t st) {
  PATHS_REQUIRE(e->have_store, "begin: no pyramid store bound");
  e->slide_ids = slide_ids;
  Level& v = e->lv[0];
  if (e->dm.precision != PREC_FP32_SIMT && e->have_params) { RUN(run_pack_jobs(e->pack_jobs, e->n_pack_jobs, st)); }
  PATHS_CK(cudaMemsetAsync(e->overflow, 0, 2 * sizeof(int), st));
  RUN(level0_rows(e->store.gw, e->store.gh, e->dm.L, slide_ids, e->B, v.cap, v.count, v.ntok, v.seg, v.row_slide, v.row_xy,
                  v.row_parent, v.parent_local, e->overflow, st));
  e->launches++;
  pdl(tseg_kernel, (e->B + 255) / 256, 256, 0, st)(v.seg, v.tseg, e->B);
  PATHS_LAUNCH_CK(); e->launches++;
  v.x_planes = false;
  if (fts_padded && N < 0) {
    PATHS_REQUIRE(-N <= v.cap, "begin: %d packed level-0 rows exceed the level's capacity %d", -N, v.cap);
    PATHS_CK(cudaMemcpyAsync(v.X, fts_padded, (size_t)(-N) * e->dm.D * sizeof(float), cudaMemcpyDeviceToDevice, st));
  } else if (fts_padded) RUN(pack_padded_rows(fts_padded, N, v.seg, v.row_slide, v.X, e->dm.D, v.cap, v.count, st));
  else {
    v.x_planes = planes_mode(e->dm);
    RUN(gather_rows(e->store.grid_ptr, e->store.gh, 0, e->dm.L, slide_ids, v.row_slide, v.row_xy, v.X, e->dm.D, v.cap, v.count, st, e->store.bf16,
                    v.x_planes ? &v.AxhP : nullptr, e->dm.precision == PREC_BF16X3_TC));
  }
  return 0;
}

int do_loss(paths_engine* e, const int32_t* ybin, const int32_t* cens, float scale, cudaStream_t st) {
  Level& v = e->lv[e->dm.L - 1];
  RUN(loss_fwd_bwd(v.logits, ybin, cens, e->hazards, e->loss_per, e->loss, e->risk, v.dlogits, e->dm.nl, e->B, e->dm.task, scale, st));
  e->launches++;
  return 0;
}

}  // namespace

// ---- C ABI ---------------------------------------------------------------------------------------------------
extern "C" {

static int register_weights(paths_engine_t* e);
static int ensure_registry(paths_engine_t* e, cudaStream_t st);

const char* paths_last_error(void) { return paths::g_err; }
int paths_abi_version(void) { return PATHS_ABI_VERSION; }

size_t paths_engine_workspace_bytes(const paths_dims_t* dims, int B, const int32_t* h_caps, const int32_t* h_maxseg) {
  if (check_dims(dims, B, h_caps, h_maxseg)) return 0;
  paths_engine tmp = paths_engine();
  tmp.dm = *dims; tmp.B = B;
  fill_level_meta(&tmp, h_caps, h_maxseg);
  return carve(&tmp, nullptr);
}

int paths_engine_create(paths_engine_t** out, const paths_dims_t* dims, int B, const int32_t* h_caps, const int32_t* h_maxseg,
                        void* workspace, size_t workspace_bytes) {
  PATHS_REQUIRE(out && workspace, "null argument");
  if (int r = check_dims(dims, B, h_caps, h_maxseg)) return r;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
    set_error("no CUDA device: paths_b200 has no CPU fallback");
    return -3;
  }
  paths_engine* e = new (std::nothrow) paths_engine();
  PATHS_REQUIRE(e, "out of host memory");
  e->dm = *dims; e->B = B;
  fill_level_meta(e, h_caps, h_maxseg);
  const size_t need = carve(e, nullptr);
  if (need > workspace_bytes || (reinterpret_cast<uintptr_t>(workspace) & 255) != 0) {
    set_error("workspace too small or misaligned: need %zu bytes (256-byte aligned), got %zu", need, workspace_bytes);
    delete e;
    return -2;
  }
  carve(e, static_cast<char*>(workspace));
  e->ws_base = static_cast<char*>(workspace); e->ws_bytes = need;
  e->site_on = getenv("PATHS_B200_PROF_DUMP") != nullptr;
  e->dec_planes = getenv("PATHS_B200_DEC_NO_PLANES") == nullptr;
  for (int l = 0; l < PATHS_MAX_LEVELS; ++l) { e->sv_prec[l] = -1; e->sv_set[l] = 0; e->lv_drop_p[l] = 0.f; e->lv_drop_seed[l] = 0; }
  e->drop_p = 0.f; e->drop_seed = 0;
  {
    // "012,34": comma-separated sets of level digits (PATHS_B200_DEC_SETS_FWD / _BWD); a forward set starts once the tokens of
    // its deepest level exist, backward sets are launched last-to-first.
    auto parse = [&](const char* spec, int* n_sets, int* set_n, int (*set_lv)[PATHS_MAX_LEVELS]) {
      const int L = e->dm.L;
      bool seen[PATHS_MAX_LEVELS] = {false};
      int ns = 0;
      bool ok = spec != nullptr && *spec != 0;
      if (ok) {
        set_n[0] = 0;
        for (const char* c = spec; *c && ok; ++c) {
          if (*c == ',') { if (set_n[ns] == 0) ok = false; else if (++ns >= PATHS_MAX_LEVELS) ok = false; else set_n[ns] = 0; continue; }
          const int l = *c - '0';
          if (l < 0 || l >= L || seen[l]) { ok = false; break; }
          seen[l] = true;
          set_lv[ns][set_n[ns]++] = l;
        }
        if (ok && set_n[ns] == 0) ok = false;
        for (int l = 0; l < L && ok; ++l) ok = seen[l];
        if (ok) ++ns;
      }
      if (!ok) {
        // default: levels 0..L-2 as one grouped pass (it overlaps the deepest level's LSTM step), the deepest level on its own
        // (forward: its tokens come last; backward: its token gradient is what the LSTM chain waits for first)
        if (L == 1) { ns = 1; set_n[0] = 1; set_lv[0][0] = 0; }
        else {
          ns = 2; set_n[0] = L - 1; set_n[1] = 1; set_lv[1][0] = L - 1;
          for (int l = 0; l < L - 1; ++l) set_lv[0][l] = l;
        }
      }
      *n_sets = ns;
    };
    parse(getenv("PATHS_B200_DEC_SETS_FWD"), &e->n_fsets, e->fset_n, e->fset_lv);
    parse(getenv("PATHS_B200_DEC_SETS_BWD"), &e->n_bsets, e->bset_n, e->bset_lv);
  }
  e->have_streams = true;
  // priorities: the chain every level waits on runs highest; among the decoders the deeper level goes first (its token
  // gradient is the first one the backward chain needs)
  int prio_least = 0, prio_greatest = 0;
  cudaDeviceGetStreamPriorityRange(&prio_least, &prio_greatest);
  {
    bool ok = cudaStreamCreateWithPriority(&e->mainp, cudaStreamNonBlocking, prio_greatest) == cudaSuccess;
    ok = ok && cudaEventCreateWithFlags(&e->ev_in, cudaEventDisableTiming) == cudaSuccess;
    ok = ok && cudaEventCreateWithFlags(&e->ev_out, cudaEventDisableTiming) == cudaSuccess;
    ok = ok && cudaEventCreateWithFlags(&e->ev_sel, cudaEventDisableTiming) == cudaSuccess;
    ok = ok && cudaHostAlloc(reinterpret_cast<void**>(&e->h_over), 2 * sizeof(int), cudaHostAllocDefault) == cudaSuccess;
    if (ok) { e->h_over[0] = e->h_over[1] = 0; e->sel_recorded = false; }
    if (!ok) {
      set_error("could not create the engine streams / events: %s", cudaGetErrorString(cudaGetLastError()));
      delete e;
      return -1;
    }
  }
  for (int l = 0; l < e->dm.L && e->have_streams; ++l) {
    int prio = prio_greatest + 1 + (e->dm.L - 1 - l);
    if (prio > prio_least) prio = prio_least;
    bool ok = cudaStreamCreateWithPriority(&e->side[l], cudaStreamNonBlocking, prio) == cudaSuccess;
    ok = ok && cudaEventCreateWithFlags(&e->ev_tok[l], cudaEventDisableTiming) == cudaSuccess;
    ok = ok && cudaEventCreateWithFlags(&e->ev_z[l], cudaEventDisableTiming) == cudaSuccess;
    ok = ok && cudaEventCreateWithFlags(&e->ev_head[l], cudaEventDisableTiming) == cudaSuccess;
    ok = ok && cudaEventCreateWithFlags(&e->ev_dtok[l], cudaEventDisableTiming) == cudaSuccess;
    ok = ok && cudaEventCreateWithFlags(&e->ev_grad[l], cudaEventDisableTiming) == cudaSuccess;
    if (!ok) {
      set_error("could not create the per-level streams / events: %s", cudaGetErrorString(cudaGetLastError()));
      delete e;
      return -1;
    }
  }
  *out = e;
  return 0;
}

void paths_engine_destroy(paths_engine_t* e) {
  if (!e) return;
  if (e->site_on && !e->sites.empty()) {
    cudaDeviceSynchronize();
    struct Acc { double ms; int n; const char* what; };
    std::vector<Acc> acc(4096, Acc{0.0, 0, nullptr});
    for (const paths_engine::Site& s : e->sites) {
      float ms = 0.f;
      if (cudaEventElapsedTime(&ms, s.a, s.b) == cudaSuccess && s.line < 4096) { acc[s.line].ms += ms; acc[s.line].n++; acc[s.line].what = s.what; }
      cudaEventDestroy(s.a); cudaEventDestroy(s.b);
    }
    if (FILE* f = fopen(getenv("PATHS_B200_PROF_DUMP"), "w")) {
      for (int i = 0; i < 4096; ++i)
        if (acc[i].n) fprintf(f, "%d\t%.4f\t%d\t%.80s\n", i, acc[i].ms, acc[i].n, acc[i].what);
      fclose(f);
    }
  }
  for (const float* src : e->packed_srcs) unregister_packed(src);
  if (e->have_streams) {
    cudaStreamSynchronize(e->mainp);
    cudaStreamDestroy(e->mainp);
    cudaEventDestroy(e->ev_in); cudaEventDestroy(e->ev_out); cudaEventDestroy(e->ev_sel);
    cudaFreeHost(e->h_over);
  }
  if (e->have_streams)
    for (int l = 0; l < e->dm.L; ++l) {
      cudaStreamSynchronize(e->side[l]);
      cudaStreamDestroy(e->side[l]);
      cudaEventDestroy(e->ev_tok[l]); cudaEventDestroy(e->ev_z[l]); cudaEventDestroy(e->ev_head[l]); cudaEventDestroy(e->ev_dtok[l]);
      cudaEventDestroy(e->ev_grad[l]);
    }
  delete e;
}

int paths_engine_refresh_weights(paths_engine_t* e, void* stream) {
  PATHS_REQUIRE(e && e->have_params, "refresh_weights: engine has no parameters bound");
  if (int r = ensure_registry(e, static_cast<cudaStream_t>(stream))) return r;
  if (e->dm.precision == PREC_FP32_SIMT) return 0;
  if (int r = run_pack_jobs(e->pack_jobs, e->n_pack_jobs, static_cast<cudaStream_t>(stream))) return r;
  e->launches++;
  return 0;
}

// (Re-)registers every contracted weight of the engine with the tensor-core path: the registry then resolves those weight
// pointers to THIS engine's packed arena.
static int register_weights(paths_engine_t* e) {
  {
    for (const float* src : e->packed_srcs) unregister_packed(src);
    e->packed_srcs.clear();
    const paths_dims_t& dm = e->dm;
    const int D = dm.D, Cm = dm.C, d = dm.d, I = dm.I, F = dm.F;
    char* cur = e->packed_arena;
    auto reg = [&](const float* src, int R, int Cc, int ld) -> int {
      if (int r = register_packed(src, R, Cc, ld, cur, 1)) return r;
      cur += packed_bytes(R, Cc);
      e->packed_srcs.push_back(src);
      return 0;
    };
    // the same weight of every level, packed back to back (member l at 2R plane rows from member l-1): the head's tensor map
    // spans the stack, so a grouped contraction reaches level l's weight by a row offset
    auto reg_stack = [&](const float* const* srcs, int n, int R, int Cc, int ld) -> int {
      const size_t member = (size_t)2 * R * ((Cc + 7) / 8 * 8) * sizeof(__nv_bfloat16);
      char* base_al = reinterpret_cast<char*>((reinterpret_cast<uintptr_t>(cur) + 255) & ~uintptr_t(255));
      for (int l = 0; l < n; ++l) {
        PATHS_REQUIRE(srcs[l], "bind_params: decoder weight of level %d missing", l);
        if (int r = register_packed(srcs[l], R, Cc, ld, base_al + (size_t)l * member, l == 0 ? n : 1)) return r;
        e->packed_srcs.push_back(srcs[l]);
      }
      cur = base_al + (size_t)n * member;
      return 0;
    };
    e->stacked = d % 8 == 0 && F % 8 == 0;                 // member stride stays 256-byte aligned, every extent TMA-legal
    int rc = 0;
    if (dm.lstm) {
      PATHS_REQUIRE(e->P.Wg && e->P.Wmo, "bind_params: LSTM weights missing");
      rc = rc ? rc : reg(e->P.Wg, 3 * Cm + D, 2 * D, 2 * D);
      rc = rc ? rc : reg(e->P.Wmo, D, Cm, Cm);
    }
    for (int l = 0; l < dm.L && !rc; ++l) {
      rc = rc ? rc : reg(e->P.level[l].Wuv, I + d, D, D);
      if (!dm.lstm && l > 0 && dm.hier_ctx) {
        PATHS_REQUIRE(e->P.level[l].Wh0 && e->P.level[l].Wh2, "bind_params: hctx_mlp weights of level %d missing", l);
        rc = rc ? rc : reg(e->P.level[l].Wh0, Cm, D, D);
        rc = rc ? rc : reg(e->P.level[l].Wh2, D, Cm, Cm);
      }
      for (int j = 0; j < dm.layers && !rc && !e->stacked; ++j) {
        const LayerP& y = e->P.level[l].layer[j];
        rc = rc ? rc : reg(y.Wqkv, 3 * d, d, d);
        rc = rc ? rc : reg(y.Wo, d, d, d);
        rc = rc ? rc : reg(y.W1, F, d, d);
        rc = rc ? rc : reg(y.W2, d, F, F);
      }
    }
    for (int j = 0; j < dm.layers && !rc && e->stacked; ++j) {
      const float* w[4][PATHS_MAX_LEVELS];
      for (int l = 0; l < dm.L; ++l) {
        const LayerP& y = e->P.level[l].layer[j];
        w[0][l] = y.Wqkv; w[1][l] = y.Wo; w[2][l] = y.W1; w[3][l] = y.W2;
      }
      rc = rc ? rc : reg_stack(w[0], dm.L, 3 * d, d, d);
      rc = rc ? rc : reg_stack(w[1], dm.L, d, d, d);
      rc = rc ? rc : reg_stack(w[2], dm.L, F, d, d);
      rc = rc ? rc : reg_stack(w[3], dm.L, d, F, F);
    }
    if (rc) return rc;
    PATHS_REQUIRE((size_t)(cur - e->packed_arena) <= e->packed_arena_bytes, "packed-weight arena overrun");
  }
  return 0;
}

// Several engines may be bound to the same parameters (the step engine, per-level engines of the level API, engines of other
// batch sizes); the weight registry is keyed by the weight pointer, so the engine bound last owns it.  Every entry point that
// launches contractions first makes sure the registry points into ITS arena -- re-registering and re-packing if another
// engine took over (or was destroyed) in the meantime -- so an engine never contracts against another engine's stale copy.
static int ensure_registry(paths_engine_t* e, cudaStream_t st) {
  if (!e->have_params || e->packed_srcs.empty()) return 0;
  const char* dst = static_cast<const char*>(packed_dst_of(e->packed_srcs[0]));
  if (dst && dst >= e->packed_arena && dst < e->packed_arena + e->packed_arena_bytes) return 0;
  if (int r = register_weights(e)) return r;
  if (e->dm.precision != PREC_FP32_SIMT) {
    if (int r = run_pack_jobs(e->pack_jobs, e->n_pack_jobs, st)) return r;
    e->launches++;
  }
  return 0;
}

int paths_engine_bind_params(paths_engine_t* e, const float* params, float* grads, int64_t n_floats, const int64_t* h_offsets,
                             const float* div) {
  PATHS_REQUIRE(e && params && h_offsets && div, "null argument");
  bind(e->P, params, h_offsets, e->dm);
  if (grads) bind(e->Gr, grads, h_offsets, e->dm);
  e->grads_base = grads;
  e->grads_floats = (size_t)n_floats;
  if (int r = register_weights(e)) return r;
  if (int r = build_pack_jobs(e->packed_srcs.data(), (int)e->packed_srcs.size(), e->pack_jobs, &e->n_pack_jobs, 0)) return r;
  e->div = div;
  e->have_params = true;
  return 0;
}

int paths_engine_bind_store(paths_engine_t* e, const paths_store_t* h_store) {
  PATHS_REQUIRE(e && h_store, "null argument");
  e->store = *h_store;
  e->have_store = true;
  return 0;
}

int paths_engine_begin(paths_engine_t* e, const int32_t* slide_ids, void* stream) {
  PATHS_REQUIRE(e, "null engine");
  if (int r = ensure_registry(e, static_cast<cudaStream_t>(stream))) return r;
  return do_begin(e, slide_ids, nullptr, 0, static_cast<cudaStream_t>(stream));
}

int paths_engine_begin_padded(paths_engine_t* e, const int32_t* slide_ids, const float* fts_padded, int N, void* stream) {
  PATHS_REQUIRE(e && fts_padded, "null argument");
  if (int r = ensure_registry(e, static_cast<cudaStream_t>(stream))) return r;
  return do_begin(e, slide_ids, fts_padded, N, static_cast<cudaStream_t>(stream));
}

int paths_engine_level_forward(paths_engine_t* e, int level, void* stream) {
  PATHS_REQUIRE(e && e->have_params, "level_forward: engine has no parameters bound");
  PATHS_REQUIRE(level >= 0 && level < e->dm.L, "level %d out of range", level);
  if (int r = ensure_registry(e, static_cast<cudaStream_t>(stream))) return r;
  return level_forward(e, level, static_cast<cudaStream_t>(stream));
}

int paths_engine_select(paths_engine_t* e, int level, void* stream) {
  PATHS_REQUIRE(e, "null engine");
  if (int r = do_select(e, level, static_cast<cudaStream_t>(stream))) return r;
  return publish_overflow(e, static_cast<cudaStream_t>(stream));
}

int paths_engine_overflow_poll(paths_engine_t* e, int32_t* h_flag, int32_t* h_rows_needed) {
  PATHS_REQUIRE(e && h_flag && h_rows_needed, "null argument");
  *h_flag = 0; *h_rows_needed = 0;
  if (!e->sel_recorded) return 0;
  PATHS_CK(cudaEventSynchronize(e->ev_sel));
  *h_flag = e->h_over[0]; *h_rows_needed = e->h_over[1];
  return 0;
}

int paths_engine_loss(paths_engine_t* e, const int32_t* ybin, const int32_t* censored, float scale, void* stream) {
  PATHS_REQUIRE(e && ybin, "null argument");
  return do_loss(e, ybin, censored, scale, static_cast<cudaStream_t>(stream));
}

int paths_engine_level_backward(paths_engine_t* e, int level, int use_dlogits, void* stream) {
  PATHS_REQUIRE(e && e->have_params && e->grads_base, "level_backward: no gradient buffer bound");
  PATHS_REQUIRE(level >= 0 && level < e->dm.L, "level %d out of range", level);
  if (int r = ensure_registry(e, static_cast<cudaStream_t>(stream))) return r;
  return level_backward(e, level, use_dlogits, static_cast<cudaStream_t>(stream));
}

// The whole-step entry points run each level's aggregator on its own stream: select(l) needs only the importances, so the
// LSTM / select chain of the deeper levels (main stream) overlaps the decoders of the shallower ones; the slide-vector chain
// (heads) joins them at the end.  Backward mirrors it: head chain first, the L decoder backwards concurrently, then the LSTM
// chain from the deepest level up, each level waiting only for its own token gradient.
static bool use_side_streams(const paths_engine* e) {
  static const bool off = getenv("PATHS_B200_NO_STREAMS") != nullptr;
  return !off && e->have_streams && planes_mode(e->dm) && e->dm.d % 8 == 0 && e->dm.F % 8 == 0;
}

static int forward_impl(paths_engine_t* e, const int32_t* slide_ids, const float* fts_padded, int N, const int32_t* ybin,
                        const int32_t* censored, float scale, int with_loss, void* stream);

int paths_engine_forward(paths_engine_t* e, const int32_t* slide_ids, const int32_t* ybin, const int32_t* censored, float scale,
                         int with_loss, void* stream) {
  return forward_impl(e, slide_ids, nullptr, 0, ybin, censored, scale, with_loss, stream);
}

int paths_engine_forward_padded(paths_engine_t* e, const int32_t* slide_ids, const float* fts_padded, int N, const int32_t* ybin,
                                const int32_t* censored, float scale, int with_loss, void* stream) {
  PATHS_REQUIRE(fts_padded && N >= 1, "forward_padded: null feature batch");
  return forward_impl(e, slide_ids, fts_padded, N, ybin, censored, scale, with_loss, stream);
}

int paths_engine_forward_packed(paths_engine_t* e, const int32_t* slide_ids, const float* fts_packed, int n_rows, const int32_t* ybin,
                                const int32_t* censored, float scale, int with_loss, void* stream) {
  PATHS_REQUIRE(fts_packed && n_rows >= 1, "forward_packed: null feature rows");
  return forward_impl(e, slide_ids, fts_packed, -n_rows, ybin, censored, scale, with_loss, stream);
}

static int forward_impl(paths_engine_t* e, const int32_t* slide_ids, const float* fts_padded, int N, const int32_t* ybin,
                        const int32_t* censored, float scale, int with_loss, void* stream) {
  PATHS_REQUIRE(e && e->have_params, "forward: engine has no parameters bound");
  cudaStream_t caller = static_cast<cudaStream_t>(stream);
  if (int r = ensure_registry(e, caller)) return r;
  cudaStream_t st = caller;
  const int L = e->dm.L;
  const bool conc = use_side_streams(e);
  if (conc) {                                            // fork: the step runs on the engine's own streams
    PATHS_CK(cudaEventRecord(e->ev_in, caller));
    PATHS_CK(cudaStreamWaitEvent(e->mainp, e->ev_in, 0));
    st = e->mainp;
  }
  if (int r = do_begin(e, slide_ids, fts_padded, N, st)) return r;
  if (L == 1)
    if (int r = publish_overflow(e, st)) return r;
  if (!conc) {
    for (int l = 0; l < L; ++l) {
      if (int r = level_forward(e, l, st)) return r;
      if (l < L - 1)
        if (int r = do_select(e, l, st)) return r;
      if (l == L - 2)
        if (int r = publish_overflow(e, st)) return r;
    }
  } else {
    for (int l = 0; l < L; ++l) {
      if (int r = lstm_forward(e, l, st)) return r;
      PATHS_CK(cudaEventRecord(e->ev_tok[l], st));
      for (int si = 0; si < e->n_fsets; ++si) {            // a set starts when the tokens of its deepest level exist
        int deepest = 0;
        for (int k = 0; k < e->fset_n[si]; ++k) deepest = e->fset_lv[si][k] > deepest ? e->fset_lv[si][k] : deepest;
        if (deepest != l) continue;
        PATHS_CK(cudaStreamWaitEvent(e->side[si], e->ev_tok[l], 0));
        if (int r = decoder_forward(e, make_set(e, e->fset_lv[si], e->fset_n[si]), e->side[si])) return r;
        PATHS_CK(cudaEventRecord(e->ev_z[si], e->side[si]));
      }
      if (l < L - 1)
        if (int r = do_select(e, l, st)) return r;
      if (l == L - 2)
        if (int r = publish_overflow(e, st)) return r;
    }
    for (int si = 0; si < e->n_fsets; ++si) PATHS_CK(cudaStreamWaitEvent(st, e->ev_z[si], 0));
    for (int l = 0; l < L; ++l)
      if (int r = head_forward(e, l, st)) return r;
  }
  if (with_loss)
    if (int r = do_loss(e, ybin, censored, scale, st)) return r;
  if (conc) {                                            // join
    PATHS_CK(cudaEventRecord(e->ev_out, st));
    PATHS_CK(cudaStreamWaitEvent(caller, e->ev_out, 0));
  }
  return 0;
}

int paths_engine_backward(paths_engine_t* e, int zero_grads, void* stream) {
  PATHS_REQUIRE(e && e->have_params && e->grads_base, "backward: no gradient buffer bound");
  cudaStream_t caller = static_cast<cudaStream_t>(stream);
  if (int r = ensure_registry(e, caller)) return r;
  const int L = e->dm.L;
  if (!use_side_streams(e)) {
    if (zero_grads) PATHS_CK(cudaMemsetAsync(e->grads_base, 0, e->grads_floats * sizeof(float), caller));
    for (int l = L - 1; l >= 0; --l) {
      if (int r = level_backward(e, l, l == L - 1, caller)) return r;
      if (e->have_streams) PATHS_CK(cudaEventRecord(e->ev_grad[l], caller));
    }
    return 0;
  }
  PATHS_CK(cudaEventRecord(e->ev_in, caller));
  PATHS_CK(cudaStreamWaitEvent(e->mainp, e->ev_in, 0));
  cudaStream_t st = e->mainp;
  if (zero_grads) PATHS_CK(cudaMemsetAsync(e->grads_base, 0, e->grads_floats * sizeof(float), st));
  for (int l = L - 1; l >= 0; --l)
    if (int r = head_backward(e, l, l == L - 1, st)) return r;
  PATHS_CK(cudaEventRecord(e->ev_head[0], st));
  int set_of[PATHS_MAX_LEVELS] = {0};
  for (int si = e->n_bsets - 1; si >= 0; --si) {           // sets holding the deeper levels first: the LSTM chain needs them first
    PATHS_CK(cudaStreamWaitEvent(e->side[si], e->ev_head[0], 0));
    if (int r = decoder_backward(e, make_set(e, e->bset_lv[si], e->bset_n[si]), si, e->side[si])) return r;
    PATHS_CK(cudaEventRecord(e->ev_dtok[si], e->side[si]));
    for (int k = 0; k < e->bset_n[si]; ++k) set_of[e->bset_lv[si][k]] = si;
  }
  for (int l = L - 1; l >= 0; --l) {
    PATHS_CK(cudaStreamWaitEvent(st, e->ev_dtok[set_of[l]], 0));
    if (int r = lstm_backward(e, l, st)) return r;
    PATHS_CK(cudaEventRecord(e->ev_grad[l], st));
  }
  PATHS_CK(cudaEventRecord(e->ev_out, st));
  PATHS_CK(cudaStreamWaitEvent(caller, e->ev_out, 0));
  return 0;
}

int paths_engine_wait_grads(paths_engine_t* e, int level, void* stream) {
  PATHS_REQUIRE(e && e->have_streams && level >= 0 && level < e->dm.L, "wait_grads: bad argument");
  PATHS_CK(cudaStreamWaitEvent(static_cast<cudaStream_t>(stream), e->ev_grad[level], 0));
  return 0;
}

int paths_engine_buffer(paths_engine_t* e, int level, int which, void** ptr, size_t* bytes) {
  PATHS_REQUIRE(e && ptr && bytes, "null argument");
  PATHS_REQUIRE(level >= 0 && level < e->dm.L, "level %d out of range", level);
  const paths_dims_t& dm = e->dm;
  Level& v = e->lv[level];
  const size_t cap = v.cap, B = e->B, f = sizeof(float), i4 = sizeof(int);
  void* p = nullptr; size_t n = 0;
  switch (which) {
    case PATHS_BUF_COUNT: p = v.count; n = i4; break;
    case PATHS_BUF_NTOK: p = v.ntok; n = i4; break;
    case PATHS_BUF_SEG: p = v.seg; n = (B + 1) * i4; break;
    case PATHS_BUF_TSEG: p = v.tseg; n = B * i4; break;
    case PATHS_BUF_ROW_SLIDE: p = v.row_slide; n = cap * i4; break;
    case PATHS_BUF_ROW_XY: p = v.row_xy; n = 2 * cap * i4; break;
    case PATHS_BUF_ROW_PARENT: p = v.row_parent; n = cap * i4; v.tables_ok = false; break;
    case PATHS_BUF_PARENT_LOCAL: p = v.parent_local; n = cap * i4; break;
    case PATHS_BUF_SEL: p = v.sel; n = B * (size_t)(v.kcap > 0 ? v.kcap : 1) * i4; break;
    case PATHS_BUF_SEL_COUNT: p = v.sel_count; n = B * i4; break;
    case PATHS_BUF_CHILD_OF: p = v.child_of; n = 4 * cap * i4; break;
    case PATHS_BUF_X: p = v.X; n = cap * dm.D * f; v.x_planes = false; break;
    case PATHS_BUF_G: p = v.G; n = cap * (3 * (size_t)dm.C + dm.D) * f; break;
    case PATHS_BUF_CN: p = v.Cn; n = cap * dm.C * f; break;
    case PATHS_BUF_PT: p = v.Pt; n = cap * dm.D * f; break;
    case PATHS_BUF_H: p = v.H; n = cap * dm.D * f; break;
    case PATHS_BUF_Y: p = v.Y; n = cap * dm.D * f; break;
    case PATHS_BUF_UV: p = v.UV; n = cap * ((size_t)dm.I + dm.d) * f; break;
    case PATHS_BUF_ALPHA: p = v.alpha; n = cap * f; break;
    case PATHS_BUF_TOK: p = v.tok; n = (cap + B) * dm.d * f; break;
    case PATHS_BUF_Z: p = v.z; n = B * dm.d * f; break;
    case PATHS_BUF_F: p = v.F; n = B * dm.d * f; break;
    case PATHS_BUF_LOGITS: p = v.logits; n = B * dm.nl * f; break;
    case PATHS_BUF_DLOGITS: p = v.dlogits; n = B * dm.nl * f; break;
    case PATHS_BUF_HAZARDS: p = e->hazards; n = B * dm.nl * f; break;
    case PATHS_BUF_LOSS: p = e->loss; n = f; break;
    case PATHS_BUF_LOSS_PER: p = e->loss_per; n = B * f; break;
    case PATHS_BUF_RISK: p = e->risk; n = B * f; break;
    case PATHS_BUF_OVERFLOW: p = e->overflow; n = 2 * i4; break;
    case PATHS_BUF_DF_IN: p = v.dFl; n = B * dm.d * f; break;
    case PATHS_BUF_DHROW: p = v.dHrow; n = cap * dm.D * f; break;
    case PATHS_BUF_DCROW: p = v.dCrow; n = cap * dm.C * f; break;
    case PATHS_BUF_HF0: p = v.layer[0].Hf; n = (dm.layers > 1 ? (size_t)v.ncap_tok : (size_t)B) * dm.F * f; break;
    default: set_error("unknown buffer id %d", which); return -2;
  }
  *ptr = p; *bytes = n;
  return 0;
}

int paths_engine_set_dropout(paths_engine_t* e, float p, uint64_t seed) {
  PATHS_REQUIRE(e && p >= 0.f && p < 1.f, "dropout probability must lie in [0, 1)");
  e->drop_p = p; e->drop_seed = seed;
  return 0;
}

int paths_engine_set_precision(paths_engine_t* e, int precision) {
  PATHS_REQUIRE(e && precision >= 0 && precision <= 2, "precision must be 0, 1 or 2");
  e->dm.precision = precision;
  return 0;
}

int64_t paths_engine_launch_count(paths_engine_t* e) { return e ? e->launches : 0; }

int paths_engine_profile(paths_engine_t* e, int on) {
  PATHS_REQUIRE(e, "null engine");
  e->prof_on = on != 0;
  return 0;
}

int paths_engine_profile_read(paths_engine_t* e, int tag, double* total_ms, int64_t* launches) {
  PATHS_REQUIRE(e && total_ms && launches && tag >= 0 && tag < PATHS_PROF__COUNT, "bad argument");
  std::vector<cudaEvent_t>& v = e->prof_ev[tag];
  double tot = 0.0;
  for (size_t i = 0; i + 1 < v.size(); i += 2) {
    float ms = 0.f;
    PATHS_CK(cudaEventSynchronize(v[i + 1]));
    PATHS_CK(cudaEventElapsedTime(&ms, v[i], v[i + 1]));
    tot += ms;
  }
  *total_ms = tot;
  *launches = (int64_t)(v.size() / 2);
  for (cudaEvent_t ev : v) cudaEventDestroy(ev);
  v.clear();
  return 0;
}

int paths_gemm_nt(int precision, const float* A, int lda, const float* B, int ldb, int b_kn, float* C, int ldc, const float* bias,
                  int M, int N, int K, int flags, void* stream) {
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  void *tmp = nullptr, *jobs = nullptr, *scr = nullptr;
  if (precision != PREC_FP32_SIMT) {          // stand-alone use: temporary packed weight + activation scratch
    const int R = b_kn ? K : N, Cc = b_kn ? N : K;
    PATHS_CK(cudaMalloc(&tmp, packed_bytes(R, Cc)));
    PATHS_CK(cudaMalloc(&jobs, 64));
    PATHS_CK(cudaMalloc(&scr, tc_scratch_bytes(M, K)));
    set_tc_scratch(scr, tc_scratch_bytes(M, K), nullptr, 0);
    if (int r = register_packed(B, R, Cc, ldb, tmp, 1)) return r;
    int nj = 0;
    if (int r = build_pack_jobs(&B, 1, jobs, &nj, st)) return r;
    if (int r = run_pack_jobs(jobs, nj, st)) return r;
  }
  int rc = gemm_nt(precision, row_src(A, lda, K), B, ldb, b_kn, C, ldc, bias, M, nullptr, N, K, flags, st);
  if (tmp) {
    cudaStreamSynchronize(st);
    unregister_packed(B);
    cudaFree(tmp); cudaFree(jobs); cudaFree(scr);
    set_tc_scratch(nullptr, 0, nullptr, 0);
  }
  return rc;
}

// Stand-alone varlen self-attention (tests, per-op use): the engine's decoder calls the same launchers.
int paths_attention_fwd(int precision, const float* qkv, const int32_t* seg, int B, float* out, float* lse, int d, int heads, int max_bag,
                        float dropout_p, uint64_t dropout_seed, void* stream) {
  PATHS_REQUIRE(qkv && seg && out && lse && B >= 1 && heads >= 1 && d % heads == 0 && precision >= 0 && precision <= 2, "attention_fwd: bad argument");
  Drop dr = no_drop();
  if (dropout_p > 0.f) { dr.thresh = drop_thresh(dropout_p); dr.inv_keep = 1.f / (1.f - dropout_p); dr.key[0] = drop_site_key(dropout_seed, 0, 0, DROP_ATTN); }
  return attn_fwd_grp(qkv, seg_single(seg, B), out, lse, d, heads, max_bag + 1, 0, static_cast<cudaStream_t>(stream), dropout_p > 0.f ? &dr : nullptr,
                      precision == PREC_BF16X3_TC ? 3 : precision == PREC_BF16_TC ? 1 : 0);
}
int paths_attention_bwd(int precision, const float* qkv, const int32_t* seg, int B, const float* out, const float* lse, const float* dout,
                        float* dqkv, float* dvec, int d, int heads, int max_bag, float dropout_p, uint64_t dropout_seed, void* stream) {
  PATHS_REQUIRE(qkv && seg && out && lse && dout && dqkv && dvec && B >= 1 && heads >= 1 && d % heads == 0 && precision >= 0 && precision <= 2,
                "attention_bwd: bad argument");
  Drop dr = no_drop();
  if (dropout_p > 0.f) { dr.thresh = drop_thresh(dropout_p); dr.inv_keep = 1.f / (1.f - dropout_p); dr.key[0] = drop_site_key(dropout_seed, 0, 0, DROP_ATTN); }
  return attn_bwd_grp(qkv, seg_single(seg, B), out, lse, dout, dqkv, dvec, d, heads, max_bag + 1, 0, static_cast<cudaStream_t>(stream),
                      dropout_p > 0.f ? &dr : nullptr, precision == PREC_BF16X3_TC ? 3 : precision == PREC_BF16_TC ? 1 : 0);
}


// ---- allocation-free variant of paths_attention_bwd ------------------------------------------------------------------------------
// For callers that cannot allocate device memory (CUDA-graph capture, serving loops, harnesses that forbid cudaMalloc): the
// [tokens, heads] scratch of the backward pass (D = dO . O per query row and head, written by the dq kernel and read by the dk/dv
// kernel) lives in a file-scope device array instead of a caller buffer.  Every GPU that loads the library holds the array until
// the process exits.  In host code the array's NAME is only a host-side stand-in: its device address comes from
// cudaGetSymbolAddress, looked up once per device with that device current.  All calls share the array (it keeps the last call's
// contents; every element a call reads is written by that call first, so nothing has to be cleared): calls that overlap on the GPU
// -- different streams, or different host threads -- race on it.
static __device__ __align__(256) float g_attn_bwd_scratch[PATHS_ATTENTION_BWD_STATIC_MAX_FLOATS];
namespace paths { namespace {
constexpr int kMaxDevices = 64;
float* g_attn_bwd_scratch_ptr[kMaxDevices] = {};
std::mutex g_attn_bwd_scratch_mu;
// device address of g_attn_bwd_scratch on the current device (no allocation: the array is part of the loaded module)
int attn_bwd_scratch(float** out) {
  int dev = -1;
  PATHS_CK(cudaGetDevice(&dev));
  PATHS_REQUIRE(dev >= 0 && dev < kMaxDevices, "attention_bwd_static: device ordinal %d outside [0, %d)", dev, kMaxDevices);
  std::lock_guard<std::mutex> lk(g_attn_bwd_scratch_mu);
  if (!g_attn_bwd_scratch_ptr[dev]) {
    float* ptr = nullptr;
    PATHS_CK(cudaGetSymbolAddress(reinterpret_cast<void**>(&ptr), g_attn_bwd_scratch));
    PATHS_REQUIRE(ptr != nullptr, "attention_bwd_static: cudaGetSymbolAddress returned a null device address");
    g_attn_bwd_scratch_ptr[dev] = ptr;
  }
  *out = g_attn_bwd_scratch_ptr[dev];
  return 0;
}
}}  // namespace paths::(anonymous)

int paths_attention_bwd_static(int precision, const float* qkv, const int32_t* seg, int B, int tokens, const float* out, const float* lse,
                               const float* dout, float* dqkv, int d, int heads, int max_bag, float dropout_p, uint64_t dropout_seed,
                               void* stream) {
  PATHS_REQUIRE(qkv && seg && out && lse && dout && dqkv && B >= 1 && tokens >= B && heads >= 1 && d % heads == 0 && precision >= 0 &&
                precision <= 2, "attention_bwd_static: bad argument");
  PATHS_REQUIRE((long long)tokens * heads <= (long long)PATHS_ATTENTION_BWD_STATIC_MAX_FLOATS,
                "attention_bwd_static: tokens * heads = %lld exceeds the static scratch of %d floats (PATHS_ATTENTION_BWD_STATIC_MAX_FLOATS); "
                "use paths_attention_bwd with a caller-provided dvec", (long long)tokens * heads, PATHS_ATTENTION_BWD_STATIC_MAX_FLOATS);
  float* dvec = nullptr;
  if (int r = paths::attn_bwd_scratch(&dvec)) return r;
  Drop dr = no_drop();
  if (dropout_p > 0.f) { dr.thresh = drop_thresh(dropout_p); dr.inv_keep = 1.f / (1.f - dropout_p); dr.key[0] = drop_site_key(dropout_seed, 0, 0, DROP_ATTN); }
  return attn_bwd_grp(qkv, seg_single(seg, B), out, lse, dout, dqkv, dvec, d, heads, max_bag + 1, 0, static_cast<cudaStream_t>(stream),
                      dropout_p > 0.f ? &dr : nullptr, precision == PREC_BF16X3_TC ? 3 : precision == PREC_BF16_TC ? 1 : 0);
}


// ---- per-op entries: the single operations behind the reference's per-module methods -------------------------
namespace paths { namespace {
__global__ void iota_kernel(int* __restrict__ p, int n) {
  pdl_grid_sync();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) p[i] = i;
}
} }

// LSTMCell.forward (model/interface.py:31-58) on M rows: x, h [M, D], c [M, C] -> h_out [M, D], c_out [M, C].  Wg = the four gate
// weights stacked [forget; remember; remember_map; out_select] ([3C + D, 2D]), bg likewise; Wmo [D, C].  Temporaries are allocated
// inside (a stand-alone entry: the engine's own path keeps everything in its workspace).
int paths_lstm_cell_fwd(int precision, const float* x, const float* h, const float* c, const float* Wg, const float* bg, const float* Wmo,
                        const float* bmo, float* h_out, float* c_out, int M, int D, int C, void* stream) {
  PATHS_REQUIRE(x && h && c && Wg && bg && Wmo && bmo && h_out && c_out && M >= 1 && D % 4 == 0 && C % 4 == 0, "lstm_cell_fwd: bad argument");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int ldg = 3 * C + D;
  float *xh = nullptr, *G = nullptr, *Pt = nullptr; int* ident = nullptr;
  PATHS_CK(cudaMalloc(&xh, (size_t)M * 2 * D * sizeof(float)));
  PATHS_CK(cudaMalloc(&G, (size_t)M * ldg * sizeof(float)));
  PATHS_CK(cudaMalloc(&Pt, (size_t)M * D * sizeof(float)));
  PATHS_CK(cudaMalloc(&ident, (size_t)M * sizeof(int)));
  PATHS_CK(cudaMemcpy2DAsync(xh, 2 * D * sizeof(float), x, D * sizeof(float), D * sizeof(float), M, cudaMemcpyDeviceToDevice, st));
  PATHS_CK(cudaMemcpy2DAsync(xh + D, 2 * D * sizeof(float), h, D * sizeof(float), D * sizeof(float), M, cudaMemcpyDeviceToDevice, st));
  pdl(iota_kernel, (M + 255) / 256, 256, 0, st)(ident, M);
  PATHS_LAUNCH_CK();
  int rc = paths_gemm_nt(precision, xh, 2 * D, Wg, 2 * D, 0, G, ldg, bg, M, ldg, 2 * D, 0, stream);          // [f | r | m | o] pre-activations
  if (!rc) rc = lstm_point1_fwd(G, c, ident, c_out, no_planes(), 0, 0, D, C, M, nullptr, st);                // activations in place, c'
  if (!rc) rc = paths_gemm_nt(precision, c_out, C, Wmo, C, 0, Pt, D, bmo, M, D, C, 0, stream);
  if (!rc) rc = lstm_point2_fwd(Pt, G, x, h_out, nullptr, no_planes(), 0, D, C, M, nullptr, st);             // h' = o * tanh(P)
  cudaStreamSynchronize(st);
  cudaFree(xh); cudaFree(G); cudaFree(Pt); cudaFree(ident);
  return rc;
}

// PreprocessedSlide.iter (data_utils/slide.py:277-360) for a batch: per-slide top-K by importance, 4-child expansion, bounds +
// occupancy filter, order-preserving compaction, no-child fallback.  Integer in, integer out; every array is caller-owned.
//   level l:   alpha [rows], seg [B + 1], row_xy [rows, 2] (patch coordinates), count [1] = seg[B]
//   outputs:   sel [B, kcap] kept rows (importance-descending, -1 padded), sel_count [B], child_of [rows, 4];
//              level l + 1 tables n_count [1], n_ntok [1], n_seg [B + 1], n_row_slide / n_row_parent / n_parent_local [n_cap],
//              n_row_xy [n_cap, 2]; overflow [2] (flag, rows needed) is raised when n_cap or n_maxseg is exceeded
//   scratch:   2 * B ints
int paths_select_expand(const float* alpha, const int32_t* seg, const int32_t* row_xy, const int32_t* count, int B, int cap, int keep, int kcap,
                        int maxseg, const paths_store_t* h_store, int level, int L, const int32_t* slide_ids, int32_t* sel, int32_t* sel_count,
                        int32_t* child_of, int32_t* n_count, int32_t* n_ntok, int32_t* n_seg, int32_t* n_row_slide, int32_t* n_row_xy,
                        int32_t* n_row_parent, int32_t* n_parent_local, int n_cap, int n_maxseg, int32_t* overflow, int32_t* scratch,
                        void* stream) {
  PATHS_REQUIRE(alpha && seg && row_xy && count && h_store && sel && sel_count && child_of && n_count && n_ntok && n_seg && n_row_slide &&
                n_row_xy && n_row_parent && n_parent_local && overflow && scratch && B >= 1 && level >= 0 && level < L - 1,
                "select_expand: bad argument");
  PATHS_REQUIRE(keep < 0 || maxseg <= 65536, "select_expand: bags above 65536 rows need keep = -1");
  SelectArgs a;
  a.alpha = alpha; a.seg = seg; a.row_xy = row_xy; a.count = count; a.B = B; a.cap = cap; a.keep = keep; a.kcap = kcap; a.maxseg = maxseg;
  a.n_maxseg = n_maxseg;
  a.gw = h_store->gw; a.gh = h_store->gh; a.occ_off = reinterpret_cast<const unsigned long long*>(h_store->occ_off); a.occ = h_store->occ;
  a.nonbg = h_store->nonbg; a.level = level; a.L = L; a.slide_ids = slide_ids;
  a.sel = sel; a.sel_count = sel_count; a.child_of = child_of;
  a.n_count = n_count; a.n_seg = n_seg; a.n_row_slide = n_row_slide; a.n_row_xy = n_row_xy; a.n_row_parent = n_row_parent;
  a.n_parent_local = n_parent_local; a.n_cap = n_cap; a.n_ntok = n_ntok; a.overflow = overflow; a.slide_cnt = scratch; a.slide_fb = scratch + B;
  return select_expand(a, static_cast<cudaStream_t>(stream));
}

// X[m, :] = grid(slide_ids[row_slide[m]], level)[row_xy[m]] for m < M: the feature gather of slide.py:331 from the resident store.
int paths_gather_rows(const paths_store_t* h_store, int level, int L, const int32_t* slide_ids, const int32_t* row_slide, const int32_t* row_xy,
                      float* X, int D, int M, const int32_t* M_dev, void* stream) {
  PATHS_REQUIRE(h_store && row_slide && row_xy && X && level >= 0 && level < L, "gather_rows: bad argument");
  return gather_rows(h_store->grid_ptr, h_store->gh, level, L, slide_ids, row_slide, row_xy, X, D, M, M_dev, static_cast<cudaStream_t>(stream), h_store->bf16);
}

// TransformerAggregator.forward (model/aggregator.py:58-76) of one level on tokens the caller has placed in the level's
// PATHS_BUF_TOK rows (token of patch row m of slide s at row m + s + 1; the special tokens are filled here): decoder layers over
// the packed bags, result in PATHS_BUF_Z ([B, d]).  Row tables (COUNT, NTOK, SEG, TSEG) must describe the bags.
int paths_engine_aggregate(paths_engine_t* e, int level, void* stream) {
  PATHS_REQUIRE(e && e->have_params && level >= 0 && level < e->dm.L, "aggregate: bad argument");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (int r = ensure_registry(e, st)) return r;
  Level& v = e->lv[level];
  RUN(special_fill(e->P.level[level].special, v.seg, v.tok, e->dm.d, e->B, st));
  return decoder_forward(e, make_set(e, &level, 1), st);
}

int paths_gemm_tn(int precision, const float* P, int ldp, const float* Q, int ldq, float* C, int ldc, int M, int NI, int NJ, void* stream) {
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  void *sa = nullptr, *sb = nullptr, *sw = nullptr;
  if (precision != PREC_FP32_SIMT) {
    PATHS_CK(cudaMalloc(&sa, tc_scratch_bytes(M, NI)));
    PATHS_CK(cudaMalloc(&sb, tc_scratch_bytes(M, NJ)));
    PATHS_CK(cudaMalloc(&sw, tn_workspace_bytes(NI, NJ, M, 1)));
    set_tc_scratch(sa, tc_scratch_bytes(M, NI), sb, tc_scratch_bytes(M, NJ));
    set_tn_workspace(sw, tn_workspace_bytes(NI, NJ, M, 1));
  }
  int rc = gemm_tn(precision, P, ldp, row_src(Q, ldq, NJ), C, ldc, M, nullptr, NI, NJ, st);
  if (sa) {
    cudaStreamSynchronize(st);
    cudaFree(sa); cudaFree(sb); cudaFree(sw);
    set_tc_scratch(nullptr, 0, nullptr, 0);
    set_tn_workspace(nullptr, 0);
  }
  return rc;
}

}  // extern "C"
