/*
 * paths_b200.h -- C ABI of libpaths_b200.so: the B200-native (sm_100a) implementation of the PATHS hierarchical
 * slide-level forward/backward pass.
 *
 * The reference (zzbuzzard/PATHS) is pure Python/PyTorch and has no plugin / FFI interface; its
 * boundary is its Python module API.  This header is the boundary *below* that API: the functions the Python mirror
 * (paths_b200/utils.py:inference_end2end, paths_b200/model/*) binds through ctypes.  Each entry cites the reference
 * code it replaces (paths relative to the reference checkout).
 *
 * Conventions
 *   - extern "C", plain C types only; every pointer is a caller-owned DEVICE pointer unless named `h_*`.
 *   - every call takes the CUDA stream it must enqueue on (`void* stream` = cudaStream_t) and never synchronises.
 *   - return 0 on success, non-zero on failure; the message is available from paths_last_error().  No exceptions
 *     cross the ABI.  There is no CPU fallback: without a CUDA device every compute entry fails.
 *   - no allocation inside: the caller provides one workspace (size from paths_engine_workspace_bytes).
 *   - row tables are "packed varlen": rows of all slides of a level concatenated, seg[B+1] offsets, no padding.
 */
#ifndef PATHS_B200_H_
#define PATHS_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PATHS_MAX_LEVELS 8
#define PATHS_MAX_LAYERS 4
#define PATHS_ABI_VERSION 5

/* model_config + the train_config fields the hot path reads (config.py:19-79). */
typedef struct paths_dims {
  int32_t D;              /* patch_embed_dim */
  int32_t C;              /* hierarchical_ctx_mlp_hidden_dim (LSTM memory width) */
  int32_t d;              /* trans_dim */
  int32_t heads;          /* trans_heads */
  int32_t layers;         /* trans_layers (decoder layers; the encoder of nn.Transformer sees an empty sequence) */
  int32_t I;              /* importance_mlp_hidden_dim */
  int32_t F;              /* dim_feedforward = 4 * trans_dim (aggregator.py:30) */
  int32_t nl;             /* logits: nbins or len(filter_to_subtypes) (paths.py:25) */
  int32_t L;              /* num_levels */
  int32_t K[PATHS_MAX_LEVELS];  /* top_k_patches[l] for l < L-1; -1 keeps all (slide.py:294) */
  int32_t imp_mul;        /* importance_mode == "mul" */
  int32_t pe_mode;        /* 1 = "1d", 2 = "2d" */
  int32_t residual;       /* slide_ctx_mode: 0 "none", 1 "residual" (paths.py:130-131), 2 "concat" (paths.py:134-137) */
  int32_t task;           /* 0 survival NLL (utils.py:283), 1 cross-entropy (utils.py:277) */
  int32_t precision;      /* 0 fp32 SIMT, 1 tcgen05 bf16x3 (fp32-parity), 2 tcgen05 bf16 (fast) */
  int32_t lstm;           /* 1: shared LSTM cell (interface.py:10-58); 0: per-level RNN MLP hctx_mlp (paths.py:46-54,101-109) */
  int32_t hier_ctx;       /* hierarchical_ctx (only read when lstm == 0; config.py:106 forces it on with the LSTM) */
} paths_dims_t;

/* Offsets (in floats) of every live parameter inside the flat parameter / gradient buffers.  The stacked matrices
 * must be contiguous: Wg = [forget_gate; remember_gate; remember_map; out_select_gate].weight ([3C+D, 2D]),
 * bg likewise, Wuv = [importance_mlp.0.weight; global_agg.proj_in.weight] ([I+d, D]).  State-dict names: SURVEY §8b.
 * With slide_ctx_mode "concat" the classifier of level l is [nl, (l+1)*d]. */
enum paths_lstm_param { PATHS_P_WG = 0, PATHS_P_BG, PATHS_P_WMO, PATHS_P_BMO, PATHS_P_LSTM_COUNT };
enum paths_level_param {
  PATHS_P_WUV = 0, PATHS_P_B1, PATHS_P_W2, PATHS_P_B2, PATHS_P_BIN, PATHS_P_SPECIAL, PATHS_P_NFW, PATHS_P_NFB,
  PATHS_P_WC, PATHS_P_BC,
  PATHS_P_WH0, PATHS_P_BH0, PATHS_P_WH2, PATHS_P_BH2,   /* hctx_mlp.{0,2} of the RNN variant (offset -1 when lstm == 1) */
  PATHS_P_LEVEL_HEAD_COUNT
};
enum paths_layer_param {
  PATHS_P_WQKV = 0, PATHS_P_BQKV, PATHS_P_WO, PATHS_P_BO, PATHS_P_BCROSS, PATHS_P_FW1, PATHS_P_FB1, PATHS_P_FW2,
  PATHS_P_FB2, PATHS_P_N1W, PATHS_P_N1B, PATHS_P_N2W, PATHS_P_N2B, PATHS_P_N3W, PATHS_P_N3B, PATHS_P_LAYER_COUNT
};
#define PATHS_LEVEL_STRIDE (PATHS_P_LEVEL_HEAD_COUNT + PATHS_MAX_LAYERS * PATHS_P_LAYER_COUNT)
#define PATHS_NUM_OFFSETS (PATHS_P_LSTM_COUNT + PATHS_MAX_LEVELS * PATHS_LEVEL_STRIDE)
/* offsets[PATHS_P_LSTM_COUNT + l * PATHS_LEVEL_STRIDE + PATHS_P_LEVEL_HEAD_COUNT + j * PATHS_P_LAYER_COUNT + k] */

/* The feature-pyramid store: every slide's L grids resident in device (or mapped pinned host) memory, as written by the
 * reference preprocessor ([W, H, D] float32, all-zero vector = background; preprocess/preprocess.py:89,105,
 * preprocess/loader.py:14-18), plus a per-grid occupancy bitmap (bit x*H+y set iff the cell is tissue; replaces the
 * `fts[x, y].sum() != 0` probe of data_utils/slide.py:324).  Arrays are indexed [slide * L + level]. */
typedef struct paths_store {
  const float* const* grid_ptr;   /* device array of device pointers */
  const int32_t* gw;              /* grid extent in x */
  const int32_t* gh;              /* grid extent in y */
  const uint32_t* occ;            /* bitmap words */
  const uint64_t* occ_off;        /* word offset of each grid's bitmap */
  const int32_t* nonbg;           /* number of set bits of each grid (fallback of slide.py:336-352) */
  int32_t bf16;                   /* 0: grids are float32 (the reference's on-disk format); 1: grids were rounded to bfloat16 at load
                                     (grid_ptr then points at bfloat16 data: half the footprint and gather traffic) */
} paths_store_t;

typedef struct paths_engine paths_engine_t;

/* named internal buffers, for tests, per-level API and teacher forcing (paths_engine_buffer) */
enum paths_buffer {
  PATHS_BUF_COUNT = 0, PATHS_BUF_NTOK, PATHS_BUF_SEG, PATHS_BUF_ROW_SLIDE, PATHS_BUF_ROW_XY, PATHS_BUF_ROW_PARENT,
  PATHS_BUF_PARENT_LOCAL, PATHS_BUF_SEL, PATHS_BUF_SEL_COUNT, PATHS_BUF_CHILD_OF, PATHS_BUF_TSEG,
  PATHS_BUF_X, PATHS_BUF_G, PATHS_BUF_CN, PATHS_BUF_PT, PATHS_BUF_H, PATHS_BUF_Y, PATHS_BUF_UV, PATHS_BUF_ALPHA,
  PATHS_BUF_TOK, PATHS_BUF_Z, PATHS_BUF_F, PATHS_BUF_LOGITS,
  PATHS_BUF_HAZARDS, PATHS_BUF_LOSS, PATHS_BUF_DLOGITS, PATHS_BUF_OVERFLOW, PATHS_BUF_DF_IN, PATHS_BUF_DHROW, PATHS_BUF_DCROW,
  PATHS_BUF_LOSS_PER,
  PATHS_BUF_HF0,              /* FFN hidden activations of the level's first aggregator layer, [tokens, 4 * trans_dim] */
  PATHS_BUF_RISK,             /* [B] survival risk score -sum_k prod_{j<=k} (1 - hazard_j) (eval.py:60-61), written with the loss */
  PATHS_BUF__COUNT
};

const char* paths_last_error(void);
int paths_abi_version(void);

/* Bytes of workspace an engine for a batch of B slides needs.  caps[l] = row capacity of level l (level 0: total grid
 * cells of the batch; l >= 1: normally 4*K[l-1]*B), maxseg[l] = upper bound on one slide's rows at level l. */
size_t paths_engine_workspace_bytes(const paths_dims_t* dims, int B, const int32_t* h_caps, const int32_t* h_maxseg);
int paths_engine_create(paths_engine_t** out, const paths_dims_t* dims, int B, const int32_t* h_caps, const int32_t* h_maxseg,
                        void* workspace, size_t workspace_bytes);
void paths_engine_destroy(paths_engine_t* e);

/* params / grads: flat fp32 buffers of n_floats floats; h_offsets: PATHS_NUM_OFFSETS host int64 values (floats).  div: sinusoid frequencies,
 * d/4 values for "2d" (utils.py:56), d/2 for "1d" (utils.py:19), computed by the caller exactly as the reference does. */
int paths_engine_bind_params(paths_engine_t* e, const float* params, float* grads, int64_t n_floats, const int64_t* h_offsets,
                             const float* div);
int paths_engine_bind_store(paths_engine_t* e, const paths_store_t* h_store);
/* Re-split the bound weights into the bf16 (hi, lo) planes the tcgen05 path reads by TMA.  paths_engine_begin does this
 * itself; call it before paths_engine_level_forward when levels are driven one by one after a weight update. */
int paths_engine_refresh_weights(paths_engine_t* e, void* stream);

/* Level-0 row tables for the batch (slide.py:258-271: every cell of grid 0, x-major) and their feature gather.
 * slide_ids[B]: index of each batch slot into the store (device int32). */
int paths_engine_begin(paths_engine_t* e, const int32_t* slide_ids, void* stream);
/* Same, but level-0 features are packed from the collated zero-padded batch["fts"] [B, N, D] (data_utils/dataset.py:217,
 * utils.py:240 from_batch) already copied to the device, instead of being read from the store. */
int paths_engine_begin_padded(paths_engine_t* e, const int32_t* slide_ids, const float* fts_padded, int N, void* stream);
/* One level: LSTM cell step, importance, token projection + positional encoding, varlen transformer aggregator,
 * residual slide context, classifier.  Replaces PATHSProcessor.process (model/paths.py:66-146), LSTMCell.forward
 * (model/interface.py:31-58), TransformerAggregator.forward (model/aggregator.py:58-76). */
int paths_engine_level_forward(paths_engine_t* e, int level, void* stream);
/* Top-K select + child expansion + gather into level+1.  Replaces PreprocessedSlide.iter (data_utils/slide.py:277-360)
 * and collate_fn (data_utils/dataset.py:206-243). */
int paths_engine_select(paths_engine_t* e, int level, void* stream);
/* hazards/loss (+ dlogits) on the last level's logits: utils.py:263-277, nll_loss utils.py:283-305, and the survival risk
 * score the evaluator derives from the hazards (eval.py:60-61; PATHS_BUF_RISK).  scale = 1 / global batch size (the loss
 * is a batch mean). */
int paths_engine_loss(paths_engine_t* e, const int32_t* ybin, const int32_t* censored, float scale, void* stream);
/* Backward of one level (autograd of the above); accumulates into the bound grad buffer.  use_dlogits: take the
 * level's PATHS_BUF_DLOGITS as upstream gradient of its logits (set for the last level by paths_engine_loss). */
int paths_engine_level_backward(paths_engine_t* e, int level, int use_dlogits, void* stream);

/* Whole step: begin, L x (level_forward, select), loss.  utils.inference_end2end (utils.py:228-279). */
int paths_engine_forward(paths_engine_t* e, const int32_t* slide_ids, const int32_t* ybin, const int32_t* censored, float scale,
                         int with_loss, void* stream);
/* The same step on a collated batch whose level-0 features arrive as the reference's padded `fts` tensor [B, N, D]
 * (data_utils/dataset.py:229-243; device memory) instead of being gathered from the resident pyramid store. */
int paths_engine_forward_padded(paths_engine_t* e, const int32_t* slide_ids, const float* fts_padded, int N, const int32_t* ybin,
                                const int32_t* censored, float scale, int with_loss, void* stream);
/* ... or as the same rows without the padding: fts_packed [n_rows, D], slide after slide, each slide's cells in the order of
 * its level-0 bag (slide.py:260-269) -- what a loader that skips dataset.py:217's zero-padding ships across PCIe. */
int paths_engine_forward_packed(paths_engine_t* e, const int32_t* slide_ids, const float* fts_packed, int n_rows, const int32_t* ybin,
                                const int32_t* censored, float scale, int with_loss, void* stream);
/* loss.backward() of train.py:65: levels L-1..0.  zero_grads: clear the grad buffer first. */
int paths_engine_backward(paths_engine_t* e, int zero_grads, void* stream);
/* Data parallelism (new work; the reference is single-process): makes `stream` wait until every gradient of procs[level]
 * written by the latest paths_engine_backward is final -- after level 0 also the shared LSTM's -- so the caller can enqueue
 * that level's all-reduce on `stream` while the backward pass of the shallower levels is still running. */
int paths_engine_wait_grads(paths_engine_t* e, int level, void* stream);

/* The no-child fallback of data_utils/slide.py:336-352 ("every non-background cell of the next grid") can give one slide a
 * bag above the capacities the engine was sized for.  The kernels stay inside their buffers and raise a device flag; this
 * call waits (host-side) until the LAST select of the latest forward pass has run -- not for the end of the pass -- and
 * returns the flag and the row count that would have been needed.  A caller that sees the flag re-runs the step on an engine
 * sized for the worst case before using any output (the Python mirror does; the reference handles the case inline).
 * The only entry point that blocks the host. */
int paths_engine_overflow_poll(paths_engine_t* e, int32_t* h_flag, int32_t* h_rows_needed);

int paths_engine_buffer(paths_engine_t* e, int level, int which, void** ptr, size_t* bytes);
int paths_engine_set_precision(paths_engine_t* e, int precision);
/* nn.Transformer(dropout=p) in training mode (model/aggregator.py:25-33, train.py:131): the aggregators of the NEXT forward
 * pass (whole step or single level) drop attention probabilities, the three sublayer outputs and the FFN hidden activations
 * with probability p and rescale the survivors by 1/(1-p); the matching backward pass repeats the decisions.  Decisions come
 * from a counter-based hash of (seed, level, layer, site, token row, column) -- torch's generator stream is not reproducible
 * -- so the caller passes a fresh seed per step.  p = 0 (the default, and model.eval()) switches dropout off. */
int paths_engine_set_dropout(paths_engine_t* e, float p, uint64_t seed);
/* number of kernel launches issued by this engine since creation (for bench.py's gpu_launches) */
int64_t paths_engine_launch_count(paths_engine_t* e);

/* Per-kernel device timing for the roofline report: while enabled, the engine brackets the tagged kernels with CUDA events
 * on the launch stream; profile_read sums (and clears) the elapsed milliseconds and launch count of one tag. */
enum paths_prof_tag { PATHS_PROF_GATES_FWD = 0, PATHS_PROF_GATES_DW, PATHS_PROF_GATES_DH,
                      /* whole regions of a step (each includes every kernel launched inside it) */
                      PATHS_PROF_LSTM_FWD, PATHS_PROF_DEC_FWD, PATHS_PROF_SELECT, PATHS_PROF_DEC_BWD, PATHS_PROF_LSTM_BWD,
                      PATHS_PROF__COUNT };
int paths_engine_profile(paths_engine_t* e, int on);
int paths_engine_profile_read(paths_engine_t* e, int tag, double* total_ms, int64_t* launches);

/* Fused AdamW step over the flat parameter buffer -- torch.optim.AdamW as used by train.py:49,66.  params/exp_avg/exp_avg_sq
 * hold n_total floats, grads n_live; elements [n_live, n_total) (parameters whose gradient is identically zero: the dead
 * encoder / cross-attention weights) only receive the decoupled weight decay.  h_skip: n_skip (<= 16) host pairs
 * [start, end) of elements left untouched, in either part (parameters whose .grad the reference leaves at None).  step counts from 1. */
int paths_adamw_step(float* params, const float* grads, float* exp_avg, float* exp_avg_sq, int64_t n_live, int64_t n_total,
                     const int64_t* h_skip, int n_skip, float lr, float beta1, float beta2, float eps, float weight_decay, int step,
                     void* stream);

/* Occupancy bitmap of one [W*H, D] grid: bit c set iff any of the D floats of cell c is non-zero; *count += popcount. */
int paths_build_occupancy(const float* grid, int cells, int D, uint32_t* bits, int32_t* count, void* stream);

/* Importance heat-map compositing (replaces the numpy block of heatmap_visualise.py:142-178).  locs_yx: [n_patches, 2] int32
 * (y, x) patch origins in the pixel space of their own level, patches of level d at rows [level_off[d], level_off[d+1]);
 * importance: [n_patches] float32.  Every patch fills its (patch_size >> d)-sized square of an [L, H, W] canvas in level-0
 * pixel space with importance + 1e-4 (later patches overwrite earlier ones), finer levels are folded into coarser ones with
 * weight 0.5 wherever they are non-zero, zeros of the result become the smallest positive value.  out: [H, W] float64,
 * alpha: [H, W] float32 (0.5 where covered, else 0).  workspace: paths_heatmap_workspace_bytes(L, H, W) device bytes. */
size_t paths_heatmap_workspace_bytes(int L, int H, int W);
int paths_heatmap_composite(const int32_t* locs_yx, const float* importance, const int32_t* level_off, int32_t n_patches, int L,
                            int patch_size, int H, int W, void* workspace, size_t workspace_bytes, double* out, float* alpha,
                            void* stream);

/* ---- per-op entries: the single operations behind the reference's per-module methods ---- */
/* LSTMCell.forward (model/interface.py:31-58) on M rows: x, h [M, D], c [M, C] -> h_out [M, D], c_out [M, C].  Wg / bg: the gate
 * parameters stacked [forget_gate; remember_gate; remember_map; out_select_gate] ([3C + D, 2D] / [3C + D]); Wmo [D, C], bmo [D]
 * (mem_to_out).  Allocates its temporaries and synchronises the stream (stand-alone use; the engine's level step is the fast path). */
int paths_lstm_cell_fwd(int precision, const float* x, const float* h, const float* c, const float* Wg, const float* bg, const float* Wmo,
                        const float* bmo, float* h_out, float* c_out, int M, int D, int C, void* stream);
/* PreprocessedSlide.iter (data_utils/slide.py:277-360) for a batch of B slides: per-slide top-`keep` rows by importance (keep = -1:
 * all, in order), their 4 children each in the reference's block order, bounds + occupancy filter, order-preserving compaction,
 * no-child fallback (slide.py:336-352).  Level l in: alpha [rows], seg [B + 1], row_xy [rows, 2] patch coordinates, count [1].
 * Out: sel [B, kcap] kept rows (importance-descending, -1 padded), sel_count [B], child_of [rows, 4]; level l + 1 tables n_count [1],
 * n_ntok [1], n_seg [B + 1], n_row_slide / n_row_parent / n_parent_local [n_cap], n_row_xy [n_cap, 2]; overflow [2] = (flag, rows needed)
 * is raised when n_cap or the per-slide bound n_maxseg is exceeded.  maxseg: per-slide bound of level l; scratch: 2 B ints. */
int paths_select_expand(const float* alpha, const int32_t* seg, const int32_t* row_xy, const int32_t* count, int B, int cap, int keep, int kcap,
                        int maxseg, const paths_store_t* h_store, int level, int L, const int32_t* slide_ids, int32_t* sel, int32_t* sel_count,
                        int32_t* child_of, int32_t* n_count, int32_t* n_ntok, int32_t* n_seg, int32_t* n_row_slide, int32_t* n_row_xy,
                        int32_t* n_row_parent, int32_t* n_parent_local, int n_cap, int n_maxseg, int32_t* overflow, int32_t* scratch,
                        void* stream);
/* The feature gather of slide.py:331: X[m, :] = grid(slide_ids[row_slide[m]], level)[row_xy[m]] for m < M (*M_dev when non-null). */
int paths_gather_rows(const paths_store_t* h_store, int level, int L, const int32_t* slide_ids, const int32_t* row_slide, const int32_t* row_xy,
                      float* X, int D, int M, const int32_t* M_dev, void* stream);
/* TransformerAggregator.forward (model/aggregator.py:58-76) of one level: the caller has placed the projected, position-encoded
 * tokens in the level's PATHS_BUF_TOK rows (patch row m of slide s at token row m + s + 1) and set COUNT / NTOK / SEG / TSEG; the
 * special tokens are filled here, the decoder runs over the packed bags, the result lands in PATHS_BUF_Z ([B, d]). */
int paths_engine_aggregate(paths_engine_t* e, int level, void* stream);

/* Stand-alone varlen softmax self-attention over packed tokens -- the self-attention of nn.TransformerDecoderLayer as
 * model/aggregator.py:70-72 runs it, the key-padding mask replaced by per-slide segments.  qkv: [tokens, 3 d] (q | k | v, `heads`
 * heads of width d / heads, query scale 1/sqrt(head_dim) applied inside); slide s owns token rows [seg[s] + s, seg[s+1] + s + 1)
 * (seg: [B + 1] patch-row offsets; each bag has one extra leading token, the aggregator's special token).  out: [tokens, d],
 * lse: [tokens, heads] log-sum-exp of the scores.  max_bag: upper bound on seg[s+1] - seg[s].  precision 1 / 2 run bags of at
 * most 128 tokens on the tensor cores (tcgen05: three-term bf16 split / plain bf16), 0 and longer bags on the fp32 SIMT kernels.
 * dropout_p > 0 drops probabilities by the hash of paths_engine_set_dropout (level 0, layer 0).
 * bwd: dqkv [tokens, 3 d] receives dq | dk | dv; dvec: [tokens, heads] scratch. */
int paths_attention_fwd(int precision, const float* qkv, const int32_t* seg, int B, float* out, float* lse, int d, int heads, int max_bag,
                        float dropout_p, uint64_t dropout_seed, void* stream);
int paths_attention_bwd(int precision, const float* qkv, const int32_t* seg, int B, const float* out, const float* lse, const float* dout,
                        float* dqkv, float* dvec, int d, int heads, int max_bag, float dropout_p, uint64_t dropout_seed, void* stream);

/* Allocation-free variant of paths_attention_bwd.  This entry point has NO counterpart in the reference: it was added
 * for callers that cannot allocate device memory (code under CUDA-graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc).  Same arguments and results as paths_attention_bwd, except that the [tokens, heads] `dvec`
 * scratch is a file-scope device array inside the library (4 MiB, held by every GPU that loads the library until the process
 * exits) and the caller states `tokens` (rows of qkv = seg[B] + B) so the size can be checked on the host: the call fails with a
 * clear error, launching nothing, when tokens * heads > PATHS_ATTENTION_BWD_STATIC_MAX_FLOATS.  It calls no cudaMalloc / cudaFree
 * or allocating library routine.  Every call on a device shares the one array (it keeps the last call's contents; each call
 * writes every element it later reads, nothing needs clearing): calls that OVERLAP on the GPU -- different streams or host
 * threads -- race on it; serialise them or use paths_attention_bwd with separate dvec buffers. */
#define PATHS_ATTENTION_BWD_STATIC_MAX_FLOATS (1 << 20)
int paths_attention_bwd_static(int precision, const float* qkv, const int32_t* seg, int B, int tokens, const float* out, const float* lse,
                               const float* dout, float* dqkv, int d, int heads, int max_bag, float dropout_p, uint64_t dropout_seed,
                               void* stream);

/* Stand-alone dense contractions (tests; the engine uses the same kernels).
 * C[M,N] (+)= act(A[M,K] . B^T + bias) with B [N,K] (b_kn = 0) or [K,N] (b_kn = 1); flags: 1 relu, 2 accumulate. */
int paths_gemm_nt(int precision, const float* A, int lda, const float* B, int ldb, int b_kn, float* C, int ldc, const float* bias,
                  int M, int N, int K, int flags, void* stream);
/* C[NI,NJ] += P[M,NI]^T . Q[M,NJ] */
int paths_gemm_tn(int precision, const float* P, int ldp, const float* Q, int ldq, float* C, int ldc, int M, int NI, int NJ, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* PATHS_B200_H_ */
