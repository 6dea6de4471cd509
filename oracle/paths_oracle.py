"""CPU ORACLE for the PATHS hierarchical forward/backward path.  TEST INFRASTRUCTURE ONLY.

This file restates, with plain torch CPU tensor arithmetic, the algorithm of the reference
(zzbuzzard/PATHS) for the one hot path this repository accelerates.  It is imported only by
`tests/`, `__graft_entry__.smoke()` and `bench.py`'s CPU-baseline / `--impl reference` legs, always as
the checker or as the timed CPU baseline, never by the product path under `paths_b200/` (which fails
loudly when its CUDA library is missing and has no CPU fallback).

The reference is pure Python on top of PyTorch, so its arithmetic *is* ATen's; the restatement below
uses the same primitive ops (matmul, sigmoid/tanh, softmax, layer_norm, topk, sin/cos) in the same
padded [B, N, ...] layout and the same per-slide CPU recursion, written functionally over a
`state_dict` whose keys are the reference's.  `nn.Transformer` is *not* used: the aggregator is written
out by hand (the encoder sees a length-0 sequence, so cross-attention collapses to its output bias).

PARITY PINNING.  The reference ships no tests, golden vectors or pretrained weights,
so the oracle is pinned against the reference itself: `tests/golden/make_golden.py` imports the
unmodified reference from /root/reference, runs it on seeded synthetic pyramids with seeded weights and
commits the outputs under `tests/golden/`; `tests/test_oracle_golden.py` checks this file against them.

Every function cites the reference lines it follows (paths relative to the reference checkout).
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch
import torch.nn.functional as F

Tensor = torch.Tensor


@dataclass
class OracleConfig:
    """The subset of config.json the hot path reads (config.py:19-79)."""
    num_levels: int = 5
    top_k_patches: Sequence[int] = (20, 20, 20, 20)
    patch_embed_dim: int = 1024
    trans_dim: int = 128
    trans_heads: int = 4
    trans_layers: int = 2
    importance_mlp_hidden_dim: int = 128
    hierarchical_ctx_mlp_hidden_dim: int = 256
    patch_size: int = 256
    lstm: bool = True
    hierarchical_ctx: bool = True
    slide_ctx_mode: str = "residual"      # residual / concat / none
    importance_mode: str = "mul"          # mul / none
    pos_encoding_mode: str = "2d"         # 1d / 2d
    task: str = "survival"
    nbins: int = 4

    @staticmethod
    def from_config(cfg) -> "OracleConfig":
        mc = cfg.model_config
        return OracleConfig(
            num_levels=cfg.num_levels, top_k_patches=list(cfg.top_k_patches),
            patch_embed_dim=mc.patch_embed_dim, trans_dim=mc.trans_dim, trans_heads=mc.trans_heads,
            trans_layers=mc.trans_layers, importance_mlp_hidden_dim=mc.importance_mlp_hidden_dim,
            hierarchical_ctx_mlp_hidden_dim=mc.hierarchical_ctx_mlp_hidden_dim, patch_size=mc.patch_size,
            lstm=mc.lstm, hierarchical_ctx=mc.hierarchical_ctx, slide_ctx_mode=mc.slide_ctx_mode,
            importance_mode=mc.importance_mode, pos_encoding_mode=mc.pos_encoding_mode, task=cfg.task,
            nbins=cfg.nbins)

    @property
    def ctx_dim(self) -> Tuple[int, int]:
        # model/paths.py:148-151
        if self.lstm:
            return self.trans_dim, self.patch_embed_dim + self.hierarchical_ctx_mlp_hidden_dim
        return self.trans_dim, self.patch_embed_dim


# ----------------------------------------------------------------------------------------------
# Numerics of one level
# ----------------------------------------------------------------------------------------------

def _lin(sd: Dict[str, Tensor], name: str, x: Tensor) -> Tensor:
    return x @ sd[name + ".weight"].t() + sd[name + ".bias"]


def lstm_cell(sd: Dict[str, Tensor], xs: Tensor, hs: Tensor, cs: Tensor) -> Tuple[Tensor, Tensor]:
    """model/interface.py:49-56.  c' = c*sig(Wf xh) + sig(Wr xh)*tanh(Wm xh); h' = sig(Wo xh)*tanh(Wmo c')."""
    xh = torch.cat((xs, hs), dim=-1)
    f = torch.sigmoid(_lin(sd, "lstm.forget_gate.0", xh))
    r = torch.sigmoid(_lin(sd, "lstm.remember_gate.0", xh))
    m = torch.tanh(_lin(sd, "lstm.remember_map.0", xh))
    c_new = cs * f + r * m
    o = torch.sigmoid(_lin(sd, "lstm.out_select_gate.0", xh))
    h_new = o * torch.tanh(_lin(sd, "lstm.mem_to_out.0", c_new))
    return h_new, c_new


def importance_scores(sd: Dict[str, Tensor], depth: int, y_valid: Tensor) -> Tensor:
    """model/paths.py:40-44,95: sigmoid(w2 . relu(W1 y + b1) + b2) on the valid rows."""
    p = f"procs.{depth}.importance_mlp"
    return torch.sigmoid(_lin(sd, p + ".2", torch.relu(_lin(sd, p + ".0", y_valid))))[..., 0]


def pe_2d(xpos: Tensor, ypos: Tensor, dim: int, dtype=torch.float32, k: float = 10000.0) -> Tensor:
    """utils.py:47-67.  Note the divisor is `dim`, and only dim//4 frequencies exist."""
    n = xpos.shape[0]
    div = torch.exp(torch.arange(0, dim // 2, 2) * (-math.log(k) / dim))[None].to(dtype)
    pe = torch.zeros(n, dim, dtype=dtype)
    xp = xpos.unsqueeze(-1)
    yp = ypos.unsqueeze(-1)
    pe[:, 0:dim // 2:2] = torch.sin(xp * div)
    pe[:, 1:dim // 2:2] = torch.cos(xp * div)
    pe[:, dim // 2::2] = torch.sin(yp * div)
    pe[:, (dim // 2) + 1::2] = torch.cos(yp * div)
    return pe


def pe_1d(length: int, dim: int, dtype=torch.float32, k: float = 10000.0) -> Tensor:
    """utils.py:16-23."""
    pos = torch.arange(length).unsqueeze(1)
    div = torch.exp(torch.arange(0, dim, 2) * (-math.log(k) / dim)).to(dtype)
    pe = torch.zeros(length, dim, dtype=dtype)
    pe[:, 0::2] = torch.sin(pos * div)
    pe[:, 1::2] = torch.cos(pos * div)
    return pe


# ---- dropout with explicit masks ---------------------------------------------------------------------------------
# nn.Transformer(dropout=p) in training mode (model/aggregator.py:25-33; train.py:131) draws its masks from torch's
# generator, which no other implementation can reproduce.  WHERE the masks act and how survivors are rescaled is
# pinned against torch itself (tests/golden/make_golden_dropout.py records the masks torch drew and
# tests/test_oracle_golden.py replays them through `decoder_stack`); WHICH elements are dropped is a convention of
# the CUDA path -- a counter-based hash (paths_b200/csrc/common.cuh: drop_keep) that `HashDropout` restates.
DROP_ATTN, DROP_SA_OUT, DROP_CROSS_OUT, DROP_FFN_INNER, DROP_FFN_OUT = range(5)


def _fmix64(x: np.ndarray) -> np.ndarray:
    x = np.asarray(x, dtype=np.uint64).copy()
    with np.errstate(over="ignore"):
        x ^= x >> np.uint64(33); x *= np.uint64(0xff51afd7ed558ccd)
        x ^= x >> np.uint64(33); x *= np.uint64(0xc4ceb9fe1a85ec53)
        x ^= x >> np.uint64(33)
    return x


class HashDropout:
    """mask(level, layer, site, rows, cols) -> float32 array of {0, 1/(1-p)}: element (row, col) is kept iff
    fmix64(site_key ^ (row << 32 | col)) >> 32 >= p * 2**32; `rows` are level-local packed token rows."""

    def __init__(self, p: float, seed: int):
        self.p = float(np.float32(p))
        self.seed = int(seed) & 0xFFFFFFFFFFFFFFFF
        t = float(np.float32(p)) * 4294967296.0
        self.thresh = 0 if t <= 0 else 4294967295 if t >= 4294967295.0 else int(t)
        self.inv_keep = np.float32(1.0) / (np.float32(1.0) - np.float32(p))

    def site_key(self, level: int, layer: int, site: int) -> np.uint64:
        with np.errstate(over="ignore"):
            mult = np.uint64(0x9E3779B97F4A7C15) * np.uint64((level * 8 + layer) * 8 + site + 1)
        return _fmix64(np.uint64(self.seed) ^ mult)

    def mask(self, level: int, layer: int, site: int, rows: np.ndarray, cols: np.ndarray) -> Tensor:
        key = self.site_key(level, layer, site)
        h = _fmix64(key ^ ((rows.astype(np.uint64) << np.uint64(32)) | cols.astype(np.uint64))) >> np.uint64(32)
        return torch.from_numpy(np.where(h >= np.uint64(self.thresh), self.inv_keep, np.float32(0.0)).astype(np.float32))

    def elem(self, level: int, layer: int, site: int, rows: Tensor, width: int) -> Tensor:
        """[B, T, width] mask of a sublayer output; rows: [B, T] packed token rows."""
        r = rows.numpy().astype(np.uint64)[:, :, None]
        return self.mask(level, layer, site, r, np.arange(width, dtype=np.uint64)[None, None, :])

    def attn(self, level: int, layer: int, rows: Tensor, nhead: int) -> Tensor:
        """[B, nhead, T, T] mask of the attention probabilities (query row, head << 24 | key index)."""
        B, T = rows.shape
        r = rows.numpy().astype(np.uint64)[:, None, :, None]
        col = (np.arange(nhead, dtype=np.uint64)[None, :, None, None] << np.uint64(24)) | np.arange(T, dtype=np.uint64)[None, None, None, :]
        return self.mask(level, layer, DROP_ATTN, r, col)


class RecordedDropout:
    """Masks recorded from torch (tests/golden/make_golden_dropout.py), replayed by site."""

    def __init__(self, masks: Dict[Tuple[int, int, int], Tensor]):
        self.masks = masks

    def elem(self, level, layer, site, rows, width):
        return self.masks[(level, layer, site)]

    def attn(self, level, layer, rows, nhead):
        return self.masks[(level, layer, DROP_ATTN)]


def _mha_self(sd: Dict[str, Tensor], p: str, t: Tensor, key_pad: Tensor, nhead: int, amask: Optional[Tensor] = None) -> Tensor:
    """Self-attention of nn.MultiheadAttention (batch_first) with a key padding mask; `amask` ([B, h, T, T] of
    {0, 1/(1-p)}) is the dropout applied to the softmax probabilities (F.scaled_dot_product_attention's dropout_p)."""
    B, T, d = t.shape
    dh = d // nhead
    qkv = t @ sd[p + ".in_proj_weight"].t() + sd[p + ".in_proj_bias"]
    q, k, v = qkv.split(d, dim=-1)
    q = q.view(B, T, nhead, dh).transpose(1, 2)
    k = k.view(B, T, nhead, dh).transpose(1, 2)
    v = v.view(B, T, nhead, dh).transpose(1, 2)
    s = (q @ k.transpose(-1, -2)) / math.sqrt(dh)
    s = s.masked_fill(key_pad[:, None, None, :], float("-inf"))
    a = torch.softmax(s, dim=-1)
    if amask is not None:
        a = a * amask.to(a.dtype)
    o = (a @ v).transpose(1, 2).reshape(B, T, d)
    return o @ sd[p + ".out_proj.weight"].t() + sd[p + ".out_proj.bias"]


def decoder_stack(sd: Dict[str, Tensor], prefix: str, t: Tensor, key_pad: Tensor, nhead: int, layers: int,
                  eps: float = 1e-5, drop=None, level: int = 0, rows: Optional[Tensor] = None) -> Tensor:
    """What nn.Transformer(d, nhead, layers, layers, 4d, batch_first=True) computes for an EMPTY
    encoder input (model/paths.py:116, model/aggregator.py:25-33,70-72): post-norm decoder layers
    whose cross-attention over zero keys yields exactly `multihead_attn.out_proj.bias`.
    Training-mode dropout (`drop`: HashDropout / RecordedDropout) acts where nn.TransformerDecoderLayer applies it:
    attention probabilities, dropout1(self-attn block), dropout2(cross-attn block = its bias), the FFN's inner
    dropout after the ReLU, dropout3(FFN block) -- each block INCLUDING its output bias."""
    d = t.shape[-1]
    for j in range(layers):
        p = f"{prefix}.decoder.layers.{j}"
        ln = lambda x, n: F.layer_norm(x, (d,), sd[f"{p}.{n}.weight"], sd[f"{p}.{n}.bias"], eps)
        if drop is None:
            t = ln(t + _mha_self(sd, p + ".self_attn", t, key_pad, nhead), "norm1")
            t = ln(t + sd[p + ".multihead_attn.out_proj.bias"], "norm2")
            ff = _lin(sd, p + ".linear2", torch.relu(_lin(sd, p + ".linear1", t)))
            t = ln(t + ff, "norm3")
            continue
        m = lambda site, width: drop.elem(level, j, site, rows, width).to(t.dtype)
        sa = _mha_self(sd, p + ".self_attn", t, key_pad, nhead, drop.attn(level, j, rows, nhead))
        t = ln(t + sa * m(DROP_SA_OUT, d), "norm1")
        cross = sd[p + ".multihead_attn.out_proj.bias"].expand_as(t)
        t = ln(t + cross * m(DROP_CROSS_OUT, d), "norm2")
        h = torch.relu(_lin(sd, p + ".linear1", t))
        h = h * m(DROP_FFN_INNER, h.shape[-1])
        ff = _lin(sd, p + ".linear2", h)
        t = ln(t + ff * m(DROP_FFN_OUT, d), "norm3")
    return F.layer_norm(t, (d,), sd[prefix + ".decoder.norm.weight"], sd[prefix + ".decoder.norm.bias"], eps)


def aggregate(sd: Dict[str, Tensor], cfg: OracleConfig, depth: int, tokens: Tensor, num_ims: Tensor, drop=None) -> Tensor:
    """model/aggregator.py:58-76: prepend the special token, mask padding keys, read token 0."""
    B, N, d = tokens.shape
    p = f"procs.{depth}.global_agg"
    special = sd[p + ".special_token"].view(1, 1, d).expand(B, 1, d)
    seq = torch.cat((special, tokens), dim=1)
    lengths = num_ims + 1
    key_pad = torch.arange(N + 1)[None] >= lengths[:, None]             # utils.py:97-103
    # packed token row of padded position (b, t): slide b's tokens follow those of slides < b, special token first
    tseg = torch.cumsum(lengths, 0) - lengths
    rows = tseg[:, None] + torch.arange(N + 1)[None]
    out = decoder_stack(sd, p + ".transformer", seq, key_pad, cfg.trans_heads, cfg.trans_layers, drop=drop, level=depth,
                        rows=rows)
    return out[:, 0]


def level_forward(sd: Dict[str, Tensor], cfg: OracleConfig, depth: int, batch: Dict[str, Tensor], drop=None) -> Dict[str, Tensor]:
    """model/paths.py:66-146 on a padded batch (keys as data_utils/patch_batch.py:20-27)."""
    fts, locs, num_ims = batch["fts"], batch["locs"], batch["num_ims"]
    ctx_slide, ctx_patch = batch["ctx_slide"], batch["ctx_patch"]
    B, N, D = fts.shape
    C = cfg.hierarchical_ctx_mlp_hidden_dim
    valid = torch.arange(N)[None] < num_ims[:, None]                       # patch_batch.py:66-70
    x = fts
    if cfg.lstm:
        if depth == 0:                                                      # paths.py:78-80
            hs = torch.zeros(B, N, D, dtype=fts.dtype)
            cs = torch.zeros(B, N, C, dtype=fts.dtype)
        else:                                                               # paths.py:84-86
            st = ctx_patch[:, :, -1]
            hs, cs = st[..., :D], st[..., D:]
        hs, cs = lstm_cell(sd, x, hs, cs)
        x = x + hs                                                          # paths.py:89
        patch_ctx = torch.cat((hs, cs), dim=-1)                             # paths.py:91
    imp = torch.zeros(B, N, dtype=fts.dtype)                                # utils.py:106-115
    imp[valid] = importance_scores(sd, depth, x[valid])
    if cfg.importance_mode == "mul":
        x = x * imp[..., None]                                              # paths.py:96-98
    if not cfg.lstm:                                                        # paths.py:101-109
        if depth > 0 and cfg.hierarchical_ctx:
            p = f"procs.{depth}.hctx_mlp"
            hctx = torch.zeros(B, N, D, dtype=fts.dtype)
            hv = ctx_patch[:, :, -1][valid]
            hctx[valid] = _lin(sd, p + ".2", torch.relu(_lin(sd, p + ".0", hv)))
            x = x + hctx
        patch_ctx = x
    d = cfg.trans_dim
    pa = f"procs.{depth}.global_agg"
    tok = _lin(sd, pa + ".proj_in", x)                                      # aggregator.py:39,48
    if cfg.pos_encoding_mode == "2d":
        pl = torch.div(locs, cfg.patch_size, rounding_mode="floor")        # paths.py:120
        tok = tok + pe_2d(pl[:, :, 0].reshape(-1), pl[:, :, 1].reshape(-1), d, fts.dtype).view(B, N, d)
    else:
        tok = tok + pe_1d(N, d, fts.dtype)[None]
    slide_ft = aggregate(sd, cfg, depth, tok, num_ims, drop)                    # paths.py:127
    if cfg.slide_ctx_mode == "residual" and ctx_slide.shape[1] > 0:         # paths.py:130-131
        slide_ft = slide_ft + ctx_slide[:, -1]
    pc = f"procs.{depth}.classification_layer"
    if cfg.slide_ctx_mode == "concat":                                      # paths.py:134-137
        logits = _lin(sd, pc, torch.cat((ctx_slide.flatten(1), slide_ft), dim=1))
    else:
        logits = _lin(sd, pc, slide_ft)
    return {"logits": logits, "ctx_slide": slide_ft, "ctx_patch": patch_ctx, "importance": imp}


# ----------------------------------------------------------------------------------------------
# Recursion between levels (CPU, per slide) and collation
# ----------------------------------------------------------------------------------------------

def level0_items(grids: Sequence[Tensor], cfg: OracleConfig, dtype=torch.float32) -> Dict[str, Tensor]:
    """data_utils/slide.py:258-271,365-384: level 0 = every grid cell (background included), x-major."""
    w, h, _ = grids[0].shape
    ii, jj = torch.meshgrid(torch.arange(w), torch.arange(h), indexing="ij")
    locs = torch.stack((ii.reshape(-1), jj.reshape(-1)), dim=1)
    return {
        "fts": grids[0].reshape(w * h, -1).to(dtype),
        "locs": locs * cfg.patch_size,
        "parent_inds": torch.arange(w * h),
        "ctx_patch": torch.zeros(w * h, 0, cfg.ctx_dim[1], dtype=dtype),
        "ctx_slide": torch.zeros(0, cfg.ctx_dim[0], dtype=dtype),
    }


def select_expand(next_grid: Tensor, n: int, locs_px: Tensor, ctx_slide: Tensor, ctx_patch: Tensor,
                  importance: Tensor, new_ctx_slide: Tensor, new_ctx_patch: Tensor, keep: int,
                  patch_size: int, record: Optional[dict] = None) -> Dict[str, Tensor]:
    """data_utils/slide.py:277-360 for one slide: top-K by importance (sorted descending), 4 children per
    kept patch laid out as four blocks, in-bounds and non-background filter, feature gather, and the
    no-child-survives fallback."""
    locs = torch.div(locs_px[:n], patch_size, rounding_mode="floor")        # slide.py:280,288
    ctx_slide = torch.cat((ctx_slide, new_ctx_slide[None]), dim=0)          # slide.py:291
    ctx_patch = torch.cat((ctx_patch[:n], new_ctx_patch[:n, None]), dim=1)  # slide.py:292
    imp = importance[:n].detach()
    keep_inds = torch.arange(n)
    if keep != -1:                                                          # slide.py:294-301
        keep_inds = torch.topk(imp, min(n, keep)).indices
        ctx_patch = ctx_patch[keep_inds]
        locs = locs[keep_inds]
    k = locs.shape[0]
    base = locs * 2                                                         # slide.py:307
    parent = torch.arange(k).repeat(4)                                      # slide.py:311
    off = torch.tensor([[0, 0], [0, 1], [1, 0], [1, 1]])
    new_locs = torch.cat([base + off[i] for i in range(4)], dim=0)          # slide.py:312-315
    ctx_patch = torch.cat((ctx_patch,) * 4, dim=0)                          # slide.py:318
    x, y, _ = next_grid.shape
    inb = torch.logical_and(new_locs[:, 0] < x, new_locs[:, 1] < y)         # slide.py:322
    probe = new_locs * inb[:, None]
    nonbg = next_grid[probe[:, 0], probe[:, 1]].sum(dim=1) != 0            # slide.py:324
    flt = torch.logical_and(inb, nonbg)
    new_locs, parent, ctx_patch = new_locs[flt], parent[flt], ctx_patch[flt]
    fallback = new_locs.shape[0] == 0
    if fallback:                                                            # slide.py:336-352
        ii, jj = torch.meshgrid(torch.arange(x), torch.arange(y), indexing="ij")
        new_locs = torch.stack((ii.reshape(-1), jj.reshape(-1)), dim=1)
        parent = torch.arange(x * y)
        ctx_patch = torch.zeros(x * y, ctx_patch.shape[1], ctx_patch.shape[2], dtype=ctx_patch.dtype)
        f2 = next_grid[new_locs[:, 0], new_locs[:, 1]].sum(dim=1) != 0
        if int(f2.count_nonzero()) == 0:
            f2[:] = True
        new_locs, parent, ctx_patch = new_locs[f2], parent[f2], ctx_patch[f2]
    fts = next_grid[new_locs[:, 0], new_locs[:, 1]].to(ctx_patch.dtype)     # slide.py:331
    if record is not None:
        record.update(keep_inds=keep_inds.clone(), child_locs=new_locs.clone(), parent_inds=parent.clone(),
                      fallback=fallback)
    return {"fts": fts, "ctx_patch": ctx_patch, "ctx_slide": ctx_slide,
            "locs": new_locs * patch_size, "parent_inds": parent}


def collate(items: List[Dict[str, Tensor]]) -> Dict[str, Tensor]:
    """data_utils/dataset.py:206-243: zero-pad the variable-length fields to the longest bag."""
    num = [it["locs"].shape[0] for it in items]
    mx = max(num)
    pad_rows = lambda t, n: F.pad(t, (0, 0) * (t.dim() - 1) + (0, mx - n))
    fts = torch.stack([pad_rows(it["fts"], n) for it, n in zip(items, num)])
    locs = torch.stack([pad_rows(it["locs"], n) for it, n in zip(items, num)])
    par = torch.stack([F.pad(it["parent_inds"], (0, mx - n)) for it, n in zip(items, num)])
    _, k, dd = items[0]["ctx_patch"].shape
    if k == 0:                                                              # dataset.py:223-227
        ctxp = torch.zeros(len(items), mx, 0, dd, dtype=items[0]["ctx_patch"].dtype)
    else:
        ctxp = torch.stack([pad_rows(it["ctx_patch"], n) for it, n in zip(items, num)])
    ctxs = torch.stack([it["ctx_slide"] for it in items])
    return {"fts": fts, "locs": locs, "parent_inds": par, "ctx_patch": ctxp, "ctx_slide": ctxs,
            "num_ims": torch.tensor(num, dtype=torch.long)}


def nll_loss(hazards: Tensor, y: Tensor, c: Tensor, alpha: float = 0.4, eps: float = 1e-7) -> Tensor:
    """utils.py:283-305 (discrete-time survival NLL, mean over the batch)."""
    B = hazards.shape[0]
    surv = torch.cumprod(1 - hazards, dim=1)
    surv_p = torch.cat([torch.ones(B, 1, dtype=hazards.dtype), surv], dim=1)
    r = torch.arange(B)
    unc = -(1 - c) * (torch.log(surv_p[r, y].clamp(min=eps)) + torch.log(hazards[r, y].clamp(min=eps)))
    cen = -c * torch.log(surv_p[r, y + 1].clamp(min=eps))
    return ((1 - alpha) * (cen + unc) + alpha * unc).mean()


@dataclass
class LevelRecord:
    num_ims: Tensor
    locs: Tensor                      # [B, N, 2] pixel units (as fed to the model)
    logits: Tensor
    ctx_slide: Tensor
    importance: Tensor                # [B, N]
    ctx_patch: Tensor                 # [B, N, Dp] the level's own new patch state
    keep_inds: List[Tensor] = field(default_factory=list)      # per slide (levels 0..L-2)
    child_locs: List[Tensor] = field(default_factory=list)
    parent_inds: List[Tensor] = field(default_factory=list)
    fallback: List[bool] = field(default_factory=list)


def end2end(sd: Dict[str, Tensor], cfg: OracleConfig, slides: Sequence[Sequence[Tensor]],
            survival_bin: Optional[Tensor] = None, censored: Optional[Tensor] = None,
            subtype: Optional[Tensor] = None, dtype=torch.float32, record: bool = False,
            forced_keep: Optional[List[List[Tensor]]] = None, drop=None):
    """utils.py:228-279.  `slides[j]` is the list of L grids of slide j.  Returns (hazards|logits, loss,
    records).  `forced_keep[l][j]`, when given, replaces the top-K indices (teacher forcing for tests)."""
    sd = {k: v.to(dtype) if v.is_floating_point() else v for k, v in sd.items()}
    batch = collate([level0_items(g, cfg, dtype) for g in slides])
    recs: List[LevelRecord] = []
    out = None
    for i in range(cfg.num_levels):
        out = level_forward(sd, cfg, i, batch, drop)
        rec = LevelRecord(batch["num_ims"].clone(), batch["locs"].clone(), out["logits"], out["ctx_slide"],
                          out["importance"], out["ctx_patch"]) if record else None
        if i != cfg.num_levels - 1:
            items = []
            for j, grids in enumerate(slides):
                n = int(batch["num_ims"][j])
                r: dict = {}
                imp = out["importance"][j]
                keep = cfg.top_k_patches[i]
                if forced_keep is not None:
                    fk = forced_keep[i][j]
                    imp = torch.zeros_like(imp[:n])
                    imp[fk] = torch.arange(len(fk), 0, -1, dtype=imp.dtype)
                    keep = len(fk)
                items.append(select_expand(grids[i + 1], n, batch["locs"][j], batch["ctx_slide"][j],
                                           batch["ctx_patch"][j], imp, out["ctx_slide"][j], out["ctx_patch"][j],
                                           keep, cfg.patch_size, r))
                if rec is not None:
                    rec.keep_inds.append(r["keep_inds"]); rec.child_locs.append(r["child_locs"])
                    rec.parent_inds.append(r["parent_inds"]); rec.fallback.append(r["fallback"])
            batch = collate(items)
        if rec is not None:
            recs.append(rec)
    logits = out["logits"]
    if cfg.task == "survival":
        hazards = torch.sigmoid(logits)
        loss = nll_loss(hazards, survival_bin, censored.to(dtype))
        return hazards, loss, recs
    loss = F.cross_entropy(logits, subtype if subtype is not None else survival_bin)      # utils.py:274-278
    return logits, loss, recs


def grads(sd: Dict[str, Tensor], cfg: OracleConfig, slides, survival_bin, censored, dtype=torch.float32,
          forced_keep=None, drop=None):
    """loss.backward() of train.py:65 on leaf copies of the state dict; returns (hazards, loss, grad dict).
    Parameters that the loss does not reach get None (the reference leaves .grad = None)."""
    leaf = {k: v.detach().to(dtype).clone().requires_grad_(True) for k, v in sd.items() if v.is_floating_point()}
    hz, loss, recs = end2end(leaf, cfg, slides, survival_bin, censored, dtype=dtype, record=True,
                             forced_keep=forced_keep, drop=drop)
    loss.backward()
    return hz.detach(), loss.detach(), {k: (v.grad if v.grad is not None else None) for k, v in leaf.items()}, recs
