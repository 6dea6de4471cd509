"""Host side of the level engine: flat parameter layout, HBM-resident feature-pyramid store, and a thin wrapper over
the C ABI (include/paths_b200.h).  PyTorch is used here only for device memory, streams and autograd plumbing.
"""
from __future__ import annotations

import ctypes as C
import math
from dataclasses import dataclass
from typing import Dict, List, Optional, Sequence, Tuple

import torch

from . import _lib as L


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def make_dims(model_config, train_config, precision: int = 0) -> L.Dims:
    """paths_dims_t from the reference's PATHSProcessorConfig / Config objects (config.py:19-79)."""
    mc, tc = model_config, train_config
    if mc.slide_ctx_mode not in ("residual", "concat", "none"):
        raise ValueError(f"unknown slide_ctx_mode {mc.slide_ctx_mode!r}")
    d = L.Dims()
    d.D, d.C, d.d, d.heads, d.layers = mc.patch_embed_dim, mc.hierarchical_ctx_mlp_hidden_dim, mc.trans_dim, mc.trans_heads, mc.trans_layers
    d.I, d.F = mc.importance_mlp_hidden_dim, 4 * mc.trans_dim
    d.nl = tc.nbins if tc.task == "survival" else len(tc.filter_to_subtypes)
    d.L = tc.num_levels
    ks = list(tc.top_k_patches) if not isinstance(tc.top_k_patches, int) else [tc.top_k_patches] * (tc.num_levels - 1)
    for i in range(L.MAX_LEVELS):
        d.K[i] = int(ks[i]) if i < len(ks) else -1
    d.imp_mul = 1 if mc.importance_mode == "mul" else 0
    d.pe_mode = 2 if mc.pos_encoding_mode == "2d" else 1
    d.residual = {"none": 0, "residual": 1, "concat": 2}[mc.slide_ctx_mode]
    d.lstm = 1 if mc.lstm else 0
    d.hier_ctx = 1 if mc.hierarchical_ctx else 0
    d.task = 0 if tc.task == "survival" else 1
    d.precision = precision
    return d


def pe_div_table(dims: L.Dims, device) -> torch.Tensor:
    """Sinusoid frequencies exactly as the reference computes them (utils.py:19 / utils.py:56), float32."""
    d = dims.d
    if dims.pe_mode == 2:
        t = torch.exp(torch.arange(0, d // 2, 2) * (-math.log(10000.0) / d))
    else:
        t = torch.exp(torch.arange(0, d, 2) * (-math.log(10000.0) / d))
    return t.float().to(device).contiguous()


# ------------------------------------------------------------------------------------------------------------------
class ParamLayout:
    """Flat fp32 parameter / gradient buffers.  Every live parameter of the model becomes a view into `flat`; the stacked
    matrices the kernels contract against ([Wf;Wr;Wm;Wo], [W1;W_in]) are contiguous by construction, so the gate GEMM
    and its weight gradient are single contractions and the DP all-reduce runs over one buffer."""

    def __init__(self, model, dims: L.Dims):
        self.dims = dims
        sd_names: List[Tuple[int, List[str]]] = []
        g = "lstm.{}.0.{}"
        gates = ["forget_gate", "remember_gate", "remember_map", "out_select_gate"]
        if dims.lstm:
            sd_names.append((L.P_WG, [g.format(n, "weight") for n in gates]))
            sd_names.append((L.P_BG, [g.format(n, "bias") for n in gates]))
            sd_names.append((L.P_WMO, ["lstm.mem_to_out.0.weight"]))
            sd_names.append((L.P_BMO, ["lstm.mem_to_out.0.bias"]))
        for l in range(dims.L):
            base = L.P_LSTM_COUNT + l * L.LEVEL_STRIDE
            p = f"procs.{l}."
            a = p + "global_agg."
            sd_names += [
                (base + L.P_WUV, [p + "importance_mlp.0.weight", a + "proj_in.weight"]),
                (base + L.P_B1, [p + "importance_mlp.0.bias"]), (base + L.P_W2, [p + "importance_mlp.2.weight"]),
                (base + L.P_B2, [p + "importance_mlp.2.bias"]), (base + L.P_BIN, [a + "proj_in.bias"]),
                (base + L.P_SPECIAL, [a + "special_token"]), (base + L.P_NFW, [a + "transformer.decoder.norm.weight"]),
                (base + L.P_NFB, [a + "transformer.decoder.norm.bias"]), (base + L.P_WC, [p + "classification_layer.weight"]),
                (base + L.P_BC, [p + "classification_layer.bias"]),
            ]
            if not dims.lstm and dims.hier_ctx and l > 0:      # RNN variant; procs.0.hctx_mlp is never evaluated (paths.py:102)
                sd_names += [(base + L.P_WH0, [p + "hctx_mlp.0.weight"]), (base + L.P_BH0, [p + "hctx_mlp.0.bias"]),
                             (base + L.P_WH2, [p + "hctx_mlp.2.weight"]), (base + L.P_BH2, [p + "hctx_mlp.2.bias"])]
            for j in range(dims.layers):
                q = a + f"transformer.decoder.layers.{j}."
                lb = base + L.P_LEVEL_HEAD_COUNT + j * L.P_LAYER_COUNT
                sd_names += [
                    (lb + L.P_WQKV, [q + "self_attn.in_proj_weight"]), (lb + L.P_BQKV, [q + "self_attn.in_proj_bias"]),
                    (lb + L.P_WO, [q + "self_attn.out_proj.weight"]), (lb + L.P_BO, [q + "self_attn.out_proj.bias"]),
                    (lb + L.P_BCROSS, [q + "multihead_attn.out_proj.bias"]),
                    (lb + L.P_FW1, [q + "linear1.weight"]), (lb + L.P_FB1, [q + "linear1.bias"]),
                    (lb + L.P_FW2, [q + "linear2.weight"]), (lb + L.P_FB2, [q + "linear2.bias"]),
                    (lb + L.P_N1W, [q + "norm1.weight"]), (lb + L.P_N1B, [q + "norm1.bias"]),
                    (lb + L.P_N2W, [q + "norm2.weight"]), (lb + L.P_N2B, [q + "norm2.bias"]),
                    (lb + L.P_N3W, [q + "norm3.weight"]), (lb + L.P_N3B, [q + "norm3.bias"]),
                ]
        named = dict(model.named_parameters())
        self.offsets = (C.c_int64 * L.NUM_OFFSETS)(*([-1] * L.NUM_OFFSETS))
        self.entries: List[Tuple[str, torch.nn.Parameter, int, int]] = []     # (name, param, offset, numel)
        off = 0
        for idx, names in sd_names:
            self.offsets[idx] = off
            for n in names:
                prm = named[n]
                self.entries.append((n, prm, off, prm.numel()))
                off += prm.numel()
            off = (off + 63) // 64 * 64                                      # 256-byte alignment of every tensor group
        self.total = off
        self.live = {n for n, _, _, _ in self.entries}
        self.dead = [(n, prm) for n, prm in named.items() if n not in self.live]
        dev = self.entries[0][1].device
        self.device = dev
        # tail of the flat buffer: parameters the hot path never reads (the encoder / cross-attention weights of
        # nn.Transformer).  They have no gradient storage; the fused AdamW only weight-decays them.
        self.dead_entries: List[Tuple[str, torch.nn.Parameter, int, int]] = []
        for n, prm in self.dead:
            self.dead_entries.append((n, prm, off, prm.numel()))
            off += prm.numel()
        self.total_all = (off + 63) // 64 * 64
        self.flat = torch.zeros(self.total_all, dtype=torch.float32, device=dev)
        self.gflat = torch.zeros(self.total, dtype=torch.float32, device=dev)
        self.views: Dict[str, torch.Tensor] = {}
        self.gviews: Dict[str, torch.Tensor] = {}
        with torch.no_grad():
            for n, prm, o, k in self.entries + self.dead_entries:
                v = self.flat[o:o + k].view(prm.shape)
                v.copy_(prm.data)
                prm.data = v
                self.views[n] = v
                if o < self.total:
                    self.gviews[n] = self.gflat[o:o + k].view(prm.shape)
        # parameters whose gradient the reference leaves at None: classifiers of levels 0..L-2 (utils.py:263 only the
        # last level's logits reach the loss)
        self.none_grad = {f"procs.{l}.classification_layer.{w}" for l in range(dims.L - 1) for w in ("weight", "bias")}
        self._dead_zero: Dict[str, torch.Tensor] = {}

    def valid(self) -> bool:
        """False if some parameter was re-allocated (e.g. model.to(other_device)); the layout must then be rebuilt."""
        # every parameter is re-allocated together by the operations that invalidate the layout (.to(), .float(), load with
        # assign=True): probing the first, middle and last tensor keeps this off the per-step critical path
        ents = self.entries
        probe = (ents[0], ents[len(ents) // 2], ents[-1]) + ((self.dead_entries[-1],) if self.dead_entries else ())
        return all(prm.data_ptr() == self.views[n].data_ptr() for n, prm, _, _ in probe)

    def prepare_grads(self) -> None:
        """Make gflat the accumulation target of every live parameter's .grad (autograd semantics: grad += new)."""
        all_none = all(prm.grad is None for _, prm, _, _ in self.entries)
        if all_none:
            self.gflat.zero_()
            return
        for n, prm, o, k in self.entries:
            gv = self.gviews[n]
            if prm.grad is None:
                gv.zero_()
            elif prm.grad.data_ptr() != gv.data_ptr():
                gv.copy_(prm.grad)

    def attach_grads(self) -> None:
        for n, prm, _, _ in self.entries:
            if n in self.none_grad and prm.grad is None:
                continue
            if prm.requires_grad:
                prm.grad = self.gviews[n]
        for n, prm in self.dead:                       # the reference produces exact zeros for the dead encoder / cross-attn
            if prm.requires_grad and prm.grad is None and ".transformer." in n:
                z = self._dead_zero.get(n)             # (never-evaluated modules such as procs.0.hctx_mlp keep grad = None)
                if z is None:
                    z = self._dead_zero[n] = torch.zeros_like(prm)     # never written by the engine: allocated once
                prm.grad = z


# ------------------------------------------------------------------------------------------------------------------
class PyramidStore:
    """Feature pyramids resident in HBM, in the reference's on-disk grid format ([W, H, D] float32, zero vector =
    background; preprocess/loader.py:14-18), plus one occupancy bitmap per grid.  Loaded once per slide; a training
    step then only names slides by index."""

    def __init__(self, num_levels: int, dim: int, device, dtype: torch.dtype = torch.float32):
        """dtype=torch.bfloat16 keeps the grids as bfloat16 (half the HBM footprint and gather traffic; the features are rounded
        once, at load -- a model fed by such a store equals the reference fed with the rounded features); the occupancy bitmaps
        are always built from the float32 grids."""
        assert dtype in (torch.float32, torch.bfloat16)
        self.dtype = dtype
        self._cap_cache = {}
        self.L, self.D, self.device = num_levels, dim, torch.device(device)
        self.grids: List[List[torch.Tensor]] = []
        self.bits: List[torch.Tensor] = []
        self.counts: List[torch.Tensor] = []
        self.shapes: List[List[Tuple[int, int]]] = []
        self._tables = None

    def __len__(self) -> int:
        return len(self.grids)

    def add_slide(self, grids: Sequence[torch.Tensor]) -> int:
        assert len(grids) >= self.L, f"slide has {len(grids)} levels, store needs {self.L}"
        dev_grids, shapes = [], []
        lib = L.lib()
        for g in grids[:self.L]:
            assert g.dim() == 3 and g.shape[2] == self.D, f"grid must be [W, H, {self.D}], got {tuple(g.shape)}"
            gd = g.to(self.device, dtype=torch.float32).contiguous()
            w, h = int(gd.shape[0]), int(gd.shape[1])
            bits = torch.zeros((w * h + 31) // 32, dtype=torch.int32, device=self.device)
            cnt = torch.zeros(1, dtype=torch.int32, device=self.device)
            L.check(lib.paths_build_occupancy(gd.data_ptr(), w * h, self.D, bits.data_ptr(), cnt.data_ptr(), _stream()),
                    "build_occupancy")
            if self.dtype == torch.bfloat16:
                torch.cuda.current_stream(self.device).synchronize()      # the bitmap kernel has read the float32 grid
                gd = gd.to(torch.bfloat16)
            dev_grids.append(gd); shapes.append((w, h)); self.bits.append(bits); self.counts.append(cnt)
        self.grids.append(dev_grids); self.shapes.append(shapes)
        self._tables = None
        return len(self.grids) - 1

    def tables(self):
        if self._tables is None:
            ptrs = torch.tensor([g.data_ptr() for gs in self.grids for g in gs], dtype=torch.int64)
            gw = torch.tensor([s[0] for ss in self.shapes for s in ss], dtype=torch.int32)
            gh = torch.tensor([s[1] for ss in self.shapes for s in ss], dtype=torch.int32)
            sizes = torch.tensor([b.numel() for b in self.bits], dtype=torch.int64)
            off = torch.cumsum(sizes, 0) - sizes
            dev = self.device
            t = {
                "grid_ptr": ptrs.to(dev), "gw": gw.to(dev), "gh": gh.to(dev), "occ": torch.cat(self.bits),
                "occ_off": off.to(dev), "nonbg": torch.cat(self.counts),
            }
            st = L.Store()
            st.grid_ptr, st.gw, st.gh = t["grid_ptr"].data_ptr(), t["gw"].data_ptr(), t["gh"].data_ptr()
            st.occ, st.occ_off, st.nonbg = t["occ"].data_ptr(), t["occ_off"].data_ptr(), t["nonbg"].data_ptr()
            st.bf16 = 1 if self.dtype == torch.bfloat16 else 0
            self._tables = (t, st)
        return self._tables

    def capacities(self, slide_ids: Sequence[int], top_k: Sequence[int], worst_case: bool = False) -> Tuple[List[int], List[int]]:
        """Row capacity and largest bag per level for a batch (bounds of slide.py:294-331: at most 4 children per kept
        row and never more than the grid has cells).  worst_case sizes for the no-child fallback (slide.py:336-352)."""
        key = (tuple(slide_ids), tuple(top_k), bool(worst_case))      # the same batches recur every epoch
        hit = self._cap_cache.get(key)
        if hit is not None:
            return list(hit[0]), list(hit[1])
        caps, maxseg = [0] * self.L, [1] * self.L
        for sid in slide_ids:
            n = self.shapes[sid][0][0] * self.shapes[sid][0][1]
            for l in range(self.L):
                cells = self.shapes[sid][l][0] * self.shapes[sid][l][1]
                if l > 0:
                    k = top_k[l - 1]
                    kept = n if k < 0 else min(k, n)
                    n = cells if worst_case else min(4 * kept, cells)
                caps[l] += n
                maxseg[l] = max(maxseg[l], n)
        if len(self._cap_cache) > 4096:
            self._cap_cache.clear()
        self._cap_cache[key] = (tuple(caps), tuple(maxseg))
        return caps, maxseg


# ------------------------------------------------------------------------------------------------------------------
class Engine:
    """One paths_engine_t: workspace + row tables for batches of exactly B slides with the given capacities."""

    def __init__(self, dims: L.Dims, B: int, caps: Sequence[int], maxseg: Sequence[int], device):
        self.lib = L.lib()
        if not torch.cuda.is_available():
            raise RuntimeError("paths_b200 needs a CUDA device (sm_100a); there is no CPU fallback")
        self.dims, self.B, self.caps, self.maxseg = dims, B, list(caps), list(maxseg)
        self.device = torch.device(device)
        n = dims.L
        self._caps = (C.c_int32 * n)(*[int(c) for c in caps])
        self._maxseg = (C.c_int32 * n)(*[int(c) for c in maxseg])
        need = self.lib.paths_engine_workspace_bytes(C.byref(dims), B, self._caps, self._maxseg)
        if need == 0:
            L.check(-1, "engine_workspace_bytes")
        self.ws = torch.empty(need + 256, dtype=torch.uint8, device=self.device)
        self._ws_off = (-self.ws.data_ptr()) % 256
        self.handle = C.c_void_p()
        L.check(self.lib.paths_engine_create(C.byref(self.handle), C.byref(dims), B, self._caps, self._maxseg,
                                             self.ws.data_ptr() + self._ws_off, need), "engine_create")
        self._buf_cache: Dict[Tuple[int, str], torch.Tensor] = {}
        self._keep = []

    def __del__(self):
        try:
            if getattr(self, "handle", None) and self.handle.value:
                self.lib.paths_engine_destroy(self.handle)
                self.handle = C.c_void_p()
        except Exception:
            pass

    def fits(self, B: int, caps: Sequence[int], maxseg: Sequence[int]) -> bool:
        return B == self.B and all(a <= b for a, b in zip(caps, self.caps)) and all(a <= b for a, b in zip(maxseg, self.maxseg))

    def bind_params(self, layout: ParamLayout, div: torch.Tensor) -> None:
        self._keep = [layout, div]
        L.check(self.lib.paths_engine_bind_params(self.handle, layout.flat.data_ptr(), layout.gflat.data_ptr(), layout.total,
                                                  layout.offsets, div.data_ptr()), "bind_params")

    def bind_store(self, store: PyramidStore) -> None:
        t, st = store.tables()
        self._store = (store, t, st)
        L.check(self.lib.paths_engine_bind_store(self.handle, C.byref(st)), "bind_store")

    def buffer(self, level: int, name: str) -> torch.Tensor:
        key = (level, name)
        if key not in self._buf_cache:
            ptr, nbytes = C.c_void_p(), C.c_size_t()
            L.check(self.lib.paths_engine_buffer(self.handle, level, L.BUF[name], C.byref(ptr), C.byref(nbytes)), "engine_buffer")
            off = ptr.value - self.ws.data_ptr()
            raw = self.ws[off:off + nbytes.value]
            self._buf_cache[key] = raw.view(torch.int32 if name in L.INT_BUFS else torch.float32)
        return self._buf_cache[key]

    # -- thin call-throughs -------------------------------------------------------------------------------------
    def begin(self, slide_ids: torch.Tensor, fts_padded: Optional[torch.Tensor] = None) -> None:
        if fts_padded is None:
            L.check(self.lib.paths_engine_begin(self.handle, slide_ids.data_ptr(), _stream()), "engine_begin")
        else:
            assert fts_padded.is_cuda and fts_padded.dtype == torch.float32 and fts_padded.is_contiguous()
            L.check(self.lib.paths_engine_begin_padded(self.handle, slide_ids.data_ptr(), fts_padded.data_ptr(),
                                                       int(fts_padded.shape[1]), _stream()), "engine_begin_padded")

    def refresh_weights(self) -> None:
        L.check(self.lib.paths_engine_refresh_weights(self.handle, _stream()), "refresh_weights")

    def level_forward(self, level: int) -> None:
        L.check(self.lib.paths_engine_level_forward(self.handle, level, _stream()), "level_forward")

    def select(self, level: int) -> None:
        L.check(self.lib.paths_engine_select(self.handle, level, _stream()), "select")

    def loss(self, ybin: torch.Tensor, cens: Optional[torch.Tensor], scale: float) -> None:
        L.check(self.lib.paths_engine_loss(self.handle, ybin.data_ptr(), cens.data_ptr() if cens is not None else None,
                                           scale, _stream()), "loss")

    def level_backward(self, level: int, use_dlogits: bool) -> None:
        L.check(self.lib.paths_engine_level_backward(self.handle, level, int(use_dlogits), _stream()), "level_backward")

    def forward(self, slide_ids: torch.Tensor, ybin: Optional[torch.Tensor], cens: Optional[torch.Tensor], scale: float,
                with_loss: bool = True, fts_padded: Optional[torch.Tensor] = None, fts_packed: Optional[torch.Tensor] = None) -> None:
        yb = ybin.data_ptr() if ybin is not None else None
        cs = cens.data_ptr() if cens is not None else None
        if fts_packed is not None:
            assert fts_packed.is_cuda and fts_packed.dtype == torch.float32 and fts_packed.is_contiguous() and fts_packed.dim() == 2
            L.check(self.lib.paths_engine_forward_packed(self.handle, slide_ids.data_ptr(), fts_packed.data_ptr(),
                                                         int(fts_packed.shape[0]), yb, cs, scale, int(with_loss), _stream()),
                    "engine_forward_packed")
        elif fts_padded is None:
            L.check(self.lib.paths_engine_forward(self.handle, slide_ids.data_ptr(), yb, cs, scale, int(with_loss), _stream()),
                    "engine_forward")
        else:
            assert fts_padded.is_cuda and fts_padded.dtype == torch.float32 and fts_padded.is_contiguous()
            L.check(self.lib.paths_engine_forward_padded(self.handle, slide_ids.data_ptr(), fts_padded.data_ptr(),
                                                         int(fts_padded.shape[1]), yb, cs, scale, int(with_loss), _stream()),
                    "engine_forward_padded")

    def backward(self, zero_grads: bool = False) -> None:
        L.check(self.lib.paths_engine_backward(self.handle, int(zero_grads), _stream()), "engine_backward")

    def wait_grads(self, level: int, stream: Optional[torch.cuda.Stream] = None) -> None:
        """Make `stream` (default: the current one) wait until the latest backward pass has finished procs[level]'s gradients."""
        s = stream.cuda_stream if stream is not None else _stream()
        L.check(self.lib.paths_engine_wait_grads(self.handle, level, s), "wait_grads")

    def set_precision(self, precision: int) -> None:
        L.check(self.lib.paths_engine_set_precision(self.handle, precision), "set_precision")
        self.dims.precision = precision

    def set_dropout(self, p: float, seed: int) -> None:
        """Dropout of the aggregators for the next forward pass (and its backward pass); p = 0 switches it off."""
        L.check(self.lib.paths_engine_set_dropout(self.handle, float(p), int(seed) & 0xFFFFFFFFFFFFFFFF), "set_dropout")

    def profile(self, on: bool) -> None:
        L.check(self.lib.paths_engine_profile(self.handle, int(on)), "profile")

    def profile_read(self, tag: int):
        ms, n = C.c_double(), C.c_int64()
        L.check(self.lib.paths_engine_profile_read(self.handle, tag, C.byref(ms), C.byref(n)), "profile_read")
        return ms.value, n.value

    def launch_count(self) -> int:
        return int(self.lib.paths_engine_launch_count(self.handle))

    def counts(self) -> List[int]:
        """Live rows per level (device -> host sync; tests / reporting only)."""
        return [int(self.buffer(l, "COUNT").item()) for l in range(self.dims.L)]

    def overflowed(self) -> Tuple[bool, int]:
        """Overflow flag read straight from the device (full synchronisation; tests / reporting)."""
        o = self.buffer(0, "OVERFLOW").tolist()
        return bool(o[0]), int(o[1])

    def overflow_poll(self) -> Tuple[bool, int]:
        """Did a slide's bag exceed the capacities this engine was sized for (no-child fallback, slide.py:336-352)?  Waits only
        until the last select of the latest forward pass has run on the device, not for the end of the pass."""
        flag, need = C.c_int32(), C.c_int32()
        L.check(self.lib.paths_engine_overflow_poll(self.handle, C.byref(flag), C.byref(need)), "overflow_poll")
        return bool(flag.value), int(need.value)
