"""Mirror of the reference's model/interface.py: LSTMCell, abstract Processor, RecursiveModel -- same constructor
signatures, attribute names and state_dict keys; forward runs on the B200 level engine."""
from abc import ABC, abstractmethod
from typing import Callable, Dict, Tuple

import torch
from torch import nn


class LSTMCell(nn.Module):
    """Parameters of the hand-written LSTM cell (reference model/interface.py:10-29).  One step of it, batched over all
    patches of a level, is the gate contraction of csrc/engine.cu:level_forward."""

    def __init__(self, input_dim: int, output_dim: int, hidden_dim: int):
        super().__init__()
        self.xdim, self.hdim, self.cdim = input_dim, output_dim, hidden_dim
        self.forget_gate = nn.Sequential(nn.Linear(self.xdim + self.hdim, self.cdim), nn.Sigmoid())
        self.remember_gate = nn.Sequential(nn.Linear(self.xdim + self.hdim, self.cdim), nn.Sigmoid())
        self.remember_map = nn.Sequential(nn.Linear(self.xdim + self.hdim, self.cdim), nn.Tanh())
        self.out_select_gate = nn.Sequential(nn.Linear(self.hdim + self.xdim, self.hdim), nn.Sigmoid())
        self.mem_to_out = nn.Sequential(nn.Linear(self.cdim, self.hdim), nn.Tanh())

    def forward(self, xs, hs, cs):
        """One LSTM step (reference model/interface.py:31-58): xs, hs [..., D], cs [..., C] -> (hs', cs').  Runs the stand-alone
        per-op entry `paths_lstm_cell_fwd` (the gate contraction and the cell's row kernels of the level engine); values only --
        gradients flow through the whole-step op of utils.inference_end2end, not through this convenience method."""
        import ctypes as C
        from .. import _lib as L
        from ..runtime import default_precision
        assert xs.shape[-1] == self.xdim and hs.shape[-1] == self.hdim and cs.shape[-1] == self.cdim      # interface.py:44-47
        assert xs.shape[:-1] == hs.shape[:-1] == cs.shape[:-1]
        if not xs.is_cuda:
            raise RuntimeError("paths_b200: LSTMCell.forward needs CUDA tensors; there is no CPU path")
        lead = xs.shape[:-1]
        x = xs.detach().reshape(-1, self.xdim).float().contiguous()
        h = hs.detach().reshape(-1, self.hdim).float().contiguous()
        c = cs.detach().reshape(-1, self.cdim).float().contiguous()
        gates = (self.forget_gate, self.remember_gate, self.remember_map, self.out_select_gate)
        with torch.no_grad():
            Wg = torch.cat([g[0].weight for g in gates], 0).float().contiguous()
            bg = torch.cat([g[0].bias for g in gates], 0).float().contiguous()
            Wmo = self.mem_to_out[0].weight.detach().float().contiguous()
            bmo = self.mem_to_out[0].bias.detach().float().contiguous()
        M = x.shape[0]
        ho = torch.empty(M, self.hdim, device=x.device); co = torch.empty(M, self.cdim, device=x.device)
        if M:
            L.check(L.lib().paths_lstm_cell_fwd(default_precision(), x.data_ptr(), h.data_ptr(), c.data_ptr(), Wg.data_ptr(), bg.data_ptr(),
                                                Wmo.data_ptr(), bmo.data_ptr(), ho.data_ptr(), co.data_ptr(), M, self.xdim, self.cdim,
                                                torch.cuda.current_stream().cuda_stream), "lstm_cell_fwd")
        return ho.view(*lead, self.hdim), co.view(*lead, self.cdim)


class Processor(ABC):
    """Abstract single-depth slide processor (reference model/interface.py:61-80)."""

    @abstractmethod
    def process(self, data) -> Dict:
        raise NotImplementedError

    @abstractmethod
    def ctx_dim(self) -> Tuple[int, int]:
        raise NotImplementedError


class RecursiveModel(nn.Module):
    """N processors + one shared LSTM (reference model/interface.py:83-99)."""

    def __init__(self, processor_constructor: Callable, config_, train_config, **kwargs):
        super().__init__()
        self.procs = nn.ModuleList([processor_constructor(config_, train_config, depth=i, **kwargs)
                                    for i in range(train_config.num_levels)])
        from ..config import PATHSProcessorConfig
        if isinstance(config_, PATHSProcessorConfig) and config_.lstm:
            self.lstm = LSTMCell(config_.patch_embed_dim, config_.patch_embed_dim, config_.hierarchical_ctx_mlp_hidden_dim)
            self.use_lstm = True
        else:
            self.use_lstm = False
        import weakref
        for proc in self.procs:
            proc._owner = weakref.ref(self)
        self.model_config = config_
        self.train_config = train_config
        self._runtime = None

    def runtime(self):
        """The CUDA runtime bound to this model's parameters (created on first use, after .to(device))."""
        from ..runtime import ModelRuntime
        if self._runtime is None or not self._runtime.layout.valid():
            self._runtime = ModelRuntime(self)
        return self._runtime

    def forward(self, depth, *args, **kwargs):
        if self.use_lstm:
            kwargs["lstm"] = self.lstm
        return self.procs[depth].process(*args, **kwargs)
