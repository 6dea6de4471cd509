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
        raise NotImplementedError("paths_b200 runs the LSTM cell inside the fused level engine (RecursiveModel.forward)")


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
