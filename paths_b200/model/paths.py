"""Mirror of the reference's PATHSProcessor (model/paths.py:12-151): same constructor, sub-module names and outputs.
`process` runs one level of the B200 engine on a padded PatchBatch and returns the reference's dict
{"logits", "ctx_slide", "ctx_patch", "importance"} (padded [B, N, ...], importance 0 at padding, utils.py:113)."""
from typing import Dict, Tuple

from torch import nn

from .aggregator import TransformerAggregator
from .interface import Processor


class PATHSProcessor(nn.Module, Processor):
    def __init__(self, config, train_config, depth: int):
        super().__init__()
        self.depth = depth
        num_logits = train_config.nbins if train_config.task == "survival" else len(train_config.filter_to_subtypes)
        self.config = config
        self.train_config = train_config
        self.dim = config.patch_embed_dim
        self.slide_ctx_dim = config.trans_dim
        if self.config.slide_ctx_mode == "concat":
            self.classification_layer = nn.Linear(self.slide_ctx_dim * (depth + 1), num_logits)
        else:
            self.classification_layer = nn.Linear(self.slide_ctx_dim, num_logits)
        self.importance_mlp = nn.Sequential(
            nn.Linear(self.dim, config.importance_mlp_hidden_dim),
            nn.ReLU(),
            nn.Linear(config.importance_mlp_hidden_dim, 1),
        )
        if config.lstm:
            self.hdim = config.hierarchical_ctx_mlp_hidden_dim
        else:
            self.hctx_mlp = nn.Sequential(
                nn.Linear(self.dim, config.hierarchical_ctx_mlp_hidden_dim),
                nn.ReLU(),
                nn.Linear(config.hierarchical_ctx_mlp_hidden_dim, self.dim),
            )
        self.global_agg = TransformerAggregator(input_dim=self.dim, model_dim=config.trans_dim, output_dim=self.dim,
                                                nhead=config.trans_heads, layers=config.trans_layers, dropout=config.dropout)
        self._owner = None
        import weakref
        self.global_agg._proc = weakref.ref(self)

    def process(self, data, lstm=None) -> Dict:
        if self.config.lstm:
            assert lstm is not None
        if self._owner is None:
            raise RuntimeError("PATHSProcessor.process must be reached through RecursiveModel.forward(depth, data)")
        return self._owner().runtime().level_api(self.depth, data)

    def ctx_dim(self) -> Tuple[int, int]:
        if self.config.lstm:
            return self.slide_ctx_dim, self.dim + self.hdim
        return self.slide_ctx_dim, self.dim
