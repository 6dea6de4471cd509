"""Parameter container mirroring the reference's TransformerAggregator (model/aggregator.py:8-35): same constructor
signature, same sub-module names, hence the same state_dict keys (a reference model.pt loads unchanged, including the
dead encoder / cross-attention weights).  The arithmetic does not live here: the level engine (csrc/engine.cu) runs
proj_in + positional encoding + the varlen decoder directly from the flat parameter buffer."""
import torch
from torch import nn


class TransformerAggregator(nn.Module):
    def __init__(self, input_dim: int, model_dim: int, output_dim: int, nhead: int, layers: int, dropout: float):
        super().__init__()
        self.dim = model_dim
        self.nhead = nhead
        self.layers = layers
        self.proj_in = nn.Linear(input_dim, model_dim)
        self.proj_out = nn.Identity()
        # nn.Transformer is instantiated only as the owner of the parameters (identical names / shapes / init order as the
        # reference); its forward is never called.
        self.transformer = nn.Transformer(model_dim, nhead=nhead, num_encoder_layers=layers, num_decoder_layers=layers,
                                          dim_feedforward=model_dim * 4, dropout=dropout, batch_first=True)
        self.special_token = nn.Parameter(torch.randn(model_dim, ), requires_grad=True)

    def forward(self, seq1, seq2, lengths1, lengths2):
        raise NotImplementedError(
            "paths_b200 runs the aggregator inside the fused level engine; call RecursiveModel.forward(depth, PatchBatch) "
            "or utils.inference_end2end instead of TransformerAggregator.forward")
