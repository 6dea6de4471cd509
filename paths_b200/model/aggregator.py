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

        self._proc = None                                   # weak reference to the owning PATHSProcessor (set by it)

    def pos_encode_1d(self, xs, project=True):
        raise NotImplementedError("paths_b200 applies proj_in and the positional encoding inside the level engine (imp_tok_fwd)")

    pos_encode_2d = pos_encode_1d

    def forward(self, seq1, seq2, lengths1, lengths2):
        """model/aggregator.py:58-76: seq2 [B, N, d] padded tokens (already projected and position-encoded), lengths2 [B] -> the
        decoder's output at the prepended special token, [B, d].  The encoder input seq1 must be the empty sequence the model
        passes (paths.py:113: [B, 0, d]) -- the reference's own use; a non-empty condition sequence is not built.  Runs the
        level's aggregator of the engine (`paths_engine_aggregate`); values only (no autograd through this method)."""
        if seq1 is not None and seq1.shape[1] != 0:
            raise NotImplementedError("paths_b200: TransformerAggregator with a non-empty condition sequence (unused by the reference's models)")
        proc = self._proc() if self._proc is not None else None
        owner = proc._owner() if proc is not None and proc._owner is not None else None
        if owner is None:
            raise RuntimeError("TransformerAggregator.forward needs the aggregator to belong to a RecursiveModel (weights live in its flat buffer)")
        return owner.runtime().aggregate(proc.depth, seq2, lengths2)
