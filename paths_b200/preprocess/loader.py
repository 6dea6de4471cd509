"""Mirror of the reference's preprocess/loader.py: where the preprocessed feature grids live on disk.
File format (unchanged): `{root}/{slide_id}_{power:.3f}.pt`, a bare float tensor [W, H, D] with all-zero vectors for
background cells (reference preprocess/preprocess.py:89,105,134)."""
from os import path

import torch

root_dir = None


def set_preprocess_dir(root: str):
    global root_dir
    root_dir = root


def load(slide_id: str, power: float) -> torch.Tensor:
    assert root_dir is not None, "set_preprocess_dir must be called before loading preprocessed slides"
    return torch.load(path.join(root_dir, f"{slide_id}_{power:.3f}.pt"))
