"""Reference-facing entry points of the hot path (reference utils.py): `inference_end2end` and helpers."""
import os
import pickle
from os.path import join
from typing import Dict, Sequence

import torch

device = torch.device("cuda") if torch.cuda.is_available() else torch.device("cpu")


def todevice(x, dev):
    """utils.py:334-341."""
    if hasattr(x, "to"):
        return x.to(dev)
    if isinstance(x, (list, tuple)):
        return [todevice(i, dev) for i in x]
    return x


def inference_end2end(num_levels: int, keep_patches: Sequence[int], model, base_power: float, batch: Dict, task: str):
    """Drop-in for the reference's utils.inference_end2end (utils.py:228-279): runs all `num_levels` magnification levels
    of the PATHS model on the collated batch -- LSTM step, importance, top-K selection, child expansion, aggregation --
    and returns (hazards, loss) for task "survival" or (logits, loss) for "subtype_classification".  `loss` carries an
    autograd node: loss.backward() fills the .grad of the model's parameters.  Nothing is computed on the CPU."""
    assert num_levels == model.train_config.num_levels, "num_levels must match the model's configuration"
    assert task == model.train_config.task
    keep = [keep_patches] * (num_levels - 1) if isinstance(keep_patches, int) else list(keep_patches)
    return model.runtime().end2end(batch, keep)


def save_state(root_path: str, model, train_stats):
    """utils.py:169-178: model.pt = state_dict (reference key names, dead encoder tensors included) + pickled train stats."""
    torch.save({k: v.detach().cpu().clone() for k, v in model.state_dict().items()}, join(root_path, "model.pt"))
    with open(join(root_path, "train_stats.pkl"), "wb") as file:
        pickle.dump(train_stats, file)


def load_state(root_path: str, model, map_location=device) -> Dict:
    """utils.py:181-198.  Parameters are views into the flat buffer, so loading copies in place and the CUDA engines see
    the new weights on their next step."""
    model_path = join(root_path, "model.pt")
    train_stats_path = join(root_path, "train_stats.pkl")
    if not os.path.isfile(model_path):
        print(f"{model_path} not found, not loading model state!")
    else:
        model.load_state_dict(torch.load(model_path, map_location=map_location))
    if not os.path.isfile(train_stats_path):
        print("No train stats found, assuming first run")
        return {"epoch": 1}
    with open(train_stats_path, "rb") as file:
        return pickle.load(file)
