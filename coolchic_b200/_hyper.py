"""Weight injection used by the hypernetwork path (reference: hypernet/common.py:127-182):
parameters are REPLACED by non-leaf tensors so that autograd flows into the hypernetwork."""

import torch
from torch import nn


def select_param_from_name(obj, name: str):
    parent = None
    for sub in name.split("."):
        parent = obj
        obj = obj[int(sub)] if sub.isdigit() else getattr(obj, sub)
    assert isinstance(obj, torch.Tensor), f"Selected object is not a tensor. Got {type(obj)} instead."
    assert isinstance(parent, nn.Module), "Selected object has no parent."
    return obj, parent


def set_hypernet_weights(obj, all_weights) -> None:
    for name, weight in all_weights.items():
        _, parent = select_param_from_name(obj, name)
        leaf = name.rsplit(".", 1)[-1]
        if leaf not in ("weight", "bias", "original"):
            raise ValueError(f"Unknown parameter name {name}")
        delattr(parent, leaf)
        setattr(parent, leaf, weight)
