"""model/utils.py of the reference: layer cloning and activation lookup for the DETR-variant layers."""
import copy

import torch.nn as nn


def _get_activation_fn(activation):
    """Only "relu" has a fused epilogue in the CUDA library; the reference's other names are rejected loudly."""
    if activation == "relu":
        return "relu"
    raise RuntimeError(f"activation should be relu (fused sm_100a epilogue), not {activation}.")


def _get_clones(module, N):
    return nn.ModuleList([copy.deepcopy(module) for _ in range(N)])
