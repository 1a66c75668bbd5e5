"""Eval-side sparsity: `calculate_sparsity` of /root/reference/ResNet/testing_suite/eval_model.py:7-20
with the per-parameter `torch.sum(torch.abs(param) < threshold).item()` replaced by ONE fused count kernel per
parameter (`wt_count_f32`: count_nonzero and |p| < eps from a single read).  Same signature and return value."""
import torch

from . import engine


def calculate_sparsity(model, threshold=1e-6):
    """eval_model.py:7-20 -- fraction of |p| < threshold over the parameters with dim() > 1"""
    dev = engine.cuda_device()
    total_params = 0
    near_zero_params = 0
    for param in model.parameters():
        if param.dim() > 1:  # only weights, not biases (eval_model.py:12)
            total_params += param.numel()
            x = param.detach()
            if x.dtype != torch.float32:
                x = x.float()
            if not x.is_cuda:
                x = x.to(dev)
            near_zero_params += engine.count_nonzero_and_below(x, threshold)[1]
    if total_params == 0:
        return 0.0
    return near_zero_params / total_params
