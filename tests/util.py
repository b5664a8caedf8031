"""Shared helpers for the parity tests."""
import torch


def seeded(shape, seed, scale=1.0, device='cpu'):
    t = torch.randn(*shape, generator=torch.Generator().manual_seed(seed)) * scale
    return t.to(device)


def rel_l2(a, b):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return float((a - b).norm() / (b.norm() + 1e-30))


def nhwc(split, C=None):
    """Split -> NCHW fp32 torch tensor (via torch ops, for checking)."""
    v = split.float().permute(0, 3, 1, 2).contiguous()
    return v if C is None else v[:, :C]


def fp32_reference_mode():
    """Make torch's own GPU ops a trustworthy fp32 reference (no TF32)."""
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
