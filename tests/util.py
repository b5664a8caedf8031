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


def record(name, value):
    """Append a measured error to gpurun_out/parity.jsonl (evidence of the margins;
    travels back from the GPU box).  Never fails a test."""
    import json
    import os
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out = os.path.join(root, 'gpurun_out')
    try:
        os.makedirs(out, exist_ok=True)
        with open(os.path.join(out, 'parity.jsonl'), 'a') as f:
            f.write(json.dumps({'name': name, 'value': float(value)}) + '\n')
    except OSError:
        pass
    return value
