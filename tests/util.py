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


def relu_decision_flips(x0, vgg_weights_cpu, n_convs):
    """ReLU decisions of VGG conv 1..n_convs where the CUDA forward and the fp32 oracle disagree, for the image x0
    (CPU tensor).  Returns (count, worst): ``worst`` = the largest |oracle value| among the flipped elements relative
    to the layer's max |value| -- a legitimate flip sits within the forward rounding error of zero.

    Why tests need this: the gradient of the loss network is DISCONTINUOUS in its pre-activations.  An element whose
    oracle value is 1e-6 from zero flips its ReLU (or a max-pool argmax) under any change of summation order -- the
    fp32 oracle itself is 4.9e-4 away from its own fp64 evaluation on the 136x200 case -- and every flipped element of
    a deep, small feature map costs a whole receptive field of the input gradient (DESIGN.md "Precision")."""
    from oracle import oracle
    from styletransfer_b200 import ops, vgg_loss
    names = [f'Conv2d_{k}' for k in range(1, n_convs + 1)]
    ref = oracle.vgg_layer_outputs(x0, vgg_weights_cpu, names)
    prefix = vgg_loss.VGGPrefix([(a.cuda(), b.cuda()) for a, b in vgg_weights_cpu])
    _, taps = prefix.run(x0.cuda(), {('C', k) for k in range(1, n_convs + 1)})
    count, worst = 0, 0.0
    for k in range(1, n_convs + 1):
        got, r = ops.to_nchw(taps[k]).cpu(), ref[f'Conv2d_{k}']
        flipped = (got > 0) != (r > 0)
        n = int(flipped.sum())
        if n:
            count += n
            worst = max(worst, float(r.abs()[flipped].max() / r.abs().max()))
    return count, worst
