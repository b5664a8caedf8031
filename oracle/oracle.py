"""CPU oracle for the stransfer.network hot path.  TEST INFRASTRUCTURE ONLY.

This file is a plain PyTorch-fp32 (CPU) restatement of the arithmetic that
tupini07/StyleTransfer's ``stransfer/network.py`` performs on its hot path.  It
is the checker for the CUDA kernels in ``styletransfer_b200/csrc``; nothing in
the product path imports it (only ``tests/``, ``__graft_entry__.smoke()`` and
``bench.py``'s ``cpu_baseline`` / ``--impl reference`` legs may).

Pinning: the reference ships no tests or golden vectors, so the
restatement is pinned against the reference ITSELF, imported from
``/root/reference`` under the shims of ``oracle/reference_shim.py`` in the
build container; ``oracle/make_golden.py`` runs both on the same seeded inputs,
asserts they agree, and writes ``tests/golden/*.npz``.  On the GPU box (no
``/root/reference``) the oracle is checked against those committed fixtures.

Every function cites the reference file:line it restates.  The restatement is
*single pass* (one VGG prefix evaluation emitting all taps); the reference
re-runs the prefix once per loss node (network.py:378-398) which is
mathematically identical -- ``make_golden.py`` proves that numerically.
"""
import math

import torch
import torch.nn.functional as F

# VGG-19 "E" prefix actually executed by the reference: the loss
# taps are Conv2d_1..Conv2d_5 = conv1_1, conv1_2, conv2_1, conv2_2, conv3_1
# (network.py:214-232, piece builder :257-314).  (cin, cout, pool_before)
VGG_PREFIX = [(3, 64, False), (64, 64, False), (64, 128, True),
              (128, 128, False), (128, 256, True)]
# torchvision ``vgg19().features`` indices of those five convs
VGG_FEATURE_IDX = [0, 2, 5, 7, 10]
CONTENT_TAP = 3          # 'Conv2d_4'   network.py:214-220
FEATURE_TAP = 3          # 'ReLU_4' = relu(Conv2d_4)  network.py:230-232


# torchvision VGG-19 configuration "E" (all 16 convs): (cin, cout, pool_after) per conv
VGG_CFG_E = [64, 64, 'M', 128, 128, 'M', 256, 256, 256, 256, 'M', 512, 512, 512, 512, 'M', 512, 512, 512, 512, 'M']


def vgg_blocks():
    out, cin = [], 3
    for i, v in enumerate(VGG_CFG_E):
        if v == 'M':
            continue
        out.append((cin, v, i + 1 < len(VGG_CFG_E) and VGG_CFG_E[i + 1] == 'M'))
        cin = v
    return out


def make_vgg_weights(seed=1, n=5):
    """Seeded random-init weights for the first ``n`` VGG-19 convs (the default loss layers execute five; the
    first five draws do not depend on n).

    torchvision's VGG init (kaiming-normal fan_out, zero bias) restated with an
    explicit generator so that no torchvision import is needed on the GPU box.
    Biases get a small seeded non-zero value so the bias path is exercised.
    """
    g = torch.Generator().manual_seed(seed)
    out = []
    for cin, cout, _ in vgg_blocks()[:n]:
        std = math.sqrt(2.0 / (cout * 9))
        w = torch.randn(cout, cin, 3, 3, generator=g) * std
        b = torch.randn(cout, generator=g) * 0.05
        out.append((w, b))
    return out


def vgg_taps(x, weights, upto=5):
    """Pre-ReLU outputs of Conv2d_1..Conv2d_upto (network.py:316-340 run over
    the pieces printed in SURVEY F1; zero padding, ReLU not in place)."""
    taps = []
    for i, ((w, b), (_, _, pool)) in enumerate(zip(weights, VGG_PREFIX)):
        if i >= upto:
            break
        if i > 0:
            x = F.relu(x)
            if pool:
                x = F.max_pool2d(x, 2, 2)
        x = F.conv2d(x, w, b, padding=1)
        taps.append(x)
    return taps


def vgg_layer_outputs(x, weights, names):
    """{name: output of that layer} for layers of vgg19().features named as the reference names them
    (network.py:257-270: 'Conv2d_k' / 'ReLU_k' / 'MaxPool2d_k', k counts the convs); ReLU not in place, zero padding."""
    want, out = set(names), {}
    for k, ((w, b), (_, _, pool)) in enumerate(zip(weights, vgg_blocks()), start=1):
        x = F.conv2d(x, w, b, padding=1)
        if f'Conv2d_{k}' in want:
            out[f'Conv2d_{k}'] = x
        x = F.relu(x)
        if f'ReLU_{k}' in want:
            out[f'ReLU_{k}'] = x
        if pool:
            x = F.max_pool2d(x, 2, 2)
            if f'MaxPool2d_{k}' in want:
                out[f'MaxPool2d_{k}'] = x
        if len(out) == len(want):
            break
    missing = want - set(out)
    assert not missing, f'layers not reached with {len(weights)} conv weights: {sorted(missing)}'
    return out


def generic_closure(x, content_image, style_image, weights, style_layers, content_layers, feature_layers=(),
                    style_weight=100_000.0, content_weight=1.0):
    """StyleNetwork.forward + the Gatys closure (network.py:366-401, :438-454) for ARBITRARY loss-layer lists
    (the reference's piece builder is generic over its class attributes, network.py:214-232, :257-314).
    Returns (total, grad of x, per-node losses [style..., content..., feature...] in network order)."""
    order = {}
    k = 0
    for i, v in enumerate(VGG_CFG_E):
        if v != 'M':
            k += 1
            order[f'Conv2d_{k}'] = 3 * i
            order[f'ReLU_{k}'] = 3 * i + 1
        else:
            order[f'MaxPool2d_{k}'] = 3 * i
    srt = lambda names: sorted(names, key=lambda n: order[n])
    style_layers, content_layers, feature_layers = srt(style_layers), srt(content_layers), srt(feature_layers)
    names = set(style_layers) | set(content_layers) | set(feature_layers)
    with torch.no_grad():
        s_out = vgg_layer_outputs(style_image, weights, style_layers) if style_layers else {}
        c_out = vgg_layer_outputs(content_image, weights, set(content_layers) | set(feature_layers))
    x = x.detach().clone().requires_grad_(True)
    out = vgg_layer_outputs(x, weights, names)
    losses = [style_loss(out[n], gram_matrix(s_out[n])) for n in style_layers]
    losses += [content_loss(out[n], c_out[n]) for n in content_layers]
    losses += [feature_loss(out[n], c_out[n]) for n in feature_layers]
    ns, nc = len(style_layers), len(content_layers)
    total = style_weight * torch.stack(losses[:ns]).sum() + (content_weight * torch.stack(losses[ns:ns + nc]).sum() if nc else 0.0)
    total.backward()
    return total.detach(), x.grad.detach(), torch.stack(losses).detach()


def gram_matrix(f):
    """network.py:92-108"""
    bs, c, h, w = f.shape
    feat = f.reshape(bs, c, h * w)
    return torch.bmm(feat, feat.transpose(1, 2)) / (c * h * w)


def style_loss(f, target_gram):
    """network.py:110-123 (target is the Gram of ONE style image, expanded)"""
    g = gram_matrix(f)
    return F.mse_loss(g, target_gram.expand_as(g))


def content_loss(f, target):
    """network.py:155-164"""
    return F.mse_loss(f, target)


def feature_loss(f, target):
    """network.py:186-201"""
    bs, c, h, w = f.shape
    return F.mse_loss(f, target).pow(2) / (bs * c * h * w)


def style_targets(style_image, weights):
    """network.py:297-301 + :125-131"""
    with torch.no_grad():
        return [gram_matrix(t) for t in vgg_taps(style_image, weights)]


def loss_network(x, content_image, targets, weights):
    """StyleNetwork.forward (network.py:366-401) as one pass.

    Returns the per-node losses: 5 style, 1 content, 1 feature, in that order
    (a length-7 tensor), plus the taps."""
    with torch.no_grad():
        ctaps = vgg_taps(content_image, weights, upto=CONTENT_TAP + 1)
        ctarget = ctaps[CONTENT_TAP]
    taps = vgg_taps(x, weights)
    losses = [style_loss(t, g) for t, g in zip(taps, targets)]
    losses.append(content_loss(taps[CONTENT_TAP], ctarget))
    losses.append(feature_loss(F.relu(taps[FEATURE_TAP]), F.relu(ctarget)))
    return torch.stack(losses), taps


def gatys_closure(x, content_image, targets, weights,
                  style_weight=100_000.0, content_weight=1.0):
    """One closure of train_gatys (network.py:438-454): total loss and its
    gradient with respect to the image."""
    x = x.detach().clone().requires_grad_(True)
    losses, _ = loss_network(x, content_image, targets, weights)
    total = style_weight * losses[:5].sum() + content_weight * losses[5]
    total.backward()
    return total.detach(), x.grad.detach(), losses.detach()


# --------------------------------------------------------------------------
# Johnson transform net (network.py:461-611, :835-883)
# --------------------------------------------------------------------------
# (state-dict prefix, kind, cin, cout, k, stride).  Indices are the
# nn.Sequential positions of the reference (network.py:521-611).
TRANSFORM_LAYERS = [
    ('0', 'conv', 3, 32, 9, 1), ('1', 'in', 32), ('2', 'relu',),
    ('3', 'conv', 32, 64, 3, 2), ('4', 'in', 64), ('5', 'relu',),
    ('6', 'conv', 64, 128, 3, 2), ('7', 'in', 128), ('8', 'relu',),
    ('9', 'res', 128), ('10', 'res', 128), ('11', 'res', 128),
    ('12', 'res', 128), ('13', 'res', 128),
    ('14', 'up',), ('15', 'conv', 128, 64, 3, 1), ('16', 'in', 64), ('17', 'relu',),
    ('18', 'up',), ('19', 'conv', 64, 32, 3, 1), ('20', 'in', 32), ('21', 'relu',),
    ('22', 'conv', 32, 3, 9, 1),
]


def make_transform_params(seed=2, in_channels=3):
    """Seeded params with the reference's state-dict key names.

    Default nn.Conv2d init (kaiming-uniform a=sqrt(5) -> U(+-1/sqrt(fan_in)) for
    weight and bias) restated with an explicit generator; InstanceNorm affine
    weights are perturbed away from (1, 0) so gamma/beta gradients are
    non-trivial.  ``in_channels=6`` gives the VideoTransformNet (network.py:850)."""
    g = torch.Generator().manual_seed(seed)
    p = {}

    def conv(prefix, cin, cout, k):
        bound = 1.0 / math.sqrt(cin * k * k)
        p[prefix + '.weight'] = (torch.rand(cout, cin, k, k, generator=g) * 2 - 1) * bound
        p[prefix + '.bias'] = (torch.rand(cout, generator=g) * 2 - 1) * bound

    def inorm(prefix, c):
        p[prefix + '.weight'] = 1.0 + 0.1 * torch.randn(c, generator=g)
        p[prefix + '.bias'] = 0.1 * torch.randn(c, generator=g)

    for spec in TRANSFORM_LAYERS:
        name, kind = spec[0], spec[1]
        if kind == 'conv':
            cin = in_channels if name == '0' else spec[2]
            conv(name, cin, spec[3], spec[4])
        elif kind == 'in':
            inorm(name, spec[2])
        elif kind == 'res':
            conv(name + '.conv1', 128, 128, 3)
            inorm(name + '.insn1', 128)
            conv(name + '.conv2', 128, 128, 3)
            inorm(name + '.insn2', 128)
    return p


def _inorm(x, p, prefix):
    # nn.InstanceNorm2d(affine=True), no running stats, eps 1e-5 (network.py:474)
    return F.instance_norm(x, None, None, p[prefix + '.weight'], p[prefix + '.bias'],
                           True, 0.0, 1e-5)


def transform_net(x, p):
    """ImageTransformNet / VideoTransformNet forward (network.py:521-611).

    padding_mode='reflection' is executed as ZERO padding by the pinned
    torch 1.1.0, so zero padding is the reference arithmetic."""
    for spec in TRANSFORM_LAYERS:
        name, kind = spec[0], spec[1]
        if kind == 'conv':
            k, s = spec[4], spec[5]
            x = F.conv2d(x, p[name + '.weight'], p[name + '.bias'], stride=s, padding=k // 2)
        elif kind == 'in':
            x = _inorm(x, p, name)
        elif kind == 'relu':
            x = F.relu(x)
        elif kind == 'up':
            x = F.interpolate(x, scale_factor=2, mode='nearest')
        elif kind == 'res':
            # ResidualBlock.forward network.py:485-506: IN after the add, no final ReLU
            r = x
            o = F.conv2d(x, p[name + '.conv1.weight'], p[name + '.conv1.bias'], padding=1)
            o = F.relu(_inorm(o, p, name + '.insn1'))
            o = F.conv2d(o, p[name + '.conv2.weight'], p[name + '.conv2.bias'], padding=1)
            o = o + r
            x = _inorm(o, p, name + '.insn2')
    return x


def tv_loss(img, factor=1e-6):
    """network.py:621-641"""
    return factor * (torch.sum(torch.abs(img[:, :, :, :-1] - img[:, :, :, 1:])) +
                     torch.sum(torch.abs(img[:, :, :-1, :] - img[:, :, 1:, :])))


def temporal_loss(old_content, old_stylized, cur_content, cur_stylized, temporal_weight=1.0):
    """network.py:885-903 (no optical flow in the reference)"""
    ds = (cur_stylized - old_stylized).norm()
    dc = (cur_content - old_content).norm()
    return (ds / (dc + 1)) * temporal_weight


def fast_st_closure(batch, params, targets, vgg_w, style_weight=100_000.0, content_weight=1.0):
    """static_train closure (network.py:690-731): loss + grads of all params."""
    ps = {k: v.detach().clone().requires_grad_(True) for k, v in params.items()}
    out = transform_net(batch, ps)
    losses, _ = loss_network(out, batch, targets, vgg_w)
    total = (style_weight * losses[:5].sum() + content_weight * losses[5] + tv_loss(out))
    total.backward()
    return total.detach(), {k: v.grad.detach() for k, v in ps.items()}, out.detach(), losses.detach()


def static_test(batches, params, targets, vgg_w, style_weight=100_000.0, feature_weight=1.0):
    """ImageTransformNet.static_test (network.py:772-796): per test batch clamp(net(x), 0, 255) -> loss network with
    the batch as content image -> style_weight * sum(style) + feature_weight * feature; the mean over the batches."""
    losses = []
    with torch.no_grad():
        for batch in batches:
            out = torch.clamp(transform_net(batch, params), min=0, max=255)
            node, _ = loss_network(out, batch, targets, vgg_w)
            losses.append(float(style_weight * node[:5].sum() + feature_weight * node[6]))
    return torch.mean(torch.Tensor(losses))


def video_closure(batch, old_content, old_stylized, params, targets, vgg_w,
                  style_weight=100_000.0, content_weight=1.0, temporal_weight=0.8):
    """video_train closure (network.py:984-1037)."""
    ps = {k: v.detach().clone().requires_grad_(True) for k, v in params.items()}
    x = torch.cat([batch, old_stylized], dim=1)
    out = transform_net(x, ps)
    losses, _ = loss_network(out, batch, targets, vgg_w)
    total = (style_weight * losses[:5].sum() + content_weight * losses[5] + tv_loss(out) +
             temporal_loss(old_content, old_stylized, batch, out, temporal_weight))
    total.backward()
    return total.detach(), {k: v.grad.detach() for k, v in ps.items()}, out.detach(), losses.detach()


def gatys_closure_as_executed(x, content_image, targets, weights,
                              style_weight=100_000.0, content_weight=1.0):
    """The closure of train_gatys exactly AS THE REFERENCE EXECUTES IT: every loss
    node re-runs the VGG prefix from the image (network.py:378-398: content +
    feature node on the input and on the content image, then five style nodes =
    7 + 2 prefix passes), and the VGG weights carry requires_grad so
    autograd also computes their (unused) gradients.  Numerically equal
    to gatys_closure; used as the timed CPU baseline."""
    ws = [(w.detach().clone().requires_grad_(True), b.detach().clone().requires_grad_(True)) for w, b in weights]
    x = x.detach().clone().requires_grad_(True)
    node_losses = []
    # content node (piece 3) and feature node (piece 4): targets reset from content_image
    ctarget = vgg_taps(content_image, ws, upto=CONTENT_TAP + 1)[CONTENT_TAP].detach()
    closs = content_loss(vgg_taps(x, ws, upto=CONTENT_TAP + 1)[CONTENT_TAP], ctarget)
    ftarget = F.relu(vgg_taps(content_image, ws, upto=FEATURE_TAP + 1)[FEATURE_TAP]).detach()
    floss = feature_loss(F.relu(vgg_taps(x, ws, upto=FEATURE_TAP + 1)[FEATURE_TAP]), ftarget)
    for i in range(5):
        node_losses.append(style_loss(vgg_taps(x, ws, upto=i + 1)[i], targets[i]))
    total = style_weight * torch.stack(node_losses).sum() + content_weight * closs
    total.backward()
    return total.detach(), x.grad.detach(), torch.stack(node_losses + [closs, floss]).detach()
