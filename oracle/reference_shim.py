"""Import the UNMODIFIED reference (``/root/reference/stransfer``) in the build
container.  TEST INFRASTRUCTURE ONLY -- used by ``oracle/make_golden.py`` and
``tests/test_oracle_vs_reference.py`` (skipped when /root/reference is absent).

Shims, none of them on the numeric path:
  1. stub modules for tensorboardX / imageio / matplotlib / colored_traceback;
  2. ``nn.Conv2d(padding_mode='reflection')`` -> ``'zeros'``: that string is not
     a valid mode; under the reference's pinned torch 1.1.0 it silently fell
     through to zero padding, under torch 2.x it raises;
  3. ``torchvision.models.vgg19(pretrained=True)`` -> random-init VGG-19 whose
     five executed convs carry ``oracle.make_vgg_weights(seed)`` (no network).
"""
import importlib.abc
import importlib.machinery
import importlib.util
import os
import sys
import types

import torch
import torch.nn as nn

_HERE = os.path.dirname(os.path.abspath(__file__))


def _reference_root():
    """The reference's sources in the build container; on the GPU box the sourceless bytecode staged by
    ``oracle/build_ref.py`` under ``oracle/_ref`` (git-ignored, travels with gpurun)."""
    for cand in (os.environ.get('STRANSFER_REFERENCE_ROOT'), '/root/reference', os.path.join(_HERE, '_ref')):
        if cand and (os.path.isfile(os.path.join(cand, 'stransfer', 'network.py')) or
                     os.path.isfile(os.path.join(cand, 'stransfer', 'network' + BYTECODE_EXT))):
            return cand
    return '/root/reference'


BYTECODE_EXT = '.refbc'


class _BytecodeFinder(importlib.abc.MetaPathFinder):
    """Imports ``stransfer`` / ``stransfer.*`` from the sourceless ``*.refbc`` files staged by oracle/build_ref.py
    (plain .pyc content under an extension the gpurun snapshot does not drop)."""

    def __init__(self, root):
        self.root = root

    def find_spec(self, name, path=None, target=None):
        if name != 'stransfer' and not name.startswith('stransfer.'):
            return None
        rel = os.path.join(self.root, *name.split('.'))
        pkg, mod = os.path.join(rel, '__init__' + BYTECODE_EXT), rel + BYTECODE_EXT
        if os.path.isfile(pkg):
            return importlib.util.spec_from_file_location(name, pkg, loader=importlib.machinery.SourcelessFileLoader(name, pkg),
                                                          submodule_search_locations=[rel])
        if os.path.isfile(mod):
            return importlib.util.spec_from_file_location(name, mod, loader=importlib.machinery.SourcelessFileLoader(name, mod))
        return None


REFERENCE_ROOT = _reference_root()


def available():
    return os.path.isdir(os.path.join(REFERENCE_ROOT, 'stransfer'))


def _stub_modules():
    tbx = types.ModuleType('tensorboardX')

    class SummaryWriter:
        def __init__(self, *a, **k): pass
        def add_scalar(self, *a, **k): pass
        def add_image(self, *a, **k): pass
    tbx.SummaryWriter = SummaryWriter
    sys.modules.setdefault('tensorboardX', tbx)

    imageio = types.ModuleType('imageio')
    core = types.ModuleType('imageio.core')
    fmt = types.ModuleType('imageio.core.format')

    class Format:
        class Reader: pass
    fmt.Format = Format
    core.format = fmt
    imageio.core = core
    sys.modules.setdefault('imageio', imageio)
    sys.modules.setdefault('imageio.core', core)
    sys.modules.setdefault('imageio.core.format', fmt)

    mpl = types.ModuleType('matplotlib')
    plt = types.ModuleType('matplotlib.pyplot')
    mpl.pyplot = plt
    sys.modules.setdefault('matplotlib', mpl)
    sys.modules.setdefault('matplotlib.pyplot', plt)

    ct = types.ModuleType('colored_traceback')
    ct.add_hook = lambda *a, **k: None
    sys.modules.setdefault('colored_traceback', ct)


_ORIG_CONV_INIT = nn.Conv2d.__init__


def _conv_init(self, *a, **k):
    if k.get('padding_mode') == 'reflection':
        k['padding_mode'] = 'zeros'
    _ORIG_CONV_INIT(self, *a, **k)


def load_reference(vgg_seed=1, workdir='/tmp/stransfer_ref_cwd', n_convs=5):
    """Returns the reference's ``stransfer.network`` module (CPU).  The first ``n_convs`` convs of its VGG-19 carry
    ``oracle.make_vgg_weights(vgg_seed, n_convs)`` (the rest keep torchvision's unseeded random init)."""
    import warnings
    import torchvision
    from oracle import oracle

    _stub_modules()
    nn.Conv2d.__init__ = _conv_init

    def vgg19(pretrained=False, **kw):
        m = torchvision.models.vgg.vgg19(weights=None)
        convs = [layer for layer in m.features if isinstance(layer, nn.Conv2d)]
        with torch.no_grad():
            for conv, (w, b) in zip(convs, oracle.make_vgg_weights(vgg_seed, n_convs)):
                conv.weight.copy_(w)
                conv.bias.copy_(b)
        return m
    torchvision.models.vgg19 = vgg19

    os.makedirs(workdir, exist_ok=True)      # the import creates ./runs/runtime.log
    cwd = os.getcwd()
    os.chdir(workdir)
    # the repo root holds a drop-in package that is ALSO called `stransfer`;
    # make sure the reference's copy is the one imported here.
    for name in [n for n in sys.modules if n == 'stransfer' or n.startswith('stransfer.')]:
        del sys.modules[name]
    sys.path.insert(0, REFERENCE_ROOT)
    finder = _BytecodeFinder(REFERENCE_ROOT) if os.path.isfile(os.path.join(REFERENCE_ROOT, 'stransfer', 'network' + BYTECODE_EXT)) else None
    if finder is not None:
        sys.meta_path.insert(0, finder)
    try:
        with warnings.catch_warnings():
            warnings.simplefilter('ignore')
            import stransfer.network as ref_network
    finally:
        if finder is not None:
            sys.meta_path.remove(finder)
        sys.path.remove(REFERENCE_ROOT)
        os.chdir(cwd)
    assert ref_network.__file__.startswith(REFERENCE_ROOT), ref_network.__file__
    mods = {n: m for n, m in sys.modules.items() if n == 'stransfer' or n.startswith('stransfer.')}
    for n in mods:                           # leave no alias behind for the drop-in package
        del sys.modules[n]
    ref_network._shim_modules = mods
    torch.set_default_dtype(torch.float32)
    return ref_network
