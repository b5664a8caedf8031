"""Constants of the application (same names as the reference's stransfer/constants.py:1-32).

Differences: IMSIZE can be overridden with the STRANSFER_IMSIZE environment variable
(the 512x512 / 1080p benchmark configs need it) and the default tensor type is NOT
switched globally (torch.set_default_tensor_type is deprecated); tensors are created
on DEVICE explicitly instead.
"""
import os

import torch

RUNS_PATH = 'runs/'
LOG_PATH = os.path.join(RUNS_PATH, 'runtime.log')

IMAGENET_MEAN = [0.485, 0.456, 0.406]
IMAGENET_STD = [0.229, 0.224, 0.225]

def _device():
    """Under torchrun every rank owns the GPU named by LOCAL_RANK (a bare ``cuda`` device would put all the
    ranks of ``torchrun -m stransfer fast_st train ...`` on cuda:0)."""
    if not torch.cuda.is_available():
        return torch.device('cpu')
    local = int(os.environ.get('LOCAL_RANK', 0))
    return torch.device('cuda', local if local < torch.cuda.device_count() else 0)


DEVICE = _device()

IMSIZE = int(os.environ.get('STRANSFER_IMSIZE', 256))

PROJECT_ROOT_PATH = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
