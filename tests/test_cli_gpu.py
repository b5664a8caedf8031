"""The reference's command line end to end on the GPU (clis/gatys_st.py:10-48): image files in, picture out."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def test_gatys_st_cli_writes_a_picture(tmp_path, monkeypatch):
    from click.testing import CliRunner
    from PIL import Image
    from stransfer.clis import cli
    from styletransfer_b200 import constants, network, optim
    rng = np.random.default_rng(0)
    content, style = tmp_path / 'content.png', tmp_path / 'style.png'
    Image.fromarray(rng.integers(0, 255, (300, 400, 3), dtype=np.uint8)).save(content)      # cropped + resized to IMSIZE
    Image.fromarray(np.kron(rng.integers(0, 255, (16, 16, 3), dtype=np.uint8), np.ones((20, 20, 1), dtype=np.uint8))).save(style)
    seeded = network.load_vgg19_prefix_weights
    monkeypatch.setattr(network, 'load_vgg19_prefix_weights', lambda seed=None: seeded(seed=1))   # no download attempt
    made = []
    real = optim.DeviceLBFGS.__init__

    def spy(self, *a, **k):
        made.append(1)
        real(self, *a, **k)
    monkeypatch.setattr(optim.DeviceLBFGS, '__init__', spy)
    name = f'pytest_gatys_{os.getpid()}.png'
    out = os.path.join(constants.PROJECT_ROOT_PATH, 'results', name)
    try:
        res = CliRunner().invoke(cli, ['gatys_st', str(content), str(style), '-s', '2', '-n', name], catch_exceptions=False)
        assert res.exit_code == 0, res.output
        assert made, 'train_gatys did not run on the device L-BFGS'
        img = np.asarray(Image.open(out))
        assert img.shape == (constants.IMSIZE, constants.IMSIZE, 3) and img.std() > 0
    finally:
        if os.path.exists(out):
            os.unlink(out)
