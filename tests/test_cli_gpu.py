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


def test_fast_st_convert_image_cli_uses_a_saved_checkpoint(tmp_path):
    """clis/fast_st.py:45-64 -> ImageTransformNet.process_image (network.py:798-832): a state dict saved the way
    static_train saves it (network.py:767-770, reference key names) is found by the 'latest weights' rule, loaded,
    and the stylised picture equals the module's own forward on the loaded image."""
    from click.testing import CliRunner
    from PIL import Image
    from oracle import oracle
    from stransfer.clis import cli
    from styletransfer_b200 import constants, img_utils, network
    style_name = f'pyteststyle{os.getpid()}'
    models = os.path.join(constants.PROJECT_ROOT_PATH, 'data', 'models')
    os.makedirs(models, exist_ok=True)
    ckpt = os.path.join(models, f'fast_st_{style_name}_epoch0.pth')
    params = oracle.make_transform_params(8, 3)
    torch.save({k: v.clone() for k, v in params.items()}, ckpt)
    picture = tmp_path / 'photo.png'
    Image.fromarray(np.random.default_rng(1).integers(0, 255, (260, 300, 3), dtype=np.uint8)).save(picture)
    out_dir = os.path.join('results', f'pytest_{os.getpid()}') + '/'
    out_file = os.path.join(constants.PROJECT_ROOT_PATH, out_dir, f'converted_fast_st_{style_name}.png')
    try:
        res = CliRunner().invoke(cli, ['fast_st', 'convert-image', str(picture), style_name, '-o', out_dir], catch_exceptions=False)
        assert res.exit_code == 0, res.output
        got = np.asarray(Image.open(out_file)).astype(np.int32)
        x = img_utils.image_loader(str(picture))
        net = network.ImageTransformNet(torch.zeros(1, 3, 8, 8, device='cuda')).cuda()
        net.load_state_dict({k: v.clone() for k, v in params.items()})
        with torch.no_grad():
            want = net(x)
        ref_path = tmp_path / 'want.png'
        img_utils.imshow(want, path=str(ref_path))
        assert got.shape == (constants.IMSIZE, constants.IMSIZE, 3)
        assert np.abs(got - np.asarray(Image.open(ref_path)).astype(np.int32)).max() == 0
    finally:
        for f in (ckpt, out_file):
            if os.path.exists(f):
                os.unlink(f)
        d = os.path.join(constants.PROJECT_ROOT_PATH, out_dir)
        if os.path.isdir(d) and not os.listdir(d):
            os.rmdir(d)


def test_process_video_recurrence_from_a_checkpoint(tmp_path):
    """VideoTransformNet.process_video (network.py:1071-1158) on an iterable of frames: loads the latest
    video_st checkpoint, feeds frame_t ++ stylised_{t-1} (the first frame with itself) and writes one PNG per
    frame; the outputs equal a hand-written recurrence on the same weights."""
    from oracle import oracle
    from styletransfer_b200 import constants, network
    style_name = f'pytestvideo{os.getpid()}'
    models = os.path.join(constants.PROJECT_ROOT_PATH, 'data', 'models')
    os.makedirs(models, exist_ok=True)
    ckpt = os.path.join(models, f'video_st_{style_name}_epoch3.pth')
    params = oracle.make_transform_params(8, 6)
    torch.save({k: v.clone() for k, v in params.items()}, ckpt)
    gen = torch.Generator().manual_seed(3)
    frames = [torch.randn(1, 3, 32, 48, generator=gen) for _ in range(3)]
    work = str(tmp_path / 'work') + '/'
    try:
        net = network.VideoTransformNet(torch.zeros(1, 3, 8, 8, device='cuda')).cuda()
        outs = net.process_video('unused.mp4', style_name=style_name, working_dir=work, out_dir=str(tmp_path / 'out'),
                                 frames=[f.clone() for f in frames])
        assert sorted(os.listdir(work), key=lambda x: int(x.split('.')[0])) == ['0.png', '1.png', '2.png']
        ref = network.VideoTransformNet(torch.zeros(1, 3, 8, 8, device='cuda')).cuda()
        ref.load_state_dict({k: v.clone() for k, v in params.items()})
        old = None
        with torch.no_grad():
            for f, got in zip(frames, outs):
                f = f.cuda()
                old = f if old is None else old
                want = ref(torch.cat([f, old], dim=1))
                assert torch.equal(want, got)
                old = want
    finally:
        if os.path.exists(ckpt):
            os.unlink(ckpt)


def test_image_boundary_kernels_match_host_arithmetic():
    """img_utils.image_loader_transform / imshow (reference img_utils.py:33-42, :105-109) with the normalisation and
    the de-normalise + clamp + uint8 conversion on the device: bit-identical to the host (numpy / torch CPU) code."""
    from styletransfer_b200 import constants, ops
    rng = np.random.default_rng(3)
    u8 = torch.from_numpy(rng.integers(0, 256, (2, 37, 53, 3), dtype=np.uint8))
    mean = torch.tensor(constants.IMAGENET_MEAN).view(1, 3, 1, 1)
    std = torch.tensor(constants.IMAGENET_STD).view(1, 3, 1, 1)
    want = (torch.from_numpy(u8.numpy().astype(np.float32) / 255.0).permute(0, 3, 1, 2) - mean) / std
    got = ops.image_u8_to_nchw(u8.cuda())
    assert torch.equal(got.cpu(), want)
    x = torch.randn(2, 3, 37, 53, generator=torch.Generator().manual_seed(5)) * 1.5
    img = torch.clamp(x * std + mean, min=0, max=255)
    want_u8 = (img.clamp(max=1.0) * 255.0 + 0.5).to(torch.uint8).permute(0, 2, 3, 1)
    assert torch.equal(ops.image_nchw_to_u8(x.cuda()).cpu(), want_u8)


def test_process_video_pipelines_frames_to_disk(tmp_path):
    """VideoTransformNet.process_video (network.py:1071-1158): the recurrent per-frame forward with the pictures
    converted on the device, downloaded asynchronously and encoded by worker threads -- the PNGs equal what the
    reference's synchronous loop (imshow of every frame) writes, and the recurrence equals cat([frame, previous])."""
    from PIL import Image
    from oracle import oracle
    from styletransfer_b200 import constants, img_utils, network
    style_name = f'pytestvideo{os.getpid()}'
    models = os.path.join(constants.PROJECT_ROOT_PATH, 'data', 'models')
    os.makedirs(models, exist_ok=True)
    ckpt = os.path.join(models, f'video_st_{style_name}_epoch0.pth')
    params = oracle.make_transform_params(8, 6)
    torch.save({k: v.clone() for k, v in params.items()}, ckpt)
    g = torch.Generator().manual_seed(9)
    frames = [torch.randn(1, 3, 40, 56, generator=g)]
    for _ in range(6):
        frames.append(frames[-1] + 0.05 * torch.randn(1, 3, 40, 56, generator=g))
    work = f'workdir_pytest_{os.getpid()}/'
    work_abs = os.path.join(constants.PROJECT_ROOT_PATH, work)
    try:
        net = network.VideoTransformNet(torch.zeros(1, 3, 8, 8, device='cuda')).cuda()
        outs = net.process_video('unused.mp4', style_name=style_name, working_dir=work, out_dir=str(tmp_path) + '/', frames=frames)
        assert len(outs) == len(frames) and sorted(os.listdir(work_abs), key=lambda x: int(x.split('.')[0])) == [f'{i}.png' for i in range(7)]
        ref = network.VideoTransformNet(torch.zeros(1, 3, 8, 8, device='cuda')).cuda()
        ref.load_state_dict({k: v.clone() for k, v in params.items()})
        old = None
        with torch.no_grad():
            for i, f in enumerate(frames):
                f = f.cuda()
                old = f if old is None else old
                out = ref(torch.cat([f, old], dim=1))
                old = out
                assert torch.equal(out, outs[i])
                want = tmp_path / f'want{i}.png'
                img_utils.imshow(out[0], path=str(want))
                assert np.array_equal(np.asarray(Image.open(want)), np.asarray(Image.open(os.path.join(work_abs, f'{i}.png'))))
    finally:
        import shutil
        shutil.rmtree(work_abs, ignore_errors=True)
        if os.path.exists(ckpt):
            os.unlink(ckpt)
