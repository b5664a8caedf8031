"""PIL <-> tensor helpers with the reference's names (stransfer/img_utils.py:13-117).
Host glue outside the hot path: centre-crop to a square, resize to IMSIZE, scale to [0,1],
normalise with the ImageNet statistics; the inverse for saving."""
import numpy as np
import torch
from PIL import Image

from . import constants


def _stats(device):
    mean = torch.tensor(constants.IMAGENET_MEAN, device=device).view(-1, 1, 1)
    std = torch.tensor(constants.IMAGENET_STD, device=device).view(-1, 1, 1)
    return mean, std


def image_loader_transform(image: Image.Image) -> torch.Tensor:
    """img_utils.py:13-44 -> [1, C, IMSIZE, IMSIZE] normalised float tensor on DEVICE"""
    w, h = image.size
    side = min(w, h)
    left, top = int(round((w - side) / 2.0)), int(round((h - side) / 2.0))
    image = image.crop((left, top, left + side, top + side)).resize((constants.IMSIZE, constants.IMSIZE), Image.BILINEAR)
    arr = np.asarray(image, dtype=np.float32) / 255.0
    if arr.ndim == 2:
        arr = arr[:, :, None]
    t = torch.from_numpy(arr).permute(2, 0, 1).unsqueeze(0).contiguous().to(constants.DEVICE, torch.float)
    mean, std = _stats(t.device)
    if t.shape[1] == 3:
        t = (t - mean) / std
    return t


def concat_images(im1, im2, dim=2) -> torch.Tensor:
    """img_utils.py:47-61"""
    return torch.cat([im1, im2], dim=dim)


def image_loader(image_path: str) -> torch.Tensor:
    """img_utils.py:64-74"""
    return image_loader_transform(Image.open(image_path))


def imshow(image_tensor: torch.Tensor, ground_truth_image: torch.Tensor = None, denormalize=True, path='out.bmp') -> None:
    """img_utils.py:77-117: optionally concatenates a ground-truth image, de-normalises,
    clamps and writes the picture to ``path``."""
    if ground_truth_image is not None:
        image_tensor = concat_images(image_tensor, ground_truth_image)
    image = image_tensor.detach().cpu().float().clone()
    if denormalize:
        mean, std = _stats(image.device)
        image = image * std + mean
    image = torch.clamp(image, min=0, max=255)
    while image.dim() > 3:
        image = image.squeeze(0)
    arr = (image.clamp(max=1.0) * 255.0 + 0.5).to(torch.uint8).permute(1, 2, 0).numpy()
    Image.fromarray(arr[:, :, 0] if arr.shape[2] == 1 else arr).save(path)
