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
    raw = np.asarray(image)
    if raw.ndim == 3 and raw.shape[2] == 3 and raw.dtype == np.uint8 and constants.DEVICE.type == 'cuda':
        # RGB picture: the uint8 pixels go to the device (a quarter of the bytes) and ToTensor + Normalize run
        # there (st_image_u8_to_nchw: bit-identical to the host arithmetic below)
        from . import ops
        return ops.image_u8_to_nchw(torch.from_numpy(np.ascontiguousarray(raw)).unsqueeze(0).to(constants.DEVICE))
    arr = raw.astype(np.float32) / 255.0
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


class FrameWriter:
    """Writes stylised frames as pictures WITHOUT stalling the device loop (reference network.py:1117-1134 calls
    imshow -> .cpu() -> PIL for every frame): the de-normalise / clamp / uint8 conversion runs on the device
    (st_image_nchw_to_u8, the arithmetic of imshow), the picture is copied to one of ``depth`` pinned host buffers
    on a side stream, and a worker thread encodes it once its copy event has fired.  ``close()`` drains the queue."""

    def __init__(self, depth=4, workers=2):
        import queue
        import threading
        self.depth = depth
        self.slots = None
        self.free = queue.Queue()
        self.jobs = queue.Queue()
        self.errors = []
        self.copy_stream = None
        self.threads = [threading.Thread(target=self._work, daemon=True) for _ in range(workers)]
        for t in self.threads:
            t.start()

    def _work(self):
        while True:
            job = self.jobs.get()
            if job is None:
                return
            slot, event, n, path = job
            try:
                event.synchronize()
                arr = self.slots[slot][1][:n].numpy()
                Image.fromarray(arr[0]).save(path)
            except Exception as exc:                      # surfaced by close()
                self.errors.append(exc)
            finally:
                self.free.put(slot)

    def submit(self, image, path):
        """image: [1, 3, H, W] (or [3, H, W]) normalised fp32 on the device."""
        from . import ops
        image = image.detach()
        if image.dim() == 3:
            image = image.unsqueeze(0)
        N, _, H, W = image.shape
        if self.slots is None or self.slots[0][0].shape[1:3] != (H, W):
            self.drain()
            dev = image.device
            self.copy_stream = torch.cuda.Stream(dev)
            self.slots = [(torch.empty((1, H, W, 3), dtype=torch.uint8, device=dev),
                           torch.empty((1, H, W, 3), dtype=torch.uint8).pin_memory()) for _ in range(self.depth)]
            while not self.free.empty():
                self.free.get()
            for i in range(self.depth):
                self.free.put(i)
        slot = self.free.get()                            # blocks only when `depth` frames are in flight
        dev_u8, host_u8 = self.slots[slot]
        ops.image_nchw_to_u8(image[:1], out=dev_u8)
        cur = torch.cuda.current_stream(image.device)
        self.copy_stream.wait_stream(cur)
        with torch.cuda.stream(self.copy_stream):
            host_u8.copy_(dev_u8, non_blocking=True)
            event = torch.cuda.Event()
            event.record(self.copy_stream)
        self.jobs.put((slot, event, 1, path))

    def drain(self):
        if self.slots is None:
            return
        held = [self.free.get() for _ in range(self.depth)]      # every slot back = every job done
        for i in held:
            self.free.put(i)

    def close(self):
        self.drain()
        for _ in self.threads:
            self.jobs.put(None)
        for t in self.threads:
            t.join()
        if self.errors:
            raise self.errors[0]
