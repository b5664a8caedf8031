"""Data feeding with the reference's entry points (stransfer/dataset.py:141-360).

The reference downloads COCO / a video set over the network and decodes them with PIL
and imageio.  Downloads and video codecs are outside the hot path (and unavailable
offline), so this module reads what is already on disk -- images under
``data/coco_dataset/images`` and frame folders under ``data/video`` -- and offers seeded
synthetic loaders with the same yield shapes for benchmarks and tests."""
import os
import random

import torch
from PIL import Image
from torch.utils.data import DataLoader, Dataset

from . import c_logging, constants, img_utils

LOGGER = c_logging.get_logger()
BASE_COCO_PATH = os.path.join(constants.PROJECT_ROOT_PATH, 'data', 'coco_dataset')
IMAGE_FOLDER_PATH = os.path.join(BASE_COCO_PATH, 'images')
VIDEO_DATA_PATH = os.path.join(constants.PROJECT_ROOT_PATH, 'data', 'video')


def make_batches(l, n):
    """dataset.py:125-138"""
    return [l[i:i + n] for i in range(0, len(l), n)]


class CocoDataset(Dataset):
    """dataset.py:141-197: items are [1, 3, IMSIZE, IMSIZE] tensors; non-RGB / unreadable
    files are replaced by another random image."""

    def __init__(self, images=None, image_limit=None):
        self.images = sorted(os.listdir(IMAGE_FOLDER_PATH)) if images is None else list(images)
        if image_limit:
            self.images = self.images[:image_limit]

    def __len__(self):
        return len(self.images)

    def __getitem__(self, idx):
        for _ in range(16):
            try:
                image = img_utils.image_loader(os.path.join(IMAGE_FOLDER_PATH, self.images[idx]))
                if image.shape[1] == 3:
                    return image
                LOGGER.warning('Discarding image with %d color channels', image.shape[1])
            except Exception:                               # unreadable file / index past the end
                pass
            idx = random.randint(0, len(self.images) - 1)
        raise RuntimeError('no readable RGB image found in ' + IMAGE_FOLDER_PATH)


class SyntheticImages(Dataset):
    """Seeded stand-in for CocoDataset: ``count`` images [1, 3, size, size] ~ N(0, 1)
    (the value range of ImageNet-normalised pictures)."""

    def __init__(self, count, size=None, seed=0):
        self.count, self.size, self.seed = count, size or constants.IMSIZE, seed

    def __len__(self):
        return self.count

    def __getitem__(self, idx):
        g = torch.Generator().manual_seed(self.seed * 1_000_003 + idx)
        return torch.randn(1, 3, self.size, self.size, generator=g)


def get_coco_loader(batch_size=4, test_split=0.10, test_limit=None, train_limit=None, synthetic=None, seed=0):
    """dataset.py:314-360 -> (test_loader, train_loader), batches [B, 1, 3, S, S].
    ``synthetic`` = number of seeded random images to use when no image folder exists.  The shuffle is seeded
    (``seed``): under data parallelism every rank draws the SAME permutation, so that ``dist.shard`` cuts disjoint
    slices of one global batch."""
    if os.path.isdir(IMAGE_FOLDER_PATH) and synthetic is None:
        all_images = sorted(os.listdir(IMAGE_FOLDER_PATH))
        split_idx = int(len(all_images) * test_split)
        test_dataset = CocoDataset(all_images[:split_idx], test_limit)
        train_dataset = CocoDataset(all_images[split_idx:], train_limit)
    else:
        n = synthetic or 64 * batch_size
        LOGGER.warning('no images under %s: using %d seeded synthetic images', IMAGE_FOLDER_PATH, n)
        n_test = max(batch_size, min(test_limit or n, int(n * test_split)))
        test_dataset, train_dataset = SyntheticImages(n_test, seed=1), SyntheticImages(min(train_limit or n, n), seed=2)
    LOGGER.info('Train set has %d entries', len(train_dataset))
    LOGGER.info('Test set has %d entries', len(test_dataset))
    mk = lambda ds, k: DataLoader(ds, batch_size=batch_size, num_workers=0, drop_last=True, shuffle=True,
                                  generator=torch.Generator().manual_seed(seed * 2 + k))
    return mk(test_dataset, 0), mk(train_dataset, 1)


class FrameFolderReader:
    """Minimal stand-in for an imageio reader: yields the image files of a folder in
    numeric order through ``get_next_data`` (IndexError at the end, like imageio)."""

    def __init__(self, folder):
        names = [n for n in os.listdir(folder) if n.lower().endswith(('.png', '.jpg', '.jpeg', '.bmp'))]
        self.paths = [os.path.join(folder, n) for n in sorted(names, key=lambda x: (len(x), x))]
        self.i = 0

    def get_next_data(self):
        if self.i >= len(self.paths):
            raise IndexError('end of clip')
        self.i += 1
        return Image.open(self.paths[self.i - 1]).convert('RGB')


class SyntheticClip:
    """Seeded clip: frame_{t+1} = frame_t + 0.05 * noise, already normalised tensors."""

    def __init__(self, frames, size, seed):
        self.g = torch.Generator().manual_seed(seed)
        self.left, self.size, self.cur = frames, size, None

    def get_next_data(self):
        if self.left <= 0:
            raise IndexError('end of clip')
        self.left -= 1
        noise = torch.randn(1, 3, self.size, self.size, generator=self.g)
        self.cur = noise if self.cur is None else self.cur + 0.05 * noise
        return self.cur.clone()


def open_video(path):
    if os.path.isdir(path):
        return FrameFolderReader(path)
    raise RuntimeError(f'{path}: decoding video containers needs imageio-ffmpeg, which is not available here; '
                       'extract the frames into a folder and pass that instead')


class VideoDataset:
    """dataset.py:200-277: iterating yields lists of ``batch_size`` readers (one per video)."""

    def __init__(self, videos=None, data_limit=None, batch_size=3, synthetic_frames=None, size=None):
        if synthetic_frames is not None or not os.path.isdir(VIDEO_DATA_PATH):
            n = (data_limit or 2 * batch_size)
            frames, size = synthetic_frames or 8, size or constants.IMSIZE
            self._make = [lambda i=i: SyntheticClip(frames, size, 1000 + i) for i in range(n)]
        else:
            vids = sorted(os.listdir(VIDEO_DATA_PATH)) if videos is None else list(videos)
            if data_limit:
                vids = vids[:data_limit]
            self._make = [lambda v=v: open_video(os.path.join(VIDEO_DATA_PATH, v)) for v in vids]
        self.batch_size = min(batch_size, len(self._make))
        self.batches = [b for b in make_batches(self._make, self.batch_size) if len(b) == self.batch_size]

    def __len__(self):
        return len(self.batches)

    def __iter__(self):
        for batch in self.batches:
            yield [mk() for mk in batch]


def iterate_on_video_batches(batch, max_frames=90 * 24):
    """dataset.py:280-311: yields [B, 3, S, S] tensors, one frame from each reader, until the
    shortest clip ends.  An iterable of ready tensors (or of (frames, flow) pairs) is passed through unchanged."""
    if not isinstance(batch, (list, tuple)) or (batch and isinstance(batch[0], (torch.Tensor, tuple))):
        yield from batch
        return
    try:
        for _ in range(max_frames):
            frames = []
            for reader in batch:
                frame = reader.get_next_data()
                frames.append(frame if isinstance(frame, torch.Tensor) else img_utils.image_loader_transform(frame))
            yield torch.cat(frames, dim=0)
    except IndexError:
        pass


def write_video(frames_dir, out_dir, name, fps):
    """The reference muxes the PNG frames into an mp4 with imageio (network.py:1139-1156).
    Without a codec the stylised PNG frames in ``frames_dir`` are the result."""
    os.makedirs(out_dir, exist_ok=True)
    try:
        import imageio
    except ImportError:
        LOGGER.warning('imageio is not installed: stylised frames are left in %s (no %s written)', frames_dir, name)
        return None
    import numpy as np
    writer = imageio.get_writer(os.path.join(out_dir, name), fps=fps)
    for frame_name in sorted(os.listdir(frames_dir), key=lambda x: int(x.split('.')[0])):
        writer.append_data(np.array(Image.open(os.path.join(frames_dir, frame_name))))
    writer.close()
    return os.path.join(out_dir, name)
