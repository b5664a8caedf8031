"""``stransfer video_st train|convert-video`` (reference: stransfer/clis/video_st.py:17-87)."""
import os

import click

from stransfer.clis import cli


@cli.group('video_st')
def video_st():
    """Video style transfer: transform net fed with the previous stylised frame."""


@video_st.command('train')
@click.argument('style-image-path')
@click.option('-e', '--epochs', default=50)
@click.option('-b', '--batch-size', default=4, help='GLOBAL number of clips per step (sharded under torchrun)')
@click.option('-cw', '--content-weight', default=1)
@click.option('-sw', '--style-weight', default=100_000)
@click.option('-tw', '--temporal-weight', default=0.8)
@click.option('--use-pretrained-fast-st', is_flag=True, help='warm start from the latest fast_st weights of this style')
def train(style_image_path, epochs, batch_size, content_weight, style_weight, temporal_weight, use_pretrained_fast_st):
    from styletransfer_b200 import c_logging, constants, dist, img_utils, network
    dist.init()
    style = img_utils.image_loader(os.path.join(constants.PROJECT_ROOT_PATH, style_image_path))
    style_name = style_image_path.split('/')[-1]
    fast_dict = None
    if use_pretrained_fast_st:
        try:
            fast_dict = network._load_latest_model_weigths('fast_st', style_name)
        except AssertionError:
            c_logging.get_logger().warning('no fast_st weights for "%s": training the video net from scratch', style_name)
    net = network.VideoTransformNet(style, batch_size, fast_transfer_dict=fast_dict).to(constants.DEVICE)
    net.video_train(style_name=style_name, epochs=epochs, temporal_weight=temporal_weight,
                    style_weight=style_weight, content_weight=content_weight)


@video_st.command('convert-video')
@click.argument('video-path')
@click.argument('style-name')
@click.option('-o', '--out-dir', default='results/')
@click.option('--fps', default=24.0)
def convert_video(video_path, style_name, out_dir, fps):
    import torch
    from styletransfer_b200 import constants, network
    net = network.VideoTransformNet(torch.rand([3, 255, 255])).to(constants.DEVICE)
    net.process_video(video_path=video_path, style_name=style_name, out_dir=out_dir, fps=fps)
