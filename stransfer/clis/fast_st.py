"""``stransfer fast_st train|convert-image`` (reference: stransfer/clis/fast_st.py:17-63)."""
import os

import click

from stransfer.clis import cli


@cli.group('fast_st')
def fast_st():
    """Fast (Johnson et al.) style transfer: train a transform net, convert images."""


@fast_st.command('train')
@click.argument('style-image-path')
@click.option('-e', '--epochs', default=50)
@click.option('-b', '--batch-size', default=4, help='GLOBAL batch size (sharded over the ranks under torchrun)')
@click.option('-cw', '--content-weight', default=1)
@click.option('-sw', '--style-weight', default=100_000)
def train(style_image_path, epochs, batch_size, content_weight, style_weight):
    from styletransfer_b200 import constants, dist, img_utils, network
    dist.init()
    style = img_utils.image_loader(os.path.join(constants.PROJECT_ROOT_PATH, style_image_path))
    style_name = style_image_path.split('/')[-1]          # extension included, as in the reference
    net = network.ImageTransformNet(style, batch_size).to(constants.DEVICE)
    net.static_train(style_name=style_name, epochs=epochs, style_weight=style_weight, content_weight=content_weight)


@fast_st.command('convert-image')
@click.argument('image-path')
@click.argument('style-name')
@click.option('-o', '--out-dir', default='results/')
def convert_image(image_path, style_name, out_dir):
    import torch
    from styletransfer_b200 import constants, network
    net = network.ImageTransformNet(torch.rand([3, 255, 255])).to(constants.DEVICE)   # style image is a dummy here
    net.process_image(image_path=image_path, style_name=style_name, out_dir=out_dir)
