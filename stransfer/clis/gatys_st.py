"""``stransfer gatys_st`` (reference: stransfer/clis/gatys_st.py:10-48)."""
import os

import click

from stransfer.clis import cli


@cli.command('gatys_st')
@click.argument('content-image-path')
@click.argument('style-image-path')
@click.option('-n', '--out-name', default='gatys_converted.png', help='name of the result image (written to results/)')
@click.option('-s', '--steps', default=300, help='L-BFGS outer steps')
@click.option('-cw', '--content-weight', default=1, help='weight of the content loss')
@click.option('-sw', '--style-weight', default=100_000, help='weight of the style loss')
def gatys_st(content_image_path, style_image_path, out_name, steps, content_weight, style_weight):
    """Gatys et al. optimisation of the image itself (slow, one image at a time)."""
    from styletransfer_b200 import constants, dist, img_utils, network
    dist.init()
    root = constants.PROJECT_ROOT_PATH
    content = img_utils.image_loader(os.path.join(root, content_image_path))
    style = img_utils.image_loader(os.path.join(root, style_image_path))
    net = network.StyleNetwork(style, content)
    result = net.train_gatys(style_image=style, content_image=content, steps=steps,
                             style_weight=style_weight, content_weight=content_weight)
    out_dir = os.path.join(root, 'results')
    os.makedirs(out_dir, exist_ok=True)
    img_utils.imshow(result, path=os.path.join(out_dir, out_name))
