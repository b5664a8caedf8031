"""``stransfer gatys_st`` (reference: stransfer/clis/gatys_st.py:10-48)."""
import os

import click

from stransfer.clis import cli


@cli.command('gatys_st')
@click.argument('content-image-path', required=False, metavar='CONTENT_IMAGE_PATH')
@click.argument('style-image-path', required=False, metavar='STYLE_IMAGE_PATH')
@click.option('-p', '--pairs-file', default=None,
              help='(extension) text file with one "CONTENT STYLE [OUT-NAME]" line per pair; under torchrun rank r of W '
                   'optimises lines r, r+W, ... -- independent pairs, no communication')
@click.option('-n', '--out-name', default='gatys_converted.png', help='name of the result image (written to results/)')
@click.option('-s', '--steps', default=300, help='L-BFGS outer steps')
@click.option('-cw', '--content-weight', default=1, help='weight of the content loss')
@click.option('-sw', '--style-weight', default=100_000, help='weight of the style loss')
def gatys_st(content_image_path, style_image_path, pairs_file, out_name, steps, content_weight, style_weight):
    """Gatys et al. optimisation of the image itself (slow, one image at a time)."""
    import torch
    from styletransfer_b200 import constants, dist, img_utils, network
    rank, world, local = dist.env_rank_world()
    if world > 1 and torch.cuda.is_available():
        torch.cuda.set_device(local % torch.cuda.device_count())     # independent pairs: no process group needed
    root = constants.PROJECT_ROOT_PATH
    if pairs_file is None:
        if not (content_image_path and style_image_path):
            raise click.UsageError('CONTENT-IMAGE-PATH and STYLE-IMAGE-PATH (or --pairs-file) are required')
        # the reference's single pair; under torchrun only rank 0 has work
        pairs = [(content_image_path, style_image_path, out_name)] if rank == 0 else []
    else:
        pairs = parse_pairs(open(os.path.join(root, pairs_file)).read(), out_name)[rank::world]
    out_dir = os.path.join(root, 'results')
    os.makedirs(out_dir, exist_ok=True)
    for content_path, style_path, name in pairs:
        content = img_utils.image_loader(os.path.join(root, content_path))
        style = img_utils.image_loader(os.path.join(root, style_path))
        net = network.StyleNetwork(style, content)
        result = net.train_gatys(style_image=style, content_image=content, steps=steps,
                                 style_weight=style_weight, content_weight=content_weight)
        img_utils.imshow(result, path=os.path.join(out_dir, name))


def parse_pairs(text, default_name='gatys_converted.png'):
    """[(content, style, out name)] from the lines of a pairs file; blank lines and # comments are skipped; a pair
    without an explicit name gets ``<stem>_<index>.png``."""
    pairs = []
    stem, ext = os.path.splitext(default_name)
    for line in text.splitlines():
        parts = line.split('#', 1)[0].split()
        if not parts:
            continue
        if len(parts) not in (2, 3):
            raise click.UsageError('pairs file: expected "CONTENT STYLE [OUT-NAME]", got: ' + line)
        name = parts[2] if len(parts) == 3 else f'{stem}_{len(pairs)}{ext or ".png"}'
        pairs.append((parts[0], parts[1], name))
    return pairs
