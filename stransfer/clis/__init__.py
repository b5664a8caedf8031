"""Command line of the reference (stransfer/clis/__init__.py:6-15): one click group with the
``gatys_st``, ``fast_st`` and ``video_st`` commands.  Under ``torchrun`` the training commands
run data parallel (styletransfer_b200/dist.py)."""
import click


@click.group()
def cli():
    """Style transfer on B200: Gatys optimisation, fast (Johnson) and video style transfer."""


from stransfer.clis import fast_st, gatys_st, video_st  # noqa: E402,F401
