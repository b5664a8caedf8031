"""``python -m stransfer`` (reference: stransfer/__main__.py:4-6)."""
from stransfer.clis import cli

if __name__ == '__main__':
    cli(prog_name='stransfer')
