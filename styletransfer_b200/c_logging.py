"""Logging set-up (mirrors reference stransfer/c_logging.py:1-68): a 'StyleTransfer'
logger whose stream handler cooperates with tqdm progress bars, plus a file handler
under runs/ (created lazily, and skipped when the directory is not writable)."""
import logging
import os

import tqdm

from . import constants

_LOGGER = logging.getLogger('StyleTransfer')
_LOGGER.setLevel(logging.INFO)
_LOGGER.handlers = []

LOGGER_FORMATTER = logging.Formatter(
    '%(asctime)s [%(levelname)s] %(module)s.%(funcName)s #%(lineno)d - %(message)s')


class TqdmLoggingHandler(logging.StreamHandler):
    """Writes records through tqdm.write so progress bars stay at the bottom."""

    def emit(self, record):
        try:
            tqdm.tqdm.write(self.format(record))
            self.flush()
        except (KeyboardInterrupt, SystemExit):
            raise
        except Exception:
            self.handleError(record)


_handler = TqdmLoggingHandler()
_handler.setFormatter(LOGGER_FORMATTER)
_LOGGER.addHandler(_handler)
try:
    os.makedirs(constants.RUNS_PATH, exist_ok=True)
    _file_handler = logging.FileHandler(constants.LOG_PATH, mode='w+')
    _file_handler.setFormatter(LOGGER_FORMATTER)
    _LOGGER.addHandler(_file_handler)
except OSError:
    pass


def get_logger() -> logging.Logger:
    return _LOGGER
