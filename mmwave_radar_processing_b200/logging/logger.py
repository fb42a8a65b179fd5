"""Logger factory with the reference's contract (reference: mmwave_radar_processing/logging/logger.py:11-65).

``get_logger(name)`` hands out stdout loggers that share one format; the first call
configures the package default lazily.
"""
import logging
import sys

_FORMAT = "%(asctime)s | %(levelname)-8s | %(name)s | %(message)s"
_DEFAULT_LOGGER = None


def setup_logger(name="mmwave_radar_processing", level=logging.DEBUG, *, set_default=True):
    global _DEFAULT_LOGGER
    log = logging.getLogger(name)
    log.setLevel(level)
    for h in list(log.handlers):
        log.removeHandler(h)
    stream = logging.StreamHandler(sys.stdout)
    stream.setFormatter(logging.Formatter(fmt=_FORMAT, datefmt="%H:%M:%S"))
    log.addHandler(stream)
    log.propagate = False
    if set_default:
        _DEFAULT_LOGGER = log
    return log


def get_logger(name=None):
    if _DEFAULT_LOGGER is None:
        setup_logger()
    if name is None:
        return _DEFAULT_LOGGER
    log = logging.getLogger(name)
    if not log.hasHandlers():
        setup_logger(name=name, level=_DEFAULT_LOGGER.level, set_default=False)
    return log
