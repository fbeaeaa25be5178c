"""`get_logger` with hivemind's behaviour (hivemind.utils.logging.get_logger, imported by
/root/reference/sahajbert/huggingface_auth.py:13): a named logger that prints time, level, origin and message to
stderr exactly once however often it is requested, at the level named by the LOGLEVEL environment variable."""
import logging
import os
import sys

_FORMAT = "[%(asctime)s.%(msecs)03d][%(levelname)s][%(name)s.%(funcName)s:%(lineno)d] %(message)s"


def get_logger(name: str = None) -> logging.Logger:
    logger = logging.getLogger(name)
    if not getattr(logger, "_dedloc_configured", False):
        logger._dedloc_configured = True
        level = os.environ.get("LOGLEVEL", "INFO").upper()
        logger.setLevel(getattr(logging, level, logging.INFO))
        root_has_handlers = bool(logging.getLogger().handlers)
        if not logger.handlers and not root_has_handlers:        # do not double-print under logging.basicConfig
            handler = logging.StreamHandler(sys.stderr)
            handler.setFormatter(logging.Formatter(_FORMAT, datefmt="%Y/%m/%d %H:%M:%S"))
            logger.addHandler(handler)
            logger.propagate = False
    return logger
