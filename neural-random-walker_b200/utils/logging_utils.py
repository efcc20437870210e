"""Logger "NRW" set-up -- same name, format and file/stream handlers as the reference's
utils/logging_utils.py:23-73; ``coloredlogs`` and NxSDK's logger are optional here (the reference
calls coloredlogs.install unconditionally, :65, and fails without the package)."""
import logging
import os
import sys

try:
    import coloredlogs
except ImportError:
    coloredlogs = None

rw_logger = logging.getLogger("NRW")

LOG_FORMAT = "%(asctime)s %(levelname)-4s:%(name)s %(relativeCreated)12d ms - %(message)s"
DATE_FORMAT = "%m/%d/%Y %I:%M:%S %p"


class CustomFormatter(logging.Formatter):
    """ANSI colour per level around the common format (logging_utils.py:23-45)."""
    COLOURS = {logging.DEBUG: "\x1b[32;1m", logging.INFO: "\x1b[38;20m",
               logging.WARNING: "\x1b[33;20m", logging.ERROR: "\x1b[31;20m",
               logging.CRITICAL: "\x1b[31;1m"}
    RESET = "\x1b[0m"

    def format(self, record):
        colour = self.COLOURS.get(record.levelno, "")
        return logging.Formatter(colour + LOG_FORMAT + self.RESET, datefmt=DATE_FORMAT).format(record)


def setup_logging(log_file='log.file', useLogFile=False, log_level=logging.DEBUG):
    os.environ["COLOREDLOGS_DATE_FORMAT"] = DATE_FORMAT
    os.environ["COLOREDLOGS_LOG_FORMAT"] = LOG_FORMAT
    logging.root.setLevel(log_level)
    for h in list(rw_logger.handlers):          # idempotent: run() may be called repeatedly
        rw_logger.removeHandler(h)
    if useLogFile:
        fh = logging.FileHandler(log_file)
        fh.setFormatter(CustomFormatter())
        fh.setLevel(log_level)
        rw_logger.addHandler(fh)
    sh = logging.StreamHandler(sys.stdout)
    sh.setFormatter(CustomFormatter())
    sh.setLevel(log_level)
    rw_logger.addHandler(sh)
    rw_logger.propagate = False
    if coloredlogs is not None:
        coloredlogs.install(level=log_level, logger=rw_logger, formatter=sh.formatter,
                            stream=sys.stdout, isatty=True)
