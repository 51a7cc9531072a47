"""Process-wide logger (API parity with reference ``metisfl/utils/metis_logger.py:35-75``)."""
import logging
import sys
import time


class MetisASCIIArt(object):
    BANNER = r"""
  __  __      _   _      ___ _       _    ___ __   __
 |  \/  | ___| |_(_)___ | __| |     | |__|_  )  \ /  \
 | |\/| |/ -_)  _| (_-< | _|| |__   | '_ \/ / () | () |
 |_|  |_|\___|\__|_/__/ |_| |____|  |_.__/___\__/ \__/   federated learning on Blackwell
"""

    @classmethod
    def print(cls):
        sys.stdout.write(cls.BANNER + "\n")
        sys.stdout.flush()


class _MicrosecondFormatter(logging.Formatter):
    def formatTime(self, record, datefmt=None):
        ct = self.converter(record.created)
        return "%s.%03d" % (time.strftime("%Y-%m-%d %H:%M:%S", ct), record.msecs)


class MetisLogger(object):
    _logger = None

    @classmethod
    def getlogger(cls):
        if cls._logger is None:
            logger = logging.getLogger("Metis")
            if not logger.handlers:
                h = logging.StreamHandler(sys.stderr)
                h.setFormatter(_MicrosecondFormatter(
                    "%(asctime)s %(levelname)s pid=%(process)d %(filename)s:%(lineno)d] %(message)s"))
                logger.addHandler(h)
                logger.setLevel(logging.INFO)
                logger.propagate = False
            cls._logger = logger
        return cls._logger

    @classmethod
    def log(cls, msg, *args, **kwargs):
        cls.getlogger().log(logging.INFO, msg, *args, stacklevel=2, **kwargs)

    @classmethod
    def debug(cls, msg, *args, **kwargs):
        cls.getlogger().debug(msg, *args, stacklevel=2, **kwargs)

    @classmethod
    def error(cls, msg, *args, **kwargs):
        cls.getlogger().error(msg, *args, stacklevel=2, **kwargs)

    @classmethod
    def fatal(cls, msg, *args, **kwargs):
        cls.getlogger().critical(msg, *args, stacklevel=2, **kwargs)

    @classmethod
    def info(cls, msg, *args, **kwargs):
        cls.getlogger().info(msg, *args, stacklevel=2, **kwargs)

    @classmethod
    def warning(cls, msg, *args, **kwargs):
        cls.getlogger().warning(msg, *args, stacklevel=2, **kwargs)
