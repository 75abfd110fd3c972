import logging
import os

_FORMAT = "%(asctime)s %(levelname)s {%(processName)s} [%(filename)s => %(lineno)d] : %(message)s"


def get_logger() -> logging.Logger:
    logger = logging.getLogger("reference")
    if not logger.handlers:
        h = logging.StreamHandler()
        h.setFormatter(logging.Formatter(_FORMAT))
        logger.addHandler(h)
        logger.setLevel(os.environ.get("REF_LOG_LEVEL", "INFO"))
        logger.propagate = False
    return logger


def add_file_handler(filename: str) -> None:
    os.makedirs(os.path.dirname(os.path.abspath(filename)), exist_ok=True)
    h = logging.FileHandler(filename)
    h.setFormatter(logging.Formatter(_FORMAT))
    get_logger().addHandler(h)


def set_level(level) -> None:
    get_logger().setLevel(level)
