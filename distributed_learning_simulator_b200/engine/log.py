"""Process/thread-aware logging.

Replaces ``cyy_naive_lib.log`` (get_logger/add_file_handler, used throughout the
reference, e.g. /root/reference/simulation_lib/training.py:13,98).  Log lines carry the
thread name because executors rename their thread to ``worker <id>`` / ``server``
(reference behaviour: simulation_lib/executor.py:25-26) and the offline log scraper keys
on that prefix (simulation_lib/analysis/analyze_log.py:53-63).
"""
import logging
import os
import threading

_LOGGER_NAME = "dls_b200"
_lock = threading.Lock()
_FORMAT = "%(asctime)s %(levelname)s {%(threadName)s} [%(filename)s => %(lineno)d] : %(message)s"


def get_logger() -> logging.Logger:
    logger = logging.getLogger(_LOGGER_NAME)
    if not logger.handlers:
        with _lock:
            if not logger.handlers:
                handler = logging.StreamHandler()
                handler.setFormatter(logging.Formatter(_FORMAT))
                logger.addHandler(handler)
                logger.setLevel(os.environ.get("DLS_LOG_LEVEL", "INFO"))
                logger.propagate = False
    return logger


def set_level(level: str | int) -> None:
    get_logger().setLevel(level)


def add_file_handler(filename: str) -> logging.Handler:
    filename = os.path.abspath(filename)
    logger = get_logger()
    with _lock:
        for h in logger.handlers:
            if isinstance(h, logging.FileHandler) and h.baseFilename == filename:
                return h
        os.makedirs(os.path.dirname(filename), exist_ok=True)
        handler = logging.FileHandler(filename, encoding="utf8")
        handler.setFormatter(logging.Formatter(_FORMAT))
        logger.addHandler(handler)
    return handler


def remove_file_handler(handler: logging.Handler) -> None:
    logger = get_logger()
    with _lock:
        if handler in logger.handlers:
            logger.removeHandler(handler)
            handler.close()
