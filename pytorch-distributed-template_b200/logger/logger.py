"""Logging setup (reference: logger/logger.py:7-22).  The json is found next to this file, so
file logging also works when the working directory is not the repository root."""
import logging
import logging.config
from pathlib import Path

from utils import read_json

_DEFAULT = Path(__file__).resolve().parent / "logger_config.json"


def setup_logging(save_dir, log_config=None, default_level=logging.INFO):
    cfg_path = Path(log_config) if log_config is not None else _DEFAULT
    if not cfg_path.is_file():
        print("Warning: logging configuration file is not found in {}.".format(cfg_path))
        logging.basicConfig(level=default_level)
        return
    cfg = read_json(cfg_path)
    for handler in cfg["handlers"].values():
        if "filename" in handler:
            handler["filename"] = str(Path(save_dir) / handler["filename"])
    logging.config.dictConfig(cfg)
