"""ConfigParser: JSON config + CLI overrides + plug-in construction by name
(reference: parse_config.py:13-156; same public surface: from_args, init_obj, init_ftn,
[] access, get_logger, save_dir / log_dir / resume)."""
import logging
from datetime import datetime
from functools import partial, reduce
from operator import getitem
from pathlib import Path

from logger import setup_logging
from utils import read_json, write_json
from utils.dist import is_main_process

_LEVELS = {0: logging.WARNING, 1: logging.INFO, 2: logging.DEBUG}


class ConfigParser:
    def __init__(self, config, resume=None, modification=None, run_id=None, training=True):
        self._config = _apply_overrides(config, modification)
        self.resume = resume
        if run_id is None:
            run_id = datetime.now().strftime(r"%m%d_%H%M%S")
        root = Path(self.config["trainer"]["save_dir"])
        self._save_dir = root / self.config["name"] / ("train" if training else "test") / run_id
        self.save_dir.mkdir(parents=True, exist_ok=True)
        if is_main_process():
            write_json(self.config, self.save_dir / "config.json")
        setup_logging(self.save_dir)
        self.log_levels = dict(_LEVELS)

    @classmethod
    def from_args(cls, args, options="", training=True):
        for opt in options:
            args.add_argument(*opt.flags, default=None, type=opt.type)
        if not isinstance(args, tuple):
            args = args.parse_args()
        if args.resume is not None:
            resume = Path(args.resume)
            cfg_fname = resume.parent / "config.json"
        else:
            assert args.config is not None, \
                "Configuration file need to be specified. Add '-c config.json', for example."
            resume = None
            cfg_fname = Path(args.config)
        config = read_json(cfg_fname)
        if args.config and resume:
            config.update(read_json(args.config))  # fine-tuning: new config over the checkpoint's
        if getattr(args, "save_dir", None) is not None:
            config["trainer"]["save_dir"] = args.save_dir
        modification = {opt.target: getattr(args, _flag_name(opt.flags)) for opt in options}
        return args, cls(config, resume, modification, training=training)

    def _spec(self, name, kwargs):
        entry = self[name]
        merged = dict(entry["args"])
        assert all(k not in merged for k in kwargs), "Overwriting kwargs given in config file is not allowed"
        merged.update(kwargs)
        return entry["type"], merged

    def init_obj(self, name, module, *args, **kwargs):
        """config.init_obj('optimizer', torch.optim, params) == torch.optim.<type>(params, **args)."""
        type_name, merged = self._spec(name, kwargs)
        return getattr(module, type_name)(*args, **merged)

    def init_ftn(self, name, module, *args, **kwargs):
        type_name, merged = self._spec(name, kwargs)
        return partial(getattr(module, type_name), *args, **merged)

    def __getitem__(self, name):
        return self.config[name]

    def get_logger(self, name, verbosity=2):
        assert verbosity in self.log_levels, \
            "verbosity option {} is invalid. Valid options are {}.".format(verbosity, self.log_levels.keys())
        logger = logging.getLogger(name)
        logger.setLevel(self.log_levels[verbosity])
        return logger

    @property
    def config(self):
        return self._config

    @property
    def save_dir(self):
        return self._save_dir

    @property
    def log_dir(self):
        return self._save_dir


def _apply_overrides(config, modification):
    for path, value in (modification or {}).items():
        if value is not None:
            keys = path.split(";")
            reduce(getitem, keys[:-1], config)[keys[-1]] = value
    return config


def _flag_name(flags):
    for flag in flags:
        if flag.startswith("--"):
            return flag[2:].replace("-", "_")
    return flags[0].lstrip("-")
