"""`.env` + environment configuration (ref include/utils/env.hpp:41-153, nn/train.hpp:80-100).

Same parser rules: ``KEY=VALUE`` lines, ``#`` comments, surrounding single/double quotes stripped,
values also exported to the process environment.  Lookup order: loaded file, then ``os.environ``."""
from __future__ import annotations

import os
import sys
from typing import Dict, TypeVar

T = TypeVar("T")
_loaded: Dict[str, str] = {}
_file_loaded = False


def load_env_file(file_path: str = "./.env") -> bool:
    global _file_loaded
    try:
        f = open(file_path)
    except OSError:
        print(f"Warning: Could not open .env file: {file_path}", file=sys.stderr)
        return False
    with f:
        for n, line in enumerate(f, 1):
            line = line.rstrip("\r\n")
            if not line or line[0] == "#":
                continue
            if "=" not in line:
                print(f"Warning: Invalid line {n} in {file_path}: {line}", file=sys.stderr)
                continue
            key, value = line.split("=", 1)
            key = key.strip(" \t")
            if len(value) >= 2 and value[0] == value[-1] and value[0] in "\"'":
                value = value[1:-1]
            _loaded[key] = value
            os.environ[key] = value
    _file_loaded = True
    print(f"Loaded {len(_loaded)} environment variables from {file_path}")
    return True


def _convert(s: str, default):
    if isinstance(default, bool):
        return s.strip().lower() in ("1", "true", "yes", "on")
    if isinstance(default, int):
        return int(s)
    if isinstance(default, float):
        return float(s)
    return s


def get_env(key: str, default: T) -> T:
    if key in _loaded:
        s = _loaded[key]
    elif key in os.environ:
        s = os.environ[key]
    else:
        return default
    try:
        return _convert(s, default)
    except ValueError:
        return default


def override(key: str, value) -> None:
    """Highest-priority setting (command-line flags): beats both the .env file and the inherited environment."""
    _loaded[key] = str(value)
    os.environ[key] = str(value)


def ensure_env_loaded(file_path: str = "./.env"):
    if not _file_loaded and os.path.exists(file_path):
        load_env_file(file_path)


def get_env_auto(key: str, default: T, env_file: str = "./.env") -> T:
    ensure_env_loaded(env_file)
    return get_env(key, default)


def clear_loaded():
    global _file_loaded
    _loaded.clear()
    _file_loaded = False
