"""`foldingdiff.utils` names the GeoBPE drivers import (bin/encode.py:7 star-import, bin/induce.py:19): argparse type
helpers and the args.txt round trip (utils.py:17-112 of the reference).  Same argument meaning and error behaviour;
host-side only."""
import argparse
import ast
import glob  # noqa: F401  (star-exported like the reference's module namespace)
import hashlib  # noqa: F401
import logging  # noqa: F401
import os
import pickle
import re
from argparse import Namespace
from typing import *  # noqa: F401,F403

import numpy as np  # noqa: F401

_TRUE, _FALSE = ("yes", "true", "t", "y", "1"), ("no", "false", "f", "n", "0")


def str2bool(v):
    """utils.py:17-25"""
    if isinstance(v, bool):
        return v
    low = v.lower()
    if low in _TRUE:
        return True
    if low in _FALSE:
        return False
    raise argparse.ArgumentTypeError("Boolean value expected.")


def str2dict(v):
    """'1-100:10-10' -> {1: 100, 10: 10} (utils.py:27-35: size step -> bins / partitions)"""
    if re.match(r"\d+-\d+(?::\d+-\d+)*$", v) is None:
        raise argparse.ArgumentTypeError("Wrong format, see help.")
    return {int(a): int(b) for a, b in re.findall(r"(\d+)-(\d+)", v)}


def str2dictorint(v):
    """utils.py:37-41"""
    return int(v) if v.isdigit() else str2dict(v)


def int_or_inf(x):
    """utils.py:43-50: an integer or (case-insensitive) 'inf' / 'infinity'"""
    if x.lower() in ("inf", "infinity"):
        return float("inf")
    try:
        return int(x)
    except ValueError:
        raise argparse.ArgumentTypeError(f"‘{x}’ is not an integer or ‘inf’")


def load_args_from_txt(args_path):
    """lines 'name: value' -> Namespace, values parsed as python literals where possible (utils.py:53-72)"""
    ns = Namespace()
    with open(args_path, "r") as f:
        for raw in f:
            line = raw.rstrip("\n")
            if not line.strip() or ": " not in line:
                continue
            name, text = line.split(": ", 1)
            try:
                value = ast.literal_eval(text)
            except Exception:
                value = text
            setattr(ns, name, value)
    return ns


def validate_args_match(current, loaded, skip=None):
    """every field of `loaded` must equal the same field of `current` (paths compared absolute), utils.py:74-112"""
    skip = set(skip or [])
    get = (lambda o, k: o[k]) if isinstance(current, dict) else getattr
    has = (lambda o, k: k in o) if isinstance(current, dict) else hasattr
    items = loaded.items() if isinstance(loaded, dict) else vars(loaded).items()
    for name, saved in items:
        if name in skip:
            continue
        if not has(current, name):
            raise AssertionError(f"Current args has no field '{name}'")
        cur = get(current, name)
        if name.endswith("_dir") or name.endswith("_path"):
            saved, cur = os.path.abspath(saved), os.path.abspath(cur)
        if cur != saved:
            raise AssertionError(f"Mismatch for '{name}': current={cur!r}  saved={saved!r}")


def pickle_copy(t):
    return pickle.loads(pickle.dumps(t, protocol=pickle.HIGHEST_PROTOCOL))
