"""`from foldingdiff.bpe import *` (bin/encode.py:5) -> the CUDA-backed BPE / Tokenizer plus the star-exports the
reference drivers rely on (bpe.py:1-4 -> tokenizer.py:2-15, utils.py:4-15, angles_and_coords.py:4-23)."""
import glob, gzip, json, logging, os, pickle, re, time  # noqa: F401,E401
from collections import defaultdict  # noqa: F401
from copy import deepcopy  # noqa: F401
from pathlib import Path  # noqa: F401

import numpy as np  # noqa: F401
import pandas as pd  # noqa: F401
import torch  # noqa: F401

from geobpe_b200.bpe import BPE, Tokenizer  # noqa: F401
from geobpe_b200.data_structures import ThresholdDict, TokenHierarchy, BinaryTreeBuilder, Node  # noqa: F401


class _Absent:
    """placeholder for the biotite names the reference star-exports (PDBFile, struc): only the drivers' unused helpers
    (encode.py:187-199 amino_acid_sequence) touch them"""

    def __init__(self, name):
        self._name = name

    def __getattr__(self, k):
        raise ImportError(f"{self._name} (biotite) is not part of the GeoBPE hot path and is not installed here")


PDBFile, struc = _Absent("PDBFile"), _Absent("biotite.structure")
logger = logging.getLogger(__name__)
