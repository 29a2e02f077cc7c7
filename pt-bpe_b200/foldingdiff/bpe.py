"""`from foldingdiff.bpe import *` (bin/encode.py:5) -> the CUDA-backed BPE / Tokenizer plus the star-exports the
reference drivers rely on."""
import glob, gzip, json, logging, re, time  # noqa: F401,E401
from pathlib import Path  # noqa: F401

import numpy as np  # noqa: F401
import torch  # noqa: F401

from geobpe_b200.bpe import BPE, Tokenizer  # noqa: F401
from geobpe_b200.data_structures import ThresholdDict, TokenHierarchy, BinaryTreeBuilder, Node  # noqa: F401
