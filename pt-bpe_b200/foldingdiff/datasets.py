"""`foldingdiff.datasets.FullCathCanonicalCoordsDataset` for the GeoBPE drivers (bin/encode.py:253-259,
bin/induce.py:179-183): PDB files -> list of structure dicts in the reference's selection and order
(datasets.py:213-288: glob of *.pdb then *.pdb.gz, optional `toy` prefix, chains shorter than min_length = 40 dropped,
then `default_rng(6489).shuffle`).  Backbone internal coordinates come from geobpe_b200.pdb (numpy); side chains,
C-beta and secondary structure are outside the hot path and stay None."""
import glob
import logging
import os
from pathlib import Path

import numpy as np
import pandas as pd

from geobpe_b200 import pdb as _pdb
from geobpe_b200.synthetic import COLUMNS


def featurize_one(fname):
    """one PDB file -> structure dict (datasets.py:108-164), None when the file holds no usable backbone"""
    try:
        xyz = _pdb.read_backbone(fname)
        if xyz.shape[0] < 2:
            logging.warning(f"Angles is None: {fname}")
            return None
        cols = _pdb.featurize_backbone(xyz)
        n = xyz.shape[0]
        return {
            "angles": pd.DataFrame({c: cols[i].copy() for i, c in enumerate(COLUMNS)}),
            "coords": xyz[:, 1].copy(),
            "c_beta": None,
            "full_coords": xyz.reshape(-1, 3).copy(),
            "full_idxes": sum([[i, i, i] for i in range(1, n + 1)], []),
            "side_chain": None,
            "aa": None,
            "fname": fname,
        }
    except Exception as e:          # the reference logs and drops files it cannot featurize (datasets.py:160-162)
        logging.warning(f"Featurization failed for {fname}: {e}")
        return None


class FullCathCanonicalCoordsDataset:
    feature_names = {"coords": list("xyz")}

    def __init__(self, pdbs="cath", split=None, pad=512, min_length=40, trim_strategy="leftalign", toy=0, zero_center=True,
                 use_cache=True, cache_dir=None, debug=False, secondary=False):
        assert pad > min_length
        self.pad, self.min_length, self.trim_strategy, self.debug, self.secondary = pad, min_length, trim_strategy, debug, secondary
        self.pdbs_src = pdbs
        if isinstance(pdbs, (list, tuple)):
            fnames = list(pdbs)
        elif Path(pdbs).is_dir():
            fnames = []
            for ext in (".pdb", ".pdb.gz"):
                fnames.extend(glob.glob(os.path.join(pdbs, f"*{ext}")))
            assert fnames, f"No PDB files found in {pdbs}"
        else:
            raise ValueError(f"Unknown pdb set: {pdbs} (this build takes a directory or a list of files)")
        if toy:
            fnames = fnames[:150 if isinstance(toy, bool) else toy]
        self.fnames = fnames
        self.structures = [s for s in (featurize_one(f) for f in fnames) if s is not None]
        if self.min_length:
            self.structures = [s for s in self.structures if s["angles"].shape[0] >= self.min_length]
        if self.trim_strategy == "discard":
            self.structures = [s for s in self.structures if s["angles"].shape[0] <= self.pad]
        self.rng = np.random.default_rng(seed=6489)
        self.rng.shuffle(self.structures)             # contiguous splits act like random splits (datasets.py:286-288)
        if split is not None:
            k = int(len(self.structures) * 0.8)
            t = int(len(self.structures) * 0.1)
            self.structures = {"train": self.structures[:k], "validation": self.structures[k:k + t],
                               "test": self.structures[k + t:]}[split]
        self.all_lengths = [s["angles"].shape[0] for s in self.structures]

    def __len__(self):
        return len(self.structures)

    def __getitem__(self, i):
        return self.structures[i]
