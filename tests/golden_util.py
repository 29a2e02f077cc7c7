"""Shared helpers for the golden-vector tests (tests/golden/*.json were produced by the real
reference through oracle/ref_shim.py + oracle/make_golden.py)."""
import json
import os

import numpy as np

from geobpe_b200 import synthetic

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
INF = float("inf")


def load_golden(name):
    with open(os.path.join(GOLDEN_DIR, f"{name}.json")) as f:
        return json.load(f)


def case_inputs(g):
    if isinstance(g["gen"], str):            # frozen real-data arrays (tests/golden/<gen>_input.npz)
        z = np.load(os.path.join(GOLDEN_DIR, f"{g['gen']}_input.npz"))
        return z["vals"], z["offsets"]
    S, n, seed, ragged = g["gen"]
    vals, offs = synthetic.generate(S, n, seed=seed, ragged=bool(ragged))
    return vals, offs


def induce_inputs(g):
    """held-out backbones of a golden's induce section"""
    if isinstance(g["induce_gen"], str):
        z = np.load(os.path.join(GOLDEN_DIR, f"{g['induce_gen']}_input.npz"))
        return z["vals"], z["offsets"]
    S2, n2, seed2 = g["induce_gen"]
    return synthetic.generate(S2, n2, seed=seed2)


def case_hparams(g):
    hp = dict(g["hp"])
    hp["bins"] = {int(k): v for k, v in hp["bins"].items()}
    if hp.get("rmsd_partition_min_size") == "inf":
        hp["rmsd_partition_min_size"] = INF
    if isinstance(hp.get("num_partitions"), dict):
        hp["num_partitions"] = {int(k): v for k, v in hp["num_partitions"].items()}
    return hp


def norm(x):
    """JSON-normalise (tuples -> lists, numpy scalars -> python) for comparisons."""
    if isinstance(x, dict):
        return {str(k): norm(v) for k, v in x.items()}
    if isinstance(x, (list, tuple)):
        return [norm(v) for v in x]
    if isinstance(x, np.integer):
        return int(x)
    if isinstance(x, np.floating):
        return float(x)
    return x
