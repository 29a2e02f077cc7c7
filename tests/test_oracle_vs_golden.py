"""The CPU oracle (oracle/geobpe_oracle.py) against golden vectors produced by the REAL reference.

This is what pins the oracle: every token table, segmentation, pop order, threshold edge, bin
count, quantised id stream and reconstructed coordinate must equal the reference's output.
"""
import numpy as np
import pytest

import geobpe_oracle as O
from geobpe_b200 import synthetic
from golden_util import load_golden, case_inputs, case_hparams, induce_inputs, norm

H_CASES = ["h4", "h6_ragged", "h5_uniform", "h5_cover", "pdb72_h"]   # pdb72_*: BASELINE configs[0] inputs (real PDBs)
M_CASES = ["m20", "m6_sub", "m20_sorted", "m6_sub_sorted", "pdb72_m", "pdb72_m50", "cfg4_k500"]


def run_oracle(g):
    vals, offs = case_inputs(g)
    strucs = synthetic.to_structures(vals, offs)
    bpe = O.BPE(strucs, res_init=True, seed=0, sorted_sets=g["sorted_sets"], **case_hparams(g))
    bpe.initialize()
    return bpe, strucs


def check_learn(g):
    bpe, strucs = run_oracle(g)
    assert norm(O.dump_state(bpe)) == g["post_init"]
    assert [bpe.quantize(t) for t in bpe.tokenizers[:3]] == g["post_init_ids"]
    bpe.bin()
    assert len(bpe._geo_dict) == g["n_keys_after_bin"]
    assert sum(len(v) for v in bpe._geo_dict.values()) == g["n_pairs_after_bin"]
    steps = max(s["step"] for s in [g["final"]])
    while bpe._step < steps:
        bpe.step()
    assert [[c, k] for c, k in bpe.pop_log] == [[p[0], p[1]] for p in g["pops"]]
    assert norm(O.dump_state(bpe)) == g["final"]
    assert [bpe.quantize(t) for t in bpe.tokenizers[:3]] == g["final_ids"]
    np.testing.assert_allclose(bpe.tokenizers[0].compute_coords(), np.array(g["final_coords0"]), rtol=0, atol=1e-9)
    assert bpe.vocab_size == g["vocab_size"]
    return bpe, strucs


@pytest.mark.parametrize("name", H_CASES)
def test_oracle_h_mode(name):
    check_learn(load_golden(name))


@pytest.mark.parametrize("name", M_CASES)
def test_oracle_m_mode(name):
    g = load_golden(name)
    bpe, strucs = check_learn(g)
    if "induce" not in g:
        return
    v2, o2 = induce_inputs(g)
    s2 = synthetic.to_structures(v2, o2)
    for src, want in zip([strucs[0], strucs[1]] + s2, g["induce"]):
        tk = O.Tokenizer.from_structure(src)
        try:
            tk, metrics = bpe.tokenize(tk)
        except ValueError:
            assert not want["ok"]
            continue
        assert want["ok"]
        assert norm(list(tk.bond_to_token.values())) == want["segmentation"]
        assert metrics["L"] == want["L"]
        assert bpe.quantize(tk) == want["ids"]
