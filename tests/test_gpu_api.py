"""GPU tests of the reference-facing API surface beyond learn / induce (b):
token_geo / set_token_geo / key_coords / windowed compute_coords against the CPU oracle, the token-stream round trip
(quantize -> dequantize -> recover -> recover_structure, the reference notebook's invariant), capacity, checkpoint resume
(bin/encode.py:239-260) in both modes, and a replay of the two drivers' call sequences through the foldingdiff.* namespace."""
import json
import os
import pickle

import numpy as np
import pytest

import geobpe_oracle as O
from geobpe_b200 import synthetic
from golden_util import norm

pytestmark = pytest.mark.gpu
INF = float("inf")
HP_H = dict(bins={1: 5}, bin_strategy="histogram", std_bonds=True, rmsd_partition_min_size=INF)
HP_M = dict(bins={1: 10}, bin_strategy="histogram", std_bonds=False, rmsd_partition_min_size=0, rmsd_super_res=True,
            num_partitions={2: 2, 3: 4, 5: 2, 6: 2, 8: 1}, max_num_strucs=50)


def learn(hp, S=24, n=40, seed=31, steps=25, cls=None):
    from foldingdiff.bpe import BPE
    vals, offs = synthetic.generate(S, n, seed=seed, ragged=True)
    strucs = synthetic.to_structures(vals, offs)
    bpe = (cls or BPE)(strucs, res_init=True, seed=0, **hp)
    bpe.initialize(); bpe.bin()
    for _ in range(steps):
        bpe.step()
    return bpe, strucs


@pytest.mark.parametrize("hp", [HP_H, HP_M], ids=["H", "M"])
def test_window_geometry_vs_oracle(hp):
    bpe, strucs = learn(hp)
    ora, _ = learn(hp, cls=O.BPE)
    rng = np.random.default_rng(0)
    for i in (0, 3, 7):
        t, to = bpe.tokenizers[i], ora.tokenizers[i]
        nb = 3 * t.n - 1
        assert t.token_geo(0, nb) == to.token_geo(0, nb)
        assert t.token_geo(0, nb, orig=True) == to.token_geo(0, nb, orig=True)
        for _ in range(6):
            idx = int(rng.integers(0, nb - 3))
            l = int(rng.integers(2, nb - idx + 1))
            assert t.token_geo(idx, l) == to.token_geo(idx, l)
            for orig in (False, True):
                np.testing.assert_allclose(t.compute_coords(idx, l, orig=orig), to.compute_coords(idx, l, orig=orig), rtol=0, atol=1e-9)
        with pytest.raises(ValueError):
            t.token_geo(nb - 1, 3)
    from foldingdiff.tokenizer import Tokenizer
    for key in list(bpe._tokens.values())[-4:]:
        if isinstance(next(iter(key.values()))[0], float):        # medoid geometries (histogram keys hold bin indices)
            np.testing.assert_allclose(Tokenizer.key_coords(key), O.key_coords(key), rtol=0, atol=1e-9)


@pytest.mark.parametrize("hp", [HP_H, HP_M], ids=["H", "M"])
def test_token_stream_round_trip(hp):
    """the reference notebook's invariant recover(t.tokenize()) == token_geo(0, 3n-1), ids -> tokens -> ids, the recovered
    structure's coordinates, capacity against the reference's per-tokenizer formula"""
    from foldingdiff.tokenizer import Tokenizer
    bpe, _ = learn(hp)
    for i in (0, 5):
        t = bpe.tokenizers[i]
        tokenized = t.tokenize()
        geo = t.token_geo(0, 3 * t.n - 1)
        if bpe._mode == "M":
            assert norm(bpe.recover(tokenized)) == norm(geo)
        ids = bpe.quantize(tokenized)
        back = bpe.dequantize(ids)
        assert [x[:2] for x in back] == [x[:2] for x in tokenized]
        for a, b in zip(back, tokenized):
            if a[0] != "MOTIF":
                assert abs(a[2] - (b[2] + 2 * np.pi) % (2 * np.pi)) < 1e-12       # glue values are bin centres already
        assert bpe.quantize(back) == ids
        if bpe._mode == "M":
            t2 = bpe.recover_structure(bpe.recover(back), back)
            assert [v[1:] for v in t2.bond_to_token.values()] == [v[1:] for v in t.bond_to_token.values()]
            want = t.compute_coords()
            got = t2.compute_coords()
            from geobpe_b200.engine import INIT_N_CA
            assert got.shape == want.shape
            # (the recovered structure starts from the default first-residue geometry like the reference's init_structure)
            if abs(t._init_n_ca - INIT_N_CA) < 1e-12:
                np.testing.assert_allclose(got, want, rtol=0, atol=1e-6)
    total = 0
    for token in bpe._tokens.values():
        n = Tokenizer.num_bonds(token)
        total += 4 * (n + n - 1 + n - 2) * 8
    assert bpe.capacity() == total
    mb, bb = np.log2(len(bpe._tokens)), np.log2(bpe.bins[1])
    for t in bpe.tokenizers:
        m = (len(t.tokenize()) + 3) // 4
        total += mb * m + 3 * (m - 1) * bb
    assert abs(bpe.capacity(tokenizer=True) - total) < 1e-6 * total
    assert set(bpe._bin_centers[1]) == set(bpe._thresholds[1]) - set() and bpe._bin_weights[1]["omega"].sum().item() == pytest.approx(1.0)


@pytest.mark.parametrize("hp", [HP_H, HP_M], ids=["H", "M"])
def test_resume_from_pickle_continues_bit_exact(hp, tmp_path):
    """bin/encode.py:239-260: the newest bpe_iter={t}.pkl is loaded and stepping continues"""
    whole, _ = learn(hp, steps=40)
    part, _ = learn(hp, steps=22)
    path = tmp_path / "bpe_iter=21.pkl"
    with open(path, "wb") as f:
        pickle.dump(part, f)
    del part
    with open(path, "rb") as f:
        bpe = pickle.load(f)
    assert bpe._step == 22
    seg_before = [list(t.bond_to_token.values()) for t in bpe.tokenizers[:3]]        # readable without the device state
    while bpe._step < whole._step:
        bpe.step()
    assert [list(p) for p in bpe.pop_log] == [list(p) for p in whole.pop_log]
    assert norm(O.dump_state(bpe, max_angles=24)) == norm(O.dump_state(whole, max_angles=24))
    assert seg_before != [list(t.bond_to_token.values()) for t in bpe.tokenizers[:3]]
    trees = lambda b: [[norm(nd.value) for nd in b.tokenizers[i].bond_to_token.tree.nodes.values()] for i in range(3)]
    assert trees(bpe) == trees(whole)


def test_driver_call_sequences(tmp_path):
    """bin/encode.py:202-458 and bin/induce.py:146-240 replayed call by call through the foldingdiff.* namespace (the drivers
    themselves import imageio / esm / matplotlib, absent here): dataset-shaped structures -> BPE -> bpe_init.pkl ->
    initialize -> bpe_post_init.pkl -> stats -> bin -> step / save_every pickles / stats -> resume; then induce with --append"""
    from foldingdiff.bpe import BPE, torch
    from foldingdiff.plotting import get_codebook_utility, plot
    from foldingdiff.tokenizer import Tokenizer
    from foldingdiff.utils import load_args_from_txt, validate_args_match, str2dict, str2dictorint, int_or_inf
    import argparse
    save = str(tmp_path)
    args = argparse.Namespace(bins=str2dict("1-10"), p_min_size=int_or_inf("0"), num_p=str2dictorint("2-2:3-4:5-2:6-2:8-1"),
                              save_dir=save, max_iter=12, save_every=5)
    with open(os.path.join(save, "args.txt"), "w") as f:
        for k, v in sorted(vars(args).items()):
            f.write(f"{k}: {v}\n")
    validate_args_match(current=args, loaded=load_args_from_txt(os.path.join(save, "args.txt")), skip=["save_dir", "max_iter"])
    vals, offs = synthetic.generate(20, 44, seed=41, ragged=True)
    strucs = synthetic.to_structures(vals, offs)
    bpe = BPE(strucs, bins=args.bins, bin_strategy="histogram", save_dir=save, rmsd_partition_min_size=args.p_min_size,
              rmsd_super_res=True, rmsd_only=False, num_partitions=args.num_p, max_num_strucs=40, compute_sec_structs=False,
              plot_iou_with_sec_structs=False, res_init=True, std_bonds=False, glue_opt=False, glue_opt_prior=0.0,
              glue_opt_method="all", glue_opt_every=10, seed=0)
    pickle.dump(bpe, open(os.path.join(save, "bpe_init.pkl"), "wb+"))
    N = len(bpe.tokenizers)
    ref_coords = [bpe.tokenizers[i].compute_coords() for i in range(3)]
    bpe.tokenizers[0].visualize(os.path.join(save, "x.png"), vis_dihedral=False)
    bpe.initialize(path=os.path.join(save, "hist_plot.png"))
    pickle.dump(bpe, open(os.path.join(save, "bpe_post_init.pkl"), "wb+"))

    def stats():
        K = len(bpe._tokens)
        Lm = np.mean([len(t.bond_to_token) for t in bpe.tokenizers])
        input_ids = bpe.quantize(bpe.tokenizers)
        utility = get_codebook_utility(torch.as_tensor(sum(input_ids, [])), bpe.vocab_size)
        return {"K": K, "L": Lm, "bpr": bpe.capacity(tokenizer=True) / (N * Lm)} | utility
    s0 = stats()
    assert s0["K"] > 0 and 0 < s0["use_ratio"] <= 1
    for i in range(3):
        assert bpe.tokenizers[i].compute_coords().shape == ref_coords[i].shape
    bpe.bin()
    for t in range(args.max_iter):
        bpe.step()
        assert bpe._step == t + 1
        bpe.tokenizers[0].bond_to_token.tree.visualize(os.path.join(save, "t.png"), horizontal_gap=0.5, font_size=6)
        if t % args.save_every == 0:
            pickle.dump(bpe.tokenizers[:3], open(os.path.join(save, f"ref_tokenizers={t}.pkl"), "wb+"))
            json.dump(stats(), open(os.path.join(save, f"stats={t}.json"), "w+"))
            bpe.plot_times(os.path.join(save, "times.png"))
            pickle.dump(bpe, open(os.path.join(save, f"bpe_iter={t}.pkl"), "wb+"))
            plot(bpe, len(bpe._tokens), ref_coords, os.path.join(save, "run.png"), prev_iter=t - 5, no_iters=t, step_iter=5, ratio=N / 1000)
    full_log = [list(p) for p in bpe.pop_log]
    # resume from the newest checkpoint (encode.py:224-245) and finish the run again
    b2 = pickle.load(open(os.path.join(save, "bpe_iter=10.pkl"), "rb"))
    for t in range(11, args.max_iter):
        b2.step()
        assert b2._step == t + 1
    assert [list(p) for p in b2.pop_log] == full_log
    # induce.py: load the pickle, tokenize held-out structures one by one, store (stats, tok), quantize, --append
    src = pickle.load(open(os.path.join(save, "bpe_iter=10.pkl"), "rb"))
    hv, ho = synthetic.generate(4, 40, seed=43)
    toks = []
    for idx, struc in enumerate(synthetic.to_structures(hv, ho)):
        try:
            tok, st = src.tokenize(Tokenizer(struc))
        except ValueError:
            continue
        pickle.dump((st, tok), open(os.path.join(save, f"{idx}.pkl"), "wb+"))
        st2, tok2 = pickle.load(open(os.path.join(save, f"{idx}.pkl"), "rb"))
        assert st2["L"][0] == tok2.n and list(tok2.bond_to_token.values()) == list(tok.bond_to_token.values())
        toks.append(tok2)
    ids = [src.quantize(t.tokenize()) for t in toks]
    if ids:
        u = get_codebook_utility(torch.as_tensor(sum(ids, [])), src.vocab_size)
        assert 0 < u["use_ratio"] <= 1
    if not isinstance(src.n, list):
        src.n = [src.n]
    src.n.append(len(toks))
    src.tokenizers.extend(toks)
    pickle.dump(src, open(os.path.join(save, "out.pkl"), "wb+"))
    out = pickle.load(open(os.path.join(save, "out.pkl"), "rb"))
    assert out.n == [N, len(toks)] and len(out.tokenizers) == N + len(toks)


def test_induce_device_replay_equals_host_replay(monkeypatch):
    """K7 (csrc/k_induce.cu: all vocabulary merges replayed on the device, in-place merges on lazily compacted arrays) against
    the host-driven replay (one K2 compaction + read-back per merge) on a held-out batch: token arrays, merge events, L trace"""
    from geobpe_b200 import engine as E
    bpe, _ = learn(HP_M, S=40, n=48, seed=51, steps=60)
    hv, ho = synthetic.generate(30, 56, seed=52, ragged=True)
    outs = []
    for host in ("0", "1"):
        monkeypatch.setenv("GEOBPE_INDUCE_HOST", host)
        ind = E.Inducer(hv, ho, bpe._tokens, bpe._sphere_dict, bpe._thresholds, bpe.std_bonds)
        ts = ind.ts
        n = ind.ev_n
        outs.append(([E._u32(ts.a[f][:ts.T]).tolist() for f in ("pos", "len", "tid", "cls", "glue")],
                     E._u32(ind.ev_left[:n]).tolist(), E._u32(ind.ev_right[:n]).tolist(), E._u32(ind.ev_p[:n]).tolist(),
                     [c for _, c in ind.replay], ind.n_steps, ind.failed.tolist()))
    assert outs[0][4] == outs[1][4], "events per vocabulary merge differ"
    assert outs[0] == outs[1]
    assert outs[0][5] > 10


@pytest.mark.parametrize("hp", [HP_H, HP_M], ids=["H", "M"])
def test_token_stream_export_equals_per_token_quantize(hp):
    """f-3: one device pass over the token arrays against the reference-shaped per-token host loop, + codebook statistics"""
    import torch
    from foldingdiff.plotting import get_codebook_utility
    bpe, _ = learn(hp, S=30, n=44, seed=61, steps=35)
    want = [bpe.quantize(t.tokenize()) for t in bpe.tokenizers]
    got = bpe.quantize(bpe.tokenizers)
    assert got == want
    ids, offs = bpe.token_stream()
    assert offs[-1] == ids.numel() == sum(len(w) for w in want)
    u = bpe.codebook_utility()
    v = get_codebook_utility(torch.as_tensor(sum(want, [])), bpe.vocab_size)
    assert u.keys() == v.keys() and all(abs(u[k] - v[k]) < 1e-5 for k in u)


@pytest.mark.parametrize("hp", [HP_H, HP_M], ids=["H", "M"])
def test_reconstruction_metrics_for_all_backbones(hp):
    """f-4: full-chain NeRF + Kabsch backbone RMSD of every backbone on the device against the per-tokenizer path (compute_coords
    of the materialised tokenizer + the oracle's Kabsch RMSD); north_star tolerance 1e-4 A, asserted at 1e-7"""
    from foldingdiff.tokenizer import Tokenizer
    bpe, strucs = learn(hp, S=16, n=36, seed=71, steps=20)
    rmsd, c_cur, c_org = bpe._learner.reconstruction()
    rmsd, c_cur, c_org = rmsd.cpu().numpy(), c_cur.cpu().numpy(), c_org.cpu().numpy()
    offs = bpe._offsets
    assert rmsd.shape == (len(bpe.tokenizers),) and np.all(np.isfinite(rmsd)) and rmsd.max() > 0
    for i in (0, 3, 9, 15):
        t = bpe.tokenizers[i]
        a, b = 3 * int(offs[i]), 3 * int(offs[i + 1])
        cur = t.compute_coords()
        org = Tokenizer(strucs[i]).compute_coords()       # the pristine structure (encode.py:308 takes ref_coords before initialize())
        np.testing.assert_allclose(c_cur[a:b], cur, rtol=0, atol=1e-7)
        np.testing.assert_allclose(c_org[a:b], org, rtol=0, atol=1e-7)
        want = O.compute_rmsd(cur, org)
        assert abs(rmsd[i] - want) < 1e-7
    assert np.array_equal(bpe.reconstruction_rmsd(), rmsd)


@pytest.mark.parametrize("hp", [HP_H, HP_M], ids=["H", "M"])
def test_lddt_ca_for_all_backbones(hp):
    """f-4: lDDT-C-alpha of every backbone on the device against the oracle's restatement of esm's compute_lddt on the same
    coordinates (float32 arithmetic on both sides; summation order differs, so 1e-5 on a value in [0, 1])"""
    bpe, strucs = learn(hp, S=12, n=70, seed=73, steps=25)
    L = bpe._learner
    rmsd, c_cur, c_org = L.reconstruction()
    got = L.lddt_ca(coords=(c_cur, c_org)).cpu().numpy()
    c_cur, c_org = c_cur.cpu().numpy(), c_org.cpu().numpy()
    offs = bpe._offsets
    assert got.shape == (len(bpe.tokenizers),)
    for i in range(len(bpe.tokenizers)):
        a, b = 3 * int(offs[i]), 3 * int(offs[i + 1])
        want = O.lddt_ca(c_cur[a:b][1::3], c_org[a:b][1::3])
        assert abs(got[i] - want) < 1e-5, (i, got[i], want)
    assert 0.0 < got.min() and got.max() <= 1.0 and got.min() < 1.0        # quantisation moved something, nothing is degenerate
    assert np.array_equal(bpe.reconstruction_lddt(), got)
    # a tighter inclusion radius changes the pair set (and is honoured)
    g8 = L.lddt_ca(cutoff=8.0, coords=L.reconstruction()[1:]).cpu().numpy()
    a, b = 3 * int(offs[0]), 3 * int(offs[1])
    assert abs(g8[0] - O.lddt_ca(c_cur[a:b][1::3], c_org[a:b][1::3], cutoff=8.0)) < 1e-5
