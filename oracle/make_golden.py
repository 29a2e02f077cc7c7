"""TEST INFRASTRUCTURE ONLY.  Generates tests/golden/*.json by running the REAL reference
(/root/reference, via oracle/ref_shim.py) on frozen synthetic inputs.  Runs only in the build
container (the GPU box has no /root/reference); the fixtures it writes are committed.

    python oracle/make_golden.py [case ...]
"""
import hashlib
import json
import os
import sys
import tempfile
import time
from collections import defaultdict

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.join(ROOT, "pt-bpe_b200"))
os.environ.setdefault("SLURM_CPUS_PER_TASK", "4")   # serial medoid path is broken in the reference

import numpy as np
import ref_shim
from geobpe_b200 import synthetic
from geobpe_oracle import dump_state, _jsonable

INF = float("inf")
NUM_P = {2: 2, 3: 5, 5: 1, 6: 2, 8: 1}
NUM_P4 = {2: 100, 3: 500, 5: 20, 6: 100, 8: 5, 9: 20, 11: 1, 12: 5, 14: 1}     # README cmd 2 (configs[3])

CASES = {
    # name: dict(gen=(S, n, seed, ragged), hparams, steps, [induce=(S, n, seed)], [sorted_sets])
    "h4": dict(gen=(40, 48, 0, False), steps=40,
               hp=dict(bins={1: 4}, bin_strategy="histogram", std_bonds=True, rmsd_partition_min_size=INF)),
    "h6_ragged": dict(gen=(30, 40, 3, True), steps=25,
                      hp=dict(bins={1: 6}, bin_strategy="histogram", std_bonds=True, rmsd_partition_min_size=INF)),
    "h5_uniform": dict(gen=(20, 32, 4, False), steps=15,
                       hp=dict(bins={1: 5}, bin_strategy="uniform", std_bonds=True, rmsd_partition_min_size=INF)),
    "h5_cover": dict(gen=(20, 32, 5, False), steps=15,
                     hp=dict(bins={1: 5}, bin_strategy="histogram-cover", std_bonds=True, rmsd_partition_min_size=INF)),
    "m20": dict(gen=(24, 40, 0, False), steps=25, induce=(6, 40, 1),
                hp=dict(bins={1: 20}, bin_strategy="histogram", std_bonds=False, rmsd_partition_min_size=0,
                        rmsd_super_res=True, num_partitions=NUM_P, max_num_strucs=100)),
    "m6_sub": dict(gen=(24, 40, 2, True), steps=20, induce=(4, 36, 7),
                   hp=dict(bins={1: 6}, bin_strategy="histogram", std_bonds=False, rmsd_partition_min_size=0,
                           rmsd_super_res=True, num_partitions={2: 2, 3: 3, 5: 2, 6: 2, 8: 1}, max_num_strucs=8)),
    "m20_sorted": dict(gen=(24, 40, 0, False), steps=25, sorted_sets=True,
                       hp=dict(bins={1: 20}, bin_strategy="histogram", std_bonds=False, rmsd_partition_min_size=0,
                               rmsd_super_res=True, num_partitions=NUM_P, max_num_strucs=100)),
    "m6_sub_sorted": dict(gen=(24, 40, 2, True), steps=20, sorted_sets=True,
                          hp=dict(bins={1: 6}, bin_strategy="histogram", std_bonds=False, rmsd_partition_min_size=0,
                                  rmsd_super_res=True, num_partitions={2: 2, 3: 3, 5: 2, 6: 2, 8: 1}, max_num_strucs=8)),
    # BASELINE configs[0] inputs: the 72 bundled smoke PDBs (15 640 residues), featurized by geobpe_b200/pdb.py; the arrays
    # are frozen in tests/golden/pdb72_input.npz because /root/reference does not travel
    "pdb72_h": dict(gen="pdb72", steps=int(os.environ.get("PDB72_H_STEPS", 500)),
                    hp=dict(bins={1: 10}, bin_strategy="histogram", std_bonds=True, rmsd_partition_min_size=INF)),
    "pdb72_m": dict(gen="pdb72", steps=int(os.environ.get("PDB72_M_STEPS", 500)), induce="pdb_heldout",
                    hp=dict(bins={1: 20}, bin_strategy="histogram", std_bonds=False, rmsd_partition_min_size=0,
                            rmsd_super_res=True, num_partitions=NUM_P, max_num_strucs=100)),
    # BASELINE configs[0] at the hyper-parameters README cmd 1 decodes to: bins {1: 50}, max_num_strucs 500
    "pdb72_m50": dict(gen="pdb72", steps=int(os.environ.get("PDB72_M50_STEPS", 500)), induce="pdb_heldout",
                      hp=dict(bins={1: 50}, bin_strategy="histogram", std_bonds=False, rmsd_partition_min_size=0,
                              rmsd_super_res=True, num_partitions=NUM_P, max_num_strucs=500)),
    # BASELINE configs[3] ("pareto") shape: bins {1: 500}, num_p 2-100:3-500:5-20:6-100:8-5:9-20:11-1:12-5:14-1; the residue
    # initialisation takes the k == N branch of k_medoids (algo.py:163-165: all 500 sampled residues are medoids), the
    # README cmd 2 stops there (max_iter 0: the reference cannot merge at these settings -- rmsd_partition needs at least
    # num_p occurrences of the popped key, bpe.py:1776 -- and induces against bpe_post_init.pkl, README.md:57)
    "cfg4_k500": dict(gen=(100, 16, 9, True), steps=0, induce=(3, 16, 10),
                      hp=dict(bins={1: 500}, bin_strategy="histogram", std_bonds=False, rmsd_partition_min_size=0,
                              rmsd_super_res=True, num_partitions=NUM_P4, max_num_strucs=500)),
}


def pdb72_inputs():
    """featurize the reference's bundled smoke PDBs and freeze the arrays as a fixture"""
    import glob
    from geobpe_b200 import pdb
    files = sorted(glob.glob("/root/reference/data/vqvae_pretrain/train/*.pdb"))
    vals, offs, names = pdb.featurize_files(files)
    path = os.path.join(ROOT, "tests", "golden", "pdb72_input.npz")
    np.savez_compressed(path, vals=vals, offsets=offs, names=np.array(names))
    return vals, offs


def pdb_heldout_inputs():
    """the reference's bundled CASP14 + CAMEO smoke PDBs (held out from training), frozen like pdb72"""
    import glob
    from geobpe_b200 import pdb
    files = sorted(glob.glob("/root/reference/data/vqvae_pretrain/CASP14/*.pdb")) + \
        sorted(glob.glob("/root/reference/data/vqvae_pretrain/CAMEO/*.pdb"))
    vals, offs, names = pdb.featurize_files(files)
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "pdb_heldout_input.npz"), vals=vals, offsets=offs, names=np.array(names))
    return vals, offs


class SortedIterSet(set):
    """set whose iteration is sorted: the disclosed 'sorted()' variant of SURVEY 7.2(1)(b),
    injected without touching reference source (list(s) at bpe.py:1744,1829,1888 iterates)."""

    def __iter__(self):
        return iter(sorted(set.__iter__(self)))


def tree_repr(node):
    if node.left is None and node.right is None:
        return _jsonable(node.value)
    return [_jsonable(node.value), tree_repr(node.left), tree_repr(node.right)]


def run_case(name, ref):
    c = CASES[name]
    if c["gen"] == "pdb72":
        vals, offs = pdb72_inputs()
    else:
        S, n, seed, ragged = c["gen"]
        vals, offs = synthetic.generate(S, n, seed=seed, ragged=ragged)
    strucs = synthetic.to_structures(vals, offs)
    d = tempfile.mkdtemp(prefix="golden_")
    hp = dict(c["hp"])
    t0 = time.time()
    bpe = ref.BPE(strucs, save_dir=d, res_init=True, seed=0, **hp)
    bpe.initialize()
    out = {"case": name, "gen": c["gen"] if isinstance(c["gen"], str) else list(c["gen"]), "hp": _jsonable({k: (str(v) if v == INF else v) for k, v in hp.items()}),
           "input_sha256": hashlib.sha256(vals.tobytes() + offs.tobytes()).hexdigest(),
           "sorted_sets": bool(c.get("sorted_sets"))}
    out["post_init"] = dump_state(bpe)
    out["post_init_ids"] = [bpe.quantize(t.tokenize()) for t in bpe.tokenizers[:3]]
    bpe.bin()
    if c.get("sorted_sets"):
        bpe._geo_dict = defaultdict(SortedIterSet, {k: SortedIterSet(v) for k, v in bpe._geo_dict.items()})
    out["n_keys_after_bin"] = len(bpe._geo_dict)
    out["n_pairs_after_bin"] = sum(len(v) for v in bpe._geo_dict.values())
    pops = []
    while bpe._step < c["steps"]:
        (flag, negc, key), _ = bpe._priority_dict.peekitem(0)
        before = bpe._step
        bpe.step()
        pops.append([int(-negc), key, bool(flag), bpe._step - before])
    out["pops"] = pops
    out["final"] = dump_state(bpe)
    out["final_ids"] = [bpe.quantize(t.tokenize()) for t in bpe.tokenizers[:3]]
    out["final_coords0"] = bpe.tokenizers[0].compute_coords().tolist()
    out["final_trees"] = [[tree_repr(nd) for nd in bpe.tokenizers[i].bond_to_token.tree.nodes.values()] for i in range(2)]
    out["vocab_size"] = int(bpe.vocab_size)
    if "induce" in c:
        if c["induce"] == "pdb_heldout":
            v2, o2 = pdb_heldout_inputs()
        else:
            S2, n2, seed2 = c["induce"]
            v2, o2 = synthetic.generate(S2, n2, seed=seed2)
        s2 = synthetic.to_structures(v2, o2)
        ind = []
        # a training backbone first (train<->induce consistency, bpe.py:1141-1143), then held-out ones
        for src in [strucs[0], strucs[1]] + s2:
            tk = ref.Tokenizer(ref.deepcopy(src))
            try:
                tk, metrics = bpe.tokenize(tk)
                ind.append({"ok": True, "segmentation": [_jsonable(v) for v in tk.bond_to_token.values()],
                            "L": [int(x) for x in metrics["L"]],
                            "ids": bpe.quantize(tk.tokenize())})
            except ValueError as e:
                ind.append({"ok": False, "error": str(e)[:60]})
        out["induce"] = ind
        out["induce_gen"] = c["induce"] if isinstance(c["induce"], str) else list(c["induce"])
    out["ref_seconds"] = round(time.time() - t0, 2)
    path = os.path.join(ROOT, "tests", "golden", f"{name}.json")
    with open(path, "w") as f:
        json.dump(out, f)
    print(name, "->", path, os.path.getsize(path) // 1024, "KiB", out["ref_seconds"], "s",
          "pops", [(p[0]) for p in pops][:12])


if __name__ == "__main__":
    ref = ref_shim.load_reference()
    # ProteinChain metrics inside BPE.tokenize are not part of parity: permissive stub
    for name in (sys.argv[1:] or list(CASES)):
        run_case(name, ref)
