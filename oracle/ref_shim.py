"""TEST INFRASTRUCTURE ONLY -- never imported by the product path.

Imports the *real* reference (read-only at /root/reference) inside THIS container so that
golden vectors can be generated from the reference's own arithmetic.
The GPU box has no /root/reference: nothing in `-m gpu` tests, smoke() or bench.py imports
this module.

What the shim does (and nothing else):
  1. registers permissive stub modules for packages the reference imports at module scope
     but that are absent here (matplotlib, seaborn, astropy, esm, biotite, imageio, ...);
  2. puts /root/reference on sys.path and imports foldingdiff.bpe;
  3. replaces the three pandas chained-assignment setters (tokenizer.py:139-167), which are
     silent no-ops under pandas>=3 copy-on-write, by single-step .iloc writes with the same
     index arithmetic;
  4. turns the visualisation methods into no-ops (no graphics stack on either box).
"""
import os
import sys
import types
import importlib.machinery

REFERENCE_ROOT = os.environ.get("GEOBPE_REFERENCE_ROOT", "/root/reference")


class _Anything:
    """Callable, attribute-permissive, subclassable placeholder."""

    def __init__(self, *a, **k):
        pass

    def __call__(self, *a, **k):
        return _Anything()

    def __getattr__(self, name):
        if name.startswith("__") and name.endswith("__"):
            raise AttributeError(name)
        return _Anything()

    def __iter__(self):
        return iter(())

    def __len__(self):
        return 0

    def __mro_entries__(self, bases):
        return (object,)


class _StubModule(types.ModuleType):
    def __getattr__(self, name):
        if name.startswith("__") and name.endswith("__"):
            raise AttributeError(name)
        if name[:1].isupper():
            # classes may be used as base classes (plotting.py:45 HandlerBase)
            return type(name, (), {"__init__": lambda self, *a, **k: None})
        return _Anything()


_STUBS = [
    "matplotlib", "matplotlib.pyplot", "matplotlib.lines", "matplotlib.colors",
    "matplotlib.patheffects", "matplotlib.patches", "matplotlib.legend_handler",
    "matplotlib.cm", "matplotlib.animation", "matplotlib.ticker", "matplotlib.gridspec",
    "mpl_toolkits", "mpl_toolkits.mplot3d", "mpl_toolkits.mplot3d.art3d",
    "seaborn", "astropy", "astropy.visualization", "astropy.visualization.mpl_normalize",
    "esm", "esm.utils", "esm.utils.structure", "esm.utils.structure.protein_chain",
    "biotite", "biotite.structure", "biotite.structure.io", "biotite.structure.io.pdb",
    "biotite.sequence", "biotite.structure.info", "biotite.application",
    "biotite.application.dssp",
    "imageio", "mpl_scatter_density", "requests", "pytorch_lightning", "Bio", "Bio.PDB",
]


def _install_stubs():
    for name in _STUBS:
        try:
            if name in sys.modules:
                continue
            __import__(name)
        except Exception:
            m = _StubModule(name)
            m.__path__ = []
            m.__spec__ = importlib.machinery.ModuleSpec(name, None, is_package=True)
            sys.modules[name] = m
            if "." in name:
                parent, child = name.rsplit(".", 1)
                if parent in sys.modules:
                    setattr(sys.modules[parent], child, m)


def load_reference():
    """Return the reference's `foldingdiff.bpe` module, patched as described above."""
    if not os.path.isdir(REFERENCE_ROOT):
        raise RuntimeError(f"reference not present at {REFERENCE_ROOT} (only in the build container)")
    # make sure OUR drop-in mirror (pt-bpe_b200/foldingdiff) is not shadowing the reference
    for k in [k for k in sys.modules if k == "foldingdiff" or k.startswith("foldingdiff.")]:
        mod = sys.modules[k]
        f = getattr(mod, "__file__", "") or ""
        if not f.startswith(REFERENCE_ROOT):
            del sys.modules[k]
    _install_stubs()
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)
    else:
        sys.path.remove(REFERENCE_ROOT)
        sys.path.insert(0, REFERENCE_ROOT)
    import warnings
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        import foldingdiff.bpe as ref_bpe
    assert ref_bpe.__file__.startswith(REFERENCE_ROOT), ref_bpe.__file__
    T = ref_bpe.Tokenizer

    # -- (3) pandas>=3 CoW: df[col].iloc[i] = v is a no-op; same index arithmetic, one-step write
    def _set_bond_length(self, idx, value):      # tokenizer.py:139-146
        if idx == 0:
            self._init_n_ca = value
        elif idx == 1:
            self._init_ca_c = value
        else:
            df = self._angles_and_dists
            df.iloc[(idx - 2) // 3, df.columns.get_loc(T.BOND_TYPES[idx % 3])] = value

    def _set_bond_angle(self, idx, value):       # tokenizer.py:155-159
        if idx == 0:
            self._init_bond_angle = value
        else:
            df = self._angles_and_dists
            df.iloc[(idx - 1) // 3, df.columns.get_loc(T.BOND_ANGLES[idx % 3])] = value

    def _set_dihedral_angle(self, idx, value):   # tokenizer.py:165-167
        df = self._angles_and_dists
        df.iloc[(idx + 1) // 3, df.columns.get_loc(T.DIHEDRAL_ANGLES[idx % 3])] = value

    T._set_bond_length = _set_bond_length
    T._set_bond_angle = _set_bond_angle
    T._set_dihedral_angle = _set_dihedral_angle

    # -- (4) visualisation -> no-ops
    noop = lambda self, *a, **k: None
    ref_bpe.BPE.visualize = noop
    T.visualize = lambda self, *a, **k: (None, None, None)
    T.visualize_bonds = noop
    ref_bpe.BinaryTreeBuilder.visualize = noop
    ref_bpe.BPE.plot_times = noop
    ref_bpe.BPE.plot_iou = noop

    # -- (5) esm.ProteinChain is absent: BPE.tokenize only uses it for reported rmsd/lddt metrics
    #        (bpe.py:1120-1139), which are not part of parity
    class _Metric(float):
        def mean(self):
            return 0.0

    class _Chain:
        @classmethod
        def from_pdb(cls, *a, **k):
            return cls()

        @classmethod
        def from_backbone_atom_coordinates(cls, *a, **k):
            return cls()

        def rmsd(self, *a, **k):
            return 0.0

        def lddt_ca(self, *a, **k):
            return _Metric(0.0)

    ref_bpe.ProteinChain = _Chain
    return ref_bpe
