from geobpe_b200.bpe import Tokenizer  # noqa: F401
from geobpe_b200.engine import BOND_TYPES, BOND_ANGLES, DIHEDRAL_ANGLES  # noqa: F401
