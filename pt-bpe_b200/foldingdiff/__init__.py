"""Drop-in namespace: the reference's import paths for the GeoBPE hot path, backed by geobpe_b200 (CUDA)."""
