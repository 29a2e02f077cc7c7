from geobpe_b200.data_structures import ThresholdDict, TokenHierarchy, BinaryTreeBuilder, Node  # noqa: F401
