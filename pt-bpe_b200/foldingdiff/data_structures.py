from geobpe_b200.data_structures import ThresholdDict, TokenHierarchy, BinaryTreeBuilder, Node, _rebuild_hierarchy  # noqa: F401
