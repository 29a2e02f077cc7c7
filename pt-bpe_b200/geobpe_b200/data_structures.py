"""Containers that are part of the reference's pickle layout (foldingdiff/data_structures.py:4-88,200-296):
same class names, attribute names and behaviour; re-implemented, plotting left out."""
import bisect


class Node:
    def __init__(self, value, left=None, right=None):
        self.value, self.left, self.right = value, left, right

    def __repr__(self):
        lv = self.left.value if self.left else None
        rv = self.right.value if self.right else None
        return f"Node({self.value}, left={lv}, right={rv})"


class BinaryTreeBuilder:
    """Merge tree of one backbone: `nodes` maps start bond -> current root covering it, `leaves` keeps the
    residue-level tokens (data_structures.py:16-60)."""

    def __init__(self):
        self.nodes = {}
        self.leaves = {}

    def add_leaf(self, key):
        if key in self.nodes:
            raise ValueError(f"Key '{key}' already exists.")
        self.nodes[key[0]] = Node(key)
        self.leaves[key[0]] = Node(key)

    def combine(self, left_key, right_key, parent_key):
        if left_key[0] not in self.nodes or right_key[0] not in self.nodes:
            raise ValueError("Both keys must exist in the nodes dictionary.")
        left, right = self.nodes.pop(left_key[0]), self.nodes.pop(right_key[0])
        assert left.value == left_key and right.value == right_key
        self.nodes[parent_key[0]] = Node(parent_key, left, right)

    def visualize(self, *a, **k):      # matplotlib output: unavailable on this box
        return None


class TokenHierarchy(dict):
    """dict start bond -> (start, id, length) that mirrors every merge into `tree` (data_structures.py:200-262)"""

    def __init__(self, parent, *args, **kwargs):
        self.parent = parent
        super().__init__(*args, **kwargs)
        self.init()

    def init(self):
        prev = None
        for n in self:
            if prev:
                assert prev[0] + prev[-1] == self[n][0]
            prev = self[n]
        self.tree = BinaryTreeBuilder()
        for n in self:
            self.tree.add_leaf(self[n])

    def __setitem__(self, k, v):
        super().__setitem__(k, v)
        node = self.tree.nodes.get(k)
        if node is None:
            return
        left_key = node.value
        right = self.tree.nodes.get(left_key[0] + left_key[2])
        if right is None:
            return
        right_key = right.value
        assert right_key[0] + right_key[2] == v[0] + v[2]
        self.tree.combine(left_key, right_key, v)

    def __reduce__(self):              # dict subclasses with a required ctor argument need help to unpickle
        return (_rebuild_hierarchy, (self.parent, dict(self), self.tree))

    def clear(self):
        raise NotImplementedError

    def setdefault(self, key, default=None):
        raise NotImplementedError


def _rebuild_hierarchy(parent, items, tree):
    th = TokenHierarchy.__new__(TokenHierarchy)
    th.parent = parent
    dict.update(th, items)
    th.tree = tree
    return th


class ThresholdDict(dict):
    """d[k] = value at the largest int key <= k (data_structures.py:264-296)"""

    def __new__(cls, *args, **kwargs):
        inst = super().__new__(cls)
        inst._int_keys = []
        return inst

    def __init__(self, *args, **kwargs):
        super().__init__(*args, **kwargs)
        self._int_keys = sorted(k for k in self if isinstance(k, int))

    def __setitem__(self, key, val):
        if isinstance(key, int) and key not in self:
            bisect.insort(self._int_keys, key)
        super().__setitem__(key, val)

    def __getitem__(self, key):
        if key in self:
            return super().__getitem__(key)
        if isinstance(key, int):
            i = bisect.bisect_right(self._int_keys, key) - 1
            if i >= 0:
                return super().__getitem__(self._int_keys[i])
        raise KeyError(key)


for _c in (Node, BinaryTreeBuilder, TokenHierarchy, ThresholdDict):
    _c.__module__ = "foldingdiff.data_structures"
_rebuild_hierarchy.__module__ = "foldingdiff.data_structures"
