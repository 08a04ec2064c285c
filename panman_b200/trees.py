"""Host-side tree builders for the synthetic workloads of BASELINE.json: flat trees in the reference's
numbering (pre-order creation order of createTreeFromNewickString, /root/reference/src/panman.cpp:300-440)."""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np


@dataclass
class FlatTree:
    parent: np.ndarray     # [V] int32, -1 for the root (node 0)
    child_off: np.ndarray  # [V+1] int32
    child_idx: np.ndarray  # [V-1] int32, children in Newick order
    leaf_row: np.ndarray   # [V] int32: MSA row of a leaf, -1 internal, -2 leaf absent from the MSA

    @property
    def n_nodes(self) -> int:
        return int(self.parent.shape[0])

    @property
    def n_leaves(self) -> int:
        return int((self.leaf_row != -1).sum())


def _preorder_binary(left: np.ndarray, right: np.ndarray, root: int, n_leaves: int) -> FlatTree:
    """left/right: children of join node k (ids >= n_leaves are join nodes, < n_leaves are leaves = MSA rows)."""
    V = 2 * n_leaves - 1
    parent = np.empty(V, np.int32)
    leaf_row = np.empty(V, np.int32)
    new_id = np.empty(V, np.int64)
    stack = [(root, -1)]
    nxt = 0
    L, R = left.tolist(), right.tolist()
    par_l, row_l = [0] * V, [0] * V
    while stack:
        v, p = stack.pop()
        i = nxt
        nxt += 1
        par_l[i] = p
        if v >= n_leaves:
            row_l[i] = -1
            k = v - n_leaves
            stack.append((R[k], i))
            stack.append((L[k], i))
        else:
            row_l[i] = v
    parent[:] = par_l
    leaf_row[:] = row_l
    internal = leaf_row == -1
    child_off = np.zeros(V + 1, np.int32)
    child_off[1:] = np.cumsum(np.where(internal, 2, 0))
    # children of i in pre-order: first child is i+1; second is the other node whose parent is i
    child_idx = np.empty(V - 1, np.int32)
    order = np.argsort(parent[1:], kind="stable") + 1  # nodes grouped by parent, ascending id inside a group
    child_idx[:] = order
    return FlatTree(parent, child_off, child_idx, leaf_row)


def random_binary_tree(n_leaves: int, seed: int = 20251106) -> FlatTree:
    """Uniform random joining of current roots; leaves are MSA rows 0..n_leaves-1."""
    assert n_leaves >= 2
    rng = np.random.default_rng(seed)
    u = rng.random((n_leaves - 1, 2))
    roots = list(range(n_leaves))
    left = np.empty(n_leaves - 1, np.int64)
    right = np.empty(n_leaves - 1, np.int64)
    for k in range(n_leaves - 1):
        m = len(roots)
        i = int(u[k, 0] * m)
        j = int(u[k, 1] * (m - 1))
        if j >= i:
            j += 1
        left[k], right[k] = roots[i], roots[j]
        roots[i] = n_leaves + k
        roots[j] = roots[-1]
        roots.pop()
    return _preorder_binary(left, right, roots[0], n_leaves)


def caterpillar_tree(n_leaves: int) -> FlatTree:
    """(t0,(t1,(t2,...(t{n-2},t{n-1})...))) -- BASELINE config 5."""
    V = 2 * n_leaves - 1
    ids = np.arange(V)
    parent = np.where(ids % 2 == 1, ids - 1, ids - 2).astype(np.int32)
    parent[0] = -1
    parent[V - 1] = V - 3 if n_leaves > 1 else -1
    leaf_row = np.where(ids % 2 == 1, ids // 2, -1).astype(np.int32)
    leaf_row[V - 1] = n_leaves - 1
    internal = leaf_row == -1
    child_off = np.zeros(V + 1, np.int32)
    child_off[1:] = np.cumsum(np.where(internal, 2, 0))
    child_idx = (np.arange(V - 1) + 1).astype(np.int32)
    return FlatTree(parent, child_off, child_idx, leaf_row)


def to_newick(tree: FlatTree, leaf_names=None) -> str:
    """Newick text (no internal labels, no branch lengths -- the shape of test/sars_20.nwk) that parses back to `tree`."""
    out = []
    stack = [(0, 0)]
    off, idx, row = tree.child_off.tolist(), tree.child_idx.tolist(), tree.leaf_row.tolist()
    while stack:
        v, k = stack.pop()
        b, e = off[v], off[v + 1]
        if b == e:
            out.append(leaf_names[v] if leaf_names is not None else (f"t{row[v]}" if row[v] >= 0 else f"absent{v}"))
            continue
        if k == 0:
            out.append("(")
        elif k < e - b:
            out.append(",")
        if k < e - b:
            stack.append((v, k + 1))
            stack.append((idx[b + k], 0))
        else:
            out.append(")")
    return "".join(out) + ";"


def collapse_edges(tree: FlatTree, frac: float, seed: int = 1) -> FlatTree:
    """Polytomies out of a binary tree: a fraction `frac` of the internal non-root nodes is contracted into its parent
    (its children become children of the nearest surviving ancestor, in place).  Mutation-annotated trees of outbreak
    data look like this: most branches have no mutation and are collapsed, some nodes have hundreds of children."""
    rng = np.random.default_rng(seed)
    V = tree.n_nodes
    internal = tree.leaf_row == -1
    drop = internal & (rng.random(V) < frac)
    drop[0] = False
    keep = ~drop
    new_id = np.cumsum(keep) - 1  # pre-order is preserved by contraction
    anc = np.arange(V)            # nearest surviving ancestor-or-self, resolved top-down (parents have smaller ids)
    par = tree.parent
    for v in range(1, V):
        if drop[v]:
            anc[v] = anc[par[v]]
    kept = np.nonzero(keep)[0]
    parent = np.full(kept.shape[0], -1, np.int32)
    parent[1:] = new_id[anc[par[kept[1:]]]]
    leaf_row = tree.leaf_row[kept].astype(np.int32)
    n = kept.shape[0]
    child_off = np.zeros(n + 1, np.int32)
    np.add.at(child_off, parent[1:] + 1, 1)
    child_off = np.cumsum(child_off).astype(np.int32)
    child_idx = (np.argsort(parent[1:], kind="stable") + 1).astype(np.int32)  # pre-order = Newick order inside a parent
    return FlatTree(parent, child_off, child_idx, leaf_row)
