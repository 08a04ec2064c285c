"""Small synthetic trees and alignments for the parity tests (numpy; CPU only).

Trees are produced directly in the reference's numbering (pre-order creation order of createTreeFromNewickString,
/root/reference/src/panman.cpp:300-440) together with the Newick text that yields them.
"""
from __future__ import annotations

import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from oracle.oracle_api import FlatTree  # noqa: E402


def _flatten(children_of_root, n_leaves_hint=None):
    """nested lists (leaf = int row id, or None for a leaf absent from the MSA) -> FlatTree in pre-order + Newick."""
    parent, leaf_row, kids = [], [], []
    newick_parts = []

    # iterative pre-order to survive deep ladders
    root = children_of_root
    stack = [(root, -1, 0)]  # (obj, parent id, state)
    out_stack = []
    # First pass: assign ids in pre-order
    ids = {}
    order = []
    work = [(root, -1)]
    while work:
        obj, par = work.pop()
        nid = len(parent)
        parent.append(par)
        kids.append([])
        if par >= 0:
            kids[par].append(nid)
        if isinstance(obj, list):
            leaf_row.append(-1)
            for ch in reversed(obj):
                work.append((ch, nid))
        else:
            leaf_row.append(-2 if obj is None else int(obj))
    V = len(parent)
    child_off = np.zeros(V + 1, dtype=np.int32)
    for i in range(V):
        child_off[i + 1] = child_off[i] + len(kids[i])
    child_idx = np.array([c for k in kids for c in k], dtype=np.int32)
    return FlatTree(np.array(parent, dtype=np.int32), child_off, child_idx, np.array(leaf_row, dtype=np.int32))


def leaf_name(tree: FlatTree, node: int) -> str:
    r = int(tree.leaf_row[node])
    return f"t{r}" if r >= 0 else f"absent{node}"


def to_newick(tree: FlatTree, names=None, branch_lengths: bool = False) -> str:
    """Newick text whose parse by the reference reproduces `tree` (internal nodes unlabeled)."""
    V = tree.n_nodes
    out = []
    # iterative: emit '(' on entry, children separated by ',', ')' on exit
    stack = [(0, 0)]
    while stack:
        v, k = stack.pop()
        b, e = int(tree.child_off[v]), int(tree.child_off[v + 1])
        if b == e:
            out.append(names[v] if names is not None else leaf_name(tree, v))
            if branch_lengths:
                out.append(":0.5")
            continue
        if k == 0:
            out.append("(")
        elif k < e - b:
            out.append(",")
        if k < e - b:
            stack.append((v, k + 1))
            stack.append((int(tree.child_idx[b + k]), 0))
        else:
            out.append(")")
            if branch_lengths and v != 0:
                out.append(":1.5")
    return "".join(out) + ";"


def random_tree(n_leaves: int, rng: np.random.Generator, max_arity: int = 2, absent_frac: float = 0.0) -> FlatTree:
    """Random tree by joining uniformly chosen current roots, `arity` at a time (2..max_arity)."""
    items = [(None if rng.random() < absent_frac else i) for i in range(n_leaves)]
    if all(x is None for x in items):
        items[0] = 0
    # re-number present rows densely
    nxt = 0
    for i, x in enumerate(items):
        if x is not None:
            items[i] = nxt
            nxt += 1
    roots = list(items)
    while len(roots) > 1:
        k = int(rng.integers(2, max_arity + 1))
        k = min(k, len(roots))
        picks = sorted(rng.choice(len(roots), size=k, replace=False).tolist(), reverse=True)
        group = [roots.pop(p) for p in picks]
        roots.append(group)
    root = roots[0]
    if not isinstance(root, list):
        root = [root]
    return _flatten(root)


def caterpillar(n_leaves: int) -> FlatTree:
    """(t0,(t1,(t2,...(t{n-2},t{n-1})...))) built without recursion."""
    V = 2 * n_leaves - 1
    parent = np.empty(V, dtype=np.int32)
    leaf_row = np.empty(V, dtype=np.int32)
    child_off = np.zeros(V + 1, dtype=np.int32)
    child_idx = []
    # pre-order: node 0 internal, 1 = t0, 2 internal, 3 = t1, ... last two are leaves
    nid = 0
    par = -1
    counts = np.zeros(V, dtype=np.int32)
    for i in range(n_leaves - 1):
        internal = nid
        parent[internal] = par
        leaf_row[internal] = -1
        leaf = nid + 1
        parent[leaf] = internal
        leaf_row[leaf] = i
        counts[internal] = 2
        nid += 2
        par = internal
    parent[nid] = par
    leaf_row[nid] = n_leaves - 1
    child_off[1:] = np.cumsum(counts)
    child_idx = np.zeros(int(child_off[-1]), dtype=np.int32)
    for i in range(n_leaves - 1):
        internal = 2 * i
        child_idx[child_off[internal]] = internal + 1
        child_idx[child_off[internal] + 1] = internal + 2
    return FlatTree(parent, child_off, child_idx, leaf_row)


def n_rows(tree: FlatTree) -> int:
    r = tree.leaf_row[tree.leaf_row >= 0]
    return int(r.max()) + 1 if r.size else 0


def simulate_msa(tree: FlatTree, n_cols: int, rng: np.random.Generator, p_sub=0.02, p_del=0.01, p_ins=0.05,
                 p_n=0.02, p_amb=0.01, root_gap_frac=0.05) -> np.ndarray:
    """Top-down simulated evolution on codes; returns [n_rows, n_cols] uint8 codes with >=1 non-gap per column."""
    V = tree.n_nodes
    bases = np.array([1, 2, 4, 8], dtype=np.uint8)
    state = np.zeros((V, n_cols), dtype=np.uint8)
    state[0] = rng.choice(bases, size=n_cols, p=[0.30, 0.18, 0.20, 0.32])
    state[0][rng.random(n_cols) < root_gap_frac] = 0
    for v in range(1, V):
        s = state[tree.parent[v]].copy()
        sub = rng.random(n_cols) < p_sub
        s[sub & (s != 0)] = rng.choice(bases, size=int((sub & (s != 0)).sum()))
        dele = rng.random(n_cols) < p_del
        ins = rng.random(n_cols) < p_ins
        was_gap = s == 0
        s[dele & ~was_gap] = 0
        s[ins & was_gap] = rng.choice(bases, size=int((ins & was_gap).sum()))
        state[v] = s
    R = n_rows(tree)
    msa = np.zeros((R, n_cols), dtype=np.uint8)
    for v in range(V):
        r = tree.leaf_row[v]
        if r >= 0:
            s = state[v].copy()
            s[rng.random(n_cols) < p_n] = 15
            amb = rng.random(n_cols) < p_amb
            s[amb] = rng.choice(np.array([3, 5, 6, 7, 9, 10, 11, 12, 13, 14], dtype=np.uint8), size=int(amb.sum()))
            msa[r] = s
    if R:
        allgap = (msa != 0).sum(axis=0) == 0
        fix = np.nonzero(allgap)[0]
        msa[fix % R, fix] = rng.choice(bases, size=fix.size)
    return msa


def consensus_first_nongap(msa_codes: np.ndarray, order=None) -> np.ndarray:
    """First non-gap code per column scanning rows in `order` (the reference scans std::map, i.e. id-sorted, order)."""
    R, L = msa_codes.shape
    rows = np.arange(R) if order is None else np.asarray(order)
    m = msa_codes[rows]
    nz = m != 0
    first = nz.argmax(axis=0)
    out = m[first, np.arange(L)]
    out[~nz.any(axis=0)] = 0
    return out.astype(np.uint8)
