"""Semantic round trip, independent of any oracle: replay the emitted mutations down the tree the way the reference
rebuilds sequences from a PanMAT (src/fasta.cpp:545-657 -- start from the block consensus, at every node apply its packed
NucMut records in order: NS / NI write nibble j of `nucs` at nucPosition + j, ND writes the gap; undo on the way back up)
and every leaf must come out as its alignment row."""
import numpy as np


def replay_leaves(tree, consensus_codes, nuc_off, nuc):
    """-> {leaf node id: codes [L]} for every leaf, by applying the per-node packed NucMut lists along the root path."""
    V = tree.n_nodes
    state = np.asarray(consensus_codes, dtype=np.uint8).copy()
    out = {}
    undo = [None] * V
    # iterative DFS with explicit undo: (node, phase)
    stack = [(0, 0)]
    while stack:
        v, phase = stack.pop()
        if phase == 1:
            pos, old = undo[v]
            state[pos] = old
            undo[v] = None
            continue
        recs = nuc[nuc_off[v]: nuc_off[v + 1]]
        pos_l, val_l = [], []
        for r in recs:
            info = int(r["mutInfo"])
            typ, ln = info & 0xF, info >> 4
            assert 1 <= ln <= 6 and typ in (0, 1, 2) and int(r["nucGapPosition"]) == -1 and int(r["primaryBlockId"]) == 0
            for j in range(ln):
                pos_l.append(int(r["nucPosition"]) + j)
                val_l.append(0 if typ == 1 else (int(r["nucs"]) >> (4 * (5 - j))) & 0xF)
        pos = np.array(pos_l, np.int64)
        old = state[pos].copy() if pos.size else np.zeros(0, np.uint8)
        if pos.size:
            assert np.unique(pos).size == pos.size  # a node never mutates a position twice
            state[pos] = np.array(val_l, np.uint8)
        undo[v] = (pos, old)
        stack.append((v, 1))
        b, e = int(tree.child_off[v]), int(tree.child_off[v + 1])
        if b == e:
            out[v] = state.copy()
        for k in range(e - 1, b - 1, -1):
            stack.append((int(tree.child_idx[k]), 0))
    return out


def assert_replay_reproduces_msa(tree, msa_codes, consensus_codes, nuc_off, nuc):
    leaves = replay_leaves(tree, consensus_codes, nuc_off, nuc)
    for v, codes in leaves.items():
        r = int(tree.leaf_row[v])
        if r >= 0:
            bad = np.nonzero(codes != msa_codes[r])[0]
            assert bad.size == 0, (v, r, bad[:8], codes[bad[:8]], msa_codes[r][bad[:8]])
