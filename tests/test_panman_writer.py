"""CPU: the .panman writer (panman_b200/csrc/host/panman_writer.cpp) against an independently written reader
(tests/panman_reader.py) on BASELINE config 1: all four CLI variants, records from the oracle, the reference's own
packing.  What comes back out of the file must be the golden tree (Newick text re-parsed by the reference's parser
rules), the golden per-node NucMut lists and the Block consensus words, with the flags the reference's reader expects."""
import os
import subprocess

import numpy as np
import pytest

import golden_util as G
from oracle import oracle_api as O
from panman_b200 import host
from panman_reader import nuc_csr, read_panman


@pytest.mark.parametrize("low,with_ref", [(0, 0), (0, 1), (1, 0), (1, 1)])
def test_written_panman_decodes_to_the_golden_tree_and_mutations(c1, tmp_path, low, with_ref):
    key = f"v{low}{with_ref}_"
    tree = G.c1_tree(c1)
    L = int(c1["n_cols"])
    msa = G.unpack_nibbles(c1["msa_nibbles"], L)
    forced = msa[list(c1["row_ids"]).index(str(c1[key + "reference"]))] if with_ref else None
    muts = O.run_port(low, tree, msa, c1[key + "consensus"], forced, threads=4)
    assert G.digest_muts(muts) == str(c1[key + "muts_sha256"])
    off, nuc = O.pack_ref(muts, tree.n_nodes)
    assert G.digest_nuc(off, nuc) == str(c1[key + "nuc_sha256"])
    # Block(0, consensusSeq) words (src/panman.cpp:236-247): 8 nibbles per word, first character in the top nibble
    cons = c1[key + "consensus"].astype(np.uint32)
    pad = np.zeros((-L) % 8, np.uint32)
    words = np.concatenate([cons, pad]).reshape(-1, 8)
    block_words = np.zeros(words.shape[0], np.uint32)
    for k in range(8):
        block_words |= words[:, k] << np.uint32(4 * (7 - k))
    names = [str(x) for x in c1["names"]]
    for compress in (True, False):
        path = str(tmp_path / f"c1_{key}{int(compress)}.panman")
        host.write_panman(tree.parent, names, off, nuc, block_words, path, compress=compress)
        if compress:  # a real .xz container
            assert subprocess.run(["xz", "-t", path]).returncode == 0
        d = read_panman(path, compressed=compress)
        assert d["n_trees"] == 1 and d["n_complex"] == 0 and d["n_gaps"] == 0 and d["blockGaps_null"]
        assert d["n_circular"] == d["n_rotation"] == d["n_inverted"] == 0
        assert len(d["nodes"]) == tree.n_nodes + 1 and d["nodes"][-1]["mutations"] == []
        got_off, got_nuc = nuc_csr(d)
        assert (got_off == off).all() and G.same_nuc(got_nuc, nuc)
        # block-level flags: the root carries the block insertion; nodes with only nucleotide mutations carry the marker
        root = d["nodes"][0]["mutations"][0]
        assert (root["blockMutExist"], root["blockMutInfo"], root["blockInversion"], root["blockId"], root["blockGapExist"]) == (1, 1, 0, 0, 0)
        for v in range(1, tree.n_nodes):
            ms = d["nodes"][v]["mutations"]
            assert len(ms) == (1 if off[v + 1] > off[v] else 0)
            if ms:
                assert (ms[0]["blockMutExist"], ms[0]["blockMutInfo"], ms[0]["blockInversion"], ms[0]["blockId"]) == (0, 1, 1, 0)
        cm = d["consensusSeqMap"]
        assert len(cm) == 1 and cm[0]["blockId"].tolist() == [0] and cm[0]["blockGapExist"] == [0]
        assert (cm[0]["consensusSeq"] == block_words).all()
        # the Newick text parses back (by the library's mirror of createTreeFromNewickString) to the same tree
        pnames, parent, nch = host.parse_newick(d["newick"])
        assert pnames == names and (parent == tree.parent).all()
        assert d["newick"].count(":0.000000") == tree.n_nodes  # std::to_string(branchLength) on every node, the root included


def test_newick_text_with_branch_lengths_and_chunks(tmp_path):
    """Branch lengths are written with std::to_string (6 decimals); a Newick text over 1 MiB is split into chunks
    (NEWICK_CHUNK_SIZE, src/panman.cpp:44) and must concatenate back."""
    n_leaves = 30000
    V = 2 * n_leaves - 1
    parent = np.full(V, -1, np.int32)
    names = []
    k = 0
    for v in range(V):  # caterpillar in pre-order: internal, leaf, internal, leaf, ...
        internal = v % 2 == 0 and v < V - 1
        if v > 0:
            parent[v] = v - 1 if v % 2 == 1 else v - 2
        if v == V - 1:
            parent[v] = v - 2 if V > 1 else -1
            internal = False
        if internal:
            k += 1
            names.append(f"node_{k}")
        else:
            names.append(f"leaf_with_a_long_name_{v}")
    bl = (np.arange(V) % 7 * 0.25).astype(np.float32)
    path = str(tmp_path / "ladder.panman")
    host.write_panman(parent, names, np.zeros(V + 1, np.int64), np.zeros(0, host.NUCMUT_DTYPE), np.zeros(0, np.uint32), path,
                      compress=False, branch_length=bl)
    d = read_panman(path, compressed=False)
    assert len(d["newick"]) > (1 << 20)
    pnames, pparent, _ = host.parse_newick(d["newick"])
    assert pnames == names and (pparent == parent).all()
    assert ":1.250000" in d["newick"] and d["newick"].endswith(";")
