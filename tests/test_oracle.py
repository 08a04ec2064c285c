"""CPU: pins the plain-C oracle (oracle/pmfs_oracle.c) to the reference.

* against the committed golden fixtures, which were produced by the reference's own fitchSankoff.cpp (oracle/_ref);
* directly against oracle/_ref on fresh random inputs wherever that library exists (this container; the prebuilt file
  also travels to the GPU box).
"""
import numpy as np
import pytest

import golden_util as G
import synth
from oracle import oracle_api as O


def test_port_matches_golden_small(small_cases):
    z = small_cases
    for name in z["case_names"]:
        tree, msa, cons, forced = G.case_inputs(z, str(name))
        for algo in (0, 1):
            got = O.run_port(algo, tree, msa, cons, forced, threads=2)
            assert G.same_muts(got, z[f"{name}/muts{algo}"]), (name, algo)
            off, nuc = O.pack_port(got, tree.n_nodes)
            assert (off == z[f"{name}/nuc_off{algo}"]).all(), (name, algo)
            assert G.same_nuc(nuc, z[f"{name}/nuc{algo}"]), (name, algo)


@pytest.mark.parametrize("low,with_ref", [(0, 0), (0, 1), (1, 0), (1, 1)])
def test_port_matches_golden_c1(c1, low, with_ref):
    key = f"v{low}{with_ref}_"
    tree = G.c1_tree(c1)
    msa = G.unpack_nibbles(c1["msa_nibbles"], int(c1["n_cols"]))
    forced = None
    if with_ref:
        row = list(c1["row_ids"]).index(str(c1[key + "reference"]))
        # Fitch: refState = 1<<code(ref[i]) (panman.cpp:1409); Sankoff: defaultState = code(ref[i]) (:1573-1580)
        forced = msa[row]
    got = O.run_port(low, tree, msa, c1[key + "consensus"], forced, threads=4)
    assert got.shape[0] == int(c1[key + "n_muts"])
    assert (np.bincount(got["type"], minlength=3) == c1[key + "type_counts"]).all()
    assert G.same_muts(got[:512], c1[key + "muts_head"])
    assert G.digest_muts(got) == str(c1[key + "muts_sha256"])
    off, nuc = O.pack_port(got, tree.n_nodes)
    assert nuc.shape[0] == int(c1[key + "n_nuc"])
    assert G.digest_nuc(off, nuc) == str(c1[key + "nuc_sha256"])


@pytest.mark.skipif(not O.ref_available(), reason="oracle/_ref not built and /root/reference absent")
def test_port_matches_reference_random():
    rng = np.random.default_rng(7)
    for trial in range(30):
        arity = 2 if trial < 10 else int(rng.integers(2, 6))
        tree = synth.random_tree(int(rng.integers(2, 90)), rng, max_arity=arity, absent_frac=0.0 if trial % 3 else 0.1)
        L = int(rng.integers(1, 300))
        msa = synth.simulate_msa(tree, L, rng, p_sub=0.1, p_del=0.05, p_ins=0.2, p_n=0.05, p_amb=0.05)
        cons = synth.consensus_first_nongap(msa)
        forced = msa[int(rng.integers(0, msa.shape[0]))] if trial % 2 else None
        for algo in (0, 1):
            a = O.run_port(algo, tree, msa, cons, forced, threads=2)
            b = O.run_ref_columns(algo, tree, O.codes_to_chars(msa), O.codes_to_chars(cons),
                                  None if forced is None else O.codes_to_chars(forced), threads=2)
            assert G.same_muts(a, b), (trial, algo)


@pytest.mark.skipif(not O.ref_available(), reason="oracle/_ref not built and /root/reference absent")
def test_port_matches_reference_deep_ladder():
    # the reference recursion needs a big stack here; ref_driver runs it on 64 MB+ threads
    tree = synth.caterpillar(20000)
    rng = np.random.default_rng(3)
    msa = rng.choice(np.array([0, 1, 2, 4, 8, 15], np.uint8), size=(20000, 4), p=[0.05, 0.3, 0.2, 0.2, 0.2, 0.05])
    msa[0, (msa != 0).sum(0) == 0] = 1
    cons = synth.consensus_first_nongap(msa)
    for algo in (0, 1):
        a = O.run_port(algo, tree, msa, cons, None, threads=2)
        b = O.run_ref_columns(algo, tree, O.codes_to_chars(msa), O.codes_to_chars(cons), None, threads=2)
        assert G.same_muts(a, b)


def test_sankoff_unforced_invalid_root_is_an_error():
    tree = synth._flatten([None, None])
    with pytest.raises(RuntimeError):
        O.run_port(1, tree, np.zeros((0, 3), np.uint8), np.zeros(3, np.uint8), None)
    # Fitch: no records at all (every node has an empty set and is skipped)
    assert O.run_port(0, tree, np.zeros((0, 3), np.uint8), np.zeros(3, np.uint8), None).shape[0] == 0


def test_block_passes_are_the_nucleotide_passes_on_three_states():
    """the reference's block-level passes (blockFitch*New, blockSankoff*; 3 states) give exactly what its
    nucleotide passes -- hence the library -- give on the state codes 0/1/2 with an all-absent parent of the root, after
    relabelling the mutation types.  Checked against the reference's own block members."""
    rng = np.random.default_rng(31)
    for trial in range(12):
        arity = 2 if trial < 6 else int(rng.integers(2, 6))
        tree = synth.random_tree(int(rng.integers(2, 90)), rng, max_arity=arity, absent_frac=0.1 if trial % 3 == 0 else 0.0)
        rows, B = synth.n_rows(tree), int(rng.integers(1, 120))
        anc = rng.integers(0, 3, B)
        states = np.where(rng.random((rows, B)) < 0.7, anc[None, :], rng.integers(0, 3, (rows, B))).astype(np.uint8)
        forced = states[int(rng.integers(0, rows))] if trial % 2 else None
        cons = np.zeros(B, np.uint8)  # parentState of the root: block absent
        for algo in (0, 1):
            want = O.run_ref_blocks(algo, tree, states, forced)
            t2, st2 = tree, states
            if algo == 1 and (tree.leaf_row == -2).any():
                # the block Sankoff pass gives a leaf without a sequence the state "absent": an all-zero row
                lr = tree.leaf_row.copy()
                n_abs = int((lr == -2).sum())
                lr[lr == -2] = rows + np.arange(n_abs)
                t2 = type(tree)(tree.parent, tree.child_off, tree.child_idx, lr)
                st2 = np.vstack([states, np.zeros((n_abs, B), np.uint8)])
            got = O.blocks_from_nuc(O.run_port(algo, t2, st2, cons, forced, threads=1))
            assert G.same_muts(got, want), (trial, algo)


def test_replaying_the_mutations_reproduces_every_leaf_row(c1):
    """The emitted mutations are a lossless encoding of the alignment given the tree and the consensus: replaying them
    from the root (the rule the reference rebuilds sequences with, src/fasta.cpp:545-657) yields every leaf's row.  Checked
    here on the oracle's output (config 1 in all four CLI variants, random polytomy trees with absent leaves and forced
    roots) -- tests/test_gpu_parity.py runs the same check on the CUDA path's output."""
    import golden_util as G
    import replay
    import synth
    tree = G.c1_tree(c1)
    L = int(c1["n_cols"])
    msa = G.unpack_nibbles(c1["msa_nibbles"], L)
    for low in (0, 1):
        for with_ref in (0, 1):
            key = f"v{low}{with_ref}_"
            forced = msa[list(c1["row_ids"]).index(str(c1[key + "reference"]))] if with_ref else None
            muts = O.run_port(low, tree, msa, c1[key + "consensus"], forced, threads=4)
            off, nuc = O.pack_port(muts, tree.n_nodes)
            replay.assert_replay_reproduces_msa(tree, msa, c1[key + "consensus"], off, nuc)
    rng = np.random.default_rng(8)
    for trial in range(6):
        t = synth.random_tree(int(rng.integers(3, 120)), rng, max_arity=5, absent_frac=0.1 if trial % 2 else 0.0)
        m = synth.simulate_msa(t, 300, rng, p_sub=0.08, p_del=0.05, p_ins=0.2, p_n=0.05, p_amb=0.05)
        cons = synth.consensus_first_nongap(m)
        forced = m[0] if trial % 3 == 0 else None
        for algo in (0, 1):
            muts = O.run_port(algo, t, m, cons, forced, threads=2)
            off, nuc = O.pack_port(muts, t.n_nodes)
            replay.assert_replay_reproduces_msa(t, m, cons, off, nuc)
