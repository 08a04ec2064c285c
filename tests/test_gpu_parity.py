"""GPU (B200): the CUDA path, called through the C ABI (include/pmfs.h), against the oracle and the golden fixtures.
Bit-exact: (node, pos, type, code) records and the packed NucMut arrays must be identical."""
import numpy as np
import pytest

import golden_util as G
import synth
from oracle import oracle_api as O
from panman_b200 import api, trees

pytestmark = pytest.mark.gpu


def run_gpu(tree, algo, msa_codes, cons, forced=None, chunk_nodes=0, tile_cols=0):
    with api.Context(tree) as ctx:
        ctx.set_tuning(chunk_nodes, tile_cols)
        nib = api.pack_nibbles(msa_codes) if msa_codes.shape[0] else np.zeros((0, 1), np.uint8)
        return ctx.run(algo, nib, int(cons.shape[0]), cons, forced)


def test_device_side_packing(small_cases, c1):
    """pmfs_pack_last: greedy <=6-run packing on the device == the reference rule (golden NucMut arrays)."""
    z = small_cases
    for name in z["case_names"]:
        tree, msa, cons, forced = G.case_inputs(z, str(name))
        for algo in (0, 1):
            with api.Context(tree) as ctx:
                ctx.run(algo, api.pack_nibbles(msa), int(cons.shape[0]), cons, forced)
                off, nuc = ctx.pack_last()
            assert (off == z[f"{name}/nuc_off{algo}"]).all(), (name, algo)
            assert G.same_nuc(nuc, z[f"{name}/nuc{algo}"]), (name, algo)
    key = "v00_"
    tree = G.c1_tree(c1)
    msa = G.unpack_nibbles(c1["msa_nibbles"], int(c1["n_cols"]))
    with api.Context(tree) as ctx:
        ctx.run(api.FITCH, api.pack_nibbles(msa), int(c1["n_cols"]), c1[key + "consensus"], None)
        off, nuc = ctx.pack_last()
    assert nuc.shape[0] == int(c1[key + "n_nuc"])
    assert G.digest_nuc(off, nuc) == str(c1[key + "nuc_sha256"])
    # no records at all: every list is empty
    tree = synth._flatten([None, None])
    with api.Context(tree) as ctx:
        ctx.run(api.FITCH, np.zeros((0, 3), np.uint8), 5, np.zeros(5, np.uint8))
        off, nuc = ctx.pack_last()
    assert nuc.shape[0] == 0 and (off == 0).all()


def test_golden_small_cases(small_cases):
    z = small_cases
    for name in z["case_names"]:
        tree, msa, cons, forced = G.case_inputs(z, str(name))
        for algo in (0, 1):
            for chunk in (0, 4):
                got = run_gpu(tree, algo, msa, cons, forced, chunk_nodes=chunk)
                assert G.same_muts(got, z[f"{name}/muts{algo}"]), (name, algo, chunk)
            off, nuc = O.pack_port(got, tree.n_nodes)
            assert (off == z[f"{name}/nuc_off{algo}"]).all() and G.same_nuc(nuc, z[f"{name}/nuc{algo}"])


@pytest.mark.parametrize("low,with_ref", [(0, 0), (0, 1), (1, 0), (1, 1)])
def test_golden_c1_sars20(c1, low, with_ref):
    """BASELINE config 1 (test/sars_20.fa padded + test/sars_20.nwk), all four CLI variants."""
    key = f"v{low}{with_ref}_"
    tree = G.c1_tree(c1)
    msa = G.unpack_nibbles(c1["msa_nibbles"], int(c1["n_cols"]))
    forced = msa[list(c1["row_ids"]).index(str(c1[key + "reference"]))] if with_ref else None
    got = run_gpu(tree, low, msa, c1[key + "consensus"], forced)
    assert got.shape[0] == int(c1[key + "n_muts"])
    assert (np.bincount(got["type"], minlength=3) == c1[key + "type_counts"]).all()
    assert G.same_muts(got[:512], c1[key + "muts_head"])
    assert G.digest_muts(got) == str(c1[key + "muts_sha256"])
    off, nuc = O.pack_port(got, tree.n_nodes)
    assert G.digest_nuc(off, nuc) == str(c1[key + "nuc_sha256"])


def test_random_trees_vs_oracle():
    rng = np.random.default_rng(101)
    for trial in range(24):
        arity = 2 if trial < 10 else int(rng.integers(2, 7))
        tree = synth.random_tree(int(rng.integers(2, 400)), rng, max_arity=arity, absent_frac=0.0 if trial % 3 else 0.08)
        L = int(rng.integers(1, 700))
        msa = synth.simulate_msa(tree, L, rng, p_sub=0.08, p_del=0.04, p_ins=0.2, p_n=0.04, p_amb=0.04)
        cons = synth.consensus_first_nongap(msa)
        forced = msa[int(rng.integers(0, msa.shape[0]))] if trial % 2 else None
        chunk = int(rng.choice([0, 2, 16, 64]))
        tile = int(rng.choice([0, 32, 4096]))
        for algo in (0, 1):
            want = O.run_port(algo, tree, msa, cons, forced, threads=4)
            got = run_gpu(tree, algo, msa, cons, forced, chunk_nodes=chunk, tile_cols=tile)
            assert G.same_muts(got, want), (trial, algo, chunk, tile)


def test_medium_binary_tree_many_columns():
    """2 000 taxa x 20 000 columns: several CTAs per chunk, ragged last group, both algorithms, tiled and untiled."""
    rng = np.random.default_rng(9)
    tree = trees.random_binary_tree(2000, seed=5)
    L = 20011
    msa = synth.simulate_msa(tree, L, rng, p_sub=0.002, p_del=0.001, p_ins=0.01, p_n=0.01, p_amb=0.001)
    cons = synth.consensus_first_nongap(msa)
    for algo in (0, 1):
        want = O.run_port(algo, tree, msa, cons, None, threads=8)
        got = run_gpu(tree, algo, msa, cons, None)
        assert G.same_muts(got, want), algo
        got2 = run_gpu(tree, algo, msa, cons, msa[3], chunk_nodes=50, tile_cols=8192)
        want2 = O.run_port(algo, tree, msa, cons, msa[3], threads=8)
        assert G.same_muts(got2, want2), algo


def test_deep_ladder():
    """Config-5 shape at test size: a 30 000-deep ladder must neither overflow a stack nor lose bit-exactness."""
    rng = np.random.default_rng(4)
    tree = trees.caterpillar_tree(30000)
    L = 96
    msa = rng.choice(np.array([0, 1, 2, 4, 8, 15], np.uint8), size=(30000, L), p=[0.03, 0.3, 0.2, 0.2, 0.22, 0.05])
    msa[0, (msa != 0).sum(0) == 0] = 1
    cons = synth.consensus_first_nongap(msa)
    for algo in (0, 1):
        want = O.run_port(algo, tree, msa, cons, None, threads=8)
        got = run_gpu(tree, algo, msa, cons, None)
        assert G.same_muts(got, want), algo


def test_error_paths():
    tree = synth._flatten([None, None])
    cons = np.zeros(5, np.uint8)
    with api.Context(tree) as ctx:
        assert ctx.run(api.FITCH, np.zeros((0, 3), np.uint8), 5, cons).shape[0] == 0
        with pytest.raises(api.PmfsError) as e:
            ctx.run(api.SANKOFF, np.zeros((0, 3), np.uint8), 5, cons)
        assert e.value.status == -4
        got = ctx.run(api.SANKOFF, np.zeros((0, 3), np.uint8), 5, cons, np.array([1, 0, 2, 0, 15], np.uint8))
        assert got.shape[0] == 3 and (got["node"] == 0).all() and (got["type"] == 2).all()
    tree = synth._flatten([[0, 1], 2])
    with api.Context(tree) as ctx:
        with pytest.raises(api.PmfsError) as e:  # a leaf refers to row 2 but only 2 rows are given
            ctx.run(api.FITCH, np.zeros((2, 3), np.uint8), 5, cons)
        assert e.value.status == -1


# getCodeFromNucleotide (src/panman.cpp:83-118) as a 256-entry table: the 15 upper-case IUPAC letters, all else 0
_CODE = np.zeros(256, np.uint8)
for _i, _ch in enumerate("ACMGRSVTWYHKDBN"):
    _CODE[ord(_ch)] = _i + 1


def test_run_chars_matches_codes_path():
    """pmfs_run_chars (characters in, conversion on the device) == pmfs_run on host-converted codes == oracle,
    on alignments that hold every byte value (lower case, '-', '\\0', >127 all map to the gap code)."""
    rng = np.random.default_rng(77)
    letters = np.frombuffer(b"ACGTACGTACGTNRYKMSWBDHV--", np.uint8)
    for trial, L in enumerate((1, 7, 8, 31, 33, 600, 4099)):
        tree = synth.random_tree(int(rng.integers(2, 120)), rng, max_arity=3, absent_frac=0.05 if trial % 2 else 0.0)
        rows = synth.n_rows(tree)
        chars = letters[rng.integers(0, len(letters), (rows, L))].copy()
        junk = rng.random((rows, L)) < 0.05
        chars[junk] = rng.integers(0, 256, int(junk.sum()), dtype=np.uint8)  # any byte value at all
        if L >= 600:
            chars[0, :256] = np.arange(256, dtype=np.uint8)
        codes = _CODE[chars]
        cons_chars = np.where(codes.max(axis=0) > 0, np.frombuffer(b"A", np.uint8)[0], np.uint8(ord("-"))).astype(np.uint8)
        cons = _CODE[cons_chars]
        forced_chars = chars[int(rng.integers(0, rows))] if trial % 2 else None
        forced = None if forced_chars is None else _CODE[forced_chars]
        for algo in (0, 1):
            want = O.run_port(algo, tree, codes, cons, forced, threads=2)
            with api.Context(tree) as ctx:
                got = ctx.run_chars(algo, chars, cons_chars, forced_chars)
            assert G.same_muts(got, want), (trial, algo)
            assert G.same_muts(got, run_gpu(tree, algo, codes, cons, forced))


def test_radix_ordering_path_in_subprocess():
    """Single-GPU runs order their records through the (node, region) cell matrix; the radix-sort path (bucketed
    emission + stable sort on the high node bits) is what column shards and very large cell matrices use.  Run a few of
    the parity tests above with that path forced (PMFS_ORDER is read once per process)."""
    import os
    import subprocess
    import sys

    env = dict(os.environ, PMFS_ORDER="radix")
    here = os.path.dirname(os.path.abspath(__file__))
    r = subprocess.run([sys.executable, "-m", "pytest", os.path.join(here, "test_gpu_parity.py"), "-q", "-x", "-m", "gpu", "-k",
                        "golden_small_cases or random_trees or deep_ladder or device_side_packing"],
                       env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]


def test_block_level_passes_through_run_codes():
    """the reference's block-level Fitch / Sankoff (3 states, one column per block) through pmfs_run_codes,
    against the reference's own blockFitch*New / blockSankoff* members.  Leaves without a sequence: empty set for Fitch
    (leaf_row = -2), an all-zero row for Sankoff (the block pass gives them the state "absent")."""
    rng = np.random.default_rng(41)
    for trial in range(10):
        arity = 2 if trial < 4 else int(rng.integers(2, 7))
        tree = synth.random_tree(int(rng.integers(2, 300)), rng, max_arity=arity, absent_frac=0.1 if trial % 3 == 0 else 0.0)
        rows, B = synth.n_rows(tree), int(rng.integers(1, 900))
        anc = rng.integers(0, 3, B)
        states = np.where(rng.random((rows, B)) < 0.8, anc[None, :], rng.integers(0, 3, (rows, B))).astype(np.uint8)
        forced = states[int(rng.integers(0, rows))] if trial % 2 else None
        cons = np.zeros(B, np.uint8)
        # Fitch: absent leaves stay absent
        want = O.run_ref_blocks(0, tree, states, forced)
        with api.Context(tree) as ctx:
            got = O.blocks_from_nuc(ctx.run_codes(api.FITCH, states, cons, forced))
        assert G.same_muts(got, want), (trial, "fitch")
        # Sankoff: absent leaves read an extra all-zero row
        lr = tree.leaf_row.copy()
        n_abs = int((lr == -2).sum())
        lr[lr == -2] = rows + np.arange(n_abs)  # (one row each: the library refuses two leaves on one row)
        t2 = type(tree)(tree.parent, tree.child_off, tree.child_idx, lr)
        st2 = np.vstack([states, np.zeros((n_abs, B), np.uint8)])
        want = O.run_ref_blocks(1, tree, states, forced)
        with api.Context(t2) as ctx:
            got = O.blocks_from_nuc(ctx.run_codes(api.SANKOFF, st2, cons, forced))
        assert G.same_muts(got, want), (trial, "sankoff")


def test_large_polytomies():
    """Nodes with tens to hundreds of children (outbreak trees are full of them): the any-arity joins of both
    algorithms, with Sankoff's counters rippling through several bit planes, against the plain-C oracle."""
    rng = np.random.default_rng(29)
    star = synth._flatten(list(range(300)))
    two_level, nxt = [], 0
    for _ in range(6):
        k = int(rng.integers(40, 70))
        two_level.append(list(range(nxt, nxt + k)))
        nxt += k
    two_level += list(range(nxt, nxt + 25))
    for tree in (star, synth._flatten(two_level)):
        L = 1100
        msa = synth.simulate_msa(tree, L, rng, p_sub=0.3, p_del=0.1, p_ins=0.2, p_n=0.05, p_amb=0.05)
        cons = synth.consensus_first_nongap(msa)
        for algo in (0, 1):
            for forced in (None, msa[3]):
                want = O.run_port(algo, tree, msa, cons, forced, threads=4)
                got = run_gpu(tree, algo, msa, cons, forced)
                assert G.same_muts(got, want), (tree.n_nodes, algo, forced is not None)


def test_split_polytomies():
    """Nodes with more children than a threshold are dealt out to part work items (partial AND / OR pairs for Fitch,
    partial bit-sliced counters for Sankoff) that the node combines; the parts also walk their children top-down.
    Small thresholds and parts so that small trees have many parts, nested splits, split roots, absent leaves; both
    algorithms alternate on ONE context (the programs of split nodes differ per algorithm: the context re-plans)."""
    rng = np.random.default_rng(31)
    for trial in range(10):
        tree = synth.random_tree(int(rng.integers(100, 1500)), rng, max_arity=int(rng.integers(4, 40)),
                                 absent_frac=0.0 if trial % 3 else 0.1)
        L = int(rng.integers(40, 900))
        msa = synth.simulate_msa(tree, L, rng, p_sub=0.1, p_del=0.05, p_ins=0.2, p_n=0.05, p_amb=0.05)
        cons = synth.consensus_first_nongap(msa)
        forced = msa[int(rng.integers(0, msa.shape[0]))] if trial % 2 else None
        nib = api.pack_nibbles(msa)
        want = [O.run_port(algo, tree, msa, cons, forced, threads=4) for algo in (0, 1)]
        with api.Context(tree) as ctx:
            min_arity = int(rng.choice([2, 5, 9]))
            ctx.set_poly_split(min_arity, int(rng.choice([1, 3, 7, 63])))
            if trial % 4 == 1:
                ctx.set_tile(1, 1)
            elif trial % 4 == 2:
                ctx.set_tile(4, 2)
            for algo in (0, 1, 1, 0):
                got = ctx.run(algo, nib, L, cons, forced)
                assert (ctx.plan_info()["n_parts"] > 0) == bool((np.diff(tree.child_off) > min_arity).any())
                assert G.same_muts(got, want[algo]), (trial, algo)
    # production thresholds (more than 128 children, 63 per part): a star of 20 000 and an outbreak-like tree -- a
    # backbone whose nodes each carry hundreds of identical-looking leaves
    star = synth._flatten(list(range(20000)))
    backbone, nxt = [], 0
    node = list(range(nxt, nxt + 400)); nxt += 400
    for _ in range(12):
        k = int(rng.integers(100, 700))
        node = [node] + list(range(nxt, nxt + k)); nxt += k
    for tree in (star, synth._flatten(node)):
        L = 2048 + 37
        msa = synth.simulate_msa(tree, L, rng, p_sub=0.02, p_del=0.01, p_ins=0.02, p_n=0.01, p_amb=0.01)
        cons = synth.consensus_first_nongap(msa)
        nib = api.pack_nibbles(msa)
        with api.Context(tree) as ctx, api.Context(tree) as whole:
            whole.set_poly_split(0, 0)  # never split: the any-arity join walks all children in one work item
            for algo in (0, 1):
                got = ctx.run(algo, nib, L, cons)
                assert ctx.plan_info()["n_parts"] >= 90
                assert G.same_muts(got, O.run_port(algo, tree, msa, cons, None, threads=8)), (tree.n_nodes, algo)
                ref = whole.run(algo, nib, L, cons)
                assert whole.plan_info()["n_parts"] == 0 and G.same_muts(got, ref)


def test_record_buffer_grows_and_passes_are_redone():
    """Far more mutations than the first call's record buffer is sized for (random codes in every cell: most cells
    mutate): the overflow is detected on the device, the buffer grows, the passes run again -- same records."""
    rng = np.random.default_rng(31)
    tree = synth.random_tree(48, rng, max_arity=2)
    L = 40000
    msa = rng.integers(0, 16, (synth.n_rows(tree), L)).astype(np.uint8)
    cons = np.maximum(msa[0], 1).astype(np.uint8)
    for algo in (0, 1):
        want = O.run_port(algo, tree, msa, cons, None, threads=4)
        assert want.shape[0] > (1 << 20)
        with api.Context(tree) as ctx:
            nib = api.pack_nibbles(msa)
            got = ctx.run(algo, nib, L, cons)
            again = ctx.run(algo, nib, L, cons)  # second call: the grown buffer is kept
        assert G.same_muts(got, want) and G.same_muts(again, want), algo


def test_no_records_then_records_on_one_context():
    """A run that emits nothing (all rows equal to the consensus), then one that does, then nothing again, on the same
    context: the per-run state that is only cleared lazily (cell counts, output buffer of the previous run) must not leak."""
    rng = np.random.default_rng(37)
    tree = synth.random_tree(200, rng, max_arity=3)
    L = 3000
    rows = synth.n_rows(tree)
    same = np.tile(rng.integers(1, 16, L).astype(np.uint8), (rows, 1))
    varied = synth.simulate_msa(tree, L, rng, p_sub=0.05, p_del=0.02, p_ins=0.05, p_n=0.02)
    with api.Context(tree) as ctx:
        for algo in (0, 1):
            for msa in (same, varied, same, varied):
                cons = msa[0].copy() if msa is same else synth.consensus_first_nongap(msa)
                want = O.run_port(algo, tree, msa, cons, None, threads=4)
                got = ctx.run(algo, api.pack_nibbles(msa), L, cons)
                assert (want.shape[0] == 0) == (msa is same)
                assert G.same_muts(got, want), (algo, msa is same)


def test_replaying_the_emitted_mutations_reproduces_every_leaf_row(c1):
    """Semantic round trip of the CUDA path's output, independent of any oracle (tests/replay.py): config 1 in two CLI
    variants with the device-side packing, and random polytomy trees with absent leaves and forced roots."""
    import replay
    tree = G.c1_tree(c1)
    L = int(c1["n_cols"])
    msa = G.unpack_nibbles(c1["msa_nibbles"], L)
    for low, with_ref in ((0, 0), (1, 1)):
        key = f"v{low}{with_ref}_"
        forced = msa[list(c1["row_ids"]).index(str(c1[key + "reference"]))] if with_ref else None
        with api.Context(tree) as ctx:
            ctx.run(low, api.pack_nibbles(msa), L, c1[key + "consensus"], forced)
            off, nuc = ctx.pack_last()
        replay.assert_replay_reproduces_msa(tree, msa, c1[key + "consensus"], off, nuc)
    rng = np.random.default_rng(18)
    for trial in range(8):
        t = synth.random_tree(int(rng.integers(3, 200)), rng, max_arity=6, absent_frac=0.1 if trial % 2 else 0.0)
        m = synth.simulate_msa(t, 700, rng, p_sub=0.08, p_del=0.05, p_ins=0.2, p_n=0.05, p_amb=0.05)
        cons = synth.consensus_first_nongap(m)
        forced = m[0] if trial % 3 == 0 else None
        for algo in (0, 1):
            with api.Context(t) as ctx:
                ctx.run(algo, api.pack_nibbles(m), 700, cons, forced)
                off, nuc = ctx.pack_last()
            replay.assert_replay_reproduces_msa(t, m, cons, off, nuc)
