"""CPU: the library's plan builder + walk interpreter (the host side of the code the CUDA kernels run per thread),
emulated one column group at a time by tests/host_emul.cpp, against the oracle and the golden fixtures."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

import golden_util as G
import synth
from oracle import oracle_api as O
from panman_b200.api import MUT_DTYPE, pack_nibbles

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "tests", "_build", "libhostemul.so")


@pytest.fixture(scope="module")
def emul():
    os.makedirs(os.path.dirname(SO), exist_ok=True)
    srcs = [os.path.join(ROOT, "tests", "host_emul.cpp"), os.path.join(ROOT, "panman_b200", "csrc", "plan.cpp")]
    deps = srcs + [os.path.join(ROOT, "panman_b200", "csrc", f) for f in ("walk.cuh", "plan.hpp")]
    if not os.path.exists(SO) or any(os.path.getmtime(d) > os.path.getmtime(SO) for d in deps):
        subprocess.run(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-Wno-unknown-pragmas",
                        "-I/usr/local/cuda/include", "-o", SO] + srcs, check=True)
    lib = C.CDLL(SO)
    lib.emul_free.argtypes = [C.c_void_p]
    lib.emul_set_split.argtypes = [C.c_int32, C.c_int32]
    return lib


def run_emul(lib, algo, tree, msa_codes, cons, forced, chunk_nodes, want_plan=False, groups=2):
    nib = pack_nibbles(msa_codes) if msa_codes.shape[0] else np.zeros((0, 1), np.uint8)
    n_cols = int(cons.shape[0])
    arrs = [np.ascontiguousarray(a, np.int32) for a in (tree.parent, tree.child_off, tree.child_idx, tree.leaf_row)]
    cons = np.ascontiguousarray(cons, np.uint8)
    frc = None if forced is None else np.ascontiguousarray(forced, np.uint8)
    out, n = C.c_void_p(), C.c_int64()
    plan = (C.c_int64 * 8)()
    rc = lib.emul_run(C.c_int(algo), C.c_int32(tree.n_nodes), *[C.c_void_p(a.ctypes.data) for a in arrs],
                      C.c_void_p(nib.ctypes.data), C.c_int64(nib.shape[1]), C.c_int64(n_cols), C.c_void_p(cons.ctypes.data),
                      None if frc is None else C.c_void_p(frc.ctypes.data), C.c_int32(chunk_nodes), C.c_int32(groups), C.byref(out), C.byref(n), plan)
    if rc != 0:
        raise RuntimeError(f"emul_run rc={rc}")
    res = np.frombuffer((C.c_char * (n.value * MUT_DTYPE.itemsize)).from_address(out.value), MUT_DTYPE).copy() if n.value \
        else np.zeros(0, MUT_DTYPE)
    lib.emul_free(out)
    return (res, list(plan)) if want_plan else res


def test_emul_matches_golden_small(emul, small_cases):
    z = small_cases
    for name in z["case_names"]:
        tree, msa, cons, forced = G.case_inputs(z, str(name))
        for algo in (0, 1):
            for chunk, groups in ((1 << 30, 1), (4, 2)):
                got = run_emul(emul, algo, tree, msa, cons, forced, chunk, groups=groups)
                assert G.same_muts(got, z[f"{name}/muts{algo}"]), (name, algo, chunk, groups)


@pytest.mark.parametrize("low,with_ref", [(0, 0), (0, 1), (1, 0), (1, 1)])
def test_emul_matches_golden_c1(emul, c1, low, with_ref):
    key = f"v{low}{with_ref}_"
    tree = G.c1_tree(c1)
    msa = G.unpack_nibbles(c1["msa_nibbles"], int(c1["n_cols"]))
    forced = msa[list(c1["row_ids"]).index(str(c1[key + "reference"]))] if with_ref else None
    got = run_emul(emul, low, tree, msa, c1[key + "consensus"], forced, 8)
    assert got.shape[0] == int(c1[key + "n_muts"])
    assert G.digest_muts(got) == str(c1[key + "muts_sha256"])


def test_emul_matches_oracle_random(emul):
    rng = np.random.default_rng(11)
    for trial in range(40):
        arity = 2 if trial < 16 else int(rng.integers(2, 7))
        tree = synth.random_tree(int(rng.integers(2, 200)), rng, max_arity=arity, absent_frac=0.0 if trial % 3 else 0.1)
        L = int(rng.integers(1, 150))
        msa = synth.simulate_msa(tree, L, rng, p_sub=0.1, p_del=0.05, p_ins=0.2, p_n=0.05, p_amb=0.05)
        cons = synth.consensus_first_nongap(msa)
        forced = msa[int(rng.integers(0, msa.shape[0]))] if trial % 2 else None
        chunk = int(rng.choice([2, 3, 8, 32, 1 << 30]))
        for algo in (0, 1):
            want = O.run_port(algo, tree, msa, cons, forced, threads=2)
            got, plan = run_emul(emul, algo, tree, msa, cons, forced, chunk, want_plan=True, groups=1 + trial % 2)
            assert G.same_muts(got, want), (trial, algo, chunk, plan)


def test_emul_ladder_and_balanced_stack_depth(emul):
    rng = np.random.default_rng(5)
    ladder = synth.caterpillar(3000)
    msa = synth.simulate_msa(ladder, 40, rng, p_sub=0.02)
    cons = synth.consensus_first_nongap(msa)
    for algo in (0, 1):
        want = O.run_port(algo, ladder, msa, cons, None)
        got, plan = run_emul(emul, algo, ladder, msa, cons, None, 1 << 30, want_plan=True)
        assert G.same_muts(got, want)
        assert plan[3] <= 1 and plan[4] <= 1  # a ladder needs no stack at all: the heavy child is walked in place
        got = run_emul(emul, algo, ladder, msa, cons, None, 100)
        assert G.same_muts(got, want)
    # perfectly balanced tree in ONE chunk request: the plan cuts wherever a walk would nest more than two saved states
    # (every stack level costs shared memory per CTA, i.e. resident CTAs), and the result does not change
    def bal(lo, hi):
        return lo if hi - lo == 1 else [bal(lo, (lo + hi) // 2), bal((lo + hi) // 2, hi)]
    tree = synth._flatten(bal(0, 256))
    msa = synth.simulate_msa(tree, 40, rng, p_sub=0.05)
    cons = synth.consensus_first_nongap(msa)
    for algo in (0, 1):
        got, plan = run_emul(emul, algo, tree, msa, cons, None, 1 << 30, want_plan=True)
        assert plan[3] <= 2 and plan[4] <= 2 and plan[0] > 1
        assert G.same_muts(got, O.run_port(algo, tree, msa, cons, None))


def test_emul_walk_stack_is_capped_on_random_trees(emul):
    rng = np.random.default_rng(17)
    for trial in range(8):
        tree = synth.random_tree(int(rng.integers(200, 1500)), rng, max_arity=2 if trial % 2 else 5)
        msa = synth.simulate_msa(tree, 33, rng, p_sub=0.05, p_del=0.02, p_ins=0.05)
        cons = synth.consensus_first_nongap(msa)
        for chunk in (24, 200, 1 << 30):
            got, plan = run_emul(emul, trial % 2, tree, msa, cons, None, chunk, want_plan=True)
            assert plan[3] <= 2 and plan[4] <= 2, (trial, chunk, plan)
            assert G.same_muts(got, O.run_port(trial % 2, tree, msa, cons, None)), (trial, chunk)


def test_emul_sankoff_invalid_root(emul):
    tree = synth._flatten([None, None])
    with pytest.raises(RuntimeError):
        run_emul(emul, 1, tree, np.zeros((0, 3), np.uint8), np.zeros(3, np.uint8), None, 1 << 30)
    assert run_emul(emul, 0, tree, np.zeros((0, 3), np.uint8), np.zeros(3, np.uint8), None, 1 << 30).shape[0] == 0
    assert run_emul(emul, 1, tree, np.zeros((0, 3), np.uint8), np.zeros(3, np.uint8), np.array([1, 0, 2], np.uint8), 1 << 30).shape[0] == 2


def test_emul_large_polytomies(emul):
    """Any-arity joins far beyond what the random trees reach: a star of 300 leaves, a two-level tree with nodes of 40-70
    children, and polytomies made by collapsing the edges of a binary tree (Sankoff's counters need up to 9 bits here,
    i.e. the flush of the 2-bit register counters into the full-width ones ripples through several planes)."""
    rng = np.random.default_rng(23)
    star = synth._flatten(list(range(300)))
    two_level, nxt = [], 0
    for _ in range(6):
        k = int(rng.integers(40, 70))
        two_level.append(list(range(nxt, nxt + k)))
        nxt += k
    two_level += list(range(nxt, nxt + 25))
    trees_ = [star, synth._flatten(two_level)]
    for tree in trees_:
        L = 70
        msa = synth.simulate_msa(tree, L, rng, p_sub=0.3, p_del=0.1, p_ins=0.2, p_n=0.05, p_amb=0.05)
        cons = synth.consensus_first_nongap(msa)
        forced = msa[3]
        for algo in (0, 1):
            for frc in (None, forced):
                want = O.run_port(algo, tree, msa, cons, frc)
                for chunk in (24, 1 << 30):
                    got = run_emul(emul, algo, tree, msa, cons, frc, chunk)
                    assert G.same_muts(got, want), (tree.n_nodes, algo, chunk)


def test_emul_split_polytomies(emul):
    """Nodes with more children than the threshold are dealt out to part chunks (partial AND / OR pairs, partial
    counters) that the node's own chunk combines; the parts also walk their children top-down.  Small thresholds and
    parts here so that small trees have many parts, mixed leaf / absent / internal children, split roots and nested splits."""
    rng = np.random.default_rng(23)
    try:
        for trial in range(36):
            arity = int(rng.integers(3, 40))
            tree = synth.random_tree(int(rng.integers(5, 400)), rng, max_arity=arity, absent_frac=0.0 if trial % 3 else 0.15)
            L = int(rng.integers(1, 100))
            msa = synth.simulate_msa(tree, L, rng, p_sub=0.1, p_del=0.05, p_ins=0.2, p_n=0.05, p_amb=0.05)
            cons = synth.consensus_first_nongap(msa)
            forced = msa[int(rng.integers(0, msa.shape[0]))] if trial % 2 else None
            chunk = int(rng.choice([2, 8, 32, 1 << 30]))
            split, part = int(rng.choice([2, 3, 5, 9])), int(rng.choice([1, 2, 3, 7, 63]))
            emul.emul_set_split(split, part)
            n_parts = 0
            for algo in (0, 1):
                want = O.run_port(algo, tree, msa, cons, forced, threads=2)
                got, plan = run_emul(emul, algo, tree, msa, cons, forced, chunk, want_plan=True, groups=1 + trial % 2)
                assert G.same_muts(got, want), (trial, algo, chunk, split, part, plan)
                n_parts = plan[6]
            arities = np.diff(tree.child_off)
            assert (n_parts > 0) == bool((arities > split).any())
        # a star and a star of stars with the production thresholds (more than 128 children, 63 per part)
        emul.emul_set_split(128, 63)
        for shape in ("star", "nested"):
            if shape == "star":
                tree = synth._flatten(list(range(700)))
            else:
                tree = synth._flatten([list(range(i * 150, (i + 1) * 150)) for i in range(4)] + list(range(600, 900)))
            msa = synth.simulate_msa(tree, 70, rng, p_sub=0.05, p_del=0.02, p_ins=0.05, p_n=0.02, p_amb=0.02)
            cons = synth.consensus_first_nongap(msa)
            for algo in (0, 1):
                got, plan = run_emul(emul, algo, tree, msa, cons, None, 64, want_plan=True)
                assert plan[6] == (12 if shape == "star" else 4 * 3 + 5), plan
                assert G.same_muts(got, O.run_port(algo, tree, msa, cons, None)), (shape, algo)
                if O.ref_available():  # ... and against the reference's own fitchSankoff.cpp, not only its restatement
                    ref = O.run_ref_columns(algo, tree, O.codes_to_chars(msa), O.codes_to_chars(cons), None, threads=2)
                    assert G.same_muts(got, ref), (shape, algo, "ref")
    finally:
        emul.emul_set_split(128, 63)
