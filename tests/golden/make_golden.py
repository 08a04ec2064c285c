"""Regenerates the golden fixtures in this directory FROM THE REFERENCE ITSELF.

Run in the build container only (needs /root/reference):  python tests/golden/make_golden.py
Every expected output below is produced by oracle/_ref/libpmref.so, i.e. the reference's own src/fitchSankoff.cpp
compiled in place, driven by oracle/ref_driver.cpp.  The reference ships no golden vectors of its own.

Outputs
  c1_sars20.npz    BASELINE config 1: test/sars_20.fa right-padded with '-' to 29 903 columns + test/sars_20.nwk, through
                   the whole construction path (Newick parse, FASTA read, consensus, column loop, sort, <=6-run packing)
                   in 4 variants {Fitch, Sankoff(--low-mem-mode)} x {no reference, --reference <first id>}.  Inputs are
                   stored nibble-packed; the 270k-record outputs as SHA-256 digests + counts + the first 512 records.
  small_cases.npz  hand-made and random small trees (arity 1..5, absent leaves, forced roots, all 16 codes) with full
                   expected (node,pos,type,code) records and packed NucMut arrays for both algorithms.
"""
import hashlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", ".."))
sys.path.insert(0, os.path.join(HERE, ".."))
import synth  # noqa: E402
from oracle import oracle_api as O  # noqa: E402


def digest_muts(m):
    a = np.stack([m["node"].astype("<i4"), m["pos"].astype("<i4"), m["type"].astype("<i4"), m["code"].astype("<i4")], 1)
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def digest_nuc(off, nuc):
    a = np.stack([nuc[k].astype("<i8") for k in nuc.dtype.names], 1)
    return hashlib.sha256(np.ascontiguousarray(off.astype("<i8")).tobytes() + np.ascontiguousarray(a).tobytes()).hexdigest()


def pack_nibbles(codes):
    c = np.asarray(codes, dtype=np.uint8)
    if c.shape[-1] % 2:
        c = np.concatenate([c, np.zeros(c.shape[:-1] + (1,), np.uint8)], -1)
    return (c[..., 0::2] | (c[..., 1::2] << 4)).astype(np.uint8)


def make_c1():
    ref = "/root/reference/test"
    seqs, order, cur = {}, [], None
    for line in open(os.path.join(ref, "sars_20.fa")):
        line = line.rstrip("\n")
        if not line:
            continue
        if line[0] == ">":
            cur = line.split(" ")[0][1:]
            seqs[cur] = ""
            order.append(cur)
        else:
            seqs[cur] += line
    L = max(len(s) for s in seqs.values())
    tmp = "/tmp/pmfs_c1"
    os.makedirs(tmp, exist_ok=True)
    msa_path = os.path.join(tmp, "sars_20.padded.msa")
    with open(msa_path, "w") as f:
        for k in order:
            f.write(">" + k + "\n" + seqs[k] + "-" * (L - len(seqs[k])) + "\n")
    nwk_path = os.path.join(ref, "sars_20.nwk")
    newick = open(nwk_path).readline().rstrip("\n")
    ids_sorted = sorted(order)  # std::map order == row order of the flat problem
    chars = np.array([np.frombuffer((seqs[k] + "-" * (L - len(seqs[k]))).encode(), np.uint8) for k in ids_sorted])
    codes = O.chars_to_codes(chars)
    out = {"newick": np.array(newick), "row_ids": np.array(ids_sorted), "fasta_order": np.array(order),
           "n_cols": np.int64(L), "msa_nibbles": pack_nibbles(codes)}
    first = True
    for low in (0, 1):
        for with_ref in (0, 1):
            r = O.ref_build(msa_path, nwk_path, order[0] if with_ref else "", bool(low), threads=8)
            if first:
                out["parent"], out["child_off"], out["child_idx"] = r["parent"], r["child_off"], r["child_idx"]
                out["names"] = np.array(r["names"])
                row_of = {k: i for i, k in enumerate(ids_sorted)}
                leaf = r["child_off"][1:] == r["child_off"][:-1]
                out["leaf_row"] = np.array([row_of.get(n, -2) if leaf[i] else -1 for i, n in enumerate(r["names"])], np.int32)
                first = False
            key = f"v{low}{with_ref}_"
            out[key + "reference"] = np.array(order[0] if with_ref else "")
            out[key + "consensus"] = O.chars_to_codes(np.frombuffer(r["consensus"], np.uint8))
            out[key + "n_muts"] = np.int64(r["muts"].shape[0])
            out[key + "muts_sha256"] = np.array(digest_muts(r["muts"]))
            out[key + "muts_head"] = r["muts"][:512]
            out[key + "type_counts"] = np.bincount(r["muts"]["type"], minlength=3)
            out[key + "n_nuc"] = np.int64(r["nuc"].shape[0])
            out[key + "nuc_sha256"] = np.array(digest_nuc(r["nuc_off"], r["nuc"]))
            out[key + "nuc_head"] = r["nuc"][:256]
            out[key + "block_words_sha256"] = np.array(hashlib.sha256(r["block_words"].astype("<u4").tobytes()).hexdigest())
            print("c1", key, r["muts"].shape[0], r["nuc"].shape[0])
    np.savez_compressed(os.path.join(HERE, "c1_sars20.npz"), **out)


def hand_cases():
    """(name, nested tree, msa rows as strings)"""
    T = synth._flatten
    cases = []
    # cherry, tie at the root: {A},{C} -> union, lowest bit wins (A)
    cases.append(("root_tie", T([0, 1]), ["ACGT-N", "CCGA-A"]))
    # 3-ary node where an absent child turns AND into 0 -> OR (Fitch) but is skipped by Sankoff
    cases.append(("absent_child_polytomy", T([0, 1, None]), ["AAAC", "AACC"]))
    cases.append(("absent_child_binary", T([[0, None], [1, 2]]), ["AC-T", "ACGT", "CC-T"]))
    # Sankoff d[p]==1 three-way tie: parent state p has cost min+1 in the child, first index wins
    cases.append(("sankoff_d1_tie", T([[0, 1], [2, [3, 4]], 5]), ["A", "C", "G", "T", "C", "G"]))
    cases.append(("sankoff_d1_tie_wide", T([[[0, 1], [2, 3]], [[4, 5], [6, 7]], 8]),
                  ["ACGTACGT", "CAGTTGCA", "GGTTAACC", "TTAACCGG", "NNRYACGT", "ACGT----", "--ACGT--", "AAAAAAAA", "CCCCCCCC"]))
    cases.append(("all_n_column", T([[0, 1], [2, 3]]), ["NNA", "NNA", "NNC", "NNA"]))
    cases.append(("gaps_indels", T([[0, 1], [2, [3, 4]]]), ["A--CG", "A-TCG", "-GTC-", "-G-C-", "AG-CT"]))
    cases.append(("unary_chain", T([[[0]], [1, [2]]]), ["ACG", "CCG", "ACT"]))
    cases.append(("all_codes", T([[0, 1], [2, 3], [4, [5, 6, 7]]]),
                  ["-ACMGRSVTWYHKDBN", "NBDKHYWTVSRGMCA-", "ACMGRSVTWYHKDBN-", "AAAAAAAAAAAAAAAA",
                   "----------------", "NNNNNNNNNNNNNNNN", "-ACMGRSVTWYHKDBN", "TTTTTTTTCCCCCCCC"]))
    cases.append(("single_leaf_root", T([0]), ["ACGT"]))
    cases.append(("absent_subtree", T([[None, None], [0, 1]]), ["AC", "AG"]))
    return cases


def make_small():
    rng = np.random.default_rng(20251106)
    out = {}
    names = []

    def add(name, tree, msa_codes, forced_row):
        cons = synth.consensus_first_nongap(msa_codes)
        forced = None if forced_row is None else msa_codes[forced_row]
        if forced is not None:
            cons = forced.copy()  # Fitch branch with --reference: consensus = the reference row (src/panman.cpp:1322)
        names.append(name)
        out[name + "/parent"], out[name + "/child_off"] = tree.parent, tree.child_off
        out[name + "/child_idx"], out[name + "/leaf_row"] = tree.child_idx, tree.leaf_row
        out[name + "/msa"] = msa_codes
        out[name + "/consensus"] = cons
        out[name + "/forced"] = np.zeros(0, np.uint8) if forced is None else forced
        for algo in (0, 1):
            m = O.run_ref_columns(algo, tree, O.codes_to_chars(msa_codes), O.codes_to_chars(cons),
                                  None if forced is None else O.codes_to_chars(forced), threads=4)
            off, nuc = O.pack_ref(m, tree.n_nodes)
            out[f"{name}/muts{algo}"] = m
            out[f"{name}/nuc_off{algo}"] = off
            out[f"{name}/nuc{algo}"] = nuc

    for name, tree, rows in hand_cases():
        codes = O.chars_to_codes(np.array([np.frombuffer(r.encode(), np.uint8) for r in rows]))
        add("hand_" + name, tree, codes, None)
        add("hand_" + name + "_forced", tree, codes, 0)
    for i in range(24):
        arity = [2, 2, 3, 5][i % 4]
        absent = 0.0 if i % 3 else 0.08
        n_leaves = int(rng.integers(2, 70))
        tree = synth.random_tree(n_leaves, rng, max_arity=arity, absent_frac=absent)
        L = int(rng.integers(1, 200))
        if i % 5 == 0:
            msa = synth.simulate_msa(tree, L, rng, p_sub=0.3, p_del=0.1, p_ins=0.3, p_n=0.1, p_amb=0.2, root_gap_frac=0.3)
        else:
            msa = synth.simulate_msa(tree, L, rng)
        add(f"rand{i:02d}", tree, msa, None if i % 2 else int(rng.integers(0, msa.shape[0])))
    deep = synth.caterpillar(300)
    add("ladder300", deep, synth.simulate_msa(deep, 64, rng, p_sub=0.05), None)
    out["case_names"] = np.array(names)
    np.savez_compressed(os.path.join(HERE, "small_cases.npz"), **out)
    print("small cases:", len(names))


if __name__ == "__main__":
    O.build(force=True)
    make_c1()
    make_small()
