"""Host C++ mirror of the reference's MSA-driven Tree constructor (panman_b200/csrc/host/): Newick parser, FASTA
reader, consensus, <=6-run packing.  CPU tests cover the device-free pieces; the GPU tests build BASELINE config 1
end to end from FASTA + Newick text and compare with the golden outputs of the reference."""
import hashlib
import os

import numpy as np
import pytest

import golden_util as G
import synth
from oracle import oracle_api as O
from panman_b200 import api, host

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module", autouse=True)
def _lib():
    api.build_library()


def test_newick_matches_reference_numbering_c1(c1):
    names, parent, nch = host.parse_newick(str(c1["newick"]))
    assert names == [str(x) for x in c1["names"]]
    assert (parent == c1["parent"]).all()
    assert (nch == np.diff(c1["child_off"])).all()


def test_newick_quirks():
    # internal labels and branch lengths are ignored for naming; quoted names keep commas; quotes are stripped
    names, parent, _ = host.parse_newick("((A:0.1,'B,b':2)lab:3,(C,D)x)root;")
    assert names == ["node_1", "node_2", "A", "B,b", "node_3", "C", "D"]
    assert parent.tolist() == [-1, 0, 1, 1, 0, 4, 4]
    names, _, nch = host.parse_newick("  (t0,t1,(t2,t3,t4),t5);  ")
    assert names == ["node_1", "t0", "t1", "node_2", "t2", "t3", "t4", "t5"] and nch[0] == 4 and nch[3] == 3
    with pytest.raises(api.PmfsError):
        host.parse_newick("((A,B);")
    with pytest.raises(api.PmfsError):
        host.parse_newick("(A,B));")


@pytest.mark.skipif(not O.ref_available(), reason="oracle/_ref absent")
def test_newick_and_fasta_reader_match_reference_driver(tmp_path):
    """Same text through oracle/ref_driver.cpp (pmref_build) and through the product parser: identical trees."""
    rng = np.random.default_rng(3)
    for trial in range(6):
        tree = synth.random_tree(int(rng.integers(3, 40)), rng, max_arity=4)
        names = {v: (f"'q{v},x'" if trial % 2 and v % 5 == 0 else f"s{v}|x/{v}") for v in range(tree.n_nodes)}
        nwk = synth.to_newick(tree, names=[names[v] for v in range(tree.n_nodes)], branch_lengths=bool(trial % 2))
        msa = synth.simulate_msa(tree, 40, rng)
        fa = tmp_path / f"m{trial}.fa"
        with open(fa, "w") as f:
            for v in range(tree.n_nodes):
                r = tree.leaf_row[v]
                if r >= 0:
                    f.write(">" + names[v].strip("'") + " some description\n" + O.codes_to_chars(msa[r]).tobytes().decode() + "\n")
        nw = tmp_path / f"t{trial}.nwk"
        nw.write_text(nwk + "\n")
        ref = O.ref_build(str(fa), str(nw))
        got_names, got_parent, got_nch = host.parse_newick(nwk)
        assert got_names == ref["names"]
        assert (got_parent == ref["parent"]).all() and (got_nch == np.diff(ref["child_off"])).all()


def test_pack_matches_oracle_rule(small_cases):
    z = small_cases
    for name in z["case_names"]:
        for algo in (0, 1):
            m = z[f"{name}/muts{algo}"]
            mine = np.zeros(m.shape[0], api.MUT_DTYPE)
            for k in ("node", "pos", "type", "code"):
                mine[k] = m[k]
            off, nuc = host.pack(mine, int(z[f"{name}/parent"].shape[0]))
            assert (off == z[f"{name}/nuc_off{algo}"]).all() and G.same_nuc(nuc, z[f"{name}/nuc{algo}"]), (name, algo)


def _write_c1(c1, tmp_path, crlf=False, description=" padded"):
    L = int(c1["n_cols"])
    msa = G.unpack_nibbles(c1["msa_nibbles"], L)
    rows = {str(k): i for i, k in enumerate(c1["row_ids"])}
    fa = tmp_path / "sars_20.padded.msa"
    eol = "\r\n" if crlf else "\n"
    with open(fa, "w", newline="") as f:
        for k in c1["fasta_order"]:
            f.write(">" + str(k) + description + eol + O.codes_to_chars(msa[rows[str(k)]]).tobytes().decode() + eol)
    nw = tmp_path / "sars_20.nwk"
    nw.write_text(str(c1["newick"]) + "\n")
    return str(fa), str(nw)


@pytest.mark.gpu
@pytest.mark.parametrize("low,with_ref", [(0, 0), (0, 1), (1, 0), (1, 1)])
def test_build_c1_end_to_end(c1, tmp_path, low, with_ref):
    """`panmanUtils -M sars_20.padded.msa -N sars_20.nwk [--reference id] [--low-mem-mode]` up to serialisation."""
    key = f"v{low}{with_ref}_"
    fa, nw = _write_c1(c1, tmp_path)
    r = host.build(fa, nw, str(c1[key + "reference"]), bool(low))
    assert r["names"] == [str(x) for x in c1["names"]]
    assert (O.chars_to_codes(np.frombuffer(r["consensus"], np.uint8)) == c1[key + "consensus"]).all()
    assert r["muts"].shape[0] == int(c1[key + "n_muts"])
    assert G.digest_muts(r["muts"]) == str(c1[key + "muts_sha256"])
    assert r["nuc"].shape[0] == int(c1[key + "n_nuc"])
    assert G.digest_nuc(r["nuc_off"], r["nuc"]) == str(c1[key + "nuc_sha256"])
    assert G.same_nuc(r["nuc"][:256], c1[key + "nuc_head"])
    assert hashlib.sha256(r["block_words"].astype("<u4").tobytes()).hexdigest() == str(c1[key + "block_words_sha256"])


@pytest.mark.gpu
def test_build_input_errors(c1, tmp_path):
    fa, nw = _write_c1(c1, tmp_path)
    with open(fa, "a") as f:
        f.write(">short\nACGT\n")
    with pytest.raises(api.PmfsError) as e:
        host.build(fa, nw)
    assert e.value.status == -11  # the reference exit(-1)s on unequal lengths (src/panman.cpp:1284-1290)
    fa, nw = _write_c1(c1, tmp_path)
    with pytest.raises(api.PmfsError) as e:
        host.build(fa, nw, "no_such_id", True)
    assert e.value.status == -13  # low-mem mode: "Reference not found" (src/panman.cpp:1582-1585)


@pytest.mark.gpu
@pytest.mark.skipif(not (O.ref_available() and os.path.exists(O.REF_SO)), reason="oracle/_ref absent")
@pytest.mark.parametrize("low", [0, 1])
@pytest.mark.parametrize("description", [" padded", ""])
def test_build_c1_with_crlf_line_ends(c1, tmp_path, low, description):
    """CRLF input.  The default reader strips '\\r' from rows and from every id but the LAST sequence's
    (src/panman.cpp:1291-1302: with no description after the id that sequence then matches no tree leaf); the low-mem
    reader (:667-714) strips nothing, so every row carries one more, all-'\\r' column.  Whatever the reference's driver
    makes of it (oracle/ref_driver.cpp restates it), the mirror must make the same."""
    fa, nw = _write_c1(c1, tmp_path, crlf=True, description=description)
    try:
        ref = O.ref_build(fa, nw, "", bool(low), threads=4)
    except RuntimeError as e:  # the reference's driver gives up on this input: so must the mirror
        with pytest.raises(api.PmfsError):
            host.build(fa, nw, "", bool(low))
        assert "rc=" in str(e)
        return
    r = host.build(fa, nw, "", bool(low))
    assert r["names"] == ref["names"] and r["consensus"] == ref["consensus"]
    assert G.same_muts(r["muts"], ref["muts"])
    assert (r["nuc_off"] == ref["nuc_off"]).all() and G.same_nuc(r["nuc"], ref["nuc"])
    assert (r["block_words"] == ref["block_words"]).all()


@pytest.mark.gpu
@pytest.mark.parametrize("low,with_ref", [(0, 0), (1, 1)])
def test_build_panman_file_c1(c1, tmp_path, low, with_ref):
    """The whole `panmanUtils -M .. -N .. [--reference id] [--low-mem-mode] -o out` on the GPU, .panman file included: what
    an independent reader decodes from the file is the golden per-node NucMut CSR."""
    from panman_reader import nuc_csr, read_panman
    key = f"v{low}{with_ref}_"
    fa, nw = _write_c1(c1, tmp_path)
    out = str(tmp_path / "c1.panman")
    host.build_panman(fa, nw, out, str(c1[key + "reference"]), bool(low))
    d = read_panman(out)
    off, nuc = nuc_csr(d)
    assert nuc.shape[0] == int(c1[key + "n_nuc"]) and G.digest_nuc(off, nuc) == str(c1[key + "nuc_sha256"])
    assert hashlib.sha256(d["consensusSeqMap"][0]["consensusSeq"].astype("<u4").tobytes()).hexdigest() == str(c1[key + "block_words_sha256"])
    names, parent, _ = host.parse_newick(d["newick"])
    assert names == [str(x) for x in c1["names"]] and (parent == c1["parent"]).all()


@pytest.mark.gpu
def test_command_line_example_writes_the_same_panman(c1, tmp_path):
    """examples/msa_to_panman.c -- the C program a user would build against libpmfs.so -- on BASELINE config 1 with
    --reference and --low-mem-mode: the file it writes decodes to the golden NucMut lists; and it reports the
    reference's own complaint (unequal row lengths) instead of writing anything."""
    import subprocess
    from panman_reader import nuc_csr, read_panman
    exe = str(tmp_path / "msa_to_panman")
    r = subprocess.run(["gcc", "-std=c99", "-Wall", "-Wextra", "-pedantic", "-Werror", "-I" + os.path.join(ROOT, "include"),
                        os.path.join(ROOT, "examples", "msa_to_panman.c"), "-L" + os.path.join(ROOT, "panman_b200"), "-lpmfs",
                        "-Wl,-rpath," + os.path.join(ROOT, "panman_b200"), "-o", exe], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    key = "v11_"
    fa, nw = _write_c1(c1, tmp_path)
    out = str(tmp_path / "cli")
    r = subprocess.run([exe, fa, nw, out, "--reference", str(c1[key + "reference"]), "--low-mem-mode"], capture_output=True, text=True)
    assert r.returncode == 0 and "wrote" in r.stdout, r.stdout + r.stderr
    off, nuc = nuc_csr(read_panman(out + ".panman"))
    assert nuc.shape[0] == int(c1[key + "n_nuc"]) and G.digest_nuc(off, nuc) == str(c1[key + "nuc_sha256"])
    with open(fa, "a") as f:
        f.write(">short\nACGT\n")
    r = subprocess.run([exe, fa, nw, str(tmp_path / "bad")], capture_output=True, text=True)
    assert r.returncode == 1 and "lengths" in r.stderr and not os.path.exists(str(tmp_path / "bad.panman"))
