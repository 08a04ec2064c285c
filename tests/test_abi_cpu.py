"""CPU: the C-ABI library loads, exports every symbol include/pmfs.h declares, validates trees without a device, and
fails loudly (no fallback) when no CUDA device is present."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import synth
from panman_b200 import api

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    api.build_library()
    return api.lib()


def test_exports_match_header(lib):
    hdr = open(os.path.join(ROOT, "include", "pmfs.h")).read() + open(os.path.join(ROOT, "include", "pmfs_host.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    hdr = re.sub(r"static inline[^{]*\{[^}]*\}", "", hdr)  # header-only helpers are not exports
    declared = set(re.findall(r"\b(pmfs_[a-z_0-9]+)\s*\(", hdr))
    assert declared, "no declarations found"
    for name in sorted(declared):
        assert hasattr(lib, name), f"{name} declared in include/pmfs.h but not exported"
    from panman_b200 import host
    assert declared == set(api.EXPORTS) | set(host.HOST_EXPORTS), declared ^ (set(api.EXPORTS) | set(host.HOST_EXPORTS))
    assert b"sm_100a" in lib.pmfs_version()


def _has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def test_malformed_trees_rejected_without_device(lib):
    good = synth._flatten([[0, 1], 2])
    bad = synth._flatten([[0, 1], 2])
    bad.parent = bad.parent.copy()
    bad.parent[2] = 3  # parent after child: not a pre-order numbering
    with pytest.raises(api.PmfsError) as e:
        api.Context(bad)
    assert e.value.status == -1
    dup = synth._flatten([[0, 1], 2])
    dup.leaf_row = dup.leaf_row.copy()
    dup.leaf_row[dup.leaf_row == 2] = 1  # two leaves on one MSA row (the reference would alias them by name)
    with pytest.raises(api.PmfsError) as e:
        api.Context(dup)
    assert e.value.status == -1
    if not _has_gpu():
        with pytest.raises(api.PmfsError) as e:
            api.Context(good)
        assert e.value.status == -2 and "no CPU path" in str(e.value)


def test_nibble_packing_roundtrip():
    rng = np.random.default_rng(0)
    for cols in (1, 2, 7, 32, 33):
        codes = rng.integers(0, 16, size=(3, cols), dtype=np.uint8)
        p = api.pack_nibbles(codes)
        assert p.shape == (3, (cols + 1) // 2)
        assert (api.unpack_nibbles(p, cols) == codes).all()
        assert (p[:, 0] & 15 == codes[:, 0]).all()  # column 0 in the LOW nibble


def _build_c_example(tmp_path):
    """examples/fitch_from_c.c: the headers as C99 (-pedantic), linked against libpmfs.so only."""
    import subprocess
    exe = str(tmp_path / "fitch_from_c")
    r = subprocess.run(["gcc", "-std=c99", "-Wall", "-Wextra", "-pedantic", "-Werror", "-I" + os.path.join(ROOT, "include"),
                        os.path.join(ROOT, "examples", "fitch_from_c.c"), "-L" + os.path.join(ROOT, "panman_b200"), "-lpmfs",
                        "-Wl,-rpath," + os.path.join(ROOT, "panman_b200"), "-o", exe], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    return exe


def test_headers_are_c99_and_cxx17_and_the_c_example_links(lib, tmp_path):
    import subprocess
    inc = "-I" + os.path.join(ROOT, "include")
    for cmd in (["gcc", "-std=c99", "-pedantic", "-Werror", "-fsyntax-only", "-x", "c", inc, os.path.join(ROOT, "include", "pmfs_host.h")],
                ["g++", "-std=c++17", "-Werror", "-fsyntax-only", "-x", "c++", inc, os.path.join(ROOT, "include", "pmfs_host.h")]):
        r = subprocess.run(cmd, capture_output=True, text=True)
        assert r.returncode == 0, r.stderr
    r = subprocess.run(["gcc", "-std=c99", "-Wall", "-Wextra", "-pedantic", "-Werror", inc, os.path.join(ROOT, "examples", "msa_to_panman.c"),
                        "-L" + os.path.join(ROOT, "panman_b200"), "-lpmfs", "-o", str(tmp_path / "msa_to_panman")], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    exe = _build_c_example(tmp_path)
    if not _has_gpu():  # no device: the product fails loudly instead of falling back to anything
        r = subprocess.run([exe], capture_output=True, text=True)
        assert r.returncode == 2 and "no CPU path" in r.stderr


@pytest.mark.gpu
def test_c_example_prints_the_oracles_records(lib, tmp_path):
    import subprocess
    from oracle import oracle_api as O
    exe = _build_c_example(tmp_path)
    r = subprocess.run([exe], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    tree = synth._flatten([[0, 1], 2])
    msa = np.array([[1, 2, 4, 8], [1, 2, 8, 8], [1, 4, 4, 0]], np.uint8)
    want = O.run_port(0, tree, msa, np.array([1, 2, 4, 8], np.uint8), None)
    lines = [f"node {m['node']} pos {m['pos']} type {m['type']} code {m['code']}" for m in want]
    assert r.stdout.strip().splitlines() == lines and len(lines) > 0
