"""Helpers shared by the parity tests: golden-case access and digests (same definitions as tests/golden/make_golden.py)."""
import hashlib

import numpy as np

from oracle.oracle_api import FlatTree


def digest_muts(m):
    a = np.stack([m["node"].astype("<i4"), m["pos"].astype("<i4"), m["type"].astype("<i4"), m["code"].astype("<i4")], 1)
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def digest_nuc(off, nuc):
    a = np.stack([nuc[k].astype("<i8") for k in nuc.dtype.names], 1)
    return hashlib.sha256(np.ascontiguousarray(off.astype("<i8")).tobytes() + np.ascontiguousarray(a).tobytes()).hexdigest()


def unpack_nibbles(packed, n_cols):
    p = np.asarray(packed, dtype=np.uint8)
    out = np.empty(p.shape[:-1] + (p.shape[-1] * 2,), np.uint8)
    out[..., 0::2] = p & 15
    out[..., 1::2] = p >> 4
    return out[..., :n_cols]


def case_tree(z, name):
    return FlatTree(z[name + "/parent"], z[name + "/child_off"], z[name + "/child_idx"], z[name + "/leaf_row"])


def case_inputs(z, name):
    forced = z[name + "/forced"]
    return case_tree(z, name), z[name + "/msa"], z[name + "/consensus"], (forced if forced.size else None)


def c1_tree(c1):
    return FlatTree(c1["parent"], c1["child_off"], c1["child_idx"], c1["leaf_row"])


def same_muts(a, b):
    return (a.shape == b.shape and (a["node"] == b["node"]).all() and (a["pos"] == b["pos"]).all()
            and (a["type"] == b["type"]).all() and (a["code"] == b["code"]).all())


def same_nuc(a, b):
    return a.shape == b.shape and all((a[k] == b[k]).all() for k in a.dtype.names)
