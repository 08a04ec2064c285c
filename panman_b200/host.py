"""ctypes bindings for the host C++ mirror of the reference's MSA-driven Tree constructor (include/pmfs_host.h)."""
from __future__ import annotations

import ctypes as C

import numpy as np

from .api import MUT_DTYPE, NUCMUT_DTYPE, PmfsError, lib

HOST_EXPORTS = ["pmfs_host_last_error", "pmfs_host_free", "pmfs_host_parse_newick", "pmfs_host_build", "pmfs_host_pack",
                "pmfs_host_writer_last_error", "pmfs_host_write_panman", "pmfs_host_build_panman"]


def _L():
    L = lib()
    L.pmfs_host_last_error.restype = C.c_char_p
    L.pmfs_host_free.argtypes = [C.c_void_p]
    L.pmfs_host_free.restype = None
    return L


def _err(rc):
    raise PmfsError(rc, _L().pmfs_host_last_error().decode())


def _take(ptr, n, dtype):
    L = _L()
    if n:
        out = np.frombuffer((C.c_char * (n * dtype.itemsize)).from_address(ptr.value), dtype=dtype).copy()
    else:
        out = np.zeros(0, dtype)
    L.pmfs_host_free(ptr)
    return out


def parse_newick(newick: str):
    """-> (names list, parent int32[V], n_children int32[V]) in the reference's creation order."""
    L = _L()
    n = L.pmfs_host_parse_newick(newick.encode(), None, None, None, C.c_int64(0))
    if n <= 0:
        _err(n)
    parent = np.zeros(n, np.int32)
    nch = np.zeros(n, np.int32)
    cap = len(newick) + 16 * n + 16
    buf = C.create_string_buffer(cap)
    rc = L.pmfs_host_parse_newick(newick.encode(), C.c_void_p(parent.ctypes.data), C.c_void_p(nch.ctypes.data), buf, C.c_int64(cap))
    if rc <= 0:
        _err(rc)
    return buf.value.decode().split("\n")[:-1], parent, nch


def build(msa_path: str, nwk_path: str, reference: str = "", low_mem: bool = False, device: int = 0) -> dict:
    L = _L()
    V, n_cols, n_muts, n_nuc, n_words = C.c_int32(), C.c_int64(), C.c_int64(), C.c_int64(), C.c_int64()
    parent, names, cons, muts, off, nuc, words = (C.c_void_p() for _ in range(7))
    rc = L.pmfs_host_build(msa_path.encode(), nwk_path.encode(), reference.encode(), int(low_mem), device, C.byref(V),
                           C.byref(parent), C.byref(names), C.byref(n_cols), C.byref(cons), C.byref(muts), C.byref(n_muts),
                           C.byref(off), C.byref(nuc), C.byref(n_nuc), C.byref(words), C.byref(n_words))
    if rc != 0:
        _err(rc)
    res = {"parent": _take(parent, V.value, np.dtype("<i4"))}
    res["names"] = C.string_at(names.value).decode().split("\n")[:-1]
    L.pmfs_host_free(names)
    res["consensus"] = C.string_at(cons.value, n_cols.value)
    L.pmfs_host_free(cons)
    res["muts"] = _take(muts, n_muts.value, MUT_DTYPE)
    res["nuc_off"] = _take(off, V.value + 1, np.dtype("<i8"))
    res["nuc"] = _take(nuc, n_nuc.value, NUCMUT_DTYPE)
    res["block_words"] = _take(words, n_words.value, np.dtype("<u4"))
    return res


def pack(muts: np.ndarray, n_nodes: int):
    L = _L()
    m = np.ascontiguousarray(muts, dtype=MUT_DTYPE)
    off, nuc, n = C.c_void_p(), C.c_void_p(), C.c_int64()
    rc = L.pmfs_host_pack(C.c_void_p(m.ctypes.data), C.c_int64(m.shape[0]), C.c_int32(n_nodes), C.byref(off), C.byref(nuc), C.byref(n))
    if rc != 0:
        _err(rc)
    return _take(off, n_nodes + 1, np.dtype("<i8")), _take(nuc, n.value, NUCMUT_DTYPE)


def write_panman(parent, names, nuc_off, nuc, block_words, out_path: str, compress: bool = True, branch_length=None):
    """pmfs_host_write_panman: the .panman writer alone (no device): tree in pre-order + per-node packed NucMut CSR."""
    L = _L()
    L.pmfs_host_writer_last_error.restype = C.c_char_p
    par = np.ascontiguousarray(parent, dtype=np.int32)
    off = np.ascontiguousarray(nuc_off, dtype=np.int64)
    nm = np.ascontiguousarray(nuc, dtype=NUCMUT_DTYPE)
    bw = np.ascontiguousarray(block_words, dtype=np.uint32)
    bl = None if branch_length is None else np.ascontiguousarray(branch_length, dtype=np.float32)
    joined = "".join(n + "\n" for n in names).encode()
    rc = L.pmfs_host_write_panman(C.c_int32(par.shape[0]), C.c_void_p(par.ctypes.data), joined,
                                  None if bl is None else C.c_void_p(bl.ctypes.data), C.c_void_p(off.ctypes.data),
                                  C.c_void_p(nm.ctypes.data) if nm.shape[0] else None, C.c_void_p(bw.ctypes.data) if bw.shape[0] else None,
                                  C.c_int64(bw.shape[0]), out_path.encode(), C.c_int(int(compress)))
    if rc != 0:
        raise PmfsError(rc, L.pmfs_host_writer_last_error().decode())


def build_panman(msa_path: str, nwk_path: str, out_path: str, reference: str = "", low_mem: bool = False, device: int = 0,
                 compress: bool = True):
    """pmfs_host_build_panman: `panmanUtils -M msa -N nwk [--reference id] [--low-mem-mode] -o out` on a CUDA device."""
    L = _L()
    rc = L.pmfs_host_build_panman(msa_path.encode(), nwk_path.encode(), reference.encode(), int(low_mem), device,
                                  out_path.encode(), int(compress))
    if rc != 0:
        _err(rc)
