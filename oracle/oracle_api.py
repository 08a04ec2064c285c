"""TEST INFRASTRUCTURE ONLY: ctypes bindings for the two CPU checkers.

* ``liboracle.so``      -- oracle/pmfs_oracle.c, the plain-C restatement ("port").
* ``_ref/libpmref.so``  -- the reference's own src/fitchSankoff.cpp compiled in place + oracle/ref_driver.cpp.

Only tests/, ``__graft_entry__.smoke()`` and bench.py's ``cpu_baseline`` / ``--impl reference`` legs may import this
module.  Nothing under panman_b200/ does; the product has no CPU path.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from dataclasses import dataclass

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
PORT_SO = os.path.join(HERE, "liboracle.so")
REF_SO = os.path.join(HERE, "_ref", "libpmref.so")

CODE_CHARS = "-ACMGRSVTWYHKDBN"  # getNucleotideFromCode, /root/reference/src/panman.cpp:46-81

MUT_DTYPE = np.dtype([("node", "<i4"), ("pos", "<i4"), ("type", "u1"), ("code", "u1"), ("pad", "u1", (2,))])
NUCMUT_DTYPE = np.dtype(
    [("nucPosition", "<i4"), ("nucGapPosition", "<i4"), ("primaryBlockId", "<i4"), ("secondaryBlockId", "<i4"),
     ("mutInfo", "<u4"), ("nucs", "<u4")]
)


@dataclass
class FlatTree:
    """Tree in the reference's numbering (pre-order creation order of createTreeFromNewickString)."""

    parent: np.ndarray     # [V] int32, -1 for the root (id 0)
    child_off: np.ndarray  # [V+1] int32
    child_idx: np.ndarray  # [E] int32, children in Newick order
    leaf_row: np.ndarray   # [V] int32: MSA row, -1 internal, -2 leaf absent from the MSA

    @property
    def n_nodes(self) -> int:
        return int(self.parent.shape[0])

    def is_leaf(self) -> np.ndarray:
        return self.child_off[1:] == self.child_off[:-1]


def build(force: bool = False) -> None:
    """Compile the checkers (make -C oracle).  _ref is only (re)built where /root/reference exists."""
    if force or not os.path.exists(PORT_SO) or (os.path.isdir("/root/reference") and not os.path.exists(REF_SO)):
        subprocess.run(["make", "-C", HERE], check=True, stdout=subprocess.DEVNULL)


def _i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)


def _ptr(a, t):
    return a.ctypes.data_as(C.POINTER(t))


_port = None
_ref = None


def port_lib():
    global _port
    if _port is None:
        build()
        _port = C.CDLL(PORT_SO)
        _port.pmo_free.argtypes = [C.c_void_p]
    return _port


def ref_available() -> bool:
    return os.path.exists(REF_SO) or os.path.isdir("/root/reference")


def ref_lib():
    global _ref
    if _ref is None:
        build()
        _ref = C.CDLL(REF_SO)
        _ref.pmref_free.argtypes = [C.c_void_p]
    return _ref


def _take(lib_free, ptr, n, dtype):
    if n == 0:
        out = np.zeros(0, dtype=dtype)
    else:
        buf = (C.c_char * (n * dtype.itemsize)).from_address(ptr.value if hasattr(ptr, "value") else ptr)
        out = np.frombuffer(buf, dtype=dtype).copy()
    lib_free(ptr)
    return out


def run_port(algo: int, tree: FlatTree, msa_codes: np.ndarray, consensus: np.ndarray, forced=None, threads: int = 1):
    """oracle/pmfs_oracle.c column loop.  msa_codes [n_rows, n_cols] uint8 codes.  Returns MUT_DTYPE array or raises."""
    lib = port_lib()
    msa = np.ascontiguousarray(msa_codes, dtype=np.uint8)
    n_cols = int(consensus.shape[0])
    assert msa.ndim == 2 and (msa.shape[1] == n_cols or msa.shape[0] == 0)
    cons = np.ascontiguousarray(consensus, dtype=np.uint8)
    frc = None if forced is None else np.ascontiguousarray(forced, dtype=np.uint8)
    par, coff, cidx, lrow = _i32(tree.parent), _i32(tree.child_off), _i32(tree.child_idx), _i32(tree.leaf_row)
    out = C.c_void_p()
    n = C.c_int64()
    rc = lib.pmo_run(
        C.c_int(algo), C.c_int32(tree.n_nodes), _ptr(par, C.c_int32), _ptr(coff, C.c_int32), _ptr(cidx, C.c_int32),
        _ptr(lrow, C.c_int32), _ptr(msa, C.c_uint8), C.c_int64(n_cols), _ptr(cons, C.c_uint8),
        None if frc is None else _ptr(frc, C.c_uint8), C.c_int(threads), C.byref(out), C.byref(n))
    if rc != 0:
        raise RuntimeError(f"pmo_run rc={rc}")
    return _take(lib.pmo_free, out, n.value, MUT_DTYPE)


class _SynthSpec(C.Structure):
    _fields_ = [("seed", C.c_uint64), ("p_sub", C.c_float), ("p_del", C.c_float), ("p_ins", C.c_float),
                ("p_n_block", C.c_float), ("p_amb", C.c_float), ("accessory_frac", C.c_float)]


def synth_columns(tree: FlatTree, spec, cols: np.ndarray, n_rows: int, threads: int = 1) -> np.ndarray:
    """CPU twin of the product's device-side synthetic alignment generator (oracle/synth_twin.c): codes
    [n_rows, len(cols)] of the given GLOBAL columns.  spec: any object with the pmfs_synth_spec fields."""
    lib = port_lib()
    s = _SynthSpec(spec.seed, spec.p_sub, spec.p_del, spec.p_ins, spec.p_n_block, spec.p_amb, spec.accessory_frac)
    c = np.ascontiguousarray(cols, dtype=np.int64)
    out = np.zeros((n_rows, c.shape[0]), np.uint8)
    par, lrow = _i32(tree.parent), _i32(tree.leaf_row)
    rc = lib.pmo_synth(C.c_int32(tree.n_nodes), _ptr(par, C.c_int32), _ptr(lrow, C.c_int32), C.c_int32(n_rows), C.byref(s),
                       _ptr(c, C.c_int64), C.c_int64(c.shape[0]), C.c_int(threads), _ptr(out, C.c_uint8))
    if rc != 0:
        raise RuntimeError(f"pmo_synth rc={rc}")
    return out


def last_ref_loop_seconds() -> float:
    """Wall time of the column loop of the last run_ref_columns call (graph construction excluded)."""
    lib = ref_lib()
    lib.pmref_last_loop_seconds.restype = C.c_double
    return float(lib.pmref_last_loop_seconds())


def codes_to_chars(codes: np.ndarray) -> np.ndarray:
    lut = np.frombuffer(CODE_CHARS.encode(), dtype=np.uint8)
    return lut[np.asarray(codes, dtype=np.uint8)]


def chars_to_codes(chars: np.ndarray) -> np.ndarray:
    lut = np.zeros(256, dtype=np.uint8)
    for i, ch in enumerate(CODE_CHARS):
        if i:
            lut[ord(ch)] = i
    return lut[np.asarray(chars, dtype=np.uint8)]


def run_ref_columns(algo: int, tree: FlatTree, msa_chars: np.ndarray, consensus_chars: np.ndarray, forced_chars=None,
                    threads: int = 1):
    """The reference's own nucFitch*/nucSankoff* members driven per column (oracle/ref_driver.cpp: pmref_columns)."""
    lib = ref_lib()
    msa = np.ascontiguousarray(msa_chars, dtype=np.uint8)
    n_cols = int(consensus_chars.shape[0])
    cons = np.ascontiguousarray(consensus_chars, dtype=np.uint8)
    frc = None if forced_chars is None else np.ascontiguousarray(forced_chars, dtype=np.uint8)
    par, coff, cidx, lrow = _i32(tree.parent), _i32(tree.child_off), _i32(tree.child_idx), _i32(tree.leaf_row)
    out = C.c_void_p()
    n = C.c_int64()
    rc = lib.pmref_columns(
        C.c_int(algo), C.c_int32(tree.n_nodes), _ptr(par, C.c_int32), _ptr(coff, C.c_int32), _ptr(cidx, C.c_int32),
        _ptr(lrow, C.c_int32), C.c_int32(msa.shape[0]), C.c_int64(n_cols), _ptr(msa, C.c_char), _ptr(cons, C.c_char),
        None if frc is None else _ptr(frc, C.c_char), C.c_int(threads), C.byref(out), C.byref(n))
    if rc != 0:
        raise RuntimeError(f"pmref_columns rc={rc}")
    return _take(lib.pmref_free, out, n.value, MUT_DTYPE)


class RefGraph:
    """The reference's Node graph for a flat tree, built once (oracle/ref_driver.cpp: pmref_graph_new) so that repeated
    column samples on a big tree do not pay the graph construction every time."""

    def __init__(self, tree: FlatTree):
        self._lib = ref_lib()
        self._lib.pmref_graph_new.restype = C.c_void_p
        self._lib.pmref_graph_free.argtypes = [C.c_void_p]
        par, coff, cidx, lrow = _i32(tree.parent), _i32(tree.child_off), _i32(tree.child_idx), _i32(tree.leaf_row)
        self._h = C.c_void_p(self._lib.pmref_graph_new(C.c_int32(tree.n_nodes), _ptr(par, C.c_int32), _ptr(coff, C.c_int32),
                                                       _ptr(cidx, C.c_int32), _ptr(lrow, C.c_int32)))

    def columns(self, algo: int, msa_chars: np.ndarray, consensus_chars: np.ndarray, forced_chars=None, threads: int = 1):
        msa = np.ascontiguousarray(msa_chars, dtype=np.uint8)
        cons = np.ascontiguousarray(consensus_chars, dtype=np.uint8)
        frc = None if forced_chars is None else np.ascontiguousarray(forced_chars, dtype=np.uint8)
        out, n = C.c_void_p(), C.c_int64()
        rc = self._lib.pmref_columns_g(self._h, C.c_int(algo), C.c_int32(msa.shape[0]), C.c_int64(cons.shape[0]), _ptr(msa, C.c_char),
                                       _ptr(cons, C.c_char), None if frc is None else _ptr(frc, C.c_char), C.c_int(threads),
                                       C.byref(out), C.byref(n))
        if rc != 0:
            raise RuntimeError(f"pmref_columns_g rc={rc}")
        return _take(self._lib.pmref_free, out, n.value, MUT_DTYPE)

    def close(self):
        if self._h:
            self._lib.pmref_graph_free(self._h)
            self._h = None


def run_ref_blocks(algo: int, tree: FlatTree, states: np.ndarray, forced=None):
    """The reference's block-level members (blockFitch*New / blockSankoff*) driven per block as the PanGraph branch does
    (oracle/ref_driver.cpp: pmref_block_columns).  states: [rows, blocks] uint8, 0 absent / 1 forward / 2 reverse;
    forced: [blocks] uint8 or None.  Records: node, pos = block, type = BlockMutationType (BI 1, BD 0), code = inverted."""
    lib = ref_lib()
    st = np.ascontiguousarray(states, dtype=np.uint8)
    frc = None if forced is None else np.ascontiguousarray(forced, dtype=np.uint8)
    par, coff, cidx, lrow = _i32(tree.parent), _i32(tree.child_off), _i32(tree.child_idx), _i32(tree.leaf_row)
    out = C.c_void_p()
    n = C.c_int64()
    rc = lib.pmref_block_columns(
        C.c_int(algo), C.c_int32(tree.n_nodes), _ptr(par, C.c_int32), _ptr(coff, C.c_int32), _ptr(cidx, C.c_int32),
        _ptr(lrow, C.c_int32), C.c_int32(st.shape[0]), C.c_int64(st.shape[1]), _ptr(st, C.c_uint8),
        None if frc is None else _ptr(frc, C.c_uint8), C.byref(out), C.byref(n))
    if rc != 0:
        raise RuntimeError(f"pmref_block_columns rc={rc}")
    return _take(lib.pmref_free, out, n.value, MUT_DTYPE)


def blocks_from_nuc(rec: np.ndarray) -> np.ndarray:
    """Block view of nucleotide-pass records computed on state codes 0/1/2 (include/pmfs.h: pmfs_block_type /
    pmfs_block_inverted): NI -> (BI, code == 2), ND -> (BD, false), NS -> (BD, true)."""
    out = rec.copy()
    out["type"] = np.where(rec["type"] == 2, 1, 0)
    out["code"] = np.where(rec["type"] == 2, rec["code"] == 2, rec["type"] == 0)
    return out


def ref_build(msa_path: str, nwk_path: str, reference: str = "", low_mem: bool = False, threads: int = 1) -> dict:
    """Whole `panmanUtils -M .. -N ..` construction path up to serialisation (oracle/ref_driver.cpp: pmref_build)."""
    lib = ref_lib()
    n_nodes = C.c_int32()
    parent, coff, cidx, names, cons = C.c_void_p(), C.c_void_p(), C.c_void_p(), C.c_void_p(), C.c_void_p()
    muts, nuc_off, nuc, words = C.c_void_p(), C.c_void_p(), C.c_void_p(), C.c_void_p()
    n_cols, n_muts, n_nuc, n_words = C.c_int64(), C.c_int64(), C.c_int64(), C.c_int64()
    rc = lib.pmref_build(
        msa_path.encode(), nwk_path.encode(), reference.encode(), C.c_int(int(low_mem)), C.c_int(threads),
        C.byref(n_nodes), C.byref(parent), C.byref(coff), C.byref(cidx), C.byref(names), C.byref(n_cols), C.byref(cons),
        C.byref(muts), C.byref(n_muts), C.byref(nuc_off), C.byref(nuc), C.byref(n_nuc), C.byref(words), C.byref(n_words))
    if rc != 0:
        raise RuntimeError(f"pmref_build rc={rc}")
    V = n_nodes.value
    i32 = np.dtype("<i4")
    res = {}
    res["parent"] = _take(lib.pmref_free, parent, V, i32)
    res["child_off"] = _take(lib.pmref_free, coff, V + 1, i32)
    res["child_idx"] = _take(lib.pmref_free, cidx, int(res["child_off"][-1]), i32)
    res["names"] = C.string_at(names.value).decode().split("\n")[:-1]
    lib.pmref_free(names)
    res["consensus"] = C.string_at(cons.value, n_cols.value)
    lib.pmref_free(cons)
    res["muts"] = _take(lib.pmref_free, muts, n_muts.value, MUT_DTYPE)
    res["nuc_off"] = _take(lib.pmref_free, nuc_off, V + 1, np.dtype("<i8"))
    res["nuc"] = _take(lib.pmref_free, nuc, n_nuc.value, NUCMUT_DTYPE)
    res["block_words"] = _take(lib.pmref_free, words, n_words.value, np.dtype("<u4"))
    return res


def as_mut(muts: np.ndarray) -> np.ndarray:
    """Any structured array with node/pos/type/code fields -> MUT_DTYPE (e.g. the product's 12-byte pmfs_mut records)."""
    if muts.dtype == MUT_DTYPE:
        return np.ascontiguousarray(muts)
    out = np.zeros(muts.shape[0], dtype=MUT_DTYPE)
    for k in ("node", "pos", "type", "code"):
        out[k] = muts[k]
    return out


def _pack(lib, fn, free, muts: np.ndarray, n_nodes: int):
    m = as_mut(muts)
    off, out, n = C.c_void_p(), C.c_void_p(), C.c_int64()
    rc = fn(C.c_void_p(m.ctypes.data), C.c_int64(m.shape[0]), C.c_int32(n_nodes), C.byref(off), C.byref(out), C.byref(n))
    assert rc == 0
    return _take(free, off, n_nodes + 1, np.dtype("<i8")), _take(free, out, n.value, NUCMUT_DTYPE)


def pack_port(muts: np.ndarray, n_nodes: int):
    lib = port_lib()
    return _pack(lib, lib.pmo_pack, lib.pmo_free, muts, n_nodes)


def pack_ref(muts: np.ndarray, n_nodes: int):
    lib = ref_lib()
    return _pack(lib, lib.pmref_pack, lib.pmref_free, muts, n_nodes)
