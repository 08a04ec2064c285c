"""ctypes bindings for libpmfs.so (include/pmfs.h).  Mirrors the C ABI one to one; no compute happens in Python."""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from dataclasses import dataclass

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
FITCH, SANKOFF = 0, 1

MUT_DTYPE = np.dtype([("node", "<i4"), ("pos", "<i4"), ("type", "u1"), ("code", "u1"), ("parent_code", "u1"), ("pad", "u1")])
NUCMUT_DTYPE = np.dtype(
    [("nucPosition", "<i4"), ("nucGapPosition", "<i4"), ("primaryBlockId", "<i4"), ("secondaryBlockId", "<i4"),
     ("mutInfo", "<u4"), ("nucs", "<u4")]
)


class PmfsError(RuntimeError):
    def __init__(self, status: int, msg: str):
        super().__init__(f"pmfs status {status}: {msg}")
        self.status = status


def library_path() -> str:
    # PMFS_LIB: another build of the library (profiling / debugging variants, e.g. -DPMFS_WAITPROF)
    return os.environ.get("PMFS_LIB") or os.path.join(HERE, "libpmfs.so")


def build_library(force: bool = False) -> str:
    """Compile libpmfs.so for sm_100a (nvcc cross-compiles without a GPU)."""
    # make is incremental: a no-op when the library is newer than its sources (the GPU box runs it on the prebuilt file)
    subprocess.run(["make", "-C", os.path.join(HERE, "csrc")] + (["-B"] if force else []), check=True,
                   stdout=subprocess.DEVNULL)
    return library_path()


class _Tree(C.Structure):
    _fields_ = [("n_nodes", C.c_int32), ("parent", C.POINTER(C.c_int32)), ("child_off", C.POINTER(C.c_int32)),
                ("child_idx", C.POINTER(C.c_int32)), ("leaf_row", C.POINTER(C.c_int32))]


class _Timing(C.Structure):
    _fields_ = [("h2d_ms", C.c_float), ("forward_ms", C.c_float), ("backward_ms", C.c_float), ("sort_ms", C.c_float),
                ("d2h_ms", C.c_float), ("total_ms", C.c_float), ("n_launches", C.c_int64), ("forward_bytes", C.c_int64),
                ("backward_bytes", C.c_int64), ("wall_ms", C.c_float), ("exchange_ms", C.c_float), ("stitch_ms", C.c_float)]


class _PlanInfo(C.Structure):
    _fields_ = [("n_nodes", C.c_int32), ("n_internal", C.c_int32), ("n_leaves", C.c_int32), ("n_chunks", C.c_int32),
                ("n_fwd_levels", C.c_int32), ("n_bwd_levels", C.c_int32), ("fwd_stack_depth", C.c_int32),
                ("bwd_stack_depth", C.c_int32), ("max_arity", C.c_int32), ("program_words", C.c_int64),
                ("edges_leaf", C.c_int64), ("edges_internal", C.c_int64), ("n_parts", C.c_int32), ("n_set_rows", C.c_int32)]


class _Synth(C.Structure):
    _fields_ = [("seed", C.c_uint64), ("p_sub", C.c_float), ("p_del", C.c_float), ("p_ins", C.c_float),
                ("p_n_block", C.c_float), ("p_amb", C.c_float), ("accessory_frac", C.c_float)]


@dataclass
class SynthSpec:
    seed: int = 20251106
    p_sub: float = 2e-5
    p_del: float = 2e-6
    p_ins: float = 1e-6
    p_n_block: float = 1e-2
    p_amb: float = 1e-4
    accessory_frac: float = 0.0


BATCH_FN = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_int64, C.c_int64, C.c_void_p, C.c_int64)  # pmfs_batch_fn

_lib = None

EXPORTS = ["pmfs_last_error", "pmfs_version", "pmfs_create", "pmfs_destroy", "pmfs_set_tuning", "pmfs_set_tile", "pmfs_set_poly_split", "pmfs_set_watchdog", "pmfs_set_stream", "pmfs_run", "pmfs_run_chars", "pmfs_run_char_rows", "pmfs_run_codes",
           "pmfs_result_free", "pmfs_run_device", "pmfs_run_batched", "pmfs_set_batch_cols", "pmfs_run_fitch_static", "pmfs_set_shard", "pmfs_get_parts", "pmfs_sort_records", "pmfs_comm_unique_id", "pmfs_comm_create", "pmfs_comm_create_local", "pmfs_comm_destroy", "pmfs_comm_node_range", "pmfs_comm_batch_cols", "pmfs_comm_transport", "pmfs_comm_run", "pmfs_multi_create", "pmfs_multi_destroy", "pmfs_multi_run", "pmfs_pack_last", "pmfs_download", "pmfs_get_timing", "pmfs_get_plan_info",
           "pmfs_synth_msa"]


def lib():
    """Loads libpmfs.so; raises if it has not been built -- there is no fallback of any kind."""
    global _lib
    if _lib is None:
        path = library_path()
        if not os.path.exists(path):
            raise ImportError(f"{path} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                              "(make -C panman_b200/csrc).  panman_b200 has no CPU implementation.")
        L = C.CDLL(path)
        L.pmfs_last_error.restype = C.c_char_p
        L.pmfs_version.restype = C.c_char_p
        L.pmfs_create.argtypes = [C.POINTER(C.c_void_p), C.POINTER(_Tree), C.c_int]
        L.pmfs_destroy.argtypes = [C.c_void_p]
        L.pmfs_destroy.restype = None
        L.pmfs_set_tuning.argtypes = [C.c_void_p, C.c_int32, C.c_int64]
        L.pmfs_set_tile.argtypes = [C.c_void_p, C.c_int32, C.c_int32]
        L.pmfs_set_poly_split.argtypes = [C.c_void_p, C.c_int32, C.c_int32]
        L.pmfs_set_watchdog.argtypes = [C.c_void_p, C.c_double]
        L.pmfs_set_stream.argtypes = [C.c_void_p, C.c_void_p]
        L.pmfs_run.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int64, C.c_int32, C.c_int64, C.c_void_p, C.c_void_p,
                               C.POINTER(C.c_void_p), C.POINTER(C.c_int64)]
        L.pmfs_run_chars.argtypes = L.pmfs_run.argtypes
        L.pmfs_run_codes.argtypes = L.pmfs_run.argtypes
        L.pmfs_result_free.argtypes = [C.c_void_p, C.c_void_p]
        L.pmfs_result_free.restype = None
        L.pmfs_run_device.argtypes = L.pmfs_run.argtypes
        L.pmfs_run_batched.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int64, C.c_int32, C.c_int64, C.c_void_p, C.c_void_p,
                                       C.c_int64, C.c_void_p, C.c_void_p, C.POINTER(C.c_int64), C.POINTER(C.c_int32)]
        L.pmfs_set_batch_cols.argtypes = [C.c_void_p, C.c_int64]
        L.pmfs_run_fitch_static.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_int64, C.c_void_p, C.c_void_p,
                                            C.c_void_p, C.c_int64, C.POINTER(C.c_int64)]
        L.pmfs_set_shard.argtypes = [C.c_void_p, C.c_int64, C.c_int64, C.c_int32, C.c_void_p]
        L.pmfs_get_parts.argtypes = [C.c_void_p, C.POINTER(C.c_void_p), C.c_void_p, C.c_void_p, C.POINTER(C.c_int)]
        L.pmfs_sort_records.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.POINTER(C.c_void_p)]
        L.pmfs_comm_unique_id.argtypes = [C.c_void_p]
        L.pmfs_comm_create.argtypes = [C.POINTER(C.c_void_p), C.c_void_p, C.c_void_p, C.c_int, C.c_int]
        L.pmfs_comm_create_local.argtypes = [C.POINTER(C.c_void_p), C.POINTER(C.c_void_p), C.c_int]
        L.pmfs_comm_destroy.argtypes = [C.c_void_p]
        L.pmfs_comm_destroy.restype = None
        L.pmfs_comm_node_range.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_int32), C.POINTER(C.c_int32)]
        L.pmfs_comm_batch_cols.argtypes = [C.c_void_p, C.c_int, C.c_int64, C.c_int64, C.POINTER(C.c_int64)]
        L.pmfs_comm_transport.argtypes = [C.c_void_p]
        L.pmfs_comm_run.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int64, C.c_int32, C.c_int64, C.c_int64, C.c_int64,
                                    C.c_void_p, C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_int64)]
        L.pmfs_multi_create.argtypes = [C.POINTER(C.c_void_p), C.POINTER(_Tree), C.c_int, C.c_void_p]
        L.pmfs_multi_destroy.argtypes = [C.c_void_p]
        L.pmfs_multi_destroy.restype = None
        L.pmfs_multi_run.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int64, C.c_int32, C.c_int64, C.c_void_p, C.c_void_p,
                                     C.c_int64, C.POINTER(C.c_void_p), C.POINTER(C.c_int64)]
        L.pmfs_pack_last.argtypes = [C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_int64), C.POINTER(C.c_void_p)]
        L.pmfs_download.argtypes = [C.c_void_p, C.c_void_p, C.c_int64]
        L.pmfs_get_timing.argtypes = [C.c_void_p, C.POINTER(_Timing)]
        L.pmfs_get_plan_info.argtypes = [C.c_void_p, C.POINTER(_PlanInfo)]
        L.pmfs_synth_msa.argtypes = [C.c_void_p, C.POINTER(_Synth), C.c_int64, C.c_int64, C.c_void_p, C.c_int64, C.c_void_p]
        _lib = L
    return _lib


def _check(rc: int):
    if rc != 0:
        raise PmfsError(rc, lib().pmfs_last_error().decode())


def pack_nibbles(codes: np.ndarray, row_stride: int | None = None) -> np.ndarray:
    """[rows, cols] codes 0..15 -> [rows, row_stride] bytes, column 2j in the low nibble of byte j."""
    c = np.asarray(codes, dtype=np.uint8)
    rows, cols = c.shape
    nb = (cols + 1) // 2
    stride = nb if row_stride is None else row_stride
    assert stride >= nb
    out = np.zeros((rows, stride), dtype=np.uint8)
    out[:, : cols // 2] = c[:, 0 : cols - (cols % 2) : 2] | (c[:, 1::2] << 4)
    if cols % 2:
        out[:, nb - 1] = c[:, cols - 1]
    return out


def unpack_nibbles(packed: np.ndarray, n_cols: int) -> np.ndarray:
    p = np.asarray(packed, dtype=np.uint8)
    out = np.empty((p.shape[0], p.shape[1] * 2), np.uint8)
    out[:, 0::2] = p & 15
    out[:, 1::2] = p >> 4
    return out[:, :n_cols]


def _i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)


class Context:
    """One tree on one GPU (pmfs_ctx).  `tree` needs .parent/.child_off/.child_idx/.leaf_row int32 arrays."""

    def __init__(self, tree, device: int = 0):
        self._L = lib()
        self._keep = [_i32(tree.parent), _i32(tree.child_off), _i32(tree.child_idx), _i32(tree.leaf_row)]
        t = _Tree(int(self._keep[0].shape[0]), *[a.ctypes.data_as(C.POINTER(C.c_int32)) for a in self._keep])
        self._h = C.c_void_p()
        _check(self._L.pmfs_create(C.byref(self._h), C.byref(t), device))
        self.n_nodes = t.n_nodes
        self.device = device

    def close(self):
        if getattr(self, "_h", None) is not None and self._h:
            self._L.pmfs_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def set_tuning(self, chunk_nodes: int = 0, tile_cols: int = 0):
        _check(self._L.pmfs_set_tuning(self._h, chunk_nodes, tile_cols))

    def set_tile(self, walking_warps: int = 0, groups_per_thread: int = 0):
        """Column-tile shape of the work items (pmfs_set_tile); 0 = automatic."""
        _check(self._L.pmfs_set_tile(self._h, walking_warps, groups_per_thread))

    def set_poly_split(self, min_arity: int = 128, part_children: int = 0):
        """Split nodes with more than `min_arity` children into parts of `part_children` (pmfs_set_poly_split); 0 = never / default."""
        _check(self._L.pmfs_set_poly_split(self._h, min_arity, part_children))

    def set_watchdog(self, seconds: float):
        _check(self._L.pmfs_set_watchdog(self._h, seconds))

    def set_stream(self, cuda_stream=None):
        """Run on the given cudaStream_t handle (e.g. torch.cuda.current_stream().cuda_stream); None = the context's own
        stream.  Handle 0 is CUDA's legacy default stream -- what torch's default stream is -- and is passed on as
        cudaStreamLegacy, NOT taken as "no stream": work queued by torch on its default stream and the library's work
        must stay ordered with each other."""
        if cuda_stream is None:
            h = None
        else:
            h = int(cuda_stream) or 1  # cudaStreamLegacy == (cudaStream_t)0x1
        _check(self._L.pmfs_set_stream(self._h, h))

    def run(self, algo: int, msa_nibbles: np.ndarray, n_cols: int, consensus: np.ndarray, forced=None,
            copy: bool = True) -> np.ndarray:
        """Host buffers in, host records out (pmfs_run).  msa_nibbles: [rows, stride] uint8.
        copy=False returns a view of the context's pinned result buffer, valid until the next run."""
        msa = np.ascontiguousarray(msa_nibbles, dtype=np.uint8)
        cons = np.ascontiguousarray(consensus, dtype=np.uint8)
        assert cons.shape[0] == n_cols
        frc = None if forced is None else np.ascontiguousarray(forced, dtype=np.uint8)
        out, n = C.c_void_p(), C.c_int64()
        stride = msa.shape[1] if msa.ndim == 2 and msa.shape[0] else (n_cols + 1) // 2
        _check(self._L.pmfs_run(self._h, algo, msa.ctypes.data, stride, msa.shape[0], n_cols, cons.ctypes.data,
                                None if frc is None else frc.ctypes.data, C.byref(out), C.byref(n)))
        if n.value:
            buf = (C.c_char * (n.value * MUT_DTYPE.itemsize)).from_address(out.value)
            res = np.frombuffer(buf, dtype=MUT_DTYPE)
            if copy:
                res = res.copy()
        else:
            res = np.zeros(0, MUT_DTYPE)
        if copy:
            self._L.pmfs_result_free(self._h, out)
        return res

    def run_codes(self, algo: int, codes: np.ndarray, consensus_codes: np.ndarray, forced_codes=None) -> np.ndarray:
        """pmfs_run_codes: [rows, cols] uint8 state codes (one byte per cell) in, host records out."""
        return self.run_chars(algo, codes, consensus_codes, forced_codes, _entry="pmfs_run_codes")

    def run_chars(self, algo: int, msa_chars: np.ndarray, consensus_chars: np.ndarray, forced_chars=None,
                  _entry: str = "pmfs_run_chars") -> np.ndarray:
        """pmfs_run_chars: [rows, cols] uint8 alignment CHARACTERS in, host records out (conversion on the device)."""
        msa = np.ascontiguousarray(msa_chars, dtype=np.uint8)
        cons = np.ascontiguousarray(consensus_chars, dtype=np.uint8)
        frc = None if forced_chars is None else np.ascontiguousarray(forced_chars, dtype=np.uint8)
        n_cols = int(cons.shape[0])
        out, n = C.c_void_p(), C.c_int64()
        stride = msa.shape[1] if msa.ndim == 2 and msa.shape[0] else n_cols
        _check(getattr(self._L, _entry)(self._h, algo, msa.ctypes.data, stride, msa.shape[0], n_cols, cons.ctypes.data,
                                      None if frc is None else frc.ctypes.data, C.byref(out), C.byref(n)))
        if not n.value:
            return np.zeros(0, MUT_DTYPE)
        return np.frombuffer((C.c_char * (n.value * MUT_DTYPE.itemsize)).from_address(out.value), dtype=MUT_DTYPE).copy()

    def run_device(self, algo: int, d_msa: int, row_stride: int, n_rows: int, n_cols: int, d_cons: int, d_forced: int = 0):
        """Device pointers (ints) in; records stay on the device.  Returns (device pointer, count)."""
        out, n = C.c_void_p(), C.c_int64()
        _check(self._L.pmfs_run_device(self._h, algo, d_msa, row_stride, n_rows, n_cols, d_cons, d_forced or None,
                                       C.byref(out), C.byref(n)))
        return out.value, n.value

    def set_batch_cols(self, batch_cols: int = 0):
        _check(self._L.pmfs_set_batch_cols(self._h, batch_cols))

    def run_batched(self, algo: int, d_msa: int, row_stride: int, n_rows: int, n_cols: int, d_cons: int, d_forced: int = 0,
                    batch_cols: int = 0, consumer=None):
        """pmfs_run_batched: the pass in column batches.  consumer(col_begin, n_cols, device pointer, n_records) is called
        after every batch (records on the device, (node,pos)-ordered, global positions; valid until it returns).
        Returns (total records, number of batches)."""
        total, nb = C.c_int64(), C.c_int32()
        cb = None
        if consumer is not None:
            def _cb(_user, col_begin, ncols, ptr, n):
                consumer(col_begin, ncols, ptr or 0, n)
                return 0
            cb = BATCH_FN(_cb)
        _check(self._L.pmfs_run_batched(self._h, algo, d_msa, row_stride, n_rows, n_cols, d_cons, d_forced or None, batch_cols,
                                        C.cast(cb, C.c_void_p) if cb is not None else None, None, C.byref(total), C.byref(nb)))
        return total.value, nb.value

    def run_fitch_static(self, d_msa: int, row_stride: int, n_rows: int, n_cols: int, d_cons: int, d_forced: int,
                         d_out: int, out_capacity: int) -> int:
        """pmfs_run_fitch_static: the Fitch pass without any allocation (scratch in the library's static device array).
        All arguments are device pointers / sizes; returns the number of records written to d_out."""
        n = C.c_int64()
        rc = self._L.pmfs_run_fitch_static(self._h, d_msa, row_stride, n_rows, n_cols, d_cons, d_forced or None, d_out,
                                           out_capacity, C.byref(n))
        if rc != 0:
            err = PmfsError(rc, self._L.pmfs_last_error().decode())
            err.n_out = n.value
            raise err
        return n.value

    def set_shard(self, pos_offset: int, total_cols: int, node_bounds=None):
        """Column-shard mode (pmfs_set_shard); node_bounds: [n_parts+1] node-id boundaries, None leaves shard mode."""
        if node_bounds is None:
            # total_cols > 0: a column shard that orders and keeps its own records (global positions, no exchange)
            _check(self._L.pmfs_set_shard(self._h, pos_offset if total_cols else 0, total_cols, 0, None))
            self._n_parts = 0
            return
        b = np.ascontiguousarray(node_bounds, dtype=np.int32)
        _check(self._L.pmfs_set_shard(self._h, pos_offset, total_cols, b.shape[0] - 1, b.ctypes.data))
        self._n_parts = b.shape[0] - 1

    def get_parts(self):
        """-> (device base pointer, offsets list, counts list, column_ordered) of the last shard-mode run: owner r's
        64-bit records are base[offsets[r] : offsets[r] + counts[r]]."""
        base, flag = C.c_void_p(), C.c_int()
        offsets = np.zeros(16, np.int64)
        counts = np.zeros(16, np.int64)
        _check(self._L.pmfs_get_parts(self._h, C.byref(base), offsets.ctypes.data, counts.ctypes.data, C.byref(flag)))
        return base.value or 0, offsets[: self._n_parts].tolist(), counts[: self._n_parts].tolist(), bool(flag.value)

    def sort_records(self, d_records: int, n: int, column_ordered: bool = False) -> int:
        """Order `n` received 64-bit records (blocks concatenated in source-rank order) by (node, pos) -> device
        pointer to pmfs_mut records."""
        out = C.c_void_p()
        _check(self._L.pmfs_sort_records(self._h, d_records or None, n, int(column_ordered), C.byref(out)))
        return out.value or 0

    def pack_last(self):
        """Device-side <=6-run packing of the last result -> (node_off int64[V+1], packed NUCMUT_DTYPE array)."""
        out, off, n = C.c_void_p(), C.c_void_p(), C.c_int64()
        _check(self._L.pmfs_pack_last(self._h, C.byref(out), C.byref(n), C.byref(off)))
        node_off = np.frombuffer((C.c_char * ((self.n_nodes + 1) * 8)).from_address(off.value), dtype="<i8").copy()
        if n.value:
            nuc = np.frombuffer((C.c_char * (n.value * NUCMUT_DTYPE.itemsize)).from_address(out.value), dtype=NUCMUT_DTYPE).copy()
        else:
            nuc = np.zeros(0, NUCMUT_DTYPE)
        return node_off, nuc

    def download(self, n: int) -> np.ndarray:
        res = np.zeros(n, MUT_DTYPE)
        _check(self._L.pmfs_download(self._h, res.ctypes.data, n))
        return res

    def download_into(self, host_ptr: int, n: int):
        """pmfs_download into a caller buffer (e.g. pinned memory) of at least n records."""
        _check(self._L.pmfs_download(self._h, host_ptr, n))

    def timing(self) -> dict:
        t = _Timing()
        _check(self._L.pmfs_get_timing(self._h, C.byref(t)))
        return {k: getattr(t, k) for k, _ in _Timing._fields_}

    def plan_info(self) -> dict:
        p = _PlanInfo()
        _check(self._L.pmfs_get_plan_info(self._h, C.byref(p)))
        return {k: getattr(p, k) for k, _ in _PlanInfo._fields_}

    def synth_msa(self, spec: SynthSpec, col_begin: int, n_cols: int, d_msa: int, row_stride: int, d_cons: int = 0):
        s = _Synth(spec.seed, spec.p_sub, spec.p_del, spec.p_ins, spec.p_n_block, spec.p_amb, spec.accessory_frac)
        _check(self._L.pmfs_synth_msa(self._h, C.byref(s), col_begin, n_cols, d_msa, row_stride, d_cons or None))


def comm_unique_id() -> bytes:
    """pmfs_comm_unique_id: the 128 bytes rank 0 creates and every rank passes to Comm (ship them as you like)."""
    buf = C.create_string_buffer(128)
    _check(lib().pmfs_comm_unique_id(buf))
    return buf.raw


class Comm:
    """One rank of a group of contexts (one GPU each) whose per-node lists are gathered by node range (pmfs_comm_*)."""

    def __init__(self, ctx: Context, unique_id: bytes | None, rank: int, world: int, _handle=None):
        self._L = lib()
        self.ctx, self.rank, self.world = ctx, rank, world
        if _handle is not None:
            self._h = _handle
            return
        self._h = C.c_void_p()
        idbuf = C.create_string_buffer(unique_id, 128) if unique_id is not None else None
        _check(self._L.pmfs_comm_create(C.byref(self._h), ctx._h, idbuf, rank, world))

    @staticmethod
    def local_group(ctxs):
        """Communicators over contexts of THIS process that exchange by device copies (pmfs_comm_create_local); every
        rank must be driven by its own host thread."""
        n = len(ctxs)
        hs = (C.c_void_p * n)()
        cs = (C.c_void_p * n)(*[c._h for c in ctxs])
        _check(lib().pmfs_comm_create_local(hs, cs, n))
        return [Comm(ctxs[r], None, r, n, _handle=C.c_void_p(hs[r])) for r in range(n)]

    def close(self):
        if getattr(self, "_h", None):
            self._L.pmfs_comm_destroy(self._h)
            self._h = None

    def node_range(self, rank: int | None = None):
        b, e = C.c_int32(), C.c_int32()
        _check(self._L.pmfs_comm_node_range(self._h, self.rank if rank is None else rank, C.byref(b), C.byref(e)))
        return b.value, e.value

    @property
    def transport(self) -> str:
        """How records travel between the ranks ("nccl", "peer-copy" through CUDA IPC, "local" inside one process)."""
        return {0: "nccl", 1: "peer-copy", 2: "local"}.get(self._L.pmfs_comm_transport(self._h), "?")

    def batch_cols(self, algo: int, cols_per_rank: int, bytes_per_col: int = 0) -> int:
        out = C.c_int64()
        _check(self._L.pmfs_comm_batch_cols(self._h, algo, cols_per_rank, bytes_per_col, C.byref(out)))
        return out.value

    def run(self, algo: int, d_msa: int, row_stride: int, n_rows: int, n_cols: int, col_offset: int, total_cols: int,
            d_cons: int, d_forced: int = 0):
        """Collective.  -> (device pointer to this rank's node range of pmfs_mut records, count)."""
        out, n = C.c_void_p(), C.c_int64()
        _check(self._L.pmfs_comm_run(self._h, algo, d_msa or None, row_stride, n_rows, n_cols, col_offset, total_cols,
                                     d_cons or None, d_forced or None, C.byref(out), C.byref(n)))
        return out.value or 0, n.value


class Multi:
    """One process, several GPUs (pmfs_multi_*): host alignment in, one (node,pos)-ordered host array out."""

    def __init__(self, tree, n_gpus: int, devices=None):
        self._L = lib()
        self._keep = [_i32(tree.parent), _i32(tree.child_off), _i32(tree.child_idx), _i32(tree.leaf_row)]
        t = _Tree(int(self._keep[0].shape[0]), *[a.ctypes.data_as(C.POINTER(C.c_int32)) for a in self._keep])
        self._h = C.c_void_p()
        dev = None if devices is None else np.ascontiguousarray(devices, dtype=np.int32)
        _check(self._L.pmfs_multi_create(C.byref(self._h), C.byref(t), n_gpus, None if dev is None else dev.ctypes.data))

    def close(self):
        if getattr(self, "_h", None):
            self._L.pmfs_multi_destroy(self._h)
            self._h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def run(self, algo: int, msa_nibbles: np.ndarray, n_cols: int, consensus: np.ndarray, forced=None, window_cols: int = 0):
        msa = np.ascontiguousarray(msa_nibbles, dtype=np.uint8)
        cons = np.ascontiguousarray(consensus, dtype=np.uint8)
        frc = None if forced is None else np.ascontiguousarray(forced, dtype=np.uint8)
        out, n = C.c_void_p(), C.c_int64()
        _check(self._L.pmfs_multi_run(self._h, algo, msa.ctypes.data, msa.shape[1], msa.shape[0], n_cols, cons.ctypes.data,
                                      None if frc is None else frc.ctypes.data, window_cols, C.byref(out), C.byref(n)))
        if not n.value:
            return np.zeros(0, MUT_DTYPE)
        return np.frombuffer((C.c_char * (n.value * MUT_DTYPE.itemsize)).from_address(out.value), dtype=MUT_DTYPE).copy()
