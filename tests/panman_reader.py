"""TEST INFRASTRUCTURE: a small Cap'n Proto reader for .panman files (schema: /root/reference/panman.capnp), written
independently of the writer in panman_b200/csrc/host/panman_writer.cpp: it follows pointers generically (struct, list,
composite list, far pointers across segments) and only knows the schema's field positions, which it takes from the
Cap'n Proto layout rules (fields placed in ordinal order into the first hole that fits).  Decodes what
panmanUtils::Tree::Tree(panman::Tree::Reader) reads back (src/panman.cpp: the PANMAT constructor) for the MSA path."""
from __future__ import annotations

import lzma
import struct

import numpy as np


class Msg:
    def __init__(self, raw: bytes):
        (nseg_m1,) = struct.unpack_from("<I", raw, 0)
        nseg = nseg_m1 + 1
        sizes = struct.unpack_from("<%dI" % nseg, raw, 4)
        off = 4 + 4 * nseg
        off += (-off) % 8
        self.segs = []
        for s in sizes:
            self.segs.append(np.frombuffer(raw, dtype="<u8", count=s, offset=off))
            off += 8 * s
        assert off == len(raw), (off, len(raw))

    def word(self, seg, idx):
        return int(self.segs[seg][idx])

    def resolve(self, seg, idx):
        """Pointer at (seg, idx) -> (kind, seg, target word, a, b) with far pointers followed.  kind 0 struct (a = data
        words, b = pointers), 1 list (a = element size code, b = count), None null."""
        w = self.word(seg, idx)
        if w == 0:
            return None
        kind = w & 3
        if kind == 2:  # far pointer
            double, off, tseg = (w >> 2) & 1, (w >> 3) & 0x1FFFFFFF, w >> 32
            if not double:
                return self.resolve_landing(tseg, off)
            pad = self.word(tseg, off)
            tag = self.word(tseg, off + 1)
            target_seg, target = pad >> 32, (pad >> 3) & 0x1FFFFFFF
            return (tag & 3, target_seg, target, (tag >> 32) & 0xFFFF if tag & 3 == 0 else (tag >> 32) & 7,
                    (tag >> 48) if tag & 3 == 0 else tag >> 35)
        return self.decode(seg, idx, w)

    def resolve_landing(self, seg, idx):
        return self.decode(seg, idx, self.word(seg, idx))

    def decode(self, seg, idx, w):
        kind = w & 3
        off = (w >> 2) & 0x3FFFFFFF
        if off & 0x20000000:
            off -= 0x40000000
        target = idx + 1 + off
        if kind == 0:
            return (0, seg, target, (w >> 32) & 0xFFFF, w >> 48)
        assert kind == 1
        return (1, seg, target, (w >> 32) & 7, w >> 35)


class Struct:
    def __init__(self, msg, seg, at, data, ptrs):
        self.m, self.seg, self.at, self.data, self.ptrs = msg, seg, at, data, ptrs

    def u64(self, word):
        return self.m.word(self.seg, self.at + word) if word < self.data else 0

    def u32(self, idx):
        return (self.u64(idx // 2) >> (32 * (idx % 2))) & 0xFFFFFFFF

    def i32(self, idx):
        v = self.u32(idx)
        return v - (1 << 32) if v & 0x80000000 else v

    def bit(self, idx):
        return (self.u64(idx // 64) >> (idx % 64)) & 1

    def ptr(self, i):
        if i >= self.ptrs:
            return None
        return self.m.resolve(self.seg, self.at + self.data + i)

    def struct_list(self, i):
        r = self.ptr(i)
        if r is None:
            return []
        kind, seg, target, elem, count = r
        assert kind == 1 and elem == 7, r
        tag = self.m.word(seg, target)
        n, d, p = (tag >> 2) & 0x3FFFFFFF, (tag >> 32) & 0xFFFF, tag >> 48
        assert n * (d + p) == count
        return [Struct(self.m, seg, target + 1 + k * (d + p), d, p) for k in range(n)]

    def prim_list(self, i, dtype):
        r = self.ptr(i)
        if r is None:
            return np.zeros(0, dtype)
        kind, seg, target, elem, count = r
        assert kind == 1
        size = {2: 1, 3: 2, 4: 4, 5: 8}[elem]
        assert np.dtype(dtype).itemsize == size
        raw = self.m.segs[seg][target: target + (count * size + 7) // 8].tobytes()
        return np.frombuffer(raw, dtype=dtype, count=count).copy()

    def bool_list(self, i):
        r = self.ptr(i)
        if r is None:
            return []
        kind, seg, target, elem, count = r
        assert kind == 1 and elem == 1
        return [(self.m.word(seg, target + k // 64) >> (k % 64)) & 1 for k in range(count)]

    def text_list(self, i):
        r = self.ptr(i)
        if r is None:
            return []
        kind, seg, target, elem, count = r
        assert kind == 1 and elem == 6
        out = []
        for k in range(count):
            t = self.m.resolve(seg, target + k)
            kk, s2, t2, e2, c2 = t
            assert kk == 1 and e2 == 2
            raw = self.m.segs[s2][t2: t2 + (c2 + 7) // 8].tobytes()[:c2]
            assert raw[-1:] == b"\0"
            out.append(raw[:-1].decode())
        return out


def read_panman(path: str, compressed: bool = True) -> dict:
    """-> dict(newick, nodes=[{block mutations, nuc mutations}], consensus_words, n_trees, n_complex, ...) of tree 0."""
    raw = open(path, "rb").read()
    if compressed:
        raw = lzma.decompress(raw)
    m = Msg(raw)
    kind, seg, at, d, p = m.resolve(0, 0)
    assert kind == 0
    tg = Struct(m, seg, at, d, p)           # TreeGroup: trees @0, complexMutations @1
    trees = tg.struct_list(0)
    tree = trees[0]                          # Tree: newick 0, nodes 1, consensusSeqMap 2, gaps 3, blockGaps 4, ... 7
    nodes = []
    for n in tree.struct_list(1):            # Node: mutations 0, annotations 1
        muts = []
        for mu in n.struct_list(0):          # Mutation: blockId word 0; bits of word 1; nucMutation ptr 0
            nm = [(x.i32(0), x.i32(1), x.bit(64), x.u32(3)) for x in mu.struct_list(0)]  # NucMut: pos, gapPos, gapExist, mutInfo
            muts.append({"blockId": mu.u64(0), "blockGapExist": mu.bit(64), "blockMutExist": mu.bit(65), "blockMutInfo": mu.bit(66),
                         "blockInversion": mu.bit(67), "nuc": nm})
        nodes.append({"mutations": muts, "annotations": n.text_list(1)})
    cmap = []
    for c in tree.struct_list(2):            # ConsensusSeqToBlockIds: blockId 0, consensusSeq 1, blockGapExist 2, chromosomeName 3
        cmap.append({"blockId": c.prim_list(0, "<i8"), "consensusSeq": c.prim_list(1, "<u4"), "blockGapExist": c.bool_list(2)})
    return {"n_trees": len(trees), "n_complex": len(tg.struct_list(1)), "newick": "".join(tree.text_list(0)), "nodes": nodes,
            "consensusSeqMap": cmap, "n_gaps": len(tree.struct_list(3)), "blockGaps_null": tree.ptr(4) is None,
            "n_circular": len(tree.struct_list(5)), "n_rotation": len(tree.struct_list(6)), "n_inverted": len(tree.struct_list(7))}


def nuc_csr(decoded: dict):
    """Per-node NucMut lists back in the in-memory form (NucMut(panman::NucMut::Reader, blockId, blockGapExist),
    src/panman.hpp:160-182): -> (nuc_off int64[V+1], structured array nucPosition/nucGapPosition/primary/secondary/mutInfo/nucs)."""
    dt = np.dtype([("nucPosition", "<i4"), ("nucGapPosition", "<i4"), ("primaryBlockId", "<i4"), ("secondaryBlockId", "<i4"),
                   ("mutInfo", "<u4"), ("nucs", "<u4")])
    rows, off = [], [0]
    for n in decoded["nodes"][:-1]:  # (the writer emits allNodes.size() + 1 entries)
        for mu in n["mutations"]:
            primary = mu["blockId"] >> 32
            secondary = (mu["blockId"] & 0xFFFFFFFF) if mu["blockGapExist"] else -1
            for pos, gpos, gexist, info in mu["nuc"]:
                mi = info & 0xFF
                nucs = ((info >> 8) << (24 - (mi >> 4) * 4)) & 0xFFFFFFFF
                rows.append((pos, gpos if gexist else -1, primary, secondary, mi, nucs))
        off.append(len(rows))
    return np.array(off, np.int64), np.array(rows, dt) if rows else np.zeros(0, dt)
