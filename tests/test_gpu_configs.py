"""GPU (B200): BASELINE configs 3, 4 and 5 at their named shapes against oracle/_ref (the reference's own
src/fitchSankoff.cpp driven per column), every column-tile shape of the walk kernels, and the column-batched entry points.

Config 3 is the 8 M-taxon tree on a 2 048-column slice (the full 30 000 columns are the same slice fifteen times over;
bench.py runs them all), configs 4 and 5 run at full size.  The CUDA result is compared with the reference on a random
column sample (the CPU path needs ~25 ns per column.node on 16 cores) and, over ALL columns, through properties that do
not need the oracle: strict (node,pos) order, record-field invariants, and device-side <=6-run packing equal to the
oracle's packing of the same records."""
import os

import numpy as np
import pytest

import golden_util as G
import synth
from oracle import oracle_api as O
from panman_b200 import api, shard, trees

pytestmark = pytest.mark.gpu

SHAPES = [(4, 2), (2, 2), (4, 1), (1, 2), (2, 1), (1, 1)]


def _torch():
    import torch
    return torch


def _ref_threads(n_nodes):
    """Threads for the reference driver: every thread holds one column's node-name-keyed hash maps (~150 B per node)."""
    try:
        avail = int([l for l in open("/proc/meminfo") if l.startswith("MemAvailable")][0].split()[1]) * 1024
    except Exception:
        avail = 32 << 30
    by_mem = max(1, int(0.4 * avail / (n_nodes * 200 + 1)))
    return max(1, min(os.cpu_count() or 1, 32, by_mem))


def _synth(ctx, spec, n_rows, n_cols):
    torch = _torch()
    stride = ((n_cols + 31) // 32) * 16
    d_msa = torch.zeros((n_rows, stride), dtype=torch.uint8, device="cuda")
    d_cons = torch.zeros(n_cols, dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()
    ctx.synth_msa(spec, 0, n_cols, d_msa.data_ptr(), stride, d_cons.data_ptr())
    return d_msa, d_cons, stride


def _sample_inputs(d_msa, d_cons, cols):
    """Codes [rows, len(cols)] and consensus codes of the sampled columns, gathered on the device."""
    torch = _torch()
    byte_cols = torch.from_numpy(cols // 2).cuda()
    sub_bytes = d_msa[:, byte_cols].cpu().numpy()
    sub = np.where(cols % 2 == 0, sub_bytes & 15, sub_bytes >> 4).astype(np.uint8)
    return np.ascontiguousarray(sub), d_cons.cpu().numpy()[cols]


def _records_on_sample(ptr, n, cols):
    """The records of the sampled columns (selected on the device), positions renumbered to sample indices."""
    torch = _torch()
    t = shard.records_as_tensor(ptr, n, "cuda")
    sel = torch.isin(t[:, 1], torch.from_numpy(cols.astype(np.int32)).cuda())
    mine = shard.records_to_numpy(t[sel])
    mine["pos"] = np.searchsorted(cols, mine["pos"])
    return mine[np.lexsort((mine["pos"], mine["node"]))]


def _check_invariants_on_device(ptr, n, n_nodes, n_cols):
    torch = _torch()
    t = shard.records_as_tensor(ptr, n, "cuda")
    node, pos, pay = t[:, 0].to(torch.int64), t[:, 1].to(torch.int64), t[:, 2]
    key = (node << 32) | pos
    assert bool((key[1:] > key[:-1]).all()), "records must be strictly sorted by (node,pos)"
    assert int(pos.min()) >= 0 and int(pos.max()) < n_cols and int(node.max()) < n_nodes and int(node.min()) >= 0
    typ, code, pcode = pay & 255, (pay >> 8) & 255, (pay >> 16) & 255
    assert bool(((typ == 1) == (code == 0)).all())    # ND <=> the node's state is a gap
    assert bool(((typ == 2) == (pcode == 0)).all())   # NI <=> the parent's state is a gap
    assert bool((code != pcode).all()) and bool((typ <= 2).all())


def _check_against_ref(ctx, tree, algo, d_msa, d_cons, stride, n_rows, n_cols, n_sample, seed, check_pack=True):
    ptr, n = ctx.run_device(algo, d_msa.data_ptr(), stride, n_rows, n_cols, d_cons.data_ptr())
    assert n > 0
    _check_invariants_on_device(ptr, n, tree.n_nodes, n_cols)
    cols = np.sort(np.random.default_rng(seed).choice(n_cols, size=n_sample, replace=False))
    sub, cons = _sample_inputs(d_msa, d_cons, cols)
    want = O.run_ref_columns(algo, tree, O.codes_to_chars(sub), O.codes_to_chars(cons), None, threads=_ref_threads(tree.n_nodes))
    mine = _records_on_sample(ptr, n, cols)
    assert G.same_muts(mine, want), (algo, mine.shape, want.shape)
    if check_pack:  # device-side <=6-run packing of ALL records == the oracle's packing of the same records
        rec = ctx.download(n)
        off, nuc = ctx.pack_last()
        want_off, want_nuc = O.pack_port(rec, tree.n_nodes)
        assert (off == want_off).all() and G.same_nuc(nuc, want_nuc)
    return n


@pytest.mark.parametrize("algo", [api.FITCH, api.SANKOFF])
def test_config4_bacterial_shape_full_size(algo):
    """BASELINE config 4: 1 000 taxa x 5 000 000 columns, high gap fraction, 35 % accessory columns; 65 536 sampled
    columns against the reference."""
    N, L = 1000, 5_000_000
    tree = trees.random_binary_tree(N)
    spec = api.SynthSpec(p_sub=2e-3, p_del=5e-3, p_ins=5e-3, p_n_block=1e-3, p_amb=1e-4, accessory_frac=0.35)
    with api.Context(tree) as ctx:
        d_msa, d_cons, stride = _synth(ctx, spec, N, L)
        n = _check_against_ref(ctx, tree, algo, d_msa, d_cons, stride, N, L, 65536, seed=4)
        assert n > 50_000_000


def test_config4_every_record_identical_across_tile_shapes_and_runs():
    """All 71.6 M records of config 4 from the narrow-tile shapes, run several times each, against the 256-group shape:
    identical down to the last record.  (This comparison is what exposed a one-in-10^7 stale operand in the shapes whose
    walking warps outrun the producer warp: a TMA refill of a ring entry overtaking the walkers' reads of it, see
    walk_kernels.cuh reuse_fence.  A column sample against the oracle cannot see a defect that rare.)"""
    torch = _torch()
    N, L = 1000, 5_000_000
    tree = trees.random_binary_tree(N)
    spec = api.SynthSpec(p_sub=2e-3, p_del=5e-3, p_ins=5e-3, p_n_block=1e-3, p_amb=1e-4, accessory_frac=0.35)
    with api.Context(tree) as ctx:
        d_msa, d_cons, stride = _synth(ctx, spec, N, L)
        for algo in (api.FITCH, api.SANKOFF):
            ctx.set_tile(4, 2)
            want = shard.records_as_tensor(*ctx.run_device(algo, d_msa.data_ptr(), stride, N, L, d_cons.data_ptr()), "cuda").clone()
            for shape in ((4, 1), (2, 1), (1, 1), (2, 2)):
                ctx.set_tile(*shape)
                for rep in range(3):
                    got = shard.records_as_tensor(*ctx.run_device(algo, d_msa.data_ptr(), stride, N, L, d_cons.data_ptr()), "cuda")
                    assert got.shape == want.shape and bool((got == want).all()), (algo, shape, rep, got.shape, want.shape)


@pytest.mark.parametrize("algo,n_sample", [(api.FITCH, 4096), (api.SANKOFF, 4096)])
def test_config5_ladder_full_size(algo, n_sample):
    """BASELINE config 5: a 100 000-taxon caterpillar x 100 000 columns (serial depth 100 000: the reference's recursion
    needs a big stack, oracle/ref_driver.cpp runs it on one); 4 096 sampled columns against the reference."""
    N, L = 100_000, 100_000
    tree = trees.caterpillar_tree(N)
    spec = api.SynthSpec(p_sub=1e-4, p_del=2e-6, p_ins=1e-6, p_n_block=1e-2, p_amb=1e-4)
    with api.Context(tree) as ctx:
        d_msa, d_cons, stride = _synth(ctx, spec, N, L)
        n = _check_against_ref(ctx, tree, algo, d_msa, d_cons, stride, N, L, n_sample, seed=5)
        assert n > 50_000_000


def test_config3_eight_million_taxa_slice():
    """BASELINE config 3's tree (8 000 000 taxa, 16 M nodes) on a 2 048-column slice, both algorithms, against the
    reference on sampled columns, with the packing of all 165 M records and the same slice in two column batches
    (pmfs_run_batched).  One column of this tree costs the reference's node-name-keyed hash maps about a minute of one
    core (and ~3 GB), so oracle/_ref itself checks as many columns as there are host threads (at most 16) and the plain-C
    port -- pinned to oracle/_ref by tests/test_oracle.py, 25x faster -- the full sample of 64."""
    torch = _torch()
    N, L = 8_000_000, 2048
    tree = trees.random_binary_tree(N)
    spec = api.SynthSpec()
    with api.Context(tree) as ctx:
        d_msa, d_cons, stride = _synth(ctx, spec, N, L)
        for algo in (api.FITCH, api.SANKOFF):
            ptr, n = ctx.run_device(algo, d_msa.data_ptr(), stride, N, L, d_cons.data_ptr())
            assert n > 100_000_000
            _check_invariants_on_device(ptr, n, tree.n_nodes, L)
            cols = np.sort(np.random.default_rng(3).choice(L, size=64, replace=False))
            sub, cons = _sample_inputs(d_msa, d_cons, cols)
            mine = _records_on_sample(ptr, n, cols)
            assert G.same_muts(mine, O.run_port(algo, tree, sub, cons, None, threads=os.cpu_count() or 1)), algo
            k = min(16, _ref_threads(tree.n_nodes))  # one round of the reference's threads
            want = O.run_ref_columns(algo, tree, O.codes_to_chars(sub[:, :k]), O.codes_to_chars(cons[:k]), None, threads=k)
            assert G.same_muts(mine[mine["pos"] < k], want), algo
            if algo == api.FITCH:  # device-side <=6-run packing of ALL records == the oracle's packing of the same records
                rec = ctx.download(n)
                off, nuc = ctx.pack_last()
                want_off, want_nuc = O.pack_port(rec, tree.n_nodes)
                assert (off == want_off).all() and G.same_nuc(nuc, want_nuc)
                del rec, off, nuc, want_off, want_nuc
            # two batches of 1 024 columns: every batch is (node,pos)-ordered with global positions, and together they
            # hold exactly the records of the one-call run (compared through an order-independent checksum)
            t = shard.records_as_tensor(*ctx.run_device(algo, d_msa.data_ptr(), stride, N, L, d_cons.data_ptr()), "cuda")
            mix = torch.tensor([0x9E3779B1, 0x85EBCA77, 0xC2B2AE3D], dtype=torch.int64, device="cuda")
            want_sum = int(((t.to(torch.int64) * mix).sum(1) & 0xFFFFFFFF).sum())
            seen = []

            def consumer(col_begin, ncols, ptr, nrec, seen=seen):
                b = shard.records_as_tensor(ptr, nrec, "cuda")
                pos = b[:, 1]
                assert int(pos.min()) >= col_begin and int(pos.max()) < col_begin + ncols
                key = (b[:, 0].to(torch.int64) << 32) | pos.to(torch.int64)
                assert bool((key[1:] > key[:-1]).all())
                seen.append((col_begin, ncols, nrec, int(((b.to(torch.int64) * mix).sum(1) & 0xFFFFFFFF).sum())))

            total, nb = ctx.run_batched(algo, d_msa.data_ptr(), stride, N, L, d_cons.data_ptr(), batch_cols=1024, consumer=consumer)
            assert nb == 2 and total == n and [s[0] for s in seen] == [0, 1024]
            assert sum(s[3] for s in seen) == want_sum
            del t


@pytest.mark.parametrize("shape", SHAPES)
def test_every_tile_shape_is_bit_exact(shape, small_cases):
    """The walk kernels are built for six column-tile shapes (walking warps x column groups per thread); the shape must
    not change a single record: golden small cases, random polytomy trees with absent leaves and forced roots, a ladder,
    and a wide alignment with a ragged last group, both algorithms."""
    z = small_cases
    for name in list(z["case_names"])[::3]:
        tree, msa, cons, forced = G.case_inputs(z, str(name))
        for algo in (0, 1):
            with api.Context(tree) as ctx:
                ctx.set_tile(*shape)
                got = ctx.run(algo, api.pack_nibbles(msa), int(cons.shape[0]), cons, forced)
            assert G.same_muts(got, z[f"{name}/muts{algo}"]), (name, algo, shape)
    rng = np.random.default_rng(1000 + 10 * shape[0] + shape[1])
    cases = []
    for trial in range(6):
        tree = synth.random_tree(int(rng.integers(2, 300)), rng, max_arity=2 if trial < 2 else int(rng.integers(3, 7)),
                                 absent_frac=0.06 if trial % 2 else 0.0)
        cases.append((tree, int(rng.integers(1, 900))))
    cases.append((trees.caterpillar_tree(3000), 200))
    cases.append((trees.random_binary_tree(700, seed=8), 9000 + 13))
    for k, (tree, L) in enumerate(cases):
        msa = synth.simulate_msa(tree, L, rng, p_sub=0.03, p_del=0.02, p_ins=0.1, p_n=0.03, p_amb=0.02)
        cons = synth.consensus_first_nongap(msa)
        forced = msa[int(rng.integers(0, msa.shape[0]))] if k % 2 else None
        for algo in (0, 1):
            want = O.run_port(algo, tree, msa, cons, forced, threads=4)
            with api.Context(tree) as ctx:
                ctx.set_tile(*shape)
                ctx.set_tuning(int(rng.choice([0, 8, 40])), int(rng.choice([0, 2048])))
                got = ctx.run(algo, api.pack_nibbles(msa), L, cons, forced)
            assert G.same_muts(got, want), (k, algo, shape)


def test_host_call_streams_column_batches():
    """pmfs_run on an alignment that is sent through the GPU in column batches (forced here with pmfs_set_batch_cols; the
    automatic trigger is an alignment that does not fit on the device next to its set arrays): upload of batch b+1 under
    the passes of batch b, per-batch records stitched by node on the host -- same records as the one-pass call."""
    N, L = 1500, 21000 + 7
    tree = trees.random_binary_tree(N, seed=6)
    rng = np.random.default_rng(12)
    msa = synth.simulate_msa(tree, L, rng, p_sub=0.002, p_del=0.001, p_ins=0.004, p_n=0.004, p_amb=0.001)
    cons = synth.consensus_first_nongap(msa)
    nib = api.pack_nibbles(msa)
    for algo in (0, 1):
        for forced in (None, msa[7]):
            with api.Context(tree) as ctx:
                want = ctx.run(algo, nib, L, cons, forced)
                for batch in (4096, 5000, 20000):
                    ctx.set_batch_cols(batch)
                    got = ctx.run(algo, nib, L, cons, forced)
                    assert G.same_muts(got, want) and (got["parent_code"] == want["parent_code"]).all(), (algo, batch)
                ctx.set_batch_cols(0)
                again = ctx.run(algo, nib, L, cons, forced)
                assert G.same_muts(again, want)
    want_o = O.run_port(0, tree, msa, cons, None, threads=8)
    with api.Context(tree) as ctx:
        ctx.set_batch_cols(3000)
        assert G.same_muts(ctx.run(0, nib, L, cons), want_o)


def test_batched_run_stitches_to_the_one_call_result():
    """pmfs_run_batched with a consumer that keeps every batch: stitching the batches' per-node slices in batch order
    gives exactly the records of the one-call run (six batches, the last one ragged)."""
    tree = trees.random_binary_tree(40, seed=2)
    rng = np.random.default_rng(3)
    L = 6000
    msa = synth.simulate_msa(tree, L, rng, p_sub=0.05)
    cons = synth.consensus_first_nongap(msa)
    nib = api.pack_nibbles(msa, row_stride=((L + 31) // 32) * 16)
    torch = _torch()
    d_msa, d_cons = torch.from_numpy(nib).cuda(), torch.from_numpy(cons).cuda()
    with api.Context(tree) as ctx:
        ptr, n = ctx.run_device(0, d_msa.data_ptr(), nib.shape[1], 40, L, d_cons.data_ptr())
        want = ctx.download(n)
        parts = []
        total, nb = ctx.run_batched(0, d_msa.data_ptr(), nib.shape[1], 40, L, d_cons.data_ptr(), batch_cols=1024,
                                    consumer=lambda c0, nc, p, k: parts.append(shard.records_to_numpy(shard.records_as_tensor(p, k, "cuda"))))
        assert nb == 6 and total == n
        assert G.same_muts(shard.stitch_by_node(parts, tree.n_nodes), want)
