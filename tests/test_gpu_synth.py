"""GPU (B200): device-side synthetic alignments (pmfs_synth_msa) and full-size checks on BASELINE config 2."""
import numpy as np
import pytest

import golden_util as G
from oracle import oracle_api as O
from panman_b200 import api, trees

pytestmark = pytest.mark.gpu


def _torch():
    import torch
    return torch


def synth_on_device(ctx, spec, n_rows, col_begin, n_cols):
    torch = _torch()
    stride = ((n_cols + 31) // 32) * 16
    d_msa = torch.zeros((n_rows, stride), dtype=torch.uint8, device="cuda")
    d_cons = torch.zeros(n_cols, dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()
    ctx.synth_msa(spec, col_begin, n_cols, d_msa.data_ptr(), stride, d_cons.data_ptr())
    return d_msa, d_cons, stride


def test_synth_is_a_pure_function_of_seed_node_column():
    tree = trees.random_binary_tree(500, seed=2)
    spec = api.SynthSpec(seed=7, p_sub=0.01, p_del=0.01, p_ins=0.05, p_n_block=0.02, p_amb=0.01, accessory_frac=0.3)
    with api.Context(tree) as ctx:
        a, ca, _ = synth_on_device(ctx, spec, 500, 0, 4096)
        b, cb, _ = synth_on_device(ctx, spec, 500, 0, 4096)
        assert bool((a == b).all()) and bool((ca == cb).all())
        full = api.unpack_nibbles(a.cpu().numpy(), 4096)
        part, cp, _ = synth_on_device(ctx, spec, 500, 1024, 512)  # any column range generates the same cells
        assert (api.unpack_nibbles(part.cpu().numpy(), 512) == full[:, 1024:1536]).all()
        assert ((full != 0).sum(0) >= 1).all()  # every column has a non-gap cell (SURVEY §8 notes: all-gap is UB upstream)
        nz = full != 0
        want_cons = full[nz.argmax(0), np.arange(4096)]
        assert (ca.cpu().numpy() == want_cons).all()
        assert (cp.cpu().numpy() == want_cons[1024:1536]).all()
        spec2 = api.SynthSpec(seed=8, p_sub=0.01, p_del=0.01, p_ins=0.05, p_n_block=0.02, p_amb=0.01, accessory_frac=0.3)
        c, _, _ = synth_on_device(ctx, spec2, 500, 0, 4096)
        assert not bool((a == c).all())
        codes = set(np.unique(full).tolist())
        assert {0, 1, 2, 4, 8, 15} <= codes


def test_cpu_twin_of_the_generator_matches_the_device_cell_for_cell():
    """oracle/synth_twin.c (what bench.py's reference arm generates its input with, without a GPU) against the device
    generator: every cell of a 700-taxon alignment with all event kinds, of a polytomy tree with absent leaves, and of
    column windows far from 0 (the hash takes the GLOBAL column)."""
    import synth as S
    specs = [api.SynthSpec(seed=7, p_sub=0.01, p_del=0.01, p_ins=0.05, p_n_block=0.02, p_amb=0.01, accessory_frac=0.3),
             api.SynthSpec()]
    rng = np.random.default_rng(3)
    for tree in (trees.random_binary_tree(700, seed=2), S.random_tree(400, rng, max_arity=5, absent_frac=0.05),
                 trees.caterpillar_tree(300)):
        rows = S.n_rows(tree)
        with api.Context(tree) as ctx:
            for spec in specs:
                for col0, L in ((0, 4096), (29984, 16), (1 << 20, 1000)):
                    d_msa, d_cons, _ = synth_on_device(ctx, spec, rows, col0, L)
                    dev = api.unpack_nibbles(d_msa.cpu().numpy(), L)
                    twin = O.synth_columns(tree, spec, np.arange(col0, col0 + L), rows, threads=4)
                    assert (dev == twin).all(), (tree.n_nodes, col0, L)
                    nz = twin != 0
                    assert (d_cons.cpu().numpy() == twin[nz.argmax(0), np.arange(L)]).all()


@pytest.mark.parametrize("algo", [api.FITCH, api.SANKOFF])
def test_config2_full_size_properties_and_subsample_parity(algo):
    """BASELINE config 2 (20 000 taxa x 30 000 columns, random binary tree) at full size:
    sortedness/uniqueness of the records, tiling- and chunking-invariance, host path == device path,
    and bit-exactness against the oracle on a 96-column subsample."""
    torch = _torch()
    N, L = 20000, 30000
    tree = trees.random_binary_tree(N)
    spec = api.SynthSpec()
    with api.Context(tree) as ctx:
        d_msa, d_cons, stride = synth_on_device(ctx, spec, N, 0, L)
        _, n = ctx.run_device(algo, d_msa.data_ptr(), stride, N, L, d_cons.data_ptr())
        rec = ctx.download(n)
        key = rec["node"].astype(np.int64) << 32 | rec["pos"].astype(np.int64)
        assert (np.diff(key) > 0).all(), "records must be strictly sorted by (node,pos)"
        assert rec["pos"].min() >= 0 and rec["pos"].max() < L and rec["node"].max() < tree.n_nodes
        assert ((rec["type"] == 1) == (rec["code"] == 0)).all()      # ND <=> state is a gap
        assert ((rec["type"] == 2) == (rec["parent_code"] == 0)).all()  # NI <=> parent is a gap
        assert (rec["code"] != rec["parent_code"]).all()
        digest = G.digest_muts(rec)
        # tiling / chunking must not change a single record
        ctx.set_tuning(chunk_nodes=97, tile_cols=7000)
        _, n2 = ctx.run_device(algo, d_msa.data_ptr(), stride, N, L, d_cons.data_ptr())
        assert n2 == n and G.digest_muts(ctx.download(n2)) == digest
        # host-buffer path
        ctx.set_tuning(0, 0)
        host = d_msa.cpu().numpy()
        cons = d_cons.cpu().numpy()
        got = ctx.run(algo, host, L, cons)
        assert G.digest_muts(got) == digest
        # oracle on a column subsample (the CPU reference needs ~0.5 ms per column.node)
        cols = np.sort(np.random.default_rng(1).choice(L, size=96, replace=False))
        sub = api.unpack_nibbles(host, L)[:, cols]
        want = O.run_port(algo, tree, sub, cons[cols], None, threads=8)
        sel = np.isin(rec["pos"], cols)
        mine = rec[sel].copy()
        mine["pos"] = np.searchsorted(cols, mine["pos"])
        mine = mine[np.lexsort((mine["pos"], mine["node"]))]
        assert G.same_muts(mine, want)


def test_shard_mode_two_contexts_equal_unsharded():
    """Column shards on ONE GPU (two contexts stand in for two ranks): shard-mode runs, the all-to-all emulated by
    slicing, pmfs_sort_records on each owner -> every owner holds exactly its node range of the unsharded result."""
    torch = _torch()
    from panman_b200 import shard

    N, L = 1500, 8000
    tree = trees.random_binary_tree(N, seed=9)
    spec = api.SynthSpec(p_sub=0.003, p_del=0.002, p_ins=0.01, p_n_block=0.01, p_amb=0.001)
    world = 2
    cols = shard.column_shards(L, world)
    nodes = shard.node_ranges(tree.n_nodes, world)
    bounds = [b for b, _ in nodes] + [tree.n_nodes]
    with api.Context(tree) as full:
        d_msa, d_cons, stride = synth_on_device(full, spec, N, 0, L)
        _, n = full.run_device(api.FITCH, d_msa.data_ptr(), stride, N, L, d_cons.data_ptr())
        want = full.download(n)
    ctxs, parts, keep = [], [], []
    for r in range(world):
        ctx = api.Context(tree)
        b, e = cols[r]
        m, c, st = synth_on_device(ctx, spec, N, b, e - b)
        assert bool((api.unpack_nibbles(m.cpu().numpy(), e - b) == api.unpack_nibbles(d_msa.cpu().numpy(), L)[:, b:e]).all())
        ctx.set_shard(b, L, bounds)
        _, n_r = ctx.run_device(api.FITCH, m.data_ptr(), st, N, e - b, c.data_ptr())
        base, offs, cnt, col_ordered = ctx.get_parts()
        assert sum(cnt) == n_r and col_ordered
        parts.append([shard._as_tensor(base + 8 * offs[o], cnt[o], "<i8", "cuda").clone() for o in range(world)])
        ctxs.append(ctx)
        keep.append((m, c))
    for r in range(world):  # what rank r would receive from everybody
        inbox = torch.cat([parts[s][r] for s in range(world)])
        sel = (want["node"] >= nodes[r][0]) & (want["node"] < nodes[r][1])
        for ordered in (True, False):  # stable sort on the node bits alone, and the full sort
            box = inbox.clone()
            ctxs[r].sort_records(box.data_ptr(), int(box.shape[0]), column_ordered=ordered)
            got = ctxs[r].download(int(box.shape[0]))
            assert G.same_muts(got, want[sel]) and (got["parent_code"] == want[sel]["parent_code"]).all(), (r, ordered)
    for ctx in ctxs:
        ctx.close()


def test_local_column_shards_stitch_to_unsharded():
    """Column shards that keep their output (pmfs_set_shard with n_parts = 0: global positions, ordered per shard, no
    exchange): stitching the shards' per-node slices on the host gives exactly the unsharded result."""
    from panman_b200 import shard

    N, L = 1200, 9000
    tree = trees.random_binary_tree(N, seed=11)
    spec = api.SynthSpec(p_sub=0.003, p_del=0.002, p_ins=0.01, p_n_block=0.01, p_amb=0.001)
    with api.Context(tree) as full:
        d_msa, d_cons, stride = synth_on_device(full, spec, N, 0, L)
        _, n = full.run_device(api.FITCH, d_msa.data_ptr(), stride, N, L, d_cons.data_ptr())
        want = full.download(n)
    for world in (2, 3):
        got = []
        for b, e in shard.column_shards(L, world):
            with api.Context(tree) as ctx:
                m, c, st = synth_on_device(ctx, spec, N, b, e - b)
                ctx.set_shard(b, L, None)
                _, n_r = ctx.run_device(api.SANKOFF if world == 3 else api.FITCH, m.data_ptr(), st, N, e - b, c.data_ptr())
                rec = ctx.download(n_r)
                assert rec.shape[0] == 0 or (rec["pos"].min() >= b and rec["pos"].max() < e)
                got.append(rec)
        if world == 2:
            st_ = shard.stitch_by_node(got, tree.n_nodes)
            assert G.same_muts(st_, want) and (st_["parent_code"] == want["parent_code"]).all()
        else:  # Sankoff agrees with Fitch on record count only up to ties: check its shards against its own full run
            with api.Context(tree) as full:
                d_msa, d_cons, stride = synth_on_device(full, spec, N, 0, L)
                _, n = full.run_device(api.SANKOFF, d_msa.data_ptr(), stride, N, L, d_cons.data_ptr())
                want_s = full.download(n)
            assert G.same_muts(shard.stitch_by_node(got, tree.n_nodes), want_s)


def test_host_buffer_call_pipelines_two_slabs():
    """pmfs_run on an alignment wide and big enough for the two-slab upload (>= 2 column tiles, >= 32 MB): same records
    as the device-resident call, for a row stride equal to and larger than the device stride."""
    N, L = 3000, 24000 + 40
    tree = trees.random_binary_tree(N, seed=4)
    spec = api.SynthSpec(p_sub=0.002, p_del=0.001, p_ins=0.004, p_n_block=0.004, p_amb=0.001)
    with api.Context(tree) as ctx:
        d_msa, d_cons, stride = synth_on_device(ctx, spec, N, 0, L)
        _, n = ctx.run_device(api.FITCH, d_msa.data_ptr(), stride, N, L, d_cons.data_ptr())
        want = ctx.download(n)
        h_msa, h_cons = d_msa.cpu().numpy(), d_cons.cpu().numpy()
        assert h_msa.nbytes >= 32 << 20
        for pad in (0, 48):
            h = np.zeros((N, stride + pad), np.uint8)
            h[:, :stride] = h_msa
            h[:, stride:] = 0xff  # bytes beyond the row must never be looked at
            for algo in (api.FITCH, api.SANKOFF):
                got = ctx.run(algo, h, L, h_cons)
                if algo == api.FITCH:
                    assert G.same_muts(got, want) and (got["parent_code"] == want["parent_code"]).all(), pad
                else:
                    _, n_s = ctx.run_device(api.SANKOFF, d_msa.data_ptr(), stride, N, L, d_cons.data_ptr())
                    assert G.same_muts(got, ctx.download(n_s)), pad
