"""GPU (B200): column shards gathered by node range (pmfs_comm_*, pmfs_multi_*).

The cut / exchange / stitch logic is independent of the transport, so most of it is checked on ONE GPU: several contexts
of this process stand in for the ranks and exchange through the library's in-process transport (device-to-device copies
at host barriers, pmfs_comm_create_local), one Python thread per rank.  The NCCL transport itself needs two GPUs; those
tests skip on a single-GPU box and run in `gpurun --gpus 2`."""
import os
import subprocess
import sys
import threading

import numpy as np
import pytest

import golden_util as G
import synth
from oracle import oracle_api as O
from panman_b200 import api, shard, trees

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _torch():
    import torch
    return torch


def _run_ranks(comms, fn):
    """fn(rank, comm) on one thread per rank (the collective blocks at host barriers); re-raises the first failure."""
    errs = [None] * len(comms)

    def work(r):
        try:
            fn(r, comms[r])
        except BaseException as e:  # noqa: BLE001
            errs[r] = e

    th = [threading.Thread(target=work, args=(r,)) for r in range(len(comms))]
    for t in th:
        t.start()
    for t in th:
        t.join()
    for e in errs:
        if e is not None:
            raise e


@pytest.mark.parametrize("world,algo", [(1, 0), (2, 0), (3, 1), (4, 0)])
def test_gather_by_node_range_equals_unsharded(world, algo):
    """Every rank ends up with exactly its node range of the unsharded result: records, order, parent codes, and the
    <=6-run packing done after the gather (runs straddle the shard boundaries)."""
    torch = _torch()
    rng = np.random.default_rng(50 + world)
    tree = synth.random_tree(900, rng, max_arity=4, absent_frac=0.03)
    L = 5000 + 17
    msa = synth.simulate_msa(tree, L, rng, p_sub=0.004, p_del=0.003, p_ins=0.02, p_n=0.02, p_amb=0.002)
    cons = synth.consensus_first_nongap(msa)
    forced = msa[2] if world == 3 else None
    want = O.run_port(algo, tree, msa, cons, forced, threads=4)
    want_off, want_nuc = O.pack_port(want, tree.n_nodes)
    R = synth.n_rows(tree)
    cols = shard.column_shards(L, world)
    if world == 4:
        e1 = cols[1][1]  # a rank without columns in this call still takes part: rank 0 takes rank 1's columns as well
        cols[0], cols[1] = (cols[0][0], e1), (e1, e1)
    ctxs = [api.Context(tree) for _ in range(world)]
    comms = api.Comm.local_group(ctxs)
    got, packs = [None] * world, [None] * world

    def work(r, comm):
        b, e = cols[r]
        if e > b:
            nib = api.pack_nibbles(msa[:, b:e], row_stride=((e - b + 31) // 32) * 16)
            d_msa = torch.from_numpy(nib).cuda()
            d_cons = torch.from_numpy(np.ascontiguousarray(cons[b:e])).cuda()
            d_forced = torch.from_numpy(np.ascontiguousarray(forced[b:e])).cuda() if forced is not None else None
            torch.cuda.synchronize()
            ptr, n = comm.run(algo, d_msa.data_ptr(), nib.shape[1], R, e - b, b, L, d_cons.data_ptr(),
                              d_forced.data_ptr() if d_forced is not None else 0)
        else:
            ptr, n = comm.run(algo, 0, 16, R, 0, b, L, 0, 0)
        got[r] = ctxs[r].download(n)
        packs[r] = ctxs[r].pack_last()

    _run_ranks(comms, work)
    for r in range(world):
        nb, ne = comms[r].node_range()
        sel = (want["node"] >= nb) & (want["node"] < ne)
        assert G.same_muts(got[r], want[sel]), (world, r)
        off, nuc = packs[r]
        lo, hi = want_off[nb], want_off[ne]
        assert G.same_nuc(nuc, want_nuc[lo:hi]), (world, r)
        assert (off[nb:ne + 1] - off[nb] == want_off[nb:ne + 1] - lo).all()
    # the same group again on other columns: buffers of the previous call must not leak into this one
    def again(r, comm):
        b, e = cols[r]
        sub = np.ascontiguousarray(msa[:, b:e][:, ::-1]) if e > b else None
        if sub is None:
            ptr, n = comm.run(algo, 0, 16, R, 0, b, L, 0, 0)
        else:
            nib = api.pack_nibbles(sub, row_stride=((e - b + 31) // 32) * 16)
            d_msa = torch.from_numpy(nib).cuda()
            d_cons = torch.from_numpy(np.ascontiguousarray(cons[b:e][::-1])).cuda()
            torch.cuda.synchronize()
            ptr, n = comm.run(algo, d_msa.data_ptr(), nib.shape[1], R, e - b, b, L, d_cons.data_ptr(), 0)
        got[r] = ctxs[r].download(n)

    _run_ranks(comms, again)
    msa2, cons2 = msa.copy(), cons.copy()
    for b, e in cols:
        msa2[:, b:e] = msa[:, b:e][:, ::-1]
        cons2[b:e] = cons[b:e][::-1]
    want2 = O.run_port(algo, tree, msa2, cons2, None, threads=4)
    for r in range(world):
        nb, ne = comms[r].node_range()
        sel = (want2["node"] >= nb) & (want2["node"] < ne)
        assert G.same_muts(got[r], want2[sel]), ("again", world, r)
    for c in comms:
        c.close()
    for c in ctxs:
        c.close()


def test_gather_on_the_radix_ordering_path():
    """The gather also works when the local ordering goes through the radix path (no cell matrix): per-node offsets then
    come from a binary search in the ordered records."""
    env = dict(os.environ, PMFS_ORDER="radix")
    r = subprocess.run([sys.executable, "-m", "pytest", os.path.abspath(__file__), "-q", "-x", "-m", "gpu", "-k",
                        "test_gather_by_node_range_equals_unsharded"], env=env, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]


def test_multi_driver_one_process_copy_transport():
    """pmfs_multi_run (one process, one host thread per rank) with the copy transport and all ranks on GPU 0: same
    contract as pmfs_run -- one (node,pos)-ordered host array -- for several rank counts, window sizes (several windows,
    ragged last window, ranks without columns) and both algorithms."""
    rng = np.random.default_rng(77)
    tree = trees.random_binary_tree(600, seed=12)
    L = 9000 + 5
    msa = synth.simulate_msa(tree, L, rng, p_sub=0.004, p_del=0.002, p_ins=0.01, p_n=0.01, p_amb=0.002)
    cons = synth.consensus_first_nongap(msa)
    nib = api.pack_nibbles(msa)
    want = {a: O.run_port(a, tree, msa, cons, None, threads=4) for a in (0, 1)}
    want_f = O.run_port(0, tree, msa, cons, msa[4], threads=4)
    os.environ["PMFS_MULTI_TRANSPORT"] = "copy"
    try:
        for n_ranks, window in ((1, 0), (2, 0), (3, 2048), (5, 4096), (8, 100000)):
            with api.Multi(tree, n_ranks, devices=[0] * n_ranks) as m:
                for algo in (0, 1):
                    got = m.run(algo, nib, L, cons, None, window_cols=window)
                    assert G.same_muts(got, want[algo]), (n_ranks, window, algo)
                got = m.run(0, nib, L, cons, msa[4], window_cols=window)
                assert G.same_muts(got, want_f), (n_ranks, window)
    finally:
        del os.environ["PMFS_MULTI_TRANSPORT"]
    # an alignment narrower than the number of ranks x 32 columns: most ranks get nothing
    narrow = msa[:, :40]
    os.environ["PMFS_MULTI_TRANSPORT"] = "copy"
    try:
        with api.Multi(tree, 4, devices=[0] * 4) as m:
            got = m.run(0, api.pack_nibbles(narrow), 40, cons[:40])
        assert G.same_muts(got, O.run_port(0, tree, narrow, cons[:40], None, threads=2))
    finally:
        del os.environ["PMFS_MULTI_TRANSPORT"]


_NCCL_WORKER = r"""
import os, sys
import numpy as np
import torch
import torch.distributed as dist
sys.path.insert(0, sys.argv[1]); sys.path.insert(0, os.path.join(sys.argv[1], "tests"))
import golden_util as G, synth
from oracle import oracle_api as O
from panman_b200 import api, shard
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(rank)
dist.init_process_group("nccl", device_id=torch.device("cuda", rank))
rng = np.random.default_rng(5)
tree = synth.random_tree(700, rng, max_arity=3)
L = 6000 + 9
msa = synth.simulate_msa(tree, L, rng, p_sub=0.004, p_del=0.003, p_ins=0.02, p_n=0.02, p_amb=0.002)
cons = synth.consensus_first_nongap(msa)
idt = torch.zeros(128, dtype=torch.uint8, device="cuda")
if rank == 0:
    idt.copy_(torch.frombuffer(bytearray(api.comm_unique_id()), dtype=torch.uint8))
dist.broadcast(idt, 0)
ctx = api.Context(tree, device=rank)
comm = api.Comm(ctx, bytes(idt.cpu().numpy().tobytes()), rank, world)
nb, ne = comm.node_range()
for algo in (0, 1):
    # a narrow call first, then the full width twice: the inboxes have to grow between calls
    for lo, hi in ((0, 640), (0, L), (0, L)):
        bb, ee = shard.column_shards(hi - lo, world)[rank]
        nib = api.pack_nibbles(msa[:, bb:ee], row_stride=((ee - bb + 31) // 32) * 16)
        d_msa, d_cons = torch.from_numpy(nib).cuda(), torch.from_numpy(np.ascontiguousarray(cons[bb:ee])).cuda()
        torch.cuda.synchronize()
        ptr, n = comm.run(algo, d_msa.data_ptr(), nib.shape[1], synth.n_rows(tree), ee - bb, bb, hi, d_cons.data_ptr(), 0)
        got = ctx.download(n)
        want = O.run_port(algo, tree, msa[:, :hi], cons[:hi], None, threads=2)
        sel = (want["node"] >= nb) & (want["node"] < ne)
        assert G.same_muts(got, want[sel]), (rank, algo, hi)
comm.close(); ctx.close()
dist.barrier(); dist.destroy_process_group()
print("rank", rank, "ok")
"""


@pytest.mark.parametrize("transport", ["p2p", "nccl"])
def test_gather_two_processes(tmp_path, transport):
    """One process per GPU under torch.distributed.run, the NCCL unique id shipped by a torch broadcast; the records
    travel by peer copies into IPC-mapped inboxes (the default) or by ncclSend / ncclRecv."""
    torch = _torch()
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    script = tmp_path / "worker.py"
    script.write_text(_NCCL_WORKER)
    env = dict(os.environ, PMFS_COMM_TRANSPORT=transport)
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
                        "127.0.0.1", "--master-port", "29731" if transport == "p2p" else "29732", str(script), ROOT],
                       capture_output=True, text=True, timeout=600, env=env)
    assert r.returncode == 0 and r.stdout.count("ok") == 2, r.stdout[-3000:] + r.stderr[-3000:]


def test_multi_driver_over_nccl_in_one_process():
    """pmfs_multi_run with ncclCommInitAll over two GPUs of this process."""
    torch = _torch()
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    rng = np.random.default_rng(78)
    tree = trees.random_binary_tree(500, seed=13)
    L = 8000 + 3
    msa = synth.simulate_msa(tree, L, rng, p_sub=0.004, p_del=0.002, p_ins=0.01, p_n=0.01, p_amb=0.002)
    cons = synth.consensus_first_nongap(msa)
    with api.Multi(tree, 2) as m:
        for algo in (0, 1):
            for window in (0, 4096):
                got = m.run(algo, api.pack_nibbles(msa), L, cons, None, window_cols=window)
                assert G.same_muts(got, O.run_port(algo, tree, msa, cons, None, threads=4)), (algo, window)
