"""GPU (B200): pmfs_run_fitch_static, the Fitch entry point that performs no device allocation.

Some users of this library cannot call the CUDA allocator at the point where they need the mutations: a serving loop
that may not stall the device (cudaMalloc / cudaFree synchronise it), a process whose HBM is owned by another allocator,
or code running where the allocator is off limits.  For them the library keeps a fixed 32 MiB array in device memory and
carves every intermediate of the Fitch pass out of it.  Because the array is shared by all calls and keeps the previous
call's contents, this test runs the entry point twice in a row on DIFFERENT alignments and requires each result to be
identical, record for record, to what the ordinary entry point (pmfs_run_device) returns for the same input and to what
the reference's own CPU code (oracle/_ref) computes.  It also checks that a problem beyond the fixed size is refused
with a message instead of being attempted.

The file name sorts last on purpose: with `pytest -x` every other test has already run if this one fails."""
import numpy as np
import pytest

import golden_util as G
import synth
from oracle import oracle_api as O
from panman_b200 import api, trees

pytestmark = pytest.mark.gpu


def _device_inputs(torch, msa_codes, cons):
    nib = api.pack_nibbles(msa_codes, row_stride=((msa_codes.shape[1] + 31) // 32) * 16)
    d_msa = torch.from_numpy(nib).cuda()
    d_cons = torch.from_numpy(np.ascontiguousarray(cons)).cuda()
    torch.cuda.synchronize()
    return d_msa, d_cons


def test_static_scratch_entry_point_matches_run_device_twice_in_a_row():
    import torch

    tree = trees.random_binary_tree(300, seed=11)
    L = 6000
    cap = 400000
    with api.Context(tree) as ctx:
        d_out = torch.zeros(cap * 12, dtype=torch.uint8, device="cuda")
        for call, seed in enumerate((1, 2)):
            rng = np.random.default_rng(seed)
            msa = synth.simulate_msa(tree, L, rng, p_sub=0.01 * (call + 1), p_del=0.005, p_ins=0.02, p_n=0.01, p_amb=0.002)
            cons = synth.consensus_first_nongap(msa)
            forced = msa[5] if call == 1 else None
            d_msa, d_cons = _device_inputs(torch, msa, cons)
            d_forced = torch.from_numpy(np.ascontiguousarray(forced)).cuda() if forced is not None else None
            torch.cuda.synchronize()
            stride = d_msa.shape[1]
            n = ctx.run_fitch_static(d_msa.data_ptr(), stride, 300, L, d_cons.data_ptr(),
                                     d_forced.data_ptr() if d_forced is not None else 0, d_out.data_ptr(), cap)
            got = d_out[: n * 12].cpu().numpy().view(api.MUT_DTYPE).copy()
            _, n_ref = ctx.run_device(api.FITCH, d_msa.data_ptr(), stride, 300, L, d_cons.data_ptr(),
                                      d_forced.data_ptr() if d_forced is not None else 0)
            want = ctx.download(n_ref)
            assert n == n_ref and n > 1000, (call, n, n_ref)
            assert G.same_muts(got, want), call
            assert (got["parent_code"] == want["parent_code"]).all(), call
            # ... and to what the reference's CPU code computes (both entry points share kernels: agreeing with each
            # other is not enough)
            ref = O.run_ref_columns(api.FITCH, tree, O.codes_to_chars(msa), O.codes_to_chars(cons),
                                    None if forced is None else O.codes_to_chars(forced), threads=4)
            assert G.same_muts(got, ref), call


def test_static_scratch_entry_point_rejects_what_does_not_fit():
    import torch

    tree = trees.random_binary_tree(3000, seed=3)  # 2999 internal nodes x 938 groups x 64 B = 180 MB > 32 MiB
    L = 30000
    with api.Context(tree) as ctx:
        stride = ((L + 31) // 32) * 16
        d_msa = torch.zeros((3000, stride), dtype=torch.uint8, device="cuda")
        d_cons = torch.zeros(L, dtype=torch.uint8, device="cuda")
        d_out = torch.zeros(1024 * 12, dtype=torch.uint8, device="cuda")
        with pytest.raises(api.PmfsError) as e:
            ctx.run_fitch_static(d_msa.data_ptr(), stride, 3000, L, d_cons.data_ptr(), 0, d_out.data_ptr(), 1024)
        assert e.value.status == -1 and "static maximum" in str(e.value)
    # a result larger than the caller's buffer is reported, with the count needed
    small = trees.random_binary_tree(50, seed=4)
    rng = np.random.default_rng(0)
    msa = synth.simulate_msa(small, 640, rng, p_sub=0.2)
    cons = synth.consensus_first_nongap(msa)
    with api.Context(small) as ctx:
        d_msa, d_cons = _device_inputs(torch, msa, cons)
        d_out = torch.zeros(16 * 12, dtype=torch.uint8, device="cuda")
        with pytest.raises(api.PmfsError) as e:
            ctx.run_fitch_static(d_msa.data_ptr(), d_msa.shape[1], 50, 640, d_cons.data_ptr(), 0, d_out.data_ptr(), 16)
        assert e.value.status == -1 and e.value.n_out > 16


def test_static_scratch_entry_point_on_a_split_polytomy():
    """A tree with a split polytomy (more than 128 children): the static entry point runs the part / combine programs
    out of its fixed array; and it refuses -- instead of re-planning, which would allocate -- when the context's plan was
    last built for Sankoff, whose partial results have another layout."""
    import torch

    tree = synth._flatten([list(range(200)), list(range(200, 230))])
    L = 700
    rng = np.random.default_rng(3)
    msa = synth.simulate_msa(tree, L, rng, p_sub=0.05, p_del=0.01, p_ins=0.02, p_n=0.01, p_amb=0.01)
    cons = synth.consensus_first_nongap(msa)
    d_msa, d_cons = _device_inputs(torch, msa, cons)
    stride, cap = d_msa.shape[1], 200000
    d_out = torch.zeros(cap * 12, dtype=torch.uint8, device="cuda")
    want = O.run_port(api.FITCH, tree, msa, cons, None, threads=2)
    with api.Context(tree) as ctx:
        n = ctx.run_fitch_static(d_msa.data_ptr(), stride, 230, L, d_cons.data_ptr(), 0, d_out.data_ptr(), cap)
        assert ctx.plan_info()["n_parts"] == 4
        assert G.same_muts(d_out[: n * 12].cpu().numpy().view(api.MUT_DTYPE).copy(), want)
        ctx.run_device(api.SANKOFF, d_msa.data_ptr(), stride, 230, L, d_cons.data_ptr())
        with pytest.raises(api.PmfsError) as e:
            ctx.run_fitch_static(d_msa.data_ptr(), stride, 230, L, d_cons.data_ptr(), 0, d_out.data_ptr(), cap)
        assert e.value.status == -5 and "Sankoff" in str(e.value)
        ctx.run_device(api.FITCH, d_msa.data_ptr(), stride, 230, L, d_cons.data_ptr())
        n = ctx.run_fitch_static(d_msa.data_ptr(), stride, 230, L, d_cons.data_ptr(), 0, d_out.data_ptr(), cap)
        assert G.same_muts(d_out[: n * 12].cpu().numpy().view(api.MUT_DTYPE).copy(), want)
