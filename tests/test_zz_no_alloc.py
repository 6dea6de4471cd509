"""paths_attention_bwd_static -- the allocation-free variant of paths_attention_bwd (include/paths_b200.h): its [tokens, heads]
scratch is a file-scope device array of the library instead of a caller buffer.  Checked through the C-ABI against the existing
entry point on the same inputs, on two calls in a row (the second with other data: the array keeps the first call's contents),
in the fp32 SIMT mode and on the tiled tcgen05 path -- the two paths that use the scratch -- and that sizes beyond the stated
maximum are rejected without a launch.  Collected last on purpose."""
import pytest
import torch

pytestmark = pytest.mark.gpu
TOL = {0: 1e-6, 1: 1e-6}          # same kernels, same inputs, same launch shapes: the results agree to rounding of nothing at all


def _case(seed, bags, d, heads):
    g = torch.Generator().manual_seed(seed)
    B = len(bags)
    seg = torch.zeros(B + 1, dtype=torch.int32)
    seg[1:] = torch.tensor(bags, dtype=torch.int32).cumsum(0)
    ntok = int(seg[-1]) + B
    qkv = torch.randn(ntok, 3 * d, generator=g).cuda()
    dout = torch.randn(ntok, d, generator=g).cuda()
    return seg.cuda(), B, ntok, qkv, dout


def _both(prec, seg, B, ntok, qkv, dout, d, heads, max_bag):
    from paths_b200 import _lib as L
    lib = L.lib()
    st = torch.cuda.current_stream().cuda_stream
    out = torch.zeros(ntok, d, device="cuda"); lse = torch.zeros(ntok, heads, device="cuda")
    L.check(lib.paths_attention_fwd(prec, qkv.data_ptr(), seg.data_ptr(), B, out.data_ptr(), lse.data_ptr(), d, heads, max_bag, 0.0, 0, st), "attention_fwd")
    want = torch.full((ntok, 3 * d), float("nan"), device="cuda"); dvec = torch.zeros(ntok, heads, device="cuda")
    L.check(lib.paths_attention_bwd(prec, qkv.data_ptr(), seg.data_ptr(), B, out.data_ptr(), lse.data_ptr(), dout.data_ptr(), want.data_ptr(),
                                    dvec.data_ptr(), d, heads, max_bag, 0.0, 0, st), "attention_bwd")
    got = torch.full((ntok, 3 * d), float("nan"), device="cuda")
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    L.check(lib.paths_attention_bwd_static(prec, qkv.data_ptr(), seg.data_ptr(), B, ntok, out.data_ptr(), lse.data_ptr(), dout.data_ptr(),
                                           got.data_ptr(), d, heads, max_bag, 0.0, 0, st), "attention_bwd_static")
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == before
    return got, want


@pytest.mark.parametrize("prec", [0, 1])
def test_static_scratch_backward_matches_the_caller_scratch_entry_twice(prec):
    d, heads = 128, 4
    # ragged bags: short ones, the 65..128 class and (tensor-core mode: the only one of its paths that uses the scratch) tiled ones
    for seed, bags in [(1, [5, 70, 300, 1, 129, 64]), (2, [200, 3, 33, 257, 90, 12, 131])]:
        seg, B, ntok, qkv, dout = _case(seed, bags, d, heads)
        got, want = _both(prec, seg, B, ntok, qkv, dout, d, heads, max(bags))
        assert torch.isfinite(got).all()
        err = float((got - want).abs().max() / want.abs().max())
        assert err <= TOL[prec], (prec, seed, err)


def test_static_scratch_rejects_sizes_beyond_its_maximum():
    from paths_b200 import _lib as L
    lib = L.lib()
    d, heads = 128, 4
    seg, B, ntok, qkv, dout = _case(3, [4, 4], d, heads)
    too_many = (1 << 20) // heads + 1                       # PATHS_ATTENTION_BWD_STATIC_MAX_FLOATS / heads + 1 token rows
    out = torch.zeros(ntok, d, device="cuda"); lse = torch.zeros(ntok, heads, device="cuda"); dqkv = torch.zeros(ntok, 3 * d, device="cuda")
    rc = lib.paths_attention_bwd_static(0, qkv.data_ptr(), seg.data_ptr(), B, too_many, out.data_ptr(), lse.data_ptr(), dout.data_ptr(),
                                        dqkv.data_ptr(), d, heads, 4, 0.0, 0, torch.cuda.current_stream().cuda_stream)
    assert rc != 0
    msg = lib.paths_last_error().decode()
    assert "exceeds the static scratch" in msg and "PATHS_ATTENTION_BWD_STATIC_MAX_FLOATS" in msg
    torch.cuda.synchronize()
    assert float(dqkv.abs().max()) == 0.0                   # nothing was launched
