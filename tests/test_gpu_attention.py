"""Varlen self-attention kernels against a float64 torch restatement of aggregator.py:70-72's attention (per slide, per head:
softmax(q k^T / sqrt(32)) v): the fp32 SIMT kernels, and the tcgen05 kernels of csrc/attention_tc.cu in both precision modes,
forward and backward, on ragged bags that cover both size classes (<= 64, 65..128 tokens) and the tiled path (> 128)."""
import ctypes as C

import pytest
import torch

from tests.common import rel_err

pytestmark = pytest.mark.gpu
TOL = {0: 2e-5, 1: 1e-4, 2: 2e-2}


def _run(prec, qkv, seg, B, d, heads, max_bag, dout=None, p=0.0, seed=0):
    from paths_b200 import _lib as L
    lib = L.lib()
    ntok = qkv.shape[0]
    out = torch.zeros(ntok, d, device="cuda"); lse = torch.zeros(ntok, heads, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    L.check(lib.paths_attention_fwd(prec, qkv.data_ptr(), seg.data_ptr(), B, out.data_ptr(), lse.data_ptr(), d, heads, max_bag, p, seed, st), "attention_fwd")
    if dout is None:
        return out, lse
    dqkv = torch.full((ntok, 3 * d), float("nan"), device="cuda"); dvec = torch.zeros(ntok, heads, device="cuda")
    L.check(lib.paths_attention_bwd(prec, qkv.data_ptr(), seg.data_ptr(), B, out.data_ptr(), lse.data_ptr(), dout.data_ptr(), dqkv.data_ptr(),
                                    dvec.data_ptr(), d, heads, max_bag, p, seed, st), "attention_bwd")
    return out, lse, dqkv


def _reference(qkv, bags, d, heads, dout):
    """float64, slide by slide; returns out, lse, dqkv."""
    q = qkv.double().clone().requires_grad_(True)
    outs, lses = [], []
    t0 = 0
    dh = d // heads
    for n in bags:
        T = n + 1
        x = q[t0:t0 + T]
        qq, kk, vv = (x[:, i * d:(i + 1) * d].view(T, heads, dh).transpose(0, 1) for i in range(3))
        s = qq @ kk.transpose(1, 2) / dh ** 0.5
        lses.append(torch.logsumexp(s, -1).transpose(0, 1))
        outs.append((torch.softmax(s, -1) @ vv).transpose(0, 1).reshape(T, d))
        t0 += T
    out = torch.cat(outs); lse = torch.cat(lses)
    out.backward(dout.double())
    return out.detach(), lse.detach(), q.grad


def _case(bags, seed=0, d=128, heads=4):
    g = torch.Generator(device="cuda").manual_seed(seed)
    ntok = sum(bags) + len(bags)
    qkv = torch.randn(ntok, 3 * d, device="cuda", generator=g)
    dout = torch.randn(ntok, d, device="cuda", generator=g)
    seg = torch.tensor([0] + list(torch.tensor(bags).cumsum(0)), dtype=torch.int32, device="cuda")
    return qkv, dout, seg


BAGS = {
    "short": [1, 7, 31, 32, 33, 56, 63, 20, 45, 60, 2, 63, 50],                     # every bag <= 64 tokens: one size class
    "mixed": [63, 64, 65, 100, 127, 5, 80, 56, 57, 126, 1, 96],                     # both classes in one launch
    "long": [40, 300, 127, 128, 129, 64, 1000, 3],                                  # + bags for the tiled kernels
}


@pytest.mark.parametrize("prec", [0, 1, 2])
@pytest.mark.parametrize("name", list(BAGS))
def test_attention_matches_float64(name, prec):
    bags = BAGS[name]
    qkv, dout, seg = _case(bags, seed=len(name))
    out, lse, dqkv = _run(prec, qkv, seg, len(bags), 128, 4, max(bags), dout)
    rout, rlse, rdqkv = _reference(qkv, bags, 128, 4, dout)
    assert torch.isfinite(dqkv).all(), "every dq / dk / dv row must be written"
    assert rel_err(out, rout) < TOL[prec]
    assert float((lse.double() - rlse).abs().max()) < TOL[prec] * 10
    for i, what in enumerate(("dq", "dk", "dv")):
        assert rel_err(dqkv[:, i * 128:(i + 1) * 128], rdqkv[:, i * 128:(i + 1) * 128]) < TOL[prec], what


@pytest.mark.parametrize("prec", [1, 2])
def test_tensor_core_attention_drops_like_the_simt_kernels(prec):
    """Dropout of the probabilities: the same hashed keep/drop decisions in both implementations."""
    bags = BAGS["mixed"]
    qkv, dout, seg = _case(bags, seed=5)
    a = _run(0, qkv, seg, len(bags), 128, 4, max(bags), dout, p=0.2, seed=1234)
    b = _run(prec, qkv, seg, len(bags), 128, 4, max(bags), dout, p=0.2, seed=1234)
    plain = _run(0, qkv, seg, len(bags), 128, 4, max(bags), dout)
    assert rel_err(a[0], plain[0]) > 1e-2                                       # dropout really acted
    for x, y in zip(a, b):
        assert rel_err(y, x) < TOL[prec]
