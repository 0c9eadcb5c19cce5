"""GPU parity of the Conformer convolution-module / rel-pos helper kernels through the C ABI against an fp32
torch reference of the same op on the same bf16-rounded inputs (floating-point kernels: SURVEY.md §8c keeps a
torch fp32 reference for those).  Tolerances are written next to each assertion: outputs are bf16, so the bound
is a few bf16 ulps of the output scale; reductions over rows (weight gradients, BN statistics) are fp32."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu
BF16 = torch.bfloat16


def _rand(shape, seed, scale=1.0):
    g = torch.Generator().manual_seed(seed)
    return (torch.randn(shape, generator=g) * scale).to(BF16)


def _close(got, ref, rel=1e-2, msg=""):
    got, ref = got.float().cpu(), ref.float().cpu()
    err = (got - ref).abs().max().item()
    assert err <= rel * (ref.abs().max().item() + 1e-6), (msg, err, ref.abs().max().item())


@pytest.mark.parametrize("B,T,C,K", [(3, 83, 256, 15), (2, 250, 512, 31), (1, 5, 64, 15), (4, 16, 176, 31)])
def test_dwconv1d_fwd_bwd(cuda_lib, B, T, C, K):
    """convolution.py:48-56 depthwise Conv1d: bf16 outputs within 1 % of max (8-bit mantissa); BN partial sums fp32."""
    from s2t_b200 import ops
    x, w, dy = _rand((B, T, C), 1), _rand((C, K), 2, 0.3), _rand((B, T, C), 3)
    xr = x.float().transpose(1, 2).requires_grad_(True)
    wr = w.float().view(C, 1, K).requires_grad_(True)
    yr = F.conv1d(xr, wr, padding=(K - 1) // 2, groups=C)
    yr.backward(dy.float().transpose(1, 2))
    part = torch.zeros((ops.dwconv1d_tiles(B, T), 2, C), device="cuda")
    y = ops.dwconv1d_fwd(x.cuda(), w.cuda(), bn_partials=part)
    _close(y, yr.detach().transpose(1, 2), msg="y")
    n = B * T
    s = part.sum(0).cpu()
    yd = yr.detach().transpose(1, 2).reshape(n, C)
    assert torch.allclose(s[0], yd.sum(0), atol=2e-3 * n ** 0.5 * yd.abs().max().item())
    assert torch.allclose(s[1], (yd * yd).sum(0), rtol=2e-3)
    dw = torch.zeros((C, K), dtype=BF16, device="cuda")
    dx = ops.dwconv1d_bwd(x.cuda(), w.cuda(), dy.cuda(), dw)
    _close(dx, xr.grad.transpose(1, 2), msg="dx")
    _close(dw, wr.grad.view(C, K), msg="dw")


@pytest.mark.parametrize("act", ["relu", "swish", "none"])
@pytest.mark.parametrize("rows,C", [(249, 256), (2000, 512), (7, 64)])
def test_batchnorm_train_fwd_bwd(cuda_lib, act, rows, C):
    """nn.BatchNorm1d training mode (batch statistics, running-stat update with the unbiased variance) + activation."""
    from s2t_b200 import ops
    code = {"relu": ops.ACT_RELU, "swish": ops.ACT_SWISH, "none": ops.ACT_NONE}[act]
    fn = {"relu": F.relu, "swish": F.silu, "none": lambda v: v}[act]
    x = _rand((rows, C), 5, 2.0) + 0.5
    x = x.to(BF16)
    gamma, beta, da = (_rand((C,), 6, 0.2) + 1).to(BF16), _rand((C,), 7, 0.2), _rand((rows, C), 8)
    bn = torch.nn.BatchNorm1d(C)
    with torch.no_grad():
        bn.weight.copy_(gamma.float()); bn.bias.copy_(beta.float())
    bn.train()
    xr = x.float().requires_grad_(True)
    yr = fn(bn(xr))
    yr.backward(da.float())
    # partial sums the way dwconv1d_fwd emits them: [n][2][C]
    xf = x.float()
    chunks = torch.stack([torch.stack([c.sum(0), (c * c).sum(0)]) for c in xf.split(16)]).cuda()
    rm, rv = torch.zeros(C, device="cuda"), torch.ones(C, device="cuda")
    nbt = torch.zeros((), dtype=torch.int64, device="cuda")
    save = torch.empty((2, C), device="cuda")
    y = ops.batchnorm_fwd(x.cuda(), gamma.cuda(), beta.cuda(), partials=chunks, running_mean=rm, running_var=rv,
                          num_batches_tracked=nbt, save_stats=save, act=code, training=True)
    _close(y, yr.detach(), msg="y")
    assert torch.allclose(rm.cpu(), bn.running_mean, atol=1e-4) and torch.allclose(rv.cpu(), bn.running_var, rtol=1e-4)
    assert int(nbt.item()) == 1
    dg, db = torch.zeros(C, dtype=BF16, device="cuda"), torch.zeros(C, dtype=BF16, device="cuda")
    dx = ops.batchnorm_bwd(da.cuda(), x.cuda(), save, gamma.cuda(), beta.cuda(), dg, db, act=code, training=True)
    _close(dx, xr.grad, rel=2e-2, msg="dx")
    _close(dg, bn.weight.grad, rel=2e-2, msg="dgamma")
    _close(db, bn.bias.grad, rel=2e-2, msg="dbeta")


def test_batchnorm_eval_uses_running_stats(cuda_lib):
    from s2t_b200 import ops
    rows, C = 100, 64
    x, gamma, beta = _rand((rows, C), 1), (_rand((C,), 2, 0.1) + 1).to(BF16), _rand((C,), 3, 0.1)
    rm, rv = torch.randn(C) * 0.1, torch.rand(C) + 0.5
    save = torch.empty((2, C), device="cuda")
    rmc, rvc = rm.cuda(), rv.cuda()
    y = ops.batchnorm_fwd(x.cuda(), gamma.cuda(), beta.cuda(), partials=None, running_mean=rmc, running_var=rvc,
                          num_batches_tracked=None, save_stats=save, act=ops.ACT_NONE, training=False)
    ref = F.batch_norm(x.float(), rm, rv, gamma.float(), beta.float(), training=False)
    _close(y, ref)
    assert torch.equal(rmc.cpu(), rm) and torch.equal(rvc.cpu(), rv)


def test_relpos_qbias_and_act_dropout(cuda_lib):
    from s2t_b200 import ops
    rows, C = 333, 256
    qkv, u, v = _rand((rows, 3 * C), 1), _rand((C,), 2), _rand((C,), 3)
    quv = ops.relpos_qbias_fwd(qkv.cuda(), u.cuda(), v.cuda(), C).cpu()
    q = qkv[:, :C].float()
    assert torch.equal(quv[:, :C], (q + u.float()).to(BF16)) and torch.equal(quv[:, C:], (q + v.float()).to(BF16))
    dquv = _rand((rows, 2 * C), 4)
    dqkv = torch.zeros((rows, 3 * C), dtype=BF16, device="cuda")
    ops.relpos_qbias_bwd(dquv.cuda(), dqkv, C)
    assert torch.equal(dqkv.cpu()[:, :C], (dquv[:, :C].float() + dquv[:, C:].float()).to(BF16))
    assert torch.count_nonzero(dqkv[:, C:]) == 0
    # act + dropout: p = 0 is plain swish; p > 0 draws the same mask as s2t_dropout on the same stream id
    x = _rand((rows, C), 5, 2.0)
    y = ops.act_dropout(x.cuda(), ops.ACT_SWISH, None)
    _close(y, F.silu(x.float()), rel=5e-3)
    spec = (0.25, 1234, 7)
    yd = ops.act_dropout(x.cuda(), ops.ACT_SWISH, spec).float().cpu()
    md = ops.dropout(torch.ones_like(x).cuda(), spec).float().cpu()
    assert torch.equal(yd != 0, (md != 0) & (y.float().cpu() != 0))
    keep = md != 0
    assert abs(keep.float().mean().item() - 0.75) < 0.01
    _close(yd[keep], (F.silu(x.float()) / 0.75)[keep], rel=1e-2)


def test_softmax_head_major_bias_matches_batch_major(cuda_lib):
    """bias_head_major only changes where the rel-pos bias rows live, not the values."""
    from s2t_b200 import ops
    B, H, T = 3, 4, 37
    ld, ldb = 40, 80
    sc = torch.zeros((B * H, T, ld), dtype=BF16); sc[..., :T] = _rand((B * H, T, T), 1)
    bd = torch.zeros((B, H, T, ldb), dtype=BF16); bd[..., : 2 * T - 1] = _rand((B, H, T, 2 * T - 1), 2)
    klen = torch.tensor([37, 20, 5], dtype=torch.int32).cuda()
    p0, _ = ops.attn_softmax_fwd(sc.cuda(), klen, H, T, bias=bd.view(B * H, T, ldb).cuda(), bias_mode=2, scale=0.3)
    bd_hm = bd.transpose(0, 1).contiguous()
    p1, _ = ops.attn_softmax_fwd(sc.cuda(), klen, H, T, bias=bd_hm.view(H * B, T, ldb).cuda(), bias_mode=2, scale=0.3,
                                 bias_head_major=True)
    assert torch.equal(p0, p1)
    dp = torch.zeros_like(sc); dp[..., :T] = _rand((B * H, T, T), 3)
    db0 = torch.zeros((B * H, T, ldb), dtype=BF16, device="cuda")
    db1 = torch.zeros((H * B, T, ldb), dtype=BF16, device="cuda")
    d0 = ops.attn_softmax_bwd(p0, dp.cuda(), klen, H, T, dbias=db0, scale=0.3)
    d1 = ops.attn_softmax_bwd(p0, dp.cuda(), klen, H, T, dbias=db1, scale=0.3, bias_head_major=True)
    assert torch.equal(d0, d1)
    assert torch.equal(db0.view(B, H, T, ldb).transpose(0, 1).contiguous().view(H * B, T, ldb), db1)


@pytest.mark.parametrize("rows,V", [(37, 200), (250, 10000), (5, 1001)])
def test_softmax_rows_fwd_bwd(cuda_lib, rows, V):
    """Adapter distribution (adapter.py:215-217) from the CTC row log-sum-exp, and its backward accumulated onto an
    existing gradient.  bf16 probabilities: 1 % of the row maximum."""
    from s2t_b200 import ops
    ld = (V + 7) // 8 * 8
    logits = torch.zeros((rows, ld), dtype=BF16)
    logits[:, :V] = _rand((rows, V), 1, 3.0)
    lg = logits.cuda()
    lse = ops.row_lse(lg, V)
    p = ops.softmax_rows_fwd(lg, lse, V)
    x = logits[:, :V].float().requires_grad_(True)
    ref = torch.softmax(x, -1)
    _close(p[:, :V], ref.detach(), msg="probs")
    assert torch.count_nonzero(p[:, V:]) == 0
    dp = torch.zeros((rows, ld), dtype=BF16)
    dp[:, :V] = _rand((rows, V), 2)
    ref.backward(dp[:, :V].float())
    base = torch.zeros((rows, ld), dtype=BF16)
    base[:, :V] = _rand((rows, V), 3, 0.01)
    out = base.clone().cuda()
    pr = torch.zeros((rows, ld), dtype=BF16)
    pr[:, :V] = ref.detach().to(BF16)
    ops.softmax_rows_bwd(pr.cuda(), dp.cuda(), out, V, accumulate=True)
    _close(out[:, :V], x.grad + base[:, :V].float(), rel=2e-2, msg="dlogits")
    out2 = torch.full((rows, ld), 7.0, dtype=BF16, device="cuda")
    ops.softmax_rows_bwd(pr.cuda(), dp.cuda(), out2, V, accumulate=False)
    _close(out2[:, :V], x.grad, rel=2e-2, msg="dlogits (overwrite)")
