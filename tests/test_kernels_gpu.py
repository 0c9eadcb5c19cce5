"""GPU parity of the HBM-bound kernels (fbank/CMVN, LayerNorm, GLU, colsum, col2im, embed, softmax, CTC,
greedy) through the C ABI, against the CPU oracle (oracle/) or — for plain floating-point elementwise
math — an fp32 torch reference.  Tolerances are written next to each assertion."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
BF16 = torch.bfloat16


def rel_l2(a, b):
    a, b = a.double().cpu(), b.double().cpu()
    return float((a - b).norm() / (b.norm() + 1e-12))


def _rand(shape, seed, scale=1.0, dtype=BF16):
    g = torch.Generator().manual_seed(seed)
    return (torch.randn(shape, generator=g) * scale).to(dtype)


# ------------------------------------------------------------------------------------------------ fbank
@pytest.mark.parametrize("lens", [[160000] * 2, [16000, 12345, 400, 399, 801]])
def test_fbank_matches_oracle(cuda_lib, lens):
    """fbank fp32: max-abs 1e-3 on log-mel values (north_star tolerance)."""
    from oracle import fbank as O
    from s2t_b200 import frontend
    B, N = len(lens), max(lens)
    g = torch.Generator().manual_seed(0)
    wave = torch.randn((B, N), generator=g) * 3000.0
    t = torch.arange(N) / 16000.0
    wave += 2000.0 * torch.sin(2 * np.pi * 440.0 * t)
    for b, n in enumerate(lens):
        wave[b, n:] = 0
    feats, frames = frontend.fbank_batch(wave.cuda(), torch.tensor(lens, dtype=torch.int32).cuda())
    feats, frames = feats.cpu().numpy(), frames.cpu().numpy()
    for b, n in enumerate(lens):
        ref = O.fbank(wave[b, :n].numpy())
        assert frames[b] == ref.shape[0] == O.num_frames(n)
        if ref.shape[0]:
            assert np.abs(feats[b, :frames[b]] - ref).max() < 1e-3
        assert np.all(feats[b, frames[b]:] == 0)


def test_fbank_utterance_cmvn_matches_oracle(cuda_lib):
    """CMVN'd features: max-abs 2e-3."""
    from oracle import fbank as O
    from s2t_b200 import frontend
    lens = [48000, 40000, 33333]
    g = torch.Generator().manual_seed(1)
    wave = torch.randn((3, 48000), generator=g) * 3000.0
    feats, frames = frontend.fbank_batch(wave.cuda(), torch.tensor(lens, dtype=torch.int32).cuda(),
                                         cmvn="utterance")
    feats, frames = feats.cpu().numpy(), frames.cpu().numpy()
    for b, n in enumerate(lens):
        ref = O.utterance_cmvn(O.fbank(wave[b, :n].numpy()))
        assert np.abs(feats[b, :frames[b]] - ref).max() < 2e-3
        assert np.all(feats[b, frames[b]:] == 0)
    # global CMVN
    mean = torch.linspace(10, 20, 80)
    std = torch.linspace(1, 3, 80)
    gfeats, _ = frontend.fbank_batch(wave.cuda(), torch.tensor(lens, dtype=torch.int32).cuda(),
                                     cmvn="global", global_stats=(mean.cuda(), std.cuda()))
    ref = O.global_cmvn(O.fbank(wave[0].numpy()), mean.numpy(), std.numpy())
    assert np.abs(gfeats[0].cpu().numpy() - ref).max() < 1e-3


# ------------------------------------------------------------------------------------------------ LayerNorm
@pytest.mark.parametrize("rows,C", [(2000, 256), (37, 512), (5, 80), (1000, 1024), (64, 176)])
def test_layernorm_fwd_bwd(cuda_lib, rows, C):
    from s2t_b200 import ops
    x, gm, bt, dy = _rand((rows, C), 1, 2.0), _rand((C,), 2), _rand((C,), 3), _rand((rows, C), 4)
    res = _rand((rows, C), 5)
    y, stats = ops.layernorm_fwd(x.cuda(), gm.cuda(), bt.cuda())
    xf = x.float().requires_grad_(True)
    gf, bf = gm.float().requires_grad_(True), bt.float().requires_grad_(True)
    ref = torch.nn.functional.layer_norm(xf, (C,), gf, bf, 1e-5)
    assert (y.float().cpu() - ref).abs().max() <= 2 ** -7 * ref.abs().max()  # one bf16 rounding
    ref.backward(dy.float())
    dg = torch.empty((C,), device="cuda", dtype=BF16)
    db = torch.empty((C,), device="cuda", dtype=BF16)
    dx = ops.layernorm_bwd(dy.cuda(), x.cuda(), gm.cuda(), stats, dg, db, dx_residual=res.cuda())
    torch.cuda.synchronize()
    want_dx = xf.grad + res.float()
    assert (dx.float().cpu() - want_dx).abs().max() <= 2 ** -6 * want_dx.abs().max()
    assert (dg.float().cpu() - gf.grad).abs().max() <= 2 ** -6 * gf.grad.abs().max() + 1e-2
    assert (db.float().cpu() - bf.grad).abs().max() <= 2 ** -6 * bf.grad.abs().max() + 1e-2


def test_layernorm_bwd_dropped_second_output(cuda_lib):
    """out_drop = drop_scale * dropout(dx) with exactly the mask s2t_dropout draws for the same (seed, stream) — the
    pre-masked gradient handed to the next residual branch; the split reduce gives the same dgamma/dbeta."""
    from s2t_b200 import ops
    rows, C = 2000, 256
    x, gm, dy = _rand((rows, C), 1, 2.0), _rand((C,), 2), _rand((rows, C), 4)
    _, stats = ops.layernorm_fwd(x.cuda(), gm.cuda(), gm.cuda())
    dg, db = (torch.empty((C,), device="cuda", dtype=BF16) for _ in range(2))
    dx = ops.layernorm_bwd(dy.cuda(), x.cuda(), gm.cuda(), stats, dg, db)
    spec = (0.1, 99, 21)
    dd = torch.empty_like(dx)
    part = torch.empty((ops._lib.load().s2t_layernorm_bwd_partials(rows, C), 2, C), device="cuda")
    dx2 = ops.layernorm_bwd(dy.cuda(), x.cuda(), gm.cuda(), stats, None, None, out_drop=dd, drop=spec, drop_scale=0.5,
                            partials=part, reduce=False)
    dg2, db2 = (torch.empty((C,), device="cuda", dtype=BF16) for _ in range(2))
    ops.layernorm_bwd_reduce(part, rows, C, dg2, db2)
    assert torch.equal(dx, dx2) and torch.equal(dg, dg2) and torch.equal(db, db2)
    want = ops.dropout(dx, spec, scale=0.5)
    assert (dd.float() - want.float()).abs().max() <= 2 ** -7 * want.float().abs().max()   # one extra bf16 rounding
    assert torch.equal(dd == 0, want == 0)


# ------------------------------------------------------------------------------------------------ elementwise
def test_glu_fwd_bwd_with_padding_mask(cuda_lib):
    from s2t_b200 import ops
    B, T, C = 3, 50, 128
    lens = torch.tensor([50, 31, 7], dtype=torch.int32)
    x, dy = _rand((B * T, 2 * C), 1), _rand((B * T, C), 2)
    y = ops.glu_fwd(x.cuda(), lengths=lens.cuda(), T=T)
    xf = x.float().requires_grad_(True)
    mask = (torch.arange(T)[None, :] < lens[:, None]).reshape(B * T, 1).float()
    ref = torch.nn.functional.glu(xf, dim=1) * mask
    assert (y.float().cpu() - ref).abs().max() < 2 ** -7 * ref.abs().max()
    ref.backward(dy.float())
    dx = ops.glu_bwd(x.cuda(), dy.cuda(), lengths=lens.cuda(), T=T)
    assert (dx.float().cpu() - xf.grad).abs().max() < 2 ** -7 * xf.grad.abs().max()
    y2 = ops.glu_fwd(x.cuda())
    assert (y2.float().cpu() - torch.nn.functional.glu(x.float(), dim=1)).abs().max() < 2 ** -7 * 10


@pytest.mark.parametrize("rows,C", [(2000, 256), (63, 2048), (1, 8), (1000, 10000)])
def test_colsum(cuda_lib, rows, C):
    from s2t_b200 import ops
    x = _rand((rows, C), 3)
    out = torch.zeros((C,), device="cuda", dtype=BF16)
    ops.colsum(x.cuda(), out)
    ref = x.float().sum(0)
    assert (out.float().cpu() - ref).abs().max() <= 2 ** -7 * ref.abs().max() + 1e-3
    ops.colsum(x.cuda(), out, accumulate=True)
    assert (out.float().cpu() - 2 * ref).abs().max() <= 2 ** -6 * 2 * ref.abs().max() + 1e-3


def test_col2im1d_is_conv1d_input_grad(cuda_lib):
    from s2t_b200 import ops
    B, T, C, k, s, p = 2, 37, 16, 5, 2, 2
    T_out = (T + 2 * p - k) // s + 1
    dcol = _rand((B, T_out, k * C), 4)
    dx = ops.col2im1d(dcol.cuda(), B, T, T_out, C, k, s, p)
    ref = torch.zeros((B, T, C))
    for to in range(T_out):
        for j in range(k):
            t = s * to + j - p
            if 0 <= t < T:
                ref[:, t] += dcol[:, to, j * C:(j + 1) * C].float()
    assert (dx.float().cpu() - ref).abs().max() <= 2 ** -7 * ref.abs().max()


def test_embed_fwd_bwd(cuda_lib):
    from s2t_b200 import ops
    B, T, C = 3, 40, 64
    lens = torch.tensor([40, 25, 3], dtype=torch.int32)
    x, dy = _rand((B, T, C), 5), _rand((B, T, C), 6)
    table = torch.randn((T + 2, C), generator=torch.Generator().manual_seed(7))
    table[1] = 0
    y = ops.embed_fwd(x.cuda(), lens.cuda(), table.cuda(), 2.0)
    valid = (torch.arange(T)[None, :] < lens[:, None])
    pos = torch.where(valid[..., None], table[torch.arange(T) + 2][None].expand(B, T, C), torch.zeros(()))
    ref = 2.0 * x.float() + pos
    assert (y.float().cpu() - ref).abs().max() <= 2 ** -7 * ref.abs().max()
    # dropout: regenerable mask, backward applies the same mask and scale
    drop = (0.25, 99, 3)
    yd = ops.embed_fwd(x.cuda(), lens.cuda(), table.cuda(), 2.0, dropout=drop).float().cpu()
    kept = yd != 0
    assert abs((~kept).float().mean().item() - 0.25) < 0.03
    assert torch.allclose(yd[kept], (ref / 0.75).to(BF16).float()[kept], rtol=2e-2, atol=1e-2)
    dx = ops.embed_bwd(dy.cuda(), 2.0, dropout=drop).float().cpu()
    want = torch.where(kept, 2.0 * dy.float() / 0.75, torch.zeros(()))
    assert (dx - want).abs().max() <= 2 ** -6 * want.abs().max()
    m = x.clone().cuda()
    ops.mask_rows_(m, lens.cuda())
    assert torch.equal(m.cpu(), torch.where(valid[..., None], x, torch.zeros((), dtype=BF16)))


# ------------------------------------------------------------------------------------------------ softmax
@pytest.mark.parametrize("T", [250, 33])
def test_attn_softmax_fwd_bwd(cuda_lib, T):
    from s2t_b200 import ops
    B, H = 2, 4
    ld = (T + 7) // 8 * 8
    klen = torch.tensor([T, T - 13], dtype=torch.int32)
    s = torch.zeros((B * H, T, ld), dtype=BF16)
    s[..., :T] = _rand((B * H, T, T), 1, 3.0)
    dp = torch.zeros((B * H, T, ld), dtype=BF16)
    dp[..., :T] = _rand((B * H, T, T), 2)
    probs, _ = ops.attn_softmax_fwd(s.cuda(), klen.cuda(), H, T)
    sf = s.float()[..., :T].view(B, H, T, T).requires_grad_(True)
    mask = torch.arange(T)[None, None, None, :] >= klen[:, None, None, None]
    ref = torch.softmax(sf.masked_fill(mask, float("-inf")), -1)
    got = probs.float().cpu().view(B, H, T, ld)
    assert (got[..., :T] - ref).abs().max() < 2 ** -8
    assert got[..., T:].abs().max() == 0
    (ref * dp.float()[..., :T].view(B, H, T, T)).sum().backward()
    # feed the kernel the exact bf16 probs it produced
    ds = ops.attn_softmax_bwd(probs, dp.cuda(), klen.cuda(), H, T)
    got_ds = ds.float().cpu().view(B, H, T, ld)[..., :T]
    assert (got_ds - sf.grad).abs().max() < 2 ** -6 * sf.grad.abs().max()


def test_attn_softmax_relshift_bias_and_dropout(cuda_lib):
    from s2t_b200 import ops
    B, H, T = 1, 2, 40
    ld, ldb = 40, 80
    klen = torch.tensor([T - 5], dtype=torch.int32)
    ac = _rand((B * H, T, ld), 3)
    bd = torch.zeros((B * H, T, ldb), dtype=BF16)
    bd[..., :2 * T - 1] = _rand((B * H, T, 2 * T - 1), 4)
    scale = 0.125
    probs, pd = ops.attn_softmax_fwd(ac.cuda(), klen.cuda(), H, T, bias=bd.cuda(), bias_mode=2, scale=scale,
                                     clamp=1e8, probs_drop=torch.empty((B * H, T, ld), device="cuda", dtype=BF16),
                                     dropout=(0.2, 5, 1))
    i = torch.arange(T)[:, None]
    j = torch.arange(T)[None, :]
    shifted = bd.float()[..., :2 * T - 1].gather(-1, (j - i + T - 1).expand(B * H, T, T))  # BD[i, j-i+T-1]
    sc = (ac.float() + shifted) * scale
    sc = sc.masked_fill(j[None] >= klen[0], float("-inf"))
    ref = torch.softmax(sc, -1)
    assert (probs.float().cpu() - ref).abs().max() < 2 ** -8
    pdc = pd.float().cpu()
    kept = pdc != 0
    live = ref > 1e-3
    assert abs(1 - kept[live].float().mean().item() - 0.2) < 0.05
    # backward with un-shift of the bias gradient
    dpd = _rand((B * H, T, ld), 5)
    dbias = torch.full((B * H, T, ldb), 7.0, device="cuda", dtype=BF16)
    ds = ops.attn_softmax_bwd(probs, dpd.cuda(), klen.cuda(), H, T, dbias=dbias, scale=scale,
                              dropout=(0.2, 5, 1)).float().cpu()
    p = probs.float().cpu()
    dpr = torch.where(kept, dpd.float() / 0.8, torch.zeros(()))
    dpr = torch.where(j[None] < klen[0], dpr, torch.zeros(()))
    want = scale * p * (dpr - (dpr * p).sum(-1, keepdim=True))
    assert (ds - want).abs().max() < 2 ** -6 * want.abs().max() + 1e-4
    db = dbias.float().cpu()
    want_db = torch.zeros((B * H, T, 2 * T - 1))
    want_db.scatter_(-1, (j - i + T - 1).expand(B * H, T, T), ds)
    assert torch.equal(db[..., :2 * T - 1], want_db)


# ------------------------------------------------------------------------------------------------ CTC
def _ctc_case(seed, B, T, V, tlens, ilens, dtype):
    rng = np.random.default_rng(seed)
    logits = torch.from_numpy(rng.standard_normal((B, T, V)).astype(np.float32) * 2.0).to(dtype)
    smax = max(max(tlens), 1)
    tg = np.full((B, smax), 1, dtype=np.int32)
    tgts = []
    for b, n in enumerate(tlens):
        row = rng.integers(1, V, size=n)
        if n >= 4:
            row[2] = row[1]  # repeated label needs a blank in between
        tg[b, :n] = row
        tgts.append(row)
    return logits, tg, tgts


@pytest.mark.parametrize("dtype", [torch.float32, BF16])
def test_ctc_loss_and_grad_match_oracle(cuda_lib, dtype):
    """CTC loss rel 1e-3 (north_star); logits-gradient max-abs 1e-4 (fp32) / 2^-8 (bf16 storage)."""
    from oracle import ctc as O
    from s2t_b200 import ops
    B, T, V = 5, 60, 104
    tlens, ilens = [12, 0, 29, 40, 7], [60, 50, 59, 30, 8]   # utt 3 infeasible (S > T) -> zero_infinity
    logits, tg, tgts = _ctc_case(0, B, T, V, tlens, ilens, dtype)
    nll, dl = ops.ctc_loss(logits.cuda(), torch.from_numpy(tg).cuda(),
                           torch.tensor(ilens, dtype=torch.int32).cuda(),
                           torch.tensor(tlens, dtype=torch.int32).cuda(), V)
    ref_nll, ref_grad = O.ctc_loss(logits.float().numpy(), tgts, ilens)
    nll = nll.cpu().numpy()
    assert ref_nll[3] == 0 and nll[3] == 0
    assert np.allclose(nll, ref_nll, rtol=1e-3, atol=1e-3), (nll, ref_nll)
    tol = 1e-4 if dtype == torch.float32 else 2 ** -8
    assert np.abs(dl.float().cpu().numpy() - ref_grad).max() < tol
    assert np.all(dl.float().cpu().numpy()[3] == 0)


@pytest.mark.parametrize("smax", [1, 14, 31, 63, 100, 127, 150])
def test_ctc_lattice_sizes(cuda_lib, smax):
    """Every lattice width class (warp-resident 1/2/4/8 states per lane, block fallback above 256 extended states),
    with empty / single-label / repeated-label targets, against the float64 oracle."""
    from oracle import ctc as O
    from s2t_b200 import ops
    B, V = 4, 56
    T = 2 * smax + 9
    tlens = [smax, max(smax - 1, 0), 0, max(smax // 2, 1)]
    ilens = [T, T - 1, 5, max(T // 2, 3)]
    logits, tg, tgts = _ctc_case(smax, B, T, V, tlens, ilens, torch.float32)
    if smax >= 3:
        tg[0, : smax] = 7                        # all-repeats target: needs a blank between every pair
        tgts[0] = tg[0, : smax].copy()
    nll, dl = ops.ctc_loss(logits.cuda(), torch.from_numpy(tg).cuda(), torch.tensor(ilens, dtype=torch.int32).cuda(),
                           torch.tensor(tlens, dtype=torch.int32).cuda(), V)
    ref_nll, ref_grad = O.ctc_loss(logits.float().numpy(), tgts, ilens)
    assert np.allclose(nll.cpu().numpy(), ref_nll, rtol=1e-3, atol=1e-3), (nll.cpu().numpy(), ref_nll)
    # fp32 log-domain lattice over up to 309 frames: |log alpha| reaches ~1e3, whose fp32 ulp is 6e-5, so occupancies
    # (values in [0, 1]) carry ~1e-3 absolute round-off against the float64 oracle (torch's fp32 CTC is the same)
    assert np.abs(dl.float().cpu().numpy() - ref_grad).max() < 3e-3


def test_ctc_grad_peaked_logits_single_rounding(cuda_lib):
    """A 'trained' model: at the label frames the softmax p and the occupancy are both near 1 and nearly cancel.  The
    bf16 gradient p - occ must be rounded ONCE from the fp32 difference (the reference: fp32 log_softmax + CTCLoss, one
    final cast): relative error of the label entries <= 2^-7, where a read-modify-write of the already rounded
    bf16(p) loses it (error 2^-9 * p against |p - occ| ~ 1e-3).  Also the device-side batch weight (dummy batch)."""
    from oracle import ctc as O
    from s2t_b200 import ops
    B, T, V, S = 3, 40, 64, 12
    rng = np.random.default_rng(5)
    tgts = [rng.integers(1, V, size=S) for _ in range(B)]
    for t in tgts:                                   # no repeats: a monotone alignment of 3 frames per label + blanks
        for i in range(1, S):
            if t[i] == t[i - 1]:
                t[i] = (t[i] % (V - 1)) + 1
    logits = rng.normal(0, 0.3, size=(B, T, V)).astype(np.float32)
    for b in range(B):
        for t in range(T):
            lab = tgts[b][min(t // 3, S - 1)] if t < 3 * S else 0
            logits[b, t, lab] += 7.0                # p(label) ~ 0.95-0.99
    x = torch.from_numpy(logits).to(BF16)
    tg = np.zeros((B, S), dtype=np.int32)
    for b in range(B):
        tg[b] = tgts[b]
    ilens, tlens = [T] * B, [S] * B
    args = (x.cuda(), torch.from_numpy(tg).cuda(), torch.tensor(ilens, dtype=torch.int32).cuda(),
            torch.tensor(tlens, dtype=torch.int32).cuda(), V)
    nll, dl = ops.ctc_loss(*args)
    ref_nll, ref_grad = O.ctc_loss(x.float().numpy(), tgts, ilens)
    assert np.allclose(nll.cpu().numpy(), ref_nll, rtol=1e-3, atol=1e-3)
    got = dl.float().cpu().numpy()
    checked = 0
    for b in range(B):
        for t in range(3 * S):
            lab = tgts[b][t // 3]
            r = ref_grad[b, t, lab]
            if abs(r) > 1e-3:
                assert abs(got[b, t, lab] - r) <= 2 ** -7 * abs(r) + 2e-6, (b, t, got[b, t, lab], r)
                checked += 1
    assert checked > 10
    # batch weight 0 on the device: gradient exactly zero, loss values unchanged
    wz = torch.zeros(1, device="cuda")
    nll0, dl0 = ops.ctc_loss(*args, grad_scale_dev=wz)
    assert torch.count_nonzero(dl0) == 0 and torch.equal(nll0, nll)
    w1 = torch.ones(1, device="cuda")
    _, dl1 = ops.ctc_loss(*args, grad_scale_dev=w1)
    assert torch.equal(dl1, dl)


def test_ctc_loss_c1_shape(cuda_lib):
    """BASELINE config-1 shape: B=8, T'=250, V=10000, S=60 — loss rel 1e-3 vs oracle on bf16 logits."""
    from oracle import ctc as O
    from s2t_b200 import ops
    B, T, V = 8, 250, 10000
    tlens, ilens = [60] * 8, [250, 250, 240, 230, 200, 190, 180, 150]
    logits, tg, tgts = _ctc_case(1, B, T, V, tlens, ilens, BF16)
    nll, dl = ops.ctc_loss(logits.cuda(), torch.from_numpy(tg).cuda(),
                           torch.tensor(ilens, dtype=torch.int32).cuda(),
                           torch.tensor(tlens, dtype=torch.int32).cuda(), V, grad_scale=0.5)
    ref_nll, ref_grad = O.ctc_loss(logits.float().numpy(), tgts, ilens)
    assert np.allclose(nll.cpu().numpy(), ref_nll, rtol=1e-3)
    assert np.abs(dl.float().cpu().numpy() - 0.5 * ref_grad).max() < 2 ** -8
    # size-independent property: every live frame's gradient sums to ~0 (softmax - occupancy)
    rowsum = dl.float().sum(-1).cpu().numpy()
    assert np.abs(rowsum).max() < 0.05


def test_ctc_greedy_bit_exact(cuda_lib):
    from oracle import ctc as O
    from s2t_b200 import ops
    B, T, V = 6, 97, 1000
    rng = np.random.default_rng(3)
    logits = rng.standard_normal((B, T, V)).astype(np.float32)
    # long runs, blanks and crafted exact ties
    for b in range(B):
        for t in range(T):
            logits[b, t, [0, 5, 5, 9, 0, 700][(t // 3 + b) % 6]] += 6.0
    logits[0, 10, :] = 0.0            # full tie -> index 0 (blank)
    logits[1, 11, 17] = logits[1, 11, 400] = 50.0  # two-way tie -> 17
    ilens = [97, 96, 50, 1, 33, 64]
    for dtype in (torch.float32, BF16):
        lg = torch.from_numpy(logits).to(dtype)
        tokens, lens, frame_ids = ops.ctc_greedy(lg.cuda(), torch.tensor(ilens, dtype=torch.int32).cuda(), V)
        ref = O.greedy_decode(lg.float().numpy(), ilens)
        tokens, lens = tokens.cpu().numpy(), lens.cpu().numpy()
        for b in range(B):
            assert lens[b] == len(ref[b])
            assert np.array_equal(tokens[b, :lens[b]], ref[b])
            assert np.all(tokens[b, lens[b]:] == -1)


# --------------------------------------------------------------------------------------------------
# Conv2dSubsampling support kernels (bit-exact data movement; fp32 partial sums)
# --------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("cin,kpad", [(1, 16), (8, 72), (24, 216)])
def test_im2col2d_and_col2im2d(cuda_lib, cin, kpad):
    """Patch matrix of Conv2d(k=3, stride=2, no padding) over channels-last x[B][T][D][C] against torch unfold
    (bit-exact copies), and col2im2d as its exact transpose (<col, im2col(x)> == <col2im(col), x>)."""
    from s2t_b200 import ops
    g = torch.Generator().manual_seed(cin)
    B, T, D, k, st = 2, 21, 14, 3, 2
    x = torch.randn((B, T, D, cin), generator=g).to(BF16)
    col = ops.im2col2d(x.cuda(), k, st, kpad).cpu()
    T2, D2 = (T - k) // st + 1, (D - k) // st + 1
    # reference: unfold over (T, D) with channels first -> [B, C*k*k, T2*D2], kernel index order (c, j, i)
    u = torch.nn.functional.unfold(x.float().permute(0, 3, 1, 2), kernel_size=k, stride=st)
    u = u.view(B, cin, k, k, T2, D2).permute(0, 4, 5, 2, 3, 1).reshape(B * T2 * D2, k * k * cin)   # (j, i, c) columns
    assert torch.equal(col[:, : k * k * cin].float(), u)
    assert torch.all(col[:, k * k * cin:] == 0)
    if cin % 8 == 0:
        dcol = torch.randn((B * T2 * D2, kpad), generator=g).to(BF16)
        dx = ops.col2im2d(dcol.cuda(), B, T, D, cin, k, st).float().cpu()
        lhs = (dcol.float()[:, : k * k * cin] * u).sum()
        rhs = (dx * x.float()).sum()
        assert abs(float(lhs - rhs)) <= 2e-2 * abs(float(lhs)) + 1e-2


def test_colstats_feed_batchnorm(cuda_lib):
    """colstats partials -> batchnorm_fwd == BatchNorm over the rows (training mode, batch statistics)."""
    from s2t_b200 import ops
    g = torch.Generator().manual_seed(5)
    rows, C = 5000, 176
    x = (torch.randn((rows, C), generator=g) * 2 + 0.5).to(BF16)
    gamma, beta = torch.rand(C, generator=g).to(BF16), torch.randn(C, generator=g).to(BF16)
    part = ops.colstats(x.cuda())
    s = part.sum(0).cpu()
    assert torch.allclose(s[0], x.float().sum(0), rtol=1e-4, atol=1e-2)
    assert torch.allclose(s[1], (x.float() ** 2).sum(0), rtol=1e-4, atol=1e-2)
    rm, rv = torch.zeros(C, device="cuda"), torch.ones(C, device="cuda")
    nbt = torch.zeros((), device="cuda", dtype=torch.int64)
    save = torch.empty((2, C), device="cuda")
    y = ops.batchnorm_fwd(x.cuda(), gamma.cuda(), beta.cuda(), partials=part, running_mean=rm, running_var=rv,
                          num_batches_tracked=nbt, save_stats=save, act=ops.ACT_SWISH, training=True)
    bn = torch.nn.functional.batch_norm(x.float(), None, None, gamma.float(), beta.float(), training=True)
    ref = bn * torch.sigmoid(bn)
    assert (y.float().cpu() - ref).abs().max() <= 3e-2


# --------------------------------------------------------------------------------------------------
# fused self-attention forward (attn.cu) against the three-launch form it replaces
# --------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("T,p", [(250, 0.0), (250, 0.1), (77, 0.1), (256, 0.0)])
def test_attn_fused_fwd_matches_unfused(cuda_lib, T, p):
    """S = q k^T / sqrt(d_k) -> key-padding mask -> fp32 softmax -> dropout -> P v.  Same Philox numbering as the
    softmax kernel: the dropped positions must coincide exactly; values within bf16 rounding of the unfused path
    (which rounds the scores to bf16 before the softmax) and of a float reference."""
    from s2t_b200 import ops
    g = torch.Generator().manual_seed(T)
    B, H, dk = 3, 4, 64
    C = H * dk
    qkv = (torch.randn((B * T, 3 * C), generator=g) * 0.7).to(BF16).cuda()
    klen = torch.tensor([T, max(T - 40, 1), max(T // 2, 1)], dtype=torch.int32).cuda()
    ld = (T + 7) // 8 * 8
    spec = (p, 1234, 5, None) if p > 0 else None
    probs = torch.zeros((B * H, T, ld), dtype=BF16, device="cuda")
    probs_d = torch.zeros_like(probs) if p > 0 else None
    ctx = torch.zeros((B * T, C), dtype=BF16, device="cuda")
    ops.attn_fused_fwd(qkv, C, klen, B, H, T, dk, probs=probs, probs_drop=probs_d, ctx=ctx, scale=dk ** -0.5, dropout=spec)
    # unfused: scores GEMM -> softmax kernel -> PV GEMM (the engine's selfattn path)
    scores = torch.zeros_like(probs)
    bs = (T * ld, H * T * ld)
    ops.gemm_raw(M=T, N=T, K=dk, a=qkv, lda=3 * C, a_bs=(dk, T * 3 * C), b=qkv, ldb=3 * C, b_off=C, b_bs=(dk, T * 3 * C),
                 out=scores, ldc=ld, c_bs=bs, batch=(H, B), alpha=dk ** -0.5)
    r_probs = torch.zeros_like(probs)
    r_probs_d = torch.zeros_like(probs) if p > 0 else None
    ops.attn_softmax_fwd(scores, klen, H, T, probs=r_probs, probs_drop=r_probs_d, dropout=spec)
    r_ctx = torch.zeros_like(ctx)
    pd = r_probs_d if p > 0 else r_probs
    ops.gemm_raw(M=T, N=dk, K=T, a=pd, lda=ld, a_bs=bs, b=qkv, ldb=3 * C, b_off=2 * C, b_mn=True, b_bs=(dk, T * 3 * C),
                 out=r_ctx, ldc=C, c_bs=(dk, T * C), batch=(H, B))
    torch.cuda.synchronize()
    assert (probs.float() - r_probs.float()).abs().max() <= 2e-2
    if p > 0:
        assert torch.equal(probs_d == 0, r_probs_d == 0) or \
            ((probs_d == 0) != (r_probs_d == 0)).float().mean() < 1e-4      # identical masks (up to p == 0 underflow)
        assert (probs_d.float() - r_probs_d.float()).abs().max() <= 3e-2
    assert rel_l2(ctx, r_ctx) <= 2e-2
    # float reference (no dropout): softmax rows sum to one over the valid keys, zero elsewhere
    row_sums = probs.float().view(B, H, T, ld).sum(-1)
    assert (row_sums - 1).abs().max() <= 2e-2
    for b in range(B):
        tail = probs.float().view(B, H, T, ld)[b, :, :, int(klen[b]):]
        assert tail.numel() == 0 or float(tail.abs().max()) == 0.0


@pytest.mark.parametrize("T,p", [(250, 0.1), (256, 0.0), (100, 0.1), (129, 0.0), (37, 0.1)])
def test_attn_fused_fwd_lse_and_fused_bwd_match_fp32_reference(cuda_lib, T, p):
    """Fused attention forward (row log-sum-exp kept instead of the probabilities) and fused backward (one 2-CTA cluster
    per (utterance, head): P recomputed from lse, dQ / dK / dV) against a plain PyTorch fp32 restatement of fairseq
    MultiheadAttention's core (multihead_attention.py:372-416) using the SAME dropout mask (read off the kernel's own
    probs / probs_drop outputs).  ctx / dq / dk / dv rel-L2 <= 2e-2 (bf16 operands), lse abs 2e-3."""
    from s2t_b200 import ops
    B, H, dk = 3, 4, 64
    C = H * dk
    g = torch.Generator().manual_seed(T)
    qkv = (torch.randn((B * T, 3 * C), generator=g) * 0.7).to(BF16).cuda()
    lens = torch.tensor([T, max(1, T - 17), max(1, T // 3)], dtype=torch.int32).cuda()
    ld = (T + 7) // 8 * 8
    spec = (p, 1234, 77, None) if p > 0 else None
    probs = torch.zeros((B * H, T, ld), device="cuda", dtype=BF16)
    probs_d = torch.zeros_like(probs)
    lse = torch.zeros((B * H * T,), device="cuda", dtype=torch.float32)
    ctx = torch.zeros((B * T, C), device="cuda", dtype=BF16)
    scale = dk ** -0.5
    ops.attn_fused_fwd(qkv, C, lens, B, H, T, dk, probs=probs, probs_drop=probs_d if p > 0 else None, lse=lse, ctx=ctx,
                       scale=scale, dropout=spec)
    # the production call: no probabilities written, same ctx and lse
    ctx2, lse2 = torch.zeros_like(ctx), torch.zeros_like(lse)
    ops.attn_fused_fwd(qkv, C, lens, B, H, T, dk, lse=lse2, ctx=ctx2, scale=scale, dropout=spec)
    torch.cuda.synchronize()
    assert torch.equal(ctx, ctx2) and torch.equal(lse, lse2)
    # fp32 reference
    x = qkv.float().view(B, T, 3, H, dk).permute(2, 0, 3, 1, 4)          # (3, B, H, T, dk)
    q, k, v = x[0], x[1], x[2]
    s = (q @ k.transpose(-1, -2)) * scale
    kmask = torch.arange(T, device="cuda")[None, :] >= lens[:, None]      # (B, T) True = padded key
    s = s.masked_fill(kmask[:, None, None, :], float("-inf"))
    P = torch.softmax(s, dim=-1)
    ref_lse = torch.logsumexp(s, dim=-1)
    assert (lse.view(B, H, T) - ref_lse).abs().max() <= 2e-3
    pr, pd = probs.float().view(B, H, T, ld)[..., :T], probs_d.float().view(B, H, T, ld)[..., :T]
    if p > 0:
        M = torch.where(pr > 0, (pd != 0).float() / (1 - p), torch.full_like(pr, 1 / (1 - p)))
        keep = float((pd[pr > 0] != 0).float().mean())
        assert abs(keep - (1 - p)) < 0.02
    else:
        M = torch.ones_like(pr)
    Pd = P * M
    ref_ctx = (Pd @ v).permute(0, 2, 1, 3).reshape(B * T, C)
    rel = lambda a, b: float((a.float() - b).norm() / (b.norm() + 1e-12))
    assert rel(ctx, ref_ctx) <= 2e-2
    d_o = (torch.randn((B * T, C), generator=g) * 0.5).to(BF16).cuda()
    go = d_o.float().view(B, T, H, dk).permute(0, 2, 1, 3)               # (B, H, T, dk)
    dV = Pd.transpose(-1, -2) @ go
    dP = (go @ v.transpose(-1, -2)) * M
    dS = P * (dP - (dP * P).sum(-1, keepdim=True))
    dQ = (dS @ k) * scale
    dK = (dS.transpose(-1, -2) @ q) * scale
    out = torch.full((B * T, 3 * C), float("nan"), device="cuda", dtype=BF16)
    ops.attn_fused_bwd(qkv, C, ctx, d_o, lse, lens, B, H, T, dk, out=out, scale=scale, dropout=spec)
    torch.cuda.synchronize()
    assert torch.isfinite(out.float()).all()
    got = out.float().view(B, T, 3, H, dk).permute(2, 0, 3, 1, 4)
    for name, mine, ref in (("dq", got[0], dQ), ("dk", got[1], dK), ("dv", got[2], dV)):
        assert rel(mine, ref) <= 2e-2, (name, rel(mine, ref))
    # padded keys receive no gradient
    assert float(got[1][2, :, int(lens[2]):].abs().max()) == 0.0 and float(got[2][2, :, int(lens[2]):].abs().max()) == 0.0


def test_layernorm_bwd_reduce_batched(cuda_lib):
    """One launch reduces the dgamma / dbeta partials of many LayerNorm instances (different widths, row counts,
    accumulate flags) — equal to the per-instance reduction."""
    from s2t_b200 import ops
    g = torch.Generator().manual_seed(3)
    jobs, want = [], []
    for i, (rows, C) in enumerate([(2000, 256), (998, 512), (7, 8), (4001, 256)] * 10):          # 40 jobs: two launches
        x = torch.randn((rows, C), generator=g).to(BF16).cuda()
        dy = torch.randn((rows, C), generator=g).to(BF16).cuda()
        gamma = torch.rand(C, generator=g).to(BF16).cuda()
        stats = torch.stack([x.float().mean(-1), 1.0 / torch.sqrt(x.float().var(-1, unbiased=False) + 1e-5)], dim=1).contiguous()
        n = ops._lib.load().s2t_layernorm_bwd_partials(rows, C)
        part = torch.empty((n, 2, C), device="cuda", dtype=torch.float32)
        ops.layernorm_bwd(dy, x, gamma, stats, None, None, partials=part, reduce=False)
        acc = i % 3 == 0
        dg0, db0 = torch.randn(C, generator=g).to(BF16).cuda(), torch.randn(C, generator=g).to(BF16).cuda()
        dg1, db1 = dg0.clone(), db0.clone()
        ops.layernorm_bwd_reduce(part, rows, C, dg0, db0, accumulate=acc)
        jobs.append((part, rows, C, dg1, db1, acc))
        want.append((dg0, db0))
    ops.layernorm_bwd_reduce_batched(jobs)
    torch.cuda.synchronize()
    for (_, _, _, dg1, db1, _), (dg0, db0) in zip(jobs, want):
        assert torch.equal(dg1, dg0) and torch.equal(db1, db0)


def test_masked_ctc_loss_matches_oracle(cuda_lib):
    """s2t_ctc_loss_masked (torch_imputer.imputer_loss, criterions/ctc.py:455-465) against oracle/ctc.py imputer_loss
    (brute-force checked, tests/test_imputer_cpu.py): loss rel 1e-3, gradient max-abs 1e-4 (fp32) / 2^-8 (bf16); states
    forced along the Viterbi alignment, one utterance with an impossible forced state (-> zero_infinity), one unforced."""
    from pathlib import Path
    _GOLDEN = Path(__file__).resolve().parent / "golden"
    from oracle import ctc as O
    from oracle import viterbi as OV
    from s2t_b200 import ops
    rng = np.random.default_rng(0)
    B, T, V = 5, 40, 56
    tlens, ilens = [8, 0, 12, 5, 7], [40, 30, 39, 12, 25]
    for dtype, tol in ((torch.float32, 1e-4), (torch.bfloat16, 2 ** -8)):
        logits = torch.from_numpy(rng.normal(0, 1.5, size=(B, T, V)).astype(np.float32)).to(dtype)
        smax = max(tlens)
        tg = np.zeros((B, smax), dtype=np.int32)
        tgts = []
        for b, n in enumerate(tlens):
            row = rng.integers(1, V, size=n)
            if n >= 3:
                row[2] = row[1]
            tg[b, :n] = row
            tgts.append(row)
        lp = O.log_softmax(logits.float().numpy())
        paths, _ = OV.best_alignment(lp.astype(np.float32), tg.astype(np.int64), ilens, tlens)
        fe = np.full((B, T), -1, dtype=np.int32)
        for b in (0, 2, 3):
            idx = rng.choice(ilens[b], size=ilens[b] // 3, replace=False)
            fe[b, idx] = np.asarray(paths[b])[idx]
        fe[3, 0] = 2 * tlens[3]                      # the last blank cannot be occupied at frame 0: infeasible
        nll, dl = ops.ctc_loss(logits.cuda(), torch.from_numpy(tg).cuda(), torch.tensor(ilens, dtype=torch.int32).cuda(),
                               torch.tensor(tlens, dtype=torch.int32).cuda(), V, force_emit=torch.from_numpy(fe).cuda())
        ref_nll, ref_grad = O.imputer_loss(logits.float().numpy(), tgts, fe, ilens)
        nll = nll.cpu().numpy()
        assert ref_nll[3] == 0 and nll[3] == 0, (ref_nll, nll)
        assert np.allclose(nll, ref_nll, rtol=1e-3, atol=1e-3), (nll, ref_nll)
        err = np.abs(dl.float().cpu().numpy() - ref_grad).max()
        assert err < tol, (str(dtype), err)
        # nothing forced == the plain loss (bit for bit: same gather / dense / occupancy kernels, other lattice kernel)
        free = torch.full((B, T), -1, dtype=torch.int32).cuda()
        nll_f, dl_f = ops.ctc_loss(logits.cuda(), torch.from_numpy(tg).cuda(), torch.tensor(ilens, dtype=torch.int32).cuda(),
                                   torch.tensor(tlens, dtype=torch.int32).cuda(), V, force_emit=free)
        nll_p, dl_p = ops.ctc_loss(logits.cuda(), torch.from_numpy(tg).cuda(), torch.tensor(ilens, dtype=torch.int32).cuda(),
                                   torch.tensor(tlens, dtype=torch.int32).cuda(), V)
        assert np.allclose(nll_f.cpu().numpy(), nll_p.cpu().numpy(), rtol=1e-4, atol=1e-4)
        assert (dl_f.float() - dl_p.float()).abs().max().item() < tol


def test_device_collate_is_bit_exact(cuda_lib):
    """s2t_collate_frames (speech_to_text_dataset.py:267-285, 411-423 on the device) against the reference's own outputs
    (tests/golden/collate.npz): packed utterances -> zero-padded, length-sorted batch, bit-exact; also into a larger plan
    buffer (row stride > T * F) and with T larger than the longest utterance (frame bucket)."""
    from pathlib import Path
    _GOLDEN = Path(__file__).resolve().parent / "golden"
    from s2t_b200 import collate as PC
    fx = np.load(_GOLDEN / "collate.npz")
    for name in ("feat80", "feat83", "wave", "single"):
        F = int(fx[f"{name}::F"])
        lens = fx[f"{name}::lengths"].tolist()
        packed = torch.from_numpy(fx[f"{name}::packed"])
        frames, o = [], 0
        for n in lens:
            frames.append(packed[o:o + n * F].view(n, F) if F > 1 else packed[o:o + n])
            o += n * F
        pk, offsets, lengths, order = PC.pack_frames(frames)
        want = torch.from_numpy(fx[f"{name}::collated"])
        B, T = want.shape[0], want.shape[1]
        out, lo = PC.collate_frames(pk.cuda(), offsets.cuda(), lengths.cuda(), order.cuda(), T, F)
        assert torch.equal(out.cpu().view(want.shape), want), name
        assert lo.cpu().tolist() == sorted(lens, reverse=True), name
        big = torch.full((B, (T + 5) * F + 4), 7.0, device="cuda")
        PC.collate_frames(pk.cuda(), offsets.cuda(), lengths.cuda(), order.cuda(), T + 5, F, out=big)
        got = big.cpu()[:, : (T + 5) * F].view(B, T + 5, F)
        assert torch.equal(got[:, :T].reshape(want.shape), want) and float(got[:, T:].abs().max()) == 0.0, name
        assert float((big.cpu()[:, (T + 5) * F:] - 7.0).abs().max()) == 0.0      # beyond the row: untouched


def test_fused_adam_matches_reference_golden(cuda_lib):
    """s2t_grad_norm + s2t_fused_adam (FusedAdamB200) against three updates of the reference's own Adam + clip_grad_norm_
    (tests/golden/fused_adam.npz): gradient norm rel 1e-5, fp32 master weights and both moments max-abs error <= 1e-5 of the
    largest element, bf16 parameters = bf16(master); with clipping, weight decay and the gradient multiplier."""
    from pathlib import Path
    _GOLDEN = Path(__file__).resolve().parent / "golden"
    from s2t_b200.optim import FusedAdamB200
    fx = np.load(_GOLDEN / "fused_adam.npz")
    for name in ("recipe", "clipped_wd", "no_clip"):
        lr, b1, b2, eps, wd, mult, max_norm = (float(v) for v in fx[f"{name}::hyper"])
        p = torch.from_numpy(fx[f"{name}::p0"]).to(torch.bfloat16).cuda()
        g = torch.zeros_like(p)
        opt = FusedAdamB200(p, g, lr=lr, betas=(b1, b2), eps=eps, weight_decay=wd)
        for step in (1, 2, 3):
            g.copy_(torch.from_numpy(fx[f"{name}::grads"][step - 1]).to(torch.bfloat16))
            tn = opt.step(grad_mult=mult, max_norm=max_norm)
            torch.cuda.synchronize()
            want = fx[f"{name}::norms"][step - 1]
            assert abs(float(tn.item()) - want) <= 1e-5 * want, (name, step, float(tn.item()), want)
            ref = torch.from_numpy(fx[f"{name}::master{step}"])
            err = float((opt.master.cpu() - ref).abs().max() / ref.abs().max())
            assert err <= 1e-5, (name, step, err)
            assert torch.equal(p.cpu(), opt.master.cpu().to(torch.bfloat16))
        # moments: error relative to the largest moment, like the master weights.  (An element-wise rtol of 1e-5 is not a
        # meaningful bound: m = b1 m + (1 - b1) g cancels over the three steps, and the clip coefficient carries the
        # norm's own 1e-5; this check failed on hardware in round 2 for that reason, not because of the kernel.)
        for mine, key in ((opt.exp_avg, "exp_avg"), (opt.exp_avg_sq, "exp_avg_sq")):
            ref = torch.from_numpy(fx[f"{name}::{key}"])
            err = float((mine.cpu() - ref).abs().max() / ref.abs().max())
            assert err <= 1e-5, (name, key, err)
    # device-side lr / step scalars give the same update as host scalars
    p1 = torch.from_numpy(fx["no_clip::p0"]).to(torch.bfloat16).cuda(); p2 = p1.clone()
    g = torch.from_numpy(fx["no_clip::grads"][0]).to(torch.bfloat16).cuda()
    o1, o2 = FusedAdamB200(p1, g, lr=5e-4), FusedAdamB200(p2, g, lr=123.0)
    o1.step()
    o2.step(lr_dev=torch.tensor([5e-4], device="cuda"), step_dev=torch.tensor([1], device="cuda", dtype=torch.int64))
    torch.cuda.synchronize()
    assert float((o1.master - o2.master).abs().max()) <= 1e-7


def test_fused_adam_skips_overflowed_and_step_zero_updates(cuda_lib):
    """The reference trainer raises FloatingPointError on a NaN / Inf gradient norm and never reaches optimizer.step
    (fairseq/trainer.py:751-753).  On the device the same update must leave master weights, moments and bf16 parameters
    bit-identical, check_finite() must raise and take the step count back; a device step counter of 0 is also skipped (the
    bias correction would be 0); the next clean update then equals the update of an optimizer that never saw the bad
    ones."""
    from s2t_b200.optim import FusedAdamB200
    g0 = torch.Generator().manual_seed(3)
    n = 4099                                                # not a multiple of 8: scalar tail of the norm kernel
    p0 = torch.randn(n, generator=g0).to(BF16)
    good = (torch.randn(n, generator=g0) * 0.1).to(BF16)
    for bad_value in (float("nan"), float("inf")):
        pa, pb = p0.clone().cuda(), p0.clone().cuda()
        ga, gb = good.clone().cuda(), good.clone().cuda()
        a, b = FusedAdamB200(pa, ga, lr=1e-2, weight_decay=1e-2), FusedAdamB200(pb, gb, lr=1e-2, weight_decay=1e-2)
        a.step(max_norm=1.0); b.step(max_norm=1.0)
        before = [t.clone() for t in (b.master, b.exp_avg, b.exp_avg_sq, pb)]
        gb[17] = bad_value
        b.step(max_norm=1.0)
        with pytest.raises(FloatingPointError):
            b.check_finite()
        assert b.step_count == 1
        for t, ref in zip((b.master, b.exp_avg, b.exp_avg_sq, pb), before):
            assert torch.equal(t, ref)
        gb.copy_(ga)
        b.step(max_norm=1.0, step_dev=torch.zeros(1, dtype=torch.int64, device="cuda"))      # clean gradient, device step 0
        b.step_count -= 1
        torch.cuda.synchronize()
        for t, ref in zip((b.master, b.exp_avg, b.exp_avg_sq, pb), before):
            assert torch.equal(t, ref)
        a.step(max_norm=1.0); b.step(max_norm=1.0)
        assert a.check_finite() == b.check_finite() and a.step_count == b.step_count == 2
        assert torch.equal(a.master, b.master) and torch.equal(pa, pb) and torch.equal(a.exp_avg_sq, b.exp_avg_sq)


def test_fused_adam_state_in_reference_format(cuda_lib):
    """FusedAdamB200.reference_state_dict / load_reference_state_dict on a real engine (Conv1d weights permuted, padded
    vocabulary bias): the exported per-parameter moments have the reference's shapes, a torch Optimizer over parameters of
    those shapes loads them, and loading them back reproduces the flat moments bit for bit on every parameter element."""
    from s2t_b200.engine import EncoderConfig, S2TEncoderEngine
    from s2t_b200.optim import FusedAdamB200
    eng = S2TEncoderEngine(EncoderConfig(layers=1, ffn_dim=256, subsampling_filter=128, vocab=90, dropout=0.0))
    S = eng.store
    S.grad.copy_((torch.randn(S.total, generator=torch.Generator().manual_seed(1)) * 0.05).to(BF16))
    opt = FusedAdamB200(S.flat, S.grad, lr=1e-3)
    opt.step(); opt.step()
    torch.cuda.synchronize()
    sd = opt.reference_state_dict(eng)
    ref_sd = eng.state_dict()
    names = S.names()
    assert len(sd["state"]) == len(names) and sd["param_groups"][0]["params"] == list(range(len(names)))
    for i, nme in enumerate(names):
        assert sd["state"][i]["exp_avg"].shape == ref_sd[nme].shape, nme
        assert sd["state"][i]["step"] == 2
    params = [torch.nn.Parameter(ref_sd[nme].clone()) for nme in names]
    torch.optim.Adam(params).load_state_dict(sd)
    other = FusedAdamB200(S.flat, S.grad, lr=1e-3)
    other.load_reference_state_dict(eng, sd)
    assert other.step_count == 2
    mine, theirs = eng.export_flat(opt.exp_avg_sq), eng.export_flat(other.exp_avg_sq)
    assert all(torch.equal(mine[nme], theirs[nme]) for nme in names)
    mine, theirs = eng.export_flat(opt.exp_avg), eng.export_flat(other.exp_avg)
    assert all(torch.equal(mine[nme], theirs[nme]) for nme in names)

