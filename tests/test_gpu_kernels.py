"""GPU numerics: every native kernel against a plain PyTorch fp32 restatement of the same op."""
import math

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

DT = [torch.float16, torch.bfloat16]


def tol(dt):
    return dict(rtol=2e-2, atol=2e-2) if dt == torch.bfloat16 else dict(rtol=4e-3, atol=4e-3)


@pytest.fixture(scope='module')
def nat():
    from hetseq_b200.ops import lib, native
    lib.load()
    return native


def gelu_ref(x):
    return x * 0.5 * (1.0 + torch.erf(x / 1.41421))


@pytest.mark.parametrize('dt', DT)
@pytest.mark.parametrize('shape', [(256, 512, 256), (200, 264, 136), (4096, 1024, 1024)])
def test_gemm_forward_bias_gelu(nat, dt, shape):
    M, N, K = shape
    torch.manual_seed(0)
    x = (torch.randn(M, K, device='cuda') * 0.5).to(dt)
    w = (torch.randn(N, K, device='cuda') * 0.05).to(dt)
    b = torch.randn(N, device='cuda').to(dt)
    y = torch.empty(M, N, dtype=dt, device='cuda')
    pre = torch.empty(M, N, dtype=dt, device='cuda')
    nat.gemm(x, w, y, M, N, K, lda=K, ldb=K, ldc=N, bias=b, act=nat.ACT_GELU, aux_out=pre, ldaux=N)
    ref_pre = x.float() @ w.float().t() + b.float()
    torch.testing.assert_close(pre.float(), ref_pre, **tol(dt))
    torch.testing.assert_close(y.float(), gelu_ref(ref_pre), **tol(dt))


@pytest.mark.parametrize('dt', DT)
def test_gemm_dgrad_wgrad(nat, dt):
    M, N, K = 512, 384, 640
    torch.manual_seed(1)
    x = (torch.randn(M, K, device='cuda') * 0.5).to(dt)
    w = (torch.randn(N, K, device='cuda') * 0.05).to(dt)
    dy = (torch.randn(M, N, device='cuda') * 0.1).to(dt)
    dx = torch.empty(M, K, dtype=dt, device='cuda')
    nat.gemm(dy, w, dx, M, K, N, lda=N, ldb=K, ldc=K, b_mn=True)
    torch.testing.assert_close(dx.float(), dy.float() @ w.float(), **tol(dt))
    dw = torch.full((N, K), 0.25, dtype=torch.float32, device='cuda')
    nat.gemm(dy, x, dw, N, K, M, lda=N, ldb=K, ldc=K, a_mn=True, b_mn=True, accumulate=True)
    torch.testing.assert_close(dw, dy.float().t() @ x.float() + 0.25, rtol=2e-3, atol=2e-3)
    g = torch.tensor([3.0], device='cuda')
    nat.gemm(dy, x, dw, N, K, M, lda=N, ldb=K, ldc=K, a_mn=True, b_mn=True, alpha_dev=g)
    torch.testing.assert_close(dw, 3 * (dy.float().t() @ x.float()), rtol=2e-3, atol=4e-3)


@pytest.mark.parametrize('dt', DT)
def test_gemm_dgelu_epilogue(nat, dt):
    M, N, K = 256, 512, 384
    torch.manual_seed(2)
    df = (torch.randn(M, K, device='cuda') * 0.1).to(dt)
    w2 = (torch.randn(K, N, device='cuda') * 0.05).to(dt)      # [H_out, I] row-major: dPre = dF @ W2
    pre = torch.randn(M, N, device='cuda').to(dt)
    out = torch.empty(M, N, dtype=dt, device='cuda')
    nat.gemm(df, w2, out, M, N, K, lda=K, ldb=N, ldc=N, b_mn=True, act=nat.ACT_DGELU_MUL, aux_in=pre, ldaux=N)
    p = pre.float().requires_grad_(True)
    gelu_ref(p).backward(df.float() @ w2.float())
    torch.testing.assert_close(out.float(), p.grad, **tol(dt))


@pytest.mark.parametrize('dt', DT)
@pytest.mark.parametrize('shape', [(256, 512, 384), (200, 264, 136), (4096, 1024, 1024)])
def test_gemm_add_aux_and_colsum_epilogues(nat, dt, shape):
    """dgrad that also adds the residual-path gradient (ACT_ADD_AUX) and the bias gradient collected in the
    epilogue (colsum_out), incl. ragged edges and the dGELU epilogue."""
    M, N, K = shape
    torch.manual_seed(21)
    dy = (torch.randn(M, K, device='cuda') * 0.1).to(dt)
    w = (torch.randn(K, N, device='cuda') * 0.05).to(dt)
    add = torch.randn(M, N, device='cuda').to(dt)
    out = torch.empty(M, N, dtype=dt, device='cuda')
    cs = torch.full((N,), 2.0, device='cuda')
    nat.gemm(dy, w, out, M, N, K, lda=K, ldb=N, ldc=N, b_mn=True, act=nat.ACT_ADD_AUX, aux_in=add, ldaux=N,
             colsum_out=cs, colsum_accumulate=True)
    ref = dy.float() @ w.float() + add.float()
    torch.testing.assert_close(out.float(), ref, **tol(dt))
    torch.testing.assert_close(cs, ref.sum(0) + 2.0, rtol=2e-2, atol=0.3)
    pre = torch.randn(M, N, device='cuda').to(dt)
    cs2 = torch.full((N,), 7.0, device='cuda')     # overwritten, not accumulated
    nat.gemm(dy, w, out, M, N, K, lda=K, ldb=N, ldc=N, b_mn=True, act=nat.ACT_DGELU_MUL, aux_in=pre, ldaux=N,
             colsum_out=cs2, colsum_accumulate=False)
    pr = pre.float().requires_grad_(True)
    gelu_ref(pr).backward(dy.float() @ w.float())
    torch.testing.assert_close(out.float(), pr.grad, **tol(dt))
    torch.testing.assert_close(cs2, pr.grad.sum(0), rtol=2e-2, atol=5e-2)


@pytest.mark.parametrize('dt', DT)
@pytest.mark.parametrize('shape', [(256, 512, 384), (200, 264, 136), (4096, 4096, 1024)])
def test_gemm_gelu_with_derivative_then_mul(nat, dt, shape):
    """Forward epilogue that stores gelu'(pre) (ACT_GELU_DAUX) and the backward epilogue that multiplies by it
    (ACT_MUL_AUX): together they must reproduce autograd through the erf GELU."""
    M, N, K = shape
    torch.manual_seed(22)
    x = (torch.randn(M, K, device='cuda') * 0.5).to(dt)
    w = (torch.randn(N, K, device='cuda') * 0.05).to(dt)
    b = torch.randn(N, device='cuda').to(dt)
    y = torch.empty(M, N, dtype=dt, device='cuda')
    dact = torch.empty(M, N, dtype=dt, device='cuda')
    nat.gemm(x, w, y, M, N, K, lda=K, ldb=K, ldc=N, bias=b, act=nat.ACT_GELU_DAUX, aux_out=dact, ldaux=N)
    pre = (x.float() @ w.float().t() + b.float()).requires_grad_(True)
    ref = gelu_ref(pre)
    torch.testing.assert_close(y.float(), ref, **tol(dt))
    ref.sum().backward()
    torch.testing.assert_close(dact.float(), pre.grad, **tol(dt))
    K2 = 256
    dy = (torch.randn(M, K2, device='cuda') * 0.1).to(dt)
    w2 = (torch.randn(K2, N, device='cuda') * 0.05).to(dt)
    out = torch.empty(M, N, dtype=dt, device='cuda')
    nat.gemm(dy, w2, out, M, N, K2, lda=K2, ldb=N, ldc=N, b_mn=True, act=nat.ACT_MUL_AUX, aux_in=dact, ldaux=N)
    torch.testing.assert_close(out.float(), (dy.float() @ w2.float()) * dact.float(), **tol(dt))


@pytest.mark.parametrize('dt', DT)
@pytest.mark.parametrize('M', [300, 4096])
def test_layernorm_bwd_many_rows(nat, dt, M):
    """More rows than ring slots x warps x CTAs: exercises the refill path of the persistent kernel."""
    H = 1024
    torch.manual_seed(33)
    M = M * 5 if M == 4096 else M
    x = torch.randn(M, H, device='cuda').to(dt)
    res = torch.randn(M, H, device='cuda').to(dt)
    gamma = (1 + 0.1 * torch.randn(H, device='cuda')).to(dt)
    beta = (0.1 * torch.randn(H, device='cuda')).to(dt)
    y, z, mean, rstd = nat.ln_fwd(x, None, res, gamma, beta, p=0.0)
    zr = z.float().requires_grad_(True)
    g32, b32 = gamma.float().requires_grad_(True), beta.float().requires_grad_(True)
    u = zr.mean(-1, keepdim=True)
    sd = (zr - u).pow(2).mean(-1, keepdim=True)
    yr = g32 * ((zr - u) / torch.sqrt(sd + 1e-12)) + b32
    dy = (torch.randn(M, H, device='cuda') * 0.1).to(dt)
    yr.backward(dy.float())
    dgamma, dbeta, dbias = (torch.zeros(H, device='cuda') for _ in range(3))
    dz, dx = nat.ln_bwd(dy, z, gamma, mean, rstd, dgamma, dbeta, False, dbias=dbias)
    t = dict(rtol=3e-2, atol=3e-2) if dt == torch.bfloat16 else dict(rtol=1e-2, atol=1e-2)
    torch.testing.assert_close(dz.float(), zr.grad, **t)
    scale = math.sqrt(M)
    torch.testing.assert_close(dgamma, g32.grad, rtol=3e-2, atol=0.02 * scale)
    torch.testing.assert_close(dbeta, b32.grad, rtol=3e-2, atol=0.02 * scale)
    torch.testing.assert_close(dbias, zr.grad.sum(0), rtol=3e-2, atol=0.02 * scale)


@pytest.mark.parametrize('dt', DT)
@pytest.mark.parametrize('H', [256, 768, 1024])
@pytest.mark.parametrize('p', [0.0, 0.1])
def test_layernorm_fwd_bwd(nat, dt, H, p):
    from hetseq_b200.ops import rng
    M = 300
    torch.manual_seed(3)
    x = torch.randn(M, H, device='cuda').to(dt)
    # with dropout the oracle recovers the mask from z, which is only exact without a residual
    res = torch.randn(M, H, device='cuda').to(dt) if p == 0 else torch.zeros(M, H, device='cuda', dtype=dt)
    gamma = (1 + 0.1 * torch.randn(H, device='cuda')).to(dt)
    beta = (0.1 * torch.randn(H, device='cuda')).to(dt)
    rng.set_step_seed(1234, 0)
    seed = rng.seed_tensor(torch.device('cuda', 0))
    y, z, mean, rstd = nat.ln_fwd(x, None, res, gamma, beta, p=p, seed=seed, site=7)
    # recover the mask from z: z = res + mask*x/(1-p)
    if p > 0:
        keep = z.float() != 0
        frac = keep.float().mean().item()
        assert abs(frac - (1 - p)) < 0.02
        y2, z2, _, _ = nat.ln_fwd(x, None, res, gamma, beta, p=p, seed=seed, site=7)
        assert torch.equal(z, z2), 'dropout mask must be a pure function of (seed, site, index)'
        _, z3, _, _ = nat.ln_fwd(x, None, res, gamma, beta, p=p, seed=seed, site=8)
        assert not torch.equal(z, z3)
        mask = keep.float() / (1 - p)
    else:
        mask = torch.ones(M, H, device='cuda')
    xr, rr = x.float().requires_grad_(True), res.float().requires_grad_(True)
    g32, b32 = gamma.float().requires_grad_(True), beta.float().requires_grad_(True)
    zr = rr + xr * mask
    u = zr.mean(-1, keepdim=True)
    s = (zr - u).pow(2).mean(-1, keepdim=True)
    yr = g32 * ((zr - u) / torch.sqrt(s + 1e-12)) + b32
    torch.testing.assert_close(y.float(), yr, **tol(dt))
    dy = torch.randn(M, H, device='cuda').to(dt)
    yr.backward(dy.float())
    dgamma = torch.zeros(H, device='cuda')
    dbeta = torch.full((H,), 0.5, device='cuda')
    dbias = torch.full((H,), 0.25, device='cuda')
    dz, dx = nat.ln_bwd(dy, z, gamma, mean, rstd, dgamma, dbeta, False, p=p, seed=seed, site=7, dbias=dbias,
                        accumulate_bias=True)
    t = dict(rtol=3e-2, atol=3e-2) if dt == torch.bfloat16 else dict(rtol=1e-2, atol=1e-2)
    # the fused bias gradient is the column sum of dx (on top of what the destination already held)
    torch.testing.assert_close(dbias, xr.grad.sum(0) + 0.25, rtol=2e-2, atol=5e-2 if dt == torch.float16 else 2e-1)
    torch.testing.assert_close(dz.float(), rr.grad, **t)
    torch.testing.assert_close(dx.float(), xr.grad, **t)
    pt = dict(rtol=2e-2, atol=5e-2) if dt == torch.float16 else dict(rtol=3e-2, atol=2e-1)
    torch.testing.assert_close(dgamma, g32.grad, **pt)
    torch.testing.assert_close(dbeta, b32.grad, **pt)


@pytest.mark.parametrize('dt', DT)
@pytest.mark.parametrize('V', [2, 1000, 30522])
def test_xent(nat, dt, V):
    rows = 37
    torch.manual_seed(4)
    ld = (V + 127) // 128 * 128 if V > 128 else 8
    logits = torch.zeros(rows, ld, dtype=dt, device='cuda')
    logits[:, :V] = (torch.randn(rows, V, device='cuda') * 3).to(dt)
    labels = torch.randint(0, V, (rows,), device='cuda', dtype=torch.int32)
    labels[::5] = -1
    ref_in = logits[:, :V].float().requires_grad_(True)
    ref = F.cross_entropy(ref_in, labels.long(), ignore_index=-1)
    ref.backward()
    loss = torch.zeros(1, device='cuda')
    nat.xent(logits, ld, labels, rows, V, loss, False)
    torch.testing.assert_close(loss[0], ref.detach(), rtol=1e-4, atol=1e-4)
    torch.testing.assert_close(logits[:, :V].float(), ref_in.grad, rtol=2e-2, atol=1e-3 if dt == torch.float16 else 4e-3)
    assert torch.count_nonzero(logits[:, V:]) == 0


def test_mlm_head_gradient_survives_fp16_at_init():
    """640 labelled rows, V = 30522, near-uniform logits (a freshly initialised model): softmax / n_valid is ~5e-8,
    far below the fp16 subnormal range. The head stores softmax - onehot and carries 1/n_valid (and the upstream
    gradient) in the fp32 alpha of the decoder GEMMs; dx and dW must match an fp32 restatement of
    /root/reference/bert_modeling.py:547,881-882."""
    from hetseq_b200.ops import bert_native as bn
    torch.manual_seed(11)
    M, K, V = 640, 256, 30522
    x = (torch.randn(M, K, device='cuda') * 0.5).half().requires_grad_(True)
    w = (torch.randn(V, K, device='cuda') * 0.02).half()
    b = torch.zeros(V, device='cuda').half()
    labels = torch.randint(0, V, (M,), device='cuda', dtype=torch.int32)
    labels[::7] = -1

    class P(object):      # stands in for the flat buffer: LinearXentFn only needs a gradient target
        pass
    wg = torch.zeros(V, K, device='cuda')
    bg = torch.zeros(V, device='cuda')
    orig_target, orig_notify = bn._wgrad_target, bn._notify
    bn._wgrad_target = lambda params: ((wg.view(-1), False) if params[0] == 'w' else (bg, False))
    bn._notify = lambda params: None
    try:
        loss = bn.LinearXentFn.apply(x, w, b, labels, ['w'], ['b'])
        (loss * 1024.0).backward()      # an upstream gradient like the loss scale
    finally:
        bn._wgrad_target, bn._notify = orig_target, orig_notify
    xr = x.detach().float().requires_grad_(True)
    wr = w.float().requires_grad_(True)
    br = b.float().requires_grad_(True)
    ref = F.cross_entropy(xr @ wr.t() + br, labels.long(), ignore_index=-1)
    (ref * 1024.0).backward()
    torch.testing.assert_close(loss, ref.detach(), rtol=2e-3, atol=2e-3)

    def rel(a, r):
        return float((a.float() - r).norm() / r.norm())
    assert rel(x.grad, xr.grad) < 2e-2, rel(x.grad, xr.grad)
    assert rel(wg, wr.grad) < 2e-2, rel(wg, wr.grad)
    assert rel(bg, br.grad) < 2e-2, rel(bg, br.grad)
    # the tail of the distribution is not flushed: columns that are nobody's label still get their share
    unlabelled = torch.ones(V, dtype=torch.bool, device='cuda')
    unlabelled[labels[labels >= 0].long()] = False
    assert rel(bg[unlabelled], br.grad[unlabelled]) < 5e-2


@pytest.mark.parametrize('dt', DT)
def test_colsum_gather_scatter(nat, dt):
    M, N = 1000, 776
    torch.manual_seed(5)
    x = torch.randn(M, N, device='cuda').to(dt)
    out = torch.ones(N, device='cuda')
    nat.colsum(x, M, N, N, out, True)
    torch.testing.assert_close(out, x.float().sum(0) + 1, rtol=1e-3, atol=2e-2)
    src = torch.randn(500, 256, device='cuda').to(dt)
    idx = torch.randperm(500, device='cuda')[:64].to(torch.int32)
    g = nat.gather_rows(src, idx, 64)
    assert torch.equal(g, src[idx.long()])
    dst = torch.zeros(500, 256, dtype=dt, device='cuda')
    nat.scatter_add_rows(g, idx, dst)
    ref = torch.zeros(500, 256, dtype=dt, device='cuda')
    ref[idx.long()] = g
    assert torch.equal(dst, ref)


@pytest.mark.parametrize('dt', DT)
def test_embedding_kernels(nat, dt):
    B, S, H, V = 4, 32, 256, 1000
    torch.manual_seed(6)
    W = torch.randn(V, H, device='cuda').to(dt)
    P = torch.randn(64, H, device='cuda').to(dt)
    T = torch.randn(2, H, device='cuda').to(dt)
    ids = torch.randint(0, V, (B * S,), device='cuda', dtype=torch.int32)
    seg = torch.randint(0, 2, (B * S,), device='cuda', dtype=torch.int32)
    z = nat.embed_gather_sum(W, P, T, ids, seg, B * S, S)
    pos = torch.arange(S, device='cuda').repeat(B)
    ref = W[ids.long()].float() + P[pos].float() + T[seg.long()].float()
    torch.testing.assert_close(z.float(), ref, **tol(dt))
    de = torch.randn(B * S, H, device='cuda').to(dt)
    dW = torch.zeros(V, H, device='cuda')
    nat.embed_scatter_word(de, ids, dW)
    refW = torch.zeros(V, H, device='cuda').index_add_(0, ids.long(), de.float())
    torch.testing.assert_close(dW, refW, rtol=1e-4, atol=1e-4)
    # deterministic form: segmented reduction over the positions sorted by token id, on top of an existing
    # gradient (the tied decoder matrix); identical bits on every run, with the plan made on either side
    from hetseq_b200.tasks.bert import LanguageModelingTask
    ids[: B * S // 3] = 7                      # one long run, like [PAD]
    refW = torch.ones(V, H, device='cuda').index_add_(0, ids.long(), de.float())
    outs = []
    for plan in (nat.embedding_scatter_plan(ids), [t.cuda() for t in LanguageModelingTask.embedding_plan(ids.cpu())],
                 nat.embedding_scatter_plan(ids)):
        dW2 = torch.ones(V, H, device='cuda')
        nat.embed_scatter_word_sorted(de, ids, plan[0], plan[1], dW2)
        torch.testing.assert_close(dW2, refW, rtol=1e-4, atol=2e-3)
        outs.append(dW2)
    assert torch.equal(outs[0], outs[1]) and torch.equal(outs[0], outs[2]), 'the sorted scatter must be bit-reproducible'
    dP = torch.zeros(S * H, device='cuda')
    nat.sum_over_batch(de, B, S * H, dP, False)
    torch.testing.assert_close(dP.view(S, H), de.float().view(B, S, H).sum(0), rtol=1e-4, atol=1e-4)
    dT = torch.zeros(2, H, device='cuda')
    for t in range(2):
        nat.colsum_rows_where(de, seg, t, dT[t], False)
    refT = torch.zeros(2, H, device='cuda').index_add_(0, seg.long(), de.float())
    torch.testing.assert_close(dT, refT, rtol=1e-4, atol=1e-3)


@pytest.mark.parametrize('dt', DT)
@pytest.mark.parametrize('M,H,V,wide', [(4096, 1024, 30522, False), (1000, 256, 500, True), (40, 64, 50, False)])
def test_embedding_backward_sorted_segments(nat, dt, M, H, V, wide):
    """Word + token-type gradient in one segmented pass (hs_embed_backward_sorted): a quarter of the positions hold
    one id ([PAD]-like run over dozens of 16-row segments), a tenth another ([MASK]-like), M need not be a multiple
    of 16, ids outside the table are skipped; against index_add_ in fp32, bit-identical on repeated calls, and the
    token-type sums with and without accumulation. CPU model of the same logic: tests/test_embed_segments_cpu.py."""
    from hetseq_b200.tasks.bert import LanguageModelingTask
    g = torch.Generator(device='cpu').manual_seed(M)
    ids = torch.randint(3, V, (M,), generator=g)
    perm = torch.randperm(M, generator=g)
    ids[perm[: M // 4]] = 0
    ids[perm[M // 4: M // 4 + M // 10]] = 2
    if M >= 1000:
        ids[perm[-3:]] = V + 7                 # not a row of the table
    idt = torch.int64 if wide else torch.int32
    seg = torch.randint(0, 2, (M,), generator=g).to(idt).cuda()
    de = torch.randn(M, H, generator=g).to(dt).cuda()
    plan = [t.cuda() for t in LanguageModelingTask.embedding_plan(ids)]
    ids = ids.to(idt).cuda()
    ok = ids < V
    refW = torch.ones(V, H, device='cuda').index_add_(0, ids[ok].long(), de[ok].float())
    refT = torch.zeros(2, H, device='cuda').index_add_(0, seg.long(), de.float())
    outs = []
    for _ in range(3):
        dW = torch.ones(V, H, device='cuda')
        dT = torch.full((2, H), 5.0, device='cuda')
        nat.embed_backward_sorted(de, ids, plan[0], plan[1], dW, seg=seg, dT=dT, accumulate_T=False)
        torch.testing.assert_close(dW, refW, rtol=1e-4, atol=2e-3)
        torch.testing.assert_close(dT, refT, rtol=1e-4, atol=5e-3)
        outs.append((dW, dT))
    for dW, dT in outs[1:]:
        assert torch.equal(dW, outs[0][0]) and torch.equal(dT, outs[0][1]), 'must be bit-reproducible'
    dT = torch.full((2, H), 5.0, device='cuda')
    dW = torch.ones(V, H, device='cuda')
    nat.embed_backward_sorted(de, ids, plan[0], plan[1], dW, seg=seg, dT=dT, accumulate_T=True)
    torch.testing.assert_close(dT, refT + 5.0, rtol=1e-4, atol=5e-3)
    dW2 = torch.ones(V, H, device='cuda')
    nat.embed_scatter_word_sorted(de, ids, plan[0], plan[1], dW2)          # word part alone
    assert torch.equal(dW2, outs[0][0])


@pytest.mark.parametrize('kind', ['adam', 'adadelta'])
@pytest.mark.parametrize('dt', [torch.float16, torch.float32])
def test_fused_optimizer_matches_per_tensor(kind, dt):
    """Flat native update == the reference's per-tensor math (optim/tensor_optimizers.py)."""
    from hetseq_b200 import options
    from hetseq_b200.optim import Adam, Adadelta, _Adam, _Adadelta
    torch.manual_seed(7)
    shapes = [(300, 70), (1000,), (64, 64), (5,)]
    ref_params = [torch.nn.Parameter(torch.randn(s, device='cuda')) for s in shapes]
    params = [torch.nn.Parameter(p.detach().clone().to(dt)) for p in ref_params]
    for p, q in zip(ref_params, params):
        p.data.copy_(q.data.float())
    args = options.make_args('mnist', kind, argv=['--lr', '0.01'] + (['--weight-decay', '0.1'] if kind == 'adam' else []))
    args.kernels = 'native'
    opt = (_Adam if kind == 'adam' else _Adadelta)(args, params)
    assert opt._native is not None
    ref = Adam(ref_params, lr=0.01, weight_decay=0.1) if kind == 'adam' else Adadelta(ref_params, lr=0.01)
    for step in range(3):
        opt.zero_grad()
        for p, q in zip(ref_params, params):
            g = torch.randn_like(p)
            p.grad = g.clone() * 0.5
            q.main_grad.copy_(g)
        opt.multiply_grads(0.5)
        gn = opt.clip_grad_norm(0.0)
        ref_norm = math.sqrt(sum(float(p.grad.pow(2).sum()) for p in ref_params))
        assert abs(float(gn) - ref_norm) / ref_norm < 1e-4
        opt.step()
        ref.step()
    for p, q in zip(ref_params, params):
        if dt == torch.float32:
            torch.testing.assert_close(q.data, p.data, rtol=1e-5, atol=1e-6)
        else:
            torch.testing.assert_close(q.data.float(), p.data, rtol=2e-3, atol=2e-3)
    # fp32 master stays exact even for fp16 params
    i = 0
    torch.testing.assert_close(opt.master[:ref_params[0].numel()].view(shapes[0]), ref_params[0].data, rtol=1e-5, atol=1e-6)


def _attn_ref(qkv, mask, B, S, heads):
    M, H3 = qkv.shape
    H = H3 // 3
    d = H // heads
    q, k, v = qkv.view(B, S, 3, heads, d).permute(2, 0, 3, 1, 4)
    scores = torch.matmul(q, k.transpose(-1, -2)) / math.sqrt(d) + mask.view(B, 1, 1, S)
    return torch.matmul(torch.softmax(scores, dim=-1), v).permute(0, 2, 1, 3).reshape(M, H)


@pytest.mark.parametrize('dt', DT)
@pytest.mark.parametrize('S', [128, 256, 512, 48, 96, 208, 400])
def test_attention_fwd_bwd(dt, S):
    """S = 128: single-tile kernels; every other multiple of 16 up to 512: multi-tile kernels (two-pass forward,
    dK/dV and dQ backward), including sequence lengths whose last 128-row tile is partial (48, 96, 208, 400: rows
    beyond S arrive as zeros through the 3-D tensor map and never reach memory)."""
    from hetseq_b200.ops import attention_native as an
    B, heads, d = 3, 4, 64
    H = heads * d
    assert an.supports(B, S, heads, d, dt)
    torch.manual_seed(8)
    qkv = (torch.randn(B * S, 3 * H, device='cuda') * 0.7).to(dt).requires_grad_(True)
    mask = torch.zeros(B, S, device='cuda')
    mask[0, S - 28:] = -10000.0
    mask[2, 17:] = -10000.0
    # (with a partial last tile a CTA that wrote rows beyond its own sequence would clobber the next one's: the
    # comparison of every row below catches that)
    out = an.attention(qkv, mask, B, S, heads, 0.0, 1)
    ref_in = qkv.detach().float().requires_grad_(True)
    ref = _attn_ref(ref_in, mask, B, S, heads)
    torch.testing.assert_close(out.float(), ref, **tol(dt))
    dout = (torch.randn(B * S, H, device='cuda') * 0.5).to(dt)
    out.backward(dout)
    ref.backward(dout.float())
    err = ((qkv.grad.float() - ref_in.grad).norm() / ref_in.grad.norm()).item()
    assert err < (0.03 if dt == torch.bfloat16 else 0.006), err
    for name, sl in (('dq', slice(0, H)), ('dk', slice(H, 2 * H)), ('dv', slice(2 * H, 3 * H))):
        e = ((qkv.grad.float()[:, sl] - ref_in.grad[:, sl]).norm() / ref_in.grad[:, sl].norm()).item()
        assert e < (0.03 if dt == torch.bfloat16 else 0.006), (name, e)
    # the q/k/v bias gradient collected by the backward kernel = column sums of dqkv
    from hetseq_b200.ops import lib
    L = lib.load()
    lse = torch.empty(B * heads * S, dtype=torch.float32, device='cuda')
    o2 = torch.empty(B * S, H, dtype=dt, device='cuda')
    q2 = qkv.detach()
    lib.check(L.hs_attention_fwd(q2.data_ptr(), mask.data_ptr(), o2.data_ptr(), lse.data_ptr(), B, S, heads, H, 0.0, None,
                                 1, lib.dtype_code(dt), lib.stream_ptr()), 'fwd')
    dq2 = torch.empty_like(q2)
    dbias = torch.full((3 * H,), 1.5, device='cuda')
    lib.check(L.hs_attention_bwd(q2.data_ptr(), mask.data_ptr(), o2.data_ptr(), dout.data_ptr(), lse.data_ptr(),
                                 dq2.data_ptr(), dbias.data_ptr(), 1, B, S, heads, H, 0.0, None, 1, lib.dtype_code(dt),
                                 lib.stream_ptr()), 'bwd')
    assert torch.equal(dq2, qkv.grad)
    torch.testing.assert_close(dbias, ref_in.grad.sum(0) + 1.5, rtol=2e-2, atol=0.05 if dt == torch.float16 else 0.3)


@pytest.mark.parametrize('S', [128, 384, 80])
def test_attention_dropout_consistency(S):
    """Dropout: same (seed, site) -> identical output; backward uses the very same mask (checked through
    linearity: with V = ones-like structure the gradient of sum(out) w.r.t. v equals column sums of P_drop)."""
    from hetseq_b200.ops import attention_native as an, rng
    B, heads, d = 2, 2, 64
    H = heads * d
    dt = torch.float16
    torch.manual_seed(9)
    rng.set_step_seed(77, 0)
    qkv = (torch.randn(B * S, 3 * H, device='cuda') * 0.5).to(dt)
    mask = torch.zeros(B, S, device='cuda')
    o1 = an.attention(qkv, mask, B, S, heads, 0.1, 5)
    o2 = an.attention(qkv, mask, B, S, heads, 0.1, 5)
    o3 = an.attention(qkv, mask, B, S, heads, 0.1, 6)
    o0 = an.attention(qkv, mask, B, S, heads, 0.0, 5)
    assert torch.equal(o1, o2) and not torch.equal(o1, o3)
    # E[dropout(P)] = P: averaged over many sites the output approaches the no-dropout one
    acc = torch.zeros_like(o0, dtype=torch.float32)
    n = 64
    for site in range(100, 100 + n):
        acc += an.attention(qkv, mask, B, S, heads, 0.1, site).float()
    rel = ((acc / n - o0.float()).norm() / o0.float().norm()).item()
    assert rel < 0.06, rel
    # finite-difference check of the backward through the SAME mask: d<out, g>/dv is linear in v
    q = qkv.clone().requires_grad_(True)
    g = torch.randn(B * S, H, device='cuda').to(dt)
    out = an.attention(q, mask, B, S, heads, 0.1, 5)
    out.backward(g)
    dv = q.grad[:, 2 * H:].float()
    delta = torch.zeros_like(qkv)
    delta[:, 2 * H:] = (torch.randn(B * S, H, device='cuda') * 0.5).to(dt)
    out2 = an.attention(qkv + delta, mask, B, S, heads, 0.1, 5)
    lhs = ((out2.float() - o1.float()) * g.float()).sum().item()
    rhs = (dv * delta[:, 2 * H:].float()).sum().item()
    assert abs(lhs - rhs) / (abs(rhs) + 1e-3) < 0.05, (lhs, rhs)
