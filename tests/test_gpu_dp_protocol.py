"""The fused data-parallel protocol (csrc/comm_sm100.cu) on ONE GPU.

``hs_dp_emulate_step`` runs W virtual ranks of the production device code -- bucket-wise reduce-scatter with
announce/wait flags, norm + sample-size exchange, sharded Adam / Adadelta with the parameter all-gather and
per-slice arrival flags, the parameter wait -- inside one cooperative launch, each "rank" with its own gradient,
parameter and signal buffers. Oracle: plain PyTorch fp32 on the summed gradients + the per-tensor reference
optimizers (hetseq_b200/optim/tensor_optimizers.py, i.e. /root/reference/optim.py:162-231,263-304 and the
multiply_grads / clip_grad_norm sequence of /root/reference/controller.py:338-347).

Runs in the driver's 1-GPU ``pytest -m gpu``; the real multi-GPU path (NVLink peer memory, multimem) is
covered by tests/test_gpu_multi.py on boxes with >= 2 GPUs.
"""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu

CHUNK = 2048


@pytest.fixture(autouse=True)
def _short_peer_timeout():
    """A protocol bug shows up as a wait that never ends: make it a trap after seconds, not after the
    production default of half an hour."""
    from hetseq_b200.ops import lib
    L = lib.load()
    lib.check(L.hs_dp_set_timeout(15.0), 'hs_dp_set_timeout')
    yield
    lib.check(L.hs_dp_set_timeout(1800.0), 'hs_dp_set_timeout')


def _ptr_array(tensors, ctype=ctypes.c_void_p):
    arr = (ctype * len(tensors))()
    for i, t in enumerate(tensors):
        arr[i] = t.data_ptr()
    return arr


class VirtualRanks(object):
    def __init__(self, W, blocks, dtype, opt='adam', seed=0):
        from hetseq_b200.ops import lib
        self.L = lib.load()
        self.lib = lib
        self.W, self.blocks, self.dtype, self.opt = W, blocks, dtype, opt
        dev = torch.device('cuda', 0)
        g = torch.Generator(device='cpu').manual_seed(seed)
        self.N = N = W * blocks * CHUNK
        p0 = (torch.randn(N, generator=g) * 0.05).to(dtype)
        self.params = [p0.clone().to(dev) for _ in range(W)]
        self.grads = [torch.zeros(N, dtype=torch.float32, device=dev) for _ in range(W)]
        words = int(self.L.hs_dp_signal_words(W, 64))
        self.sigs = [torch.zeros(words, dtype=torch.int32, device=dev) for _ in range(W)]
        self.sumsq = [torch.zeros(1, dtype=torch.float32, device=dev) for _ in range(W)]
        self.ss = [torch.zeros(1, dtype=torch.float32, device=dev) for _ in range(W)]
        self.scalars = [torch.zeros(8, dtype=torch.float32, device=dev) for _ in range(W)]
        self.skip = [torch.zeros(1, dtype=torch.int32, device=dev) for _ in range(W)]
        self.pending = [torch.zeros(1, dtype=torch.int32, device=dev) for _ in range(W)]
        self.ngather_req = 4
        self.ngather = int(self.L.hs_dp_gather_count(blocks, self.ngather_req))      # non-empty gather buckets
        self.counters = [torch.zeros(self.ngather + 1, dtype=torch.int32, device=dev) for _ in range(W)]
        self.barrier = torch.zeros(W, dtype=torch.int32, device=dev)

        def owned(flat, r):
            return flat.view(blocks, W, CHUNK)[:, r, :].reshape(-1)
        self.owned = owned
        self.master = [owned(self.params[r], r).float().contiguous() for r in range(W)]
        self.s1 = [torch.zeros(blocks * CHUNK, dtype=torch.float32, device=dev) for _ in range(W)]
        self.s2 = [torch.zeros(blocks * CHUNK, dtype=torch.float32, device=dev) for _ in range(W)]
        self.epoch = 0
        # oracle state: one fp32 parameter vector + the reference optimizer
        from hetseq_b200.optim import tensor_optimizers as to
        self.ref_p = torch.nn.Parameter(p0.float().to(dev))
        self.hyper = dict(lr=1e-2, eps=1e-8 if opt == 'adam' else 1e-6, weight_decay=0.01)
        if opt == 'adam':
            self.ref_opt = to.Adam([self.ref_p], lr=self.hyper['lr'], betas=(0.9, 0.98), eps=self.hyper['eps'],
                                   weight_decay=self.hyper['weight_decay'])
        else:
            self.ref_opt = to.Adadelta([self.ref_p], lr=self.hyper['lr'], rho=0.9, eps=self.hyper['eps'],
                                       weight_decay=self.hyper['weight_decay'])
        self.t = 0

    def step(self, grads, sample_sizes, mult_host=1.0, max_norm=1.0, chunks_per_bucket=3, blocks_per_rank=3, split=1):
        import math
        W = self.W
        for r in range(W):
            self.grads[r].copy_(grads[r])
            self.ss[r].fill_(float(sample_sizes[r]))
        self.epoch += 1
        a, b = (0.9, 0.98) if self.opt == 'adam' else (0.9, 0.0)
        t = self.t + 1
        step_factor = math.sqrt(1 - b ** t) / (1 - a ** t) if self.opt == 'adam' else 1.0
        f = ctypes.c_float
        rc = self.L.hs_dp_emulate_step(
            W, blocks_per_rank, _ptr_array(self.grads), _ptr_array(self.params), _ptr_array(self.sigs),
            _ptr_array(self.sumsq), _ptr_array(self.ss), _ptr_array(self.scalars), _ptr_array(self.skip),
            _ptr_array(self.master), _ptr_array(self.s1), _ptr_array(self.s2), _ptr_array(self.counters),
            _ptr_array(self.pending), self.barrier.data_ptr(), W * self.blocks, chunks_per_bucket, CHUNK, self.ngather_req,
            0 if self.opt == 'adam' else 1, self.lib.dtype_code(self.dtype), self.hyper['lr'], a, b, self.hyper['eps'],
            self.hyper['weight_decay'], step_factor, mult_host, max_norm, self.epoch, int(split), self.lib.stream_ptr())
        self.lib.check(rc, 'hs_dp_emulate_step')
        torch.cuda.synchronize()
        # ---- oracle ------------------------------------------------------------------------------------
        gsum = torch.stack([g.double() for g in grads]).sum(0)
        ss = float(sum(sample_sizes))
        mult = mult_host * (1.0 / ss if ss > 0 else 1.0 / W)
        norm = float(gsum.pow(2).sum().sqrt()) * abs(mult)
        bad = not math.isfinite(norm)
        coef = 1.0 if bad or max_norm <= 0 else min(max_norm / (norm + 1e-6), 1.0)
        if not bad:
            self.t += 1
            self.ref_p.grad = (gsum * (mult * coef)).float()
            self.ref_opt.step()
        return gsum, norm, bad


def _check(vr, gsum, norm, bad, rtol_norm=2e-5):
    W = vr.W
    for r in range(W):
        sc = vr.scalars[r].tolist()
        assert bool(sc[2]) == bad and int(vr.skip[r]) == int(bad)
        assert int(vr.pending[r]) == 0, 'the update kernel must consume the pending flag'
        assert int(vr.counters[r].abs().sum()) == 0, 'completion counters must be back at zero'
        if not bad:
            assert abs(sc[0] - norm) <= rtol_norm * max(1.0, norm), (sc[0], norm)
            # the owned chunks hold the sum over ranks
            torch.testing.assert_close(vr.owned(vr.grads[r], r), vr.owned(gsum.float(), r), rtol=1e-5, atol=1e-5)
    ref16 = vr.ref_p.detach().to(vr.dtype)
    for r in range(W):
        # every rank ends with the same parameters, and they are the oracle's (one 16-bit ulp: the kernels
        # multiply the gradient by the clip coefficient in fp32, the oracle in fp64)
        assert torch.equal(vr.params[r], vr.params[0]), 'replicas differ'
        diff = (vr.params[r].float() - ref16.float()).abs()
        tol = 2.0 ** -8 if vr.dtype == torch.bfloat16 else 2.0 ** -11
        assert float((diff / ref16.float().abs().clamp(min=1e-2)).max()) <= 2 * tol, float(diff.max())
        torch.testing.assert_close(vr.master[r], vr.owned(vr.ref_p.detach(), r), rtol=2e-4, atol=2e-6)


@pytest.mark.parametrize('W,dtype,opt', [(2, torch.float16, 'adam'), (4, torch.bfloat16, 'adam'),
                                         (8, torch.float16, 'adam'), (4, torch.float16, 'adadelta')])
def test_virtual_ranks_match_reference_update(W, dtype, opt):
    vr = VirtualRanks(W, blocks=7, dtype=dtype, opt=opt, seed=W)
    g = torch.Generator(device='cpu').manual_seed(100 + W)
    for it in range(3):
        grads = [(torch.randn(vr.N, generator=g) * (0.5 + it)).cuda() for _ in range(W)]
        sizes = [4 + r for r in range(W)]              # uneven sample sizes, one rank may even hold none
        if it == 1:
            sizes[0] = 0
        gsum, norm, bad = vr.step(grads, sizes, mult_host=1.0 / 128.0, max_norm=0.05 if it != 2 else 0.0,
                                  chunks_per_bucket=3 + it, blocks_per_rank=1 + 2 * it, split=it % 2 == 0)
        assert not bad
        _check(vr, gsum, norm, bad)


def test_virtual_ranks_overflow_skips_everywhere():
    W = 4
    vr = VirtualRanks(W, blocks=5, dtype=torch.float16, seed=9)
    g = torch.Generator(device='cpu').manual_seed(5)
    grads = [torch.randn(vr.N, generator=g).cuda() for _ in range(W)]
    gsum, norm, bad = vr.step(grads, [8] * W)
    _check(vr, gsum, norm, bad)
    before = [p.clone() for p in vr.params]
    m_before = [m.clone() for m in vr.master]
    s1_before = [s.clone() for s in vr.s1]
    # one element of ONE rank's gradient overflows: every rank must skip, nothing may change
    grads = [torch.randn(vr.N, generator=g).cuda() for _ in range(W)]
    grads[2][12345] = float('inf')
    gsum, norm, bad = vr.step(grads, [8] * W)
    assert bad
    _check(vr, gsum, norm, bad)
    for r in range(W):
        assert torch.equal(vr.params[r], before[r]) and torch.equal(vr.master[r], m_before[r])
        assert torch.equal(vr.s1[r], s1_before[r])
    # and the protocol keeps working afterwards (epochs moved on, flags were still published)
    grads = [torch.randn(vr.N, generator=g).cuda() for _ in range(W)]
    gsum, norm, bad = vr.step(grads, [8] * W)
    assert not bad
    _check(vr, gsum, norm, bad)


def test_wait_params_and_flags_are_epoch_numbered():
    """Replaying an OLD epoch must not satisfy waits of a newer one: flags only ever grow."""
    from hetseq_b200.ops import lib
    L = lib.load()
    W = 2
    vr = VirtualRanks(W, blocks=3, dtype=torch.float16, seed=1)
    g = torch.Generator(device='cpu').manual_seed(2)
    for _ in range(2):
        grads = [torch.randn(vr.N, generator=g).cuda() for _ in range(W)]
        vr.step(grads, [1] * W)
    flags = vr.sigs[0].view(-1)
    words_param = [3 * W + j * W + src for j in range(vr.ngather) for src in range(W)]
    assert all(int(flags[w]) == 2 for w in words_param)
    # a wait for epoch 2 on rank 0 returns immediately (would trap after the timeout otherwise)
    lib.check(L.hs_dp_wait_params(_ptr_array(vr.sigs), 0, W, 0, vr.ngather - 1, 2, None, 0, lib.stream_ptr()),
              'hs_dp_wait_params')
    torch.cuda.synchronize()
