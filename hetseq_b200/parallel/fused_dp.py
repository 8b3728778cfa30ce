"""Fused NVLink data-parallel backend (``--dp-backend fused``): the gradient path without NCCL.

Counterpart of DDP's bucketed all-reduce + the reference's optimizer loop (controller.py:74-89,
338-347; optim.py:59-70,162-231,263-304). The flat gradient and parameter buffers of all
ranks of one node live in *symmetric memory*: every rank maps every peer's copy and, where the NVSwitch
offers it, one NVLS multicast address per buffer. The kernels of ``csrc/comm_sm100.cu`` then do, per update:

  backward ............ as soon as a bucket's gradients are final (gradient-ready callbacks from autograd /
                        the native wgrad ops) its ``hs_dp_reduce_scatter`` is queued on a side stream:
                        announce to peers -> wait for peers -> reduce the W copies of the owned chunks (one
                        ``multimem.ld_reduce`` per vector, or W-1 peer loads) -> norm partial.
  clip_grad_norm ...... ``hs_dp_norm_exchange``: 2 W floats through the signal pads; identical norm, gradient
                        denominator (sum of the ranks' sample sizes), clip coefficient and overflow flag on
                        all ranks, no host round trip and no NCCL call.
  step ................ ``hs_dp_update_allgather``: Adam / Adadelta on the 1/W owned chunks, new 16-bit
                        parameters stored into every rank's buffer (``multimem.st`` or W peer stores), one
                        arrival flag per slice of the buffer ("gather bucket").

The update is *rotated* into the next step: ``queue_update`` only marks it pending; ``begin_update`` of the next
step launches it on a low-priority side stream and the forward pass waits per gather bucket
(``gate``/``hs_dp_wait_params``) in front of the first layer that reads it, so the all-gather of layer k+1..L
travels while layers 0..k compute. ``before_backward`` joins the side stream (the update reads the gradient
buffer that backward is about to overwrite). ``flush()`` applies a pending update immediately -- needed before
anything else looks at the parameters (checkpoints, validation, consistency checks, the end of training).
Eager execution and CUDA-graph replay queue exactly the same cross-rank operations in the same order, so ranks
may mix the two freely.

torch.distributed is used only at start-up (symmetric-memory rendezvous or IPC handle exchange) and, in the
Controller, for the 48-byte statistics all-reduce, which the north star allows to stay on NCCL.
"""
from __future__ import annotations

import ctypes
import os
from typing import List, Optional

import torch
import torch.distributed as dist

from hetseq_b200.ops import lib
from hetseq_b200.parallel.data_parallel import AllReduceBackend
from hetseq_b200.parallel.flat_buffer import CHUNK, FlatParamBuffer

MAX_BUCKETS = 1024
OPT_KIND = {'adam': 0, 'adadelta': 1}
# one dedicated VMM block per symmetric buffer (must be set before torch's symmetric memory is first used)
os.environ.setdefault('TORCH_SYMMMEM_IMPLICIT_POOL', '0')


# ---------------------------------------------------------------------------------------------------------
# symmetric allocations
# ---------------------------------------------------------------------------------------------------------
class _IpcBlock(object):
    """Owner object behind a tensor that aliases an hs_ipc_alloc block (CUDA array interface)."""

    def __init__(self, ptr: int, nbytes: int):
        self.ptr, self.nbytes = ptr, nbytes
        self.__cuda_array_interface__ = {'shape': (nbytes,), 'typestr': '|u1', 'data': (ptr, False), 'version': 3,
                                         'strides': None}


class SymmetricPool(object):
    """Allocates tensors that every rank of the group can map, and resolves them to per-rank device pointers.

    ``kind == 'symm'``: ``torch.distributed._symmetric_memory`` (CUDA VMM allocations, exchanged over the
    process group's store; when the fabric supports NVLS the block is also bound to a multicast object and
    ``multicast_ptr`` is non-zero). ``kind == 'ipc'``: plain ``cudaMalloc`` + legacy CUDA IPC handles (peer
    pointers only). Which one is used is decided collectively by ``probe``."""

    def __init__(self, kind: str):
        self.kind = kind
        self._blocks = []     # (tensor, device pointer, ipc handle bytes)            [ipc]
        self._handles = {}    # data_ptr -> symmetric-memory handle                   [symm]
        self._opened = []     # peer mappings opened by this rank                     [ipc]

    # -- allocation ---------------------------------------------------------------------------------
    def alloc(self, numel: int, dtype: torch.dtype, device) -> torch.Tensor:
        device = torch.device(device)
        if self.kind == 'symm':
            import torch.distributed._symmetric_memory as symm_mem
            with torch.cuda.device(device):
                t = symm_mem.empty(max(numel, 64), dtype=dtype, device=device)
            t.zero_()
            return t[:numel]
        L = lib.load()
        nbytes = max(numel * torch.empty(0, dtype=dtype).element_size(), 256)
        ptr = ctypes.c_void_p()
        handle = ctypes.create_string_buffer(64)
        with torch.cuda.device(device):
            lib.check(L.hs_ipc_alloc(nbytes, ctypes.byref(ptr), handle), 'hs_ipc_alloc')
        owner = _IpcBlock(ptr.value, nbytes)
        raw = torch.as_tensor(owner, device=device)
        t = raw.view(dtype)[:numel]
        self._blocks.append((t, ptr.value, handle.raw))
        return t

    __call__ = alloc     # FlatParamBuffer's allocator signature

    # -- resolution ---------------------------------------------------------------------------------
    def resolve(self, t: torch.Tensor, group, rank: int, world: int):
        """(ctypes array of W device pointers indexed by rank, multicast pointer or None)."""
        arr = (ctypes.c_void_p * world)()
        if self.kind == 'symm':
            import torch.distributed._symmetric_memory as symm_mem
            hdl = symm_mem.rendezvous(t, group if group is not None else dist.group.WORLD)
            self._handles[t.data_ptr()] = hdl
            for p in range(world):
                arr[p] = t.data_ptr() if p == rank else hdl.get_buffer(p, (1,), torch.uint8).data_ptr()
            mc = int(getattr(hdl, 'multicast_ptr', 0) or 0)
            if mc:
                # the tensor may sit at an offset inside the mapped block; the multicast mapping mirrors the block
                mc += t.data_ptr() - int(hdl.buffer_ptrs[rank])
            return arr, (mc or None)
        L = lib.load()
        mine = None
        for tt, ptr, handle in self._blocks:
            if tt.data_ptr() == t.data_ptr():
                mine = handle
        if mine is None:
            raise RuntimeError('tensor was not allocated from this symmetric pool')
        everyone: List[Optional[bytes]] = [None] * world
        dist.all_gather_object(everyone, mine, group=group)
        for p in range(world):
            if p == rank:
                arr[p] = t.data_ptr()
            else:
                ptr = ctypes.c_void_p()
                lib.check(L.hs_ipc_open(everyone[p], ctypes.byref(ptr)), 'hs_ipc_open')
                self._opened.append(ptr.value)
                arr[p] = ptr.value
        return arr, None

    def release(self):
        """Unmap the peers' blocks and free our own (call after a barrier: nobody may still be using them)."""
        L = lib.load() if lib.available() else None
        for p in self._opened:
            L.hs_ipc_close(p)
        self._opened = []
        for _, ptr, _h in self._blocks:
            L.hs_ipc_free(ptr)
        self._blocks = []
        self._handles = {}


_probe_result = {}


def probe(group=None) -> str:
    """Collective: 'symm' when every rank can allocate and rendezvous symmetric memory, else 'ipc'.
    ``HETSEQ_DP_ALLOC=ipc|symm`` overrides the probe."""
    forced = os.environ.get('HETSEQ_DP_ALLOC', '').lower()
    if forced in ('ipc', 'symm'):
        return forced
    key = id(group)
    if key in _probe_result:
        return _probe_result[key]
    ok = 1
    try:
        import torch.distributed._symmetric_memory as symm_mem
        dev = torch.device('cuda', torch.cuda.current_device())
        t = symm_mem.empty(1024, dtype=torch.float32, device=dev)
        symm_mem.rendezvous(t, group if group is not None else dist.group.WORLD)
    except Exception as e:     # noqa: BLE001 -- any failure means "use the IPC path"
        print('| symmetric memory unavailable ({}: {}); using CUDA IPC peer mappings'.format(
            type(e).__name__, str(e)[:120]), flush=True)
        ok = 0
    flag = torch.tensor([ok], dtype=torch.int32, device=torch.device('cuda', torch.cuda.current_device()))
    dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=group)
    _probe_result[key] = 'symm' if int(flag.item()) == 1 else 'ipc'
    return _probe_result[key]


def is_supported(args) -> bool:
    """Single node, CUDA, all ranks' GPUs peer-accessible, kernel library present, 16-bit params."""
    if not torch.cuda.is_available() or getattr(args, 'cpu', False) or not lib.available():
        return False
    world = getattr(args, 'distributed_world_size', 1)
    if world < 2 or world > min(16, torch.cuda.device_count()):
        return False
    if getattr(args, 'optimizer', 'adam') not in OPT_KIND or not (getattr(args, 'fp16', False) or getattr(args, 'bf16', False)):
        return False
    if getattr(args, 'kernels', 'auto') == 'torch':
        return False
    dev = torch.cuda.current_device()
    return all(d == dev or torch.cuda.can_device_access_peer(dev, d) for d in range(world))


def symmetric_allocator(args, group=None) -> SymmetricPool:
    return SymmetricPool(probe(group))


# ---------------------------------------------------------------------------------------------------------
class FusedBackend(AllReduceBackend):
    sharded = True
    fused = True
    name = 'fused-nvlink'

    def __init__(self, args, buffer: FlatParamBuffer, group=None, pool: Optional[SymmetricPool] = None):
        super().__init__(args, buffer, group)
        self.L = lib.load()
        if buffer.flat_grad.dtype != torch.float32:
            raise RuntimeError('fused backend expects fp32 main gradients')
        if len(self.buckets) > MAX_BUCKETS:
            raise RuntimeError('too many gradient buckets ({}); raise --bucket-cap-mb'.format(len(self.buckets)))
        self.pool = pool or getattr(buffer, 'allocator', None)
        if not isinstance(self.pool, SymmetricPool):
            raise RuntimeError('the flat buffers were not allocated from a symmetric pool')
        dev = buffer.device
        W, r = self.world_size, self.rank
        self.sig = self.pool.alloc(int(self.L.hs_dp_signal_words(W, MAX_BUCKETS)), torch.int32, dev)
        self.grad_ptrs, self.grad_mc = self.pool.resolve(buffer.flat_grad, self.group, r, W)
        self.param_ptrs, self.param_mc = self.pool.resolve(buffer.flat_param, self.group, r, W)
        self.sig_ptrs, _ = self.pool.resolve(self.sig, self.group, r, W)
        transport = os.environ.get('HETSEQ_DP_TRANSPORT', getattr(args, 'fused_transport', 'auto') or 'auto').lower()
        if transport == 'p2p':
            self.grad_mc = self.param_mc = None
        elif transport == 'multimem' and not (self.grad_mc and self.param_mc):
            raise RuntimeError('--fused-transport multimem: no NVLS multicast mapping on this system')
        elif transport == 'auto' and W < 4:
            # measured on 2 GPUs: 15.6 ms/step with multimem vs 14.5 with peer loads. A switch-side reduction also
            # pulls this GPU's OWN copy over NVLink, which doubles the egress when there is only one peer.
            self.grad_mc = self.param_mc = None
        self.transport = 'multimem' if (self.grad_mc and self.param_mc) else 'p2p'
        self.sumsq = torch.zeros(1, dtype=torch.float32, device=dev)
        self.scalars = torch.zeros(8, dtype=torch.float32, device=dev)
        self.scalars[2] = 1.0    # "skip": nothing to apply yet
        self.skip = torch.zeros(1, dtype=torch.int32, device=dev)
        self.pending_dev = torch.zeros(1, dtype=torch.int32, device=dev)
        self.ss_local = torch.zeros(1, dtype=torch.float32, device=dev)
        self.ngather_req = max(1, min(int(os.environ.get('HETSEQ_DP_GATHER_BUCKETS', '32')),
                                      int(self.L.hs_dp_max_gather_buckets()), buffer.num_blocks))
        #: gather buckets that actually hold blocks (only these are published and may be waited for)
        self.ngather = int(self.L.hs_dp_gather_count(buffer.num_blocks, self.ngather_req))
        self.counters = torch.zeros(self.ngather + 1, dtype=torch.int32, device=dev)
        # default priority (the lowest): the reduce / update kernels must never delay the placement of the
        # forward's and backward's CTAs; they run in whatever the GEMMs leave free
        prio = int(os.environ.get('HETSEQ_COMM_PRIORITY', '0'))
        self.comm_stream = torch.cuda.Stream(device=dev, priority=prio)
        self.update_stream = torch.cuda.Stream(device=dev, priority=prio)
        # the update number lives in device memory and is bumped by a tiny device op per update, so the
        # very same kernel arguments stay valid when the step is replayed from a CUDA graph
        self.epoch_t = torch.zeros(1, dtype=torch.int32, device=dev)
        # Grids of the kernels that run NEXT TO forward / backward: the fewer CTAs cover the NVLink latency, the
        # less they take from the GEMMs (which deal their tiles statically: one slowed SM slows the whole GEMM).
        # Measured (bench.py, BERT-large, graph; (reduce-scatter, gather) CTAs -> ms per step; profiles/scaling.md):
        #   8 GPUs multimem: (148,148) 14.38  (74,148) 13.56  (37,148) 13.46  (74,37) 13.15  (37,37) 14.41
        #   4 GPUs multimem: (74,37) 14.84  (37,37) 13.75  (37,16) 13.82  (24,24) 14.54  (16,16) 15.03
        #   2 GPUs peer loads: (148,148) 14.46  (74,148) 14.48  (37,37) 15.63;  8 GPUs peer loads (296,148) 14.56
        # i.e. per rank the reduce-scatter wants roughly one CTA per 4-5 MB of owned gradient, and it must not fall
        # behind backward. HETSEQ_RS_BLOCKS / HETSEQ_GATHER_BLOCKS pin other sizes (RS: -1 = one CTA per owned chunk).
        if self.transport == 'multimem':
            default_rb, default_gb = (74 if W >= 8 else 37), 37
        else:
            default_rb, default_gb = (148 if W <= 2 else 296), 148
        rb = int(os.environ.get('HETSEQ_RS_BLOCKS', '0') or getattr(args, 'fused_rs_blocks', 0) or default_rb)
        self.rs_blocks = 0 if rb < 0 else rb
        self.update_blocks = int(os.environ.get('HETSEQ_UPDATE_BLOCKS', '0') or 0)
        self.gather_blocks = int(os.environ.get('HETSEQ_GATHER_BLOCKS', '0') or default_gb)
        self.overlap_update = os.environ.get('HETSEQ_DP_ROTATE', '1') != '0'
        # split: Adam / Adadelta at the end of the step at full speed, only the gather rides next to the forward
        # pass (a pure copy fits beside the GEMM CTAs; the full update does not get enough of an SM there)
        self.split_update = os.environ.get('HETSEQ_DP_SPLIT', '1') != '0'
        self._opt = None            # the optimizer whose state the update kernel works on (bind_optimizer)
        self._joined = True
        self._gated = self.ngather  # gather buckets already waited for in this update (all: nothing outstanding)
        timeout = float(getattr(args, 'distributed_timeout', 0) or 0)
        lib.check(self.L.hs_dp_set_timeout(timeout if timeout > 0 else 1800.0), 'hs_dp_set_timeout')
        buffer.param_gate = self.gate
        buffer.before_backward = self.before_backward
        torch.cuda.synchronize()
        dist.barrier(group=self.group, device_ids=[dev.index])

    # -- protocol --------------------------------------------------------------------------------------
    def broadcast_params(self):
        dist.broadcast(self.buffer.flat_param, src=0, group=self.group)
        torch.cuda.synchronize()

    def begin_update(self, num_micro_batches: int):
        self.epoch_t.add_(1)
        self._handles = []
        self._joined = False
        self._gated = 0
        # the previous step's update, rotated into this step: runs next to the forward pass
        self._launch_update(epoch_offset=0, flush=False)
        if not getattr(self.buffer, 'gated_forward', False) or not self.overlap_update:
            self.gate(self.buffer.total)      # this model's forward has no gates: wait for everything up front

    def bind_optimizer(self, opt):
        self._opt = opt

    def _opt_args(self, opt):
        h = opt.hyper
        kind = OPT_KIND[opt.kind]
        if kind == 0:
            a, b = h['betas']
            s1, s2 = opt.state['exp_avg'], opt.state['exp_avg_sq']
        else:
            a, b = h['rho'], 0.0
            s1, s2 = opt.state['square_avg'], opt.state['acc_delta']
        return kind, float(a), float(b), s1, s2

    def _launch_update(self, epoch_offset: int, flush: bool):
        """Side-stream half of the update. Split mode (default): the all-gather of this rank's freshly updated
        chunks (``hs_dp_allgather``; the Adam / Adadelta math already ran at the end of the previous step). Fused
        mode (``HETSEQ_DP_SPLIT=0``): update and gather in one kernel, which then reads the PREVIOUS step's
        [lr, step factor] (``hyper_prev_t``) in-step and the current ones from ``flush``. Either kernel acts only
        if the device-side ``pending`` word says something is waiting (nothing on the first step or right after
        a flush) and publishes this epoch's parameter flags in any case."""
        opt = self._opt
        if opt is None:
            raise RuntimeError('fused backend: no optimizer bound (bind_optimizer)')
        main = torch.cuda.current_stream()
        self.update_stream.wait_stream(main)
        buf = self.buffer
        if self.split_update:
            lib.check(self.L.hs_dp_allgather(
                self.param_ptrs, self.param_mc, self.sig_ptrs, self.rank, self.world_size, buf.num_blocks, CHUNK,
                self.scalars.data_ptr(), 0, self.epoch_t.data_ptr(), int(epoch_offset), self.counters.data_ptr(),
                self.pending_dev.data_ptr(), self.ngather_req, self.gather_blocks, self.update_stream.cuda_stream),
                'hs_dp_allgather')
            return
        h = opt.hyper
        kind, a, b, s1, s2 = self._opt_args(opt)
        lib.check(self.L.hs_dp_update_allgather(
            kind, buf.flat_grad.data_ptr(), opt.master.data_ptr(), s1.data_ptr(), s2.data_ptr(), self.param_ptrs,
            self.param_mc, self.sig_ptrs, self.rank, self.world_size, buf.num_blocks, CHUNK,
            lib.dtype_code(buf.param_dtype), float(h['lr']), a, b, float(h['eps']),
            float(h['weight_decay']), 1.0, self.scalars.data_ptr(),
            (opt.hyper_t if flush else opt.hyper_prev_t).data_ptr(), 0,
            self.epoch_t.data_ptr(), int(epoch_offset), self.counters.data_ptr(), self.pending_dev.data_ptr(),
            self.ngather_req, self.update_blocks, self.update_stream.cuda_stream), 'hs_dp_update_allgather')

    def gate(self, flat_end: int):
        """Forward-side wait: returns (on the current stream) once every rank's slice of the parameters up to
        flat offset ``flat_end`` has arrived. Monotone within an update; a no-op for slices already waited for."""
        if self._gated >= self.ngather:
            return
        block = max(0, (int(flat_end) - 1) // (self.world_size * CHUNK))
        j = int(self.L.hs_dp_gather_bucket_of(self.buffer.num_blocks, self.ngather_req, block))
        if j < self._gated:
            return
        lib.check(self.L.hs_dp_wait_params(self.sig_ptrs, self.rank, self.world_size, self._gated, j, 0,
                                           self.epoch_t.data_ptr(), 0, lib.stream_ptr()), 'hs_dp_wait_params')
        self._gated = j + 1

    def before_backward(self):
        """The rotated update reads the gradients of the previous step: it must be done before backward
        overwrites them (and every gate must have been passed: recomputation in backward reads parameters)."""
        if self._joined:
            return
        self._joined = True
        self.gate(self.buffer.total)
        torch.cuda.current_stream().wait_stream(self.update_stream)

    def _launch(self, b: int):
        """Queue bucket b's in-kernel reduce-scatter behind everything enqueued so far on the main stream."""
        s, e = self.buckets[b]
        self._launched[b] = True
        ev = torch.cuda.Event()
        ev.record()
        self.comm_stream.wait_event(ev)
        lib.check(self.L.hs_dp_reduce_scatter(self.grad_ptrs, self.grad_mc, self.sig_ptrs, self.rank, self.world_size,
                                              s // CHUNK, e // CHUNK, CHUNK, b, 0, self.epoch_t.data_ptr(),
                                              self.sumsq.data_ptr(), self.rs_blocks, self.comm_stream.cuda_stream),
                  'hs_dp_reduce_scatter')

    def finish_gradients(self, optimizer):
        self.before_backward()      # an update without any backward pass (all micro-batches empty) still joins
        for b in range(len(self.buckets)):
            if not self._launched[b]:
                self._launch(b)
        self._armed = False
        torch.cuda.current_stream().wait_stream(self.comm_stream)
        # no 1/W here: the norm exchange divides the SUM over ranks by the global sample size

    def norm_exchange(self, mult_host: float, mult_dev: Optional[torch.Tensor], max_norm: float,
                      sample_size: Optional[torch.Tensor] = None):
        """Global gradient norm + gradient denominator + clip coefficient + overflow flag, identical on every
        rank. ``sample_size`` (0-dim / 1-element tensor, this rank's share) switches on the in-kernel
        ``1 / sum of sample sizes``; without it the multiplier is used as given. Returns the 0-dim norm tensor;
        ``self.scalars`` / ``self.skip`` feed the update kernel."""
        md = None if mult_dev is None else mult_dev.to(torch.float32).reshape(1).contiguous()
        self._mult_keepalive = md
        ss = None
        if sample_size is None:
            mult_host = float(mult_host) / self.world_size     # plain mean over ranks, like the other backends
        else:
            self.ss_local.copy_(sample_size.reshape(1).to(torch.float32))
            ss = self.ss_local
        lib.check(self.L.hs_dp_norm_exchange(self.sig_ptrs, self.rank, self.world_size, 0, self.epoch_t.data_ptr(),
                                             self.sumsq.data_ptr(), float(mult_host), lib.ptr(md), lib.ptr(ss),
                                             float(max_norm), self.scalars.data_ptr(), self.skip.data_ptr(),
                                             self.pending_dev.data_ptr(), lib.stream_ptr()), 'hs_dp_norm_exchange')
        return self.scalars[0]

    def queue_update(self, opt):
        """Split mode: run the update on the owned chunks now (main stream) and leave the gather to the start of
        the next step; fused mode: leave both. ``flush`` brings either forward."""
        self._opt = opt
        if self.split_update:
            h = opt.hyper
            kind, a, b, s1, s2 = self._opt_args(opt)
            buf = self.buffer
            lib.check(self.L.hs_dp_update_local(
                kind, buf.flat_grad.data_ptr(), opt.master.data_ptr(), s1.data_ptr(), s2.data_ptr(),
                buf.flat_param.data_ptr(), self.rank, self.world_size, buf.num_blocks, CHUNK,
                lib.dtype_code(buf.param_dtype), float(h['lr']), a, b, float(h['eps']), float(h['weight_decay']), 1.0,
                self.scalars.data_ptr(), opt.hyper_t.data_ptr(), self.update_blocks, lib.stream_ptr()),
                'hs_dp_update_local')
        if not self.overlap_update:
            self._apply_pending()

    def _apply_pending(self):
        """Bring the pending gather (fused mode: update + gather) forward and wait for every rank's parameters;
        stream-ordered only, so it is safe inside a graph capture."""
        self._launch_update(epoch_offset=1, flush=True)
        lib.check(self.L.hs_dp_wait_params(self.sig_ptrs, self.rank, self.world_size, 0, self.ngather - 1, 0,
                                           self.epoch_t.data_ptr(), 1, self.update_stream.cuda_stream),
                  'hs_dp_wait_params')
        torch.cuda.current_stream().wait_stream(self.update_stream)

    def flush(self, sync: bool = True):
        """Make a pending update visible now and wait until every rank's parameters have arrived. Collective in
        effect (every rank reaches the same call sites: checkpoints, validation, consistency checks, close)."""
        if self._opt is None or self._capturing():
            return
        self._apply_pending()
        if sync:
            torch.cuda.current_stream().synchronize()

    @staticmethod
    def _capturing() -> bool:
        return torch.cuda.is_current_stream_capturing()

    def check_params_consistent(self) -> bool:
        self.flush()
        return super().check_params_consistent()

    # -- sharded-optimizer interface (checkpoint gathering etc.) ---------------------------------------------
    def all_gather_compact(self, compact: torch.Tensor) -> torch.Tensor:
        compact = compact.contiguous()
        out = torch.empty(self.world_size, compact.numel(), dtype=compact.dtype, device=compact.device)
        dist.all_gather_into_tensor(out.view(-1), compact, group=self.group)
        return out

    def all_gather_params(self, buffer: FlatParamBuffer):
        mine = buffer.owned(buffer.flat_param).reshape(-1).contiguous()
        buffer.flat_param.copy_(buffer.compact_to_flat(self.all_gather_compact(mine)))

    def close(self):
        try:
            self.flush()
            torch.cuda.synchronize()
            if dist.is_initialized():
                dist.barrier(group=self.group, device_ids=[self.buffer.device.index])
            self.buffer.param_gate = None
            self.buffer.before_backward = None
            self.pool.release()
        except Exception:
            pass
