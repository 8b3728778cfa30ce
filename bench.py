#!/usr/bin/env python
"""Headline benchmark: BERT pre-training throughput (sequences/s), seq 128, on N B200s of one node.

    python bench.py --gpus 1 --steps 20 --warmup 5
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 \
        --master-port 29500 bench.py --gpus 8 --steps 20 --warmup 5
    python bench.py --impl reference ...      # the unmodified reference from baseline/_ref

Config (BASELINE.json): BERT-large phase-1 (seq 128, 20 predictions), 32 sequences per GPU (global batch
256 on 8 GPUs -> weak scaling), random-init weights, synthetic HDF5-shaped batches, fp16 with fp32
master weights / Adam state and dynamic loss scaling; every step = forward + backward + gradient
reduction + clip + Adam update, dropout on. Prints ONE JSON line on rank 0.

Timing: W >= 3 untimed warm-up steps; exactly K timed steps bracketed by barrier + synchronize on both
sides, CUDA events on the training stream, max over ranks. "value" feeds device-resident batches;
"e2e" goes through the same public call (Controller.train_step) but every step copies its batch from
pinned host memory and reads the loss back to the host. The per-step working set (fp16 weights +
fp32 gradients + master/Adam state + activations, > 5 GB for BERT-large) is far larger than the 126 MB
L2, so no explicit flush is needed between iterations; "config.l2" says so.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

MODELS = {
    'bert-large': dict(hidden_size=1024, num_hidden_layers=24, num_attention_heads=16, intermediate_size=4096),
    'bert-base': dict(hidden_size=768, num_hidden_layers=12, num_attention_heads=12, intermediate_size=3072),
    'bert-tiny': dict(hidden_size=128, num_hidden_layers=2, num_attention_heads=2, intermediate_size=512),
}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=20)
    ap.add_argument('--warmup', type=int, default=5)
    ap.add_argument('--impl', default='ours', choices=['ours', 'reference'])
    ap.add_argument('--model', default='bert-large', choices=sorted(MODELS))
    ap.add_argument('--batch', type=int, default=32, help='sequences per GPU per step')
    ap.add_argument('--seq-len', type=int, default=128)
    ap.add_argument('--max-predictions', type=int, default=None)
    ap.add_argument('--dtype', default='fp16', choices=['fp16', 'bf16', 'fp32'])
    ap.add_argument('--dp-backend', default='auto', choices=['auto', 'fused', 'nccl'])
    ap.add_argument('--kernels', default='auto', choices=['auto', 'native', 'torch'])
    ap.add_argument('--cuda-graph', dest='cuda_graph', action='store_true', default=True,
                    help='replay the captured update from a CUDA graph (default)')
    ap.add_argument('--no-cuda-graph', dest='cuda_graph', action='store_false')
    ap.add_argument('--bucket-cap-mb', type=int, default=64)
    ap.add_argument('--no-e2e', action='store_true')
    ap.add_argument('--no-dp-check', action='store_true',
                    help='skip the untimed multi-GPU correctness check (replica equality, fused vs NCCL backend)')
    return ap.parse_args()


def write_config(model_name, tmpdir, dropout=0.1, name='bert_config.json'):
    cfg = dict(vocab_size=30522, hidden_act='gelu', hidden_dropout_prob=dropout, attention_probs_dropout_prob=dropout,
               max_position_embeddings=512, type_vocab_size=2, initializer_range=0.02)
    cfg.update(MODELS[model_name])
    path = os.path.join(tmpdir, name)
    with open(path, 'w') as f:
        json.dump(cfg, f)
    vocab = os.path.join(tmpdir, 'vocab.txt')
    with open(vocab, 'w') as f:
        f.write('\n'.join('tok%d' % i for i in range(64)))
    return path, vocab


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons with nvidia-smi while the timed region runs."""

    FIELDS = ('clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,'
              'clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,'
              'clocks_event_reasons.sw_power_cap')

    def __init__(self, index=0):
        super().__init__(daemon=True)
        self.index = index
        self.samples = []
        self._halt = threading.Event()

    def run(self):
        while not self._halt.is_set():
            try:
                out = subprocess.run(['nvidia-smi', '-i', str(self.index), '--query-gpu=' + self.FIELDS,
                                      '--format=csv,noheader,nounits'], capture_output=True, text=True, timeout=5).stdout
                parts = [p.strip() for p in out.strip().split(',')]
                if len(parts) >= 7:
                    self.samples.append(parts)
            except Exception:
                pass
            self._halt.wait(0.2)

    def stop(self):
        self._halt.set()
        self.join(timeout=6)

    def summary(self):
        if not self.samples:
            return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': ['unsampled']}
        sm = [float(s[0]) for s in self.samples if s[0].replace('.', '').isdigit()]
        mx = [float(s[1]) for s in self.samples if s[1].replace('.', '').isdigit()]
        names = ['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']
        reasons = [n for k, n in enumerate(names) if any(s[3 + k].lower().startswith('active') for s in self.samples)]
        return {'sm_mhz': statistics.median(sm) if sm else None, 'sm_max_mhz': max(mx) if mx else None,
                'reasons': reasons, 'samples': len(self.samples)}


def dist_env(a):
    rank = int(os.environ.get('RANK', 0))
    world = int(os.environ.get('WORLD_SIZE', 1))
    local = int(os.environ.get('LOCAL_RANK', 0))
    if world != a.gpus and world > 1:
        a.gpus = world
    return rank, world, local


def max_over_ranks(ms, world, device):
    import torch
    import torch.distributed as dist
    if world <= 1:
        return ms
    t = torch.tensor([ms], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def barrier_sync(world):
    import torch
    import torch.distributed as dist
    if world > 1:
        dist.barrier(device_ids=[torch.cuda.current_device()])
    torch.cuda.synchronize()


def make_batches(a, n, vocab, seed):
    """n collated host batches (list of five int64 tensors), pinned."""
    from hetseq_b200.data.bert_dataset import SyntheticBertData
    maxp = a.max_predictions or (20 if a.seq_len <= 128 else 80)
    ds = SyntheticBertData(n * a.batch, a.seq_len, maxp, vocab, seed=seed)
    out = []
    for i in range(n):
        b = ds.fetch_rows(list(range(i * a.batch, (i + 1) * a.batch)))
        out.append([t.pin_memory() for t in b])
    return out


def timed_loop(step_fn, batches, warmup, steps, world, device, after_warmup=None):
    import torch
    for i in range(warmup):
        step_fn(batches[i % len(batches)])
    if after_warmup is not None:
        after_warmup()
    barrier_sync(world)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps):
        step_fn(batches[(warmup + i) % len(batches)])
    e1.record()
    barrier_sync(world)
    return max_over_ranks(e0.elapsed_time(e1), world, device)


# --------------------------------------------------------------------------------------------------
def run_ours(a):
    import torch
    rank, world, local = dist_env(a)
    torch.cuda.set_device(local)
    device = torch.device('cuda', local)
    tmp = tempfile.mkdtemp(prefix='hsbench')
    cfg_path, vocab_path = write_config(a.model, tmp)

    from hetseq_b200 import options, registry
    from hetseq_b200.controller import Controller
    from hetseq_b200.ops import lib
    from hetseq_b200.parallel import distributed_utils

    def build(cfg, dp_backend, cuda_graph):
        argv = ['--config_file', cfg, '--max-sentences', str(a.batch), '--synthetic-data',
                '--synthetic-seq-len', str(a.seq_len), '--lr', '1e-4', '--warmup-updates', '100',
                '--total-num-update', '100000', '--weight-decay', '0.01', '--clip-norm', '1.0',
                '--adam-betas', '(0.9, 0.999)', '--no-save', '--log-format', 'none', '--fast-stat-sync',
                '--distributed-world-size', str(world), '--distributed-rank', str(rank), '--device-id', str(local),
                '--dp-backend', dp_backend, '--kernels', a.kernels, '--save-dir', tmp,
                '--bucket-cap-mb', str(a.bucket_cap_mb)]
        if a.dtype == 'fp16':
            argv.append('--fp16')
        elif a.dtype == 'bf16':
            argv.append('--bf16')
        if cuda_graph:
            argv.append('--cuda-graph')
        args_ = options.make_args('bert', 'adam', 'PolynomialDecayScheduler', argv)
        if world > 1:
            args_.distributed_init_method = 'env://'
            if not torch.distributed.is_initialized():
                distributed_utils.distributed_init(args_)
        torch.manual_seed(args_.seed)
        task_ = registry.lookup('task', 'bert').setup_task(args_)
        model_ = task_.build_model(args_)
        ctl = Controller(args_, task_, model_)
        ctl.optimizer  # build engine
        return args_, task_, model_, ctl

    args, task, model, controller = build(cfg_path, a.dp_backend, a.cuda_graph)

    nb = 8
    host = make_batches(a, nb, model.config.vocab_size, seed=1234 + rank)
    dev = [task.prepare_sample(b, device) for b in host]
    # what actually crosses PCIe per step: the five batch tensors narrowed to int32 + the MLM row index/labels
    h2d_bytes = sum(t.numel() * t.element_size() for t in dev[0])

    def step_device(batch):
        # device-resident batch: _prepare_sample's copy is a no-op
        controller.train_step([batch])

    # end to end = what hetseq_b200.train does on a GPU: host batches flow through the framework's PrefetchIterator
    # (Task.host_prepare in its worker thread: MLM row index, int32 narrowing, pinning), then Controller.train_step
    # copies the pinned batch to the device (every step, inside the timed region) and reads the result back
    import itertools
    from hetseq_b200.data.iterators import PrefetchIterator
    e2e_source = PrefetchIterator(itertools.cycle(host), depth=3, transform=task.host_prepare)

    def step_e2e(_unused):
        out = controller.train_step([next(e2e_source)])
        return out

    # kernels of ours per update: counted on one eager update (a CUDA-graph replay re-issues exactly the
    # kernels that were captured, without going through the launchers' counter)
    step_device(dev[0])          # first update: lazy initialisation, not counted
    torch.cuda.synchronize()
    prof = None
    try:                         # every kernel of one eager update, whoever launched it (CUPTI)
        from torch.profiler import ProfilerActivity, profile
        prof = profile(activities=[ProfilerActivity.CUDA])
        prof.__enter__()
    except Exception:            # noqa: BLE001 -- no profiler: only the library's own counter is reported
        prof = None
    c0 = lib.launch_count() if lib.available() else 0
    step_device(dev[1 % len(dev)])
    torch.cuda.synchronize()
    launches_per_step = (lib.launch_count() - c0) if lib.available() else 0
    other_launches = None
    if prof is not None:
        try:
            prof.__exit__(None, None, None)
            total = sum(int(k.count) for k in prof.key_averages()
                        if k.device_time_total > 0 and 'Memcpy' not in k.key and 'Memset' not in k.key)
            other_launches = max(0, total - int(launches_per_step))
            dump = os.environ.get('HETSEQ_BENCH_KERNEL_LIST')      # diagnosis: every kernel of that update, by name
            if dump and int(os.environ.get('RANK', '0')) == 0:
                rows = sorted(((k.key, int(k.count), float(k.device_time_total)) for k in prof.key_averages()
                               if k.device_time_total > 0), key=lambda r: -r[2])
                with open(dump, 'w') as f:
                    f.write('count\tdevice_us\tkernel (one eager update, CUPTI)\n')
                    for name, cnt, us in rows:
                        f.write('%d\t%.1f\t%s\n' % (cnt, us, name[:160]))
        except Exception:        # noqa: BLE001
            other_launches = None

    e2e_first = os.environ.get('HETSEQ_BENCH_E2E_FIRST', '0') == '1' and not a.no_e2e
    if e2e_first:      # diagnosis: does the order of the two timed loops matter (power / thermal state)?
        ms_e2e = timed_loop(step_e2e, host, a.warmup, a.steps, world, device)
    sampler = ClockSampler(local) if rank == 0 else None
    if sampler:
        sampler.start()
    # updates skipped for a non-finite gradient norm are ~1.6 ms cheaper than real ones: count them (the dynamic
    # loss scale backs off during warm-up; any skip inside the timed region is reported as overflow_steps)
    mark = {}
    ms = timed_loop(step_device, dev, a.warmup, a.steps, world, device,
                    after_warmup=lambda: mark.update(overflow=controller.overflow_updates))
    overflow_before = mark['overflow']
    if sampler:
        sampler.stop()
    launches_timed = int(launches_per_step * a.steps)
    overflow_timed = controller.overflow_updates - overflow_before

    e2e = None
    if not a.no_e2e:
        if not e2e_first:
            ms_e2e = timed_loop(step_e2e, host, 3, a.steps, world, device)
        e2e = {'value': round(world * a.batch * a.steps / (ms_e2e / 1e3), 2), 'unit': 'sequences/s',
               'ms_per_step': round(ms_e2e / a.steps, 3), 'h2d_bytes_per_step': int(h2d_bytes),
               'd2h_bytes_per_step': 8 * 8}

    # ---- untimed: is the multi-GPU product correct? ----------------------------------------------------------
    dp_check = None
    if world > 1:
        controller.dp.flush()
        dp_check = {'params_consistent': bool(controller.dp.check_params_consistent()),
                    'transport': getattr(controller.dp, 'transport', None)}
        if not a.no_dp_check and controller.dp.name == 'fused-nvlink':
            # the same 6 updates (3 eager, capture, replays) from the same initial weights through the fused
            # NVLink path and through NCCL all-reduce + replicated Adam; dropout off so the two runs see the same
            # function (dropout sites are numbered per model instance)
            cfg0, _ = write_config(a.model, tmp, dropout=0.0, name='bert_config_nodrop.json')
            finals = []
            for backend, graph in (('nccl', False), (a.dp_backend, a.cuda_graph)):
                _, t2, _, c2 = build(cfg0, backend, graph)
                for i in range(6):
                    c2.train_step([t2.prepare_sample(host[i % len(host)], device)])
                c2.dp.flush()
                finals.append((c2.buffer.flat_param[:c2.buffer.used].float().clone(), c2.dp.name,
                               bool(c2.dp.check_params_consistent())))
                c2.close()
                del c2
                torch.cuda.empty_cache()
            diff = (finals[0][0] - finals[1][0]).abs()
            dp_check.update({'max_abs_param_diff_vs_nccl_backend_after_6_updates': float(diff.max()),
                             'mean_abs_param_diff': float(diff.mean()), 'compared': [finals[0][1], finals[1][1]],
                             'both_consistent': finals[0][2] and finals[1][2]})

    if rank == 0:
        value = world * a.batch * a.steps / (ms / 1e3)
        baseline = published_baseline(world)
        out = {
            'metric': 'BERT seq%d sequences/sec' % a.seq_len, 'value': round(value, 2), 'unit': 'sequences/s',
            'n_gpus': world, 'steps': a.steps, 'warmup': a.warmup, 'ms_per_step': round(ms / a.steps, 3),
            'higher_is_better': True, 'scaling': 'weak',
            'vs_baseline': round(value / baseline, 3) if baseline else None,
            'dtype': a.dtype, 'data': 'synthetic (HDF5 phase-%d shaped batches, random-init weights)' % (1 if a.seq_len <= 128 else 2),
            'impl': 'ours',
            'config': {'model': a.model, 'global_batch': world * a.batch, 'per_gpu_batch': a.batch, 'seq_len': a.seq_len,
                       'parallelism': 'dp%d' % world, 'dp_backend': controller.dp.name, 'optimizer': 'adam',
                       'kernels': 'native sm_100a' if controller._model.native is not None else 'torch',
                       'l2': 'per-step working set (weights+grads+Adam state+activations) >> 126 MB L2; no flush needed',
                       'cuda_graph': bool(a.cuda_graph)},
            'clocks': sampler.summary() if sampler else None,
            'gpu_launches': launches_timed, 'gpu_launches_per_step': launches_per_step,
            'overflow_steps': int(overflow_timed),
            # kernels of one eager update that are NOT this library's (ATen elementwise / reductions / copies, NCCL)
            'other_kernel_launches_per_step': other_launches,
        }
        if dp_check is not None:
            out['dp_check'] = dp_check
        if e2e:
            out['e2e'] = e2e
        print(json.dumps(out), flush=True)
    controller.close()
    if world > 1:
        torch.distributed.destroy_process_group()


def published_baseline(world):
    """seq/s derived from the reference README's table (batch / s-per-step); only 4..32 GPUs exist,
    hardware unspecified. For other GPU counts there is no published number."""
    table = {4: 49.2, 8: 95.2, 16: 183.3, 32: 327.6}
    return table.get(world)


# --------------------------------------------------------------------------------------------------
def run_reference(a):
    """The unmodified reference (baseline/_ref): its options parser, LanguageModelingTask, BertForPreTraining,
    Controller.train_step (DDP + its Adam). fp32 -- the reference has no fp16 path (options.py:11)."""
    ref = os.path.join(ROOT, 'baseline', '_ref')
    if not os.path.exists(os.path.join(ref, 'controller.py')):
        print(json.dumps({'impl': 'reference', 'unavailable': 'baseline/_ref not installed (run baseline/install_reference.sh)'}))
        return
    try:
        import types
        import torch
        # import-time dependencies of the reference that are not in this image and are never exercised here
        for name in ('boto3', 'botocore', 'botocore.exceptions', 'h5py'):
            if name not in sys.modules:
                try:
                    __import__(name)
                except Exception:
                    m = types.ModuleType(name)
                    if name == 'botocore.exceptions':
                        m.ClientError = type('ClientError', (Exception,), {})
                    sys.modules[name] = m
        sys.path.insert(0, ref)
        rank, world, local = dist_env(a)
        torch.cuda.set_device(local)
        device = torch.device('cuda', local)
        tmp = tempfile.mkdtemp(prefix='hsref')
        cfg_path, vocab_path = write_config(a.model, tmp)

        import options as ref_options        # noqa: E402  (reference modules, flat imports)
        import tasks as ref_tasks            # noqa: E402
        import distributed_utils as ref_dist  # noqa: E402
        from controller import Controller as RefController  # noqa: E402

        parser = ref_options.get_training_parser(task='bert', optimizer='adam', lr_scheduler='PolynomialDecayScheduler')
        argv = ['--task', 'bert', '--data', tmp, '--dict', vocab_path, '--config_file', cfg_path,
                '--max-sentences', str(a.batch), '--lr', '1e-4', '--warmup-updates', '100',
                '--total-num-update', '100000', '--weight-decay', '0.01', '--clip-norm', '1.0',
                '--adam-betas', '(0.9, 0.999)', '--no-save', '--fast-stat-sync', '--save-dir', tmp,
                '--distributed-world-size', str(world), '--distributed-rank', str(rank), '--device-id', str(local)]
        args = ref_options.parse_args_and_arch(parser, argv)
        if world > 1:
            args.distributed_init_method = 'env://'
            ref_dist.distributed_init(args)
        torch.manual_seed(args.seed)
        task = ref_tasks.LanguageModelingTask.setup_task(args)
        model = task.build_model(args)
        controller = RefController(args, task, model)

        nb = 8
        host = make_batches(a, nb, model.config.vocab_size, seed=1234 + rank)
        dev = [[t.cuda(non_blocking=True) for t in b] for b in host]
        h2d_bytes = sum(t.numel() * t.element_size() for t in host[0])

        def step(batch):
            controller.train_step([batch])

        sampler = ClockSampler(local) if rank == 0 else None
        if sampler:
            sampler.start()
        ms = timed_loop(step, dev, a.warmup, a.steps, world, device)
        if sampler:
            sampler.stop()
        e2e = None
        if not a.no_e2e:
            ms_e2e = timed_loop(step, host, 2, a.steps, world, device)
            e2e = {'value': round(world * a.batch * a.steps / (ms_e2e / 1e3), 2), 'unit': 'sequences/s',
                   'ms_per_step': round(ms_e2e / a.steps, 3), 'h2d_bytes_per_step': int(h2d_bytes),
                   'd2h_bytes_per_step': 6 * 8}
        if rank == 0:
            value = world * a.batch * a.steps / (ms / 1e3)
            baseline = published_baseline(world)
            out = {'metric': 'BERT seq%d sequences/sec' % a.seq_len, 'value': round(value, 2), 'unit': 'sequences/s',
                   'n_gpus': world, 'steps': a.steps, 'warmup': a.warmup, 'ms_per_step': round(ms / a.steps, 3),
                   'higher_is_better': True, 'scaling': 'weak',
                   'vs_baseline': round(value / baseline, 3) if baseline else None, 'dtype': 'fp32',
                   'data': 'synthetic (HDF5 phase-1 shaped batches, random-init weights)', 'impl': 'reference',
                   'config': {'model': a.model, 'global_batch': world * a.batch, 'per_gpu_batch': a.batch,
                              'seq_len': a.seq_len, 'parallelism': 'dp%d' % world, 'dp_backend': 'torch DDP + NCCL',
                              'note': 'unmodified reference; it has no fp16 path, runs fp32',
                              'l2': 'per-step working set >> 126 MB L2; no flush needed'},
                   'clocks': sampler.summary() if sampler else None, 'gpu_launches': 0}
            if e2e:
                out['e2e'] = e2e
            # the reference silences print on non-master ranks only; rank 0 prints normally
            print(json.dumps(out), flush=True)
        if world > 1:
            torch.distributed.destroy_process_group()
    except Exception as e:  # the contract: report unavailability in one line, exit 0
        import traceback
        traceback.print_exc(file=sys.stderr)
        if int(os.environ.get('RANK', 0)) == 0:
            print(json.dumps({'impl': 'reference', 'unavailable': '%s: %s' % (type(e).__name__, str(e)[:200])}))


if __name__ == '__main__':
    a = parse()
    if a.impl == 'reference':
        run_reference(a)
    else:
        run_ours(a)
