#!/usr/bin/env python
"""Flattened-gradient "all-reduce + Adam" sweep (BASELINE.json config 5): for buffer sizes from 1 MB to
1.3 GB of fp32 gradient, time on the device (CUDA events, max over ranks)

  nccl   : NCCL all-reduce of the flat fp32 gradient + grad-norm + replicated fused Adam (the baseline path;
           what the reference does with DDP + its per-tensor optimizer loop, minus the launch overhead)
  fused  : in-kernel reduce-scatter + norm exchange + sharded Adam whose epilogue stores the fp16 parameters
           into every rank (no NCCL call), once with NVLS (multimem.ld_reduce / multimem.st through the switch)
           and once with plain peer loads / stores

and report the fused path against its roofline = max(bytes over NVLink / link bandwidth, shard HBM bytes /
measured HBM bandwidth). Launch: python -m torch.distributed.run --nproc-per-node N tools/allreduce_adam_sweep.py
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--sizes-mb', default='1,4,16,64,256,1024,1345')
    ap.add_argument('--iters', type=int, default=10)
    ap.add_argument('--link-gbs', type=float, default=770.0, help='measured NVLink peer bandwidth per direction')
    ap.add_argument('--hbm-gbs', type=float, default=6533.5)
    ap.add_argument('--out', default=None)
    ap.add_argument('--arms', default='nccl,fused-multimem,fused-p2p')
    a = ap.parse_args()
    rank, world, local = int(os.environ['RANK']), int(os.environ['WORLD_SIZE']), int(os.environ['LOCAL_RANK'])
    torch.cuda.set_device(local)
    dist.init_process_group('nccl', device_id=torch.device('cuda', local))
    from hetseq_b200 import options
    from hetseq_b200.optim import _Adam
    from hetseq_b200.parallel import data_parallel, fused_dp
    from hetseq_b200.parallel.flat_buffer import FlatParamBuffer
    dev = torch.device('cuda', local)
    rows = []
    for mb in [float(x) for x in a.sizes_mb.split(',')]:
        n = int(mb * (1 << 20) / 4)
        res = {'grad_mb': mb, 'n_params': n, 'world': world}
        for backend in a.arms.split(','):
            fused = backend.startswith('fused')
            if fused:
                os.environ['HETSEQ_DP_TRANSPORT'] = backend.split('-')[1]
            args = options.make_args('mnist', 'adam', argv=['--lr', '1e-4', '--fp16', '--distributed-world-size', str(world),
                                                              '--distributed-rank', str(rank), '--dp-backend', 'fused' if fused else 'nccl',
                                                              '--bucket-cap-mb', '64'])
            p = torch.nn.Parameter(torch.zeros(n, dtype=torch.float16, device=dev))
            buf = FlatParamBuffer([p], world_size=world, rank=rank,
                                  allocator=fused_dp.symmetric_allocator(args) if fused else None)
            dp = fused_dp.FusedBackend(args, buf) if fused else data_parallel.AllReduceBackend(args, buf)
            opt = _Adam(args, buf, dp=dp)
            ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            eva, evb = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            times, parts = [], []
            for it in range(a.iters + 3):
                opt.zero_grad()
                buf.flat_grad.normal_(0, 1e-3)                 # "backward": every rank has its own gradient
                dp.begin_update(1)
                dp.begin_micro_batch(0, True)
                torch.cuda.synchronize()
                dist.barrier(device_ids=[local])
                ev0.record()
                dp.finish_gradients(opt)
                eva.record()
                opt.clip_grad_norm(1.0)
                evb.record()
                opt.launch_update()
                if fused:
                    dp.flush(sync=False)          # the update normally rides next to the next forward pass
                ev1.record()
                torch.cuda.synchronize()
                opt.finish_update(False)
                if it >= 3:
                    times.append(ev0.elapsed_time(ev1))
                    parts.append((ev0.elapsed_time(eva), eva.elapsed_time(evb), evb.elapsed_time(ev1)))
            t = torch.tensor([min(times)], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            res[backend + '_ms'] = round(float(t.item()), 4)
            best = min(range(len(times)), key=lambda k: times[k])
            res[backend + '_parts_ms'] = [round(x, 3) for x in parts[best]]   # [reduce, norm, update(+gather)] on rank 0
            dp.close()
            del opt, dp, buf, p
            torch.cuda.empty_cache()
        # roofline of the fused path per GPU: wire = (W-1)/W * N * (4 in + 2 out) bytes per direction-bound stream
        nbytes = n * 4
        wire_in = (world - 1) / world * nbytes                       # fp32 gradients pulled from peers
        wire_out = (world - 1) / world * n * 2                       # fp16 parameters pushed to peers
        t_link = max(wire_in, wire_out) / (a.link_gbs * 1e9) * 1e3    # in and out travel opposite directions
        hbm = n / world * (4 * 4 + 3 * 4 + 2) + nbytes / world       # m,v,master,grad read; m,v,master write; param; RS local
        t_hbm = hbm / (a.hbm_gbs * 1e9) * 1e3
        res['roofline_ms'] = round(max(t_link, t_hbm), 4)
        for arm in a.arms.split(','):
            if arm.startswith('fused'):
                res[arm + '_roofline_frac'] = round(res['roofline_ms'] / res[arm + '_ms'], 3)
                if 'nccl_ms' in res:
                    res[arm + '_speedup_vs_nccl'] = round(res['nccl_ms'] / res[arm + '_ms'], 3)
        rows.append(res)
        if rank == 0:
            print(json.dumps(res), flush=True)
    if rank == 0 and a.out:
        with open(a.out, 'w') as f:
            json.dump(rows, f, indent=1)
    dist.destroy_process_group()


if __name__ == '__main__':
    main()
