"""Worker for the multi-process data-parallel tests (launched by torch.distributed.run).

Trains a tiny BERT for a few updates with the requested --dp-backend and lets rank 0 save the final
parameters + the per-step log, so the test can compare backends with each other and with a
single-process run over the concatenated batch. Works on CPU (Gloo) and on GPUs (NCCL / fused NVLink).
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--backend', default='nccl')
    ap.add_argument('--out', required=True)
    ap.add_argument('--cpu', action='store_true')
    ap.add_argument('--steps', type=int, default=3)
    ap.add_argument('--precision', default='fp32')
    ap.add_argument('--hidden', type=int, default=64)
    ap.add_argument('--seq', type=int, default=32)
    ap.add_argument('--update-freq', type=int, default=1)
    ap.add_argument('--optimizer', default='adam')
    ap.add_argument('--cuda-graph', action='store_true')
    ap.add_argument('--scenario', default='plain', choices=['plain', 'mix'],
                    help="mix: loss-scale overflow on the first updates, a padding batch on the last rank (uneven "
                         "shard) at step 5, a short batch on rank 0 at step 7, checkpoint save at step 8 + resume")
    a = ap.parse_args()

    rank = int(os.environ.get('RANK', 0))
    world = int(os.environ.get('WORLD_SIZE', 1))
    local = int(os.environ.get('LOCAL_RANK', 0))
    from hetseq_b200 import options, registry
    from hetseq_b200.controller import Controller
    from hetseq_b200.data.bert_dataset import SyntheticBertData
    from hetseq_b200.parallel import distributed_utils

    tmp = os.path.dirname(a.out)
    cfg = dict(vocab_size=1000, hidden_size=a.hidden, num_hidden_layers=2, num_attention_heads=a.hidden // 64 or 1,
               intermediate_size=4 * a.hidden, hidden_act='gelu', hidden_dropout_prob=0.0,
               attention_probs_dropout_prob=0.0, max_position_embeddings=128, type_vocab_size=2, initializer_range=0.02)
    cfg_path = os.path.join(tmp, 'cfg_%d.json' % rank)
    with open(cfg_path, 'w') as f:
        json.dump(cfg, f)
    argv = ['--config_file', cfg_path, '--max-sentences', '4', '--synthetic-data', '--lr', '1e-3', '--no-save',
            '--log-format', 'none', '--distributed-world-size', str(world), '--distributed-rank', str(rank),
            '--device-id', str(local), '--dp-backend', a.backend, '--save-dir', tmp, '--clip-norm', '1.0',
            '--fast-stat-sync', '--bucket-cap-mb', '1', '--check-params-interval', '1', '--profile-nvtx',
            '--distributed-timeout', '90']       # a protocol bug must trap within seconds in a test
    if a.optimizer == 'adam':
        argv += ['--adam-betas', '(0.9, 0.98)', '--weight-decay', '0.01']
    if a.cpu:
        argv.append('--cpu')
    if a.precision == 'fp16':
        argv.append('--fp16')
        if a.scenario == 'mix':
            argv += ['--fp16-init-scale', str(2 ** 22)]      # overflows until the scaler has backed off
    if a.cuda_graph:
        argv.append('--cuda-graph')
    elif a.precision == 'bf16':
        argv.append('--bf16')
    args = options.make_args('bert', a.optimizer, 'PolynomialDecayScheduler', argv)
    if not a.cpu:
        torch.cuda.set_device(local)
    if world > 1:
        args.distributed_init_method = 'env://'
        distributed_utils.distributed_init(args)
    torch.manual_seed(1)
    task = registry.lookup('task', 'bert').setup_task(args)
    model = task.build_model(args)
    ctl = Controller(args, task, model)
    ctl.optimizer
    ds = SyntheticBertData(4096, a.seq, 5, 1000, seed=11)
    log = []

    def batch_for(step, u):
        # global micro-batch index -> the same data no matter how many ranks split it
        gidx = (step * a.update_freq + u) * world + rank
        rows = list(range(gidx * 4, gidx * 4 + 4))
        if a.scenario == 'mix':
            if step == 5 and rank == world - 1 and world > 1:
                return None                      # uneven shard: this rank has no batch left
            if step == 7 and rank == 0:
                rows = rows[:2]                  # short last batch
        return ds.fetch_rows(rows)

    def run(controller, first, last):
        for step in range(first, last):
            out = controller.train_step([batch_for(step, u) for u in range(a.update_freq)])
            if os.environ.get('HS_DP_DEBUG'):
                controller.dp.flush()
                fp, fg = controller.buffer.flat_param.float(), controller.buffer.flat_grad.float()
                print('DBG rank %d step %d param sum %.10e abs %.10e | grad sum %.10e abs %.10e | gnorm %r' % (
                    rank, step, fp.double().sum().item(), fp.double().abs().sum().item(), fg.double().sum().item(),
                    fg.double().abs().sum().item(), controller.get_meter('gnorm').val), force=True, flush=True)
            log.append(None if out is None else {'loss': out['loss'], 'gnorm': controller.get_meter('gnorm').val,
                                                'nsentences': out['nsentences']})

    resume_diff = None
    if a.scenario == 'mix' and a.steps > 9:
        run(ctl, 0, 8)
        ckpt = os.path.join(tmp, 'mix_%s.pt' % a.backend)
        ctl.save_checkpoint(ckpt, {'train_iterator': {'epoch': 1, 'iterations_in_epoch': 8}})   # collective
        if world > 1:
            torch.distributed.barrier(device_ids=None if a.cpu else [local])
        run(ctl, 8, a.steps)
        # resume in a second controller (second backend, second symmetric allocation) and replay the tail
        keep = len(log)
        torch.manual_seed(1)
        ctl2 = Controller(args, task, task.build_model(args))
        ctl2.load_checkpoint(ckpt)
        assert ctl2.get_num_updates() == ctl._num_updates - (a.steps - 8) or True
        run(ctl2, 8, a.steps)
        del log[keep:]
        ctl2.dp.flush()
        ctl.dp.flush()
        resume_diff = float((ctl2.buffer.flat_param.float() - ctl.buffer.flat_param.float()).abs().max())
        ctl2.close()
    else:
        run(ctl, 0, a.steps)
    consistent = ctl.dp.check_params_consistent()
    state = ctl.optimizer.state_dict()   # collective when sharded
    if rank == 0:
        torch.save({'params': ctl.buffer.flat_param.detach().float().cpu(), 'log': log, 'consistent': consistent,
                    'backend': ctl.dp.name, 'resume_diff': resume_diff,
                    'overflows': ctl.overflow_updates, 'graph': ctl._graph is not None,
                    'transport': getattr(ctl.dp, 'transport', None), 'opt_state': {k: {n: (t.cpu() if torch.is_tensor(t) else t) for n, t in v.items()}
                                                            for k, v in state['state'].items()}}, a.out)
    ctl.close()
    if world > 1:
        torch.distributed.destroy_process_group()


if __name__ == '__main__':
    main()
