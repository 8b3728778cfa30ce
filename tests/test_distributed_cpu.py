"""Multi-process data parallelism on CPU (Gloo): baseline all-reduce backend, sharded-protocol backend,
equivalence with a single process over the concatenated batch, uneven shards, stats exchange."""
import os
import subprocess
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
WORKER = os.path.join(ROOT, 'tests', 'dp_worker.py')


def _run(tmp_path, nproc, backend, tag, extra=()):
    out = os.path.join(str(tmp_path), 'out_%s.pt' % tag)
    port = 29500 + (os.getpid() * 7 + hash(tag)) % 2000
    cmd = [sys.executable]
    if nproc > 1:
        cmd += ['-m', 'torch.distributed.run', '--nnodes=1', '--nproc-per-node', str(nproc), '--master-addr',
                '127.0.0.1', '--master-port', str(port)]
    cmd += [WORKER, '--backend', backend, '--out', out, '--cpu', *extra]
    env = dict(os.environ, OMP_NUM_THREADS='2')
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    return torch.load(out, weights_only=False)


def test_allreduce_and_sharded_backends_match_single_process(tmp_path):
    """2 ranks x 4 sequences == 1 process x (2 micro-batches of 4): same parameters after 3 updates."""
    single = _run(tmp_path, 1, 'nccl', 'single', ['--update-freq', '2'])
    ar = _run(tmp_path, 2, 'nccl', 'ar')
    sh = _run(tmp_path, 2, 'sharded', 'sh')
    assert ar['backend'] == 'nccl' and sh['backend'] == 'sharded-allreduce'
    assert ar['consistent'] and sh['consistent']
    torch.testing.assert_close(ar['params'], single['params'], rtol=1e-4, atol=2e-5)
    torch.testing.assert_close(sh['params'], ar['params'], rtol=1e-5, atol=1e-6)
    for a, b in zip(ar['log'], single['log']):
        assert abs(a['loss'] - b['loss']) < 1e-4 and abs(a['gnorm'] - b['gnorm']) < 1e-3 * max(1.0, b['gnorm'])
        assert a['nsentences'] == 8
    # the sharded optimizer's gathered state equals the replicated one (checkpoint interchangeability)
    for k in ar['opt_state']:
        for name in ('exp_avg', 'exp_avg_sq'):
            torch.testing.assert_close(sh['opt_state'][k][name], ar['opt_state'][k][name], rtol=1e-5, atol=1e-7)


def test_adadelta_sharded(tmp_path):
    ar = _run(tmp_path, 2, 'nccl', 'ar_ad', ['--optimizer', 'adadelta'])
    sh = _run(tmp_path, 2, 'sharded', 'sh_ad', ['--optimizer', 'adadelta'])
    torch.testing.assert_close(sh['params'], ar['params'], rtol=1e-5, atol=1e-6)


def test_all_gather_list_and_uneven_shards(tmp_path):
    """3 ranks over Gloo: pickled stats exchange keeps the reference's contract (fixed-size slots, ValueError
    beyond max_size) and a rank whose shard ran out of batches trains on the dummy batch with zeroed gradients
    so that nobody hangs (reference controller.py:237-241)."""
    script = os.path.join(str(tmp_path), 'w.py')
    with open(script, 'w') as f:
        f.write('''
import os, sys, json
sys.path.insert(0, %r)
import torch, torch.distributed as dist
from hetseq_b200 import options, registry
from hetseq_b200.controller import Controller
from hetseq_b200.data.bert_dataset import SyntheticBertData
from hetseq_b200.parallel import distributed_utils as du
rank, world = int(os.environ['RANK']), int(os.environ['WORLD_SIZE'])
tmp = %r
cfg = dict(vocab_size=500, hidden_size=32, num_hidden_layers=1, num_attention_heads=2, intermediate_size=64, hidden_act='gelu',
           hidden_dropout_prob=0.0, attention_probs_dropout_prob=0.0, max_position_embeddings=64, type_vocab_size=2, initializer_range=0.02)
p = os.path.join(tmp, 'c%%d.json' %% rank); json.dump(cfg, open(p, 'w'))
args = options.make_args('bert', 'adam', argv=['--config_file', p, '--max-sentences', '4', '--synthetic-data', '--lr', '1e-3', '--no-save',
    '--log-format', 'none', '--cpu', '--distributed-world-size', str(world), '--distributed-rank', str(rank), '--save-dir', tmp])
args.distributed_init_method = 'env://'
du.distributed_init(args)
got = du.all_gather_list({'rank': rank, 'payload': 'x' * (10 * rank)})
assert [g['rank'] for g in got] == list(range(world)), got
try:
    du.all_gather_list('y' * 20000)
    raise SystemExit('expected ValueError')
except ValueError:
    pass
torch.manual_seed(0)
task = registry.lookup('task', 'bert').setup_task(args)
ctl = Controller(args, task, task.build_model(args))
ds = SyntheticBertData(64, 16, 3, 500, seed=2)
out1 = ctl.train_step([ds.fetch_rows(list(range(rank * 4, rank * 4 + 4)))])
# second update: rank 2 has no batch left (padding batch = empty list)
sample = [] if rank == 2 else ds.fetch_rows(list(range(20 + rank * 4, 24 + rank * 4)))
out2 = ctl.train_step([sample])
assert out1['nsentences'] == 12 and out2['nsentences'] == 8, (out1, out2)
assert ctl.dp.check_params_consistent()
if rank == 0:
    open(os.path.join(tmp, 'ok'), 'w').write('ok')
dist.destroy_process_group()
''' % (ROOT, str(tmp_path)))
    port = 29500 + (os.getpid() * 13) % 2000
    r = subprocess.run([sys.executable, '-m', 'torch.distributed.run', '--nnodes=1', '--nproc-per-node', '3', '--master-addr',
                        '127.0.0.1', '--master-port', str(port), script], capture_output=True, text=True, timeout=600,
                       env=dict(os.environ, OMP_NUM_THREADS='2'))
    assert r.returncode == 0 and os.path.exists(os.path.join(str(tmp_path), 'ok')), r.stdout[-2000:] + r.stderr[-4000:]


def test_cli_single_node_spawn_mode(tmp_path):
    """Launch mode B of the reference (train.py:226-236): --distributed-world-size 2 without an init method
    spawns the workers itself; here on CPU over Gloo, MNIST, with checkpoints written by rank 0 only."""
    from hetseq_b200.data.mnist_dataset import make_synthetic_mnist
    data = os.path.join(str(tmp_path), 'data')
    make_synthetic_mnist(data, 64, 16)
    save = os.path.join(str(tmp_path), 'ckpt')
    r = subprocess.run([sys.executable, '-m', 'hetseq_b200.train', '--task', 'mnist', '--optimizer', 'adadelta', '--data', data,
                        '--cpu', '--max-sentences', '8', '--max-epoch', '1', '--lr', '1.0', '--save-dir', save,
                        '--valid-subset', 'test', '--distributed-world-size', '2', '--log-interval', '1'],
                       capture_output=True, text=True, timeout=600, cwd=ROOT, env=dict(os.environ, OMP_NUM_THREADS='2'))
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    assert '| training on 2 GPUs' in r.stdout and '| done training' in r.stdout
    assert r.stdout.count('| epoch 001:') == 3          # 4 updates per rank, rank 0 prints (not the first one)
    assert sorted(os.listdir(save)) == ['checkpoint1.pt', 'checkpoint_last.pt']
