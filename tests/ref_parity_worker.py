"""Runs in its own interpreter (the reference uses flat module names such as ``utils``, ``data``, ``options`` that must
not leak into the test process): imports the UNMODIFIED reference from baseline/_ref and compares this framework
against it, component by component, on CPU. Prints one JSON object {check: "ok" | error text}.

The reference modules are used as oracles only; nothing here is timed."""
import json
import os
import sys
import traceback
import types
import warnings

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(ROOT, 'baseline', '_ref')
warnings.filterwarnings('ignore')


def _stub_missing():
    for name in ('boto3', 'botocore', 'botocore.exceptions', 'h5py'):
        try:
            __import__(name)
        except Exception:
            m = types.ModuleType(name)
            if name == 'botocore.exceptions':
                m.ClientError = type('ClientError', (Exception,), {})
            sys.modules[name] = m


def check_lr_schedule():
    import argparse
    import lr_scheduler as ref_lr
    from hetseq_b200.optim import lr_scheduler as my_lr

    class Opt(object):
        def __init__(self):
            self.lr = None

        def set_lr(self, lr):
            self.lr = lr

        def get_lr(self):
            return self.lr

    for warm, total, power, end, fa in ((0, 50, 1.0, 0.0, None), (10, 100, 1.0, 0.0, None), (7, 60, 2.0, 1e-5, None),
                                        (5, 40, 0.5, 1e-6, 2)):
        ns = dict(lr=[1e-3, 5e-4], warmup_updates=warm, total_num_update=total, power=power, end_learning_rate=end,
                  force_anneal=fa)
        import optim as ref_optim

        class RefOpt(ref_optim._Optimizer):      # the reference insists on its own wrapper type
            def __init__(self):
                self.lr = None

            def set_lr(self, lr):
                self.lr = lr

            def get_lr(self):
                return self.lr

        a, b = RefOpt(), Opt()
        r = ref_lr.PolynomialDecayScheduler(argparse.Namespace(**ns), a)
        m = my_lr.PolynomialDecayScheduler(argparse.Namespace(**ns), b)
        assert abs(a.get_lr() - b.get_lr()) < 1e-15, ('init', a.get_lr(), b.get_lr())
        for epoch in range(1, 4):
            r.step(epoch)
            m.step(epoch)
            for n in range((epoch - 1) * 30, epoch * 30):
                x, y = r.step_update(n), m.step_update(n)
                assert abs(x - y) <= 1e-12 * max(1.0, abs(x)), (ns, epoch, n, x, y)


def check_optimizers():
    import torch
    import optim as ref_optim
    from hetseq_b200.optim import tensor_optimizers as mine
    torch.manual_seed(0)
    shapes = [(7, 5), (13,), (3, 4, 2)]
    for kind in ('adam', 'adadelta'):
        p_ref = [torch.nn.Parameter(torch.randn(s)) for s in shapes]
        p_my = [torch.nn.Parameter(p.detach().clone()) for p in p_ref]
        if kind == 'adam':
            o_ref = ref_optim.Adam(p_ref, lr=1e-2, betas=(0.9, 0.98), eps=1e-6, weight_decay=0.01)
            o_my = mine.Adam(p_my, lr=1e-2, betas=(0.9, 0.98), eps=1e-6, weight_decay=0.01)
        else:
            o_ref = ref_optim.Adadelta(p_ref, lr=1.0, rho=0.9, eps=1e-6, weight_decay=0.01)
            o_my = mine.Adadelta(p_my, lr=1.0, rho=0.9, eps=1e-6, weight_decay=0.01)
        for step in range(6):
            for a, b in zip(p_ref, p_my):
                g = torch.randn_like(a)
                a.grad, b.grad = g.clone(), g.clone()
            o_ref.step()
            o_my.step()
            for a, b in zip(p_ref, p_my):
                assert torch.allclose(a, b, rtol=1e-6, atol=1e-7), (kind, step, (a - b).abs().max().item())


def check_iterators():
    import data.iterators as ref_it
    from hetseq_b200.data import iterators as my_it
    items = list(range(23))
    for shards in (1, 2, 3, 8):
        for sid in range(shards):
            a = list(ref_it.ShardedIterator(items, shards, sid, fill_value=[]))
            b = list(my_it.ShardedIterator(items, shards, sid, fill_value=[]))
            assert a == b, (shards, sid, a, b)
    for chunk in (1, 2, 4, 5):
        a = list(ref_it.GroupedIterator(ref_it.CountingIterator(items), chunk))
        b = list(my_it.GroupedIterator(my_it.CountingIterator(items), chunk))
        assert a == b, (chunk, a, b)
    a, b = ref_it.CountingIterator(items, start=3), my_it.CountingIterator(items, start=3)
    for _ in range(5):
        assert next(a) == next(b)
    assert a.count == b.count and a.has_next() == b.has_next()
    assert list(a.skip(4)) == list(b.skip(4)) and a.count == b.count


def check_batch_by_size():
    import numpy as np
    from data import data_utils as ref_du
    from hetseq_b200.data import data_utils as my_du
    rng = np.random.RandomState(0)
    for trial in range(12):
        n = int(rng.randint(1, 300))
        sizes = rng.randint(1, 64, size=n).astype(np.int64)
        idx = rng.permutation(n).astype(np.int64)
        max_tokens = int(rng.choice([64, 128, 500, 4096]))
        max_sent = int(rng.choice([1, 3, 8, 32, 1000]))
        mult = int(rng.choice([1, 2, 8]))
        ref = ref_du.batch_by_size(idx, lambda i: int(sizes[i]), max_tokens=max_tokens, max_sentences=max_sent,
                                   required_batch_size_multiple=mult)
        got = my_du.batch_by_size(idx, lambda i: int(sizes[i]), max_tokens=max_tokens, max_sentences=max_sent,
                                  required_batch_size_multiple=mult)
        ref = [list(map(int, b)) for b in ref]
        got = [list(map(int, b)) for b in got]
        assert ref == got, (trial, n, max_tokens, max_sent, mult)


def check_bert_loss_and_grads():
    """Same weights, same batch, fp32 on CPU: the reference's BertForPreTraining against this framework's PyTorch
    path (all positions through the MLM head, as in the reference, and the gathered-rows default)."""
    import torch
    import bert_modeling as ref_bm
    from hetseq_b200.models import bert as my_bm
    cfg = dict(vocab_size=120, hidden_size=64, num_hidden_layers=2, num_attention_heads=4, intermediate_size=128,
               hidden_act='gelu', hidden_dropout_prob=0.0, attention_probs_dropout_prob=0.0,
               max_position_embeddings=32, type_vocab_size=2, initializer_range=0.02)
    torch.manual_seed(0)
    ref = ref_bm.BertForPreTraining(ref_bm.BertConfig.from_dict(cfg)).float()
    mine = my_bm.BertForPreTraining(my_bm.BertConfig.from_dict(cfg)).float()
    assert list(ref.state_dict().keys()) == list(mine.state_dict().keys()), 'state-dict key order/names differ'
    assert sum(p.numel() for p in ref.parameters()) == sum(p.numel() for p in mine.parameters())
    mine.load_state_dict(ref.state_dict())
    B, S = 4, 16
    g = torch.Generator().manual_seed(1)
    ids = torch.randint(0, 120, (B, S), generator=g)
    seg = torch.randint(0, 2, (B, S), generator=g)
    mask = torch.ones(B, S, dtype=torch.long)
    mask[1, 11:] = 0
    labels = torch.full((B, S), -1, dtype=torch.long)
    labels[0, 3], labels[0, 9], labels[2, 5], labels[3, 1] = 7, 11, 13, 17
    nsp = torch.tensor([0, 1, 1, 0])
    ref.train()
    mine.train()
    lr = ref(ids, seg, mask, labels, nsp)
    lm = mine(ids, seg, mask, labels, nsp)
    assert abs(lr.item() - lm.item()) < 1e-5, (lr.item(), lm.item())
    lr.backward()
    lm.backward()
    gr = dict(ref.named_parameters())
    for name, p in mine.named_parameters():
        a, b = gr[name].grad, p.grad
        assert a is not None and b is not None, name
        assert torch.allclose(a, b, rtol=1e-4, atol=1e-6), (name, (a - b).abs().max().item())
    mine.mlm_all_positions = True
    mine.zero_grad()
    lm2 = mine(ids, seg, mask, labels, nsp)
    assert abs(lr.item() - lm2.item()) < 1e-5


def check_all_heads():
    """Every model class the reference exports (BertModel and the six task heads, bert_modeling.py:736-1290): same
    state-dict keys, and with the reference's weights loaded the same outputs -- logits without labels, loss with
    labels, and the gradient of that loss for every parameter."""
    import torch
    import bert_modeling as ref_bm
    from hetseq_b200.models import bert as my_bm
    cfg = dict(vocab_size=90, hidden_size=32, num_hidden_layers=2, num_attention_heads=4, intermediate_size=64,
               hidden_act='gelu', hidden_dropout_prob=0.0, attention_probs_dropout_prob=0.0,
               max_position_embeddings=24, type_vocab_size=2, initializer_range=0.02)
    B, S, C = 3, 12, 4
    g = torch.Generator().manual_seed(3)
    ids = torch.randint(0, 90, (B, S), generator=g)
    seg = torch.randint(0, 2, (B, S), generator=g)
    mask = torch.ones(B, S, dtype=torch.long)
    mask[2, 8:] = 0
    mlm = torch.full((B, S), -1, dtype=torch.long)
    mlm[0, 2], mlm[1, 7], mlm[2, 3] = 5, 9, 11
    cases = [
        ('BertForMaskedLM', (), dict(masked_lm_labels=mlm)),
        ('BertForNextSentencePrediction', (), dict(next_sentence_label=torch.tensor([1, 0, 1]))),
        ('BertForSequenceClassification', (5,), dict(labels=torch.tensor([4, 0, 2]))),
        ('BertForTokenClassification', (6,), dict(labels=torch.randint(0, 6, (B, S), generator=g))),
        ('BertForQuestionAnswering', (), dict(start_positions=torch.tensor([1, 5, 40]), end_positions=torch.tensor([3, 9, 2]))),
    ]
    for name, extra, labels in cases:
        torch.manual_seed(1)
        ref = getattr(ref_bm, name)(ref_bm.BertConfig.from_dict(cfg), *extra).float()
        mine = getattr(my_bm, name)(my_bm.BertConfig.from_dict(cfg), *extra).float()
        assert list(ref.state_dict().keys()) == list(mine.state_dict().keys()), name + ': state-dict keys differ'
        mine.load_state_dict(ref.state_dict())
        ref.eval()
        mine.eval()
        with torch.no_grad():
            a, b = ref(ids, seg, mask), mine(ids, seg, mask)
        a, b = (a, b) if isinstance(a, tuple) else ((a,), (b,))
        for x, y in zip(a, b):
            assert x.shape == y.shape and torch.allclose(x, y, rtol=1e-4, atol=1e-5), name + ': logits differ'
        la, lb = ref(ids, seg, mask, **labels), mine(ids, seg, mask, **labels)
        assert abs(la.item() - lb.item()) < 1e-5, (name, la.item(), lb.item())
        la.backward()
        lb.backward()
        gr = dict(ref.named_parameters())
        for pn, p in mine.named_parameters():
            ga, gb = gr[pn].grad, p.grad
            assert (ga is None) == (gb is None), (name, pn)
            if ga is not None:
                assert torch.allclose(ga, gb, rtol=1e-4, atol=1e-6), (name, pn, (ga - gb).abs().max().item())
    # multiple choice: inputs are [B, choices, S]
    torch.manual_seed(1)
    ref = ref_bm.BertForMultipleChoice(ref_bm.BertConfig.from_dict(cfg), C).float().eval()
    mine = my_bm.BertForMultipleChoice(my_bm.BertConfig.from_dict(cfg), C).float().eval()
    mine.load_state_dict(ref.state_dict())
    ids4 = torch.randint(0, 90, (B, C, S), generator=g)
    seg4 = torch.randint(0, 2, (B, C, S), generator=g)
    mask4 = torch.ones(B, C, S, dtype=torch.long)
    lab = torch.tensor([3, 0, 1])
    with torch.no_grad():
        assert torch.allclose(ref(ids4, seg4, mask4), mine(ids4, seg4, mask4), rtol=1e-4, atol=1e-5)
        assert abs(ref(ids4, seg4, mask4, lab).item() - mine(ids4, seg4, mask4, lab).item()) < 1e-5
    # the trunk: all encoder layers and the pooled output
    torch.manual_seed(1)
    ref = ref_bm.BertModel(ref_bm.BertConfig.from_dict(cfg)).float().eval()
    mine = my_bm.BertModel(my_bm.BertConfig.from_dict(cfg)).float().eval()
    mine.load_state_dict(ref.state_dict())
    with torch.no_grad():
        (la, pa), (lb, pb) = ref(ids, seg, mask), mine(ids, seg, mask)
    assert len(la) == len(lb) == 2 and torch.allclose(pa, pb, rtol=1e-4, atol=1e-5)
    for x, y in zip(la, lb):
        assert torch.allclose(x, y, rtol=1e-4, atol=1e-5)


def check_meters_and_format():
    import meters as ref_m
    import progress_bar as ref_pb
    from hetseq_b200.utils import meters as my_m, progress_bar as my_pb
    a, b = ref_m.AverageMeter(), my_m.AverageMeter()
    for v, n in ((1.5, 2), (3.0, 1), (0.25, 5)):
        a.update(v, n)
        b.update(v, n)
    assert (a.val, a.sum, a.count, a.avg) == (b.val, b.sum, b.count, b.avg)
    assert ref_pb.format_stat(a) == my_pb.format_stat(b)
    for x in (3, 0.000123456, 12345.678):
        assert ref_pb.format_stat(x) == my_pb.format_stat(x)


def check_flag_defaults():
    """Every flag of the reference's training parser exists here with the same default."""
    import options as ref_opt
    from hetseq_b200 import options as my_opt
    for task, opt in (('bert', 'adam'), ('mnist', 'adadelta')):
        rp = ref_opt.get_training_parser(task=task, optimizer=opt, lr_scheduler='PolynomialDecayScheduler')
        mp = my_opt.get_training_parser(task=task, optimizer=opt)
        mine = {}
        for act in mp._actions:
            for s in act.option_strings:
                mine[s] = act
        skip = {'--distributed-world-size'}       # default = number of visible GPUs in both; environment dependent
        for act in rp._actions:
            for s in act.option_strings:
                if s in ('-h', '--help') or s in skip:
                    continue
                assert s in mine, 'reference flag {} missing'.format(s)
                rd, md = act.default, mine[s].default
                if s in ('--distributed-backend',):
                    continue                      # nccl there; nccl-or-gloo by device here
                same = rd == md or str(rd) == str(md)
                if not same:
                    try:
                        same = float(rd) == float(md)
                    except (TypeError, ValueError):
                        pass
                assert same, (s, rd, md)


def check_checkpoint_interchange():
    """Files written by the reference's save_state load into this framework's Controller (weights, Adam state in
    the per-parameter layout, update counter, LR-scheduler state); files written here feed the reference's own
    optimizer wrapper. MNIST net on CPU."""
    import argparse
    import tempfile
    import torch
    import checkpoint_utils as ref_ck
    import lr_scheduler as ref_lr
    import optim as ref_optim
    import tasks as ref_tasks
    from hetseq_b200 import options, registry
    from hetseq_b200.controller import Controller
    from hetseq_b200.data.mnist_dataset import make_synthetic_mnist

    tmp = tempfile.mkdtemp(prefix='hsparity')
    data = os.path.join(tmp, 'data')
    make_synthetic_mnist(data, 32, 16)
    argv = ['--data', data, '--cpu', '--max-sentences', '16', '--lr', '0.01', '--save-dir', os.path.join(tmp, 'ckpt'),
            '--distributed-world-size', '1', '--log-format', 'none', '--adam-betas', '(0.9, 0.98)', '--adam-eps', '1e-6',
            '--weight-decay', '0.01']
    a = options.make_args('mnist', 'adam', argv=argv)
    task = registry.lookup('task', 'mnist').setup_task(a)
    ctl = Controller(a, task, task.build_model(a))

    # ---- reference side: its MNISTNet, its _Adam wrapper, two updates, its save_state
    torch.manual_seed(5)
    ref_model = ref_tasks.MNISTNet()
    assert list(ref_model.state_dict().keys()) == list(ctl.get_model().state_dict().keys())
    rargs = argparse.Namespace(lr=[0.01], adam_betas='(0.9, 0.98)', adam_eps=1e-6, weight_decay=0.01, warmup_updates=0,
                               total_num_update=100, power=1.0, end_learning_rate=0.0, force_anneal=None,
                               no_save_optimizer_state=False)
    ref_opt = ref_optim._Adam(rargs, list(ref_model.parameters()))
    ref_sched = ref_lr.PolynomialDecayScheduler(rargs, ref_opt)
    for _ in range(2):
        for p in ref_model.parameters():
            p.grad = torch.randn_like(p) * 0.1
        ref_opt.step()
    f1 = os.path.join(tmp, 'from_reference.pt')
    ref_ck.save_state(f1, rargs, ref_model.state_dict(), None, ref_opt, ref_sched, 2)
    # ---- this framework reads it
    ctl.load_checkpoint(f1)
    assert ctl.get_num_updates() == 2
    for (k, v), (_, w) in zip(ref_model.state_dict().items(), ctl.get_model().state_dict().items()):
        assert torch.equal(v, w), k
    mine_state = ctl.optimizer.state_dict()
    ref_state = ref_opt.state_dict()
    assert set(mine_state['state'].keys()) == set(ref_state['state'].keys())
    for idx, st in ref_state['state'].items():
        for key in ('exp_avg', 'exp_avg_sq'):
            assert torch.allclose(st[key].float(), mine_state['state'][idx][key].float(), rtol=0, atol=0), (idx, key)
        assert int(st['step']) == int(mine_state['state'][idx]['step'])
    # one more identical update on both sides keeps them in lock-step (same Adam math on the restored state)
    grads = [torch.randn_like(p) * 0.1 for p in ref_model.parameters()]
    for p, g in zip(ref_model.parameters(), grads):
        p.grad = g.clone()
    ref_opt.step()
    ctl.zero_grad()
    for p, g in zip(ctl.get_model().parameters(), grads):
        p.grad = g.clone() if p.grad is None else p.grad.copy_(g)
    ctl.optimizer.step()
    for (k, v), (_, w) in zip(ref_model.state_dict().items(), ctl.get_model().state_dict().items()):
        assert torch.allclose(v, w, rtol=1e-6, atol=1e-7), (k, (v - w).abs().max().item())
    # ---- and back: a file written here feeds the reference's optimizer wrapper
    f2 = os.path.join(tmp, 'from_here.pt')
    ctl.save_checkpoint(f2, {'train_iterator': {'epoch': 1, 'iterations_in_epoch': 0}})
    state = torch.load(f2, map_location='cpu', weights_only=False)    # (the reference's own loader predates weights_only)
    assert set(state) >= {'args', 'model', 'optimizer_history', 'extra_state', 'last_optimizer_state'}
    assert state['optimizer_history'][-1]['optimizer_name'] == '_Adam'
    ref_model2 = ref_tasks.MNISTNet()
    ref_model2.load_state_dict(state['model'])
    ref_opt2 = ref_optim._Adam(rargs, list(ref_model2.parameters()))
    ref_opt2.load_state_dict(state['last_optimizer_state'])
    st2 = ref_opt2.state_dict()['state']
    for idx, st in ref_opt.state_dict()['state'].items():
        assert torch.allclose(st['exp_avg'], st2[idx]['exp_avg'], rtol=1e-6, atol=1e-8), idx
        assert int(st2[idx]['step']) == 3


def check_mnist_training_curve():
    """Whole trainer against whole trainer: the reference's Controller.train_step and this framework's, same
    initial weights, same synthetic MNIST files, 8 updates on CPU. Identical batch order (seeded shuffle), identical
    dropout masks (seed + num_updates), identical Adadelta + clipping => the same loss curve. The reference's
    single-process path logs the raw loss, this framework logs base 2 like the reference's fast-stat-sync path."""
    import math
    import tempfile
    import torch
    import options as ro
    import tasks as rt
    from controller import Controller as RefController
    from hetseq_b200 import options, registry
    from hetseq_b200.controller import Controller
    from hetseq_b200.data.mnist_dataset import make_synthetic_mnist
    tmp = tempfile.mkdtemp(prefix='hscurve')
    data = os.path.join(tmp, 'data')
    make_synthetic_mnist(data, 128, 16)
    parser = ro.get_training_parser(task='mnist', optimizer='adadelta', lr_scheduler='PolynomialDecayScheduler')
    args = ro.parse_args_and_arch(parser, ['--task', 'mnist', '--optimizer', 'adadelta', '--data', data, '--cpu',
                                           '--max-sentences', '16', '--lr', '0.5', '--save-dir', tmp + '/ck',
                                           '--distributed-world-size', '1', '--no-save', '--clip-norm', '25'])
    torch.manual_seed(args.seed)
    task = rt.MNISTTask.setup_task(args)
    task.load_dataset('train')
    model = task.build_model(args)
    init = {k: v.clone() for k, v in model.state_dict().items()}
    ctl = RefController(args, task, model)
    ref_losses = []
    for sample in ctl.get_train_iterator(epoch=0).next_epoch_itr(shuffle=True):
        ref_losses.append(float(ctl.train_step([sample])['loss']))
    a = options.make_args('mnist', 'adadelta', argv=['--data', data, '--cpu', '--max-sentences', '16', '--lr', '0.5',
                                                    '--save-dir', tmp + '/ck2', '--distributed-world-size', '1',
                                                    '--log-format', 'none', '--no-save', '--sample-size-mode',
                                                    'reference', '--clip-norm', '25'])
    torch.manual_seed(a.seed)
    t2 = registry.lookup('task', 'mnist').setup_task(a)
    t2.load_dataset('train')
    m2 = t2.build_model(a)
    m2.load_state_dict(init)
    c2 = Controller(a, t2, m2)
    mine = []
    for sample in c2.get_train_iterator(epoch=0).next_epoch_itr(shuffle=True):
        mine.append(float(c2.train_step([sample])['loss']) * math.log(2))
    assert len(mine) == len(ref_losses) == 8
    assert ref_losses[-1] < ref_losses[0] - 0.5, ref_losses          # it does learn
    for i, (x, y) in enumerate(zip(ref_losses, mine)):
        assert abs(x - y) <= 2e-5 * max(1.0, abs(x)), (i, x, y)
    for (k, v), (_, w) in zip(model.state_dict().items(), m2.state_dict().items()):
        assert torch.allclose(v, w, rtol=1e-4, atol=1e-6), (k, (v - w).abs().max().item())


def check_bert_training_curve():
    """BERT pre-training, whole trainer against whole trainer on CPU (fp32, dropout off): the reference's
    LanguageModelingTask + Controller (DDP-free single process) + its Adam, and this framework with
    --sample-size-mode reference. Same weights, same synthetic batches, 6 updates with warm-up + polynomial decay
    + clipping + weight decay: same losses, same weights afterwards."""
    import json as _json
    import math
    import tempfile
    import torch
    import options as ro
    import tasks as rt
    from controller import Controller as RefController
    from hetseq_b200 import options, registry
    from hetseq_b200.controller import Controller
    from hetseq_b200.data.bert_dataset import SyntheticBertData
    tmp = tempfile.mkdtemp(prefix='hsbertcurve')
    cfg = dict(vocab_size=300, hidden_size=64, num_hidden_layers=2, num_attention_heads=4, intermediate_size=128,
               hidden_act='gelu', hidden_dropout_prob=0.0, attention_probs_dropout_prob=0.0,
               max_position_embeddings=32, type_vocab_size=2, initializer_range=0.02)
    cfg_path = os.path.join(tmp, 'bert_config.json')
    with open(cfg_path, 'w') as f:
        _json.dump(cfg, f)
    vocab = os.path.join(tmp, 'vocab.txt')
    with open(vocab, 'w') as f:
        f.write('\n'.join('tok%d' % i for i in range(16)))
    ds = SyntheticBertData(6 * 4, 32, 5, 300, seed=11)
    batches = [ds.fetch_rows(list(range(i * 4, i * 4 + 4))) for i in range(6)]
    common = ['--config_file', cfg_path, '--max-sentences', '4', '--lr', '1e-3', '--warmup-updates', '2',
              '--total-num-update', '20', '--weight-decay', '0.01', '--clip-norm', '1.0', '--adam-betas', '(0.9, 0.98)',
              '--adam-eps', '1e-6', '--no-save', '--distributed-world-size', '1', '--cpu']
    parser = ro.get_training_parser(task='bert', optimizer='adam', lr_scheduler='PolynomialDecayScheduler')
    args = ro.parse_args_and_arch(parser, ['--task', 'bert', '--data', tmp, '--dict', vocab, '--save-dir', tmp + '/ck'] + common)
    torch.manual_seed(args.seed)
    task = rt.LanguageModelingTask.setup_task(args)
    model = task.build_model(args)
    init = {k: v.clone() for k, v in model.state_dict().items()}
    ctl = RefController(args, task, model)
    ref_losses = [float(ctl.train_step([[t.clone() for t in b]])['loss']) for b in batches]

    a = options.make_args('bert', 'adam', argv=['--synthetic-data', '--save-dir', tmp + '/ck2', '--log-format', 'none',
                                                '--sample-size-mode', 'reference'] + common)
    torch.manual_seed(a.seed)
    t2 = registry.lookup('task', 'bert').setup_task(a)
    m2 = t2.build_model(a)
    m2.load_state_dict(init)
    c2 = Controller(a, t2, m2)
    # logged here: sum / (sample_size * ln 2) like the reference's fast-stat-sync path, and in reference mode
    # sample_size is the sequence length (32); the reference's single-process path logs the raw loss
    mine = [float(c2.train_step([[t.clone() for t in b]])['loss']) * math.log(2) * 32 for b in batches]
    for i, (x, y) in enumerate(zip(ref_losses, mine)):
        assert abs(x - y) <= 5e-5 * max(1.0, abs(x)), (i, ref_losses, mine)
    assert abs(ctl.get_lr() - c2.get_lr()) < 1e-12, (ctl.get_lr(), c2.get_lr())
    worst = max((v - w).abs().max().item() for (k, v), (_, w) in zip(model.state_dict().items(), m2.state_dict().items()))
    assert worst < 2e-5, worst


CHECKS = [check_bert_training_curve, check_mnist_training_curve, check_checkpoint_interchange, check_lr_schedule, check_optimizers, check_iterators, check_batch_by_size, check_bert_loss_and_grads,
          check_all_heads, check_meters_and_format, check_flag_defaults]


def main():
    out = {}
    if not os.path.exists(os.path.join(REF, 'controller.py')):
        print(json.dumps({'unavailable': 'baseline/_ref not installed'}))
        return
    _stub_missing()
    sys.path.insert(0, ROOT)
    sys.path.insert(0, REF)
    for fn in CHECKS:
        try:
            fn()
            out[fn.__name__] = 'ok'
        except BaseException:
            out[fn.__name__] = traceback.format_exc()[-1500:]
    print('RESULT ' + json.dumps(out))


if __name__ == '__main__':
    main()
