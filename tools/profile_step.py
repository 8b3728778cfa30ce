"""Per-kernel device-time breakdown of one training step (torch.profiler; run on the GPU box)."""
import json, os, sys, tempfile
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from torch.profiler import profile, ProfilerActivity
import bench

a = bench.parse()
torch.cuda.set_device(0)
tmp = tempfile.mkdtemp()
cfg_path, _ = bench.write_config(a.model, tmp)
from hetseq_b200 import options, registry
from hetseq_b200.controller import Controller
argv = ['--config_file', cfg_path, '--max-sentences', str(a.batch), '--synthetic-data', '--lr', '1e-4', '--no-save',
        '--log-format', 'none', '--distributed-world-size', '1', '--save-dir', tmp, '--fp16', '--clip-norm', '1.0',
        '--kernels', a.kernels]
args = options.make_args('bert', 'adam', 'PolynomialDecayScheduler', argv)
task = registry.lookup('task', 'bert').setup_task(args)
ctl = Controller(args, task, task.build_model(args))
ctl.optimizer
host = bench.make_batches(a, 4, 30522, 1)
dev = [task.prepare_sample(b, ctl.device) for b in host]
for i in range(4):
    ctl.train_step([dev[i % 4]])
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
    for i in range(3):
        ctl.train_step([dev[i % 4]])
    torch.cuda.synchronize()
ka = prof.key_averages()
rows = [(k.key, k.device_time_total / 3e3, k.count // 3) for k in ka if k.device_time_total > 0]
rows.sort(key=lambda r: -r[1])
tot = sum(r[1] for r in rows)
print('total device ms/step: %.3f' % tot)
for name, ms, cnt in rows[:40]:
    print('%8.3f ms  x%-4d %s' % (ms, cnt, name[:110]))
