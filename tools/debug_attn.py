import sys, torch
sys.path.insert(0, '.')
from hetseq_b200.ops import attention_native as an
S = int(sys.argv[1]); what = sys.argv[2]
B, heads, d = 3, 4, 64; H = heads * d
dt = torch.float16
torch.manual_seed(8)
qkv = (torch.randn(B * S, 3 * H, device='cuda') * 0.7).to(dt).requires_grad_(True)
mask = torch.zeros(B, S, device='cuda')
print('start', S, what, flush=True)
out = an.attention(qkv, mask, B, S, heads, 0.0, 1)
torch.cuda.synchronize(); print('fwd done', float(out.float().abs().mean()), flush=True)
if what == 'bwd':
    out.backward(torch.randn_like(out))
    torch.cuda.synchronize(); print('bwd done', float(qkv.grad.float().abs().mean()), flush=True)
