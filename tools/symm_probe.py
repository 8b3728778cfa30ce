import os, torch, torch.distributed as dist
import torch.distributed._symmetric_memory as symm_mem
rank, world, local = int(os.environ['RANK']), int(os.environ['WORLD_SIZE']), int(os.environ['LOCAL_RANK'])
torch.cuda.set_device(local)
dist.init_process_group('nccl', device_id=torch.device('cuda', local))
try:
    t = symm_mem.empty(1 << 20, dtype=torch.float32, device=torch.device('cuda', local))
    hdl = symm_mem.rendezvous(t, dist.group.WORLD)
    t.fill_(rank + 1)
    torch.cuda.synchronize(); dist.barrier(device_ids=[local])
    attrs = {k: getattr(hdl, k, None) for k in ('rank', 'world_size', 'buffer_size', 'signal_pad_size', 'multicast_ptr')}
    print(rank, 'OK', attrs, 'bufs', [hex(p) for p in hdl.buffer_ptrs][:4], 'sig', [hex(p) for p in hdl.signal_pad_ptrs][:2], flush=True)
    peer = hdl.get_buffer((rank + 1) % world, (4,), torch.float32)
    print(rank, 'peer value', peer.tolist(), flush=True)
except Exception as e:
    import traceback; traceback.print_exc()
    print(rank, 'FAILED', repr(e)[:300], flush=True)
dist.barrier(device_ids=[local])
dist.destroy_process_group()
