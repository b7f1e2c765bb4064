import os, torch, torch.distributed as dist
rank, world, local = int(os.environ['RANK']), int(os.environ['WORLD_SIZE']), int(os.environ['LOCAL_RANK'])
torch.cuda.set_device(local)
dist.init_process_group('nccl', device_id=torch.device('cuda', local))
import torch.distributed._symmetric_memory as symm
t = symm.empty((1024,), dtype=torch.float32, device=torch.device('cuda', local))
t.fill_(rank + 1)
h = symm.rendezvous(t, dist.group.WORLD.group_name)
print(rank, 'buffer_ptrs', [hex(p) for p in h.buffer_ptrs], 'signal', [hex(p) for p in h.signal_pad_ptrs], 'pad size', getattr(h, 'signal_pad_size', None), flush=True)
h.barrier()
peer = h.get_buffer((rank + 1) % world, (1024,), torch.float32)
print(rank, 'peer value', float(peer[0]), flush=True)
h.barrier()
dist.destroy_process_group()
