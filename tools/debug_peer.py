import os, sys, ctypes
import torch, torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from rafiki_b200 import _native
from rafiki_b200.distributed import PeerReducer

rank, world, local = int(os.environ['RANK']), int(os.environ['WORLD_SIZE']), int(os.environ['LOCAL_RANK'])
torch.cuda.set_device(local)
dev = torch.device('cuda', local)
dist.init_process_group('nccl', device_id=dev)
red = PeerReducer(1024, 10, dev)
part = red.next_partial()
part.fill_(float(rank + 1))
torch.cuda.synchronize()
dist.barrier()
print(rank, 'launching reduce', flush=True)
res = red.reduce()
torch.cuda.synchronize()
print(rank, 'reduce done', float(res.sum()) if rank == 0 else None, flush=True)
dist.barrier()
dist.destroy_process_group()
