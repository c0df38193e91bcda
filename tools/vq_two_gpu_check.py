"""Each rank builds the codebook plan on ITS device and searches the same rows; the indices must be identical on every
rank (python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tools/vq_two_gpu_check.py)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch, torch.distributed as dist
from xtts_b200 import Quantize, get_codebook_indices
rank, world, local = (int(os.environ.get(k, d)) for k, d in (("RANK", "0"), ("WORLD_SIZE", "1"), ("LOCAL_RANK", "0")))
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
g = torch.Generator().manual_seed(3)
q = Quantize(512, 8192)
q.embed.copy_(torch.randn(512, 8192, generator=g))
q = q.eval().to(dev)                                   # note: torch.cuda.current_device() stays 0 on every rank
z = (q.embed.T[torch.randint(0, 8192, (5000,), generator=g).to(dev)] + 0.5 * torch.randn(5000, 512, generator=g).to(dev))
quant, diff, ind = q(z)
ind2 = get_codebook_indices(q, z.view(5, 1000, 512).transpose(1, 2).contiguous().transpose(1, 2)).reshape(-1)
q.plan(dev).status()
assert torch.equal(ind, ind2)
if world > 1:
    gathered = [torch.empty_like(ind) for _ in range(world)]
    dist.all_gather(gathered, ind)
    assert all(torch.equal(gathered[0], t) for t in gathered)
    dist.destroy_process_group()
print(f"rank {rank} on {dev}: ok, diff {float(diff):.6f}, first indices {ind[:4].tolist()}")
