"""Worker for the replicated-delivery test: every rank in turn writes a pattern into all copies."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from ddl_b200.parallel import World  # noqa: E402
from ddl_b200.parallel.multicast import MulticastBuffer, query  # noqa: E402

world = World.from_environment()
dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0")))
torch.cuda.set_device(dev)
caps = query(dev)
buf = MulticastBuffer(world.comm, dev, 1 << 20)
if world.rank == 0:
    print("CAPS", caps, "MODE", buf.mode, flush=True)
for owner in range(world.size):
    pattern = (torch.arange(1 << 20, device=dev, dtype=torch.int64) * (owner + 3) % 251).to(torch.uint8)
    if world.rank == owner:
        buf.put(pattern)
    torch.cuda.synchronize(dev)
    world.comm.barrier()
    assert torch.equal(buf.local, pattern), f"rank {world.rank}: copy differs after owner {owner} wrote"
    world.comm.barrier()
buf.close()
print(f"rank {world.rank} ok", flush=True)
