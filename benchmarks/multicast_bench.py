"""Replicated delivery: one owner writes a buffer into every GPU's copy.  NVLS multicast
(multimem.st through a multicast mapping) against one peer copy per consumer."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ddl_b200.parallel import World  # noqa: E402
from ddl_b200.parallel.multicast import MulticastBuffer  # noqa: E402

world = World.from_environment()
dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0")))
torch.cuda.set_device(dev)
nbytes = 256 << 20
src = (torch.arange(nbytes, device=dev, dtype=torch.int64) * 7 % 253).to(torch.uint8)  # same on every rank
out = {}
for mode in ("multicast", "unicast"):
    buf = MulticastBuffer(world.comm, dev, nbytes)
    if mode == "unicast" and buf.mode == "multicast":
        buf.close()
        import ddl_b200.parallel.multicast as m

        orig = m.query
        m.query = lambda d: {"vmm": False, "posix_fd": False, "multicast": False}
        buf = MulticastBuffer(world.comm, dev, nbytes)
        m.query = orig
    elif mode == "multicast" and buf.mode != "multicast":
        out[mode] = None
        buf.close()
        continue
    for _ in range(2):
        if world.rank == 0:
            buf.put(src)
        torch.cuda.synchronize(dev)
        world.comm.barrier()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if world.rank == 0:
        a.record()
        for _ in range(5):
            buf.put(src)
        b.record()
    torch.cuda.synchronize(dev)
    world.comm.barrier()
    ok = bool(torch.equal(buf.local[:4096], src[:4096]))
    oks = world.comm.allgather_obj(ok)
    if world.rank == 0:
        ms = a.elapsed_time(b) / 5
        out[mode] = {"ms": round(ms, 3), "owner_egress_gbs_payload": round(nbytes / ms / 1e6, 1),
                     "delivered_gbs": round(nbytes * world.size / ms / 1e6, 1), "all_copies_ok": all(oks)}
    buf.close()
if world.rank == 0:
    print(json.dumps({"bench": "replicated_delivery_256MiB", "n_gpus": world.size, **out}), flush=True)
