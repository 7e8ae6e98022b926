"""Zero-copy pull over PCIe: gather kernel reading a cudaHostAlloc'ed (torch pinned) shard."""
import json, os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ddl_b200.ops import Convert, GatherPlan, SourceTable

dev = torch.device("cuda:0")
SB = 150528
rows, B = 4096, 256
host = torch.randint(0, 255, (rows * SB,), dtype=torch.uint8).pin_memory()
img = Convert.normalize_u8([0.485, 0.456, 0.406], [0.229, 0.224, 0.225], 224 * 224)
for name, cv in (("copy", Convert.copy()), ("u8_bf16", img)):
    for variant, tune in ((1, 0), (2, 0), (2, 2), (2, 3)):
        plan = GatherPlan(SourceTable([host.data_ptr()], dev), SB, SB, cv, variant=variant, tune=tune)
        slots = torch.randperm(rows, device=dev)[: 8 * B].to(torch.int64).view(8, B).contiguous()
        out = plan.alloc_batch(B)
        plan.run(slots[0], out)
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for i in range(1, 8):
            plan.run(slots[i], out)
        b.record()
        torch.cuda.synchronize()
        ms = a.elapsed_time(b) / 7
        print(json.dumps({"case": name, "variant": variant, "tune": tune, "ms": round(ms, 3),
                          "pcie_gbs": round(B * SB / ms / 1e6, 1)}), flush=True)
