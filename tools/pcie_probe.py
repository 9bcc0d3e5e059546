"""Pinned host <-> device copy rate on this box for the e2e arm's transfer sizes (0.8, 2.4, 3.2, 4.0 MB) and one large one."""
import json, torch
for mb in (0.8, 2.4, 3.2, 4.0, 64.0):
    n = int(mb * 1e6 / 8)
    h = torch.empty(n, dtype=torch.float64).pin_memory(); d = torch.empty(n, dtype=torch.float64, device="cuda")
    out = {"MB": mb}
    for name, f in (("h2d", lambda: d.copy_(h, non_blocking=True)), ("d2h", lambda: h.copy_(d, non_blocking=True))):
        for _ in range(3): f()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(20): f()
        b.record(); torch.cuda.synchronize()
        us = a.elapsed_time(b) / 20 * 1e3
        out[name + "_us"] = round(us, 1); out[name + "_GBs"] = round(mb * 1e6 / us / 1e3, 1)
    print(json.dumps(out))
