#!/usr/bin/env python
"""Times the onesweep radix sort through the C ABI (measurement aid for the ranking / tile variants).

    QDSP_RS_RANK=0|1|2 python tools/sort_probe.py [log2_n=26] [reps=5] [configs=kv32,k64v32,k32]

One JSON line per config: ms per sort, ms per digit pass, algorithmic GB/s (N*kb + P*2*N*(kb+vb))."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch

import quadrants_b200 as qd
from quadrants_b200 import _lib


def main():
    log2n = int(sys.argv[1]) if len(sys.argv) > 1 else 26
    reps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
    configs = (sys.argv[3] if len(sys.argv) > 3 else "kv32,k64v32,k32").split(",")
    qd.init(arch=qd.cuda, device=0)
    lib = _lib.load()
    stream = torch.cuda.current_stream().cuda_stream
    n = 1 << log2n
    g = torch.Generator(device="cuda").manual_seed(4)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for cfg in configs:
        kb, vb, kind = {"kv32": (4, 4, _lib.KEY_I32), "k32": (4, 0, _lib.KEY_I32), "k64v32": (8, 4, _lib.KEY_U64)}[cfg]
        if kb == 4:
            src = torch.randint(-2**31, 2**31, (n,), dtype=torch.int64, device="cuda", generator=g).to(torch.int32)
        else:
            src = torch.randint(-2**63, 2**63 - 1, (n,), dtype=torch.int64, device="cuda", generator=g)
        k = torch.empty_like(src)
        v = torch.arange(n, dtype=torch.int32, device="cuda") if vb else None
        tmp = torch.empty(lib.qdsp_sort_tmp_bytes(n, kind, vb), dtype=torch.uint8, device="cuda")
        run = lambda: _lib.check(lib.qdsp_sort_pairs(k.data_ptr(), v.data_ptr() if vb else None, n, kind, vb,
                                                     tmp.data_ptr(), tmp.numel(), stream))
        total = 0.0
        for it in range(2 + reps):
            k.copy_(src)
            torch.cuda.synchronize()
            e0.record(); run(); e1.record()
            torch.cuda.synchronize()
            if it >= 2:
                total += e0.elapsed_time(e1)
        ms = total / reps
        ok = bool((k[1:] >= k[:-1]).all().item())
        passes = kb
        nbytes = n * kb + passes * 2 * n * (kb + vb)
        print(json.dumps({"config": cfg, "log2_n": log2n, "rank_variant": os.environ.get("QDSP_RS_RANK", "default"),
                          "ms": round(ms, 4), "ms_per_pass": round(ms / passes, 4), "gbs": round(nbytes / ms / 1e6, 1),
                          "gitems_per_s": round(n / ms / 1e6, 2), "sorted": ok}), flush=True)
        del src, k, v, tmp


if __name__ == "__main__":
    main()
