#!/usr/bin/env python
"""GVX_SHARD=z with ONE GPU PER RANK: tests/cpp/shard_example (reference-API caller code) as 1 process and as N
processes (N = number of visible GPUs, or argv[1]); the N outputs must equal the single-process output, with the
map x map counts all-reduced inside the collide kernel over NVLink peer memory (fused) and through the host segment."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import torch  # noqa: E402

from test_gpu_shard_api import run_spmd  # noqa: E402
from test_shard_host import build_shard_example  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else torch.cuda.device_count()
exe = build_shard_example()
single, _ = run_spmd(exe, 1, False)
for fused in (True, False):
    outs, errs = run_spmd(exe, n, fused)
    for rank, o in enumerate(outs):
        assert o == single[0], (fused, rank, o, single[0])
        assert ("in-kernel all-reduce %d" % int(fused)) in errs[rank], errs[rank]
print("shard_spmd_check ok: %d ranks == 1 process (fused in-kernel all-reduce and host segment)\n%s" % (n, single[0]))
