"""Multi-rank process plumbing shared by bench.py and the sweep."""
from __future__ import annotations

import gc
import os
import sys
import threading


def teardown_process_group(pipe, ok: bool, watchdog_s: float = 60.0) -> int:
    """Leave a torchrun job cleanly.  The captured CUDA graphs hold the pair communicator (the per-step all-gather is
    inside the graph), so they are reset first; then the device is drained, the ranks meet, and the process group is
    destroyed under a watchdog (a wedged NCCL finalizer must not keep torchrun alive -- the first 2-GPU run of the bench
    printed its JSON line and never returned).  Returns the exit code: non-zero when this rank failed."""
    import torch
    import torch.distributed as dist
    if pipe is not None:
        for pl in list(pipe._plans.values()):
            if pl.graph is not None:
                pl.graph.reset()
                pl.graph = None
        pipe._plans.clear()
    gc.collect()
    torch.cuda.synchronize()
    code = 0 if ok else 1

    def _bail():
        sys.stdout.flush()
        sys.stderr.flush()
        os._exit(code)

    wd = threading.Timer(watchdog_s, _bail)
    wd.daemon = True
    wd.start()
    try:
        if ok:
            dist.barrier()
            torch.cuda.synchronize()
        dist.destroy_process_group()
    finally:
        wd.cancel()
    return code
