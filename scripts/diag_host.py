#!/usr/bin/env python
"""Host-side facts that bound the e2e path: cores, and the multi-threaded copy bandwidth of host memory (numpy copies
release the GIL).  Diagnostic only."""
import os
import threading
import time

import numpy as np

print("cpus:", os.cpu_count(), "affinity:", len(os.sched_getaffinity(0)))
N = 64 << 20
for T in (1, 2, 4, 8, 16, 32, 64):
    if T > 2 * len(os.sched_getaffinity(0)):
        break
    src = [np.ones(N, dtype=np.uint8) for _ in range(T)]
    dst = [np.empty(N, dtype=np.uint8) for _ in range(T)]
    for d in dst:
        d[:] = 0

    def work(i):
        for _ in range(4):
            np.copyto(dst[i], src[i])
    th = [threading.Thread(target=work, args=(i,)) for i in range(T)]
    t0 = time.perf_counter()
    [t.start() for t in th]; [t.join() for t in th]
    dt = time.perf_counter() - t0
    print(f"threads={T}: copy {T * 4 * N / dt / 1e9:.1f} GB/s (read+write {2 * T * 4 * N / dt / 1e9:.1f})", flush=True)
