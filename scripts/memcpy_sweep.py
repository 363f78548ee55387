"""What pageable <-> pinned staging can reach on this host: parallel memcpy (numpy slices on T threads, the GIL is released
inside the copy) beside the driver's own pageable cudaMemcpy."""
import threading
import time

import numpy as np
import torch

NB = 256 << 20
src = np.random.randint(0, 255, NB, dtype=np.uint8)
pin = torch.empty(NB, dtype=torch.uint8).pin_memory()
pin_np = pin.numpy()
dev = torch.empty(NB, dtype=torch.uint8, device="cuda:0")
out = np.empty(NB, dtype=np.uint8)


def par_copy(dst, s, T):
    step = (NB + T - 1) // T
    ts = [threading.Thread(target=lambda a=a: np.copyto(dst[a:a + step], s[a:a + step])) for a in range(0, NB, step)]
    [t.start() for t in ts]
    [t.join() for t in ts]


for T in (1, 2, 4, 6, 8, 12):
    par_copy(pin_np, src, T)
    t0 = time.perf_counter()
    for _ in range(4):
        par_copy(pin_np, src, T)
    dt = (time.perf_counter() - t0) / 4
    t0 = time.perf_counter()
    for _ in range(4):
        par_copy(out, pin_np, T)
    dt2 = (time.perf_counter() - t0) / 4
    print(f"{T:2d} threads: pageable -> pinned {NB / dt / 1e9:5.1f} GB/s, pinned -> pageable {NB / dt2 / 1e9:5.1f} GB/s")
src_t, out_t = torch.from_numpy(src), torch.from_numpy(out)
for name, fn in (("driver pageable H2D", lambda: dev.copy_(src_t)), ("driver pageable D2H", lambda: out_t.copy_(dev)),
                 ("pinned H2D", lambda: dev.copy_(pin, non_blocking=True)), ("pinned D2H", lambda: pin.copy_(dev, non_blocking=True))):
    fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(4):
        fn()
    torch.cuda.synchronize()
    print(f"{name}: {4 * NB / (time.perf_counter() - t0) / 1e9:.1f} GB/s")
