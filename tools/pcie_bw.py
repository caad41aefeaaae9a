#!/usr/bin/env python3
"""Pinned host <-> device copy bandwidth of the box (one direction, then both at once): the ceiling of bench.py's e2e.

    python tools/pcie_bw.py            one GPU (cuda:0)
    python tools/pcie_bw.py --all      every visible GPU AT THE SAME TIME, one process per GPU: the aggregate ceiling the
                                       multi-GPU e2e numbers are held against (host memory / root-complex bound)
"""
import multiprocessing as mp
import sys
import time


def worker(dev, barrier, out, reps=8):
    import torch
    torch.cuda.set_device(dev)
    n = 64 << 20                       # 64 Mi floats = 256 MiB
    h_in, h_out = torch.empty(n).pin_memory(), torch.empty(n).pin_memory()
    d_in, d_out = torch.empty(n, device="cuda"), torch.empty(n, device="cuda")
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()

    def run(up, down, reps):
        torch.cuda.synchronize()
        if barrier is not None:
            barrier.wait()             # every GPU copies over the same interval
        t0 = time.perf_counter()
        for _ in range(reps):
            if up:
                with torch.cuda.stream(s1):
                    d_in.copy_(h_in, non_blocking=True)
            if down:
                with torch.cuda.stream(s2):
                    h_out.copy_(d_out, non_blocking=True)
        torch.cuda.synchronize()
        return reps * n * 4 / (time.perf_counter() - t0) / 1e9

    for _ in range(2):
        run(True, True, 2)
    res = (run(True, False, reps), run(False, True, reps), run(True, True, reps))
    if out is not None:
        out.put((dev, res))
    return res


def main():
    if "--all" not in sys.argv:
        up, down, both = worker(0, None, None)
        print(f"H2D alone   {up:6.1f} GB/s")
        print(f"D2H alone   {down:6.1f} GB/s")
        print(f"both at once {both:6.1f} GB/s each direction")
        return
    import torch
    n_gpu = torch.cuda.device_count()
    ctx = mp.get_context("spawn")
    barrier, out = ctx.Barrier(n_gpu), ctx.Queue()
    procs = [ctx.Process(target=worker, args=(d, barrier, out)) for d in range(n_gpu)]
    for p in procs:
        p.start()
    res = dict(out.get() for _ in range(n_gpu))
    for p in procs:
        p.join()
    print(f"{n_gpu} GPUs copying at the same time (one process per GPU, 256 MiB pinned buffers), GB/s per GPU:")
    for d in sorted(res):
        print(f"  cuda:{d}  H2D alone {res[d][0]:6.1f}   D2H alone {res[d][1]:6.1f}   both at once {res[d][2]:6.1f} each direction")
    tot = [sum(res[d][i] for d in res) for i in range(3)]
    print(f"aggregate: H2D alone {tot[0]:.1f} GB/s, D2H alone {tot[1]:.1f} GB/s, both at once {tot[2]:.1f} GB/s each direction")


if __name__ == "__main__":
    main()
