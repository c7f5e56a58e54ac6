"""Host-side cost of one pre-training step (the kernel-launch path): cProfile over 20 steps at a batch small enough that the GPU
never is the bottleneck.   python tools/host_profile.py [--batch 2]"""
import argparse
import cProfile
import os
import pstats
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import bench


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=2)
    ap.add_argument("--steps", type=int, default=20)
    a = ap.parse_args()
    args = argparse.Namespace(workload="pretrain", batch=a.batch, dropout=0.3, feat_bf16=False)
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    wl = bench.Workload(args, dev, 0, 1)
    batches = [{k: v.to(dev) for k, v in hb.items()} for hb in wl.host]
    for i in range(8):
        wl.step(batches[i % 4])
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for i in range(a.steps):
        wl.step(batches[i % 4])
    t1 = time.perf_counter()
    torch.cuda.synchronize()
    t2 = time.perf_counter()
    print("enqueue %.2f ms/step, with drain %.2f ms/step" % ((t1 - t0) / a.steps * 1e3, (t2 - t0) / a.steps * 1e3))
    pr = cProfile.Profile()
    pr.enable()
    for i in range(a.steps):
        wl.step(batches[i % 4])
    pr.disable()
    torch.cuda.synchronize()
    st = pstats.Stats(pr)
    st.sort_stats("tottime").print_stats(28)


if __name__ == "__main__":
    main()
