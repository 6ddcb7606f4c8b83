"""Host time spent inside svo_gpu_step_submit / _wait for one stream (is the single-stream loop CPU-bound?)."""
import sys, time, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from svo_b200 import capi, synth
from svo_b200.frontend import Frontend
import bench
wl = synth.WORKLOADS["C1"]; w, h = wl["w"], wl["h"]; pitch = 1280
frames = bench.make_frames(wl, [0], 11000)
dev = torch.zeros((bench.RING, 1, h, pitch), dtype=torch.uint8, device="cuda"); dev[..., :w] = torch.from_numpy(frames).cuda()
cells = (h // 32 + 1) * (w // 32 + 1) * 5
for S in (1,):
    with capi.Context(max_w=w, max_h=h, max_levels=4, max_points=32768, max_batch=S) as ctx:
        fe = Frontend(ctx, w, h, S, bucket_capacity=cells, **bench.step_params(wl))
        ts = tw = 0.0; n = 0
        for i in range(400):
            a = time.perf_counter()
            fe.submit(dev[i % bench.RING].data_ptr(), None, capi.MEM_DEVICE, pitch, h * pitch)
            b = time.perf_counter()
            if fe.outstanding == 2:
                fe.wait()
            c = time.perf_counter()
            if i >= 100:
                ts += b - a; tw += c - b; n += 1
        while fe.outstanding: fe.wait()
        print("streams %d: submit %.1f us, wait %.1f us per frame (host wall clock)" % (S, 1e6 * ts / n, 1e6 * tw / n))
