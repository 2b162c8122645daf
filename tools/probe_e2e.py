"""GPU probe: where the end-to-end arm's time goes.  Replays bench.py's e2e loop (4 workspaces, pinned host
buffers, regions out) with host timestamps around submit / wait, and times the same loop with pieces removed."""
import os, sys, time, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "x-edr-pointcloud-segmentation_b200"))
sys.path.insert(0, ROOT)
import numpy as np, torch
import pcseg, bench
F, N = 256, 480 * 640
host = torch.from_numpy(bench.gen_frames(list(range(F)))).pin_memory()
host_np = host.numpy()
dev = host.cuda()
keep = []
def pinned(shape, dtype):
    n = int(np.prod(shape)) * np.dtype(dtype).itemsize
    t = torch.empty(max(n, 1), dtype=torch.uint8).pin_memory(); keep.append(t)
    return t.numpy()[:n].view(dtype).reshape(shape)
ctx = pcseg.Context(0)
D = 4
ws = [pcseg.Batch(ctx, F, 480, 640) for _ in range(D)]
lab = [pinned((F, N), np.int32) for _ in range(D)]
res = [pcseg.BatchResults(F, N, hull=True, alloc=pinned) for _ in range(D)]
res_dev = [pcseg.BatchResults(F, N, hull=True, alloc=lambda s, d: None) for _ in range(D)]
params = pcseg.pipeline_params(seed_mode=pcseg.SEEDS_ALL_DESCENDING, run_cluster=True)
modes = {
    "e2e (H2D + labels + regions out)": lambda j: ws[j].submit_regions(host_np, params, res[j], labels_out=lab[j]),
    "H2D only (results stay on the device)": lambda j: ws[j].submit_regions(host_np, params, res_dev[j]),
    "D2H only (input resident)": lambda j: ws[j].submit_regions(dev.data_ptr(), params, res[j], labels_out=lab[j]),
    "no copies": lambda j: ws[j].submit_regions(dev.data_ptr(), params, res_dev[j]),
}
out = {}
for name, sub in modes.items():
    for j in range(D):
        sub(j)
    for j in range(D):
        ws[j].wait()
    steps = 12
    t_sub = t_wait = 0.0
    torch.cuda.synchronize(); t0 = time.perf_counter()
    busy = [False] * D
    for i in range(steps):
        j = i % D
        if busy[j]:
            a = time.perf_counter(); ws[j].wait(); t_wait += time.perf_counter() - a
        a = time.perf_counter(); sub(j); t_sub += time.perf_counter() - a
        busy[j] = True
    for j in range(D):
        a = time.perf_counter(); ws[j].wait(); t_wait += time.perf_counter() - a
    torch.cuda.synchronize()
    tot = time.perf_counter() - t0
    out[name] = {"ms_per_step": round(1e3 * tot / steps, 2), "host_ms_in_submit_per_step": round(1e3 * t_sub / steps, 2),
                 "host_ms_in_wait_per_step": round(1e3 * t_wait / steps, 2)}
print(json.dumps(out, indent=1))
