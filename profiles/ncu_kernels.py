"""The launches behind bench.py's `kernels` rooflines, eager and a few times each, for an `ncu --set full` capture:

  ncu --set full --clock-control none --import-source on -k regex:'igemm_fwd|wgrad_kernel|adam_kernel|sumsq_kernel|mse_multi' \
      -f -o gpurun_out/r2_kernels_full python profiles/ncu_kernels.py

Same shapes, same rotating buffers as bench.kernel_rooflines (it is that function with the timing replaced by plain
launches); profiles/make_traffic.py turns the report into profiles/r2_traffic.json (`roofline.traffic`)."""
import importlib
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import bench  # noqa: E402


def eager(fns, n=0, reps=0):
    if callable(fns):
        fns = [fns]
    for _ in range(2):
        for f in fns:
            f()
    torch.cuda.synchronize()
    return 1.0


if __name__ == "__main__":
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    L = importlib.import_module(bench.PKG + "._lib")
    layers = importlib.import_module(bench.PKG + ".layers")
    bench.kernel_rooflines(torch, L, layers, dev, timeit=eager)
    torch.cuda.synchronize()
    print("ncu_kernels: done, launches", L.launch_count())
