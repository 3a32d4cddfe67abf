"""Times initialize() = ls_init with srt = .true. (device radix sort + gather) on data already in HBM,
and the O(n) stages, against the HBM roofline.   python scripts/time_init.py [n ...]"""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
import numpy as np
import torch

from curvefit_b200 import _lib, datagen

lib = _lib.load()
peak = 6455.3
try:
    import json
    peak = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    pass
for n in [int(float(a)) for a in (sys.argv[1:] or ["1e6", "1e7", "1e8"])]:
    x = torch.from_numpy(datagen.random_x(n, 1)).cuda()
    y = torch.from_numpy(datagen.noisy_y(np.zeros(1), 0).repeat(n)).cuda()
    h = C.c_void_p()
    assert lib.cf_lowess_create(C.byref(h), -1, None) == 0
    best = 1e30
    for rep in range(4):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        rc = lib.cf_lowess_init_device(h, C.c_void_p(x.data_ptr()), C.c_void_p(y.data_ptr()), n, 1)
        e1.record()
        torch.cuda.synchronize()
        assert rc == 0, _lib.last_error()
        best = min(best, e0.elapsed_time(e1))
    # algorithmic bytes: an 8-pass LSD sort of (8-byte key, 4-byte index) records, per pass
    # histogram read 8 + scatter read 12 + write 12, plus key transform and the gather of y.  (The kernel
    # carries y itself as an 8-byte payload: its real traffic is 8 x 40 + 48 B per element.)
    alg = n * (8 * (8 + 12 + 12) + 20 + 36)
    xs = torch.empty(n, dtype=torch.float64, device="cuda")
    ys = torch.empty(n, dtype=torch.float64, device="cuda")
    lib.cf_lowess_fetch_xy  # (sorted x stays on the device; check order through torch)
    buf = lib.cf_lowess_buffer(h, _lib.CF_BUF_X)
    class A:  # zero-copy view
        __cuda_array_interface__ = {"shape": (n,), "typestr": "<f8", "data": (int(buf), False), "version": 3, "strides": None}
    xs = torch.as_tensor(A(), device="cuda")
    ok = bool(torch.all(xs[1:] >= xs[:-1]).item())
    print(f"n={n}: initialize(srt) {best:.3f} ms  -> {n / best / 1e3:.1f} Melem/s, algorithmic {alg / 1e9:.2f} GB "
          f"= {alg / best / 1e6:.0f} GB/s = {alg / best / 1e6 / peak:.2f} of HBM peak ({peak:.0f} GB/s); sorted={ok}", flush=True)
    lib.cf_lowess_destroy(h)
