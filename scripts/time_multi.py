"""Times cf_lowess_smooth_multi (one process, all devices) against cf_lowess_smooth on pinned buffers.
    python scripts/time_multi.py n f nsteps [ndev]      (CF_DEV_LIB=1 CF_MULTI_TIMES=1: per-rank phase times)"""
import ctypes as C, os, sys, time
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
import numpy as np, torch
from curvefit_b200 import _lib
if os.environ.get("CF_DEV_LIB"):
    _lib.LIB_PATH = os.path.join(os.path.dirname(__file__), "dev", "libcurvefit_cuda.so")
from curvefit_b200 import datagen, regression
n = int(float(sys.argv[1])); f = float(sys.argv[2]); nsteps = int(sys.argv[3])
ndev = int(sys.argv[4]) if len(sys.argv) > 4 else torch.cuda.device_count()
lib = _lib.load()
x = datagen.grid_x(n); y = datagen.noisy_y(x)
pin = lambda a: torch.from_numpy(a).pin_memory().numpy()
xp, yp = pin(x), pin(y)
outs = [torch.empty(n, dtype=torch.float64).pin_memory().numpy() for _ in range(3)]
lib.cf_set_default_method(_lib.CF_METHOD_TREE)
p = regression._dp
def multi(k):
    devs = (C.c_int32 * k)(*range(k))
    rc = lib.cf_lowess_smooth_multi(p(xp), p(yp), n, f, nsteps, devs, k, p(outs[0]), p(outs[1]), p(outs[2]))
    assert rc == 0, _lib.last_error()
for k in sorted({1, ndev}):
    multi(k)
    t0 = time.perf_counter()
    for _ in range(3): multi(k)
    print(f"cf_lowess_smooth_multi ndev={k}: {(time.perf_counter() - t0) / 3 * 1e3:.2f} ms", flush=True)
lib.cf_lowess_multi_release()
