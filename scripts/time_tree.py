"""Times both evaluators (direct kernel / moment tree) on one resident series.
    python scripts/time_tree.py n f nsteps [direct_too] [delta]      (CF_TT_X=random: sorted random x)"""
import sys, os
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
import numpy as np
from curvefit_b200 import _lib
if os.environ.get("CF_DEV_LIB"):        # the -DCF_DEVELOPMENT build (make DEV=1): reads CF_TREE_EVAL, CF_TREE_BITS, ...
    _lib.LIB_PATH = os.path.join(os.path.dirname(__file__), "dev", "libcurvefit_cuda.so")
from curvefit_b200 import datagen, lowess_smoothing

n = int(float(sys.argv[1])); f = float(sys.argv[2]); nsteps = int(sys.argv[3])
both = len(sys.argv) > 4 and sys.argv[4] == "1"
delta = float(sys.argv[5]) if len(sys.argv) > 5 else 0.0
x = np.sort(datagen.random_x(n)) if os.environ.get("CF_TT_X") == "random" else datagen.grid_x(n)
y = datagen.noisy_y(x)
res = {}
for m in (["direct"] if both else []) + ["tree"]:
    fit = lowess_smoothing(); fit.set_method(m); fit.initialize(x, y, srt=False)
    if delta > 0: fit.set_delta(delta)
    fit.run_resident(f, nsteps)
    best = None
    for _ in range(3):
        st = fit.run_resident(f, nsteps)
        if best is None or st.ms_total < best.ms_total: best = st
    if delta > 0: print(f"delta={delta}: {fit.fitted_points().size} fitted points")
    print(f"{m:7s} n={n} f={f} nsteps={nsteps} ns={best.ns}: total {best.ms_total:.3f} ms  windows {best.ms_windows:.3f}  "
          f"fit {best.ms_fit:.3f} (tree build {best.ms_tree_build:.3f})  robust {best.ms_robust:.3f}  launches {best.launches}  -> {n / best.ms_total * 1e3:.4g} points/s", flush=True)
    res[m] = fit.smooth(f, nsteps=nsteps)
if both:
    d = np.max(np.abs(res["tree"] - res["direct"])); print("max |tree - direct| =", d, "rel", d / np.max(np.abs(res["direct"])))
