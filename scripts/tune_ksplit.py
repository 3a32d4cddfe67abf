"""Development helper (DEV library, CF_FIT_KSPLIT): time of ONE direct fit pass for several window
splits K, on the whole series (nranks = 1) and on one interior shard of 8 -- the geometry a rank of an
8-GPU run sees.  python scripts/tune_ksplit.py [n] [f]"""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from curvefit_b200 import _lib
_lib.LIB_PATH = os.path.join(os.path.dirname(__file__), "dev", "libcurvefit_cuda.so")
from curvefit_b200 import datagen, lowess_smoothing


def main():
    n = int(float(sys.argv[1])) if len(sys.argv) > 1 else 1_000_000
    f = float(sys.argv[2]) if len(sys.argv) > 2 else 0.05
    x = datagen.grid_x(n)
    y = datagen.noisy_y(x)
    ns = int(f * n)
    lib = _lib.load()
    for nranks, ks_list in [(1, [0, 3, 5, 6, 7, 8, 9, 10, 12, 13]), (8, [0, 12, 24, 29, 31, 33, 36, 41, 43, 45, 48, 49]),
                            (4, [0, 12, 18, 21, 24]), (2, [0, 6, 9, 12, 18])]:
        for ks in ks_list:
            os.environ["CF_FIT_KSPLIT"] = str(ks)
            fit = lowess_smoothing()
            fit.initialize(x, y, srt=False)
            if nranks > 1:
                assert lib.cf_lowess_set_shard(fit._h, nranks // 2, nranks) == 0
            pts = n // nranks
            best = 1e30
            for _ in range(4):
                st = fit.run_resident(f, 0)              # one pass (a shard without peers may only run one)
                best = min(best, st.ms_fit)
            tf = 17.0 * pts * ns / (best * 1e-3) / 1e12
            print(f"nranks={nranks} ksplit={ks:2d} fit_ms={best:.3f} TF={tf:.2f}", flush=True)
            fit.close()


if __name__ == "__main__":
    main()
