"""Development helper: time one fit pass of a workload for several kernel variants
(CF_FIT_VARIANT / CF_FIT_KSPLIT are read when a handle is created)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

from curvefit_b200 import datagen, lowess_smoothing, regression


def main():
    n = int(float(sys.argv[1])) if len(sys.argv) > 1 else 1_000_000
    f = float(sys.argv[2]) if len(sys.argv) > 2 else 0.05
    x = datagen.grid_x(n)
    y = datagen.noisy_y(x)
    ns = int(f * n)
    print("fp64 peak", regression.fp64_peak(-1, 200.0))
    # variant 4: cluster/DSMEM reduction (K <= 8); 0: partial moments through HBM
    for variant, ks in [(4, 0), (4, 2), (4, 4), (4, 8), (0, 0), (0, 4), (0, 6), (0, 8), (0, 10), (0, 12), (0, 16)]:
        if True:
            os.environ["CF_FIT_VARIANT"] = str(variant)
            os.environ["CF_FIT_KSPLIT"] = str(ks)
            fit = lowess_smoothing()
            fit.initialize(x, y, srt=False)
            fit.run_resident(f, 1)
            best = 1e30
            for _ in range(3):
                st = fit.run_resident(f, 1)
                best = min(best, st.ms_fit / st.fit_launches)
            tf = 17.0 * n * ns / (best * 1e-3) / 1e12
            print(f"variant={variant} ksplit={ks} fit_ms={best:.3f} TF={tf:.2f} total_ms={st.ms_total:.3f}", flush=True)
            fit.close()


if __name__ == "__main__":
    main()
