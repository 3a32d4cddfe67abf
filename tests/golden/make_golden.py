"""Generates tests/golden/*.npz with the C oracle (oracle/lowess_oracle.c).

PARITY UNPINNED: the reference has no golden vectors for LOWESS and cannot be
compiled here (no Fortran compiler), so these are OUR frozen vectors: outputs of
the line-by-line C restatement on seeded inputs.  They pin the oracle against
accidental change and travel to the GPU box (which has no /root/reference).

    python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from curvefit_b200 import datagen  # noqa: E402
from oracle import oracle as o  # noqa: E402
from tests import cases  # noqa: E402


def main():
    o.build(force=True)
    # C1: README Example 2 / test_lowess recipe, n = 100, f = 0.2 and 0.8, nsteps = 2
    x, y = datagen.example2(100)
    out = {"x": x, "y": y}
    for f in (0.2, 0.8):
        r = o.lowess(x, y, f, 2, want_windows=True)
        tag = f"f{int(f * 10):02d}"
        for k in ("ys", "rw", "res", "cmad", "nleft", "nright", "nrt"):
            out[f"{tag}_{k}"] = r[k]
    np.savez_compressed(os.path.join(HERE, "c1_example2.npz"), **out)

    # 4096-point case with outliers and non-uniform x, nsteps = 3
    x = cases.make_x("uniform", 4096, 7)
    y = cases.make_y(x, 7)
    r = o.lowess(x, y, 0.1, 3, want_windows=True)
    np.savez_compressed(os.path.join(HERE, "n4096_uniform.npz"), x=x, y=y, f=0.1, nsteps=3,
                        **{k: r[k] for k in ("ys", "rw", "res", "cmad", "nleft", "nright", "nrt")})

    # heavy ties, n = 1000
    x = cases.make_x("ties", 1000, 3)
    y = cases.make_y(x, 3)
    r = o.lowess(x, y, 0.3, 2, want_windows=True)
    np.savez_compressed(os.path.join(HERE, "n1000_ties.npz"), x=x, y=y, f=0.3, nsteps=2,
                        **{k: r[k] for k in ("ys", "rw", "res", "cmad", "nleft", "nright", "nrt")})
    print("golden vectors written to", HERE)


if __name__ == "__main__":
    main()
