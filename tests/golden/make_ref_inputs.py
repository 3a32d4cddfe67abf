"""Writes the binary inputs oracle/ref/ref_driver.f90 reads (see its header): the cases the
oracle is to be pinned on against the untouched reference wherever a Fortran compiler exists.
    python tests/golden/make_ref_inputs.py OUTDIR
Seeded (splitmix64 data generator of curvefit_b200/datagen.py), so the inputs are reproducible."""
import os
import struct
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from tests import cases                      # noqa: E402

# name -> (kind, n, f, srt, seed); nsteps = 2, delta = 0 are fixed by the reference's smooth()
CASES = {
    "c1_f02": ("grid", 100, 0.2, 0, 0),              # BASELINE configs[0], README Example 2
    "c1_f08": ("grid", 100, 0.8, 0, 0),
    "n4096_uniform": ("uniform", 4096, 0.3, 0, 1),
    "n1000_ties": ("ties", 1000, 0.25, 0, 2),
    "n1500_clustered_srt": ("clustered", 1500, 0.1, 1, 3),   # unsorted input, srt = .true.
}


def make(name):
    kind, n, f, srt, seed = CASES[name]
    x = cases.make_x(kind, n, seed)
    y = cases.make_y(x, seed)
    if srt:
        p = np.random.default_rng(77 + seed).permutation(n)
        x, y = x[p].copy(), y[p].copy()
    return x, y, f, srt


def main(outdir):
    os.makedirs(outdir, exist_ok=True)
    for name in CASES:
        x, y, f, srt = make(name)
        with open(os.path.join(outdir, f"in_{name}.bin"), "wb") as fh:
            fh.write(struct.pack("<idi", x.size, f, srt))
            fh.write(x.astype("<f8").tobytes())
            fh.write(y.astype("<f8").tobytes())


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "oracle", "_ref"))
