#!/usr/bin/env python3
"""Golden vectors for BASELINE.json configs[1] (lid-driven cavity Re=1000 at 1024x1024, fields vs the CPU reference).

The oracle (unmodified reference + PETSc stand-in, rtol 1e-12) needs about a minute per step at this size, so it cannot run
inside the GPU tests; this script runs it once in this container and commits a SAMPLE of the fields after STEPS steps from rest:
every 16th cell in both directions, six full lines (first / middle / last column and row: the walls, the lid and the corner
singularities are where the fields are least smooth), the min/max the reference prints every step, and four functionals
of each full field (sum, sum of squares, max |.|, a fixed oscillating weight) so that cells outside the sample are pinned too.

  python tests/golden/make_golden_1024.py [steps]
"""
import os, sys, tempfile, time
import numpy as np
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import refapi
from refapi import CASES, U, V, PHI

N = 1024
STEPS = int(sys.argv[1]) if len(sys.argv) > 1 else 3
STRIDE = 16
LINES = (0, N // 2, N - 1)


def functionals(a):
    w = np.cos(0.37 * np.arange(a.size) ** 0.5)
    return np.array([a.sum(), (a * a).sum(), np.abs(a).max(), (a * w).sum()])


def sample(a):
    f = a.reshape(N, N)                       # id = i * ny + j  (G.cpp:149-161): row index = x, column index = y
    return dict(sub=f[::STRIDE, ::STRIDE].copy(), cols=f[list(LINES), :].copy(), rows=f[:, list(LINES)].copy(), fun=functionals(a))


if __name__ == "__main__":
    with tempfile.TemporaryDirectory() as d:
        os.chdir(d)
        r = refapi.RefSolver(f"{CASES}/cavity_{N}.grid", f"{CASES}/cavity_re1000_{N}.sim")
        r.set_rtol(1e-12)
        mm = []
        t0 = time.time()
        for s in range(STEPS):
            mm.append(r.step(1))
            print("step", s + 1, "minmax", mm[-1], "%.0f s" % (time.time() - t0), flush=True)
        out = dict(steps=STEPS, stride=STRIDE, lines=np.array(LINES), minmax=np.array(mm), dt=r.dt, re=r.re)
        for name, k in (("u", U), ("v", V), ("phi", PHI)):
            a = r.get(k)
            if name == "phi":
                a = a - a.mean()
            for key, val in sample(a).items():
                out[f"{name}_{key}"] = val
        np.savez_compressed(os.path.join(HERE, f"cavity{N}_re1000_{STEPS}steps.npz"), **out)
    print("written", f"cavity{N}_re1000_{STEPS}steps.npz")
