#!/usr/bin/env python3
"""Author the canonical input files (the reference ships none: SURVEY.md App. B).

Grid file format: reference SRC/Grid.cpp:234-290; simulation file: SRC/FluidSolver.cpp:626-669.
"""
import os, sys

def cavity_grid(n, path):
    with open(path, "w") as f:
        f.write("Vertices {\n0 0\n0 1\n1 1\n1 0\n}\nNx {\n0 1 %d -1\n}\nNy {\n0 1 %d -1\n}\n" % (n, n))

def cavity_sim(path, dt, re, final_time, save_iter=1000000, lid=1.0):
    with open(path, "w") as f:
        f.write("BC {\n2 0\n2 %g\n2 0\n2 0\n}\ndt %.17g\nfinal_time %.17g\nre %g\nsaveIter %d\n" % (lid, dt, final_time, re, save_iter))

def main(out):
    os.makedirs(out, exist_ok=True)
    for n in (16, 32, 64, 128, 256, 512, 1024):
        cavity_grid(n, os.path.join(out, "cavity_%d.grid" % n))
    # "config 1" of this repo (our choice, the reference has no default case): 128^2, Re 100, dt 0.005
    cavity_sim(os.path.join(out, "cavity_re100.sim"), 0.005, 100, 10.0, 50)
    for n in (16, 32, 64, 128, 256, 512, 1024):
        cavity_sim(os.path.join(out, "cavity_re1000_%d.sim" % n), 0.25 / n, 1000, 20 * 0.25 / n + 1e-12)
    # channel: inlet (left, u=1) / wall / outflow (right) / wall ; non-uniform x with a geometric segment
    with open(os.path.join(out, "channel.grid"), "w") as f:
        f.write("Vertices {\n0 0\n0 1\n8 1\n8 0\n}\nNx {\n0 2 48 -1\n2 8 0 1.03\n}\nNy {\n0 1 32 -1\n}\n")
    with open(os.path.join(out, "channel.sim"), "w") as f:
        f.write("BC {\n0 1\n2 0\n4 0\n2 0\n}\ndt 0.01\nfinal_time 0.2\nre 100\nsaveIter 1000000\n")
    # backward-facing step: L-shaped polygon (masked cells), inlet on the upper-left edge
    with open(os.path.join(out, "step.grid"), "w") as f:
        f.write("Vertices {\n0 0.5\n0 1\n4 1\n4 0\n1 0\n1 0.5\n}\nNx {\n0 1 24 -1\n1 4 72 -1\n}\nNy {\n0 0.5 16 -1\n0.5 1 16 -1\n}\n")
    with open(os.path.join(out, "step.sim"), "w") as f:
        f.write("BC {\n0 1\n2 0\n4 0\n2 0\n2 0\n2 0\n}\ndt 0.005\nfinal_time 0.1\nre 100\nsaveIter 1000000\n")
    # outflow on the other three sides (the F_OW / F_OS / F_ON rows of LHS_phi): flow towards -x, -y and +y
    with open(os.path.join(out, "channel_west.grid"), "w") as f:
        f.write("Vertices {\n0 0\n0 1\n3 1\n3 0\n}\nNx {\n0 3 48 -1\n}\nNy {\n0 1 24 -1\n}\n")
    with open(os.path.join(out, "channel_west.sim"), "w") as f:   # edges: left, top, right, bottom
        f.write("BC {\n4 0\n2 0\n0 -1\n2 0\n}\ndt 0.01\nfinal_time 0.2\nre 100\nsaveIter 1000000\n")
    with open(os.path.join(out, "channel_south.grid"), "w") as f:
        f.write("Vertices {\n0 0\n0 3\n1 3\n1 0\n}\nNx {\n0 1 24 -1\n}\nNy {\n0 1 16 -1\n1 3 0 1.04\n}\n")
    with open(os.path.join(out, "channel_south.sim"), "w") as f:
        f.write("BC {\n2 0\n0 -1\n2 0\n4 0\n}\ndt 0.01\nfinal_time 0.2\nre 100\nsaveIter 1000000\n")
    with open(os.path.join(out, "channel_north.grid"), "w") as f:
        f.write("Vertices {\n0 0\n0 3\n1 3\n1 0\n}\nNx {\n0 1 24 -1\n}\nNy {\n0 3 48 -1\n}\n")
    with open(os.path.join(out, "channel_north.sim"), "w") as f:
        f.write("BC {\n2 0\n4 0\n2 0\n0 1\n}\ndt 0.01\nfinal_time 0.2\nre 100\nsaveIter 1000000\n")
    # wall-refined lid-driven cavity: geometric stretching towards all four walls (ratio 1.04, h_max / h_min = 3.4)
    with open(os.path.join(out, "cavity_stretched_64.grid"), "w") as f:
        f.write("Vertices {\n0 0\n0 1\n1 1\n1 0\n}\nNx {\n0 0.5 32 1.04\n0.5 1 32 0.9615384615\n}\nNy {\n0 0.5 32 1.04\n0.5 1 32 0.9615384615\n}\n")
    cavity_sim(os.path.join(out, "cavity_stretched_64.sim"), 0.005, 100, 1.0, 1000000)

if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else os.path.dirname(os.path.abspath(__file__)))
