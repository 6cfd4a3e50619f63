"""Krylov route (BiCGStab on the bordered phi system) with device-side control: ms/step as CUDA graphs and with direct launches, next to a
cavity of the same cell count on the multigrid route.   python tools/krylov_bench.py [nx] [ny] [steps]"""
import os, sys, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import navierstokessolver_b200 as ns

nx = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
ny = int(sys.argv[2]) if len(sys.argv) > 2 else 256
steps = int(sys.argv[3]) if len(sys.argv) > 3 else 10
h = 1.0 / ny


def channel():   # inlet on the left (u = 1), walls top and bottom, outflow on the right; edges in the reference's order left, top, right, bottom
    return ns.GridDesc(np.full(nx, h), np.full(ny, h), [-1, 0, 1, 0], [0, 1, 0, -1], [0, 2, 4, 2], [1.0, 0.0, 0.0, 0.0])


def run(name, desc, **kw):
    s = ns.GpuFluidSolver(desc, 0.25 * h, 100.0, **kw)
    s.step(3)
    t0 = time.perf_counter()
    info = s.step(steps)
    wall = (time.perf_counter() - t0) / steps * 1e3
    rec = {"case": name, "cells": desc.n_cells, "ms_per_step_device": round(info.ms_total / steps, 3), "ms_per_step_wall": round(wall, 3),
           "iterations_last_step": [info.it_u, info.it_v, info.it_phi], "status": info.status, "flags": info.flags,
           "launches_per_step": info.kernel_launches // steps, "phases_ms": [round(x, 3) for x in s.phase_ms()]}
    print(json.dumps(rec), flush=True)
    s.close()
    return rec


a = run(f"channel {nx}x{ny}, graphs", channel(), use_graph=0)
b = run(f"channel {nx}x{ny}, direct launches", channel(), use_graph=-1)
c = run(f"cavity {nx}x{ny} (multigrid route)", ns.GridDesc.cavity(ny, nx=nx, lx=nx / float(ny)), use_graph=0)
print(json.dumps({"channel_over_cavity_per_cell": round(a["ms_per_step_device"] / c["ms_per_step_device"], 2),
                  "graphs_over_direct": round(a["ms_per_step_wall"] / b["ms_per_step_wall"], 2)}))
