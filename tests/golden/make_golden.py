#!/usr/bin/env python3
"""Writes the committed golden vectors from the oracle (unmodified reference + PETSc stand-in, exact mode rtol 1e-12).
Run in this container only (needs oracle/_ref/libnsref.so, which needs /root/reference):  python tests/golden/make_golden.py"""
import os, sys, tempfile
import numpy as np
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import refapi
from refapi import CASES, U, V, PHI, RHS_U, RHS_V, RHS_PHI, CONV_U0, CONV_V0, GRADPHI_X, GRADPHI_Y

with tempfile.TemporaryDirectory() as d:
    os.chdir(d)
    r = refapi.RefSolver(f"{CASES}/cavity_32.grid", f"{CASES}/cavity_re100.sim")
    r.set_rtol(1e-12)
    mm = r.step(5)
    np.savez_compressed(os.path.join(HERE, "cavity32_re100_5steps.npz"), u=r.get(U), v=r.get(V), phi=r.get(PHI), minmax=mm,
                        conv_u0=r.get(CONV_U0), conv_v0=r.get(CONV_V0), gpx=r.get(GRADPHI_X), gpy=r.get(GRADPHI_Y))
    # solver-free phases on the state after 5 steps, for all three geometries
    for case, sim in (("cavity_32", "cavity_re100"), ("channel", "channel"), ("step", "step")):
        r = refapi.RefSolver(f"{CASES}/{case}.grid", f"{CASES}/{sim}.sim")
        rng = np.random.default_rng(2024)
        st = {k: 0.3 * rng.standard_normal(r.n) for k in (U, V, CONV_U0, CONV_V0, GRADPHI_X, GRADPHI_Y)}
        for k, a in st.items():
            r.set(k, a)
        r.construct_rhs_v()
        out = dict(u=st[U], v=st[V], conv_u0=st[CONV_U0], conv_v0=st[CONV_V0], gpx=st[GRADPHI_X], gpy=st[GRADPHI_Y],
                   rhs_u=r.get(RHS_U), rhs_v=r.get(RHS_V), conv_u1=r.get(CONV_U0), conv_v1=r.get(CONV_V0))
        np.savez_compressed(os.path.join(HERE, f"phases_{case}.npz"), **out)
print("golden vectors written to", HERE)
