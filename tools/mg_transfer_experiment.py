"""CPU experiment on the host emulation (tests/emul): cycle counts of the phi multigrid with other grid transfers.  Needs a build of
tests/emul/emul.cpp with the prolongation loop of vcycle() switched by EMUL_PROLONG (0: piecewise constant, as shipped; 1: bilinear,
weights 9/16 3/16 3/16 1/16, clamped at the walls) -- see DESIGN.md section 7 for the numbers; kept as the record of how they were made.
  EMUL_PROLONG=1 NU1=2 NU2=2 python tools/mg_transfer_experiment.py 512 6"""
import sys, os, numpy as np
sys.path.insert(0,'/root/repo/tests')
import refapi
refapi.EMUL_SO = '/tmp/libnsemul_alpha.so'
n = int(sys.argv[1]); steps = int(sys.argv[2])
em = refapi.Emul(np.full(n,1.0/n), np.full(n,1.0/n), [-1,0,1,0],[0,1,0,-1],[2,2,2,2],[0,1.0,0,0], 0.25/n, 1000.0, nu_pre=int(os.environ.get("NU1","2")), nu_post=int(os.environ.get("NU2","2")))
cyc=[]
for s in range(steps):
    mm, worst, c = em.step(1, rtol=1e-8)
    cyc.append(c)
print("alpha", os.environ.get("EMUL_ALPHA","1"), "nu", os.environ.get("NU1","2"), os.environ.get("NU2","2"), "n", n, "cycles (u,v,phi) per step:", cyc)
