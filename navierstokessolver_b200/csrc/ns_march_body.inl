// ns_march_body.inl -- the march of one warp over its chunk (body of march_run; included, not called, so that a kernel can also hold it
// directly).  In scope: template constants NSW, PROLONG, RES, PF, SOR and `P`; g, mcy, xin, b, xout, xf, s_c, ib, ie, n0, nend, r0, lane, wj.
// Leaves the Marcher `m` behind (m.acc = this lane's share of |r|^2).
  using MB = Marcher<NSW, PROLONG, RES, PF, SOR, P>;
  constexpr int LB = MB::L;
  constexpr int HJB = march_halo(NSW, RES);
  constexpr int UJ = 64 - 2 * HJB;
  constexpr int HL = HJB / 2;
  // warps past the last strip run with every predicate false (no divergent exit: the shuffles stay convergent)
  MB m;
  m.xf = xf;
  m.ld = g.ld; m.ny = g.ny; m.ib = ib; m.ie = ie; m.nxl = g.nx;
  m.pfo = (size_t)MARCH_PFD * g.ld;
  m.pushL = P && xf.dL != 0 && ib < xf.zone;
  m.pushR = P && xf.dR != 0 && ie > g.nx - xf.zone;
  m.j0 = wj * UJ - HJB + 2 * lane;                            // even, >= -HJB >= -PJ
  const int jl = min(m.j0, (g.ny + PJ - 2) & ~1);            // lanes past the column end: an even (16-byte aligned) zero pair
  {
    const int a = jl + PJ;                                   // entries outside the domain have cs = cn = 0
    m.k.csa = mcy[3 * a]; m.k.cna = mcy[3 * a + 1]; m.k.dya = mcy[3 * a + 2];
    m.k.csb = mcy[3 * a + 3]; m.k.cnb = mcy[3 * a + 4]; m.k.dyb = mcy[3 * a + 5];
  }
  const size_t joff = (size_t)(jl + PJ);
  m.useful = lane >= HL && lane < 32 - HL;
  m.jstore2 = m.useful && (m.j0 + 1 < g.ny);
  m.jstore1 = m.useful && (m.j0 < g.ny) && !m.jstore2;
  m.cc = s_c + 3 * (n0 - r0);
  m.px = xin ? xin + (size_t)(n0 + HX) * g.ld + joff : nullptr;
  m.pb = b + (size_t)(n0 + HX) * g.ld + joff;
  m.po = xout + (size_t)(n0 - LB + HX) * g.ld + joff;        // only dereferenced inside [ib, ie)
  m.pbc = nullptr;
  if (RES == 1) m.pbc = xf.bc + (size_t)((xf.rcx == 2 ? (ib >> 1) : ib) + HX) * xf.ldbc + (xf.rcy == 2 ? (size_t)((m.j0 >> 1) + PJ) : joff);
  m.pxc = nullptr;
  if (PROLONG) m.pxc = xf.xc + (size_t)((xf.pcy == 2 ? ((jl + 2 * PJ) >> 1) - PJ : jl) + PJ);
  m.racc = make_double2(0.0, 0.0);
  m.acc = 0.0;
  m.xshift = xf.shift ? *xf.shift * xf.shift_scale : 0.0;
#pragma unroll
  for (int q = 0; q < MB::RL; q++) { m.X[q] = make_double2(0.0, 0.0); m.Bq[q] = make_double2(0.0, 0.0); }
#pragma unroll
  for (int p = 0; p < PF; p++) m.load_col(n0 + p, m.X[PF - 1 - p], m.Bq[PF - 1 - p]);

#pragma unroll 2
  for (int n = n0; n < nend; n += 2) {
    m.template step<true>(n);
    m.template step<false>(n + 1);
  }
