// TEST INFRASTRUCTURE ONLY -- host compile of the per-cell DEVICE functions (csrc/ns_cell.cuh) and of the
// host-side geometry / multigrid plan (csrc/ns_hostgeom.h), so that the CPU-only test run can check the very
// arithmetic the sm_100a kernels execute against the oracle.  Nothing in navierstokessolver_b200/ or
// libnsgpu.so links, loads or calls this file; it is not a fallback path.  The loops below mirror the kernel
// launch structure of csrc/ns_solver.inl one to one (same red-black order, same restriction/prolongation).
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <vector>
#include <string>
#include "ns_hostgeom.h"

using namespace nsgpu;

namespace {

struct ELevel {
  GeomArrays A;
  LevelGeom g;
  int cx = 1, cy = 1;
  std::vector<double> x, xt, b;
};

struct Emul {
  HostPlan plan;
  std::vector<ELevel> lv;
  double dt, re;
  int nu_pre = 2, nu_post = 2;
  std::vector<double> f[13];   // NSGPU_* ids 0..11, 12 = tmp
  std::string err;
  // outflow domains: w = ILU(0)(LHS_phi) * 1 in the padded layout, (w, w), sum of phi before the last mean removal
  bool aug = false;       // outflow edge: Krylov on the augmented operator
  bool krylov_phi = false;   // phi solve by BiCGStab (aug, or a single-level grid)
  bool aug_rhs = false;   // no outflow, stretched grid: b - c w with c = (vol, b) / (vol, w), then the ordinary multigrid
  std::vector<double> aug_w;
  double aug_ww = 0.0, aug_shift = 0.0, aug_yw = 0.0;
};

void bind(ELevel& L, const HostPlan& P) {
  geom_scalars(L.A, P, L.g);
  L.g.hx = L.A.hx.data(); L.g.hy = L.A.hy.data(); L.g.rhx = L.A.rhx.data(); L.g.rhy = L.A.rhy.data();
  L.g.rsx = L.A.rsx.data(); L.g.rsy = L.A.rsy.data(); L.g.ax = L.A.ax.data(); L.g.ay = L.A.ay.data();
  L.g.bx = L.A.bx.data(); L.g.by = L.A.by.data();
  L.g.flags = L.A.flags.empty() ? nullptr : L.A.flags.data();
  L.g.cedges = L.A.cedges.empty() ? nullptr : L.A.cedges.data();
}

template <bool RECT> unsigned fl(const LevelGeom& g, int i, int j) { return cell_flags<RECT>(g, i, j); }

struct Coef { int op; double alpha, beta; };

template <int OP, bool RECT>
void smoothphase_t(const LevelGeom& g, std::vector<double>& x, std::vector<double>& xt, const std::vector<double>& b, double alpha, double beta, int n) {
  for (int s = 0; s < n; s++) {
    xt = x;   // xin stays untouched: outflow rows read their second inward cell from it
    for (int colour = 0; colour < 2; colour++)
      for (int i = 0; i < g.nx; i++)
        for (int j = 0; j < g.ny; j++) {
          if (((g.gi0 + i + j) & 1) != colour) continue;
          const unsigned f = fl<RECT>(g, i, j);
          if (f == F_OUTSIDE) continue;
          const size_t id = cidx(g, i, j);
          xt[id] = relax_cell<OP>(g, f, i, j, xt[id - g.ld], xt[id + g.ld], xt[id - 1], xt[id + 1], x.data() + id, b[id], alpha, beta);
        }
    x.swap(xt);
  }
}
void smooth(const Emul& E, const ELevel& L, Coef c, std::vector<double>& x, std::vector<double>& xt, const std::vector<double>& b, int n) {
  if (c.op == 0) { if (E.plan.rect) smoothphase_t<0, true>(L.g, x, xt, b, c.alpha, c.beta, n); else smoothphase_t<0, false>(L.g, x, xt, b, c.alpha, c.beta, n); }
  else if (c.op == 2) { if (E.plan.rect) smoothphase_t<2, true>(L.g, x, xt, b, c.alpha, c.beta, n); else smoothphase_t<2, false>(L.g, x, xt, b, c.alpha, c.beta, n); }
  else { if (E.plan.rect) smoothphase_t<1, true>(L.g, x, xt, b, c.alpha, c.beta, n); else smoothphase_t<1, false>(L.g, x, xt, b, c.alpha, c.beta, n); }
}

template <int OP, bool RECT>
double apply_at(const LevelGeom& g, int i, int j, const std::vector<double>& x, double alpha, double beta) {
  return apply_cell<OP>(g, fl<RECT>(g, i, j), i, j, x.data() + cidx(g, i, j), alpha, beta);
}
double apply_any(const Emul& E, const LevelGeom& g, Coef c, int i, int j, const std::vector<double>& x) {
  if (c.op == 0) return E.plan.rect ? apply_at<0, true>(g, i, j, x, c.alpha, c.beta) : apply_at<0, false>(g, i, j, x, c.alpha, c.beta);
  if (c.op == 2) return E.plan.rect ? apply_at<2, true>(g, i, j, x, c.alpha, c.beta) : apply_at<2, false>(g, i, j, x, c.alpha, c.beta);
  return E.plan.rect ? apply_at<1, true>(g, i, j, x, c.alpha, c.beta) : apply_at<1, false>(g, i, j, x, c.alpha, c.beta);
}
unsigned flags_any(const Emul& E, const LevelGeom& g, int i, int j) { return E.plan.rect ? fl<true>(g, i, j) : fl<false>(g, i, j); }

double resnorm2(const Emul& E, const ELevel& L, Coef c, const std::vector<double>& x, const std::vector<double>& b) {
  double s = 0;
  for (int i = 0; i < L.g.nx; i++)
    for (int j = 0; j < L.g.ny; j++) {
      if (flags_any(E, L.g, i, j) == F_OUTSIDE) continue;
      const double r = b[cidx(L.g, i, j)] - apply_any(E, L.g, c, i, j, x);
      s += r * r;
    }
  return s;
}

void vcycle(Emul& E, int l, Coef c, std::vector<double>& x, std::vector<double>& xt, const std::vector<double>& b) {
  ELevel& L = E.lv[l];
  if (l == (int)E.lv.size() - 1) {
    smooth(E, L, c, x, xt, b, E.lv.size() == 1 ? E.nu_pre + E.nu_post : coarsest_sweeps(L.g.gnx, L.g.ny));
    return;
  }
  ELevel& C = E.lv[l + 1];
  smooth(E, L, c, x, xt, b, E.nu_pre);
  for (int I = 0; I < C.g.nx; I++)
    for (int J = 0; J < C.g.ny; J++) {
      double acc = 0;
      for (int di = 0; di < L.cx; di++)
        for (int dj = 0; dj < L.cy; dj++) {
          const int i = I * L.cx + di, j = J * L.cy + dj;
          if (flags_any(E, L.g, i, j) == F_OUTSIDE) continue;
          acc += b[cidx(L.g, i, j)] - apply_any(E, L.g, c, i, j, x);
        }
      C.b[cidx(C.g, I, J)] = flags_any(E, C.g, I, J) == F_OUTSIDE ? 0.0 : acc / (double)(L.cx * L.cy);
    }
  std::fill(C.x.begin(), C.x.end(), 0.0);
  vcycle(E, l + 1, c, C.x, C.xt, C.b);
  for (int i = 0; i < L.g.nx; i++)
    for (int j = 0; j < L.g.ny; j++)
      if (flags_any(E, L.g, i, j) != F_OUTSIDE) x[cidx(L.g, i, j)] += C.x[cidx(C.g, i / L.cx, j / L.cy)];
  smooth(E, L, c, x, xt, b, E.nu_post);
}

int solve_mg(Emul& E, Coef c, std::vector<double>& x, const std::vector<double>& b, double rtol, int max_cycles, double* rel) {
  ELevel& L = E.lv[0];
  double bn = 0;
  for (double v : b) bn += v * v;   // padding and masked cells hold zeros
  bn = std::sqrt(bn);
  double rn = std::sqrt(resnorm2(E, L, c, x, b));
  int cycles = 0;
  if (bn == 0) { std::fill(x.begin(), x.end(), 0.0); *rel = 0; return 0; }
  while (rn > rtol * bn && cycles < max_cycles) {
    vcycle(E, 0, c, x, E.f[12], b);
    cycles++;
    rn = std::sqrt(resnorm2(E, L, c, x, b));
  }
  *rel = rn / bn;
  return cycles;
}

// ---- outflow domains (mirrors solve_bicgstab's augmented branch in csrc/ns_solver.inl) ------------------------------
// Right-preconditioned BiCGStab on  A^ x = A x + mean(x) w  (nonsingular; x - mean(x) is the reference's phi, mean(x) its c),
// preconditioner  P v = V_walls(v - mean v), mean removed, + (v.w)/(w.w) * 1  (V_walls: one V-cycle of the wall-only L_phi, op 2).
double fluid_sum(const Emul& E, const std::vector<double>& a) { double s = 0; for (double v : a) s += v; (void)E; return s; }
void add_const(const Emul& E, std::vector<double>& a, double c) {
  const LevelGeom& g = E.lv[0].g;
  for (int i = 0; i < g.nx; i++)
    for (int j = 0; j < g.ny; j++)
      if (flags_any(E, g, i, j) != F_OUTSIDE) a[cidx(g, i, j)] += c;
}
void aug_apply(const Emul& E, const std::vector<double>& x, std::vector<double>& y) {
  const LevelGeom& g = E.lv[0].g;
  const Coef c{1, 0.0, -1.0};
  const double m = fluid_sum(E, x) / (double)E.plan.n_cells_global;
  std::fill(y.begin(), y.end(), 0.0);
  for (int i = 0; i < g.nx; i++)
    for (int j = 0; j < g.ny; j++)
      if (flags_any(E, g, i, j) != F_OUTSIDE) { const size_t id = cidx(g, i, j); y[id] = apply_any(E, g, c, i, j, x) + (E.aug ? m * E.aug_w[id] : 0.0); }
}
void aug_precond(Emul& E, const std::vector<double>& v, std::vector<double>& z) {
  const double n = (double)E.plan.n_cells_global;
  double vw = 0; if (E.aug) for (size_t k = 0; k < v.size(); k++) vw += v[k] * E.aug_w[k];
  std::vector<double> t(v);
  add_const(E, t, -fluid_sum(E, v) / n);
  std::fill(z.begin(), z.end(), 0.0);
  vcycle(E, 0, Coef{2, 0.0, -1.0}, z, E.f[12], t);
  add_const(E, z, -fluid_sum(E, z) / n + (E.aug ? vw / E.aug_ww : 0.0));
}
double dotv(const std::vector<double>& a, const std::vector<double>& b) { double s = 0; for (size_t k = 0; k < a.size(); k++) s += a[k] * b[k]; return s; }

int solve_phi_aug(Emul& E, std::vector<double>& x, const std::vector<double>& b, double rtol, int max_it, double* rel) {
  const size_t S = x.size();
  std::vector<double> r(S), r0, p(S, 0.0), v(S, 0.0), s(S), t(S), ph(S), sh(S), y(S);
  add_const(E, x, E.aug_shift / (double)E.plan.n_cells_global);   // warm start: last step's phi plus last step's mean(x)
  const double bn = std::sqrt(dotv(b, b));
  if (bn == 0) { std::fill(x.begin(), x.end(), 0.0); *rel = 0; return 0; }
  aug_apply(E, x, y);
  for (size_t k = 0; k < S; k++) r[k] = b[k] - y[k];
  r0 = r;
  double rho = dotv(r0, r), rho_old = 1, alpha = 1, omega = 1, rn = std::sqrt(dotv(r, r));
  int it = 0;
  while (rn > rtol * bn && it < max_it) {
    const double beta = (rho / rho_old) * (alpha / omega);
    for (size_t k = 0; k < S; k++) p[k] = r[k] + beta * (p[k] - omega * v[k]);
    aug_precond(E, p, ph);
    aug_apply(E, ph, v);
    alpha = rho / dotv(r0, v);
    for (size_t k = 0; k < S; k++) s[k] = r[k] - alpha * v[k];
    aug_precond(E, s, sh);
    aug_apply(E, sh, t);
    omega = dotv(t, s) / dotv(t, t);
    for (size_t k = 0; k < S; k++) { x[k] += alpha * ph[k] + omega * sh[k]; r[k] = s[k] - omega * t[k]; }
    rho_old = rho; rho = dotv(r0, r);
    rn = std::sqrt(dotv(r, r));
    it++;
    if (!(rn == rn)) break;
  }
  aug_apply(E, x, y);
  double tr = 0; for (size_t k = 0; k < S; k++) { const double d = b[k] - y[k]; tr += d * d; }
  *rel = std::sqrt(tr) / bn;
  E.aug_shift = fluid_sum(E, x);
  return it;
}

template <bool RECT>
void rhsphase_t(Emul& E) {
  const LevelGeom& g = E.lv[0].g;
  const StepConst sc{E.dt, E.re};
  std::vector<double> nu(E.f[NSGPU_CONV_U0]), nv(E.f[NSGPU_CONV_V0]);
  for (int i = 0; i < g.nx; i++)
    for (int j = 0; j < g.ny; j++) {
      if (fl<RECT>(g, i, j) == F_OUTSIDE) continue;
      const size_t id = cidx(g, i, j);
      RhsOut o = rhs_cell<RECT>(g, E.plan.edges.data(), sc, i, j, E.f[NSGPU_U].data() + id, E.f[NSGPU_V].data() + id, nu[id], nv[id],
                                E.f[NSGPU_GRADPHI_X].data() + id, E.f[NSGPU_GRADPHI_Y].data() + id);
      E.f[NSGPU_RHS_U][id] = o.rhs_u; E.f[NSGPU_RHS_V][id] = o.rhs_v;
      E.f[NSGPU_CONV_U0][id] = o.nu; E.f[NSGPU_CONV_V0][id] = o.nv;
    }
}
template <bool RECT>
void divphase_t(Emul& E, bool remove_mean) {
  const LevelGeom& g = E.lv[0].g;
  const StepConst sc{E.dt, E.re};
  double sum = 0;
  for (int i = 0; i < g.nx; i++)
    for (int j = 0; j < g.ny; j++) {
      if (fl<RECT>(g, i, j) == F_OUTSIDE) continue;
      const size_t id = cidx(g, i, j);
      const double d = div_cell<RECT>(g, E.plan.edges.data(), sc, i, j, E.f[NSGPU_USTAR].data() + id, E.f[NSGPU_VSTAR].data() + id);
      E.f[NSGPU_RHS_PHI][id] = d;
      sum += d;
    }
  if (remove_mean) {
    const double m = sum / (double)E.plan.n_cells_global;
    for (int i = 0; i < g.nx; i++)
      for (int j = 0; j < g.ny; j++)
        if (fl<RECT>(g, i, j) != F_OUTSIDE) E.f[NSGPU_RHS_PHI][cidx(g, i, j)] -= m;
  }
}
template <bool RECT>
void correctphase_t(Emul& E, double* mm) {
  const LevelGeom& g = E.lv[0].g;
  double umin = 1e300, umax = -1e300, vmin = 1e300, vmax = -1e300;
  for (int i = 0; i < g.nx; i++)
    for (int j = 0; j < g.ny; j++) {
      const unsigned f = fl<RECT>(g, i, j);
      if (f == F_OUTSIDE) continue;
      const size_t id = cidx(g, i, j);
      const UV gr = grad_cell(g, f, i, j, E.f[NSGPU_PHI].data() + id);
      E.f[NSGPU_GRADPHI_X][id] = gr.u; E.f[NSGPU_GRADPHI_Y][id] = gr.v;
      const double un = E.f[NSGPU_USTAR][id] - E.dt * gr.u, vn = E.f[NSGPU_VSTAR][id] - E.dt * gr.v;
      E.f[NSGPU_U][id] = un; E.f[NSGPU_V][id] = vn;
      umin = std::min(umin, un); umax = std::max(umax, un); vmin = std::min(vmin, vn); vmax = std::max(vmax, vn);
    }
  if (mm) { mm[0] = umin; mm[1] = umax; mm[2] = vmin; mm[3] = vmax; }
}

void phi_mean_zero(Emul& E) {
  const LevelGeom& g = E.lv[0].g;
  double s = 0;
  for (double v : E.f[NSGPU_PHI]) s += v;
  const double m = s / (double)E.plan.n_cells_global;
  for (int i = 0; i < g.nx; i++)
    for (int j = 0; j < g.ny; j++)
      if (flags_any(E, g, i, j) != F_OUTSIDE) E.f[NSGPU_PHI][cidx(g, i, j)] -= m;
}

}  // namespace

extern "C" {

void* emul_create_rank(const nsgpu_grid* G, double dt, double re, int nu_pre, int nu_post, int max_levels, int rank, int world, char* errbuf, int errlen);
void* emul_create(const nsgpu_grid* G, double dt, double re, int nu_pre, int nu_post, int max_levels, char* errbuf, int errlen) {
  return emul_create_rank(G, dt, re, nu_pre, nu_post, max_levels, 0, 1, errbuf, errlen);
}
// x-slab `rank` of `world` (the host plan libnsgpu.so uses for several GPUs)
void* emul_create_rank(const nsgpu_grid* G, double dt, double re, int nu_pre, int nu_post, int max_levels, int rank, int world, char* errbuf, int errlen) {
  Emul* E = new Emul;
  E->dt = dt; E->re = re;
  if (nu_pre > 0) E->nu_pre = nu_pre;
  if (nu_post > 0) E->nu_post = nu_post;
  int rc = make_host_plan(G, rank, world, max_levels, E->plan, E->err);
  if (rc) {
    if (errbuf && errlen > 0) { strncpy(errbuf, E->err.c_str(), errlen - 1); errbuf[errlen - 1] = 0; }
    delete E;
    return nullptr;
  }
  E->lv.resize(E->plan.levels.size());
  for (size_t l = 0; l < E->lv.size(); l++) {
    ELevel& L = E->lv[l];
    make_geom_arrays(E->plan.levels[l], E->plan.lgi0[l], E->plan.lnx[l], (l == 0 && !E->plan.rect) ? &E->plan.cedges0 : nullptr, L.A);
    bind(L, E->plan);
    L.cx = E->plan.lcx[l]; L.cy = E->plan.lcy[l];
    if (l > 0) { L.x.assign(L.A.stride, 0.0); L.xt.assign(L.A.stride, 0.0); L.b.assign(L.A.stride, 0.0); }
  }
  for (auto& v : E->f) v.assign(E->lv[0].A.stride, 0.0);
  const bool outflow = plan_has_outflow(E->plan);
  if (outflow || !plan_is_uniform(E->plan)) {
    std::vector<double> w;
    rc = phi_ilu0_image_of_ones(G, E->plan, w, E->err);
    if (rc) { if (errbuf && errlen > 0) { strncpy(errbuf, E->err.c_str(), errlen - 1); errbuf[errlen - 1] = 0; } delete E; return nullptr; }
    const LevelGeom& g = E->lv[0].g;
    E->aug_w.assign(E->lv[0].A.stride, 0.0);
    for (int i = 0; i < g.nx; i++)
      for (int j = 0; j < g.ny; j++) {
        const long id = E->plan.rect ? (long)i * g.ny + j : (long)E->plan.local_id[(size_t)i * g.ny + j];
        if (id >= 0) { const double wv = w[(size_t)(E->plan.id0 + id)]; E->aug_w[cidx(g, i, j)] = wv; }
      }
    for (double wv : w) E->aug_ww += wv * wv;
    E->aug_yw = volume_dot(G, E->plan, w);
    // Stretched grids go to the Krylov solve on the augmented operator as well (plain multigrid diverges beyond a stretch of ~5), and so
    // does a grid that cannot be coarsened at all.  EMUL_RHS_CONSISTENT=1 (NSGPU_RHS_CONSISTENT=1 in the library) keeps the multigrid
    // iteration on b - c w for mildly stretched grids (stretch <= 2).
    const bool mg_on_corrected_rhs = getenv("EMUL_RHS_CONSISTENT") && atoi(getenv("EMUL_RHS_CONSISTENT")) && !outflow && E->lv.size() > 1 &&
                                     (plan_stretch(E->plan) <= 2.0 || atoi(getenv("EMUL_RHS_CONSISTENT")) > 1);
    E->aug = !mg_on_corrected_rhs;
    E->aug_rhs = mg_on_corrected_rhs;
  }
  E->krylov_phi = E->aug || E->lv.size() == 1;   // uniform and uncoarsenable: plain preconditioned BiCGStab on the consistent system
  return E;
}
void emul_destroy(void* h) { delete (Emul*)h; }
// w = ILU(0)(LHS_phi) * 1 in compact local id order (zeros when the domain has no outflow edge); returns (w, w) over the global domain
double emul_phi_ilu_w(void* h, double* out) {
  Emul* E = (Emul*)h;
  const LevelGeom& g = E->lv[0].g;
  for (int i = 0; i < g.nx; i++)
    for (int j = 0; j < g.ny; j++) {
      const long id = E->plan.rect ? (long)i * g.ny + j : (long)E->plan.local_id[(size_t)i * g.ny + j];
      if (id >= 0) out[id] = (E->aug || E->aug_rhs) ? E->aug_w[cidx(g, i, j)] : 0.0;
    }
  return E->aug_ww;
}
int emul_rect(void* h) { return ((Emul*)h)->plan.rect ? 1 : 0; }
int emul_num_levels(void* h) { return (int)((Emul*)h)->lv.size(); }
long emul_n_local(void* h) { return (long)((Emul*)h)->plan.n_local; }
void emul_level_shape(void* h, int l, int* gnx, int* ny, int* cx, int* cy) {
  Emul* E = (Emul*)h; *gnx = E->lv[l].g.gnx; *ny = E->lv[l].g.ny; *cx = E->lv[l].cx; *cy = E->lv[l].cy;
}

void emul_set(void* h, int which, const double* in) {
  Emul* E = (Emul*)h;
  const LevelGeom& g = E->lv[0].g;
  std::fill(E->f[which].begin(), E->f[which].end(), 0.0);
  for (int i = 0; i < g.nx; i++)
    for (int j = 0; j < g.ny; j++) {
      const long id = E->plan.rect ? (long)i * g.ny + j : (long)E->plan.local_id[(size_t)i * g.ny + j];
      if (id >= 0) E->f[which][cidx(g, i, j)] = in[id];
    }
}
void emul_get(void* h, int which, double* out) {
  Emul* E = (Emul*)h;
  const LevelGeom& g = E->lv[0].g;
  for (int i = 0; i < g.nx; i++)
    for (int j = 0; j < g.ny; j++) {
      const long id = E->plan.rect ? (long)i * g.ny + j : (long)E->plan.local_id[(size_t)i * g.ny + j];
      if (id >= 0) out[id] = E->f[which][cidx(g, i, j)];
    }
}

void emul_rhs_build(void* h) { Emul* E = (Emul*)h; if (E->plan.rect) rhsphase_t<true>(*E); else rhsphase_t<false>(*E); }
void emul_div_rhs(void* h, int remove_mean) { Emul* E = (Emul*)h; if (E->plan.rect) divphase_t<true>(*E, remove_mean != 0); else divphase_t<false>(*E, remove_mean != 0); }
void emul_correct(void* h, double* mm) { Emul* E = (Emul*)h; if (E->plan.rect) correctphase_t<true>(*E, mm); else correctphase_t<false>(*E, mm); }

void emul_apply(void* h, int op, const double* x, double* y) {
  Emul* E = (Emul*)h;
  emul_set(h, 12, x);
  const LevelGeom& g = E->lv[0].g;
  Coef c = op == 0 ? Coef{0, 1.0, E->dt / (2.0 * E->re)} : Coef{op, 0.0, -1.0};   // 1 = LHS_phi, 2 = its wall-only variant
  for (int i = 0; i < g.nx; i++)
    for (int j = 0; j < g.ny; j++) {
      const long id = E->plan.rect ? (long)i * g.ny + j : (long)E->plan.local_id[(size_t)i * g.ny + j];
      if (id >= 0) y[id] = apply_any(*E, g, c, i, j, E->f[12]);
    }
}

// which: 0 = u* from rhs_u (initial guess u), 1 = v*, 2 = phi (warm start)
int emul_solve(void* h, int which, double rtol, int max_cycles, double* rel) {
  Emul* E = (Emul*)h;
  if (which == 2) {
    int n;
    if (E->krylov_phi) n = solve_phi_aug(*E, E->f[NSGPU_PHI], E->f[NSGPU_RHS_PHI], rtol, max_cycles, rel);
    else if (E->aug_rhs) {   // mirrors phase_phi (csrc/ns_solver.inl): one volume-weighted sum, one axpy, then the multigrid iteration
      const LevelGeom& g = E->lv[0].g;
      const std::vector<double>& b = E->f[NSGPU_RHS_PHI];
      double yb = 0.0;
      for (int i = 0; i < g.nx; i++)
        for (int j = 0; j < g.ny; j++)
          if (flags_any(*E, g, i, j) != F_OUTSIDE) yb += g.hx[i + HX] * g.hy[j + PJ] * b[cidx(g, i, j)];
      const double c = yb / E->aug_yw;
      std::vector<double> b2(b);
      for (size_t k = 0; k < b2.size(); k++) b2[k] -= c * E->aug_w[k];
      n = solve_mg(*E, Coef{1, 0.0, -1.0}, E->f[NSGPU_PHI], b2, rtol, max_cycles, rel);
    } else n = solve_mg(*E, Coef{1, 0.0, -1.0}, E->f[NSGPU_PHI], E->f[NSGPU_RHS_PHI], rtol, max_cycles, rel);
    phi_mean_zero(*E);
    return n;
  }
  const Coef c{0, 1.0, E->dt / (2.0 * E->re)};
  if (which == 0) { E->f[NSGPU_USTAR] = E->f[NSGPU_U]; return solve_mg(*E, c, E->f[NSGPU_USTAR], E->f[NSGPU_RHS_U], rtol, max_cycles, rel); }
  E->f[NSGPU_VSTAR] = E->f[NSGPU_V];
  return solve_mg(*E, c, E->f[NSGPU_VSTAR], E->f[NSGPU_RHS_V], rtol, max_cycles, rel);
}

// slab plan: out = {i0, i1, id0, n_local, n_levels, rep_start}; per level lnx[l], lgi0[l]
void emul_plan_info(void* h, long* out, int* lnx, int* lgi0) {
  Emul* E = (Emul*)h;
  out[0] = E->plan.i0; out[1] = E->plan.i1; out[2] = (long)E->plan.id0; out[3] = (long)E->plan.n_local;
  out[4] = (long)E->plan.levels.size(); out[5] = E->plan.rep_start;
  for (size_t l = 0; l < E->plan.levels.size(); l++) { lnx[l] = E->plan.lnx[l]; lgi0[l] = E->plan.lgi0[l]; }
}
int emul_ld(void* h) { return ((Emul*)h)->lv[0].g.ld; }
// `width` owned columns next to side 0 (left) / 1 (right) -> out[width * ld]
void emul_halo_pack(void* h, int which, int side, int width, double* out) {
  Emul* E = (Emul*)h;
  const LevelGeom& g = E->lv[0].g;
  const size_t first = side == 0 ? (size_t)HX * g.ld : (size_t)(HX + g.nx - width) * g.ld;
  std::copy(E->f[which].begin() + first, E->f[which].begin() + first + (size_t)width * g.ld, out);
}
// in[width * ld] -> the halo columns on that side
void emul_halo_unpack(void* h, int which, int side, int width, const double* in) {
  Emul* E = (Emul*)h;
  const LevelGeom& g = E->lv[0].g;
  const size_t first = side == 0 ? (size_t)(HX - width) * g.ld : (size_t)(HX + g.nx) * g.ld;
  std::copy(in, in + (size_t)width * g.ld, E->f[which].begin() + first);
}
// one colour of a red-black Gauss-Seidel sweep of L_phi on the slab, in place (halo columns are read, never written)
void emul_phi_half_sweep(void* h, int colour) {
  Emul* E = (Emul*)h;
  const LevelGeom& g = E->lv[0].g;
  std::vector<double>& x = E->f[NSGPU_PHI];
  const std::vector<double>& b = E->f[NSGPU_RHS_PHI];
  for (int i = 0; i < g.nx; i++)
    for (int j = 0; j < g.ny; j++) {
      if (((g.gi0 + i + j) & 1) != colour) continue;
      const unsigned f = flags_any(*E, g, i, j);
      if (f == F_OUTSIDE) continue;
      const size_t id = cidx(g, i, j);
      x[id] = E->plan.rect ? relax_cell<1>(g, f, i, j, x[id - g.ld], x[id + g.ld], x[id - 1], x[id + 1], x.data() + id, b[id], 0.0, -1.0)
                           : relax_cell<1>(g, f, i, j, x[id - g.ld], x[id + g.ld], x[id - 1], x[id + 1], x.data() + id, b[id], 0.0, -1.0);
    }
}

void emul_poisson_sweeps(void* h, int n, double* resid) {
  Emul* E = (Emul*)h;
  const Coef c{1, 0.0, -1.0};
  smooth(*E, E->lv[0], c, E->f[NSGPU_PHI], E->f[12], E->f[NSGPU_RHS_PHI], n);
  if (resid) *resid = std::sqrt(resnorm2(*E, E->lv[0], c, E->f[NSGPU_PHI], E->f[NSGPU_RHS_PHI]));
}

// n passes of the Solve() loop body (FS.cpp:546-557); returns the worst relative residual
double emul_step(void* h, int n, double rtol, int max_cycles, double* mm, int* cycles3) {
  Emul* E = (Emul*)h;
  double worst = 0, rel;
  for (int s = 0; s < n; s++) {
    emul_rhs_build(h);
    int cu = emul_solve(h, 0, rtol, max_cycles, &rel); worst = std::max(worst, rel);
    int cv = emul_solve(h, 1, rtol, max_cycles, &rel); worst = std::max(worst, rel);
    emul_div_rhs(h, 1);
    int cp = emul_solve(h, 2, rtol, max_cycles, &rel); worst = std::max(worst, rel);
    emul_correct(h, mm);
    if (cycles3) { cycles3[0] = cu; cycles3[1] = cv; cycles3[2] = cp; }
  }
  (void)E;
  return worst;
}

}  // extern "C"
