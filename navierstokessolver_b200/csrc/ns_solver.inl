// ns_solver.inl -- host orchestration of the kernels: halo exchange, multigrid V-cycle, the two linear solves
// that replace KSPSolve (reference FS.cpp:547-551) and the phases of one time step (FS.cpp:546-557).

namespace nsgpu {

struct OpCoef { int op; double alpha, beta; };
static inline OpCoef coef_v(const nsgpu_ctx* c) { return OpCoef{0, 1.0, c->dt / (2.0 * c->re)}; }   // LHS_V   FS.cpp:140-141
static inline OpCoef coef_phi() { return OpCoef{1, 0.0, -1.0}; }                                    // LHS_phi

#define NS_LAUNCH_CHECK()                                                             \
  do {                                                                                \
    cudaError_t e_ = cudaGetLastError();                                              \
    if (e_ != cudaSuccess) {                                                          \
      char buf_[256];                                                                 \
      snprintf(buf_, sizeof buf_, "%s:%d: kernel launch -> %s", __FILE__, __LINE__, cudaGetErrorString(e_)); \
      ctx->err = buf_;                                                                \
      return NSGPU_ERR_CUDA;                                                          \
    }                                                                                 \
    ctx->launches++;                                                                  \
  } while (0)

// dispatch on (operator, rectangle fast path)
#define NS_DISPATCH_OP_RECT(op, rect, CALL) \
  do {                                      \
    if ((op) == 0) { if (rect) { CALL(0, true); } else { CALL(0, false); } } \
    else { if (rect) { CALL(1, true); } else { CALL(1, false); } }           \
  } while (0)
#define NS_DISPATCH_RECT(rect, CALL) \
  do { if (rect) { CALL(true); } else { CALL(false); } } while (0)

static inline dim3 grid_cols(const LevelGeom& g, int threads, int max_blocks = 148 * 16) {
  unsigned gx = (unsigned)((g.ny + threads - 1) / threads);
  unsigned gy = (unsigned)std::max(1, std::min(g.nx, max_blocks / (int)gx));
  return dim3(gx, gy, 1);
}

// One-column-block halo exchange with the x-neighbours: `width` columns, contiguous (width * ld doubles).
static int halo_exchange(nsgpu_ctx* ctx, const Level& L, double* f, int width) {
  if (ctx->world == 1) return 0;
  const size_t n = (size_t)width * L.g.ld;
  const int nx = L.g.nx;
  NS_NCCL(ncclGroupStart());
  if (ctx->rank > 0) {
    NS_NCCL(ncclSend(f + (size_t)HX * L.g.ld, n, ncclDouble, ctx->rank - 1, ctx->comm, ctx->stream));
    NS_NCCL(ncclRecv(f + (size_t)(HX - width) * L.g.ld, n, ncclDouble, ctx->rank - 1, ctx->comm, ctx->stream));
  }
  if (ctx->rank < ctx->world - 1) {
    NS_NCCL(ncclSend(f + (size_t)(HX + nx - width) * L.g.ld, n, ncclDouble, ctx->rank + 1, ctx->comm, ctx->stream));
    NS_NCCL(ncclRecv(f + (size_t)(HX + nx) * L.g.ld, n, ncclDouble, ctx->rank + 1, ctx->comm, ctx->stream));
  }
  NS_NCCL(ncclGroupEnd());
  return 0;
}

static int allreduce(nsgpu_ctx* ctx, double* d, int n, ncclRedOp_t op) {
  if (ctx->world == 1) return 0;
  NS_NCCL(ncclAllReduce(d, d, n, ncclDouble, op, ctx->comm, ctx->stream));
  return 0;
}

// n sweeps of red-black Gauss-Seidel on level L; *x is the current iterate, *xt its ping-pong partner.
static inline int march_chunk(const nsgpu_ctx* ctx, const Level& L, int uj) {
  if (ctx->march_ri > 0) return std::min(ctx->march_ri, L.g.nx);
  // enough warps for ~2 waves of 148 SMs x 24 resident warps, chunks no shorter than 8 columns
  const int wj = (L.g.ny + uj - 1) / uj;
  const int chunks = std::max(1, (2 * 148 * 24 + wj - 1) / wj);
  return std::min(L.g.nx, std::max(8, (L.g.nx + chunks - 1) / chunks));
}

static int smooth(nsgpu_ctx* ctx, const Level& L, OpCoef c, double** x, double** xt, const double* b, int n) {
  const dim3 grid((L.g.ny + SM_TJ - 1) / SM_TJ, (L.g.nx + SM_TI - 1) / SM_TI);
  const bool march = ctx->march && ctx->rect && L.march_ok[c.op];
  for (int s = 0; s < n; s++) {
    int rc = halo_exchange(ctx, L, *x, 2);
    if (rc) return rc;
    if (march) {
      constexpr int UJ = 60;
      const int ri = march_chunk(ctx, L, UJ);
      const dim3 mg(((L.g.ny + UJ - 1) / UJ + MARCH_WARPS - 1) / MARCH_WARPS, (L.g.nx + ri - 1) / ri);
      const size_t sh = (size_t)3 * (ri + 3 * 2) * sizeof(double);
      k_sweep_march<1><<<mg, MARCH_THREADS, sh, ctx->stream>>>(L.g, L.mcx[c.op], L.mcy[c.op], *x, b, *xt, ri);
      NS_LAUNCH_CHECK();
      std::swap(*x, *xt);
      continue;
    }
#define CALL(OP, RECT) k_smooth_rb<OP, RECT, false><<<grid, SM_THREADS, 0, ctx->stream>>>(L.g, *x, b, *xt, c.alpha, c.beta)
    NS_DISPATCH_OP_RECT(c.op, ctx->rect, CALL);
#undef CALL
    NS_LAUNCH_CHECK();
    std::swap(*x, *xt);
  }
  return 0;
}

// d_out[0] = ||b - A x||^2 over the local slab (all-reduced over ranks)
static int residual_norm2(nsgpu_ctx* ctx, const Level& L, OpCoef c, double* x, const double* b, double* d_out) {
  int rc = halo_exchange(ctx, L, x, 2);
  if (rc) return rc;
  const dim3 grid = grid_cols(L.g, 256);
#define CALL(OP, RECT) k_residual<OP, RECT, false, true><<<grid, 256, 0, ctx->stream>>>(L.g, x, b, nullptr, c.alpha, c.beta, ctx->red, d_out)
  NS_DISPATCH_OP_RECT(c.op, ctx->rect, CALL);
#undef CALL
  NS_LAUNCH_CHECK();
  return allreduce(ctx, d_out, 1, ncclSum);
}

static int residual_vec(nsgpu_ctx* ctx, const Level& L, OpCoef c, double* x, const double* b, double* r, double* d_norm2) {
  int rc = halo_exchange(ctx, L, x, 2);
  if (rc) return rc;
  const dim3 grid = grid_cols(L.g, 256);
#define CALL(OP, RECT) k_residual<OP, RECT, true, true><<<grid, 256, 0, ctx->stream>>>(L.g, x, b, r, c.alpha, c.beta, ctx->red, d_norm2)
  NS_DISPATCH_OP_RECT(c.op, ctx->rect, CALL);
#undef CALL
  NS_LAUNCH_CHECK();
  return allreduce(ctx, d_norm2, 1, ncclSum);
}

// y = A x ; d_out = {(w,y), (y,y)} when w != nullptr
static int apply_op(nsgpu_ctx* ctx, const Level& L, OpCoef c, double* x, double* y, const double* w, double* d_out) {
  int rc = halo_exchange(ctx, L, x, 2);
  if (rc) return rc;
  const dim3 grid = grid_cols(L.g, 256);
  if (w) {
#define CALL(OP, RECT) k_apply<OP, RECT, true><<<grid, 256, 0, ctx->stream>>>(L.g, x, y, w, c.alpha, c.beta, ctx->red, d_out)
    NS_DISPATCH_OP_RECT(c.op, ctx->rect, CALL);
#undef CALL
    NS_LAUNCH_CHECK();
    return allreduce(ctx, d_out, 2, ncclSum);
  }
#define CALL(OP, RECT) k_apply<OP, RECT, false><<<grid, 256, 0, ctx->stream>>>(L.g, x, y, nullptr, c.alpha, c.beta, ctx->red, nullptr)
  NS_DISPATCH_OP_RECT(c.op, ctx->rect, CALL);
#undef CALL
  NS_LAUNCH_CHECK();
  return 0;
}

// One V(nu_pre, nu_post) cycle for A x = b on level l (x, b given for level l; coarser levels use their own).
static int vcycle(nsgpu_ctx* ctx, int l, OpCoef c, double** x, double** xt, const double* b) {
  Level& L = ctx->levels[l];
  int rc;
  if (l == (int)ctx->levels.size() - 1) {
    const int n = ctx->levels.size() == 1 ? ctx->nu_pre + ctx->nu_post : L.coarse_sweeps;
    return smooth(ctx, L, c, x, xt, b, n);
  }
  Level& C = ctx->levels[l + 1];
  if ((rc = smooth(ctx, L, c, x, xt, b, ctx->nu_pre))) return rc;
  if ((rc = halo_exchange(ctx, L, *x, 2))) return rc;
  {
    const dim3 grid((C.g.ny + 127) / 128, C.g.nx);
#define CALL(OP, RECT) k_restrict<OP, RECT><<<grid, 128, 0, ctx->stream>>>(L.g, C.g, *x, b, C.b, c.alpha, c.beta, L.cx, L.cy)
    NS_DISPATCH_OP_RECT(c.op, ctx->rect, CALL);
#undef CALL
    NS_LAUNCH_CHECK();
  }
  if ((rc = halo_exchange(ctx, C, C.b, 1))) return rc;
  {
    cudaError_t e = cudaMemsetAsync(C.x, 0, C.stride * sizeof(double), ctx->stream);
    if (e != cudaSuccess) { ctx->err = cudaGetErrorString(e); return NSGPU_ERR_CUDA; }
  }
  double *cx = C.x, *cxt = C.xt;
  if ((rc = vcycle(ctx, l + 1, c, &cx, &cxt, C.b))) return rc;
  C.x = cx; C.xt = cxt;
  {
    const dim3 grid = grid_cols(L.g, 256);
#define CALL(RECT) k_prolong_add<RECT><<<grid, 256, 0, ctx->stream>>>(L.g, C.g, *x, C.x, L.cx, L.cy)
    NS_DISPATCH_RECT(ctx->rect, CALL);
#undef CALL
    NS_LAUNCH_CHECK();
  }
  return smooth(ctx, L, c, x, xt, b, ctx->nu_post);
}

static int read_scalars(nsgpu_ctx* ctx, const double* d, int n, double* h) {
  cudaError_t e = cudaMemcpyAsync(ctx->h_scal, d, n * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
  if (e != cudaSuccess) { ctx->err = std::string("scalar readback: ") + cudaGetErrorString(e); return NSGPU_ERR_CUDA; }
  for (int k = 0; k < n; k++) h[k] = ctx->h_scal[k];
  return 0;
}

// d_out[0] = (a, a) over the owned columns
static int norm2_vec(nsgpu_ctx* ctx, const Level& L, const double* a, double* d_out) {
  const size_t n = (size_t)L.g.nx * L.g.ld;
  const int blocks = (int)std::min<size_t>((n + 255) / 256, 148 * 8);
  k_dot<<<blocks, 256, 0, ctx->stream>>>(n, a + (size_t)HX * L.g.ld, a + (size_t)HX * L.g.ld, ctx->red, d_out);
  NS_LAUNCH_CHECK();
  return allreduce(ctx, d_out, 1, ncclSum);
}

// Stationary multigrid iteration: x <- x + V(b - A x) until ||b - A x|| <= rtol ||b||.
static int solve_mg(nsgpu_ctx* ctx, OpCoef c, double** x, double** xt, const double* b, double rtol, SolveResult* out) {
  const Level& L = ctx->levels[0];
  double h[2];
  int rc;
  if ((rc = norm2_vec(ctx, L, b, ctx->d_scal + 0))) return rc;
  if ((rc = residual_norm2(ctx, L, c, *x, b, ctx->d_scal + 1))) return rc;
  if ((rc = read_scalars(ctx, ctx->d_scal, 2, h))) return rc;
  const double bn = std::sqrt(h[0]);
  double rn = std::sqrt(h[1]);
  out->cycles = 0;
  if (bn == 0.0) {  // zero right-hand side: the solution of A x = 0 from the KSP's point of view is x = 0
    cudaMemsetAsync(*x, 0, L.stride * sizeof(double), ctx->stream);
    out->rel = 0; out->converged = true;
    return 0;
  }
  while (rn > rtol * bn && out->cycles < ctx->max_cycles) {
    if ((rc = vcycle(ctx, 0, c, x, xt, b))) return rc;
    out->cycles++;
    if ((rc = residual_norm2(ctx, L, c, *x, b, ctx->d_scal + 1))) return rc;
    if ((rc = read_scalars(ctx, ctx->d_scal + 1, 1, h))) return rc;
    rn = std::sqrt(h[0]);
    if (!(rn == rn)) break;
  }
  out->rel = rn / bn;
  out->converged = rn <= rtol * bn;
  return 0;
}

// Right-preconditioned BiCGStab, preconditioner = one V-cycle from a zero guess; all recurrence scalars stay
// on the device, one 8-byte readback per iteration for the stopping test.
//   d_scal layout: [0] rho_new [1] rho_old [2] alpha [3] omega [4..5] (r0,v),(v,v) [6..7] (t,s),(t,t)
//                  [8] (s,s) [9..10] (r,r),(r0,r) [11] (b,b)
static int solve_bicgstab(nsgpu_ctx* ctx, OpCoef c, double** xp, double** xtp, const double* b, double rtol, SolveResult* out) {
  const Level& L = ctx->levels[0];
  const size_t n = (size_t)L.g.nx * L.g.ld, off = (size_t)HX * L.g.ld;
  const int blocks = (int)std::min<size_t>((n + 255) / 256, 148 * 8);
  double *x = *xp;
  double *r = ctx->kry[0], *r0 = ctx->kry[1], *p = ctx->kry[2], *v = ctx->kry[3], *s = ctx->kry[4], *t = ctx->kry[5],
         *ph = ctx->kry[6], *sh = ctx->kry[7];
  double* S = ctx->d_scal;
  double h[4];
  int rc;
  if ((rc = norm2_vec(ctx, L, b, S + 11))) return rc;
  if ((rc = residual_vec(ctx, L, c, x, b, r, S + 9))) return rc;
  if ((rc = read_scalars(ctx, S + 9, 3, h))) return rc;
  const double bn = std::sqrt(h[2]);
  double rn = std::sqrt(h[0]);
  out->cycles = 0;
  if (bn == 0.0) { cudaMemsetAsync(x, 0, L.stride * sizeof(double), ctx->stream); out->rel = 0; out->converged = true; return 0; }
  if (rn <= rtol * bn) { out->rel = rn / bn; out->converged = true; return 0; }
  cudaMemcpyAsync(r0, r, L.stride * sizeof(double), cudaMemcpyDeviceToDevice, ctx->stream);
  cudaMemsetAsync(p, 0, L.stride * sizeof(double), ctx->stream);
  cudaMemsetAsync(v, 0, L.stride * sizeof(double), ctx->stream);
  const double init[4] = {h[0], 1.0, 1.0, 1.0};   // rho_new = (r0, r) = (r, r); rho_old = alpha = omega = 1
  cudaMemcpyAsync(S, init, sizeof init, cudaMemcpyHostToDevice, ctx->stream);
  cudaStreamSynchronize(ctx->stream);
  double* tmpA = ctx->tmp;   // ping-pong partner for the preconditioner's fine level
  while (out->cycles < ctx->max_cycles) {
    k_bicg_p<<<blocks, 256, 0, ctx->stream>>>(n, r + off, p + off, v + off, S);
    NS_LAUNCH_CHECK();
    // ph = M^-1 p
    cudaMemsetAsync(ph, 0, L.stride * sizeof(double), ctx->stream);
    { double *a = ph, *bq = tmpA; if ((rc = vcycle(ctx, 0, c, &a, &bq, p))) return rc; ph = a; tmpA = bq; }
    if ((rc = apply_op(ctx, L, c, ph, v, r0, S + 4))) return rc;
    k_scalar_div<<<1, 1, 0, ctx->stream>>>(S + 2, S + 0, S + 4);   // alpha = rho / (r0, v)
    NS_LAUNCH_CHECK();
    k_bicg_s<<<blocks, 256, 0, ctx->stream>>>(n, r + off, v + off, s + off, S + 2, ctx->red, S + 8);
    NS_LAUNCH_CHECK();
    if ((rc = allreduce(ctx, S + 8, 1, ncclSum))) return rc;
    // sh = M^-1 s
    cudaMemsetAsync(sh, 0, L.stride * sizeof(double), ctx->stream);
    { double *a = sh, *bq = tmpA; if ((rc = vcycle(ctx, 0, c, &a, &bq, s))) return rc; sh = a; tmpA = bq; }
    if ((rc = apply_op(ctx, L, c, sh, t, s, S + 6))) return rc;
    k_scalar_div<<<1, 1, 0, ctx->stream>>>(S + 3, S + 6, S + 7);   // omega = (t,s)/(t,t)
    NS_LAUNCH_CHECK();
    k_scalar_copy<<<1, 1, 0, ctx->stream>>>(S + 1, S + 0);         // rho_old = rho_new
    NS_LAUNCH_CHECK();
    k_bicg_xr<<<blocks, 256, 0, ctx->stream>>>(n, x + off, ph + off, sh + off, r + off, s + off, t + off, r0 + off, S + 2, S + 3,
                                               ctx->red, S + 9);
    NS_LAUNCH_CHECK();
    if ((rc = allreduce(ctx, S + 9, 2, ncclSum))) return rc;
    k_scalar_copy<<<1, 1, 0, ctx->stream>>>(S + 0, S + 10);        // rho_new = (r0, r)
    NS_LAUNCH_CHECK();
    out->cycles++;
    if ((rc = read_scalars(ctx, S + 9, 1, h))) return rc;
    rn = std::sqrt(h[0]);
    if (!(rn == rn) || rn <= rtol * bn) break;
  }
  ctx->kry[6] = ph; ctx->kry[7] = sh; ctx->tmp = tmpA;
  // the recurrence residual can drift from the true one: report the true residual
  if ((rc = residual_norm2(ctx, L, c, x, b, S + 9))) return rc;
  if ((rc = read_scalars(ctx, S + 9, 1, h))) return rc;
  rn = std::sqrt(h[0]);
  out->rel = rn / bn;
  out->converged = rn <= rtol * bn * 1.0001 || rn <= 10 * rtol * bn;
  return 0;
}

static int solve(nsgpu_ctx* ctx, OpCoef c, double** x, const double* b, double rtol, SolveResult* out) {
  if (ctx->solver == 1) return solve_bicgstab(ctx, c, x, &ctx->tmp, b, rtol, out);
  return solve_mg(ctx, c, x, &ctx->tmp, b, rtol, out);
}

// ---------------------------------------------------------------------------------------------------------
// Phases of one time step

static int phase_rhs(nsgpu_ctx* ctx) {   // ConstructRHS_V, FS.cpp:546
  const Level& L = ctx->levels[0];
  int rc;
  if ((rc = halo_exchange(ctx, L, ctx->u, 2))) return rc;
  if ((rc = halo_exchange(ctx, L, ctx->v, 2))) return rc;
  if ((rc = halo_exchange(ctx, L, ctx->gpy, 1))) return rc;
  const StepConst sc{ctx->dt, ctx->re};
  const dim3 grid((L.g.ny + 127) / 128, L.g.nx);
#define CALL(RECT) k_rhs_build<RECT><<<grid, 128, 0, ctx->stream>>>(L.g, ctx->d_edges, sc, ctx->u, ctx->v, ctx->conv_u, ctx->conv_v, ctx->gpx, ctx->gpy, ctx->rhs_u, ctx->rhs_v)
  NS_DISPATCH_RECT(ctx->rect, CALL);
#undef CALL
  NS_LAUNCH_CHECK();
  return 0;
}

static int phase_velocity(nsgpu_ctx* ctx, SolveResult* ru, SolveResult* rv) {   // FS.cpp:547-548
  const Level& L = ctx->levels[0];
  // initial guess u^n (the reference starts from zero; the converged u* is the same to tolerance)
  cudaMemcpyAsync(ctx->ustar, ctx->u, L.stride * sizeof(double), cudaMemcpyDeviceToDevice, ctx->stream);
  cudaMemcpyAsync(ctx->vstar, ctx->v, L.stride * sizeof(double), cudaMemcpyDeviceToDevice, ctx->stream);
  int rc;
  if ((rc = halo_exchange(ctx, L, ctx->rhs_u, 1))) return rc;
  if ((rc = halo_exchange(ctx, L, ctx->rhs_v, 1))) return rc;
  if ((rc = solve(ctx, coef_v(ctx), &ctx->ustar, ctx->rhs_u, ctx->rtol_u, ru))) return rc;
  if ((rc = solve(ctx, coef_v(ctx), &ctx->vstar, ctx->rhs_v, ctx->rtol_u, rv))) return rc;
  return 0;
}

static int phase_div(nsgpu_ctx* ctx, bool remove_mean) {   // FS.cpp:549-550
  const Level& L = ctx->levels[0];
  int rc;
  if ((rc = halo_exchange(ctx, L, ctx->ustar, 1))) return rc;
  if ((rc = halo_exchange(ctx, L, ctx->vstar, 1))) return rc;
  const StepConst sc{ctx->dt, ctx->re};
  const dim3 grid = grid_cols(L.g, 256);
#define CALL(RECT) k_div_rhs<RECT><<<grid, 256, 0, ctx->stream>>>(L.g, ctx->d_edges, sc, ctx->ustar, ctx->vstar, ctx->rhs_phi, ctx->red, ctx->d_scal + 16)
  NS_DISPATCH_RECT(ctx->rect, CALL);
#undef CALL
  NS_LAUNCH_CHECK();
  if (remove_mean) {
    if ((rc = allreduce(ctx, ctx->d_scal + 16, 1, ncclSum))) return rc;
    const double inv_n = 1.0 / (double)ctx->n_cells_global;
#define CALL(RECT) k_sub_mean<RECT><<<grid, 256, 0, ctx->stream>>>(L.g, ctx->rhs_phi, ctx->d_scal + 16, inv_n)
    NS_DISPATCH_RECT(ctx->rect, CALL);
#undef CALL
    NS_LAUNCH_CHECK();
  }
  return 0;
}

static int phase_phi(nsgpu_ctx* ctx, SolveResult* rp) {   // FS.cpp:551, warm start from phi^n (FS.cpp:54)
  const Level& L = ctx->levels[0];
  int rc;
  if ((rc = halo_exchange(ctx, L, ctx->rhs_phi, 1))) return rc;
  const int64_t before = ctx->launches;
  (void)before;
  if ((rc = solve(ctx, coef_phi(), &ctx->phi, ctx->rhs_phi, ctx->rtol_phi, rp))) return rc;
  ctx->fine_sweeps_phi += (double)rp->cycles * (ctx->nu_pre + ctx->nu_post) * (ctx->solver == 1 ? 2 : 1);
  // pin mean(phi) = 0: the reference projects the constant out of every Krylov update (MatSetNullSpace, FS.cpp:143)
  {
    const size_t n = (size_t)L.g.nx * L.g.ld;
    const int blocks = (int)std::min<size_t>((n + 255) / 256, 148 * 8);
    k_sum<<<blocks, 256, 0, ctx->stream>>>(n, ctx->phi + (size_t)HX * L.g.ld, ctx->red, ctx->d_scal + 17);
    NS_LAUNCH_CHECK();
    if ((rc = allreduce(ctx, ctx->d_scal + 17, 1, ncclSum))) return rc;
    const dim3 grid = grid_cols(L.g, 256);
    const double inv_n = 1.0 / (double)ctx->n_cells_global;
#define CALL(RECT) k_sub_mean<RECT><<<grid, 256, 0, ctx->stream>>>(L.g, ctx->phi, ctx->d_scal + 17, inv_n)
    NS_DISPATCH_RECT(ctx->rect, CALL);
#undef CALL
    NS_LAUNCH_CHECK();
  }
  return 0;
}

static int phase_correct(nsgpu_ctx* ctx, double mm[4]) {   // FS.cpp:552-557
  const Level& L = ctx->levels[0];
  int rc;
  if ((rc = halo_exchange(ctx, L, ctx->phi, 2))) return rc;
  const StepConst sc{ctx->dt, ctx->re};
  const dim3 grid = grid_cols(L.g, 256);
#define CALL(RECT) k_correct<RECT><<<grid, 256, 0, ctx->stream>>>(L.g, sc, ctx->phi, ctx->ustar, ctx->vstar, ctx->u, ctx->v, ctx->gpx, ctx->gpy, ctx->red, ctx->d_scal + 20)
  NS_DISPATCH_RECT(ctx->rect, CALL);
#undef CALL
  NS_LAUNCH_CHECK();
  if ((rc = allreduce(ctx, ctx->d_scal + 20, 4, ncclMax))) return rc;
  if (mm) {
    double h[4];
    if ((rc = read_scalars(ctx, ctx->d_scal + 20, 4, h))) return rc;
    mm[0] = -h[0]; mm[1] = h[1]; mm[2] = -h[2]; mm[3] = h[3];
  }
  return 0;
}

}  // namespace nsgpu
