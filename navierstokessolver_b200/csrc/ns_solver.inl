// ns_solver.inl -- host orchestration of the kernels: halo exchange, multigrid V-cycle, the two linear solves
// that replace KSPSolve (reference FS.cpp:547-551) and the phases of one time step (FS.cpp:546-557).

namespace nsgpu {

struct OpCoef { int op; double alpha, beta; };
static inline OpCoef coef_v(const nsgpu_ctx* c) { return OpCoef{0, 1.0, c->dt / (2.0 * c->re)}; }   // LHS_V   FS.cpp:140-141
static inline OpCoef coef_phi() { return OpCoef{1, 0.0, -1.0}; }                                    // LHS_phi
static inline OpCoef coef_phi_walls() { return OpCoef{2, 0.0, -1.0}; }   // L_phi with wall ghosts on every side (preconditioner only)

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
    else if ((op) == 1) { if (rect) { CALL(1, true); } else { CALL(1, false); } } \
    else { if (rect) { CALL(2, true); } else { CALL(2, false); } }           \
  } while (0)
#define NS_DISPATCH_RECT(rect, CALL) \
  do { if (rect) { CALL(true); } else { CALL(false); } } while (0)

static inline dim3 grid_cols(const LevelGeom& g, int threads, int max_blocks = 148 * 16) {
  unsigned gx = (unsigned)((g.ny + threads - 1) / threads);
  unsigned gy = (unsigned)std::max(1, std::min(g.nx, max_blocks / (int)gx));
  return dim3(gx, gy, 1);
}

// Halo exchange of up to HXP_MAX fields in ONE kernel, through peer memory (k_halo_push): no NCCL call, capturable in a graph.
// Safety of the one-sided stores.  A neighbour may be ahead of this rank: its stores of exchange k can land here while this rank is still
// between its own exchanges k - 1 and k (not earlier: the neighbour finishes exchange k - 1 only after this rank has signalled it).  So no
// kernel that runs between two consecutive exchanges may read the halo columns of a field of the LATER exchange.  The sequence of one
// time step on this route is
//   X(u, v, divPhi_y) K_rhs  X(rhs_u, rhs_v) X(u) solve_u X(v) solve_v  X(u*, v*) K_div  X(rhs_phi) X(phi) solve_phi  X(phi) K_correct
// and every kernel between two exchanges reads only halos exchanged BEFORE it and none of the next exchange's fields' halos (the solves'
// own passes exchange through the separate march protocol; the X(phi) behind solve_phi rewrites values its last pass has pushed already).
static int halo_exchange_push(nsgpu_ctx* ctx, const Level& L, double* const* f, const int* width, int nf) {
  HaloPushArgs a{};
  a.nf = 0; a.ld = L.g.ld; a.nx = L.g.nx;
  size_t nmax = 0;
  for (int k = 0; k < nf; k++) {
    const int w = std::min(width[k], std::min(HX, L.g.nx));
    if (w <= 0) continue;
    a.f[a.nf] = f[k]; a.width[a.nf] = w; a.nf++;
    nmax = std::max(nmax, (size_t)w * L.g.ld / 2);
    ctx->halo_w[f[k]] = w;
  }
  if (!a.nf) return 0;
  a.offL = ctx->peerL ? (long long)((ctx->peerL - ctx->arena) / (ptrdiff_t)sizeof(double)) : 0;
  a.offR = ctx->peerR ? (long long)((ctx->peerR - ctx->arena) / (ptrdiff_t)sizeof(double)) : 0;
  a.hdr = (unsigned long long*)ctx->arena;
  a.err = ctx->d_errflag;
  const dim3 grid((unsigned)std::max<size_t>(1, std::min<size_t>(16, (nmax + 255) / 256)), (unsigned)a.nf);
  k_halo_push<<<grid, 256, 0, ctx->stream>>>(a);
  NS_LAUNCH_CHECK();
  return 0;
}
static inline bool hx_push_ok(const nsgpu_ctx* ctx, const Level& L, const double* f) {
  return ctx->hx_push && &L == &ctx->levels[0] && !L.replicated && ctx->arena && (const char*)f >= ctx->arena && (const char*)f < ctx->arena + ctx->arena_bytes;
}
// several fields at once (one kernel on the peer-store route, one NCCL group per field otherwise)
static int halo_exchange(nsgpu_ctx* ctx, const Level& L, double* f, int width, bool march_consumer = false);
static int halo_exchange_n(nsgpu_ctx* ctx, const Level& L, std::initializer_list<std::pair<double*, int>> fw) {
  if (ctx->world == 1 || L.replicated) return 0;
  bool all = fw.size() <= (size_t)HXP_MAX;
  for (auto& e : fw) all = all && hx_push_ok(ctx, L, e.first);
  if (all) {
    double* f[HXP_MAX]; int w[HXP_MAX]; int n = 0;
    for (auto& e : fw) { f[n] = e.first; w[n] = e.second; n++; }
    return halo_exchange_push(ctx, L, f, w, n);
  }
  for (auto& e : fw) { int rc = halo_exchange(ctx, L, e.first, e.second, false); if (rc) return rc; }
  return 0;
}

// One-column-block halo exchange with the x-neighbours: `width` columns, contiguous (width * ld doubles).
// march_consumer: the reader is a marching pass, which waits for the neighbours' pushes itself -- if the buffer's halo
// was filled by a pushing pass (ctx->halo_w) there is nothing to send.
static int halo_exchange(nsgpu_ctx* ctx, const Level& L, double* f, int width, bool march_consumer) {
  if (ctx->world == 1 || L.replicated) return 0;
  width = std::min(width, std::min(HX, L.g.nx));
  if (march_consumer) {
    auto it = ctx->halo_w.find(f);
    if (it != ctx->halo_w.end() && it->second >= width) return 0;
  }
  if (hx_push_ok(ctx, L, f)) return halo_exchange_push(ctx, L, &f, &width, 1);
  if (ctx->capturing) ctx->capture_nccl = true;   // solve graphs hold this library's kernels only
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

// tracing (NSGPU_TRACE=1): key = level * 1000 + nsw * 100 + res * 10 + prolong * 2 + push; 99999 = coarse tail
static inline void trace_begin(nsgpu_ctx* ctx, int key) {
  if (!ctx->trace || ctx->capturing) return;
  nsgpu_ctx::TraceEv t; t.key = key;
  cudaEventCreate(&t.a); cudaEventCreate(&t.b);
  cudaEventRecord(t.a, ctx->stream);
  ctx->trace_ev.push_back(t);
}
static inline void trace_end(nsgpu_ctx* ctx) {
  if (!ctx->trace || ctx->capturing || ctx->trace_ev.empty()) return;
  cudaEventRecord(ctx->trace_ev.back().b, ctx->stream);
}
static void trace_collect(nsgpu_ctx* ctx) {   // after a stream synchronize
  for (auto& t : ctx->trace_ev) {
    float ms = 0;
    if (cudaEventElapsedTime(&ms, t.a, t.b) == cudaSuccess) { auto& e = ctx->trace_acc[t.key]; e.first += ms; e.second += 1; }
    cudaEventDestroy(t.a); cudaEventDestroy(t.b);
  }
  ctx->trace_ev.clear();
  cudaGetLastError();
}

static int allreduce(nsgpu_ctx* ctx, double* d, int n, ncclRedOp_t op) {
  if (ctx->world == 1) return 0;
  if (ctx->capturing) ctx->capture_nccl = true;
  NS_NCCL(ncclAllReduce(d, d, n, ncclDouble, op, ctx->comm, ctx->stream));
  return 0;
}

// ---------------------------------------------------------------------------------------------------------
// Fused marching passes (rectangle fast path, ns_march.cuh)

// Folded tables of the marching kernels: the tables of op 1 never contain the outflow rows (they are not separable, march_ok[1] is
// false on outflow domains), i.e. they ARE the wall-only L_phi -- op 2, the preconditioner of the outflow solve, marches on them.
static inline int op_tab(OpCoef c) { return c.op == 2 ? 1 : c.op; }
static inline bool march_tab_ok(const Level& L, OpCoef c) { return c.op == 2 ? L.mcx[1] != nullptr : L.march_ok[c.op]; }

static inline bool march_level_ok(const nsgpu_ctx* ctx, const Level& L, OpCoef c) {
  // with several ranks the 5-column halo of a fused pass has to come from ONE neighbour
  return ctx->march && ctx->rect && march_tab_ok(L, c) && (ctx->world == 1 || L.replicated || L.g.nx >= HX);
}

typedef void (*MarchKernel)(LevelGeom, const double*, const double*, const double*, const double*, double*, int, MarchXfer);

// CTAs of `kern` that fit on the GPU at once (one wave), cached per kernel
static int march_wave(MarchKernel kern, size_t smem = 4096) {
  static thread_local std::vector<std::pair<MarchKernel, int>> cache;   // (per host thread: one thread drives one GPU)
  for (auto& e : cache) if (e.first == kern) return e.second;
  int per_sm = 0, dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  if (smem > 48 * 1024) cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);   // ring kernels (ns_march_ring.cuh)
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, MARCH_THREADS, smem) != cudaSuccess || per_sm < 1) per_sm = 2;
  cache.push_back({kern, per_sm * sms});
  return per_sm * sms;
}

// Chunk length (columns per CTA).  Every warp marches `ri + ovh` columns and the CTAs run in ceil(ctas / wave) waves
// of equal work, so the estimated time is waves * (ri + ovh): exact multiples of a wave win on the big levels, the
// shortest chunks (most warps) on the small, latency-bound ones.
static inline int march_chunk(const nsgpu_ctx* ctx, const Level& L, int uj, int ovh, int wave, int rb = 0) {
  // rb > 0: two short boundary chunks of rb columns each come on top of the chunks of the interior
  const int gx = ((L.g.ny + uj - 1) / uj + MARCH_WARPS - 1) / MARCH_WARPS;
  const int nxi = L.g.nx - 2 * rb, extra = rb ? 2 : 0;
  int ri = 0;
  if (ctx->march_ri > 0) ri = std::min(ctx->march_ri + (ctx->march_ri & 1), nxi + (nxi & 1));
  else {
    long long best = -1;
    const int hi = std::min(nxi + (nxi & 1), 1024);
    for (int r = 4; r <= hi; r += 2) {
      const long long ctas = (long long)gx * (extra + (nxi + r - 1) / r);
      const long long cost = ((ctas + wave - 1) / wave) * (r + ovh);
      if (best < 0 || cost < best) { best = cost; ri = r; }
    }
    if (ri == 0) ri = 4;
  }
  while ((long long)gx * (extra + (nxi + ri - 1) / ri) > RED_MAX_BLOCKS) ri *= 2;
  return ri;
}

// One pass: xout <- NSW red-black sweeps of (xin [+ P xc]) on level L; optional fused restriction / residual norm.
//   xin == nullptr: zero initial guess;  Cp/xc: prolong from;  Cr: restrict the residual to Cr->b;  norm_out/ctl: |r|^2
static int march_pass(nsgpu_ctx* ctx, const Level& L, OpCoef c, int nsw, double* xin, const double* b, double* xout,
                      const Level* Cp, double* xc, const Level* Cr, double* norm_out, SolveCtl* ctl, double omega = 0.0) {
  const int res = Cr ? 1 : (norm_out ? 2 : 0);
  const int lead = march_lead(nsw, res);
  int rc;
  if (xin && (rc = halo_exchange(ctx, L, xin, lead, true))) return rc;
  if (Cp && (rc = halo_exchange(ctx, *Cp, xc, (lead + 1) / 2 + 1, true))) return rc;
  MarchXfer xf{};
  // a replicated coarse level holds the whole domain: this rank's part starts at its global coarse column
  const size_t rep_off_p = (Cp && Cp->replicated && !L.replicated) ? (size_t)(L.g.gi0 / L.cx) * Cp->g.ld : 0;
  const size_t rep_off_r = (Cr && Cr->replicated && !L.replicated) ? (size_t)(L.g.gi0 / L.cx) * Cr->g.ld : 0;
  if (Cp) { xf.xc = xc + rep_off_p; xf.ldc = Cp->g.ld; xf.pcx = L.cx; xf.pcy = L.cy; }
  if (Cr) { xf.bc = Cr->b + rep_off_r; xf.ldbc = Cr->g.ld; xf.rcx = L.cx; xf.rcy = L.cy; xf.rscale = 1.0 / (double)(L.cx * L.cy); }
  if (res == 2) { xf.red = ctx->red; xf.out = norm_out; xf.ctl = ctl; }
  xf.omega = omega;
  if (ctx->next_pass_shift && xin && res == 1) { xf.shift = ctx->next_pass_shift; xf.shift_scale = 1.0 / (double)ctx->n_cells_global; }
  else if (ctx->next_pass_shift) { ctx->err = "internal: phi's pending mean has to be removed by a restricting pass"; return NSGPU_ERR_ARG; }
  ctx->next_pass_shift = nullptr;
  const int uj = march_uj(nsw, res);
  // several ranks: push the boundary columns into the neighbours' halos from inside the pass (peer memory over NVLink)
  static const bool nopush = getenv("NSGPU_NOPUSH") != nullptr;   // debugging: arena mapped, halo columns through NCCL
  // (a buffer outside the arena has no image on the neighbour: `xout + offset` would be a wild peer pointer -- NCCL then)
  auto in_arena = [&](const double* q) { return (const char*)q >= ctx->arena && (const char*)q < ctx->arena + ctx->arena_bytes; };
  const bool selfpush = ctx->selfpush && L.g.nx >= 2 * HX;
  const bool push = selfpush || (ctx->world > 1 && ctx->use_p2p && ctx->arena && !L.replicated && (ctx->peerL || ctx->peerR) && !nopush &&
                                 in_arena(xout) && (!Cr || in_arena(Cr->b)));
  const bool push_b = push && Cr && (selfpush ? Cr->g.nx >= 2 * HX : !Cr->replicated);
  MarchKernel kern = nullptr;
#define NS_MARCH(NSW, PRO, RES, PF) do { if (push) kern = k_march<NSW, PRO, RES, PF, false, true>; else kern = k_march<NSW, PRO, RES, PF, false, false>; } while (0)
#define NS_MARCH_RES(NSW, PRO) do { if (res == 0) NS_MARCH(NSW, PRO, 0, MARCH_PF); else if (res == 1) NS_MARCH(NSW, PRO, 1, MARCH_PF); else NS_MARCH(NSW, PRO, 2, MARCH_PF); } while (0)
  if (omega > 0.0) {                   // red-black SOR double sweep (velocity solves)
    if (nsw != 2 || Cp || res == 1) { ctx->err = "SOR passes are double sweeps without transfers"; return NSGPU_ERR_ARG; }
    if (res == 0) kern = push ? k_march<2, false, 0, MARCH_PF, true, true> : k_march<2, false, 0, MARCH_PF, true, false>;
    else kern = push ? k_march<2, false, 2, MARCH_PF, true, true> : k_march<2, false, 2, MARCH_PF, true, false>;
  } else if (nsw == 1 && !Cp && res == 0) {   // the plain sweep has its own prefetch depth
    if (push) kern = k_march<1, false, 0, 6, false, true>;
    else switch (ctx->march_pf) {
      case 2: kern = k_march<1, false, 0, 2, false, false>; break;
      case 3: kern = k_march<1, false, 0, 3, false, false>; break;
      case 4: kern = k_march<1, false, 0, 4, false, false>; break;
      default: kern = k_march<1, false, 0, 6, false, false>; break;
    }
  } else if (nsw == 1) { if (Cp) NS_MARCH_RES(1, true); else NS_MARCH_RES(1, false); }
  else { if (Cp) NS_MARCH_RES(2, true); else NS_MARCH_RES(2, false); }
#undef NS_MARCH_RES
#undef NS_MARCH
  // one rank, no transfer from below: the columns of x and b through the shared-memory ring (bulk copies) instead of the register ring
  bool ring = false;
  if (ctx->march_ring && !push && nsw == 2 && !Cp && L.g.nx >= 64) {
    ring = true;
    if (omega > 0.0 && res == 2) kern = k_march_ring<2, false, 2, true>;
    else if (omega > 0.0) ring = false;                                   // the plain SOR double sweep is at its roofline already
    else if (res == 1) kern = k_march_ring<2, false, 1, false>;
    else if (res == 2) kern = k_march_ring<2, false, 2, false>;
    else ring = false;
  }
  // Several ranks: the first and the last chunk of the slab run the pushing loop (march_run<..., true>), which costs about 1.5 x the
  // plain loop per column; every other chunk runs the plain one.  The two boundary chunks are therefore made SHORTER, so that all CTAs
  // of the launch take the same time: rb columns with 1.5 (rb + ovh) = ri + ovh, the interior shared evenly between the other rows.
  const int ovh = 2 * lead + 2 * nsw + 2, wave = march_wave(kern, ring ? march_ring_smem(256, nsw, res) : 4096);
  int rb = 0, ri = march_chunk(ctx, L, uj, ovh, wave, 0);
  if (push && ctx->march_rb != 0 && !(L.g.nx & 1) && L.g.nx >= ctx->march_rb_min_nx && !(Cr && Cr->replicated && !selfpush)) {
    const int rows = (L.g.nx + ri - 1) / ri, rb_min = 2 * (HX - 1);   // >= the push zone of a restricting pass
    if (ctx->march_rb > 0) { rb = std::max(ctx->march_rb, rb_min); rb += rb & 1; }
    else if (rows >= 3) {
      const double pc = (Cp && res == 2) ? ctx->rb_cost_up : ctx->rb_cost;   // cost of a column in the pushing loop / in the plain loop
      double best = 1e300;
      for (int r = rb_min; 2 * r + 4 <= L.g.nx; r += 2) {
        int rin = (L.g.nx - 2 * r + rows - 3) / (rows - 2);
        rin += rin & 1;
        const double cost = std::max(pc * (r + ovh), (double)(rin + ovh));
        if (cost < best) { best = cost; rb = r; }
        if (pc * (r + ovh) > rin + ovh) break;
      }
      if (rb) { ri = (L.g.nx - 2 * rb + rows - 3) / (rows - 2); ri += ri & 1; ri = std::max(ri, 2); }
    }
    if (rb && (L.g.nx < 2 * rb + 4)) rb = 0;
    if (!rb) ri = march_chunk(ctx, L, uj, ovh, wave, 0);
    else if (ctx->march_rb > 0) ri = march_chunk(ctx, L, uj, ovh, wave, rb);
  }
  dim3 grid(((L.g.ny + uj - 1) / uj + MARCH_WARPS - 1) / MARCH_WARPS, (L.g.nx + ri - 1) / ri);
  if (rb) grid.y = 2 + (unsigned)((L.g.nx - 2 * rb + ri - 1) / ri);
  const size_t sh = ring ? march_ring_smem(ri, nsw, res) : (size_t)3 * (std::max(ri, rb) + 2 * lead + 2 * nsw + 8) * sizeof(double);   // >= nrows of k_march
  const bool gather = push && !selfpush && Cr && Cr->replicated;   // the slab of the replicated coarse right-hand side goes to every rank
  if (push) {
    const long long offL = ctx->peerL ? (long long)((ctx->peerL - ctx->arena) / (ptrdiff_t)sizeof(double)) : 0;
    const long long offR = ctx->peerR ? (long long)((ctx->peerR - ctx->arena) / (ptrdiff_t)sizeof(double)) : 0;
    xf.pushw = std::min(HX - 1, L.g.nx);
    xf.dL = (ctx->peerL || selfpush) ? offL + (long long)L.g.nx * L.g.ld : 0;
    xf.dR = (ctx->peerR || selfpush) ? offR - (long long)L.g.nx * L.g.ld : 0;
    if (push_b) {
      xf.nxc = Cr->g.nx;
      xf.dbL = (ctx->peerL || selfpush) ? offL + (long long)Cr->g.nx * Cr->g.ld : 0;
      xf.dbR = (ctx->peerR || selfpush) ? offR - (long long)Cr->g.nx * Cr->g.ld : 0;
    }
    xf.zone = xf.pushw * (push_b ? L.cx : 1);
    xf.pt = ctx->d_pt;
    xf.hdr = (unsigned long long*)ctx->arena; xf.offL = offL; xf.offR = offR;
    if (selfpush) { xf.hdr = ctx->selfpush_hdr; xf.offL = xf.offR = 0; }
    xf.gather = gather ? 1 : 0;
    // chunks that start inside the left zone / end inside the right zone signal once per CTA
    const int gy = (int)grid.y, zone = std::min(xf.zone, L.g.nx);
    xf.rb = rb;
    if (rb) xf.incL = xf.incR = grid.x;   // rb >= zone: one boundary chunk per side
    else {
      xf.incL = grid.x * (unsigned)((zone + ri - 1) / ri);
      xf.incR = grid.x * (unsigned)(gy - (L.g.nx - zone) / ri);
    }
    if (gather) ctx->last_gather_ctas = grid.x * grid.y;
  } else if (res == 2 && ctx->d_pt) xf.pt = ctx->d_pt;   // the residual norm is summed over the ranks in the epilogue
  const int trace_key = (int)(&L - ctx->levels.data()) * 1000 + nsw * 100 + res * 10 + (Cp ? 2 : 0) + (push ? 1 : 0) + (omega > 0 ? 50000 : 0);
  xf.stamp = ctx->stamp_key == -2 || ctx->stamp_key == trace_key;
  trace_begin(ctx, trace_key);
  kern<<<grid, MARCH_THREADS, sh, ctx->stream>>>(L.g, L.mcx[op_tab(c)], L.mcy[op_tab(c)], xin, b, xout, ri, xf);
  trace_end(ctx);
  if (push) {
    ctx->halo_w[xout] = xf.pushw;
    if (push_b) ctx->halo_w[Cr->b] = xf.pushw;
  } else {
    ctx->halo_w.erase(xout);
    if (Cr) ctx->halo_w.erase(Cr->b);
  }
  NS_LAUNCH_CHECK();
  if (gather) {
    k_wait_gather<<<1, 1, 0, ctx->stream>>>(ctx->d_pt, ctx->last_gather_ctas);
    NS_LAUNCH_CHECK();
    return 0;
  }
  if (Cr && Cr->replicated && !L.replicated && ctx->world > 1) {
    // every rank contributes its slab of the coarse right-hand side; in place: the slabs are contiguous column ranges
    const size_t count = (size_t)(L.g.nx / L.cx) * Cr->g.ld;
    NS_NCCL(ncclAllGather(Cr->b + (size_t)HX * Cr->g.ld + rep_off_r, Cr->b + (size_t)HX * Cr->g.ld, count, ncclDouble, ctx->comm, ctx->stream));
  } else if (Cr && (rc = halo_exchange(ctx, *Cr, Cr->b, HX - 1, true))) return rc;
  return 0;
}

// `n` sweeps as passes of two (and one single if n is odd).  first_* apply to the first pass, last_* to the last.
// src0: the first pass reads its iterate from there instead of *x (initial guess kept in another array).
static int march_sweeps(nsgpu_ctx* ctx, const Level& L, OpCoef c, int n, double** x, double** xt, const double* b, bool xzero,
                        const Level* Cp, double* xc, const Level* Cr, double* norm_out, SolveCtl* ctl, double* src0 = nullptr) {
  for (int k = 0; k < n;) {
    const int nsw = (n - k >= 2) ? 2 : 1;
    const bool first = k == 0, last = k + nsw >= n;
    int rc = march_pass(ctx, L, c, nsw, (xzero && first) ? nullptr : (first && src0 ? src0 : *x), b, *xt, first ? Cp : nullptr, first ? xc : nullptr,
                        last ? Cr : nullptr, last ? norm_out : nullptr, last ? ctl : nullptr);
    if (rc) return rc;
    std::swap(*x, *xt);
    k += nsw;
  }
  return 0;
}

// One V(nu_pre, nu_post) cycle with every transfer fused into a smoothing pass: two passes per level.
static int vcycle_fused(nsgpu_ctx* ctx, int l, OpCoef c, double** x, double** xt, const double* b, bool xzero, double* norm_out,
                        SolveCtl* ctl, double* src0 = nullptr) {
  Level& L = ctx->levels[l];
  if (ctx->ctail_start > 0 && l == ctx->ctail_start && march_tab_ok(L, c)) {   // the rest of the cycle in one kernel on a cluster of CT_C CTAs
    const CtDesc& T = ctx->ctail[op_tab(c)];
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(ctx->ctail_c); cfg.blockDim = dim3(CT_THREADS);
    cfg.dynamicSmemBytes = (size_t)T.smem_doubles * sizeof(double);
    cfg.stream = ctx->stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = ctx->ctail_c; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    trace_begin(ctx, 99998);
    cudaError_t e = cudaLaunchKernelEx(&cfg, k_cluster_tail, T, (const double*)ctx->ctail_img[op_tab(c)], b, *x, ctx->tail_dbg);
    trace_end(ctx);
    if (e != cudaSuccess) { ctx->err = std::string("cluster tail launch: ") + cudaGetErrorString(e); return NSGPU_ERR_CUDA; }
    NS_LAUNCH_CHECK();
    return 0;
  }
  if (ctx->tail_start > 0 && l == ctx->tail_start && march_tab_ok(L, c)) {   // the rest of the cycle in one single-CTA kernel
    const TailDesc& T = ctx->tail[op_tab(c)];
    trace_begin(ctx, 99999);
    k_coarse_tail<<<1, TAIL_THREADS, (size_t)T.smem_doubles * sizeof(double), ctx->stream>>>(T, ctx->tail_img[op_tab(c)], b, *x, nullptr, 0u, ctx->tail_dbg);
    trace_end(ctx);
    NS_LAUNCH_CHECK();
    return 0;
  }
  if (l == (int)ctx->levels.size() - 1) {
    const int n = ctx->levels.size() == 1 ? ctx->nu_pre + ctx->nu_post : L.coarse_sweeps;
    return march_sweeps(ctx, L, c, n, x, xt, b, xzero, nullptr, nullptr, nullptr, norm_out, ctl, src0);
  }
  Level& C = ctx->levels[l + 1];
  int rc;
  if ((rc = march_sweeps(ctx, L, c, ctx->nu_pre, x, xt, b, xzero, nullptr, nullptr, &C, nullptr, nullptr, src0))) return rc;
  double *cx = C.x, *cxt = C.xt;
  if ((rc = vcycle_fused(ctx, l + 1, c, &cx, &cxt, C.b, true, nullptr, nullptr))) return rc;
  C.x = cx; C.xt = cxt;
  return march_sweeps(ctx, L, c, ctx->nu_post, x, xt, b, false, &C, C.x, nullptr, norm_out, ctl);
}

static inline bool all_levels_march(const nsgpu_ctx* ctx, OpCoef c) {
  for (const Level& L : ctx->levels) if (!march_level_ok(ctx, L, c)) return false;
  if (ctx->world > 1 && ctx->levels.size() > 1 && ctx->tail_start == 0) return false;   // slabs thinner than HX below
  return true;
}

// n sweeps of red-black Gauss-Seidel on level L; *x is the current iterate, *xt its ping-pong partner.
static int smooth(nsgpu_ctx* ctx, const Level& L, OpCoef c, double** x, double** xt, const double* b, int n) {
  const dim3 grid((L.g.ny + SM_TJ - 1) / SM_TJ, (L.g.nx + SM_TI - 1) / SM_TI);
  for (int s = 0; s < n; s++) {
    if (march_level_ok(ctx, L, c)) {   // plain single sweep through the marching kernel: 24 B per cell-update
      int rc = march_pass(ctx, L, c, 1, *x, b, *xt, nullptr, nullptr, nullptr, nullptr, nullptr);
      if (rc) return rc;
      std::swap(*x, *xt);
      continue;
    }
    int rc = halo_exchange(ctx, L, *x, 2);
    if (rc) return rc;
    ctx->halo_w.clear();
#define CALL(OP, RECT) k_smooth_rb<OP, RECT, false><<<grid, SM_THREADS, 0, ctx->stream>>>(L.g, *x, b, *xt, c.alpha, c.beta)
    NS_DISPATCH_OP_RECT(c.op, ctx->rect, CALL);
#undef CALL
    NS_LAUNCH_CHECK();
    std::swap(*x, *xt);
  }
  return 0;
}

// d_out[0] = ||b - A x||^2 over the local slab (all-reduced over ranks)
static int residual_norm2(nsgpu_ctx* ctx, const Level& L, OpCoef c, double* x, const double* b, double* d_out, SolveCtl* ctl = nullptr) {
  int rc = halo_exchange(ctx, L, x, 2, ctl != nullptr);   // device-controlled solves exchange x before their head
  if (rc) return rc;
  const dim3 grid = grid_cols(L.g, 256);
#define CALL(OP, RECT) k_residual<OP, RECT, false, true><<<grid, 256, 0, ctx->stream>>>(L.g, x, b, nullptr, c.alpha, c.beta, ctx->red, d_out, ctl, ctx->d_pt)
  NS_DISPATCH_OP_RECT(c.op, ctx->rect, CALL);
#undef CALL
  NS_LAUNCH_CHECK();
  return ctx->d_pt ? 0 : allreduce(ctx, d_out, 1, ncclSum);   // peer map: summed over the ranks in the kernel's epilogue
}

static int residual_vec(nsgpu_ctx* ctx, const Level& L, OpCoef c, double* x, const double* b, double* r, double* d_norm2) {
  int rc = halo_exchange(ctx, L, x, 2);
  if (rc) return rc;
  const dim3 grid = grid_cols(L.g, 256);
#define CALL(OP, RECT) k_residual<OP, RECT, true, true><<<grid, 256, 0, ctx->stream>>>(L.g, x, b, r, c.alpha, c.beta, ctx->red, d_norm2, nullptr, nullptr)
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

// n <= 32 device scalars -> host.  Written into mapped page-locked memory by a one-warp kernel (k_publish) rather than copied: the copy
// engines may be busy with hundreds of megabytes of field copies (nsgpu_step_host, nsgpu_export_begin) that a small copy would queue behind.
static int read_scalars(nsgpu_ctx* ctx, const double* d, int n, double* h) {
  k_publish<<<1, 32, 0, ctx->stream>>>(ctx->h_scal, d, n);
  cudaError_t e = cudaGetLastError();
  if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
  if (e != cudaSuccess) { ctx->err = std::string("scalar readback: ") + cudaGetErrorString(e); return NSGPU_ERR_CUDA; }
  for (int k = 0; k < n; k++) h[k] = ctx->h_scal[k];
  return 0;
}

// d_out[0] = (a, a) over the owned columns
static int norm2_vec(nsgpu_ctx* ctx, const Level& L, const double* a, double* d_out) {
  const size_t n = (size_t)L.g.nx * L.g.ld;
  const int blocks = (int)std::min<size_t>((n + 255) / 256, 148 * 8);
  k_dot<<<blocks, 256, 0, ctx->stream>>>(n, a + (size_t)HX * L.g.ld, a + (size_t)HX * L.g.ld, ctx->red, d_out, ctx->d_pt);
  NS_LAUNCH_CHECK();
  return ctx->d_pt ? 0 : allreduce(ctx, d_out, 1, ncclSum);
}

// V-cycle for the host-driven solvers: fused marching passes where every level supports them, generic kernels otherwise.
static int vcycle_any(nsgpu_ctx* ctx, OpCoef c, double** x, double** xt, const double* b, bool xzero) {
  if (all_levels_march(ctx, c)) {
    // the fused passes recompute halo columns and read b there: a right-hand side written by a vector kernel (owned columns only,
    // the Krylov recurrences) gets its halo columns first, or the preconditioner would differ from the one-rank one at the slab edges
    int rc = halo_exchange(ctx, ctx->levels[0], const_cast<double*>(b), HX - 1);
    if (rc) return rc;
    return vcycle_fused(ctx, 0, c, x, xt, b, xzero, nullptr, nullptr);
  }
  if (xzero) {
    cudaError_t e = cudaMemsetAsync(*x, 0, ctx->levels[0].stride * sizeof(double), ctx->stream);
    if (e != cudaSuccess) { ctx->err = cudaGetErrorString(e); return NSGPU_ERR_CUDA; }
  }
  return vcycle(ctx, 0, c, x, xt, b);
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
  // several ranks on a rectangle: the same fused passes as the device-controlled path, the residual norm comes out of
  // the last pass and is all-reduced; the host reads one scalar per cycle
  const bool sor = c.op == 0 && ctx->vel_omega > 0.0 && march_level_ok(ctx, L, c);
  const bool fused = sor || all_levels_march(ctx, c);
  while (rn > rtol * bn && out->cycles < ctx->max_cycles) {
    if (sor) {
      if ((rc = march_pass(ctx, L, c, 2, *x, b, *xt, nullptr, nullptr, nullptr, nullptr, nullptr, ctx->vel_omega))) return rc;
      std::swap(*x, *xt);
      if ((rc = march_pass(ctx, L, c, 2, *x, b, *xt, nullptr, nullptr, nullptr, ctx->d_scal + 1, nullptr, ctx->vel_omega))) return rc;
      std::swap(*x, *xt);
    } else if (fused) {
      if ((rc = vcycle_fused(ctx, 0, c, x, xt, b, false, ctx->d_scal + 1, nullptr))) return rc;
    } else {
      if ((rc = vcycle_any(ctx, c, x, xt, b, false))) return rc;
    }
    out->cycles++;
    if (fused) { if (!ctx->d_pt && (rc = allreduce(ctx, ctx->d_scal + 1, 1, ncclSum))) return rc; }
    else if ((rc = residual_norm2(ctx, L, c, *x, b, ctx->d_scal + 1))) return rc;
    if ((rc = read_scalars(ctx, ctx->d_scal + 1, 1, h))) return rc;
    rn = std::sqrt(h[0]);
    if (!(rn == rn)) break;
  }
  out->rel = rn / bn;
  out->converged = rn <= rtol * bn;
  return 0;
}

// ---------------------------------------------------------------------------------------------------------
// Outflow domains (ctx->aug): the phi solve is  A^ x = A x + mean(x) w = b  (nonsingular); x - mean(x) is the phi the reference's
// KSP converges to and mean(x) its c (ns_hostgeom.h).  d_scal[41..46]: sum x, (v, w), sum z, {scratch, |r|^2}, saved sum of x.

static int vec_sum(nsgpu_ctx* ctx, const Level& L, const double* a, double* d_out) {
  const size_t n = (size_t)L.g.nx * L.g.ld;
  const int blocks = (int)std::min<size_t>((n + 255) / 256, 148 * 8);
  k_sum<<<blocks, 256, 0, ctx->stream>>>(n, a + (size_t)HX * L.g.ld, ctx->red, d_out);
  NS_LAUNCH_CHECK();
  return allreduce(ctx, d_out, 1, ncclSum);
}

// field += gamma * (*d_s) on the fluid cells
static int add_scaled_const(nsgpu_ctx* ctx, const Level& L, double* field, const double* d_s, double gamma) {
  const dim3 grid = grid_cols(L.g, 256);
#define CALL(RECT) k_sub_mean<RECT><<<grid, 256, 0, ctx->stream>>>(L.g, field, d_s, -gamma)
  NS_DISPATCH_RECT(ctx->rect, CALL);
#undef CALL
  NS_LAUNCH_CHECK();
  ctx->halo_w.clear();
  return 0;
}

// y += sign * mean(x) * w;  d_out = {(dotwith, y), (y, y)}
static int aug_rank_one(nsgpu_ctx* ctx, const Level& L, const double* x, double* y, double sign, const double* dotwith, double* d_out) {
  int rc;
  if ((rc = vec_sum(ctx, L, x, ctx->d_scal + 41))) return rc;
  const size_t n = (size_t)L.g.nx * L.g.ld, off = (size_t)HX * L.g.ld;
  const int blocks = (int)std::min<size_t>((n + 255) / 256, 148 * 8);
  k_aug_axpy<<<blocks, 256, 0, ctx->stream>>>(n, y + off, ctx->aug_w + off, ctx->d_scal + 41, sign / (double)ctx->n_cells_global,
                                              dotwith ? dotwith + off : nullptr, ctx->red, d_out);
  NS_LAUNCH_CHECK();
  ctx->halo_w.clear();
  return allreduce(ctx, d_out, 2, ncclSum);
}

static int vcycle_any(nsgpu_ctx* ctx, OpCoef c, double** x, double** xt, const double* b, bool xzero);

// z = V_walls(v - mean v), its mean removed, + (v, w) / (w, w): the wall-only L_phi is symmetric and its V-cycle converges; the
// constant carries the component of v along w, which only the constant mode of x can produce (A^ 1 = w).
static int aug_precond(nsgpu_ctx* ctx, const Level& L, const double* v, double** z, double** zt) {
  int rc;
  const size_t n = (size_t)L.g.nx * L.g.ld, off = (size_t)HX * L.g.ld;
  const int blocks = (int)std::min<size_t>((n + 255) / 256, 148 * 8);
  const double inv_n = 1.0 / (double)ctx->n_cells_global;
  if ((rc = vec_sum(ctx, L, v, ctx->d_scal + 41))) return rc;
  k_dot<<<blocks, 256, 0, ctx->stream>>>(n, v + off, ctx->aug_w + off, ctx->red, ctx->d_scal + 42, nullptr);
  NS_LAUNCH_CHECK();
  if ((rc = allreduce(ctx, ctx->d_scal + 42, 1, ncclSum))) return rc;
  cudaError_t e = cudaMemcpyAsync(ctx->aug_t, v, L.stride * sizeof(double), cudaMemcpyDeviceToDevice, ctx->stream);
  if (e != cudaSuccess) { ctx->err = cudaGetErrorString(e); return NSGPU_ERR_CUDA; }
  if ((rc = add_scaled_const(ctx, L, ctx->aug_t, ctx->d_scal + 41, -inv_n))) return rc;
  if ((rc = vcycle_any(ctx, coef_phi_walls(), z, zt, ctx->aug_t, true))) return rc;
  if ((rc = vec_sum(ctx, L, *z, ctx->d_scal + 43))) return rc;
  if ((rc = add_scaled_const(ctx, L, *z, ctx->d_scal + 43, -inv_n))) return rc;
  return add_scaled_const(ctx, L, *z, ctx->d_scal + 42, 1.0 / ctx->aug_ww);
}

// Right-preconditioned BiCGStab, preconditioner = one V-cycle from a zero guess; all recurrence scalars stay
// on the device, one 8-byte readback per iteration for the stopping test.
//   d_scal layout: [0] rho_new [1] rho_old [2] alpha [3] omega [4..5] (r0,v),(v,v) [6..7] (t,s),(t,t)
//                  [8] (s,s) [9..10] (r,r),(r0,r) [11] (b,b)
// The pieces of a solve are launches only (no readback), so that they can be captured: bicg_begin -> WHILE { bicg_iteration } -> bicg_end.
struct BicgVecs { double *r, *r0, *p, *v, *s, *t; };
static inline BicgVecs bicg_vecs(nsgpu_ctx* ctx) { return {ctx->kry[0], ctx->kry[1], ctx->kry[2], ctx->kry[3], ctx->kry[4], ctx->kry[5]}; }

// r <- b - A x (the TRUE residual; on augmented systems minus mean(x) w), S[9] = (r, r)
static int bicg_true_residual(nsgpu_ctx* ctx, OpCoef c, double* x, const double* b, bool aug) {
  const Level& L = ctx->levels[0];
  double* S = ctx->d_scal;
  const BicgVecs V = bicg_vecs(ctx);
  int rc;
  if ((rc = residual_vec(ctx, L, c, x, b, V.r, S + 9))) return rc;
  if (aug) {
    if ((rc = aug_rank_one(ctx, L, x, V.r, -1.0, nullptr, S + 44))) return rc;
    k_scalar_copy<<<1, 1, 0, ctx->stream>>>(S + 9, S + 45);
    NS_LAUNCH_CHECK();
  }
  return 0;
}

// (re)start from the true residual in r: r0 = r, p = v = 0, recurrence scalars, loop condition
static int bicg_begin(nsgpu_ctx* ctx, SolveCtl* ctl, unsigned long long handle) {
  const Level& L = ctx->levels[0];
  const BicgVecs V = bicg_vecs(ctx);
  cudaError_t e = cudaMemcpyAsync(V.r0, V.r, L.stride * sizeof(double), cudaMemcpyDeviceToDevice, ctx->stream);
  if (e == cudaSuccess) e = cudaMemsetAsync(V.p, 0, L.stride * sizeof(double), ctx->stream);
  if (e == cudaSuccess) e = cudaMemsetAsync(V.v, 0, L.stride * sizeof(double), ctx->stream);
  if (e != cudaSuccess) { ctx->err = cudaGetErrorString(e); return NSGPU_ERR_CUDA; }
  k_kry_begin<<<1, 1, 0, ctx->stream>>>(ctx->d_scal, ctl, handle);
  NS_LAUNCH_CHECK();
  return 0;
}

// One iteration.  ph, sh and the preconditioner's ping-pong partner carry nothing from one iteration to the next, so every iteration starts
// from the same three arrays: the launch sequence is the same each time, which is what lets it be the body of a graph's WHILE node.
static int bicg_iteration(nsgpu_ctx* ctx, OpCoef c, double* x, bool aug, SolveCtl* ctl) {
  const Level& L = ctx->levels[0];
  const size_t n = (size_t)L.g.nx * L.g.ld, off = (size_t)HX * L.g.ld;
  const int blocks = (int)std::min<size_t>((n + 255) / 256, 148 * 8);
  const BicgVecs V = bicg_vecs(ctx);
  double *r = V.r, *r0 = V.r0, *p = V.p, *v = V.v, *s = V.s, *t = V.t;
  double *ph = ctx->kry[6], *sh = ctx->kry[7], *tmpA = ctx->tmp;
  double* S = ctx->d_scal;
  int rc;
  k_bicg_p<<<blocks, 256, 0, ctx->stream>>>(n, r + off, p + off, v + off, S);
  NS_LAUNCH_CHECK();
  // ph = M^-1 p
  { double *a = ph, *bq = tmpA; if ((rc = aug ? aug_precond(ctx, L, p, &a, &bq) : vcycle_any(ctx, c, &a, &bq, p, true))) return rc; ph = a; tmpA = bq; }
  if (aug) {
    if ((rc = apply_op(ctx, L, c, ph, v, nullptr, nullptr))) return rc;
    if ((rc = aug_rank_one(ctx, L, ph, v, 1.0, r0, S + 4))) return rc;
  } else if ((rc = apply_op(ctx, L, c, ph, v, r0, S + 4))) return rc;
  k_scalar_div<<<1, 1, 0, ctx->stream>>>(S + 2, S + 0, S + 4);   // alpha = rho / (r0, v)
  NS_LAUNCH_CHECK();
  k_bicg_s<<<blocks, 256, 0, ctx->stream>>>(n, r + off, v + off, s + off, S + 2, ctx->red, S + 8);
  NS_LAUNCH_CHECK();
  if ((rc = allreduce(ctx, S + 8, 1, ncclSum))) return rc;
  // sh = M^-1 s
  { double *a = sh, *bq = tmpA; if ((rc = aug ? aug_precond(ctx, L, s, &a, &bq) : vcycle_any(ctx, c, &a, &bq, s, true))) return rc; sh = a; tmpA = bq; }
  if (aug) {
    if ((rc = apply_op(ctx, L, c, sh, t, nullptr, nullptr))) return rc;
    if ((rc = aug_rank_one(ctx, L, sh, t, 1.0, s, S + 6))) return rc;
  } else if ((rc = apply_op(ctx, L, c, sh, t, s, S + 6))) return rc;
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
  k_kry_ctl<<<1, 1, 0, ctx->stream>>>(S, ctl);                   // stopping test of the recurrence residual, on the device
  NS_LAUNCH_CHECK();
  return 0;
}

static int bicg_end(nsgpu_ctx* ctx, int slot, OpCoef c, double* x, const double* b, bool aug) {
  int rc = bicg_true_residual(ctx, c, x, b, aug);
  if (rc) return rc;
  k_kry_end<<<1, 1, 0, ctx->stream>>>(ctx->d_scal, ctx->d_ctl + slot, ctx->h_ctl + slot);   // h_ctl: mapped host memory
  NS_LAUNCH_CHECK();
  return 0;
}

#define NS_GRAPH_TRY(call)                                                                          \
  do {                                                                                              \
    cudaError_t e_ = (call);                                                                        \
    if (e_ != cudaSuccess) {                                                                        \
      fprintf(stderr, "nsgpu: %s -> %s; falling back to direct launches\n", #call, cudaGetErrorString(e_)); \
      cudaGetLastError();                                                                           \
      return 1;                                                                                     \
    }                                                                                               \
  } while (0)

// begin -> WHILE (go) { iteration } -> true residual + publish, as one graph.  0 ok, 1 unsupported, < 0 error.
static int build_kry_graph_(nsgpu_ctx* ctx, int slot, OpCoef c, double* x, const double* b, bool aug) {
  SolveGraph& G = ctx->kgraph[slot];
  if (G.exec) { cudaGraphExecDestroy(G.exec); G.exec = nullptr; }
  if (G.graph) { cudaGraphDestroy(G.graph); G.graph = nullptr; }
  const cudaStreamCaptureMode mode = cudaStreamCaptureModeThreadLocal;
  SolveCtl* ctl = ctx->d_ctl + slot;
  NS_GRAPH_TRY(cudaGraphCreate(&G.graph, 0));
  cudaGraphConditionalHandle handle;
  NS_GRAPH_TRY(cudaGraphConditionalHandleCreate(&handle, G.graph, 0, cudaGraphCondAssignDefault));
  const int64_t l0 = ctx->launches;
  NS_GRAPH_TRY(cudaStreamBeginCaptureToGraph(ctx->stream, G.graph, nullptr, nullptr, 0, mode));
  int rc = bicg_begin(ctx, ctl, (unsigned long long)handle);
  cudaStreamCaptureStatus st;
  const cudaGraphNode_t* deps = nullptr;
  size_t ndeps = 0;
  cudaGraph_t gcap = nullptr;
  cudaError_t e = cudaStreamGetCaptureInfo(ctx->stream, &st, nullptr, &gcap, &deps, &ndeps);
  std::vector<cudaGraphNode_t> leaf(deps, deps + (e == cudaSuccess ? ndeps : 0));
  cudaGraph_t gout = nullptr;
  cudaError_t e2 = cudaStreamEndCapture(ctx->stream, &gout);
  if (rc) return rc;
  NS_GRAPH_TRY(e);
  NS_GRAPH_TRY(e2);
  const int64_t l1 = ctx->launches;
  cudaGraphNodeParams p = {cudaGraphNodeTypeConditional};
  p.type = cudaGraphNodeTypeConditional;
  p.conditional.handle = handle;
  p.conditional.type = cudaGraphCondTypeWhile;
  p.conditional.size = 1;
  cudaGraphNode_t loop;
  NS_GRAPH_TRY(cudaGraphAddNode(&loop, G.graph, leaf.data(), leaf.size(), &p));
  cudaGraph_t body = p.conditional.phGraph_out[0];
  NS_GRAPH_TRY(cudaStreamBeginCaptureToGraph(ctx->stream, body, nullptr, nullptr, 0, mode));
  rc = bicg_iteration(ctx, c, x, aug, ctl);
  e2 = cudaStreamEndCapture(ctx->stream, &gout);
  if (rc) return rc;
  NS_GRAPH_TRY(e2);
  const int64_t l2 = ctx->launches;
  NS_GRAPH_TRY(cudaStreamBeginCaptureToGraph(ctx->stream, G.graph, &loop, nullptr, 1, mode));
  rc = bicg_end(ctx, slot, c, x, b, aug);
  e2 = cudaStreamEndCapture(ctx->stream, &gout);
  if (rc) return rc;
  NS_GRAPH_TRY(e2);
  const int64_t l3 = ctx->launches;
  NS_GRAPH_TRY(cudaGraphInstantiate(&G.exec, G.graph, 0));
  ctx->launches = l0;   // capture is not execution
  G.launches_fixed = (int)((l1 - l0) + (l3 - l2));
  G.launches_body = (int)(l2 - l1);
  G.x = x; G.b = b; G.xt = ctx->tmp; G.x0 = nullptr; G.rtol = (double)(c.op * 2 + (aug ? 1 : 0));   // (operator, augmented) the graph was built for
  return 0;
}
static int build_kry_graph(nsgpu_ctx* ctx, int slot, OpCoef c, double* x, const double* b, bool aug) {
  ctx->capturing = true; ctx->capture_nccl = false;
  int rc = build_kry_graph_(ctx, slot, c, x, b, aug);
  ctx->capturing = false;
  if (rc == 0 && ctx->capture_nccl) rc = 1;   // several ranks without the peer map: NCCL all-reduces inside the iteration
  if (rc > 0) {
    cudaStreamCaptureStatus st = cudaStreamCaptureStatusNone;
    cudaStreamIsCapturing(ctx->stream, &st);
    if (st != cudaStreamCaptureStatusNone) { cudaGraph_t g = nullptr; cudaStreamEndCapture(ctx->stream, &g); }
    cudaGetLastError();
    SolveGraph& G = ctx->kgraph[slot];
    if (G.exec) { cudaGraphExecDestroy(G.exec); G.exec = nullptr; }
  }
  return rc;
}

// BiCGStab with the stopping test on the device.  With graphs: one graph launch and one readback per (re)start; without (NSGPU_GRAPH=0,
// several ranks, or a capture the driver refuses): the same kernels launched directly, the loop condition read back once per iteration.
static int solve_bicgstab(nsgpu_ctx* ctx, int slot, OpCoef c, double** xp, double** xtp, const double* b, double rtol, SolveResult* out) {
  (void)xtp;
  const Level& L = ctx->levels[0];
  double* x = *xp;
  double* S = ctx->d_scal;
  SolveCtl* ctl = ctx->d_ctl + slot;
  int rc;
  const bool aug = ctx->aug && c.op == 1;
  if ((rc = norm2_vec(ctx, L, b, S + 11))) return rc;
  if (aug && (rc = add_scaled_const(ctx, L, x, S + 46, 1.0 / (double)ctx->n_cells_global))) return rc;   // warm start: phi^n + the last mean(x)
  if ((rc = bicg_true_residual(ctx, c, x, b, aug))) return rc;
  k_kry_first<<<1, 1, 0, ctx->stream>>>(S, ctl, rtol * rtol, ctx->max_cycles);
  NS_LAUNCH_CHECK();
  // The recurrence residual can drift from the true one.  After the recurrence says "converged" the true residual is recomputed (into
  // r: it is the restart vector if one is needed); if it misses rtol the iteration restarts from the current x (r0 = r = true residual)
  // -- at most MAX_RESTARTS times; a solve is reported as converged only when the TRUE residual meets rtol (include/nsgpu.h).
  constexpr int MAX_RESTARTS = 3;
  // (several ranks: the all-reduces and halo exchanges of this route are NCCL calls -- no capture is attempted, the loop condition is read back)
  bool use_graph = ctx->use_graph != 0 && ctx->world == 1 && !ctx->kry_graph_off;
  double bn = 0.0, rn = 0.0;
  out->cycles = 0;
  for (int restart = 0;; restart++) {
    if (use_graph) {
      SolveGraph& G = ctx->kgraph[slot];
      const double key = (double)(c.op * 2 + (aug ? 1 : 0));
      if (!G.exec || G.x != x || G.b != b || G.xt != ctx->tmp || G.rtol != key) {
        rc = build_kry_graph(ctx, slot, c, x, b, aug);
        if (rc < 0) return rc;
        if (rc > 0) { use_graph = false; ctx->kry_graph_off = true; if (ctx->world == 1) ctx->graph_fallback = true; }
      }
    }
    const int cycles_before = out->cycles;
    if (use_graph) {
      SolveGraph& G = ctx->kgraph[slot];
      cudaError_t e = cudaGraphLaunch(G.exec, ctx->stream);
      if (e != cudaSuccess) { ctx->err = std::string("cudaGraphLaunch (Krylov solve): ") + cudaGetErrorString(e); return NSGPU_ERR_CUDA; }
      ctx->launches += G.launches_fixed;
    } else {
      if ((rc = bicg_begin(ctx, ctl, 0ull))) return rc;
      for (;;) {
        k_publish_ctl<<<1, 1, 0, ctx->stream>>>(ctx->h_ctl + slot, ctl);
        cudaError_t e = cudaGetLastError();
        if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
        if (e != cudaSuccess) { ctx->err = std::string("ctl readback: ") + cudaGetErrorString(e); return NSGPU_ERR_CUDA; }
        if (!ctx->h_ctl[slot].go) break;
        if ((rc = bicg_iteration(ctx, c, x, aug, ctl))) return rc;
      }
      if ((rc = bicg_end(ctx, slot, c, x, b, aug))) return rc;
    }
    {
      cudaError_t e = cudaStreamSynchronize(ctx->stream);
      if (e != cudaSuccess) { ctx->err = std::string("Krylov solve: ") + cudaGetErrorString(e); return NSGPU_ERR_CUDA; }
    }
    const SolveCtl& h = ctx->h_ctl[slot];
    out->cycles = h.cycles;
    if (use_graph) ctx->launches += (int64_t)(out->cycles - cycles_before) * ctx->kgraph[slot].launches_body;
    bn = std::sqrt(h.bn2);
    rn = std::sqrt(h.rn2);
    if (!(h.bn2 > 0.0)) {   // KSPSolve returns x = 0 for b = 0
      cudaMemsetAsync(x, 0, L.stride * sizeof(double), ctx->stream);
      out->rel = 0; out->converged = true;
      return 0;
    }
    if (restart == 0 && out->cycles == 0 && rn <= rtol * bn) { out->rel = rn / bn; out->converged = true; return 0; }   // the guess was the answer
    if (!(rn == rn) || rn <= rtol * bn || out->cycles >= ctx->max_cycles || restart >= MAX_RESTARTS) break;
    ctx->krylov_restarts++;
  }
  if (aug && (rc = vec_sum(ctx, L, x, S + 46))) return rc;   // mean(x) * n: next step's warm start (phase_phi removes it from phi)
  out->rel = rn / bn;
  out->converged = rn <= rtol * bn;
  return 0;
}

// ---------------------------------------------------------------------------------------------------------
// Device-controlled multigrid solve (one rank, rectangle): the stopping test runs in the epilogue of the
// residual reductions (CtlEpilogue), so the host never has to look at a residual between cycles.

// One iteration of a device-controlled solve: a V-cycle, or (velocity, SOR mode) four over-relaxed sweeps; the last
// pass reduces the residual norm and its epilogue updates the control block.
// check == false: the iteration without its stopping test (no residual reduction in the last pass: a cheaper kernel variant, and no
// cross-rank sum on several ranks) -- solve_head runs such iterations when the test could not succeed yet
static int solve_body(nsgpu_ctx* ctx, int slot, OpCoef c, double** x, double** xt, const double* b, double* src0 = nullptr, bool check = true) {
  SolveCtl* ctl = check ? ctx->d_ctl + slot : nullptr;
  double* norm = check ? &ctx->d_ctl[slot].rn2 : nullptr;
  if (c.op == 0 && ctx->vel_omega > 0.0) {
    const Level& L = ctx->levels[0];
    int rc = march_pass(ctx, L, c, 2, src0 ? src0 : *x, b, *xt, nullptr, nullptr, nullptr, nullptr, nullptr, ctx->vel_omega);
    if (rc) return rc;
    std::swap(*x, *xt);
    rc = march_pass(ctx, L, c, 2, *x, b, *xt, nullptr, nullptr, nullptr, norm, ctl, ctx->vel_omega);
    if (rc) return rc;
    std::swap(*x, *xt);
    return 0;
  }
  return vcycle_fused(ctx, 0, c, x, xt, b, false, norm, ctl, src0);
}

// Head of a solve: |b|^2, then the FIRST iteration, unconditionally, reading the initial guess from x0 (which may be
// another array: no copy of the guess, no separate residual pass for it).  Its residual reduction sets the loop
// condition.  The reference's KSP tests the initial residual first; here a guess that is already converged costs one
// iteration more and comes out more accurate.
static int solve_head(nsgpu_ctx* ctx, int slot, OpCoef c, double* x, double* xt, double* x0, const double* b, double rtol,
                      unsigned long long handle, bool bn2_pre, int k0 = 1) {
  const Level& L = ctx->levels[0];
  SolveCtl* ctl = ctx->d_ctl + slot;
  const bool iterate = (c.op == 0 && ctx->vel_omega > 0.0) || ctx->nu_pre >= 1 || ctx->levels.size() == 1;
  if (!iterate) k0 = 1;
  k_ctl_init<<<1, 1, 0, ctx->stream>>>(ctl, rtol * rtol, ctx->max_cycles, handle, iterate ? k0 - 1 : -1, bn2_pre ? ctx->d_scal + 30 + slot : nullptr, k0 - 1);
  NS_LAUNCH_CHECK();
  int rc;
  if (!bn2_pre && (rc = norm2_vec(ctx, L, b, &ctl->bn2))) return rc;
  if (iterate) {
    double *xa = x, *xb = xt;
    if (slot == 2 && ctx->lazy_mean && ctx->levels.size() > 1 && ctx->nu_pre >= 1 && ctx->rect && ctx->phys_march) ctx->next_pass_shift = &ctl->xsum;   // the first pass (a restricting one) removes the mean the last solve left in phi
    // the first k0 - 1 iterations without a stopping test (it could not succeed: the last time step's solve needed more), then the first tested one
    for (int it = 0; it < k0; it++) {
      if ((rc = solve_body(ctx, slot, c, &xa, &xb, b, (it == 0 && x0 && x0 != x) ? x0 : nullptr, it == k0 - 1))) return rc;
      if (xa != x) { ctx->err = "internal: an iteration does not return to its buffer"; return NSGPU_ERR_ARG; }
    }
    return 0;
  }
  if (x0 && x0 != x) {
    cudaError_t e = cudaMemcpyAsync(x, x0, L.stride * sizeof(double), cudaMemcpyDeviceToDevice, ctx->stream);
    if (e != cudaSuccess) { ctx->err = cudaGetErrorString(e); return NSGPU_ERR_CUDA; }
  }
  return residual_norm2(ctx, L, c, x, b, &ctl->rn2, ctl);
}

static int solve_tail(nsgpu_ctx* ctx, int slot, double* x) {
  const Level& L = ctx->levels[0];
  ctx->halo_w.erase(x);
  k_zero_if<<<148 * 4, 256, 0, ctx->stream>>>(L.stride, x, ctx->d_ctl + slot, ctx->h_ctl + slot);   // h_ctl: mapped host memory
  NS_LAUNCH_CHECK();
  return 0;
}

// Builds head -> WHILE(go){ V-cycle + residual norm } -> tail for (x, xt, b).  Returns 0 ok, 1 unsupported, < 0 error.
static int build_solve_graph_(nsgpu_ctx* ctx, int slot, OpCoef c, double* x, double* xt, double* x0, const double* b, double rtol, bool bn2_pre, int k0);
static int build_solve_graph(nsgpu_ctx* ctx, int slot, OpCoef c, double* x, double* xt, double* x0, const double* b, double rtol, bool bn2_pre, int k0) {
  ctx->capturing = true; ctx->capture_nccl = false;
  int rc = build_solve_graph_(ctx, slot, c, x, xt, x0, b, rtol, bn2_pre, k0);
  ctx->capturing = false;
  if (rc == 0 && ctx->capture_nccl) {
    fprintf(stderr, "nsgpu: a solve would need NCCL inside its loop; direct launches instead of graphs\n");
    rc = 1;
  }
  return rc;
}
static int build_solve_graph_(nsgpu_ctx* ctx, int slot, OpCoef c, double* x, double* xt, double* x0, const double* b, double rtol, bool bn2_pre, int k0) {
  SolveGraph& G = ctx->sgraph[slot];
  if (G.exec) { cudaGraphExecDestroy(G.exec); G.exec = nullptr; }
  if (G.graph) { cudaGraphDestroy(G.graph); G.graph = nullptr; }
  const cudaStreamCaptureMode mode = cudaStreamCaptureModeThreadLocal;
  NS_GRAPH_TRY(cudaGraphCreate(&G.graph, 0));
  cudaGraphConditionalHandle handle;
  NS_GRAPH_TRY(cudaGraphConditionalHandleCreate(&handle, G.graph, 0, cudaGraphCondAssignDefault));
  const int64_t l0 = ctx->launches;
  int rc;
  // head
  NS_GRAPH_TRY(cudaStreamBeginCaptureToGraph(ctx->stream, G.graph, nullptr, nullptr, 0, mode));
  rc = solve_head(ctx, slot, c, x, xt, x0, b, rtol, (unsigned long long)handle, bn2_pre, k0);
  cudaStreamCaptureStatus st;
  const cudaGraphNode_t* deps = nullptr;
  size_t ndeps = 0;
  cudaGraph_t gcap = nullptr;
  cudaError_t e = cudaStreamGetCaptureInfo(ctx->stream, &st, nullptr, &gcap, &deps, &ndeps);
  std::vector<cudaGraphNode_t> leaf(deps, deps + (e == cudaSuccess ? ndeps : 0));
  cudaGraph_t gout = nullptr;
  cudaError_t e2 = cudaStreamEndCapture(ctx->stream, &gout);
  if (rc) return rc;
  NS_GRAPH_TRY(e);
  NS_GRAPH_TRY(e2);
  const int64_t l1 = ctx->launches;
  // the loop
  cudaGraphNodeParams p = {cudaGraphNodeTypeConditional};
  p.type = cudaGraphNodeTypeConditional;
  p.conditional.handle = handle;
  p.conditional.type = cudaGraphCondTypeWhile;
  p.conditional.size = 1;
  cudaGraphNode_t loop;
  NS_GRAPH_TRY(cudaGraphAddNode(&loop, G.graph, leaf.data(), leaf.size(), &p));
  cudaGraph_t body = p.conditional.phGraph_out[0];
  NS_GRAPH_TRY(cudaStreamBeginCaptureToGraph(ctx->stream, body, nullptr, nullptr, 0, mode));
  double *xa = x, *xb = xt;
  rc = solve_body(ctx, slot, c, &xa, &xb, b);
  e2 = cudaStreamEndCapture(ctx->stream, &gout);
  if (rc) return rc;
  NS_GRAPH_TRY(e2);
  if (xa != x) { fprintf(stderr, "nsgpu: V-cycle does not return to its buffer (odd number of passes); no graph\n"); return 1; }
  const int64_t l2 = ctx->launches;
  // tail
  NS_GRAPH_TRY(cudaStreamBeginCaptureToGraph(ctx->stream, G.graph, &loop, nullptr, 1, mode));
  rc = solve_tail(ctx, slot, x);
  e2 = cudaStreamEndCapture(ctx->stream, &gout);
  if (rc) return rc;
  NS_GRAPH_TRY(e2);
  const int64_t l3 = ctx->launches;
  NS_GRAPH_TRY(cudaGraphInstantiate(&G.exec, G.graph, 0));
  ctx->launches = l0;   // capture is not execution
  G.launches_fixed = (int)((l1 - l0) + (l3 - l2));
  G.launches_body = (int)(l2 - l1);
  G.x = x; G.xt = xt; G.b = b; G.x0 = x0; G.rtol = rtol; G.bn2_pre = bn2_pre;
  G.head_iterations = ((c.op == 0 && ctx->vel_omega > 0.0) || ctx->nu_pre >= 1 || ctx->levels.size() == 1) ? k0 : 0;
  G.k0 = k0;
  return 0;
}

// Issues the whole solve on the stream; results land in h_ctl[slot] (decode_solve after a synchronize).
// x0: where the initial guess is (nullptr or *x: in place)
static int solve_mg_dev(nsgpu_ctx* ctx, int slot, OpCoef c, double** x, double** xt, double* x0, const double* b, double rtol) {
  int rc;
  if (!x0) x0 = *x;
  const bool bn2_pre = ctx->bn2_ready[slot];
  ctx->bn2_ready[slot] = false;
  if (ctx->world > 1) {   // the only NCCL step of the solve: the initial guess' halo columns, as wide as the passes push
    if ((rc = halo_exchange(ctx, ctx->levels[0], x0, HX - 1))) return rc;
    ctx->halo_w[x0] = std::min(HX - 1, ctx->levels[0].g.nx);
  }
  // (velocity solves only: their iteration count is the same for hundreds of steps, so the graph is rebuilt once; the phi solve's count
  // moves every few steps from rest -- each change costs a graph capture -- and its last pass is hardly cheaper without the reduction:
  // measured 1.807 -> 1.801 ms per solve, against 1.029 -> 0.968 ms for the two velocity solves)
  const int k0 = (ctx->skip_checks && slot < 2) ? std::max(1, std::min(ctx->solve_k0[slot], ctx->max_cycles)) : 1;
  if (ctx->use_graph) {
    SolveGraph& G = ctx->sgraph[slot];
    if (!G.exec || G.x != *x || G.xt != *xt || G.b != b || G.x0 != x0 || G.rtol != rtol || G.bn2_pre != bn2_pre || G.k0 != k0) {
      rc = build_solve_graph(ctx, slot, c, *x, *xt, x0, b, rtol, bn2_pre, k0);
      if (rc < 0) return rc;
      if (rc > 0) {
        cudaStreamCaptureStatus st = cudaStreamCaptureStatusNone;
        cudaStreamIsCapturing(ctx->stream, &st);
        if (st != cudaStreamCaptureStatusNone) { cudaGraph_t g = nullptr; cudaStreamEndCapture(ctx->stream, &g); }
        cudaGetLastError();
        ctx->use_graph = 0;
        ctx->graph_fallback = true;   // reported in nsgpu_step_info.flags
      }
    }
    if (ctx->use_graph) {
      cudaError_t e = cudaGraphLaunch(G.exec, ctx->stream);
      if (e != cudaSuccess) { ctx->err = std::string("cudaGraphLaunch: ") + cudaGetErrorString(e); return NSGPU_ERR_CUDA; }
      ctx->launches += G.launches_fixed;
      ctx->graph_cycles_pending[slot] += 1;
      return 0;
    }
  }
  if ((rc = solve_head(ctx, slot, c, *x, *xt, x0, b, rtol, 0ull, bn2_pre, k0))) return rc;
  for (;;) {
    k_publish_ctl<<<1, 1, 0, ctx->stream>>>(ctx->h_ctl + slot, ctx->d_ctl + slot);
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    if (e != cudaSuccess) { ctx->err = std::string("ctl readback: ") + cudaGetErrorString(e); return NSGPU_ERR_CUDA; }
    if (!ctx->h_ctl[slot].go) break;
    if ((rc = solve_body(ctx, slot, c, x, xt, b))) return rc;
  }
  return solve_tail(ctx, slot, *x);
}

// After a stream synchronize: turn h_ctl[slot] into a SolveResult and book the loop-body launches of graph solves.
static void decode_solve(nsgpu_ctx* ctx, int slot, SolveResult* out) {
  const SolveCtl& h = ctx->h_ctl[slot];
  out->cycles = h.cycles;
  out->rel = h.bn2 > 0 ? std::sqrt(h.rn2 / h.bn2) : 0.0;
  out->converged = !(h.bn2 > 0) || h.rn2 <= h.tol2 * h.bn2;
  const int new_cycles = h.total_cycles - ctx->seen_cycles[slot];
  if (ctx->graph_cycles_pending[slot]) {
    const int in_heads = ctx->graph_cycles_pending[slot] * ctx->sgraph[slot].head_iterations;   // already in launches_fixed
    ctx->launches += (int64_t)std::max(new_cycles - in_heads, 0) * ctx->sgraph[slot].launches_body;
    ctx->graph_cycles_pending[slot] = 0;
  }
  // the next solves of this slot test from the iteration before the smallest count of the last four solves: a count that hovers between two
  // values (a residual at the tolerance) does not change k0 -- every change of k0 costs a graph capture
  {
    int* hst = ctx->solve_hist[slot];
    hst[ctx->solve_hist_n[slot]++ & 3] = h.cycles;
    int lo = h.cycles;
    for (int q = 0; q < std::min(ctx->solve_hist_n[slot], 4); q++) lo = std::min(lo, hst[q]);
    ctx->solve_k0[slot] = std::max(1, lo - 1);   // (results are decoded once per nsgpu_step call: within a call of many steps k0 does not change)
  }
  if (slot == 2) ctx->fine_sweeps_phi += (double)new_cycles * (ctx->nu_pre + ctx->nu_post);
  ctx->seen_cycles[slot] = h.total_cycles;
}

static inline bool dev_solve_ok(const nsgpu_ctx* ctx, OpCoef c) {
  if (ctx->solver != 0 || (ctx->world != 1 && !ctx->d_pt)) return false;
  if (c.op == 0 && ctx->vel_omega > 0.0) return march_level_ok(ctx, ctx->levels[0], c);
  return all_levels_march(ctx, c);
}

// phi <- phi - mean(phi) now, if the marching solve left it pending (ctx->phi_shift); see nsgpu_ctx::lazy_mean
static int phi_normalize(nsgpu_ctx* ctx) {
  if (!ctx->phi_shift) return 0;
  const Level& L = ctx->levels[0];
  const dim3 grid = grid_cols(L.g, 256);
  const double inv_n = 1.0 / (double)ctx->n_cells_global;
  double* xsum = &ctx->d_ctl[2].xsum;
  if (!ctx->phi_sum_known) {   // the projection that reduces sum(phi) has not run since the solve
    const size_t n = (size_t)L.g.nx * L.g.ld;
    const int blocks = (int)std::min<size_t>((n + 255) / 256, 148 * 8);
    k_sum<<<blocks, 256, 0, ctx->stream>>>(n, ctx->phi + (size_t)HX * L.g.ld, ctx->red, xsum);
    NS_LAUNCH_CHECK();
    int rc = allreduce(ctx, xsum, 1, ncclSum);
    if (rc) return rc;
  }
#define CALL(RECT) k_sub_mean<RECT><<<grid, 256, 0, ctx->stream>>>(L.g, ctx->phi, xsum, inv_n)
  NS_DISPATCH_RECT(ctx->rect, CALL);
#undef CALL
  NS_LAUNCH_CHECK();
  k_scalar_set<<<1, 1, 0, ctx->stream>>>(xsum, 0.0);
  NS_LAUNCH_CHECK();
  ctx->phi_shift = false;
  ctx->phi_sum_known = false;
  ctx->halo_w.clear();
  return 0;
}

// slot: 0 = u, 1 = v, 2 = phi.  On the device-controlled path the result is pending until finish_results().
static int solve(nsgpu_ctx* ctx, int slot, OpCoef c, double** x, const double* b, double rtol, SolveResult* out, double* x0 = nullptr) {
  const bool krylov = ctx->solver == 1 || (ctx->krylov_phi && c.op == 1);
  if (!krylov && dev_solve_ok(ctx, c)) { out->pending = true; return solve_mg_dev(ctx, slot, c, x, &ctx->tmp, x0, b, rtol); }
  out->pending = false;
  ctx->bn2_ready[slot] = false;
  if (slot == 2) { int rcn = phi_normalize(ctx); if (rcn) return rcn; }
  if (x0 && x0 != *x) {   // host-driven solvers iterate in place
    cudaError_t e = cudaMemcpyAsync(*x, x0, ctx->levels[0].stride * sizeof(double), cudaMemcpyDeviceToDevice, ctx->stream);
    if (e != cudaSuccess) { ctx->err = cudaGetErrorString(e); return NSGPU_ERR_CUDA; }
  }
  if (krylov) return solve_bicgstab(ctx, slot, c, x, &ctx->tmp, b, rtol, out);
  return solve_mg(ctx, c, x, &ctx->tmp, b, rtol, out);
}

static void finish_results(nsgpu_ctx* ctx, SolveResult* ru, SolveResult* rv, SolveResult* rp) {
  SolveResult* r[3] = {ru, rv, rp};
  for (int k = 0; k < 3; k++)
    if (r[k] && r[k]->pending) { decode_solve(ctx, k, r[k]); r[k]->pending = false; }
}

// ---------------------------------------------------------------------------------------------------------
// Phases of one time step

struct NvtxRange {   // host-side range around the launches of one phase (the kernels inherit it in nsys / ncu --nvtx views)
  explicit NvtxRange(const char* name) { nvtxRangePushA(name); }
  ~NvtxRange() { nvtxRangePop(); }
};

// Launch shape of the projection kernels of ns_phys.cuh: gx CTAs of 4 warps cover the rows, the columns are cut into chunks of *ri
// so that the CTAs fill whole waves (same estimate as march_chunk: waves x (columns per chunk + start-up)).
static dim3 phys_grid(const void* kern, int fields, const Level& L, int* ri, size_t* smem) {
  static thread_local std::vector<std::pair<const void*, int>> cache;
  int wave = 0;
  for (auto& e : cache) if (e.first == kern) wave = e.second;
  if (!wave) {
    int per_sm = 0, dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)phys_smem(1024, fields));
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, PH_THREADS, phys_smem(128, fields)) != cudaSuccess || per_sm < 1) per_sm = 4;
    wave = per_sm * sms;
    cache.push_back({kern, wave});
  }
  const int gx = ((L.g.ny + PH_UJ - 1) / PH_UJ + PH_WARPS - 1) / PH_WARPS;
  long long best = -1;
  int r_best = L.g.nx;
  for (int r = 8; r <= std::min(L.g.nx, 1024); r += 4) {
    const long long ctas = (long long)gx * ((L.g.nx + r - 1) / r);
    const long long cost = ((ctas + wave - 1) / wave) * (r + 2 * PH_D);
    if (best < 0 || cost < best) { best = cost; r_best = r; }
  }
  if (L.g.nx < 8) r_best = L.g.nx;
  while ((long long)gx * ((L.g.nx + r_best - 1) / r_best) > RED_MAX_BLOCKS) r_best *= 2;
  *ri = r_best;
  *smem = phys_smem(r_best, fields);
  return dim3(gx, (L.g.nx + r_best - 1) / r_best);
}

// launch shape of k_rhs_march over `ncols` columns: whole waves of CTAs, like phys_grid
static dim3 rhs_grid(const Level& L, int ncols, int* ri, size_t* smem) {
  static thread_local int wave = 0;
  if (!wave) {
    int per_sm = 0, dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    cudaFuncSetAttribute(k_rhs_march, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)rhs_smem(1024));
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_rhs_march, RHS_THREADS, rhs_smem(128)) != cudaSuccess || per_sm < 1) per_sm = 2;
    wave = per_sm * sms;
  }
  const int gx = ((L.g.ny + RHS_UJ - 1) / RHS_UJ + RHS_WARPS - 1) / RHS_WARPS;
  long long best = -1;
  int r_best = ncols;
  for (int r = 8; r <= std::min(ncols, 1024); r += 4) {
    const long long ctas = (long long)gx * ((ncols + r - 1) / r);
    const long long cost = ((ctas + wave - 1) / wave) * (r + RHS_D + 4);
    if (best < 0 || cost < best) { best = cost; r_best = r; }
  }
  if (ncols < 8) r_best = ncols;
  while ((long long)gx * ((ncols + r_best - 1) / r_best) > RED_MAX_BLOCKS) r_best *= 2;
  *ri = r_best;
  *smem = rhs_smem(r_best);
  return dim3(gx, (ncols + r_best - 1) / r_best);
}

static int phase_rhs(nsgpu_ctx* ctx) {   // ConstructRHS_V, FS.cpp:546
  NvtxRange nvtx_("nsgpu:ConstructRHS_V");
  ctx->halo_w.clear();   // every field may have been rewritten by a non-marching kernel since
  const Level& L = ctx->levels[0];
  int rc;
  if ((rc = halo_exchange_n(ctx, L, {{ctx->u, 2}, {ctx->v, 2}, {ctx->gpy, 1}}))) return rc;
  const StepConst sc{ctx->dt, ctx->re};
  if (ctx->rect && ctx->rhs_march && L.g.nx >= 2) {   // rectangles: marching kernel, every slope and face flux once
    int ri; size_t sh;
    const dim3 grid = rhs_grid(L, L.g.nx, &ri, &sh);
    k_rhs_march<<<grid, RHS_THREADS, sh, ctx->stream>>>(L.g, ctx->d_edges, sc, ctx->u, ctx->v, ctx->conv_u, ctx->conv_v, ctx->gpx, ctx->gpy,
                                                        ctx->rhs_u, ctx->rhs_v, ri, ctx->red, ctx->fuse_bn2 ? ctx->d_scal + 30 : nullptr, 0, L.g.nx, ctx->d_pt);
    NS_LAUNCH_CHECK();
    if (ctx->fuse_bn2) {
      if (!ctx->d_pt && (rc = allreduce(ctx, ctx->d_scal + 30, 2, ncclSum))) return rc;
      ctx->bn2_ready[0] = ctx->bn2_ready[1] = true;
    }
    return 0;
  }
  ctx->bn2_ready[0] = ctx->bn2_ready[1] = false;
  const dim3 grid((L.g.ny + 127) / 128, L.g.nx);
#define CALL(RECT) k_rhs_build<RECT><<<grid, 128, 0, ctx->stream>>>(L.g, ctx->d_edges, sc, ctx->u, ctx->v, ctx->conv_u, ctx->conv_v, ctx->gpx, ctx->gpy, ctx->rhs_u, ctx->rhs_v)
  NS_DISPATCH_RECT(ctx->rect, CALL);
#undef CALL
  NS_LAUNCH_CHECK();
  return 0;
}

// ConstructRHS_V on the columns [c0, c1) only (one rank, rectangle): nsgpu_step_host runs it chunk by chunk behind the uploads
static int phase_rhs_cols(nsgpu_ctx* ctx, int c0, int c1) {
  const Level& L = ctx->levels[0];
  const StepConst sc{ctx->dt, ctx->re};
  int ri; size_t sh;
  const dim3 grid = rhs_grid(L, c1 - c0, &ri, &sh);
  k_rhs_march<<<grid, RHS_THREADS, sh, ctx->stream>>>(L.g, ctx->d_edges, sc, ctx->u, ctx->v, ctx->conv_u, ctx->conv_v, ctx->gpx, ctx->gpy,
                                                      ctx->rhs_u, ctx->rhs_v, ri, ctx->red, nullptr, c0, c1, nullptr);
  NS_LAUNCH_CHECK();
  return 0;
}

static int phase_velocity(nsgpu_ctx* ctx, SolveResult* ru, SolveResult* rv) {   // FS.cpp:547-548
  NvtxRange nvtx_("nsgpu:KSPSolve_u_v");
  ctx->halo_w.clear();   // every field may have been rewritten by a non-marching kernel since
  const Level& L = ctx->levels[0];
  // initial guess u^n (the reference starts from zero; the converged u* is the same to tolerance)
  int rc;
  if ((rc = halo_exchange_n(ctx, L, {{ctx->rhs_u, HX - 1}, {ctx->rhs_v, HX - 1}}))) return rc;   // the fused passes recompute up to 4 halo columns
  double *gu = ctx->vel_guess_rhs ? ctx->rhs_u : ctx->u, *gv = ctx->vel_guess_rhs ? ctx->rhs_v : ctx->v;
  if ((rc = solve(ctx, 0, coef_v(ctx), &ctx->ustar, ctx->rhs_u, ctx->rtol_u, ru, gu))) return rc;
  if ((rc = solve(ctx, 1, coef_v(ctx), &ctx->vstar, ctx->rhs_v, ctx->rtol_u, rv, gv))) return rc;
  return 0;
}

static int phase_div(nsgpu_ctx* ctx, bool remove_mean) {   // FS.cpp:549-550
  NvtxRange nvtx_("nsgpu:ConstructRHS_phi");
  ctx->halo_w.clear();   // every field may have been rewritten by a non-marching kernel since
  const Level& L = ctx->levels[0];
  int rc;
  if ((rc = halo_exchange_n(ctx, L, {{ctx->ustar, 1}, {ctx->vstar, 1}}))) return rc;
  const StepConst sc{ctx->dt, ctx->re};
  const dim3 grid = grid_cols(L.g, 256);
  if (ctx->rect && ctx->phys_march) {
    int ri; size_t sh;
    const dim3 pg = phys_grid((const void*)k_div_march, 2, L, &ri, &sh);
    k_div_march<<<pg, PH_THREADS, sh, ctx->stream>>>(L.g, ctx->d_edges, sc, ctx->ustar, ctx->vstar, ctx->rhs_phi, ri, ctx->red, ctx->d_scal + 28, ctx->d_pt);
  } else {
#define CALL(RECT) k_div_rhs<RECT><<<grid, 256, 0, ctx->stream>>>(L.g, ctx->d_edges, sc, ctx->ustar, ctx->vstar, ctx->rhs_phi, ctx->red, ctx->d_scal + 28)
    NS_DISPATCH_RECT(ctx->rect, CALL);
#undef CALL
  }
  NS_LAUNCH_CHECK();
  if (!(ctx->rect && ctx->phys_march && ctx->d_pt) && (rc = allreduce(ctx, ctx->d_scal + 28, 2, ncclSum))) return rc;   // (k_div_march sums over the ranks itself)
  k_bn2_from_sums<<<1, 1, 0, ctx->stream>>>(ctx->d_scal + 32, ctx->d_scal + 28, remove_mean ? 1.0 / (double)ctx->n_cells_global : 0.0);
  NS_LAUNCH_CHECK();
  ctx->bn2_ready[2] = ctx->fuse_bn2 != 0;
  if (remove_mean) {
    const double inv_n = 1.0 / (double)ctx->n_cells_global;
#define CALL(RECT) k_sub_mean<RECT><<<grid, 256, 0, ctx->stream>>>(L.g, ctx->rhs_phi, ctx->d_scal + 28, inv_n)
    NS_DISPATCH_RECT(ctx->rect, CALL);
#undef CALL
    NS_LAUNCH_CHECK();
  }
  return 0;
}

static int phase_phi(nsgpu_ctx* ctx, SolveResult* rp) {   // FS.cpp:551, warm start from phi^n (FS.cpp:54)
  NvtxRange nvtx_("nsgpu:KSPSolve_phi");
  if (ctx->phi_shift && !ctx->phi_sum_known) { int rcn = phi_normalize(ctx); if (rcn) return rcn; }   // two phi solves without a projection in between
  ctx->halo_w.clear();   // every field may have been rewritten by a non-marching kernel since
  const Level& L = ctx->levels[0];
  int rc;
  const double* b = ctx->rhs_phi;
  if (ctx->aug_rhs) {
    // stretched grid, no outflow edge: b'' = b - c w with c = (vol, b) / (vol, w) is the right-hand side whose solution the reference's
    // KSP converges to (ns_hostgeom.h); rhs_phi itself stays what ConstructRHS_phi + MatNullSpaceRemove made it
    const dim3 grid = grid_cols(L.g, 256);
#define CALL(RECT) k_vol_dot<RECT><<<grid, 256, 0, ctx->stream>>>(L.g, ctx->rhs_phi, ctx->red, ctx->d_scal + 41)
    NS_DISPATCH_RECT(ctx->rect, CALL);
#undef CALL
    NS_LAUNCH_CHECK();
    if ((rc = allreduce(ctx, ctx->d_scal + 41, 1, ncclSum))) return rc;
    cudaError_t e = cudaMemcpyAsync(ctx->aug_t, ctx->rhs_phi, L.stride * sizeof(double), cudaMemcpyDeviceToDevice, ctx->stream);
    if (e != cudaSuccess) { ctx->err = cudaGetErrorString(e); return NSGPU_ERR_CUDA; }
    const size_t n = (size_t)L.g.nx * L.g.ld, off = (size_t)HX * L.g.ld;
    const int blocks = (int)std::min<size_t>((n + 255) / 256, 148 * 8);
    k_aug_axpy<<<blocks, 256, 0, ctx->stream>>>(n, ctx->aug_t + off, ctx->aug_w + off, ctx->d_scal + 41, -1.0 / ctx->aug_yw, nullptr, ctx->red, ctx->d_scal + 44);
    NS_LAUNCH_CHECK();
    ctx->halo_w.erase(ctx->aug_t);
    ctx->bn2_ready[2] = false;   // |b''|^2 is reduced by the solve head
    b = ctx->aug_t;
  }
  if ((rc = halo_exchange(ctx, L, const_cast<double*>(b), HX - 1))) return rc;
  if ((rc = solve(ctx, 2, coef_phi(), &ctx->phi, b, ctx->rtol_phi, rp))) return rc;
  if (!rp->pending) ctx->fine_sweeps_phi += (double)rp->cycles * (ctx->nu_pre + ctx->nu_post) * (ctx->solver == 1 ? 2 : 1);
  // pin mean(phi) = 0: the reference projects the constant out of every Krylov update (MatSetNullSpace, FS.cpp:143)
  const bool lazy = rp->pending && ctx->lazy_mean && ctx->nu_pre >= 1 && ctx->levels.size() > 1 && ctx->rect && ctx->phys_march;   // the next solve's head starts with a restricting pass (solve_head), k_correct_march reduces the sum
  if (lazy) { ctx->phi_shift = true; ctx->phi_sum_known = false; }   // sum(phi) is in the solve's control block; removed by the next solve's first pass, or by phi_normalize()
  else {
    const size_t n = (size_t)L.g.nx * L.g.ld;
    const int blocks = (int)std::min<size_t>((n + 255) / 256, 148 * 8);
    ctx->halo_w.clear();
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
  NvtxRange nvtx_("nsgpu:CorrectVelocities");
  ctx->halo_w.clear();   // every field may have been rewritten by a non-marching kernel since
  const Level& L = ctx->levels[0];
  int rc;
  if ((rc = halo_exchange(ctx, L, ctx->phi, 2))) return rc;
  const StepConst sc{ctx->dt, ctx->re};
  if (ctx->rect && ctx->phys_march) {
    int ri; size_t sh;
    const dim3 pg = phys_grid((const void*)k_correct_march, 3, L, &ri, &sh);
    // sum(phi) rides along when phi's mean is pending; with the peer map the four maxima and the sum are combined over the ranks in the
    // kernel's epilogue (no NCCL call in this phase)
    k_correct_march<<<pg, PH_THREADS, sh, ctx->stream>>>(L.g, sc, ctx->phi, ctx->ustar, ctx->vstar, ctx->u, ctx->v, ctx->gpx, ctx->gpy, ri, ctx->red,
                                                         ctx->d_scal + 20, ctx->d_pt, ctx->phi_shift ? &ctx->d_ctl[2].xsum : nullptr);
    NS_LAUNCH_CHECK();
    if (ctx->phi_shift) {
      ctx->phi_sum_known = true;
      if (!ctx->d_pt && (rc = allreduce(ctx, &ctx->d_ctl[2].xsum, 1, ncclSum))) return rc;
    }
    if (!ctx->d_pt && (rc = allreduce(ctx, ctx->d_scal + 20, 4, ncclMax))) return rc;
    if (mm) {
      double h[4];
      if ((rc = read_scalars(ctx, ctx->d_scal + 20, 4, h))) return rc;
      mm[0] = -h[0]; mm[1] = h[1]; mm[2] = -h[2]; mm[3] = h[3];
    }
    return 0;
  } else {
    const dim3 grid = grid_cols(L.g, 256);
#define CALL(RECT) k_correct<RECT><<<grid, 256, 0, ctx->stream>>>(L.g, sc, ctx->phi, ctx->ustar, ctx->vstar, ctx->u, ctx->v, ctx->gpx, ctx->gpy, ctx->red, ctx->d_scal + 20)
    NS_DISPATCH_RECT(ctx->rect, CALL);
#undef CALL
  }
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
