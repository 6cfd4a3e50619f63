// ns_setup.inl -- nsgpu_create / nsgpu_destroy: slab descriptor, matrix-free operator coefficients and the
// multigrid hierarchy.  Stands in for SolverInitialize + SolverSetup (reference FS.cpp:16-58, 84-164).

namespace nsgpu {

static thread_local std::string g_create_error;

#define NS_CUDA(call)                                                                                   \
  do {                                                                                                  \
    cudaError_t e_ = (call);                                                                            \
    if (e_ != cudaSuccess) {                                                                            \
      char buf_[512];                                                                                   \
      snprintf(buf_, sizeof buf_, "%s:%d: %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_)); \
      ctx->err = buf_;                                                                                  \
      return NSGPU_ERR_CUDA;                                                                            \
    }                                                                                                   \
  } while (0)

#define NS_NCCL(call)                                                                                   \
  do {                                                                                                  \
    ncclResult_t e_ = (call);                                                                           \
    if (e_ != ncclSuccess) {                                                                            \
      char buf_[512];                                                                                   \
      snprintf(buf_, sizeof buf_, "%s:%d: %s -> %s", __FILE__, __LINE__, #call, ncclGetErrorString(e_)); \
      ctx->err = buf_;                                                                                  \
      return NSGPU_ERR_NCCL;                                                                            \
    }                                                                                                   \
  } while (0)

template <class T>
static int dev_upload(nsgpu_ctx* ctx, std::vector<void*>& owned, const std::vector<T>& h, const T** out) {
  void* p = nullptr;
  NS_CUDA(cudaMalloc(&p, std::max<size_t>(h.size() * sizeof(T), 16)));
  owned.push_back(p);
  if (!h.empty()) NS_CUDA(cudaMemcpy(p, h.data(), h.size() * sizeof(T), cudaMemcpyHostToDevice));
  *out = (const T*)p;
  return 0;
}

static int dev_zeros(nsgpu_ctx* ctx, std::vector<void*>& owned, size_t n_doubles, double** out) {
  const size_t bytes = (n_doubles * sizeof(double) + 255) & ~(size_t)255;
  // Large arrays do not all start on the same 2 MB boundary (cudaMalloc's granularity) / at multiples of a column: the k-th one is shifted by
  // k x field_skew bytes, so that the streams a pass reads and writes at the same column do not meet in the same memory channels
  const size_t skew = (ctx->field_skew && bytes >= ((size_t)1 << 20)) ? ((size_t)ctx->field_skew * (size_t)(ctx->alloc_k++)) % ((size_t)1 << 21) : 0;
  if (ctx->arena && ctx->arena_used + skew + bytes <= ctx->arena_bytes) {   // peer-visible arena (already zeroed)
    ctx->arena_used += skew;
    *out = (double*)(ctx->arena + ctx->arena_used);
    ctx->arena_used += bytes;
    return 0;
  }
  void* p = nullptr;
  NS_CUDA(cudaMalloc(&p, n_doubles * sizeof(double) + skew));
  owned.push_back(p);
  NS_CUDA(cudaMemset(p, 0, n_doubles * sizeof(double) + skew));
  *out = (double*)((char*)p + skew);
  return 0;
}

// Device copy of the tables of one level.
static int build_level_geom(nsgpu_ctx* ctx, Level& L, const HostPlan& P, size_t l) {
  GeomArrays A;
  make_geom_arrays(P.levels[l], P.lgi0[l], P.lnx[l], (l == 0 && !P.rect) ? &P.cedges0 : nullptr, A);
  make_march_coefs(P.levels[l], P, ctx->dt, ctx->re, A);
  LevelGeom& g = L.g;
  geom_scalars(A, P, g);
  L.gny = A.ny;
  L.stride = A.stride;
  int rc;
  if ((rc = dev_upload(ctx, L.owned, A.hx, &g.hx))) return rc;
  if ((rc = dev_upload(ctx, L.owned, A.hy, &g.hy))) return rc;
  if ((rc = dev_upload(ctx, L.owned, A.rhx, &g.rhx))) return rc;
  if ((rc = dev_upload(ctx, L.owned, A.rhy, &g.rhy))) return rc;
  if ((rc = dev_upload(ctx, L.owned, A.rsx, &g.rsx))) return rc;
  if ((rc = dev_upload(ctx, L.owned, A.rsy, &g.rsy))) return rc;
  if ((rc = dev_upload(ctx, L.owned, A.ax, &g.ax))) return rc;
  if ((rc = dev_upload(ctx, L.owned, A.ay, &g.ay))) return rc;
  if ((rc = dev_upload(ctx, L.owned, A.bx, &g.bx))) return rc;
  if ((rc = dev_upload(ctx, L.owned, A.by, &g.by))) return rc;
  if (l == 0 && !A.mcx[0].empty()) {
    // Jacobi spectral-radius bound of LHS_V, max over cells of sum|off-diagonal| / diagonal (distinct columns x distinct rows)
    std::vector<std::pair<double, double>> ux, uy;
    for (int a = HX; a < HX + A.nx; a++) ux.push_back({std::fabs(A.mcx[0][3 * a]) + std::fabs(A.mcx[0][3 * a + 1]), A.mcx[0][3 * a + 2]});
    for (int a = PJ; a < PJ + A.ny; a++) uy.push_back({std::fabs(A.mcy[0][3 * a]) + std::fabs(A.mcy[0][3 * a + 1]), A.mcy[0][3 * a + 2]});
    std::sort(ux.begin(), ux.end()); ux.erase(std::unique(ux.begin(), ux.end()), ux.end());
    std::sort(uy.begin(), uy.end()); uy.erase(std::unique(uy.begin(), uy.end()), uy.end());
    double rho = 0;
    for (auto& p : ux) for (auto& q : uy) rho = std::max(rho, (p.first + q.first) / (p.second + q.second));
    ctx->vel_rho = rho;
  }
  for (int op = 0; op < 2; op++) {
    L.march_ok[op] = A.march_ok[op];
    if (!A.mcx[op].empty()) {
      if ((rc = dev_upload(ctx, L.owned, A.mcx[op], &L.mcx[op]))) return rc;
      if ((rc = dev_upload(ctx, L.owned, A.mcy[op], &L.mcy[op]))) return rc;
    }
  }
  g.flags = nullptr;
  g.cedges = nullptr;
  if (!A.flags.empty() && (rc = dev_upload(ctx, L.owned, A.flags, &g.flags))) return rc;
  if (!A.cedges.empty() && (rc = dev_upload(ctx, L.owned, A.cedges, &g.cedges))) return rc;
  return 0;
}

static void destroy_ctx(nsgpu_ctx* ctx) {
  if (!ctx) return;
  cudaSetDevice(ctx->device);
  if (ctx->stream) cudaStreamSynchronize(ctx->stream);
  for (int r = 0; r < MAX_RANKS; r++) if (ctx->peers[r] && !ctx->peer_inproc[r]) cudaIpcCloseMemHandle(ctx->peers[r]);
  if (ctx->comm) ncclCommDestroy(ctx->comm);
  if (ctx->arena) cudaFree(ctx->arena);
  for (auto& L : ctx->levels)
    for (void* p : L.owned) cudaFree(p);
  for (void* p : ctx->owned) cudaFree(p);
  for (auto& sg : ctx->sgraph) { if (sg.exec) cudaGraphExecDestroy(sg.exec); if (sg.graph) cudaGraphDestroy(sg.graph); }
  for (auto& sg : ctx->kgraph) { if (sg.exec) cudaGraphExecDestroy(sg.exec); if (sg.graph) cudaGraphDestroy(sg.graph); }
  if (ctx->s_in) { cudaStreamDestroy(ctx->s_in); cudaStreamDestroy(ctx->s_out); for (auto& e : ctx->ev_io) if (e) cudaEventDestroy(e); }
  for (auto& e : ctx->ev_chunk) if (e) cudaEventDestroy(e);
  for (auto& q : ctx->exp_host) if (q) cudaFreeHost(q);
  if (ctx->s_exp) cudaStreamDestroy(ctx->s_exp);
  if (ctx->exp_snap) cudaEventDestroy(ctx->exp_snap);
  for (auto& e : ctx->exp_home) if (e) cudaEventDestroy(e);
  for (auto& e : ctx->exp_t0) if (e) cudaEventDestroy(e);
  for (auto& e : ctx->exp_t1) if (e) cudaEventDestroy(e);
  if (ctx->h_scal) cudaFreeHost(ctx->h_scal);
  if (ctx->h_ctl) cudaFreeHost(ctx->h_ctl);
  if (ctx->h_stage) cudaFreeHost(ctx->h_stage);
  if (ctx->ev0) cudaEventDestroy(ctx->ev0);
  if (ctx->ev1) cudaEventDestroy(ctx->ev1);
  for (auto& e : ctx->evp) if (e) cudaEventDestroy(e);
  if (ctx->stream) cudaStreamDestroy(ctx->stream);
  delete ctx;
}

static int create_impl(nsgpu_ctx* ctx, const nsgpu_grid* G, const nsgpu_params* P) {
  if (!G || !P) { ctx->err = "null grid or params"; return NSGPU_ERR_ARG; }
  if (!(P->dt > 0) || !(P->re > 0)) { ctx->err = "dt and re must be positive"; return NSGPU_ERR_ARG; }   // FS.cpp:17-28
  ctx->dt = P->dt; ctx->re = P->re;
  ctx->device = P->device; ctx->rank = P->rank; ctx->world = P->world > 0 ? P->world : 1;
  ctx->rtol_u = P->rtol_u > 0 ? P->rtol_u : 1e-8;
  ctx->rtol_phi = P->rtol_phi > 0 ? P->rtol_phi : 1e-8;
  ctx->max_cycles = P->max_cycles > 0 ? P->max_cycles : 200;
  ctx->nu_pre = P->nu_pre > 0 ? P->nu_pre : 2;
  ctx->nu_post = P->nu_post > 0 ? P->nu_post : 2;
  ctx->solver = P->solver;
  ctx->use_graph = P->use_graph < 0 ? 0 : 1;
  if (const char* e = getenv("NSGPU_MARCH")) ctx->march = atoi(e);
  if (const char* e = getenv("NSGPU_RI")) ctx->march_ri = atoi(e);
  if (const char* e = getenv("NSGPU_PF")) ctx->march_pf = atoi(e);
  if (const char* e = getenv("NSGPU_WARPS_PER_SM")) ctx->march_wps = std::max(1, atoi(e));
  if (const char* e = getenv("NSGPU_GRAPH")) ctx->use_graph = atoi(e);
  HostPlan plan;
  {
    long long rep_cells = 262144;
    if (const char* e = getenv("NSGPU_REP_CELLS")) rep_cells = atoll(e);
    int rc = make_host_plan(G, ctx->rank, ctx->world, P->max_levels, plan, ctx->err, rep_cells);
    if (rc) return rc;
  }
  ctx->gnx = G->nx; ctx->ny = G->ny;
  ctx->rect = plan.rect;
  ctx->n_cells_global = plan.n_cells_global;
  ctx->i0 = plan.i0; ctx->i1 = plan.i1; ctx->id0 = plan.id0; ctx->n_local = plan.n_local;
  ctx->local_id = plan.local_id;

  // ---- device
  int ndev = 0;
  NS_CUDA(cudaGetDeviceCount(&ndev));
  if (ctx->device < 0 || ctx->device >= ndev) { ctx->err = "no such CUDA device"; return NSGPU_ERR_CUDA; }
  NS_CUDA(cudaSetDevice(ctx->device));
  NS_CUDA(cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking));
  NS_CUDA(cudaEventCreate(&ctx->ev0));
  NS_CUDA(cudaEventCreate(&ctx->ev1));
  for (auto& e : ctx->evp) NS_CUDA(cudaEventCreate(&e));
  if (ctx->world > 1) {
    if (!P->nccl_id) { ctx->err = "world > 1 needs params.nccl_id"; return NSGPU_ERR_ARG; }
    ncclUniqueId id;
    memcpy(&id, P->nccl_id, sizeof id);
    NS_NCCL(ncclCommInitRank(&ctx->comm, ctx->world, id, ctx->rank));
  }

  // ---- several ranks: one arena for everything the marching passes write, opened by both neighbours (CUDA IPC)
  if (const char* e = getenv("NSGPU_P2P")) ctx->use_p2p = atoi(e);
  if (const char* e = getenv("NSGPU_TRACE")) ctx->trace = atoi(e);
  ctx->march_rb = -1;   // several ranks: boundary chunks (pushing loop) sized automatically; NSGPU_RB=0: uniform chunks, > 0: that width
  if (const char* e = getenv("NSGPU_RB")) ctx->march_rb = atoi(e);
  if (const char* e = getenv("NSGPU_XRED_DBG")) { const int v = atoi(e); cudaMemcpyToSymbol(g_ns_xdbg, &v, sizeof v); }   // 1: reduction tails, 2: CTA stamps
  if (const char* e = getenv("NSGPU_MARCH_RING")) ctx->march_ring = atoi(e);
  if (const char* e = getenv("NSGPU_SKIP_CHECKS")) ctx->skip_checks = atoi(e);
  if (const char* e = getenv("NSGPU_STAMP_KEY")) ctx->stamp_key = atoi(e);
  else if (getenv("NSGPU_XRED_DBG") && (atoi(getenv("NSGPU_XRED_DBG")) & 2)) ctx->stamp_key = -2;
  if (const char* e = getenv("NSGPU_FIELD_SKEW")) ctx->field_skew = (atoi(e) + 255) & ~255;
  if (const char* e = getenv("NSGPU_RB_COST")) ctx->rb_cost = ctx->rb_cost_up = std::max(1.0, atof(e));
  if (const char* e = getenv("NSGPU_RB_COST_UP")) ctx->rb_cost_up = std::max(1.0, atof(e));
  if (const char* e = getenv("NSGPU_RHS_MARCH")) ctx->rhs_march = atoi(e);
  if (const char* e = getenv("NSGPU_PHYS_MARCH")) ctx->phys_march = atoi(e);
  if (const char* e = getenv("NSGPU_LAZY_MEAN")) ctx->lazy_mean = atoi(e);
  if (const char* e = getenv("NSGPU_FUSE_BN2")) ctx->fuse_bn2 = atoi(e);
  if (const char* e = getenv("NSGPU_VEL_GUESS")) ctx->vel_guess_rhs = atoi(e);
  if (const char* e = getenv("NSGPU_IO_CHUNKS")) ctx->io_chunks = std::max(1, atoi(e));
  if (getenv("NSGPU_TAIL_DBG")) { void* q = nullptr; if (cudaMalloc(&q, 64 * sizeof(long long)) == cudaSuccess) { cudaMemset(q, 0, 64 * sizeof(long long)); ctx->owned.push_back(q); ctx->tail_dbg = (long long*)q; } }
  if (const char* e = getenv("NSGPU_RB_MINNX")) ctx->march_rb_min_nx = atoi(e);
  if (ctx->world > 1 && ctx->world <= MAX_RANKS && ctx->use_p2p && plan.rect && plan.levels[0].gnx % ctx->world == 0) {   // equal slabs: equal arena layouts
    // Everything a marching pass may write has to lie inside the arena (the pass pushes `xout + neighbour offset`): the 13 state
    // fields, and on the Krylov routes (solver == 1, outflow / stretched / uncoarsenable grids -- decided below from the same plan)
    // the 8 BiCGStab vectors plus aug_w, aug_t.  march_pass checks the range again before it enables the push.
    int rhs_mode_a = 0;
    if (const char* e = getenv("NSGPU_RHS_CONSISTENT")) rhs_mode_a = atoi(e);
    const bool need_aug = plan_has_outflow(plan) || (rhs_mode_a >= 0 && !plan_is_uniform(plan));
    const bool need_kry = ctx->solver == 1 || need_aug || plan.levels.size() == 1;
    const size_t fine_fields = 13 + (need_kry ? 8 : 0) + (need_aug ? 2 : 0);
    size_t doubles = 0;
    for (size_t l = 0; l < plan.levels.size(); l++) {
      const size_t ld = ((plan.levels[l].ny + 3) / 4) * 4 + 2 * PJ;
      const size_t stride = (size_t)(plan.lnx[l] + 2 * HX + SLACK_COLS) * ld + 32;
      doubles += (l == 0 ? fine_fields : 3) * stride + 64;   // + 256-byte rounding of every allocation
    }
    ctx->arena_bytes = doubles * sizeof(double) + (1 << 20);
    NS_CUDA(cudaMalloc((void**)&ctx->arena, ctx->arena_bytes));
    NS_CUDA(cudaMemset(ctx->arena, 0, ctx->arena_bytes));
    ctx->arena_used = HDR_WORDS * sizeof(unsigned long long);   // control header: arrival / sent counters, reduction slots
    // exchange the IPC handles through the communicator that exists anyway.  Ranks of ONE process (FluidSolver with NSGPU_GPUS=N: one
    // host thread per GPU) cannot open each other's IPC handles; they pass the arena's address itself and enable peer access.
    struct PeerInfo { cudaIpcMemHandle_t h; long long pid; unsigned long long ptr; int dev, pad; };
    static_assert(sizeof(cudaIpcMemHandle_t) == 64 && sizeof(PeerInfo) == 88, "peer record size");
    PeerInfo mine{};
    NS_CUDA(cudaIpcGetMemHandle(&mine.h, ctx->arena));
    mine.pid = (long long)getpid(); mine.ptr = (unsigned long long)(uintptr_t)ctx->arena; mine.dev = ctx->device;
    char* d_h = nullptr;
    NS_CUDA(cudaMalloc((void**)&d_h, sizeof(PeerInfo) * (size_t)ctx->world));
    NS_CUDA(cudaMemcpy(d_h + sizeof(PeerInfo) * (size_t)ctx->rank, &mine, sizeof mine, cudaMemcpyHostToDevice));
    NS_NCCL(ncclAllGather(d_h + sizeof(PeerInfo) * (size_t)ctx->rank, d_h, sizeof(PeerInfo), ncclChar, ctx->comm, ctx->stream));
    NS_CUDA(cudaStreamSynchronize(ctx->stream));
    std::vector<PeerInfo> all(ctx->world);
    NS_CUDA(cudaMemcpy(all.data(), d_h, sizeof(PeerInfo) * (size_t)ctx->world, cudaMemcpyDeviceToHost));
    cudaFree(d_h);
    bool ok = true;
    for (int r = 0; r < ctx->world; r++) {
      if (r == ctx->rank) continue;
      if (all[r].pid == mine.pid) {   // same process: direct peer access
        int can = 0;
        cudaError_t e = all[r].dev == ctx->device ? cudaSuccess : cudaDeviceCanAccessPeer(&can, ctx->device, all[r].dev);
        if (e == cudaSuccess && all[r].dev != ctx->device && can) {
          e = cudaDeviceEnablePeerAccess(all[r].dev, 0);
          if (e == cudaErrorPeerAccessAlreadyEnabled) { e = cudaSuccess; cudaGetLastError(); }
        } else if (all[r].dev != ctx->device) e = cudaErrorInvalidDevice;
        if (e == cudaSuccess) { ctx->peers[r] = (char*)(uintptr_t)all[r].ptr; ctx->peer_inproc[r] = true; }
        else { cudaGetLastError(); ok = false; }
      } else if (cudaIpcOpenMemHandle((void**)&ctx->peers[r], all[r].h, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { ctx->peers[r] = nullptr; ok = false; }
    }
    if (ok) { ctx->peerL = ctx->rank > 0 ? ctx->peers[ctx->rank - 1] : nullptr; ctx->peerR = ctx->rank < ctx->world - 1 ? ctx->peers[ctx->rank + 1] : nullptr; }
    // every rank must agree, or the halo protocols would differ: all-reduce the outcome
    int* d_ok = nullptr;
    NS_CUDA(cudaMalloc((void**)&d_ok, sizeof(int)));
    const int okv = ok ? 1 : 0;
    NS_CUDA(cudaMemcpy(d_ok, &okv, sizeof(int), cudaMemcpyHostToDevice));
    NS_NCCL(ncclAllReduce(d_ok, d_ok, 1, ncclInt, ncclMin, ctx->comm, ctx->stream));
    NS_CUDA(cudaStreamSynchronize(ctx->stream));
    int all_ok = 0;
    NS_CUDA(cudaMemcpy(&all_ok, d_ok, sizeof(int), cudaMemcpyDeviceToHost));
    cudaFree(d_ok);
    cudaGetLastError();
    if (!all_ok) {
      fprintf(stderr, "nsgpu: CUDA IPC mapping of the neighbour slabs failed; halo columns go through NCCL\n");
      for (int r = 0; r < MAX_RANKS; r++) if (ctx->peers[r]) { if (!ctx->peer_inproc[r]) cudaIpcCloseMemHandle(ctx->peers[r]); ctx->peers[r] = nullptr; }
      ctx->peerL = ctx->peerR = nullptr;
      ctx->use_p2p = 0;
    }
  } else ctx->use_p2p = 0;

  // ---- level hierarchy
  ctx->levels.resize(plan.levels.size());
  for (size_t l = 0; l < plan.levels.size(); l++) {
    Level& L = ctx->levels[l];
    L.cx = plan.lcx[l]; L.cy = plan.lcy[l];
    L.replicated = plan.rep_start > 0 && (int)l >= plan.rep_start;
    int rc = build_level_geom(ctx, L, plan, l);
    if (rc) return rc;
    if (l > 0) {
      if ((rc = dev_zeros(ctx, L.owned, L.stride, &L.x))) return rc;
      if ((rc = dev_zeros(ctx, L.owned, L.stride, &L.xt))) return rc;
      if ((rc = dev_zeros(ctx, L.owned, L.stride, &L.b))) return rc;
    }
  }
  ctx->levels.back().coarse_sweeps = coarsest_sweeps(ctx->levels.back().g.gnx, ctx->levels.back().gny);

  // ---- velocity solves: LHS_V = I - dt/(2Re) L is diagonally dominant; when red-black SOR with the optimal factor
  // reaches rtol in a few sweeps it beats multigrid (no coarse levels, every pass is a fused double sweep)
  ctx->vel_omega = 0.0;
  {
    int mode = 2;   // 0 = multigrid, 1 = SOR, 2 = auto
    if (const char* e = getenv("NSGPU_VEL_SOLVER")) mode = !strcmp(e, "mg") ? 0 : !strcmp(e, "sor") ? 1 : 2;
    if (P->vel_solver == 1) mode = 0; else if (P->vel_solver == 2) mode = 1;
    const double rho = ctx->vel_rho;
    if (mode != 0 && ctx->rect && ctx->solver == 0 && rho > 0 && rho < 1) {
      const double omega = 2.0 / (1.0 + std::sqrt(1.0 - rho * rho));
      const double sweeps = std::log(ctx->rtol_u) / std::log(std::max(omega - 1.0, 1e-3));
      if (mode == 1 || sweeps <= 28.0) ctx->vel_omega = omega;
    }
  }

  // ---- coarse tail: every level of at most TAIL_MAX_CELLS cells (never level 0) in one single-CTA kernel
  if (const char* e = getenv("NSGPU_TAIL")) ctx->use_tail = atoi(e);
  ctx->tail_start = 0;
  if ((ctx->use_tail || plan.rep_start > 0) && (ctx->world == 1 || plan.rep_start > 0) && ctx->rect) {
    const int nl = (int)ctx->levels.size();
    int ts = 0;
    for (int l = 1; l < nl; l++)
      if ((long long)ctx->levels[l].g.gnx * ctx->levels[l].g.ny <= TAIL_MAX_CELLS) { ts = l; break; }
    if (plan.rep_start > 0 && (ts == 0 || ts < plan.rep_start)) { ctx->err = "internal: the coarse tail starts on a distributed level"; return NSGPU_ERR_ARG; }
    if (ts > 0 && nl - ts <= TAIL_MAX_LEVELS) {
      for (int op = 0; op < 2; op++) {
        TailDesc& T = ctx->tail[op];
        T.nlev = nl - ts; T.nu_pre = ctx->nu_pre; T.nu_post = ctx->nu_post;
        T.coarse_sweeps = ctx->levels.back().coarse_sweeps; T.gi0 = 0;
        // The coarsest grid of a long domain is long too (y cannot be coarsened any further): plain Gauss-Seidel would need
        // O(n^2) sweeps there.  Optimal SOR for the Neumann problem on n cells, Jacobi radius cos(pi / n).
        T.coarse_omega = 1.0;
        if (op == 1) {
          const int n = std::max(ctx->levels.back().g.gnx, ctx->levels.back().g.ny);
          if (n >= 4) T.coarse_omega = 2.0 / (1.0 + std::sin(3.14159265358979323846 / n));
          if (const char* e = getenv("NSGPU_COARSE_OMEGA")) T.coarse_omega = atof(e);
        }
        int off = 0;
        for (int k = 0; k < T.nlev; k++) {
          const Level& L = ctx->levels[ts + k];
          TailLevel& t = T.lv[k];
          t.nx = L.g.nx; t.ny = L.g.ny; t.cx = L.cx; t.cy = L.cy; t.ld = L.g.ld;
          t.mcx = L.mcx[op]; t.mcy = L.mcy[op];
          const int cells = (t.nx + 2) * (t.ny + 2);
          t.xoff = off; off += cells;
          t.boff = off; off += cells;
          t.rdoff = off; off += cells;
          t.sxoff = off; off += 3 * t.nx;
          t.syoff = off; off += 3 * t.ny;
        }
        T.smem_doubles = off + (off & 1);   // even: the image is copied as double2
      }
      const size_t bytes = (size_t)ctx->tail[0].smem_doubles * sizeof(double);
      if (plan.rep_start > 0 && (bytes > 200 * 1024 || nl - ts > TAIL_MAX_LEVELS)) { ctx->err = "internal: coarse tail does not fit"; return NSGPU_ERR_ARG; }
      if (bytes <= 200 * 1024) {
        NS_CUDA(cudaFuncSetAttribute(k_coarse_tail, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
        for (int op = 0; op < 2; op++) {
          void* q = nullptr;
          NS_CUDA(cudaMalloc(&q, bytes));
          ctx->owned.push_back(q);
          ctx->tail_img[op] = (double*)q;
          k_tail_image<<<1, TAIL_THREADS, 0, ctx->stream>>>(ctx->tail[op], ctx->tail_img[op]);
          NS_CUDA(cudaGetLastError());
        }
        ctx->tail_start = ts;
      }
    }
  }

  // ---- cluster tail: every level from the first one of at most CT_FIRST_CELLS cells down, in one kernel on a cluster of CT_C CTAs
  if (const char* e = getenv("NSGPU_CTAIL")) ctx->use_ctail = atoi(e);
  ctx->ctail_start = 0;
  if (ctx->use_ctail && ctx->rect && (ctx->world == 1 || plan.rep_start > 0)) {
    long long first_cells = CT_FIRST_CELLS;
    if (const char* e = getenv("NSGPU_CTAIL_CELLS")) first_cells = atoll(e);
    int C = CT_C;
    if (const char* e = getenv("NSGPU_CTAIL_C")) C = atoi(e) == 16 ? 16 : CT_C;
    ctx->ctail_c = C;
    const int nl = (int)ctx->levels.size();
    int ts = 0;
    for (int l = 1; l < nl; l++)
      if ((long long)ctx->levels[l].g.gnx * ctx->levels[l].g.ny <= first_cells) { ts = l; break; }
    bool ok = ts > 0 && nl - ts <= CT_MAX_LEVELS && (plan.rep_start == 0 || ts >= plan.rep_start);
    for (int op = 0; op < 2 && ok; op++) {
      CtDesc& T = ctx->ctail[op];
      T.nlev = nl - ts; T.nu_pre = ctx->nu_pre; T.nu_post = ctx->nu_post;
      T.coarse_sweeps = ctx->levels.back().coarse_sweeps;
      T.coarse_omega = 1.0;
      if (op == 1) {   // as for the one-CTA tail: optimal SOR for the Neumann problem on the long coarsest grid
        const int n = std::max(ctx->levels.back().g.gnx, ctx->levels.back().g.ny);
        if (n >= 4) T.coarse_omega = 2.0 / (1.0 + std::sin(3.14159265358979323846 / n));
        if (const char* e = getenv("NSGPU_COARSE_OMEGA")) T.coarse_omega = atof(e);
      }
      int off = 0;
      bool dist = true;
      for (int k = 0; k < T.nlev; k++) {
        const Level& L = ctx->levels[ts + k];
        CtLevel& t = T.lv[k];
        t.nx = L.g.nx; t.ny = L.g.ny; t.cx = L.cx; t.cy = L.cy; t.ld = L.g.ld;
        t.mcx = L.mcx[op]; t.mcy = L.mcy[op];
        // a level stays distributed while its columns split evenly and every coarse column has both children in one CTA
        const bool last = k == T.nlev - 1;
        dist = dist && t.nx % C == 0 && (last || t.cx == 1 || (t.nx / C) % 2 == 0);
        t.dist = dist ? 1 : 0;
        t.nc = dist ? t.nx / C : t.nx;
        if ((long long)t.nc * ((t.ny + 1) / 2) > (long long)CT_KC * CT_THREADS) ok = false;
        t.xoff = off; off += (t.nc + 2) * (t.ny + 2);
        t.boff = off; off += t.nc * t.ny;
        t.rdoff = off; off += t.nc * t.ny;
        t.sxoff = off; off += 3 * t.nc;
        t.syoff = off; off += 3 * t.ny;
      }
      if (!T.lv[0].dist) ok = false;
      T.smem_doubles = off + (off & 1);
      if ((size_t)T.smem_doubles * sizeof(double) > 200 * 1024) ok = false;
    }
    if (ok) {
      const size_t bytes = (size_t)ctx->ctail[0].smem_doubles * sizeof(double);
      ok = cudaFuncSetAttribute(k_cluster_tail, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes) == cudaSuccess;
      if (ok && C > 8) {   // a cluster of 16 CTAs is a non-portable size: allowed explicitly, and checked against what the GPU can co-schedule
        ok = cudaFuncSetAttribute(k_cluster_tail, cudaFuncAttributeNonPortableClusterSizeAllowed, 1) == cudaSuccess;
        if (ok) {
          cudaLaunchConfig_t cfg = {};
          cfg.gridDim = dim3(C); cfg.blockDim = dim3(CT_THREADS); cfg.dynamicSmemBytes = bytes;
          cudaLaunchAttribute at[1];
          at[0].id = cudaLaunchAttributeClusterDimension;
          at[0].val.clusterDim.x = C; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
          cfg.attrs = at; cfg.numAttrs = 1;
          int nclusters = 0;
          ok = cudaOccupancyMaxActiveClusters(&nclusters, k_cluster_tail, &cfg) == cudaSuccess && nclusters >= 1;
        }
        if (!ok) fprintf(stderr, "nsgpu: a 16-CTA cluster cannot be scheduled here; coarse tail without clusters\n");
        cudaGetLastError();
      }
      for (int op = 0; op < 2 && ok; op++) {
        void* q = nullptr;
        NS_CUDA(cudaMalloc(&q, bytes * C));
        ctx->owned.push_back(q);
        ctx->ctail_img[op] = (double*)q;
        k_ctail_image<<<C, CT_THREADS, 0, ctx->stream>>>(ctx->ctail[op], ctx->ctail_img[op]);
        NS_CUDA(cudaGetLastError());
      }
      cudaGetLastError();
      if (ok) ctx->ctail_start = ts;
    }
  }

  // ---- fields
  const size_t S = ctx->levels[0].stride;
  double** fields[] = {&ctx->u, &ctx->v, &ctx->phi, &ctx->conv_u, &ctx->conv_v, &ctx->gpx, &ctx->gpy,
                       &ctx->rhs_u, &ctx->rhs_v, &ctx->rhs_phi, &ctx->ustar, &ctx->vstar, &ctx->tmp};
  for (double** f : fields) { int rc = dev_zeros(ctx, ctx->owned, S, f); if (rc) return rc; }
  const bool outflow = plan_has_outflow(plan);
  // NSGPU_RHS_CONSISTENT: -1 treats a stretched grid like a uniform one (the pre-round-1 behaviour: the solve stagnates at the level of the
  // inconsistency); 0 (default): stretched grids take the Krylov solve on the augmented operator; 1: mildly stretched grids (<= 2) keep the
  // multigrid iteration, on b - c w; 2: every stretched grid does
  int rhs_mode = 0;
  if (const char* e = getenv("NSGPU_RHS_CONSISTENT")) rhs_mode = atoi(e);
  if (outflow || (rhs_mode >= 0 && !plan_is_uniform(plan))) {   // LHS_phi x = b is inconsistent there; see ns_hostgeom.h for what the reference returns
    std::vector<double> w;
    int rc = phi_ilu0_image_of_ones(G, plan, w, ctx->err);
    if (rc) return rc;
    const LevelGeom& g0 = ctx->levels[0].g;
    std::vector<double> wp(S, 0.0);
    for (int i = 0; i < g0.nx; i++)
      for (int j = 0; j < g0.ny; j++) {
        const int64_t id = plan.rect ? (int64_t)i * g0.ny + j : (int64_t)plan.local_id[(size_t)i * g0.ny + j];
        if (id >= 0) wp[cidx(g0, i, j)] = w[(size_t)(plan.id0 + id)];
      }
    for (double x : w) ctx->aug_ww += x * x;
    if ((rc = dev_zeros(ctx, ctx->owned, S, &ctx->aug_w))) return rc;
    if ((rc = dev_zeros(ctx, ctx->owned, S, &ctx->aug_t))) return rc;
    NS_CUDA(cudaMemcpy(ctx->aug_w, wp.data(), S * sizeof(double), cudaMemcpyHostToDevice));
    ctx->aug_yw = volume_dot(G, plan, w);
    // stretched grids go to the Krylov solve as well (plain multigrid diverges beyond a stretch of ~5, ns_hostgeom.h plan_stretch), and so
    // does a grid that cannot be coarsened at all
    const bool mg_on_corrected_rhs = rhs_mode >= 1 && !outflow && ctx->levels.size() > 1 && ctx->aug_yw != 0.0 &&
                                     (rhs_mode >= 2 || plan_stretch(plan) <= 2.0);
    ctx->aug = !mg_on_corrected_rhs && ctx->aug_ww > 0.0;
    ctx->aug_rhs = mg_on_corrected_rhs;
  }
  ctx->krylov_phi = ctx->aug || ctx->levels.size() == 1;   // uniform and uncoarsenable: plain preconditioned BiCGStab on the consistent system
  if (ctx->solver == 1 || ctx->krylov_phi)
    for (int k = 0; k < 8; k++) { int rc = dev_zeros(ctx, ctx->owned, S, &ctx->kry[k]); if (rc) return rc; }
  { const EdgeDev* p; int rc = dev_upload(ctx, ctx->owned, plan.edges, &p); if (rc) return rc; ctx->d_edges = const_cast<EdgeDev*>(p); }
  { int rc = dev_zeros(ctx, ctx->owned, (size_t)5 * RED_MAX_BLOCKS, &ctx->red.partials); if (rc) return rc; }
  { double* c; int rc = dev_zeros(ctx, ctx->owned, 2, &c); if (rc) return rc; ctx->red.counter = (unsigned*)c; }
  { int rc = dev_zeros(ctx, ctx->owned, 8, &ctx->red.tmp); if (rc) return rc; }
  { int rc = dev_zeros(ctx, ctx->owned, 64, &ctx->d_scal); if (rc) return rc; }
  ctx->d_errflag = (int*)(ctx->d_scal + 48);
  NS_CUDA(cudaMallocHost((void**)&ctx->h_scal, 64 * sizeof(double)));
  { void* p = nullptr; NS_CUDA(cudaMalloc(&p, 3 * sizeof(SolveCtl))); ctx->owned.push_back(p); NS_CUDA(cudaMemset(p, 0, 3 * sizeof(SolveCtl))); ctx->d_ctl = (SolveCtl*)p; }
  NS_CUDA(cudaMallocHost((void**)&ctx->h_ctl, 3 * sizeof(SolveCtl)));
  memset(ctx->h_ctl, 0, 3 * sizeof(SolveCtl));
  if (ctx->world > 1 && ctx->use_p2p) {   // the peer map the kernels use
    PeerTable t{};
    t.world = ctx->world; t.rank = ctx->rank;
    for (int r = 0; r < ctx->world; r++) t.off[r] = r == ctx->rank ? 0 : (long long)((ctx->peers[r] - ctx->arena) / (ptrdiff_t)sizeof(double));
    t.hdr = (unsigned long long*)ctx->arena;
    t.err = ctx->d_errflag;
    t.nowait = getenv("NSGPU_XRED_NOWAIT") ? atoi(getenv("NSGPU_XRED_NOWAIT")) : 0;
    if (getenv("NSGPU_NOWAIT")) { const int one = atoi(getenv("NSGPU_NOWAIT")); t.nowait = one; NS_CUDA(cudaMemcpyToSymbol(g_ns_nowait, &one, sizeof one)); }
    void* p = nullptr;
    NS_CUDA(cudaMalloc(&p, sizeof(PeerTable)));
    ctx->owned.push_back(p);
    NS_CUDA(cudaMemcpy(p, &t, sizeof t, cudaMemcpyHostToDevice));
    ctx->d_pt = (PeerTable*)p;
    // The exchanges around the solves go through peer stores too on the all-multigrid route (uniform rectangles): there the order of
    // kernels and exchanges is fixed and audited (halo_exchange_push); the Krylov routes keep NCCL's rendezvous.
    ctx->hx_push = ctx->rect && ctx->solver == 0 && !ctx->krylov_phi;
    if (const char* e = getenv("NSGPU_HX_PUSH")) ctx->hx_push = ctx->hx_push && atoi(e);
  }
  if (ctx->world == 1 && getenv("NSGPU_SELFPUSH") && atoi(getenv("NSGPU_SELFPUSH")) && ctx->rect) {
    // profiling aid: the PUSH variants of k_march on ONE GPU (ncu cannot follow several ranks): the rank is its own left and right
    // neighbour, pushes land in its own halo columns (results are wrong), waits are disabled
    PeerTable t{};
    t.world = 1; t.rank = 0; t.nowait = 1;
    void* h = nullptr;
    NS_CUDA(cudaMalloc(&h, HDR_WORDS * sizeof(unsigned long long)));
    NS_CUDA(cudaMemset(h, 0, HDR_WORDS * sizeof(unsigned long long)));
    ctx->owned.push_back(h);
    t.hdr = (unsigned long long*)h;
    ctx->selfpush_hdr = t.hdr;
    t.err = ctx->d_errflag;
    void* p = nullptr;
    NS_CUDA(cudaMalloc(&p, sizeof(PeerTable)));
    ctx->owned.push_back(p);
    NS_CUDA(cudaMemcpy(p, &t, sizeof t, cudaMemcpyHostToDevice));
    ctx->d_pt = (PeerTable*)p;
    const int one = 1;
    NS_CUDA(cudaMemcpyToSymbol(g_ns_nowait, &one, sizeof one));
    ctx->selfpush = 1;
  }
  NS_CUDA(cudaDeviceSynchronize());
  if (ctx->world > 1) {   // nobody pushes into an arena that is still being zeroed
    NS_NCCL(ncclAllReduce(ctx->d_scal + 40, ctx->d_scal + 40, 1, ncclDouble, ncclSum, ctx->comm, ctx->stream));
    NS_CUDA(cudaStreamSynchronize(ctx->stream));
  }
  return 0;
}

}  // namespace nsgpu
