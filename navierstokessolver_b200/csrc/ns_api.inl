// ns_api.inl -- the extern "C" entry points declared in include/nsgpu.h.

using namespace nsgpu;

namespace nsgpu {

static double** field_slot(nsgpu_ctx* ctx, int which) {
  switch (which) {
    case NSGPU_U: return &ctx->u;
    case NSGPU_V: return &ctx->v;
    case NSGPU_PHI: return &ctx->phi;
    case NSGPU_RHS_U: return &ctx->rhs_u;
    case NSGPU_RHS_V: return &ctx->rhs_v;
    case NSGPU_RHS_PHI: return &ctx->rhs_phi;
    case NSGPU_CONV_U0: return &ctx->conv_u;
    case NSGPU_CONV_V0: return &ctx->conv_v;
    case NSGPU_GRADPHI_X: return &ctx->gpx;
    case NSGPU_GRADPHI_Y: return &ctx->gpy;
    case NSGPU_USTAR: return &ctx->ustar;
    case NSGPU_VSTAR: return &ctx->vstar;
  }
  return nullptr;
}

static int ensure_stage(nsgpu_ctx* ctx) {
  if (ctx->h_stage) return 0;
  NS_CUDA(cudaMallocHost((void**)&ctx->h_stage, ctx->levels[0].stride * sizeof(double)));
  return 0;
}

// host compact-id order -> padded device layout
static int upload_field(nsgpu_ctx* ctx, double* dev, const double* host) {
  if (dev == ctx->phi) {   // a new phi replaces whatever mean was pending
    ctx->phi_shift = false;
    k_scalar_set<<<1, 1, 0, ctx->stream>>>(&ctx->d_ctl[2].xsum, 0.0);
    NS_LAUNCH_CHECK();
  }
  ctx->halo_w.clear();
  ctx->bn2_ready[0] = ctx->bn2_ready[1] = ctx->bn2_ready[2] = false;   // a right-hand side may be what is being replaced
  const Level& L = ctx->levels[0];
  const LevelGeom& g = L.g;
  NS_CUDA(cudaSetDevice(ctx->device));
  if (ctx->rect) {
    NS_CUDA(cudaMemcpy2DAsync(dev + cidx(g, 0, 0), (size_t)g.ld * sizeof(double), host, (size_t)g.ny * sizeof(double),
                              (size_t)g.ny * sizeof(double), (size_t)g.nx, cudaMemcpyHostToDevice, ctx->stream));
    NS_CUDA(cudaStreamSynchronize(ctx->stream));
    return 0;
  }
  int rc = ensure_stage(ctx);
  if (rc) return rc;
  memset(ctx->h_stage, 0, L.stride * sizeof(double));
  for (int i = 0; i < g.nx; i++)
    for (int j = 0; j < g.ny; j++) {
      const int32_t id = ctx->local_id[(size_t)i * g.ny + j];
      if (id >= 0) ctx->h_stage[cidx(g, i, j)] = host[id];
    }
  NS_CUDA(cudaMemcpyAsync(dev, ctx->h_stage, L.stride * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
  NS_CUDA(cudaStreamSynchronize(ctx->stream));
  return 0;
}

static int download_field(nsgpu_ctx* ctx, const double* dev, double* host) {
  const Level& L = ctx->levels[0];
  const LevelGeom& g = L.g;
  NS_CUDA(cudaSetDevice(ctx->device));
  if (dev == ctx->phi) { int rcn = phi_normalize(ctx); if (rcn) return rcn; }
  if (ctx->rect) {
    NS_CUDA(cudaMemcpy2DAsync(host, (size_t)g.ny * sizeof(double), dev + cidx(g, 0, 0), (size_t)g.ld * sizeof(double),
                              (size_t)g.ny * sizeof(double), (size_t)g.nx, cudaMemcpyDeviceToHost, ctx->stream));
    NS_CUDA(cudaStreamSynchronize(ctx->stream));
    return 0;
  }
  int rc = ensure_stage(ctx);
  if (rc) return rc;
  NS_CUDA(cudaMemcpyAsync(ctx->h_stage, dev, L.stride * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
  NS_CUDA(cudaStreamSynchronize(ctx->stream));
  for (int i = 0; i < g.nx; i++)
    for (int j = 0; j < g.ny; j++) {
      const int32_t id = ctx->local_id[(size_t)i * g.ny + j];
      if (id >= 0) host[id] = ctx->h_stage[cidx(g, i, j)];
    }
  return 0;
}

static int sync_check(nsgpu_ctx* ctx) {
  NS_CUDA(cudaStreamSynchronize(ctx->stream));
  NS_CUDA(cudaGetLastError());
  return 0;
}

static void fill_info(nsgpu_step_info* info, const SolveResult& ru, const SolveResult& rv, const SolveResult& rp) {
  info->it_u = ru.cycles; info->it_v = rv.cycles; info->it_phi = rp.cycles;
  info->res_u = ru.rel; info->res_v = rv.rel; info->res_phi = rp.rel;
}

}  // namespace nsgpu

extern "C" {

int nsgpu_nccl_unique_id(void* out128) {
  ncclUniqueId id;
  if (ncclGetUniqueId(&id) != ncclSuccess) return NSGPU_ERR_NCCL;
  static_assert(sizeof(id) == 128, "ncclUniqueId is 128 bytes");
  memcpy(out128, &id, sizeof id);
  return 0;
}

int nsgpu_create(const nsgpu_grid* grid, const nsgpu_params* params, nsgpu_ctx** out) {
  if (!out) return NSGPU_ERR_ARG;
  *out = nullptr;
  nsgpu_ctx* ctx = new nsgpu_ctx;
  int rc = create_impl(ctx, grid, params);
  if (rc) {
    g_create_error = ctx->err;
    destroy_ctx(ctx);
    return rc;
  }
  *out = ctx;
  return 0;
}

void nsgpu_destroy(nsgpu_ctx* ctx) {
  if (ctx && getenv("NSGPU_XRED_DBG")) {
    long long h[8];
    cudaSetDevice(ctx->device);
    if (cudaMemcpyFromSymbol(h, g_ns_xdbg_acc, sizeof h) == cudaSuccess && h[0] > 0) {
      const double us = 1.0 / 1965.0;   // cycles -> microseconds at the B200's 1965 MHz
      fprintf(stderr, "nsgpu xred rank %d: %lld reductions: last block's sum %.2f us, epilogue %.2f us", ctx->rank, h[0], us * h[1] / h[0], us * h[2] / h[0]);
      if (h[7] > 0) fprintf(stderr, "; %lld cross-rank sums: stores %.2f, fence %.2f, flags %.2f, wait %.2f us", h[7], us * h[3] / h[7], us * h[4] / h[7], us * h[5] / h[7], us * h[6] / h[7]);
      fprintf(stderr, "\n");
    }
    unsigned gd[2] = {0, 0};
    if ((atoi(getenv("NSGPU_XRED_DBG")) & 2) && cudaMemcpyFromSymbol(gd, g_ns_cta_grid, sizeof gd) == cudaSuccess && gd[0] * gd[1] > 0 && gd[0] * gd[1] <= NS_DBG_CTAS) {
      std::vector<unsigned long long> t(4 * NS_DBG_CTAS);
      if (cudaMemcpyFromSymbol(t.data(), g_ns_cta_t, t.size() * sizeof t[0]) == cudaSuccess) {
        unsigned long long t0 = ~0ull;
        for (unsigned c = 0; c < gd[0] * gd[1]; c++) t0 = std::min(t0, t[4 * c]);
        fprintf(stderr, "nsgpu cta stamps rank %d, last k_march launch, grid %u x %u; per chunk row y: start / after wait / after march / end (us after the first CTA's start; min..max over x)\n", ctx->rank, gd[0], gd[1]);
        // rows 0 and 1 are the boundary chunks of a pushing pass (when it has them); the other rows together
        for (int part = 0; part < 3; part++) {
          double lo[4] = {1e30, 1e30, 1e30, 1e30}, hi[4] = {0, 0, 0, 0};
          const unsigned y0 = part < 2 ? part : 2, y1 = part < 2 ? part + 1 : gd[1];
          for (unsigned y = y0; y < y1 && y < gd[1]; y++)
            for (unsigned x = 0; x < gd[0]; x++)
              for (int k = 0; k < 4; k++) { const double v = 1e-3 * (double)(t[4 * (x + gd[0] * y) + k] - t0); lo[k] = std::min(lo[k], v); hi[k] = std::max(hi[k], v); }
          if (y0 < gd[1])
            fprintf(stderr, "nsgpu cta rank %d rows %u..%u  start %6.1f..%6.1f  waited %6.1f..%6.1f  marched %6.1f..%6.1f  end %6.1f..%6.1f\n", ctx->rank, y0, std::min(y1, gd[1]) - 1, lo[0], hi[0], lo[1], hi[1], lo[2], hi[2], lo[3], hi[3]);
        }
      }
    }
  }
  if (ctx && ctx->tail_dbg) {
    long long h[64];
    if (cudaMemcpy(h, ctx->tail_dbg, sizeof h, cudaMemcpyDeviceToHost) == cudaSuccess)
      for (int k = 1; k < 64 && h[k]; k++) fprintf(stderr, "nsgpu tail stamp %2d  +%8lld cycles (total %8lld)\n", k, h[k] - h[k - 1], h[k] - h[0]);
  }
  if (ctx && ctx->trace && !ctx->trace_acc.empty()) {
    std::vector<std::pair<int, std::pair<double, long>>> rows(ctx->trace_acc.begin(), ctx->trace_acc.end());
    std::sort(rows.begin(), rows.end(), [](const auto& a, const auto& b) { return a.first < b.first; });
    for (auto& r : rows)
      fprintf(stderr, "nsgpu trace rank %d key %6d  n %6ld  avg %9.2f us  total %9.3f ms\n", ctx->rank, r.first, r.second.second,
              1e3 * r.second.first / (double)r.second.second, r.second.first);
  }
  destroy_ctx(ctx);
}

const char* nsgpu_last_error(const nsgpu_ctx* ctx) { return ctx ? ctx->err.c_str() : g_create_error.c_str(); }

int nsgpu_local_range(const nsgpu_ctx* ctx, int32_t* i0, int32_t* i1, int64_t* id0, int64_t* n_local) {
  if (!ctx) return NSGPU_ERR_ARG;
  if (i0) *i0 = ctx->i0;
  if (i1) *i1 = ctx->i1;
  if (id0) *id0 = ctx->id0;
  if (n_local) *n_local = ctx->n_local;
  return 0;
}

int nsgpu_phase_ms(const nsgpu_ctx* ctx, float* out5) {
  if (!ctx || !out5) return NSGPU_ERR_ARG;
  for (int k = 0; k < 5; k++) out5[k] = ctx->phase_ms[k];
  return 0;
}
int nsgpu_num_levels(const nsgpu_ctx* ctx) { return ctx ? (int)ctx->levels.size() : 0; }
int64_t nsgpu_kernel_launches(const nsgpu_ctx* ctx) { return ctx ? ctx->launches : 0; }
int nsgpu_synchronize(nsgpu_ctx* ctx) { if (!ctx) return NSGPU_ERR_ARG; cudaSetDevice(ctx->device); return sync_check(ctx); }

int nsgpu_get_array(nsgpu_ctx* ctx, int32_t which, double* out) {
  if (!ctx || !out) return NSGPU_ERR_ARG;
  double** s = field_slot(ctx, which);
  if (!s) { ctx->err = "unknown array id"; return NSGPU_ERR_ARG; }
  return download_field(ctx, *s, out);
}
int nsgpu_set_array(nsgpu_ctx* ctx, int32_t which, const double* in) {
  if (!ctx || !in) return NSGPU_ERR_ARG;
  double** s = field_slot(ctx, which);
  if (!s) { ctx->err = "unknown array id"; return NSGPU_ERR_ARG; }
  return upload_field(ctx, *s, in);
}
int nsgpu_get_fields(nsgpu_ctx* ctx, double* u, double* v, double* phi) {
  int rc = 0;
  if (u && (rc = nsgpu_get_array(ctx, NSGPU_U, u))) return rc;
  if (v && (rc = nsgpu_get_array(ctx, NSGPU_V, v))) return rc;
  if (phi && (rc = nsgpu_get_array(ctx, NSGPU_PHI, phi))) return rc;
  return 0;
}
int nsgpu_set_fields(nsgpu_ctx* ctx, const double* u, const double* v, const double* phi) {
  int rc = 0;
  if (u && (rc = nsgpu_set_array(ctx, NSGPU_U, u))) return rc;
  if (v && (rc = nsgpu_set_array(ctx, NSGPU_V, v))) return rc;
  if (phi && (rc = nsgpu_set_array(ctx, NSGPU_PHI, phi))) return rc;
  return 0;
}

namespace {
// version 2: dt, re and the warm-start shift of the augmented phi solve (d_scal[46] = n * mean(x), outflow / stretched domains) are part
// of the state, so a restarted run continues bit for bit there too
struct CkptHeader { char magic[8]; int64_t nx, ny, n_cells, id0, n_local, step; double dt, re, aug_c, reserved; };
static_assert(sizeof(CkptHeader) == 88, "checkpoint header is 88 bytes");
const int kCkptFields[7] = {NSGPU_U, NSGPU_V, NSGPU_PHI, NSGPU_CONV_U0, NSGPU_CONV_V0, NSGPU_GRADPHI_X, NSGPU_GRADPHI_Y};
std::string ckpt_name(const nsgpu_ctx* ctx, const char* path) {
  return ctx->world > 1 ? std::string(path) + "." + std::to_string(ctx->rank) : std::string(path);
}
}  // namespace

int nsgpu_checkpoint_save(nsgpu_ctx* ctx, const char* path, int64_t step) {
  if (!ctx || !path) return NSGPU_ERR_ARG;
  cudaSetDevice(ctx->device);
  const std::string name = ckpt_name(ctx, path), tmp = name + ".tmp";
  FILE* f = fopen(tmp.c_str(), "wb");
  if (!f) { ctx->err = "cannot open " + tmp + " for writing"; return NSGPU_ERR_ARG; }
  CkptHeader h{};
  memcpy(h.magic, "NSGPUCK2", 8);
  h.nx = ctx->gnx; h.ny = ctx->ny; h.n_cells = ctx->n_cells_global; h.id0 = ctx->id0; h.n_local = ctx->n_local; h.step = step;
  h.dt = ctx->dt; h.re = ctx->re;
  if (cudaStreamSynchronize(ctx->stream) != cudaSuccess ||
      cudaMemcpy(&h.aug_c, ctx->d_scal + 46, sizeof(double), cudaMemcpyDeviceToHost) != cudaSuccess) {
    fclose(f); remove(tmp.c_str()); ctx->err = "checkpoint: device read failed"; return NSGPU_ERR_CUDA;
  }
  bool ok = fwrite(&h, sizeof h, 1, f) == 1;
  std::vector<double> buf((size_t)ctx->n_local);
  for (int k = 0; k < 7 && ok; k++) {
    const int rc = nsgpu_get_array(ctx, kCkptFields[k], buf.data());
    if (rc) { fclose(f); remove(tmp.c_str()); return rc; }
    ok = fwrite(buf.data(), sizeof(double), buf.size(), f) == buf.size();
  }
  ok = (fclose(f) == 0) && ok;
  if (ok) ok = rename(tmp.c_str(), name.c_str()) == 0;   // a crash while writing never leaves a truncated checkpoint under the final name
  if (!ok) { remove(tmp.c_str()); ctx->err = "short write to " + name; return NSGPU_ERR_ARG; }
  return 0;
}

int nsgpu_checkpoint_load(nsgpu_ctx* ctx, const char* path, int64_t* step) {
  if (!ctx || !path) return NSGPU_ERR_ARG;
  cudaSetDevice(ctx->device);
  const std::string name = ckpt_name(ctx, path);
  FILE* f = fopen(name.c_str(), "rb");
  if (!f) { ctx->err = "cannot open " + name; return NSGPU_ERR_ARG; }
  CkptHeader h{};
  if (fread(&h, sizeof h, 1, f) != 1 || memcmp(h.magic, "NSGPUCK2", 8) != 0) { fclose(f); ctx->err = name + " is not an nsgpu checkpoint (version 2)"; return NSGPU_ERR_ARG; }
  if (h.nx != ctx->gnx || h.ny != ctx->ny || h.n_cells != ctx->n_cells_global || h.id0 != ctx->id0 || h.n_local != ctx->n_local) {
    fclose(f); ctx->err = name + " was written for another grid or another slab decomposition"; return NSGPU_ERR_ARG;
  }
  if (h.dt != ctx->dt || h.re != ctx->re) {   // the history terms and divPhi in the file belong to that dt and re
    fclose(f); ctx->err = name + " was written with another dt or re"; return NSGPU_ERR_ARG;
  }
  // read everything before touching the device state: a truncated file leaves the context as it was
  std::vector<double> buf((size_t)7 * (size_t)ctx->n_local);
  const bool ok = fread(buf.data(), sizeof(double), buf.size(), f) == buf.size();
  fclose(f);
  if (!ok) { ctx->err = name + " is truncated"; return NSGPU_ERR_ARG; }
  for (int k = 0; k < 7; k++) {
    const int rc = nsgpu_set_array(ctx, kCkptFields[k], buf.data() + (size_t)k * (size_t)ctx->n_local);
    if (rc) return rc;
  }
  NS_CUDA(cudaMemcpy(ctx->d_scal + 46, &h.aug_c, sizeof(double), cudaMemcpyHostToDevice));
  if (step) *step = h.step;
  return 0;
}

int nsgpu_rhs_build(nsgpu_ctx* ctx) {
  if (!ctx) return NSGPU_ERR_ARG;
  cudaSetDevice(ctx->device);
  int rc = phase_rhs(ctx);
  return rc ? rc : sync_check(ctx);
}

int nsgpu_solve_velocity(nsgpu_ctx* ctx, nsgpu_step_info* info) {
  if (!ctx) return NSGPU_ERR_ARG;
  cudaSetDevice(ctx->device);
  SolveResult ru, rv, rp;
  int rc = phase_velocity(ctx, &ru, &rv);
  if (rc) return rc;
  if ((rc = sync_check(ctx))) return rc;
  finish_results(ctx, &ru, &rv, nullptr);
  if (info) { memset(info, 0, sizeof *info); fill_info(info, ru, rv, rp); }
  return (ru.converged && rv.converged) ? 0 : NSGPU_ERR_NOT_CONVERGED;
}

int nsgpu_div_rhs(nsgpu_ctx* ctx, int32_t remove_mean) {
  if (!ctx) return NSGPU_ERR_ARG;
  cudaSetDevice(ctx->device);
  int rc = phase_div(ctx, remove_mean != 0);
  return rc ? rc : sync_check(ctx);
}

int nsgpu_solve_phi(nsgpu_ctx* ctx, nsgpu_step_info* info) {
  if (!ctx) return NSGPU_ERR_ARG;
  cudaSetDevice(ctx->device);
  SolveResult ru, rv, rp;
  int rc = phase_phi(ctx, &rp);
  if (rc) return rc;
  if ((rc = sync_check(ctx))) return rc;
  finish_results(ctx, nullptr, nullptr, &rp);
  if (info) { memset(info, 0, sizeof *info); fill_info(info, ru, rv, rp); }
  return rp.converged ? 0 : NSGPU_ERR_NOT_CONVERGED;
}

int nsgpu_correct(nsgpu_ctx* ctx, nsgpu_step_info* info) {
  if (!ctx) return NSGPU_ERR_ARG;
  cudaSetDevice(ctx->device);
  double mm[4];
  int rc = phase_correct(ctx, mm);
  if (rc) return rc;
  if (info) { memset(info, 0, sizeof *info); info->umin = mm[0]; info->umax = mm[1]; info->vmin = mm[2]; info->vmax = mm[3]; }
  return 0;
}

}  // extern "C"

namespace nsgpu {

// async copies of one field between the compact host order and the padded device layout (rectangles only)
static cudaError_t copy_in(nsgpu_ctx* ctx, double* dev, const double* host, cudaStream_t st) {
  const LevelGeom& g = ctx->levels[0].g;
  return cudaMemcpy2DAsync(dev + cidx(g, 0, 0), (size_t)g.ld * sizeof(double), host, (size_t)g.ny * sizeof(double),
                           (size_t)g.ny * sizeof(double), (size_t)g.nx, cudaMemcpyHostToDevice, st);
}
// columns [c0, c1) only
static cudaError_t copy_in_cols(nsgpu_ctx* ctx, double* dev, const double* host, int c0, int c1, cudaStream_t st) {
  const LevelGeom& g = ctx->levels[0].g;
  return cudaMemcpy2DAsync(dev + cidx(g, c0, 0), (size_t)g.ld * sizeof(double), host + (size_t)c0 * g.ny, (size_t)g.ny * sizeof(double),
                           (size_t)g.ny * sizeof(double), (size_t)(c1 - c0), cudaMemcpyHostToDevice, st);
}
static cudaError_t copy_out_cols(nsgpu_ctx* ctx, const double* dev, double* host, int c0, int c1, cudaStream_t st) {
  const LevelGeom& g = ctx->levels[0].g;
  return cudaMemcpy2DAsync(host + (size_t)c0 * g.ny, (size_t)g.ny * sizeof(double), dev + cidx(g, c0, 0), (size_t)g.ld * sizeof(double),
                           (size_t)g.ny * sizeof(double), (size_t)(c1 - c0), cudaMemcpyDeviceToHost, st);
}
static cudaError_t copy_out(nsgpu_ctx* ctx, const double* dev, double* host, cudaStream_t st) {
  const LevelGeom& g = ctx->levels[0].g;
  return cudaMemcpy2DAsync(host, (size_t)g.ny * sizeof(double), dev + cidx(g, 0, 0), (size_t)g.ld * sizeof(double),
                           (size_t)g.ny * sizeof(double), (size_t)g.nx, cudaMemcpyDeviceToHost, st);
}

// n_steps time steps; io != nullptr (one step): the state comes from / goes back to host buffers, see nsgpu_step_host
static int step_impl(nsgpu_ctx* ctx, int32_t n_steps, nsgpu_step_info* last, const nsgpu_host_state* io) {
  if (!ctx || n_steps < 0) return NSGPU_ERR_ARG;
  cudaSetDevice(ctx->device);
  int io_k = 1, nx0 = 0;   // nsgpu_step_host: number of column chunks of the pipelined upload (1: none)
  if (io) {
    if (n_steps != 1 || !ctx->rect) { ctx->err = "nsgpu_step_host: one step, rectangular domains"; return NSGPU_ERR_ARG; }
    if (!ctx->s_in) {
      NS_CUDA(cudaStreamCreateWithFlags(&ctx->s_in, cudaStreamNonBlocking));
      NS_CUDA(cudaStreamCreateWithFlags(&ctx->s_out, cudaStreamNonBlocking));
      for (auto& e : ctx->ev_io) NS_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    }
      ctx->halo_w.clear();
    ctx->bn2_ready[0] = ctx->bn2_ready[1] = ctx->bn2_ready[2] = false;
    ctx->phi_shift = false;   // phi arrives from the host
    k_scalar_set<<<1, 1, 0, ctx->stream>>>(&ctx->d_ctl[2].xsum, 0.0);
    NS_LAUNCH_CHECK();
    // five-array form on one rank: the uploads of u, v, N_u, N_v are cut into column chunks (at least 8 columns each)
    nx0 = ctx->levels[0].g.nx;
    io_k = (!io->gradphi_x && ctx->world == 1 && ctx->rhs_march && nx0 >= 16) ? std::min(ctx->io_chunks, nx0 / 8) : 1;
    // upload stream: everything ConstructRHS_V reads first, phi (needed by the third solve only) behind it
    NS_CUDA(cudaEventRecord(ctx->ev_io[0], ctx->stream));          // earlier work on the solver's stream reads these arrays
    NS_CUDA(cudaStreamWaitEvent(ctx->s_in, ctx->ev_io[0], 0));
    if (io->gradphi_x) {
      NS_CUDA(copy_in(ctx, ctx->u, io->u, ctx->s_in));
      NS_CUDA(copy_in(ctx, ctx->v, io->v, ctx->s_in));
      NS_CUDA(copy_in(ctx, ctx->conv_u, io->conv_u0, ctx->s_in));
      NS_CUDA(copy_in(ctx, ctx->conv_v, io->conv_v0, ctx->s_in));
      NS_CUDA(copy_in(ctx, ctx->gpx, io->gradphi_x, ctx->s_in));
      NS_CUDA(copy_in(ctx, ctx->gpy, io->gradphi_y, ctx->s_in));
      NS_CUDA(cudaEventRecord(ctx->ev_io[1], ctx->s_in));
      NS_CUDA(copy_in(ctx, ctx->phi, io->phi, ctx->s_in));
      NS_CUDA(cudaEventRecord(ctx->ev_io[2], ctx->s_in));
      NS_CUDA(cudaStreamWaitEvent(ctx->stream, ctx->ev_io[1], 0));
    } else {
      // divPhi = GradP(phi^n) is not shipped: phi goes first and the gradient is rebuilt while u, v and the convective terms arrive
      NS_CUDA(copy_in(ctx, ctx->phi, io->phi, ctx->s_in));
      NS_CUDA(cudaEventRecord(ctx->ev_io[2], ctx->s_in));
      if (io_k > 1) {   // column chunks: ConstructRHS_V follows the uploads chunk by chunk, its convective terms go home at once
        if ((int)ctx->ev_chunk.size() < 2 * io_k) {
          const size_t have = ctx->ev_chunk.size();
          ctx->ev_chunk.resize((size_t)2 * io_k, nullptr);
          for (size_t k = have; k < ctx->ev_chunk.size(); k++) NS_CUDA(cudaEventCreateWithFlags(&ctx->ev_chunk[k], cudaEventDisableTiming));
        }
        for (int c = 0; c < io_k; c++) {
          const int c0 = (int)((long long)nx0 * c / io_k), c1 = (int)((long long)nx0 * (c + 1) / io_k);
          NS_CUDA(copy_in_cols(ctx, ctx->u, io->u, c0, c1, ctx->s_in));
          NS_CUDA(copy_in_cols(ctx, ctx->v, io->v, c0, c1, ctx->s_in));
          NS_CUDA(copy_in_cols(ctx, ctx->conv_u, io->conv_u0, c0, c1, ctx->s_in));
          NS_CUDA(copy_in_cols(ctx, ctx->conv_v, io->conv_v0, c0, c1, ctx->s_in));
          NS_CUDA(cudaEventRecord(ctx->ev_chunk[c], ctx->s_in));
        }
      } else {
        NS_CUDA(copy_in(ctx, ctx->u, io->u, ctx->s_in));
        NS_CUDA(copy_in(ctx, ctx->v, io->v, ctx->s_in));
        NS_CUDA(copy_in(ctx, ctx->conv_u, io->conv_u0, ctx->s_in));
        NS_CUDA(copy_in(ctx, ctx->conv_v, io->conv_v0, ctx->s_in));
      }
      NS_CUDA(cudaEventRecord(ctx->ev_io[1], ctx->s_in));
      NS_CUDA(cudaStreamWaitEvent(ctx->stream, ctx->ev_io[2], 0));
      const Level& L = ctx->levels[0];
      int rc0;
      if ((rc0 = halo_exchange(ctx, L, ctx->phi, 2))) return rc0;
      const dim3 grid = grid_cols(L.g, 256);
      k_gradphi<true><<<grid, 256, 0, ctx->stream>>>(L.g, ctx->phi, ctx->gpx, ctx->gpy);
      NS_LAUNCH_CHECK();
      ctx->halo_w.clear();
      if (io_k <= 1) NS_CUDA(cudaStreamWaitEvent(ctx->stream, ctx->ev_io[1], 0));
    }
  }
  SolveResult ru, rv, rp;
  double mm[4] = {0, 0, 0, 0};
  const int64_t l0 = ctx->launches;
  const double s0 = ctx->fine_sweeps_phi;
  ctx->krylov_restarts = 0;
  int status = 0, rc;
  NS_CUDA(cudaEventRecord(ctx->ev0, ctx->stream));
  for (int s = 0; s < n_steps; s++) {
    const bool last = s == n_steps - 1;   // the last step of the call carries the phase timers (events are free)
    if (last) cudaEventRecord(ctx->evp[0], ctx->stream);
    if (io && io_k > 1) {
      ctx->halo_w.clear();
      ctx->bn2_ready[0] = ctx->bn2_ready[1] = false;   // |rhs|^2 is reduced by the solve heads here
      for (int c = 0; c < io_k; c++) {
        const int c0 = (int)((long long)nx0 * c / io_k), c1 = (int)((long long)nx0 * (c + 1) / io_k);
        NS_CUDA(cudaStreamWaitEvent(ctx->stream, ctx->ev_chunk[std::min(c + 1, io_k - 1)], 0));   // radius-2 stencil: the next chunk too
        if ((rc = phase_rhs_cols(ctx, c0, c1))) return rc;
        NS_CUDA(cudaEventRecord(ctx->ev_chunk[io_k + c], ctx->stream));
        NS_CUDA(cudaStreamWaitEvent(ctx->s_out, ctx->ev_chunk[io_k + c], 0));
        NS_CUDA(copy_out_cols(ctx, ctx->conv_u, io->conv_u0, c0, c1, ctx->s_out));
        NS_CUDA(copy_out_cols(ctx, ctx->conv_v, io->conv_v0, c0, c1, ctx->s_out));
      }
    } else if ((rc = phase_rhs(ctx))) return rc;
    if (last) cudaEventRecord(ctx->evp[1], ctx->stream);
    if (io && io_k <= 1) {   // the convective terms are final: they go home while the solves run
      NS_CUDA(cudaEventRecord(ctx->ev_io[3], ctx->stream));
      NS_CUDA(cudaStreamWaitEvent(ctx->s_out, ctx->ev_io[3], 0));
      NS_CUDA(copy_out(ctx, ctx->conv_u, io->conv_u0, ctx->s_out));
      NS_CUDA(copy_out(ctx, ctx->conv_v, io->conv_v0, ctx->s_out));
    }
    if ((rc = phase_velocity(ctx, &ru, &rv))) return rc;
    if (last) cudaEventRecord(ctx->evp[2], ctx->stream);
    if ((rc = phase_div(ctx, true))) return rc;
    if (last) cudaEventRecord(ctx->evp[3], ctx->stream);
    if (io) NS_CUDA(cudaStreamWaitEvent(ctx->stream, ctx->ev_io[2], 0));   // phi has arrived
    if ((rc = phase_phi(ctx, &rp))) return rc;
    if (last) cudaEventRecord(ctx->evp[4], ctx->stream);
    if (io) {   // phi is final (mean pinned): out, under the correction step
      if ((rc = phi_normalize(ctx))) return rc;
      NS_CUDA(cudaEventRecord(ctx->ev_io[4], ctx->stream));
      NS_CUDA(cudaStreamWaitEvent(ctx->s_out, ctx->ev_io[4], 0));
      NS_CUDA(copy_out(ctx, ctx->phi, io->phi, ctx->s_out));
    }
    if ((rc = phase_correct(ctx, last ? mm : nullptr))) return rc;
    if (last) cudaEventRecord(ctx->evp[5], ctx->stream);
    if (io) {
      NS_CUDA(cudaEventRecord(ctx->ev_io[5], ctx->stream));
      NS_CUDA(cudaStreamWaitEvent(ctx->s_out, ctx->ev_io[5], 0));
      NS_CUDA(copy_out(ctx, ctx->u, io->u, ctx->s_out));
      NS_CUDA(copy_out(ctx, ctx->v, io->v, ctx->s_out));
      if (io->gradphi_x) {
        NS_CUDA(copy_out(ctx, ctx->gpx, io->gradphi_x, ctx->s_out));
        NS_CUDA(copy_out(ctx, ctx->gpy, io->gradphi_y, ctx->s_out));
      }
    }
    if (!((ru.pending || ru.converged) && (rv.pending || rv.converged) && (rp.pending || rp.converged))) status = NSGPU_ERR_NOT_CONVERGED;   // silent in the reference
  }
  NS_CUDA(cudaEventRecord(ctx->ev1, ctx->stream));
  if ((rc = sync_check(ctx))) return rc;
  if (io) { NS_CUDA(cudaStreamSynchronize(ctx->s_out)); NS_CUDA(cudaStreamSynchronize(ctx->s_in)); }
  trace_collect(ctx);
  if (n_steps > 0) {
    const bool pend[3] = {ru.pending, rv.pending, rp.pending};
    finish_results(ctx, &ru, &rv, &rp);
    for (int k = 0; k < 3; k++)   // device-controlled solves count their failures on the device, over all the steps of this call
      if (pend[k]) { if (ctx->h_ctl[k].fails != ctx->seen_fails[k]) status = NSGPU_ERR_NOT_CONVERGED; ctx->seen_fails[k] = ctx->h_ctl[k].fails; }
  }
  float ms = 0;
  NS_CUDA(cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1));
  if (n_steps > 0) for (int k = 0; k < 5; k++) cudaEventElapsedTime(&ctx->phase_ms[k], ctx->evp[k], ctx->evp[k + 1]);
  if (ctx->world > 1 && ctx->use_p2p) {
    double fl = 0.0;   // d_errflag lives in d_scal[48] (an int in a double's slot): published like every other scalar
    int rcf = read_scalars(ctx, ctx->d_scal + 48, 1, &fl);
    if (rcf) return rcf;
    int flag = 0;
    memcpy(&flag, &fl, sizeof flag);
    if (flag) { ctx->err = "a marching pass timed out waiting for a neighbour's halo push"; return NSGPU_ERR_NCCL; }
  }
  if (n_steps > 0 && !(std::isfinite(mm[0]) && std::isfinite(mm[1]) && std::isfinite(mm[2]) && std::isfinite(mm[3]))) status = NSGPU_ERR_NAN;
  if (last) {
    memset(last, 0, sizeof *last);
    fill_info(last, ru, rv, rp);
    last->umin = mm[0]; last->umax = mm[1]; last->vmin = mm[2]; last->vmax = mm[3];
    last->fine_sweeps = ctx->fine_sweeps_phi - s0;
    last->ms_total = ms;
    last->kernel_launches = (int32_t)(ctx->launches - l0);
    last->status = status;
    last->flags = (ctx->graph_fallback ? NSGPU_FLAG_GRAPH_FALLBACK : 0) | (ctx->krylov_phi ? NSGPU_FLAG_KRYLOV_PHI : 0) |
                  ((ctx->world > 1 && ctx->use_p2p && ctx->d_pt) ? NSGPU_FLAG_P2P : 0);
    last->krylov_restarts = ctx->krylov_restarts;
  }
  return status;
}

}  // namespace nsgpu

extern "C" {

int nsgpu_step(nsgpu_ctx* ctx, int32_t n_steps, nsgpu_step_info* last) { return nsgpu::step_impl(ctx, n_steps, last, nullptr); }

int nsgpu_step_host(nsgpu_ctx* ctx, const nsgpu_host_state* io, nsgpu_step_info* info) {
  if (!io || !io->u || !io->v || !io->phi || !io->conv_u0 || !io->conv_v0 || ((io->gradphi_x == nullptr) != (io->gradphi_y == nullptr))) return NSGPU_ERR_ARG;
  return nsgpu::step_impl(ctx, 1, info, io);
}

int nsgpu_poisson_sweeps(nsgpu_ctx* ctx, int32_t n_sweeps, double* residual_l2, float* ms) {
  if (!ctx || n_sweeps < 0) return NSGPU_ERR_ARG;
  cudaSetDevice(ctx->device);
  const Level& L = ctx->levels[0];
  int rc;
  if ((rc = phi_normalize(ctx))) return rc;
  ctx->halo_w.clear();
  if ((rc = halo_exchange(ctx, L, ctx->rhs_phi, 1))) return rc;
  NS_CUDA(cudaEventRecord(ctx->ev0, ctx->stream));
  if ((rc = smooth(ctx, L, coef_phi(), &ctx->phi, &ctx->tmp, ctx->rhs_phi, n_sweeps))) return rc;
  NS_CUDA(cudaEventRecord(ctx->ev1, ctx->stream));
  if ((rc = sync_check(ctx))) return rc;
  if (ms) NS_CUDA(cudaEventElapsedTime(ms, ctx->ev0, ctx->ev1));
  if (residual_l2) {
    double h;
    if ((rc = residual_norm2(ctx, L, coef_phi(), ctx->phi, ctx->rhs_phi, ctx->d_scal + 24))) return rc;
    if ((rc = read_scalars(ctx, ctx->d_scal + 24, 1, &h))) return rc;
    *residual_l2 = std::sqrt(h);
  }
  return 0;
}

// The passes the time step really runs on the fine level, timed alone (bench.py "roofline_inloop"): kind 1 = first pass of a phi V-cycle
// (two fused sweeps + residual restricted to level 1), 2 = last pass (coarse correction added, two fused sweeps, |r|^2 reduced),
// 3 = one red-black SOR double sweep of LHS_V on (u*, RHS_u); kind 0 = the plain sweep of nsgpu_poisson_sweeps.  Work on copies of
// nothing: phi / u* are smoothed in place n times (harmless for a benchmark context, wrong inside a time loop).
int nsgpu_time_passes(nsgpu_ctx* ctx, int32_t kind, int32_t n, float* ms) {
  if (!ctx || n < 0 || kind < 0 || kind > 3) return NSGPU_ERR_ARG;
  cudaSetDevice(ctx->device);
  if (kind == 0) return nsgpu_poisson_sweeps(ctx, n, nullptr, ms);
  const Level& L = ctx->levels[0];
  const OpCoef c = kind == 3 ? coef_v(ctx) : coef_phi();
  if (!march_level_ok(ctx, L, c) || (kind != 3 && ctx->levels.size() < 2)) { ctx->err = "nsgpu_time_passes: rectangle fast path with at least two levels only"; return NSGPU_ERR_ARG; }
  int rc;
  if ((rc = phi_normalize(ctx))) return rc;
  ctx->halo_w.clear();
  double** x = kind == 3 ? &ctx->ustar : &ctx->phi;
  const double* b = kind == 3 ? ctx->rhs_u : ctx->rhs_phi;
  if ((rc = halo_exchange(ctx, L, const_cast<double*>(b), HX - 1))) return rc;
  if ((rc = halo_exchange(ctx, L, *x, HX - 1))) return rc;
  ctx->halo_w[*x] = std::min(HX - 1, L.g.nx);
  if (kind == 2) {   // inside a cycle the level-1 pass has pushed the halo of the coarse correction: exchange it once, outside the timed launches
    const Level& C = ctx->levels[1];
    if ((rc = halo_exchange(ctx, C, C.x, HX - 1))) return rc;
    ctx->halo_w[C.x] = std::min(HX - 1, C.g.nx);
  }
  NS_CUDA(cudaEventRecord(ctx->ev0, ctx->stream));
  for (int k = 0; k < n; k++) {
    if (kind == 1) rc = march_pass(ctx, L, c, 2, *x, b, ctx->tmp, nullptr, nullptr, &ctx->levels[1], nullptr, nullptr);
    else if (kind == 2) rc = march_pass(ctx, L, c, 2, *x, b, ctx->tmp, &ctx->levels[1], ctx->levels[1].x, nullptr, ctx->d_scal + 24, nullptr);
    else rc = march_pass(ctx, L, c, 2, *x, b, ctx->tmp, nullptr, nullptr, nullptr, nullptr, nullptr, ctx->vel_omega > 0 ? ctx->vel_omega : 1.0);
    if (rc) return rc;
    std::swap(*x, ctx->tmp);
  }
  NS_CUDA(cudaEventRecord(ctx->ev1, ctx->stream));
  if ((rc = sync_check(ctx))) return rc;
  if (ms) NS_CUDA(cudaEventElapsedTime(ms, ctx->ev0, ctx->ev1));
  ctx->halo_w.clear();
  return 0;
}

int nsgpu_apply_operator(nsgpu_ctx* ctx, int32_t op, const double* x, double* y) {
  if (!ctx || !x || !y || (op != 0 && op != 1)) return NSGPU_ERR_ARG;
  cudaSetDevice(ctx->device);
  const Level& L = ctx->levels[0];
  // scratch: the smoother's ping-pong partner holds x, rhs_phi's slot is NOT touched: use two lazily made arrays
  static const int SCR = 2;
  if (!ctx->kry[0]) for (int k = 0; k < SCR; k++) { int rc = dev_zeros(ctx, ctx->owned, L.stride, &ctx->kry[k]); if (rc) return rc; }
  int rc;
  if ((rc = upload_field(ctx, ctx->kry[0], x))) return rc;
  if ((rc = apply_op(ctx, L, op == 0 ? coef_v(ctx) : coef_phi(), ctx->kry[0], ctx->kry[1], nullptr, nullptr))) return rc;
  if ((rc = sync_check(ctx))) return rc;
  return download_field(ctx, ctx->kry[1], y);
}

// ExportData without stalling the loop (FS.cpp:561, 569-615).  begin: one kernel on the solver's stream takes the snapshot -- copies of
// u, v, phi, plus lapPhi = LHS_phi phi and Pr = phi - dt/(2Re) lapPhi -- into five staging arrays, a copy stream brings them into one of
// two pinned buffers.  The call returns at once and the next time steps run under the copies with no dependence on them (an earlier
// version copied u, v, phi from the live arrays behind event fences: at 1024^2 the next step then waited for PCIe, 0.84 -> 1.12 ms).
// wait: blocks until the OLDEST pending snapshot is home and hands out pointers into the pinned buffer (valid until the second
// nsgpu_export_begin after it).
int nsgpu_export_begin(nsgpu_ctx* ctx, int64_t tag) {
  if (!ctx) return NSGPU_ERR_ARG;
  cudaSetDevice(ctx->device);
  if (ctx->exp_pending >= 2) { ctx->err = "nsgpu_export_begin: two snapshots are already pending (call nsgpu_export_wait)"; return NSGPU_ERR_ARG; }
  const Level& L = ctx->levels[0];
  const size_t n_host = ctx->rect ? (size_t)ctx->n_local : L.stride;
  if (!ctx->s_exp) {
    NS_CUDA(cudaStreamCreateWithFlags(&ctx->s_exp, cudaStreamNonBlocking));
    NS_CUDA(cudaEventCreateWithFlags(&ctx->exp_snap, cudaEventDisableTiming));
    for (auto& e : ctx->exp_home) NS_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    for (auto& e : ctx->exp_t0) NS_CUDA(cudaEventCreate(&e));
    for (auto& e : ctx->exp_t1) NS_CUDA(cudaEventCreate(&e));
    for (auto& q : ctx->exp_dev) { void* p = nullptr; NS_CUDA(cudaMalloc(&p, L.stride * sizeof(double))); ctx->owned.push_back(p); NS_CUDA(cudaMemset(p, 0, L.stride * sizeof(double))); q = (double*)p; }
    for (auto& q : ctx->exp_host) NS_CUDA(cudaMallocHost((void**)&q, 5 * n_host * sizeof(double)));
    for (auto& e : ctx->exp_home) NS_CUDA(cudaEventRecord(e, ctx->s_exp));
  }
  int rc;
  if ((rc = phi_normalize(ctx))) return rc;
  if ((rc = halo_exchange(ctx, L, ctx->phi, 2))) return rc;
  ctx->halo_w.clear();
  // the staging arrays are reused by every snapshot: the previous snapshot's copies must have left them
  NS_CUDA(cudaStreamWaitEvent(ctx->stream, ctx->exp_home[(ctx->exp_head + 1) & 1], 0));
  const dim3 grid = grid_cols(L.g, 256);
  const double half = ctx->dt / (2.0 * ctx->re);
  NS_CUDA(cudaEventRecord(ctx->exp_t0[ctx->exp_head], ctx->stream));
#define CALL(RECT) k_export_snapshot<RECT><<<grid, 256, 0, ctx->stream>>>(L.g, half, ctx->u, ctx->v, ctx->phi, ctx->exp_dev[0], ctx->exp_dev[1], ctx->exp_dev[2], ctx->exp_dev[3], ctx->exp_dev[4])
  NS_DISPATCH_RECT(ctx->rect, CALL);
#undef CALL
  NS_LAUNCH_CHECK();
  NS_CUDA(cudaEventRecord(ctx->exp_t1[ctx->exp_head], ctx->stream));
  NS_CUDA(cudaEventRecord(ctx->exp_snap, ctx->stream));
  NS_CUDA(cudaStreamWaitEvent(ctx->s_exp, ctx->exp_snap, 0));
  const int slot = ctx->exp_head;
  double* h = ctx->exp_host[slot];
  for (int k = 0; k < 5; k++) {   // host order: u, v, phi, Pr, lapPhi
    if (ctx->rect) {
      const LevelGeom& g = L.g;
      NS_CUDA(cudaMemcpy2DAsync(h + (size_t)k * n_host, (size_t)g.ny * sizeof(double), ctx->exp_dev[k] + cidx(g, 0, 0), (size_t)g.ld * sizeof(double),
                                (size_t)g.ny * sizeof(double), (size_t)g.nx, cudaMemcpyDeviceToHost, ctx->s_exp));
    } else {
      NS_CUDA(cudaMemcpyAsync(h + (size_t)k * n_host, ctx->exp_dev[k], L.stride * sizeof(double), cudaMemcpyDeviceToHost, ctx->s_exp));
    }
  }
  NS_CUDA(cudaEventRecord(ctx->exp_home[slot], ctx->s_exp));
  ctx->exp_tag[slot] = tag;
  ctx->exp_head ^= 1;
  ctx->exp_pending++;
  return 0;
}

int nsgpu_export_wait(nsgpu_ctx* ctx, const double** u, const double** v, const double** phi, const double** pr, const double** lap_phi, int64_t* tag) {
  if (!ctx) return NSGPU_ERR_ARG;
  if (ctx->exp_pending <= 0) { ctx->err = "nsgpu_export_wait: no snapshot pending"; return NSGPU_ERR_ARG; }
  cudaSetDevice(ctx->device);
  const int slot = (ctx->exp_head + 2 - ctx->exp_pending) & 1;
  NS_CUDA(cudaEventSynchronize(ctx->exp_home[slot]));
  ctx->exp_pending--;
  if (cudaEventElapsedTime(&ctx->exp_last_ms, ctx->exp_t0[slot], ctx->exp_t1[slot]) != cudaSuccess) { ctx->exp_last_ms = 0.0f; cudaGetLastError(); }
  const Level& L = ctx->levels[0];
  const double* base = ctx->exp_host[slot];
  size_t n_host = (size_t)ctx->n_local;
  if (!ctx->rect) {   // general domains: padded layout -> compact ids on the host
    const LevelGeom& g = L.g;
    std::vector<double>& c = ctx->exp_compact[slot];
    c.assign((size_t)5 * (size_t)ctx->n_local, 0.0);
    for (int k = 0; k < 5; k++)
      for (int i = 0; i < g.nx; i++)
        for (int j = 0; j < g.ny; j++) {
          const int32_t id = ctx->local_id[(size_t)i * g.ny + j];
          if (id >= 0) c[(size_t)k * ctx->n_local + id] = base[(size_t)k * L.stride + cidx(g, i, j)];
        }
    base = c.data();
  }
  const double** outp[5] = {u, v, phi, pr, lap_phi};
  for (int k = 0; k < 5; k++) if (outp[k]) *outp[k] = base + (size_t)k * n_host;
  if (tag) *tag = ctx->exp_tag[slot];
  return 0;
}

int nsgpu_export_last_ms(const nsgpu_ctx* ctx, float* ms) {
  if (!ctx || !ms) return NSGPU_ERR_ARG;
  *ms = ctx->exp_last_ms;
  return 0;
}

// physical pressure, synchronous form (kept for callers that want it now): one snapshot, waited for at once
int nsgpu_pressure(nsgpu_ctx* ctx, double* pr, double* lap_phi) {
  if (!ctx || !pr) return NSGPU_ERR_ARG;
  if (ctx->exp_pending) { ctx->err = "nsgpu_pressure: an asynchronous export is pending"; return NSGPU_ERR_ARG; }
  int rc = nsgpu_export_begin(ctx, -1);
  if (rc) return rc;
  const double *p = nullptr, *l = nullptr;
  if ((rc = nsgpu_export_wait(ctx, nullptr, nullptr, nullptr, &p, &l, nullptr))) return rc;
  memcpy(pr, p, (size_t)ctx->n_local * sizeof(double));
  if (lap_phi) memcpy(lap_phi, l, (size_t)ctx->n_local * sizeof(double));
  return 0;
}

}  // extern "C"
