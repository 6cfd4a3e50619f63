// FluidSolver.cpp -- host driver of the GPU time-step loop; behaviour modelled on the reference's
// FluidSolver (SRC/FluidSolver.cpp), implementation new.  What stays on the host: the simulation-file parser
// (FS.cpp:626-669), parameter validation (FS.cpp:16-32), ghost-stencil bookkeeping on the borrowed Grid
// (FS.cpp:84-103), the per-step stdout table (FS.cpp:559-560) and the CSV export (FS.cpp:569-615).
// Everything numerical goes through include/nsgpu.h.
#include "FluidSolver.h"

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <iostream>
#include <sstream>

GpuTeam::GpuTeam(int n) {
  for (int r = 0; r < n; r++) workers.emplace_back([this, r] { loop(r); });
}
GpuTeam::~GpuTeam() {
  { std::lock_guard<std::mutex> l(m); stop = true; }
  cv_go.notify_all();
  for (auto& t : workers) t.join();
}
void GpuTeam::loop(int r) {
  long seen = 0;
  for (;;) {
    std::function<void(int)> f;
    {
      std::unique_lock<std::mutex> l(m);
      cv_go.wait(l, [&] { return stop || gen != seen; });
      if (stop) return;
      seen = gen;
      f = job;
    }
    f(r);
    { std::lock_guard<std::mutex> l(m); done++; }
    cv_done.notify_one();
  }
}
void GpuTeam::run(const std::function<void(int)>& f) {
  std::unique_lock<std::mutex> l(m);
  job = f; done = 0; gen++;
  cv_go.notify_all();
  cv_done.wait(l, [&] { return done == (int)workers.size(); });
}

template <class F> int FluidSolver::onAll(F f) {
  if (!team) return f(0);
  std::vector<int> rc(gpus.size(), 0);
  team->run([&](int r) { rc[r] = f(r); });
  for (int c : rc) if (c) return c;
  return 0;
}

double minmode(double a, double b) {   // FS.cpp:671-674 (kept as a free function of the public surface)
  if (a * b > 0) return a * std::min(1.0, std::abs(b / a));
  return 0;
}

FluidSolver::FluidSolver(char* fname, Grid* g) : grid(g) {
  ifstream in{fname};
  if (readSimulationFile(in) && checkParameters()) {
    buildGhostStencils();
    if (createDeviceSolver()) {
      cout << "Solver Setup Complete!\n";
      setup = true;
    }
  }
  in.close();
}

FluidSolver::~FluidSolver() {
  if (writer.joinable()) writer.join();
  if (team) {
    onAll([&](int r) { if (gpus[r]) nsgpu_destroy(gpus[r]); return 0; });
    delete team;
  } else if (gpu) nsgpu_destroy(gpu);
}

// Keywords: BC { one "<type> [info]" line per edge }, dt, final_time, re, saveIter.  As in the reference every
// item must be followed by more input (a number that ends exactly at end-of-file is rejected, FS.cpp:662).
bool FluidSolver::readSimulationFile(ifstream& in) {
  if (!in) { cout << "Simulation data file not found!\n"; return false; }
  bool ok = true;
  string key;
  while (!in.eof()) {
    in >> key;
    if (in.eof()) break;
    if (key == "BC") {
      if ((in >> ws).get() != '{') { ok = false; break; }
      size_t k = 0;
      while (in.good()) {
        if ((in >> ws).peek() == '}') {
          if (k == grid->edges.size()) in.ignore(); else ok = false;
          break;
        }
        string line;
        getline(in, line);
        if (in.eof() || k == grid->edges.size()) { ok = false; break; }
        istringstream ls(line);
        ls >> grid->edges[k].bcType >> grid->edges[k].bcInfo;   // a missing info keeps Edge's default 1.0
        k++;
      }
      if (!ok) break;
    }
    else if (key == "dt") in >> dt;
    else if (key == "final_time") in >> finalTime;
    else if (key == "re") in >> re;
    else if (key == "saveIter") in >> saveIter;
    else { ok = false; break; }
    if (!in.good()) { ok = false; break; }
  }
  if (!ok) cout << "Invalid data file format!\n";
  return ok;
}

bool FluidSolver::checkParameters() {   // same order and messages as FS.cpp:17-32
  if (dt <= 0) { cout << "Time step should be positive\n"; return false; }
  if (dt > finalTime) { cout << "Time step should be less than final time!\n"; return false; }
  if (re <= 0) { cout << "Reynolds number should be positive\n"; return false; }
  if (saveIter <= 0) { cout << "saveIter must be greater than zero!\n"; return false; }
  return true;
}

// The reference stores the ghost-cell closures on the Grid's edges (FS.cpp:84-103) where user code can see them;
// the device builds its own copy from bcType / bcInfo, this one serves ExportData and API compatibility.
void FluidSolver::buildGhostStencils() {
  for (Edge& e : grid->edges) {
    if (e.bcType <= WALL) {
      Stencil vel, phi;
      vel.weights = {-1}; vel.support = {{0, 0}};
      phi.weights = {1};  phi.support = {{0, 0}};
      const double c = 2 * e.bcInfo;
      if (e.bcType == INLET_UNI) vel.constant = (e.nx == 0) ? vector<double>{0, c} : vector<double>{c, 0};
      else if (e.bcType == WALL) vel.constant = (e.nx != 0) ? vector<double>{0, c} : vector<double>{c, 0};
      e.ghost.push_back(vel);
      e.ghost.push_back(phi);
    } else if (e.bcType == NEUMANN) {
      Stencil vel, phi;
      vel.weights = {1}; vel.support = {{0, 0}}; vel.constant = {0, 0};
      phi.weights = {2.5, -2.0, 0.5};
      phi.support = {{0, 0}, {-e.nx, -e.ny}, {-2 * e.nx, -2 * e.ny}};
      e.ghost.push_back(vel);
      e.ghost.push_back(phi);
    }
  }
}

static int env_int(const char* name, int dflt) { const char* v = std::getenv(name); return v ? std::atoi(v) : dflt; }
static double env_double(const char* name, double dflt) { const char* v = std::getenv(name); return v ? std::atof(v) : dflt; }

bool FluidSolver::createDeviceSolver() {
  const int nx = (int)grid->hx.size(), ny = (int)grid->hy.size();
  vector<int32_t> ids((size_t)nx * ny), tags((size_t)nx * ny * 4);
  for (int i = 0; i < nx; i++)
    for (int j = 0; j < ny; j++) {
      const Cell& c = grid->cells[i][j];
      ids[(size_t)i * ny + j] = c.id;
      for (int k = 0; k < 4; k++) tags[((size_t)i * ny + j) * 4 + k] = c.edges[k];
    }
  const int ne = (int)grid->edges.size();
  vector<int32_t> enx(ne), eny(ne), etype(ne);
  vector<double> einfo(ne);
  for (int k = 0; k < ne; k++) {
    enx[k] = grid->edges[k].nx; eny[k] = grid->edges[k].ny;
    etype[k] = grid->edges[k].bcType; einfo[k] = grid->edges[k].bcInfo;
  }
  nsgpu_grid g{};
  g.nx = nx; g.ny = ny; g.n_cells = grid->N;
  g.hx = grid->hx.data(); g.hy = grid->hy.data();
  g.cell_id = ids.data(); g.cell_edges = tags.data();
  g.n_edges = ne; g.edge_nx = enx.data(); g.edge_ny = eny.data(); g.edge_bc_type = etype.data(); g.edge_bc_info = einfo.data();
  nsgpu_params p{};
  p.dt = dt; p.re = re;
  p.device = env_int("NSGPU_DEVICE", 0);
  p.rank = 0; p.world = 1;
  p.rtol_u = env_double("NSGPU_RTOL_U", 1e-8);       // KSPSetTolerances(..., 1.e-8, ...) FS.cpp:68,80
  p.rtol_phi = env_double("NSGPU_RTOL_PHI", 1e-8);
  p.solver = env_int("NSGPU_SOLVER", grid->N == nx * ny ? 0 : 1);
  p.max_cycles = env_int("NSGPU_MAX_CYCLES", 0);
  const int ngpu = env_int("NSGPU_GPUS", 1);
  if (ngpu > 1) {   // one process, ngpu x-slabs: one context per GPU, created concurrently (the communicator needs every rank)
    unsigned char id[128];
    if (nsgpu_nccl_unique_id(id) != 0) { cout << "GPU solver setup failed: no NCCL id\n"; return false; }
    team = new GpuTeam(ngpu);
    gpus.assign(ngpu, nullptr);
    std::vector<std::string> msg(ngpu);
    std::vector<int> rcs(ngpu, 0);
    const int dev0 = p.device;
    team->run([&](int r) {
      nsgpu_params q = p;
      q.rank = r; q.world = ngpu; q.device = dev0 + r; q.nccl_id = id;
      rcs[r] = nsgpu_create(&g, &q, &gpus[r]);
      if (rcs[r]) { msg[r] = nsgpu_last_error(nullptr); gpus[r] = nullptr; }
    });
    for (int r = 0; r < ngpu; r++)
      if (rcs[r]) {
        cout << "GPU solver setup failed (" << rcs[r] << ") on GPU " << dev0 + r << ": " << msg[r] << "\n";
        onAll([&](int q) { if (gpus[q]) nsgpu_destroy(gpus[q]); gpus[q] = nullptr; return 0; });
        delete team; team = nullptr; gpus.clear();
        return false;
      }
    gpu = gpus[0];
    slab_id0.assign(ngpu, 0); slab_n.assign(ngpu, 0);
    for (int r = 0; r < ngpu; r++) nsgpu_local_range(gpus[r], nullptr, nullptr, &slab_id0[r], &slab_n[r]);
    return true;
  }
  int rc = nsgpu_create(&g, &p, &gpu);
  if (rc != 0) {
    cout << "GPU solver setup failed (" << rc << "): " << nsgpu_last_error(nullptr) << "\n";
    gpu = nullptr;
    return false;
  }
  gpus.assign(1, gpu);
  slab_id0.assign(1, 0); slab_n.assign(1, (int64_t)grid->N);
  return true;
}

// Same loop control as FS.cpp:536-567: iter starts at 1, the first step always runs, stop when dt*iter > finalTime.
bool FluidSolver::SaveCheckpoint(const char* path) {
  if (!gpu || onAll([&](int r) { return nsgpu_checkpoint_save(gpus[r], path, lastIter); }) != 0) { cout << "Checkpoint not written: " << nsgpu_last_error(gpu) << "\n"; return false; }
  return true;
}

bool FluidSolver::LoadCheckpoint(const char* path) {
  int64_t step = 0;
  std::vector<int64_t> steps(gpus.size(), 0);
  const int rcl = gpu ? onAll([&](int r) { return nsgpu_checkpoint_load(gpus[r], path, &steps[r]); }) : -1;
  if (!steps.empty()) step = steps[0];
  if (rcl != 0) { cout << "Checkpoint not loaded: " << nsgpu_last_error(gpu) << "\n"; return false; }
  lastIter = (int)step;
  firstIter = lastIter + 1;
  return true;
}

void FluidSolver::Solve() {
  int iter = firstIter;
  cout << "Initiating Solver..\n";
  if (firstIter > 1 && dt * iter > finalTime) {   // restarted from a checkpoint of the last step: nothing left to do
    cout << "Solution Complete!\n";
    return;
  }
  do {
    nsgpu_step_info info;
    std::vector<nsgpu_step_info> infos(gpus.size());
    int rc = onAll([&](int r) { return nsgpu_step(gpus[r], 1, &infos[r]); });   // min / max are combined over the ranks on the device: every rank reports the same four
    info = infos[0];
    if (rc != 0 && rc != NSGPU_ERR_NOT_CONVERGED && rc != NSGPU_ERR_NAN) {
      cout << "GPU time step failed (" << rc << "): " << nsgpu_last_error(gpu) << "\n";
      return;
    }
    if ((iter - 1) % 10 == 0) printf("iter\tumin\t\tumax\t\tvmin\t\tvmax\n");
    printf("%d\t%lf\t%lf\t%lf\t%lf\n", iter, info.umin, info.umax, info.vmin, info.vmax);
    if (saveIter > 0 && iter % saveIter == 0) ExportData(iter);
    lastIter = iter;
    if (checkpointIter > 0 && iter % checkpointIter == 0) SaveCheckpoint("checkpoint.nsgpu");
    iter++;
  } while (dt * iter <= finalTime);
  if (writer.joinable()) writer.join();   // the last export is on disk before "Solution Complete!"
  cout << "Solution Complete!\n";
}

double FluidSolver::ghostVelocity(const double* V, int i, int j, int e, int d) {   // EvaluateGhostStencil_V, FS.cpp:166-173
  const Stencil& s = grid->edges[e].ghost[0];
  double r = 0.0;
  for (size_t l = 0; l < s.weights.size(); l++) r += s.weights[l] * V[grid->cells[i + s.support[l][0]][j + s.support[l][1]].id];
  return r + s.constant[d];
}

double FluidSolver::ghostPhi(const double* P, int i, int j, int e) {                // EvaluateGhostStencil_P, FS.cpp:175-181
  const Stencil& s = grid->edges[e].ghost[1];
  double r = 0.0;
  for (size_t l = 0; l < s.weights.size(); l++) r += s.weights[l] * P[grid->cells[i + s.support[l][0]][j + s.support[l][1]].id];
  return r;
}

// FlowData_<iter>.csv: one row per cell centre followed by one row per adjacent boundary face (FS.cpp:588-605).
// The reference computes lapPhi and Pr on the host and writes the file inside the time loop (FS.cpp:561, 569-615).  Here the snapshot
// (u, v, phi, Pr, lapPhi) is taken by one kernel, travels home on a copy stream and is formatted by a writer thread while the next
// time steps run (nsgpu_export_begin / nsgpu_export_wait); NSGPU_EXPORT_SYNC=1 writes the file before the loop goes on.
void FluidSolver::ExportData(int iter) {
  if (writer.joinable()) writer.join();   // one file at a time; its pinned buffer is free again for the snapshot after this one
  if (onAll([&](int r) { return nsgpu_export_begin(gpus[r], iter); }) != 0) {
    cout << "GPU export failed: " << nsgpu_last_error(gpu) << "\n";
    return;
  }
  writer = std::thread([this] { writeSnapshot(); });
  if (env_int("NSGPU_EXPORT_SYNC", 0)) writer.join();
}

void FluidSolver::writeSnapshot() {
  const double *u, *v, *phi, *pr, *lap;
  int64_t iter = 0;
  std::vector<double> all;   // several GPUs: the slabs' snapshots side by side in compact-id order
  if (gpus.size() > 1) {
    const size_t n = (size_t)grid->N;
    all.resize(5 * n);
    for (size_t r = 0; r < gpus.size(); r++) {
      const double* q[5];
      if (nsgpu_export_wait(gpus[r], &q[0], &q[1], &q[2], &q[3], &q[4], &iter) != 0) {
        cout << "GPU export failed: " << nsgpu_last_error(gpus[r]) << "\n";
        return;
      }
      for (int k = 0; k < 5; k++) std::copy(q[k], q[k] + slab_n[r], all.begin() + k * n + slab_id0[r]);
    }
    u = all.data(); v = u + n; phi = v + n; pr = phi + n; lap = pr + n;
  } else if (nsgpu_export_wait(gpu, &u, &v, &phi, &pr, &lap, &iter) != 0) {
    cout << "GPU export failed: " << nsgpu_last_error(gpu) << "\n";
    return;
  }
  char name[255];
  snprintf(name, sizeof name, "FlowData_%d.csv", (int)iter);
  vector<char> iobuf(1 << 20);
  ofstream out;
  out.rdbuf()->pubsetbuf(iobuf.data(), (std::streamsize)iobuf.size());
  out.open(name);
  out << "Point_X,Point_Y,Point_Z,U,V,Pr\n";
  const double half = dt / (2 * re);
  for (size_t i = 0; i < grid->cells.size(); i++)
    for (size_t j = 0; j < grid->cells[i].size(); j++) {
      const Cell& c = grid->cells[i][j];
      if (c.id < 0) continue;
      out << c.x << "," << c.y << "," << 0.0 << "," << u[c.id] << "," << v[c.id] << "," << pr[c.id] << "\n";
      for (int k : c.edges) {
        if (k < 0) continue;
        const Edge& e = grid->edges[k];
        out << c.x + e.nx * grid->hx[i] / 2 << "," << c.y + e.ny * grid->hy[j] / 2 << "," << 0.0 << ","
            << 0.5 * (u[c.id] + ghostVelocity(u, (int)i, (int)j, k, 0)) << ","
            << 0.5 * (v[c.id] + ghostVelocity(v, (int)i, (int)j, k, 1)) << ","
            << 0.5 * (phi[c.id] + ghostPhi(phi, (int)i, (int)j, k)) - half * lap[c.id] << "\n";
      }
    }
  out.close();
}
