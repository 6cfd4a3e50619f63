// nshost_capi.cpp -- plain-C access to the host Grid / FluidSolver classes, for the Python mirror
// (navierstokessolver_b200/host.py) and the host-side tests.  Not part of the reference's surface.
#include <iostream>
#include <sstream>
#include "FluidSolver.h"

namespace {
struct Quiet {
  std::streambuf* old = nullptr;
  std::ostringstream sink;
  explicit Quiet(bool on) { if (on) old = std::cout.rdbuf(sink.rdbuf()); }
  ~Quiet() { if (old) std::cout.rdbuf(old); }
};
struct HostHandle { Grid* grid = nullptr; FluidSolver* fs = nullptr; std::string log; };
}

extern "C" {

void* nshost_grid_create(const char* gridfile, int quiet) {
  HostHandle* h = new HostHandle;
  Quiet q(quiet != 0);
  h->grid = new Grid(const_cast<char*>(gridfile));
  h->log = q.sink.str();
  return h;
}
const char* nshost_log(void* p) { return ((HostHandle*)p)->log.c_str(); }
int nshost_grid_setup(void* p) { return ((HostHandle*)p)->grid->setup ? 1 : 0; }
int nshost_n(void* p) { return ((HostHandle*)p)->grid->N; }
int nshost_nx(void* p) { return (int)((HostHandle*)p)->grid->hx.size(); }
int nshost_ny(void* p) { return (int)((HostHandle*)p)->grid->hy.size(); }
int nshost_n_edges(void* p) { return (int)((HostHandle*)p)->grid->edges.size(); }
void nshost_grid_arrays(void* p, double* hx, double* hy, int* cell_id, int* cell_edges, double* xc, double* yc) {
  Grid* g = ((HostHandle*)p)->grid;
  const int nx = (int)g->hx.size(), ny = (int)g->hy.size();
  for (int i = 0; i < nx; i++) hx[i] = g->hx[i];
  for (int j = 0; j < ny; j++) hy[j] = g->hy[j];
  for (int i = 0; i < nx; i++)
    for (int j = 0; j < ny; j++) {
      const Cell& c = g->cells[i][j];
      cell_id[i * ny + j] = c.id;
      for (int k = 0; k < 4; k++) cell_edges[(i * ny + j) * 4 + k] = c.edges[k];
      if (xc) xc[i * ny + j] = c.x;
      if (yc) yc[i * ny + j] = c.y;
    }
}
void nshost_edges(void* p, int* enx, int* eny, int* type, double* info, double* loc) {
  Grid* g = ((HostHandle*)p)->grid;
  for (size_t k = 0; k < g->edges.size(); k++) {
    enx[k] = g->edges[k].nx; eny[k] = g->edges[k].ny; type[k] = g->edges[k].bcType; info[k] = g->edges[k].bcInfo;
    if (loc) for (int q = 0; q < 3; q++) loc[3 * k + q] = g->edges[k].loc[q];
  }
}
// FluidSolver on an existing grid handle; returns 1 when setup succeeded
int nshost_solver_create(void* p, const char* simfile, int quiet) {
  HostHandle* h = (HostHandle*)p;
  Quiet q(quiet != 0);
  h->fs = new FluidSolver(const_cast<char*>(simfile), h->grid);
  h->log = q.sink.str();
  return h->fs->setup ? 1 : 0;
}
void nshost_solve(void* p) { ((HostHandle*)p)->fs->Solve(); }
nsgpu_ctx* nshost_device(void* p) { HostHandle* h = (HostHandle*)p; return h->fs ? h->fs->device() : nullptr; }
void nshost_destroy(void* p) {
  HostHandle* h = (HostHandle*)p;
  delete h->fs;
  delete h->grid;
  delete h;
}
double nshost_minmode(double a, double b) { return minmode(a, b); }

}  // extern "C"
