// FluidSolver.h -- host class with the public surface of the reference's FluidSolver (SRC/FluidSolver.h:20-61):
//   FluidSolver(char* fname, Grid* grid);  bool setup;  void Solve();  double minmode(double, double);
// The time-step loop itself (reference FS.cpp:536-567 and everything it calls) runs on the GPU behind the C ABI of
// include/nsgpu.h; this class parses the simulation file, flattens the borrowed Grid into an nsgpu_grid, drives
// nsgpu_step() and writes the same stdout lines and FlowData_<iter>.csv files as the reference.
#ifndef NSB200_FLUIDSOLVER_H
#define NSB200_FLUIDSOLVER_H

#include "compat/petsc.h"   // the reference header pulls in petsc.h (FS.h:3); MAIN_Solver.cpp relies on it
#include "Grid.h"
#include "nsgpu.h"
#include <condition_variable>
#include <functional>
#include <mutex>
#include <thread>
#include <vector>

// NSGPU_GPUS=N (N > 1): one process drives N GPUs -- the domain is cut into N x-slabs (rank r on device NSGPU_DEVICE + r), one
// persistent host thread per GPU runs that rank's C-ABI calls; the calls of one time step must run concurrently (the kernels exchange
// halo columns with each other over NVLink), which is what run() does: the same job on every worker, return when all are done.
class GpuTeam {
  std::vector<std::thread> workers;
  std::mutex m;
  std::condition_variable cv_go, cv_done;
  std::function<void(int)> job;
  long gen = 0;
  int done = 0;
  bool stop = false;
  void loop(int r);
 public:
  explicit GpuTeam(int n);
  ~GpuTeam();
  int size() const { return (int)workers.size(); }
  void run(const std::function<void(int)>& f);
};

class FluidSolver {
 private:
  Grid* grid;                 // borrowed; its edges receive bcType / bcInfo / ghost stencils like in the reference
  nsgpu_ctx* gpu = nullptr;   // device-resident prevField, RHS vectors, operators (FS.h:22-33); rank 0 of gpus when there are several
  std::vector<nsgpu_ctx*> gpus;        // NSGPU_GPUS > 1: one context per GPU (gpus[0] == gpu)
  std::vector<int64_t> slab_id0, slab_n;   // compact-id range of every rank's x-slab
  GpuTeam* team = nullptr;
  template <class F> int onAll(F f);   // f(rank) on every GPU concurrently; first non-zero return code
  double dt;                  // deliberately uninitialised when the file omits them, like FS.h:24-26
  double finalTime;
  double re;
  int saveIter = 50;
  int firstIter = 1;          // 1, or the step after a loaded checkpoint
  int lastIter = 0;           // last completed step
  bool readSimulationFile(ifstream& in);
  bool checkParameters();
  void buildGhostStencils();
  bool createDeviceSolver();
  double ghostVelocity(const double* V, int i, int j, int e, int d);
  double ghostPhi(const double* P, int i, int j, int e);
  void ExportData(int iter);
  void writeSnapshot();       // formats the oldest pending device snapshot into FlowData_<iter>.csv (runs on `writer`)
  std::thread writer;         // ExportData does not stall the loop: the CSV of step n is written while steps n+1... run

 public:
  bool setup = false;
  FluidSolver(char* fname, Grid* grid);
  ~FluidSolver();
  void Solve();
  // extras (not in the reference): access for tests, and checkpoint / restart of the time loop -- Solve() resumes after the
  // step a loaded checkpoint was written at (nsgpu_checkpoint_save / _load, include/nsgpu.h)
  nsgpu_ctx* device() { return gpu; }
  bool SaveCheckpoint(const char* path);
  bool LoadCheckpoint(const char* path);
  int checkpointIter = 0;     // > 0: Solve() writes "checkpoint.nsgpu" every checkpointIter steps
};

double minmode(double a, double b);

#endif
