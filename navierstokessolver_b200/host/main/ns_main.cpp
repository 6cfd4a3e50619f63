// ns_main.cpp -- command-line driver with the reference's interface: main <grid data file> <simulation data file>
// (reference SRC/MAIN_Solver.cpp:11-18).  The reference's own MAIN_Solver.cpp also compiles unmodified against
// host/FluidSolver.h + libnsgpu.so; this file exists so the repository builds without it.
// Extra arguments (the reference hands them to PETSc as options, MAIN:10): -checkpoint_every <n> writes checkpoint.nsgpu every
// n steps, -restart <file> resumes the loop after the step that file was written at.
#include <cstdlib>
#include <cstring>
#include <iostream>
#include "FluidSolver.h"

static void run(char* gridFile, char* simFile, int every, const char* restart) {
  Grid grid{gridFile};
  if (!grid.setup) {
    std::cout << "Grid setup failed!\n";
    return;
  }
  FluidSolver solver{simFile, &grid};
  if (!solver.setup) return;
  solver.checkpointIter = every;
  if (restart && !solver.LoadCheckpoint(restart)) return;
  solver.Solve();
}

int main(int argc, char* argv[]) {
  int every = 0;
  const char* restart = nullptr;
  for (int k = 3; k + 1 < argc; k++) {
    if (!strcmp(argv[k], "-checkpoint_every")) every = atoi(argv[++k]);
    else if (!strcmp(argv[k], "-restart")) restart = argv[++k];
  }
  if (argc >= 3) run(argv[1], argv[2], every, restart);
  else std::cout << "Grid data file or simulation data file not provided!\n";
  return 0;   // like the reference, failures are reported on stdout only
}
