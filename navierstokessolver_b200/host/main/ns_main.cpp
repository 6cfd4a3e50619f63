// ns_main.cpp -- command-line driver with the reference's interface: main <grid data file> <simulation data file>
// (reference SRC/MAIN_Solver.cpp:11-18).  The reference's own MAIN_Solver.cpp also compiles unmodified against
// host/FluidSolver.h + libnsgpu.so; this file exists so the repository builds without it.
#include <iostream>
#include "FluidSolver.h"

static void run(char* gridFile, char* simFile) {
  Grid grid{gridFile};
  if (!grid.setup) {
    std::cout << "Grid setup failed!\n";
    return;
  }
  FluidSolver solver{simFile, &grid};
  if (solver.setup) solver.Solve();
}

int main(int argc, char* argv[]) {
  if (argc >= 3) run(argv[1], argv[2]);
  else std::cout << "Grid data file or simulation data file not provided!\n";
  return 0;   // like the reference, failures are reported on stdout only
}
