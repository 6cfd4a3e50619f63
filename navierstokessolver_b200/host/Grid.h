// Grid.h -- host-side mesh description with the public surface of the reference's Grid class
// (reference SRC/Grid.h:10-71): same public member names, types and meaning, so code written against the
// reference (including FluidSolver) compiles and behaves the same.  The implementation (Grid.cpp) is new.
#ifndef NSB200_GRID_H
#define NSB200_GRID_H

#include <cstdint>
#include <fstream>
#include <set>
#include <string>
#include <vector>

using namespace std;   // the reference header exports std into its includers (Grid.h:8); kept for drop-in use

enum bcTypes : int8_t { INLET_UNI, INLET_PARABOLIC, WALL, PRESSURE, NEUMANN };   // Grid.h:10
const double TOL = 1E-8;                                                         // Grid.h:11

struct Stencil {                 // Grid.h:13-18
  vector<double> weights;
  vector<vector<int>> support;
  vector<double> constant;
};
typedef vector<Stencil> StencilList;

struct Edge {                    // Grid.h:22-28
  int nx, ny;                    // outward normal of the boundary segment
  vector<double> loc{0, 0, 0};   // {coordinate of the line, lower end, upper end}
  int bcType{-1};
  double bcInfo{1.0};
  StencilList ghost;             // ghost-cell stencils: [0] velocities, [1] phi
};
typedef vector<Edge> EdgeList;

struct Cell {                    // Grid.h:32-38
  int id{-1};                    // compact cell number, -1 outside the domain
  double x, y;                   // centre
  vector<double> X{0, 0};        // x of the west / east face
  vector<double> Y{0, 0};        // y of the south / north face
  vector<int> edges{-1, -1, -1, -1};   // adjacent boundary edge on the -x, +x, -y, +y side
};
typedef vector<vector<Cell>> CellList;

class Grid {
 private:
  vector<double> xrange{1E15, -1E15};
  vector<double> yrange{1E15, -1E15};
  vector<double> AR{1E15, -1E15};
  vector<vector<double>> vertices;
  bool readSpec(ifstream& in);
  bool readBlock(ifstream& in, int kind);
  bool edgesFromPolygon();
  bool layFaces();
  bool expandAxis(vector<vector<double>>& segs, double lo, vector<double>& faces, vector<double>& widths, double& carry);
  void makeCells();
  void numberCells();
  bool classify(Cell& c);
  void report();
  void writeCellCenters();

 public:
  int N;                          // number of fluid cells
  vector<vector<double>> Nx;      // x segments: {start, end, cells, ratio}
  vector<vector<double>> Ny;
  vector<double> X, Y;            // face coordinates
  vector<double> hx, hy;          // cell widths
  CellList cells;
  EdgeList edges;
  bool setup = false;
  void ShowEdges(bool BC = false);
  bool inDomain(int, int);
  Grid(char* fname);
};

bool equals(double, double);

#endif
