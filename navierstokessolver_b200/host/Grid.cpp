// Grid.cpp -- rectilinear-polygon mesh generator, written from scratch to the behaviour of the reference's
// Grid class (SRC/Grid.cpp): same input format (Grid.cpp:234-290), same face/width arithmetic so that hx, hy,
// X, Y come out bit-identical (Grid.cpp:75-128), same compact cell numbering (i outer, j inner,
// Grid.cpp:149-161), same edge tags per cell (Grid.cpp:165-185), same stdout summary and CellCenters.csv.
#include "Grid.h"

#include <algorithm>
#include <cmath>
#include <iomanip>
#include <iostream>
#include <sstream>

bool equals(double a, double b) { return !(std::abs(a - b) > TOL); }

Grid::Grid(char* fname) {
  ifstream in{fname};
  if (readSpec(in) && edgesFromPolygon() && layFaces()) {
    makeCells();
    numberCells();
    report();
    setup = true;
  }
  in.close();
  writeCellCenters();   // the reference exports even when the setup failed (Grid.cpp:13-15)
}

// ---- input -------------------------------------------------------------------------------------------------
// kind 0: "x y" pairs (free format); kind 1 / 2: one segment per line into Nx / Ny
bool Grid::readBlock(ifstream& in, int kind) {
  if ((in >> ws).get() != '{') return false;
  while (in.good()) {
    if ((in >> ws).peek() == '}') { in.ignore(); break; }
    if (kind == 0) {
      vector<double> p(2, 0.0);
      in >> p[0] >> p[1];
      vertices.push_back(p);
      if (in.eof()) return false;
    } else {
      string line;
      getline(in, line);
      vector<double> seg(4, -1.0);   // missing fields stay -1 (Grid.cpp:260)
      istringstream ls(line);
      for (int k = 0; k < 4 && (ls >> seg[k]); k++) {}
      (kind == 1 ? Nx : Ny).push_back(seg);
      if (in.eof()) return false;
    }
  }
  return true;
}

bool Grid::readSpec(ifstream& in) {
  if (!in) { cout << "Grid data file not found!\n"; return false; }
  bool ok = true;
  string key;
  while (!in.eof()) {
    in >> key;
    if (in.eof()) break;
    if (key == "Vertices") ok = readBlock(in, 0);
    else if (key == "Nx") ok = readBlock(in, 1);
    else if (key == "Ny") ok = readBlock(in, 2);
    else ok = false;
    if (!ok || !in.good()) { ok = false; break; }
  }
  if (!ok) cout << "Invalid data file format!\n";
  return ok;
}

// ---- boundary edges: vertex k -> k+1, clockwise polygon (Grid.cpp:28-72) --------------------------------------
bool Grid::edgesFromPolygon() {
  cout << "Generating Edges...\n";
  vertices.push_back(vertices[0]);
  for (size_t k = 0; k + 1 < vertices.size(); k++) {
    const vector<double>&a = vertices[k], &b = vertices[k + 1];
    edges.push_back({});
    Edge& e = edges.back();
    if (b[0] == a[0]) {                       // vertical segment
      e.loc[0] = b[0];
      e.loc[1] = std::min(a[1], b[1]);
      e.loc[2] = std::max(a[1], b[1]);
      if (b[1] > a[1]) { e.nx = -1; if (b[0] < xrange[0]) xrange[0] = b[0]; }   // going up: fluid on the right
      else { e.nx = 1; if (b[0] > xrange[1]) xrange[1] = b[0]; }
    } else if (b[1] == a[1]) {                // horizontal segment
      e.loc[0] = b[1];
      e.loc[1] = std::min(a[0], b[0]);
      e.loc[2] = std::max(a[0], b[0]);
      if (b[0] > a[0]) { e.ny = 1; if (b[1] > yrange[1]) yrange[1] = b[1]; }    // going right: fluid below
      else { e.ny = -1; if (b[1] < yrange[0]) yrange[0] = b[1]; }
    } else {
      cout << "Edges should be parallel to the x-axis or y-axis\n";
      return false;
    }
  }
  return true;
}

// ---- faces: piecewise uniform / geometric segments (Grid.cpp:75-128) -------------------------------------------
// `carry` is the last width produced; it seeds segments given without a cell count and, like in the
// reference, carries over from the x pass into the y pass.
bool Grid::expandAxis(vector<vector<double>>& segs, double lo, vector<double>& faces, vector<double>& widths, double& carry) {
  faces.push_back(lo);
  for (auto& s : segs) {
    if (!equals(s[0], faces.back()) || (faces.size() == 1 && s[2] <= 0)) return false;
    const double len = s[1] - s[0], r = s[3];
    double x = s[0];
    if (r > 0) {
      if (s[2] <= 0) s[2] = ceil(log(len * (r - 1) / carry + 1) / log(r));
      carry = len * (r - 1) / (pow(r, s[2]) - 1);
      for (int k = 0; k < s[2]; k++) { faces.push_back(x += carry); widths.push_back(carry); carry *= r; }
    } else if (r == -1) {
      if (s[2] <= 0) s[2] = ceil(len / carry);
      carry = len / s[2];
      for (int k = 0; k < s[2]; k++) { faces.push_back(x += carry); widths.push_back(carry); }
    } else {
      return false;
    }
  }
  return true;
}

bool Grid::layFaces() {
  cout << "Generating Faces...\n";
  double carry = 0.0;
  // the reference seeds both face lists before the first check (Grid.cpp:78-79)
  bool ok = expandAxis(Nx, xrange[0], X, hx, carry);
  if (!ok) Y.push_back(yrange[0]);
  ok = ok && expandAxis(Ny, yrange[0], Y, hy, carry);
  ok = ok && equals(xrange[1], X.back()) && equals(yrange[1], Y.back());
  if (!ok) cout << "Invalid specification for number of cells\n";
  return ok;
}

// ---- cells -----------------------------------------------------------------------------------------------------
void Grid::makeCells() {
  cout << "Generating Cells...\n";
  const size_t nx = X.size() - 1, ny = Y.size() - 1;
  cells.assign(nx, vector<Cell>(ny));
  for (size_t i = 0; i < nx; i++)
    for (size_t j = 0; j < ny; j++) {
      Cell& c = cells[i][j];
      c.X[0] = X[i]; c.X[1] = X[i + 1];
      c.Y[0] = Y[j]; c.Y[1] = Y[j + 1];
      c.x = 0.5 * (c.X[0] + c.X[1]);
      c.y = 0.5 * (c.Y[0] + c.Y[1]);
    }
}

// Ray cast towards +x over the vertical edges; also tags the sides lying on a boundary edge (Grid.cpp:165-185).
bool Grid::classify(Cell& c) {
  int crossings = 0;
  for (size_t k = 0; k < edges.size(); k++) {
    const Edge& e = edges[k];
    if (e.nx != 0) {
      if (!(c.y > e.loc[1] && c.y < e.loc[2])) continue;
      if (e.loc[0] > c.x) crossings++;
      if (e.nx == -1 && equals(c.X[0], e.loc[0])) c.edges[0] = (int)k;
      if (e.nx == 1 && equals(c.X[1], e.loc[0])) c.edges[1] = (int)k;
    } else {
      if (!(c.x > e.loc[1] && c.x < e.loc[2])) continue;
      if (e.ny == -1 && equals(c.Y[0], e.loc[0])) c.edges[2] = (int)k;
      if (e.ny == 1 && equals(c.Y[1], e.loc[0])) c.edges[3] = (int)k;
    }
  }
  return crossings % 2 == 1;
}

void Grid::numberCells() {
  int next = 0;
  for (auto& column : cells)
    for (auto& c : column) {
      if (!classify(c)) continue;
      c.id = next++;
      const double ar = (c.X[1] - c.X[0]) / (c.Y[1] - c.Y[0]);
      if (AR[0] > ar) AR[0] = ar;
      if (AR[1] < ar) AR[1] = ar;
    }
  N = next;
}

bool Grid::inDomain(int i, int j) {
  return i >= 0 && i < (int)hx.size() && j >= 0 && j < (int)hy.size() && cells[i][j].id >= 0;
}

// ---- output (same text as the reference: Grid.cpp:188-231) ------------------------------------------------------
void Grid::ShowEdges(bool BC) {
  cout << endl;
  for (size_t k = 0; k < edges.size(); k++) {
    const Edge& e = edges[k];
    cout << "Edge\t" << k << ":\t" << "n\t" << e.nx << "\t" << e.ny << "\t";
    cout << (e.nx != 0 ? "x\t" : "y\t") << e.loc[0] << (e.nx != 0 ? "\ty\t" : "\tx\t") << e.loc[1] << "\t" << e.loc[2] << endl;
    if (BC) cout << "BC\t" << e.bcType << "\t" << e.bcInfo << endl;
  }
}

void Grid::report() {
  cout << setprecision(2);
  ShowEdges();
  cout << endl;
  cout << "X:\t" << xrange[0] << "\t" << xrange[1] << endl;
  cout << "Y:\t" << yrange[0] << "\t" << yrange[1] << endl;
  cout << "AR:\t" << AR[0] << "\t" << AR[1] << endl;
  cout << "hx:\t" << *min_element(hx.begin(), hx.end()) << "\t" << *max_element(hx.begin(), hx.end()) << endl;
  cout << "hy:\t" << *min_element(hy.begin(), hy.end()) << "\t" << *max_element(hy.begin(), hy.end()) << endl << endl;
}

void Grid::writeCellCenters() {
  ofstream out{"CellCenters.csv"};
  for (auto& column : cells)
    for (auto& c : column)
      if (c.id != -1) out << c.x << "," << c.y << endl;
  out.close();
}
