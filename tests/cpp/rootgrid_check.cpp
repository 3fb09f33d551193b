// rootgrid_check.cpp — the sparse-grid bridge against the REFERENCE's real
// sparse::RootGrid, not a stand-in.  Compiled only where the reference tree is mounted
// (tests/test_cpp_host.py: -I/root/reference/lib/sparse/include, with oracle/tbb_shim standing in
// for oneTBB, which the reference downloads at configure time); the binary travels to the GPU box.
//
//   rootgrid_check cells   CPU only: enumerates RootGrid::foreach (RootGrid.hpp:20-22, 62-75) for the
//                          reference's own line-writing workload (test_hpc_benchmark.cpp:859-880) and
//                          prints the (x, y, value) triples — the golden fixture
//                          tests/golden/rootgrid_cells.json was made from this output
//   rootgrid_check gpu     sparse::csr_from_grid on the true tree: rows must be the enumerated cells in
//                          x-major order, x0 the cells' values, and A*x0 must equal, bit for bit, the
//                          5-point operator evaluated through RootGrid::read
// the reference headers rely on transitive includes of their own build (std::intptr_t, std::vector, ...)
#include <cstdint>
#include <functional>
#include <memory>
#include <optional>
#include <utility>
#include <vector>

#include <SparseDS.hpp>

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <map>
#include <mutex>
#include <string>
#include <tuple>
#include <vector>

#include <hpcx/CsrMatrix.hpp>

using Grid = sparse::RootGrid<float, sparse::HashBlock<sparse::DenseBlock<16, float>>>;
using Grid2 = sparse::RootGrid<float, sparse::HashBlock<sparse::PointerBlock<8, sparse::DenseBlock<4, float>>>>;

template <typename G>
static void fill(G &grid, int n) {
  // the reference's workload: a particle on a straight line (test_hpc_benchmark.cpp:862-869)
  float px = -100.f, py = 100.f, vx = 0.2f, vy = -0.6f;
  for (int time = 0; time < n; ++time)
    grid.write(static_cast<std::intptr_t>(std::floor(px + vx * time)), static_cast<std::intptr_t>(std::floor(py + vy * time)),
               1.0f + 0.001f * (float)(time % 97));
}

template <typename G>
static std::vector<std::tuple<long, long, float>> enumerate(G &grid) {
  std::vector<std::tuple<long, long, float>> cells;
  std::mutex mu;
  grid.foreach ([&](auto x, auto y, auto &value) {
    std::lock_guard<std::mutex> lock(mu);
    cells.emplace_back((long)x, (long)y, (float)value);
  });
  std::sort(cells.begin(), cells.end());
  return cells;
}

template <typename G>
static int check_gpu(hpcx::Context &ctx, const char *name, int n, bool via_read) {
  G grid;
  fill(grid, n);
  const auto cells = enumerate(grid);
  sparse::ActiveCells ac;
  std::vector<double> x0, y;
  auto dG = sparse::csr_from_grid<double>(ctx, grid, &ac, &x0);
  int ok = dG.nrows() == cells.size() && ac.row_x.size() == cells.size();
  for (std::size_t r = 0; ok && r < cells.size(); ++r)  // rows = enumerated cells, (x, y) lexicographic
    ok = ac.row_x[r] == std::get<0>(cells[r]) && ac.row_y[r] == std::get<1>(cells[r]) && x0[r] == (double)std::get<2>(cells[r]);
  sparse::spmv(dG, x0, y);
  // A cell is active iff foreach enumerates it.  (For a single-level layout — HashBlock over
  // DenseBlock — RootGrid::read agrees with that, which `via_read` asserts.  In a multi-level tree
  // the reference's non-leaf blocks carry BShift = log2(their own grid size), not the cumulative
  // shift (BaseBlock.hpp:45-47, HashBlock.hpp:50-51, RootGrid.hpp:69-71,79-84), so the coordinates
  // foreach reconstructs are not the ones write / read address; the bridge takes what foreach
  // reports, which is its contract.)
  std::map<std::pair<long, long>, double> active;
  for (auto &c : cells) active[{std::get<0>(c), std::get<1>(c)}] = (double)std::get<2>(c);
  auto at = [&](long cx, long cy) -> std::pair<bool, double> {
    auto it = active.find({cx, cy});
    return {it != active.end(), it != active.end() ? it->second : 0.0};
  };
  if (via_read)
    for (auto &c : cells) {
      auto v = grid.read(std::get<0>(c), std::get<1>(c));
      ok = ok && v.has_value() && *v == std::get<2>(c) && !grid.read(std::get<0>(c) + 100000, std::get<1>(c)).has_value();
    }
  std::size_t bad = 0;
  for (std::size_t r = 0; r < cells.size(); ++r) {
    const long cx = std::get<0>(cells[r]), cy = std::get<1>(cells[r]);
    double s = 0.0;  // ascending columns: (x-1,y) (x,y-1) (x,y) (x,y+1) (x+1,y), each product rounded, then added
    const std::pair<long, long> nb[5] = {{cx - 1, cy}, {cx, cy - 1}, {cx, cy}, {cx, cy + 1}, {cx + 1, cy}};
    for (int k = 0; k < 5; ++k) {
      auto [active, v] = at(nb[k].first, nb[k].second);
      if (!active) continue;
      const double prod = (k == 2 ? 4.0 : -1.0) * v;
      s = s + prod;
    }
    bad += std::memcmp(&s, &y[r], 8) != 0;
  }
  ok = ok && bad == 0;
  std::printf("rootgrid %s cells %zu rows %zu product_mismatches %zu ok %d\n", name, cells.size(), dG.nrows(), bad, ok);
  return !ok;
}

int main(int argc, char **argv) {
  const std::string mode = argc > 1 ? argv[1] : "cells";
  try {
    if (mode == "cells") {
      Grid grid;
      fill(grid, 400);
      for (auto &c : enumerate(grid)) std::printf("cell %ld %ld %.9g\n", std::get<0>(c), std::get<1>(c), std::get<2>(c));
      return 0;
    }
    hpcx::Context ctx(0);
    int bad = check_gpu<Grid>(ctx, "hash_dense16", 5000, true);
    bad += check_gpu<Grid2>(ctx, "hash_pointer8_dense4", 3000, false);
    std::printf(bad ? "ROOTGRID FAILED\n" : "ROOTGRID OK\n");
    return bad ? 3 : 0;
  } catch (const std::exception &e) {
    std::printf("exception %s\n", e.what());
    return 2;
  }
}
