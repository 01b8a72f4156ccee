// Host cost of build_plan for a backend / loop-closure sized graph:
//   g++ -O3 -std=c++17 -o tools/_build/plan_bench tools/plan_bench.cpp cut3r_slam_b200/csrc/dba_plan.cpp && tools/_build/plan_bench 128 2000 3
#include "../cut3r_slam_b200/csrc/dba_plan.h"
#include <chrono>
#include <cstdio>
#include <random>
#include <set>
using namespace dba;
int main(int argc, char** argv) {
  int P = argc > 1 ? atoi(argv[1]) : 128, E = argc > 2 ? atoi(argv[2]) : 2000, r = argc > 3 ? atoi(argv[3]) : 3;
  std::vector<int64_t> ii, jj; std::set<std::pair<int,int>> seen;
  for (int i = 0; i < P; ++i) for (int j = 0; j < P; ++j) if (i != j && abs(i - j) <= r && (int)ii.size() < E) { ii.push_back(i); jj.push_back(j); seen.insert({i, j}); }
  std::mt19937 g(1);
  while ((int)ii.size() < E) { int i = g() % P, j = g() % P; if (abs(i - j) <= r || seen.count({i, j})) continue; ii.push_back(i); jj.push_back(j); seen.insert({i, j}); if ((int)ii.size() < E && !seen.count({j, i})) { ii.push_back(j); jj.push_back(i); seen.insert({j, i}); } }
  Plan pl; double best = 1e9;
  for (int rep = 0; rep < 30; ++rep) {
    auto t = std::chrono::steady_clock::now();
    build_plan(pl, ii.data(), jj.data(), E, P, 2, P, false, -1, 4);
    double us = std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now() - t).count();
    if (us < best) best = us;
  }
  printf("P %d E %d: build_plan %.0f us, items %zu + %zu, nSblk %d\n", P, E, best, pl.sblk_items.size(), pl.rhs_items.size(), pl.nSblk);
}
