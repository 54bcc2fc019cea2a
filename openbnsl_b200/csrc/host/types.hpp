// types.hpp -- tag types of the reference's score / citest interfaces.
//   include/score/score_type.h:9-63      BDeu{ess}, K2, BIC, AIC (only BDeu has arithmetic)
//   include/citest/citest_type.h:11-65   ChiSquare{level}, GSquare{level}, OracleGraph{graph}
#pragma once
#include <variant>

#include "graph.hpp"

namespace obn {

struct BDeu {
  double ess;
  explicit BDeu(double ess = 1.0) : ess(ess) {}
};
struct K2 {};
struct BIC {};
struct AIC {};
using ScoreType = std::variant<BDeu, K2, BIC, AIC>;

struct ChiSquare {
  double level;
  explicit ChiSquare(double level = 0.05) : level(level) {}
};
struct GSquare {
  double level;
  explicit GSquare(double level = 0.05) : level(level) {}
};
struct OracleGraph {
  PDAG graph;
  explicit OracleGraph(const PDAG& g) : graph(g) {}
};
using CITestType = std::variant<ChiSquare, GSquare, OracleGraph>;

}  // namespace obn
