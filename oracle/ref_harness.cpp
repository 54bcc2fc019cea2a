// oracle/ref_harness.cpp -- TEST INFRASTRUCTURE, not product code.
//
// A small pybind11 module (`obn_refharness`) compiled TOGETHER WITH THE
// REFERENCE'S OWN HEADERS (from /root/reference, where they lie) into
// oracle/_ref/.  It does two things the unmodified reference cannot do by
// itself:
//   1. run the reference's ContingencyTable<true> + citest<true> /
//      calculate_local_score<double,true> over a LIST of tuples under
//      `#pragma omp parallel for` -- the "constructed OpenMP CPU baseline" of
//      BASELINE.md section 3 (the reference has no OpenMP on this path);
//   2. read back the statistic and dof that citest<>() keeps as locals
//      (citest.h:86-87,139) through the recorder in the Boost shim;
//   3. call PDAGwithAdjMat::apply_meeks_rules (src/graph/pdag_with_adjmat.cpp:215-279),
//      which the reference compiles but neither binds nor calls (pc.cpp:84), on an
//      arbitrary adjacency matrix.
// Nothing here re-implements reference arithmetic; it only calls it.
#include <omp.h>
#include <pybind11/numpy.h>
#include <pybind11/pybind11.h>
#include <pybind11/stl.h>

#include <chrono>
#include <cmath>
#include <limits>
#include <vector>

#include "base/contingency_table.h"
#include "base/dataframe_wrapper.h"
#include "citest/citest.h"
#include "graph/pdag_with_adjmat.h"
#include "score/local_score.h"

namespace py = pybind11;
using i64 = long long;

static double now_s() {
  return std::chrono::duration<double>(
             std::chrono::steady_clock::now().time_since_epoch())
      .count();
}

// CI tests: tuple t is (x[t], y[t] | z_flat[z_off[t] .. z_off[t+1])).
static py::dict ci_batch(const DataframeWrapper& df,
                         py::array_t<i64, py::array::c_style> x,
                         py::array_t<i64, py::array::c_style> y,
                         py::array_t<i64, py::array::c_style> z_flat,
                         py::array_t<i64, py::array::c_style> z_off,
                         int kind, double level, int threads) {
  const size_t T = x.size();
  const i64 *px = x.data(), *pyv = y.data(), *pz = z_flat.data(),
            *po = z_off.data();
  py::array_t<bool> dec(T);
  py::array_t<double> stat(T), pval(T);
  py::array_t<i64> dof(T);
  bool* pdec = dec.mutable_data();
  double *pstat = stat.mutable_data(), *pp = pval.mutable_data();
  i64* pdof = dof.mutable_data();
  CITestType test = (kind == 0) ? CITestType(ChiSquare(level))
                                : CITestType(GSquare(level));
  if (threads <= 0) threads = omp_get_max_threads();
  double t0, t1;
  {
    py::gil_scoped_release rel;
    t0 = now_s();
#pragma omp parallel for schedule(dynamic) num_threads(threads)
    for (size_t t = 0; t < T; ++t) {
      std::vector<size_t> Z(pz + po[t], pz + po[t + 1]);
      std::vector<size_t> vars = Z;
      vars.push_back((size_t)px[t]);
      vars.push_back((size_t)pyv[t]);
      std::sort(vars.begin(), vars.end());  // as pc.cpp:34-37
      ContingencyTable<true> ct(vars, df);  // pc.cpp:38
      boost::math::obn_shim::last_k = -1.0;
      boost::math::obn_shim::last_x = -1.0;
      const bool d = citest<true>((size_t)px[t], (size_t)pyv[t], Z, ct, test);
      pdec[t] = d;
      if (boost::math::obn_shim::last_k < 0) {  // dof == 0 branch of pchisq
        pdof[t] = 0;
        pstat[t] = d ? 0.0 : std::numeric_limits<double>::quiet_NaN();
        pp[t] = d ? 1.0 : 0.0;
      } else {
        pdof[t] = (i64)boost::math::obn_shim::last_k;
        pstat[t] = boost::math::obn_shim::last_x;
        pp[t] = pchisq(pstat[t], (size_t)pdof[t]);
      }
    }
    t1 = now_s();
  }
  py::dict out;
  out["decision"] = dec;
  out["stat"] = stat;
  out["dof"] = dof;
  out["p"] = pval;
  out["seconds"] = t1 - t0;
  out["threads"] = threads;
  return out;
}

// Local scores: tuple t is (child[t] | pa_flat[pa_off[t] .. pa_off[t+1])).
static py::dict score_batch(const DataframeWrapper& df,
                            py::array_t<i64, py::array::c_style> child,
                            py::array_t<i64, py::array::c_style> pa_flat,
                            py::array_t<i64, py::array::c_style> pa_off,
                            double ess, int threads) {
  const size_t T = child.size();
  const i64 *pc = child.data(), *pf = pa_flat.data(), *po = pa_off.data();
  py::array_t<double> sc(T);
  double* ps = sc.mutable_data();
  ScoreType st = BDeu(ess);
  if (threads <= 0) threads = omp_get_max_threads();
  double t0, t1;
  {
    py::gil_scoped_release rel;
    t0 = now_s();
#pragma omp parallel for schedule(dynamic) num_threads(threads)
    for (size_t t = 0; t < T; ++t) {
      std::vector<size_t> pa(pf + po[t], pf + po[t + 1]);
      std::vector<size_t> vars = pa;
      vars.push_back((size_t)pc[t]);
      std::sort(vars.begin(), vars.end());
      ContingencyTable<true> ct(vars, df);
      ps[t] = calculate_local_score<double, true>((size_t)pc[t], pa, ct, st);
    }
    t1 = now_s();
  }
  py::dict out;
  out["score"] = sc;
  out["seconds"] = t1 - t0;
  out["threads"] = threads;
  return out;
}

// One table per subset serving |W| scores, the reference's own sharing in
// exhaustive_search.cpp:55-72.  Tuple t is the subset w_flat[w_off[t]..).
// Returns the scores flattened in (subset, ascending child) order.
static py::dict subset_score_batch(const DataframeWrapper& df,
                                   py::array_t<i64, py::array::c_style> w_flat,
                                   py::array_t<i64, py::array::c_style> w_off,
                                   double ess, int threads) {
  const size_t T = w_off.size() - 1;
  const i64 *pf = w_flat.data(), *po = w_off.data();
  py::array_t<double> sc(w_flat.size());
  double* ps = sc.mutable_data();
  ScoreType st = BDeu(ess);
  if (threads <= 0) threads = omp_get_max_threads();
  double t0, t1;
  {
    py::gil_scoped_release rel;
    t0 = now_s();
#pragma omp parallel for schedule(dynamic) num_threads(threads)
    for (size_t t = 0; t < T; ++t) {
      std::vector<size_t> vars(pf + po[t], pf + po[t + 1]);
      ContingencyTable<true> ct(vars, df);
      for (size_t c = 0; c < vars.size(); ++c) {
        std::vector<size_t> pa;
        for (auto v : vars)
          if (v != vars[c]) pa.push_back(v);
        ps[po[t] + c] = calculate_local_score<double, true>(vars[c], pa, ct, st);
      }
    }
    t1 = now_s();
  }
  py::dict out;
  out["score"] = sc;
  out["seconds"] = t1 - t0;
  out["threads"] = threads;
  return out;
}

// adj[u][v] != 0 <=> arc u -> v (undirected = both).  Built by deleting arcs from the complete
// graph the constructor makes, since that is the only public way in.
static py::array_t<uint8_t> meek(py::array_t<uint8_t, py::array::c_style> adj, bool apply_r4) {
  const size_t n = (size_t)adj.shape(0);
  auto a = adj.unchecked<2>();
  PDAGwithAdjMat g(n);
  for (size_t u = 0; u < n; ++u)
    for (size_t v = 0; v < n; ++v)
      if (u != v && !a(u, v)) g.remove_directed_edge(u, v);
  g.apply_meeks_rules(apply_r4);
  py::array_t<uint8_t> out({n, n});
  auto o = out.mutable_unchecked<2>();
  for (size_t u = 0; u < n; ++u)
    for (size_t v = 0; v < n; ++v) o(u, v) = (u != v && g.has_directed_edge(u, v)) ? 1 : 0;
  return out;
}

PYBIND11_MODULE(obn_refharness, m) {
  m.doc() = "OpenMP batch harness around the unmodified reference functions";
  m.def("ci_batch", &ci_batch);
  m.def("score_batch", &score_batch);
  m.def("subset_score_batch", &subset_score_batch);
  m.def("meek", &meek, py::arg("adj"), py::arg("apply_r4") = false);
  m.def("max_threads", []() { return omp_get_max_threads(); });
  m.def("pchisq", [](double x, size_t dof) { return pchisq(x, dof); });
}
