// bindings.cpp -- pybind11 module `openbnsllib` with the reference's Python
// surface (bindings/pybind_module.cpp:13-20 and bindings/submodules/*.cpp):
// submodules base / score / citest / structure_learning, same class and
// function names, argument names, defaults and exception classes.  The data
// path of every function runs on the GPU through the C ABI.
#include <deque>
#include <pybind11/functional.h>
#include <pybind11/numpy.h>
#include <pybind11/pybind11.h>
#include <pybind11/stl.h>

#include <cstdlib>
#include <cstring>

#include "dataframe_wrapper.hpp"
#include "structure_learning.hpp"

namespace py = pybind11;
using namespace obn;

namespace {

PcStats g_last_pc_stats;
EsStats g_last_es_stats;

template <typename T>
py::array_t<T> view(const T* p, size_t n) {   // zero-copy numpy view over C++ memory
  return py::array_t<T>({(py::ssize_t)n}, {(py::ssize_t)sizeof(T)}, p, py::none());
}

// Python object with .rank, .world, .allreduce_min_i32(arr), .allgather(send, recv)
struct NcclCommTag {};   // Python-side marker: "use the communicator the library owns" (openbnsllib.nccl_comm())

Comm make_comm(const py::object& o) {
  Comm c;
  if (o.is_none()) return c;
  if (py::isinstance<NcclCommTag>(o)) return native_comm();
  c.rank = o.attr("rank").cast<int>();
  c.world = o.attr("world").cast<int>();
  if (c.world < 1 || c.rank < 0 || c.rank >= c.world) throw std::invalid_argument("bad comm rank/world");
  py::object red = o.attr("allreduce_min_i32"), gat = o.attr("allgather");
  c.allreduce_min_i32 = [red](int32_t* p, size_t n) { red(view<int32_t>(p, n)); };
  if (py::hasattr(o, "allgather_dev")) {
    py::object gd = o.attr("allgather_dev");
    c.allgather_dev = [gd](uintptr_t p, size_t bytes) { gd((size_t)p, bytes); };
  }
  c.allgather = [gat, w = c.world](const void* s, void* r, size_t bytes) {
    gat(view<uint8_t>(static_cast<const uint8_t*>(s), bytes), view<uint8_t>(static_cast<uint8_t*>(r), bytes * w));
  };
  return c;
}

// test seam: an evaluator implemented in Python (tests/ inject the oracle here
// to run the drivers and the sharding logic on machines without a GPU)
struct PyEvaluator : Evaluator {
  py::object obj;
  explicit PyEvaluator(py::object o) : obj(std::move(o)) {}
  void ci_pvalues(const TupleBatch& b, size_t begin, size_t end, int kind, double* p) override {
    if (end <= begin) return;
    std::vector<uint64_t> off(end - begin + 1);
    for (size_t i = begin; i <= end; ++i) off[i - begin] = b.offsets[i] - b.offsets[begin];
    py::array_t<double> r = obj.attr("ci_pvalues")(
        view<uint32_t>(b.var_ids.data() + b.offsets[begin], b.offsets[end] - b.offsets[begin]),
        view<uint64_t>(off.data(), off.size()), view<uint32_t>(b.x.data() + begin, end - begin),
        view<uint32_t>(b.y.data() + begin, end - begin), kind);
    if ((size_t)r.size() != end - begin) throw std::runtime_error("evaluator returned a wrong-sized array");
    std::memcpy(p, r.data(), (end - begin) * sizeof(double));
  }
  // "asynchronous" rounds for the test seam: the answer is computed at submit and handed out at collect (FIFO of
  // two, like the library), so the driver's pipelining runs on machines without a GPU.  `chain` is honoured the
  // way the device does it: tests of edges the previously submitted round decided are not evaluated.  A Python
  // evaluator with a `submitted` attribute counts the rounds, one with a `skipped` attribute the tests left out.
  std::deque<std::vector<int32_t>> rounds_;
  std::vector<int32_t> last_first_;
  bool pc_round_submit(const TupleBatch& b, const std::vector<uint32_t>& slot_of_test, size_t n_slots, int kind,
                       double level, bool shard, bool chain) override {
    if (shard) return false;                    // the seam has no device-side communicator
    if (rounds_.size() == 2) throw std::runtime_error("three rounds in flight");
    if (chain && last_first_.size() != n_slots) throw std::runtime_error("chain: no previous round over the same edge slots");
    TupleBatch sub;
    sub.clear();
    std::vector<uint32_t> ids;
    for (size_t t = 0; t < b.size(); ++t) {
      if (chain && last_first_[slot_of_test[t]] != INT32_MAX) continue;
      sub.var_ids.insert(sub.var_ids.end(), b.var_ids.begin() + b.offsets[t], b.var_ids.begin() + b.offsets[t + 1]);
      sub.offsets.push_back(sub.var_ids.size());
      sub.x.push_back(b.x[t]);
      sub.y.push_back(b.y[t]);
      ids.push_back((uint32_t)t);
    }
    std::vector<double> p(sub.size());
    ci_pvalues(sub, 0, sub.size(), kind, p.data());
    std::vector<int32_t> first(n_slots, INT32_MAX);
    for (size_t i = 0; i < ids.size(); ++i)
      if (p[i] >= level) {                      // citest.h:147
        int32_t& f = first[slot_of_test[ids[i]]];
        if ((int32_t)ids[i] < f) f = (int32_t)ids[i];
      }
    last_first_ = first;
    rounds_.push_back(std::move(first));
    if (py::hasattr(obj, "submitted")) obj.attr("submitted") = obj.attr("submitted").cast<size_t>() + 1;
    if (py::hasattr(obj, "skipped")) obj.attr("skipped") = obj.attr("skipped").cast<size_t>() + (b.size() - ids.size());
    return true;
  }
  void pc_round_collect(std::vector<int32_t>& first) override {
    if (rounds_.empty()) throw std::runtime_error("collect without a round in flight");
    first = std::move(rounds_.front());
    rounds_.pop_front();
  }
  void subset_scores(const TupleBatch& b, size_t begin, size_t end, double ess, int exact, double* out) override {
    if (end <= begin) return;
    std::vector<uint64_t> off(end - begin + 1);
    for (size_t i = begin; i <= end; ++i) off[i - begin] = b.offsets[i] - b.offsets[begin];
    const size_t nv = b.offsets[end] - b.offsets[begin];
    py::array_t<double> r = obj.attr("subset_scores")(view<uint32_t>(b.var_ids.data() + b.offsets[begin], nv),
                                                      view<uint64_t>(off.data(), off.size()), ess, exact);
    if ((size_t)r.size() != nv) throw std::runtime_error("evaluator returned a wrong-sized array");
    std::memcpy(out, r.data(), nv * sizeof(double));
  }
  void child_scores(const TupleBatch& b, const uint32_t* child, double ess, int exact, double* out) override {
    if (!b.size()) return;
    py::array_t<double> r = obj.attr("child_scores")(view<uint32_t>(b.var_ids.data(), b.var_ids.size()),
                                                     view<uint64_t>(b.offsets.data(), b.offsets.size()),
                                                     view<uint32_t>(child, b.size()), ess, exact);
    if ((size_t)r.size() != b.size()) throw std::runtime_error("evaluator returned a wrong-sized array");
    std::memcpy(out, r.data(), b.size() * sizeof(double));
  }
};

PDAG run_pc(size_t n_vars, Evaluator* ev, const CITestType& test, size_t max_cond_vars, bool stable,
            const py::object& comm, size_t target_batch, bool record = false, size_t shard_min_slots = 0,
            bool overlap = false) {
  g_last_pc_stats.record = record;
  if (std::holds_alternative<OracleGraph>(test))
    return pc_oracle(n_vars, std::get<OracleGraph>(test).graph, max_cond_vars, stable);
  PcOptions o;
  o.max_cond_vars = max_cond_vars;
  o.stable = stable;
  if (std::holds_alternative<ChiSquare>(test)) {
    o.kind = OBN_CI_CHI2;
    o.level = std::get<ChiSquare>(test).level;
  } else {
    o.kind = OBN_CI_G2;
    o.level = std::get<GSquare>(test).level;
  }
  if (target_batch) o.target_batch = target_batch;
  o.shard_min_slots = shard_min_slots;
  o.overlap = overlap;
  const Comm c = make_comm(comm);
  return pc_skeleton_and_colliders(n_vars, *ev, c, o, &g_last_pc_stats);
}

PDAG run_es(size_t n_vars, Evaluator* ev, const ScoreType& score, size_t max_parents, int exact_order,
            const py::object& comm) {
  if (!std::holds_alternative<BDeu>(score)) throw std::invalid_argument("Unsupported score type.");
  EsOptions o;
  o.ess = std::get<BDeu>(score).ess;
  o.max_parents = max_parents;
  o.exact_order = exact_order;
  const Comm c = make_comm(comm);
  g_last_es_stats = EsStats{};
  return exhaustive_search_driver(n_vars, *ev, c, o, &g_last_es_stats);
}

SaStats g_last_sa_stats;
PDAG run_sa(size_t n_vars, Evaluator* ev, const ScoreType& score, size_t max_parents, size_t max_iters, double init_temp,
            double cooling_rate, bool is_deterministic, uint64_t seed, size_t num_chains, const py::object& comm) {
  if (max_parents >= n_vars) throw std::invalid_argument("max_parents out of range");     // simulated_annealing.cpp:109-111
  if (!std::holds_alternative<BDeu>(score)) throw std::invalid_argument("Unsupported score type.");   // local_score.h:43
  SaOptions o;
  o.ess = std::get<BDeu>(score).ess;
  o.max_parents = max_parents;
  o.max_iters = max_iters;
  o.init_temp = init_temp;
  o.cooling_rate = cooling_rate;
  o.deterministic = is_deterministic;
  o.seed = seed;
  o.num_chains = num_chains;
  const Comm c = make_comm(comm);
  g_last_sa_stats = SaStats{};
  return simulated_annealing_driver(n_vars, *ev, c, o, &g_last_sa_stats);
}

}  // namespace

PYBIND11_MODULE(openbnsllib, m) {
  m.doc() = "B200-native drop-in for OpenBNSL's contingency-counting hot path.";
  m.attr("cuda_enabled") = true;
  m.def("device", []() { return Engine::instance().device(); }, "CUDA device of this process (creates the context)");
  m.def("stream_handle", []() { return (uintptr_t)obn_ctx_stream(Engine::instance().ctx()); },
        "cudaStream_t of the library's main stream (for CUDA-event timing from Python, e.g. torch.cuda.ExternalStream)");

  // ---- the library's own communicator (NCCL below the C boundary) ----
  m.def("comm_unique_id", []() {
    uint8_t id[OBN_COMM_ID_BYTES];
    check(obn_comm_unique_id(id));
    return py::bytes(reinterpret_cast<const char*>(id), sizeof(id));
  }, "rank 0: the id every rank passes to comm_init (hand it over by any host-side means)");
  m.def("comm_init", [](const py::bytes& id, int rank, int world) {
    const std::string s = id;
    if (s.size() != OBN_COMM_ID_BYTES) throw std::invalid_argument("id must be OBN_COMM_ID_BYTES bytes");
    py::gil_scoped_release nogil;
    check(obn_comm_init(Engine::instance().ctx(), reinterpret_cast<const uint8_t*>(s.data()), rank, world));
  }, py::arg("id"), py::arg("rank"), py::arg("world"), "collective: create the NCCL communicator of this process's ctx");
  m.def("comm_destroy", []() { obn_comm_destroy(Engine::instance().ctx()); });
  py::class_<NcclCommTag>(m, "NcclComm", "comm= argument selecting the library's own NCCL communicator")
      .def_property_readonly("rank", [](const NcclCommTag&) { return obn_comm_rank(Engine::instance().ctx()); })
      .def_property_readonly("world", [](const NcclCommTag&) { return obn_comm_world(Engine::instance().ctx()); });
  m.def("nccl_comm", []() { return NcclCommTag{}; });
  m.def("set_option", [](const std::string& name, double value) {
    check(obn_ctx_set_option(Engine::instance().ctx(), name.c_str(), value));
  }, py::arg("name"), py::arg("value"), "obn_ctx_set_option on this process's ctx");

  m.def("profile_enable", [](bool on) { check(obn_profile_enable(Engine::instance().ctx(), on ? 1 : 0)); });
  m.def("profile_read", [](bool reset) {
    obn_profile p;
    check(obn_profile_read(Engine::instance().ctx(), &p, reset ? 1 : 0));
    py::dict d;
    d["count_kernel_ms"] = p.count_kernel_ms;
    d["count_launches"] = p.count_launches;
    d["total_launches"] = p.total_launches;
    d["algorithmic_bytes"] = p.algorithmic_bytes;
    d["tuples"] = p.tuples;
    d["host_prep_ms"] = p.host_prep_ms;
    d["host_wait_ms"] = p.host_wait_ms;
    d["h2d_bytes"] = p.h2d_bytes;
    d["d2h_bytes"] = p.d2h_bytes;
    d["scanned_bytes"] = p.scanned_bytes;
    d["increments"] = p.increments;
    d["tables"] = p.tables;
    d["collectives"] = p.collectives;
    d["collective_ms"] = p.collective_ms;
    d["exact_fallbacks"] = p.exact_fallbacks;
    return d;
  }, py::arg("reset") = true);

  // ---------------- base (bindings/submodules/base.cpp:7-51) ----------------
  auto base = m.def_submodule("base", "Base submodule");
  py::class_<DataframeWrapper>(base, "DataframeWrapper")
      .def(py::init<const py::object&>())
      .def_static("from_codes",
                  [](const py::array& codes, const std::vector<std::string>& names, const std::vector<size_t>& cards) {
                    return new DataframeWrapper(codes, names, cards);
                  },
                  py::arg("codes"), py::arg("names"), py::arg("cards"),
                  "Extension: wrap an already-coded uint8 [n_rows, n_vars] array")
      .def_readonly("num_of_vars", &DataframeWrapper::num_of_vars)
      .def_readonly("num_of_datapoints", &DataframeWrapper::num_of_datapoints)
      .def_readonly("col_idx2str", &DataframeWrapper::col_idx2str)
      .def_readonly("col_str2idx", &DataframeWrapper::col_str2idx)
      .def_readonly("val_idx2str", &DataframeWrapper::val_idx2str)
      .def_readonly("val_str2idx", &DataframeWrapper::val_str2idx)
      .def_readonly("num_of_values", &DataframeWrapper::num_of_values)
      .def_property_readonly("data_column_major", &DataframeWrapper::data_column_major)
      .def_property_readonly("data_row_major", &DataframeWrapper::data_row_major)
      .def("to_device", [](const DataframeWrapper& df) { df.device(); }, "Upload to HBM now (otherwise lazy)")
      .def("release_device", &DataframeWrapper::release_device)
      .def("__repr__", [](const DataframeWrapper& df) {
        std::string r = "DataframeWrapper(\n  num_of_datapoints: " + std::to_string(df.num_of_datapoints) +
                        ",\n  num_of_vars: " + std::to_string(df.num_of_vars) + ",\n  columns: [";
        for (size_t i = 0; i < df.col_idx2str.size(); ++i) r += (i ? ", " : "") + df.col_idx2str[i];
        return r + "]\n)";
      });
  py::class_<ContingencyTable>(base, "ContingencyTable")
      .def(py::init<const std::vector<size_t>&, const DataframeWrapper&>(), py::arg("var_ids"), py::arg("df"),
           py::keep_alive<1, 3>())
      .def_readonly("var_ids", &ContingencyTable::var_ids)
      .def_readonly("cardinalities", &ContingencyTable::cardinalities)
      .def_readonly("counts", &ContingencyTable::counts);
  py::class_<PDAG>(base, "PDAG")
      .def(py::init<size_t>())
      .def(py::init<const PDAG&>())
      .def_readonly("num_vars", &PDAG::num_vars)
      .def("has_edge", &PDAG::has_edge, py::arg("from"), py::arg("to"))
      .def("add_edge", &PDAG::add_edge, py::arg("from"), py::arg("to"))
      .def("remove_edge", &PDAG::remove_edge, py::arg("from"), py::arg("to"));

  // ---------------- score (bindings/submodules/score.cpp:6-29) --------------
  auto score = m.def_submodule("score", "Score submodule");
  py::class_<BDeu>(score, "BDeu").def(py::init<double>(), py::arg("ess") = 1.0)
      .def_readonly("ess", &BDeu::ess)
      .def("__repr__", [](const BDeu& b) { return "BDeu(" + std::to_string(b.ess) + ")"; });
  py::class_<K2>(score, "K2").def(py::init<>()).def("__repr__", [](const K2&) { return "K2()"; });
  py::class_<BIC>(score, "BIC").def(py::init<>()).def("__repr__", [](const BIC&) { return "BIC()"; });
  py::class_<AIC>(score, "AIC").def(py::init<>()).def("__repr__", [](const AIC&) { return "AIC()"; });
  score.def(
      "calculate_local_score",
      [](size_t child_var, const std::vector<size_t>& /*parent_set*/, const ContingencyTable& ct,
         const ScoreType& st) {
        // local_score.h:27-44: the child must be in the table; only BDeu has arithmetic
        if (std::find(ct.var_ids.begin(), ct.var_ids.end(), child_var) == ct.var_ids.end())
          throw std::invalid_argument("child_var not present in contingency table");
        if (!std::holds_alternative<BDeu>(st)) throw std::invalid_argument("Unsupported score type.");
        std::vector<uint32_t> v32(ct.var_ids.begin(), ct.var_ids.end());
        const uint64_t offs[2] = {0, (uint64_t)v32.size()};
        const uint32_t child = (uint32_t)child_var;
        double out = 0;
        check(obn_bdeu_scores(Engine::instance().ctx(), ct.df->device(), v32.data(), offs, 1, &child,
                              std::get<BDeu>(st).ess, /*exact_order=*/2, &out));
        return out;
      },
      py::arg("child_var"), py::arg("parent_set"), py::arg("ct"), py::arg("score_type"));

  // ---------------- citest (bindings/submodules/citest.cpp:8-33) ------------
  auto ci = m.def_submodule("citest", "CITests submodule");
  py::class_<ChiSquare>(ci, "ChiSquare").def(py::init<double>(), py::arg("level") = 0.05)
      .def_readonly("level", &ChiSquare::level)
      .def("__repr__", [](const ChiSquare& c) { return "ChiSquare(level=" + std::to_string(c.level) + ")"; });
  py::class_<GSquare>(ci, "GSquare").def(py::init<double>(), py::arg("level") = 0.05)
      .def_readonly("level", &GSquare::level)
      .def("__repr__", [](const GSquare& c) { return "GSquare(level=" + std::to_string(c.level) + ")"; });
  py::class_<OracleGraph>(ci, "OracleGraph").def(py::init<const PDAG&>(), py::arg("pdag"))
      .def("__repr__", [](const OracleGraph&) { return "<OracleGraph>"; });
  auto ci_one = [](size_t x, size_t y, const std::vector<size_t>& sepset, const ContingencyTable& ct,
                   const CITestType& test, double* stat, uint64_t* dof, double* p) {
    if (x == y) throw std::invalid_argument("x and y must differ");            // citest.h:47
    if (x >= y) std::swap(x, y);
    if (std::holds_alternative<OracleGraph>(test)) {
      *p = is_d_separated(std::get<OracleGraph>(test).graph, x, y, sepset) ? 1.0 : -1.0;
      return 0.0;
    }
    const bool chi = std::holds_alternative<ChiSquare>(test);
    const double level = chi ? std::get<ChiSquare>(test).level : std::get<GSquare>(test).level;
    std::vector<uint32_t> v32(ct.var_ids.begin(), ct.var_ids.end());
    const uint64_t offs[2] = {0, (uint64_t)v32.size()};
    const uint32_t xs = (uint32_t)x, ys = (uint32_t)y;
    check(obn_ci_tests(Engine::instance().ctx(), ct.df->device(), v32.data(), offs, 1, &xs, &ys,
                       chi ? OBN_CI_CHI2 : OBN_CI_G2, stat, dof, p));
    return level;
  };
  ci.def("citest",
         [ci_one](size_t x, size_t y, const std::vector<size_t>& sepset, const ContingencyTable& ct,
                  const CITestType& test) {
           double stat = 0, p = 0;
           uint64_t dof = 0;
           const double level = ci_one(x, y, sepset, ct, test, &stat, &dof, &p);
           return p >= level;                                                   // citest.h:147
         },
         py::arg("x"), py::arg("y"), py::arg("sepset_candidate"), py::arg("ct"), py::arg("citest_type"));
  ci.def("citest_statistic",
         [ci_one](size_t x, size_t y, const std::vector<size_t>& sepset, const ContingencyTable& ct,
                  const CITestType& test) {
           double stat = 0, p = 0;
           uint64_t dof = 0;
           ci_one(x, y, sepset, ct, test, &stat, &dof, &p);
           return py::make_tuple(stat, dof, p);
         },
         py::arg("x"), py::arg("y"), py::arg("sepset_candidate"), py::arg("ct"), py::arg("citest_type"),
         "Extension: (statistic, dof, p) that the reference keeps internal (citest.h:86-87,139)");

  // ------- structure_learning (bindings/submodules/structure_learning.cpp:9-44) -------
  auto sl = m.def_submodule("structure_learning", "Structure learning submodule");
  sl.def("exhaustive_search",
         [](const DataframeWrapper& df, const ScoreType& st, size_t max_parents, bool /*is_deterministic*/,
            int exact_order, const py::object& comm, bool host_dp, const py::object& lattice, double shard_min_bytes,
            size_t dp_shard_min) {
           if (!std::holds_alternative<BDeu>(st)) throw std::invalid_argument("Unsupported score type.");
           if (df.num_of_vars && max_parents > df.num_of_vars - 1)
             throw std::invalid_argument("max_parents must be in [0, num_of_vars-1]");
           if (host_dp) {                 // scores on the GPU, DP tables on the host (A/B and debugging)
             GpuEvaluator ev(df.device());
             return run_es(df.num_of_vars, &ev, st, max_parents, exact_order, comm);
           }
           EsOptions o;
           o.ess = std::get<BDeu>(st).ess;
           o.max_parents = max_parents;
           o.exact_order = exact_order;
           o.cards.assign(df.num_of_values.begin(), df.num_of_values.end());
           o.shard_min_bytes = shard_min_bytes;
           o.dp_shard_min_subsets = dp_shard_min;
           o.force_lattice = lattice.is_none() ? -1 : (lattice.cast<bool>() ? 1 : 0);
           const Comm c = make_comm(comm);
           if (c.world > 1 && !c.allgather_dev) throw std::invalid_argument("comm lacks allgather_dev");
           g_last_es_stats = EsStats{};
           return exhaustive_search_device(df.device(), c, o, &g_last_es_stats);
         },
         "Run exhaustive search algorithm", py::arg("df"), py::arg("score_type"), py::arg("max_parents"),
         py::arg("is_deterministic") = false, py::kw_only(), py::arg("exact_order") = 2,
         py::arg("comm") = py::none(), py::arg("host_dp") = false, py::arg("lattice") = py::none(),
         py::arg("shard_min_bytes") = 4e9, py::arg("dp_shard_min") = 4096);
  sl.def("pc",
         [](const DataframeWrapper& df, const CITestType& test, size_t max_cond_vars, bool stable,
            const py::object& comm, size_t target_batch, bool record_tests, bool apply_meek, bool complete_cpdag,
            bool overlap, double shard_min_bytes) {
           // apply_meek: what un-commenting pc.cpp:84 would do; complete_cpdag: the published rules
           auto finish = [&](PDAG g) {
             if (apply_meek) g = apply_meeks_rules(g, false, false);
             return complete_cpdag ? apply_meeks_rules(g, false, true) : g;
           };
           if (std::holds_alternative<OracleGraph>(test))
             return finish(run_pc(df.num_of_vars, nullptr, test, max_cond_vars, stable, comm, target_batch));
           GpuEvaluator ev(df.device());
           // shard a round over the GPUs only when it is worth a collective: >= shard_min_bytes of
           // algorithmic bytes (default 1 GB, about 0.15 ms of one B200)
           const size_t min_slots = (size_t)(shard_min_bytes / (double)std::max<size_t>(1, df.num_of_datapoints));
           return finish(run_pc(df.num_of_vars, &ev, test, max_cond_vars, stable, comm, target_batch, record_tests,
                                min_slots, overlap));
         },
         "Run PC algorithm", py::arg("df"), py::arg("ci_test_type"), py::arg("max_cond_vars"),
         py::arg("stable") = true, py::kw_only(), py::arg("comm") = py::none(), py::arg("target_batch") = 0,
         py::arg("record_tests") = false, py::arg("apply_meek") = false, py::arg("complete_cpdag") = false,
         py::arg("overlap") = true, py::arg("shard_min_bytes") = 1e9);
  // the reference's Meek rules (pdag_with_adjmat.cpp:215-279) exist but pc() leaves them
  // switched off (pc.cpp:84) and PDAGwithAdjMat is not bound; here they are reachable, off by default.
  sl.def("apply_meeks_rules",
         [](const PDAG& g, bool apply_r4, bool textbook) { return apply_meeks_rules(g, apply_r4, textbook); },
         "Meek's orientation rules as PDAGwithAdjMat::apply_meeks_rules states them (textbook=False), or as "
         "published, which completes a pattern to its CPDAG (textbook=True); returns a new PDAG",
         py::arg("pdag"), py::arg("apply_r4") = false, py::arg("textbook") = false);
  sl.def("last_pc_tests", []() {
    const TupleBatch& r = g_last_pc_stats.recorded;
    auto cp = [](const auto& v) { using T = typename std::decay_t<decltype(v)>::value_type;
                                  py::array_t<T> a(v.size()); if (v.size()) std::memcpy(a.mutable_data(), v.data(), v.size() * sizeof(T)); return a; };
    return py::make_tuple(cp(r.var_ids), cp(r.offsets), cp(r.x), cp(r.y));
  }, "(var_ids, offsets, x, y) of every CI test the last pc(record_tests=True) evaluated");
  sl.def("simulated_annealing",
         [](const DataframeWrapper& df, const ScoreType& st, size_t max_parents, size_t max_iters, double init_temp,
            double cooling_rate, bool is_deterministic, uint64_t seed, size_t num_chains, const py::object& comm) {
           if (max_parents >= df.num_of_vars) throw std::invalid_argument("max_parents out of range");
           if (!std::holds_alternative<BDeu>(st)) throw std::invalid_argument("Unsupported score type.");
           GpuEvaluator ev(df.device());
           return run_sa(df.num_of_vars, &ev, st, max_parents, max_iters, init_temp, cooling_rate, is_deterministic, seed,
                         num_chains, comm);
         },
         "Run simulated annealing algorithm", py::arg("df"), py::arg("score_type"), py::arg("max_parents"),
         py::arg("max_iters") = 10000, py::arg("init_temp") = 1.0, py::arg("cooling_rate") = 0.9995,
         py::arg("is_deterministic") = false, py::arg("seed") = 0, py::arg("num_chains") = 0, py::kw_only(),
         py::arg("comm") = py::none());
  sl.def("last_sa_stats", []() {
    py::dict d;
    d["chains"] = g_last_sa_stats.chains;
    d["proposals"] = g_last_sa_stats.proposals;
    d["evaluated"] = g_last_sa_stats.evaluated;
    d["batches"] = g_last_sa_stats.batches;
    d["best_chain"] = g_last_sa_stats.best_chain;
    d["best_score"] = g_last_sa_stats.best_score;
    d["evaluate_ms"] = g_last_sa_stats.evaluate_ms;
    return d;
  });
  sl.def("rai", [](py::args, py::kwargs) -> PDAG {
    throw std::runtime_error("rai is an unimplemented stub upstream and outside the B200 hot-path scope");
  });
  sl.def("last_pc_stats", []() {
    py::dict d;
    d["tests_per_depth"] = g_last_pc_stats.tests_per_depth;
    d["batches_per_depth"] = g_last_pc_stats.batches_per_depth;
    d["total_tests"] = g_last_pc_stats.total_tests();
    d["enumerate_ms"] = g_last_pc_stats.enumerate_ms;
    d["evaluate_ms"] = g_last_pc_stats.evaluate_ms;
    d["resolve_ms"] = g_last_pc_stats.resolve_ms;
    d["exposed_ms"] = g_last_pc_stats.exposed_ms;
    d["collect_wait_ms"] = g_last_pc_stats.collect_wait_ms;
    return d;
  });
  sl.def("last_es_stats", []() {
    py::dict d;
    d["tables"] = g_last_es_stats.tables;
    d["scores"] = g_last_es_stats.scores;
    d["best_score"] = g_last_es_stats.best_score;
    d["enumerate_ms"] = g_last_es_stats.enumerate_ms;
    d["score_ms"] = g_last_es_stats.score_ms;
    d["gather_ms"] = g_last_es_stats.gather_ms;
    d["dp_ms"] = g_last_es_stats.dp_ms;
    d["lattice"] = g_last_es_stats.lattice;
    return d;
  });

  // ---------------- test seam (drivers over an injected evaluator) -----------
  auto ts = m.def_submodule("_testing", "Driver entry points over an injected evaluator (CPU test-suite only)");
  ts.def("pc", [](size_t n_vars, py::object evaluator, const CITestType& test, size_t max_cond_vars, bool stable,
                  const py::object& comm, size_t target_batch, bool record_tests, bool overlap) {
    PyEvaluator ev(std::move(evaluator));
    return run_pc(n_vars, &ev, test, max_cond_vars, stable, comm, target_batch, record_tests, 0, overlap);
  }, py::arg("n_vars"), py::arg("evaluator"), py::arg("ci_test_type"), py::arg("max_cond_vars"),
     py::arg("stable") = true, py::arg("comm") = py::none(), py::arg("target_batch") = 0,
     py::arg("record_tests") = false, py::arg("overlap") = false);
  ts.def("exhaustive_search", [](size_t n_vars, py::object evaluator, const ScoreType& st, size_t max_parents,
                                 int exact_order, const py::object& comm) {
    PyEvaluator ev(std::move(evaluator));
    return run_es(n_vars, &ev, st, max_parents, exact_order, comm);
  }, py::arg("n_vars"), py::arg("evaluator"), py::arg("score_type"), py::arg("max_parents"),
     py::arg("exact_order") = 2, py::arg("comm") = py::none());
  ts.def("simulated_annealing", [](size_t n_vars, py::object evaluator, const ScoreType& st, size_t max_parents,
                                   size_t max_iters, double init_temp, double cooling_rate, bool is_deterministic,
                                   uint64_t seed, size_t num_chains, const py::object& comm) {
    PyEvaluator ev(std::move(evaluator));
    return run_sa(n_vars, &ev, st, max_parents, max_iters, init_temp, cooling_rate, is_deterministic, seed, num_chains, comm);
  }, py::arg("n_vars"), py::arg("evaluator"), py::arg("score_type"), py::arg("max_parents"), py::arg("max_iters") = 10000,
     py::arg("init_temp") = 1.0, py::arg("cooling_rate") = 0.9995, py::arg("is_deterministic") = false,
     py::arg("seed") = 0, py::arg("num_chains") = 0, py::arg("comm") = py::none());
}
