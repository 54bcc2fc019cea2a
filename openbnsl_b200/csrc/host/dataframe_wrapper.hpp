// dataframe_wrapper.hpp -- host mirror of the reference's data layer:
//   DataframeWrapper   include/base/dataframe_wrapper.h:24-46, src/base/dataframe_wrapper.cpp:5-98
//   ContingencyTable   include/base/contingency_table.h:22-86
// Same attribute names and meanings; the sample matrix additionally lives in
// HBM as ONE padded column-major uint8 matrix (uploaded lazily, once).
#pragma once
#include <pybind11/numpy.h>
#include <pybind11/pybind11.h>

#include <map>
#include <memory>
#include <string>
#include <vector>

#include "engine.hpp"

namespace py = pybind11;

namespace obn {

using dtype = uint8_t;

struct DataframeWrapper {
  explicit DataframeWrapper(const py::object& dataframe);
  // pre-coded fast path: codes[n_rows][n_vars] uint8 (any strides), optional names / cards
  DataframeWrapper(const py::array& codes, const std::vector<std::string>& names,
                   const std::vector<size_t>& cards);

  size_t num_of_vars = 0;
  size_t num_of_datapoints = 0;
  std::vector<std::string> col_idx2str;
  std::map<std::string, size_t> col_str2idx;
  std::vector<std::vector<std::string>> val_idx2str;
  std::vector<std::map<std::string, dtype>> val_str2idx;
  std::vector<size_t> num_of_values;

  // contiguous column-major codes: column v at colmajor[v * num_of_datapoints ...)
  std::vector<dtype> colmajor;

  // views with the reference's attribute shapes (built on demand for Python)
  std::vector<std::vector<dtype>> data_column_major() const;
  std::vector<std::vector<dtype>> data_row_major() const;

  // device residency (lazy; throws if no GPU is usable -- no CPU fallback)
  obn_dataset* device() const;
  void release_device() const { dev_.reset(); }
  ~DataframeWrapper();
  DataframeWrapper(const DataframeWrapper&) = delete;
  DataframeWrapper& operator=(const DataframeWrapper&) = delete;

 private:
  mutable std::unique_ptr<DeviceDataset> dev_;
  mutable bool pinned_ = false;   // colmajor is page-locked (obn_host_pin) from the first upload on
};

struct ContingencyTable {
  std::vector<size_t> var_ids;
  std::vector<size_t> cardinalities;
  std::vector<size_t> radix_weights;
  std::map<size_t, size_t> counts;       // linear index -> frequency, observed cells only
  const DataframeWrapper* df;            // kept alive by the binding (keep_alive)

  ContingencyTable(const std::vector<size_t>& var_ids, const DataframeWrapper& df);
};

}  // namespace obn
