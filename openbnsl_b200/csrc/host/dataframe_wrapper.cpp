#include "dataframe_wrapper.hpp"

#include <algorithm>
#include <set>
#include <stdexcept>

namespace obn {

namespace {

// assign codes = rank among lexicographically sorted unique strings
// (src/base/dataframe_wrapper.cpp:36-52,63-70)
void finish_column(DataframeWrapper& w, size_t v, const std::set<std::string>& uniq) {
  if (uniq.size() > 256)
    throw std::invalid_argument("column '" + w.col_idx2str[v] + "' has more than 256 distinct values (uint8 codes)");
  for (const auto& s : uniq) {
    w.val_str2idx[v][s] = (dtype)w.val_idx2str[v].size();
    w.val_idx2str[v].push_back(s);
  }
  w.num_of_values[v] = w.val_idx2str[v].size();
}

void read_columns(DataframeWrapper& w, const py::object& dataframe) {
  if (!py::hasattr(dataframe, "columns")) throw std::invalid_argument("Input must be a pandas dataframe");
  for (auto column : dataframe.attr("columns")) {
    std::string s;
    try {
      s = column.cast<std::string>();
    } catch (const std::exception&) {
      throw std::invalid_argument("Failed to cast column name to string");
    }
    w.col_str2idx[s] = w.col_idx2str.size();
    w.col_idx2str.push_back(s);
  }
}

void alloc(DataframeWrapper& w) {
  w.val_str2idx.assign(w.num_of_vars, {});
  w.val_idx2str.assign(w.num_of_vars, {});
  w.num_of_values.assign(w.num_of_vars, 0);
  w.colmajor.assign(w.num_of_vars * w.num_of_datapoints, 0);
}

template <typename Int>
void ingest_integers(DataframeWrapper& w, const py::array& arr) {
  const auto* base = static_cast<const unsigned char*>(arr.data());
  const ssize_t s0 = arr.strides(0), s1 = arr.strides(1);
  const size_t n = w.num_of_datapoints;
  for (size_t v = 0; v < w.num_of_vars; ++v) {
    std::set<long long> vals;
    for (size_t r = 0; r < n; ++r) vals.insert((long long)*reinterpret_cast<const Int*>(base + r * s0 + v * s1));
    std::set<std::string> uniq;
    for (auto x : vals) uniq.insert(std::to_string(x));
    finish_column(w, v, uniq);
    std::map<long long, dtype> code;
    for (auto x : vals) code[x] = w.val_str2idx[v][std::to_string(x)];
    dtype* col = w.colmajor.data() + v * n;
    for (size_t r = 0; r < n; ++r) col[r] = code[(long long)*reinterpret_cast<const Int*>(base + r * s0 + v * s1)];
  }
}

}  // namespace

DataframeWrapper::DataframeWrapper(const py::object& dataframe) {
  read_columns(*this, dataframe);
  if (!py::hasattr(dataframe, "values")) throw std::invalid_argument("Input must be a pandas dataframe");
  py::array arr = py::array::ensure(dataframe.attr("values"));
  if (!arr) throw std::invalid_argument("Input must be a 2D array");
  if (arr.ndim() != 2) throw std::invalid_argument("Input must be a 2D array");
  num_of_datapoints = (size_t)arr.shape(0);
  num_of_vars = (size_t)arr.shape(1);
  if (col_idx2str.size() != num_of_vars) throw std::invalid_argument("columns / values shape mismatch");
  alloc(*this);
  const char kind = arr.dtype().kind();
  if (kind == 'O') {
    // Python str cells (dataframe_wrapper.cpp:42,67).  Strides are honoured, so
    // C- and Fortran-ordered object arrays both work.
    const auto* base = static_cast<const unsigned char*>(arr.data());
    const ssize_t s0 = arr.strides(0), s1 = arr.strides(1);
    const size_t n = num_of_datapoints;
    for (size_t v = 0; v < num_of_vars; ++v) {
      std::set<std::string> uniq;
      std::vector<std::string> cells(n);
      for (size_t r = 0; r < n; ++r) {
        PyObject* o = *reinterpret_cast<PyObject* const*>(base + r * s0 + v * s1);
        try {
          cells[r] = py::handle(o).cast<std::string>();
        } catch (const std::exception&) {
          throw std::invalid_argument("Failed to cast value to string");
        }
        uniq.insert(cells[r]);
      }
      finish_column(*this, v, uniq);
      dtype* col = colmajor.data() + v * n;
      for (size_t r = 0; r < n; ++r) col[r] = val_str2idx[v][cells[r]];
    }
  } else if (kind == 'i' || kind == 'u' || kind == 'b') {
    // extension: integer-coded frames; values are ranked by their decimal
    // strings, i.e. exactly what the reference computes after .astype(str)
    switch (arr.itemsize()) {
      case 1: kind == 'i' ? ingest_integers<int8_t>(*this, arr) : ingest_integers<uint8_t>(*this, arr); break;
      case 2: kind == 'i' ? ingest_integers<int16_t>(*this, arr) : ingest_integers<uint16_t>(*this, arr); break;
      case 4: kind == 'i' ? ingest_integers<int32_t>(*this, arr) : ingest_integers<uint32_t>(*this, arr); break;
      case 8: kind == 'i' ? ingest_integers<int64_t>(*this, arr) : ingest_integers<uint64_t>(*this, arr); break;
      default: throw std::invalid_argument("unsupported integer width");
    }
  } else {
    throw std::invalid_argument("Failed to cast value to string");
  }
}

DataframeWrapper::DataframeWrapper(const py::array& codes, const std::vector<std::string>& names,
                                   const std::vector<size_t>& cards) {
  if (codes.ndim() != 2) throw std::invalid_argument("Input must be a 2D array");
  if (codes.dtype().kind() != 'u' || codes.itemsize() != 1)
    throw std::invalid_argument("from_codes expects a uint8 array");
  num_of_datapoints = (size_t)codes.shape(0);
  num_of_vars = (size_t)codes.shape(1);
  if (names.size() != num_of_vars || cards.size() != num_of_vars)
    throw std::invalid_argument("names / cards length mismatch");
  col_idx2str = names;
  for (size_t v = 0; v < num_of_vars; ++v) col_str2idx[names[v]] = v;
  alloc(*this);
  const auto* base = static_cast<const unsigned char*>(codes.data());
  const ssize_t s0 = codes.strides(0), s1 = codes.strides(1);
  for (size_t v = 0; v < num_of_vars; ++v) {
    if (cards[v] == 0 || cards[v] > 256) throw std::invalid_argument("cardinality must be in [1,256]");
    num_of_values[v] = cards[v];
    const int width = (int)std::to_string(cards[v] - 1).size();
    for (size_t s = 0; s < cards[v]; ++s) {
      std::string d = std::to_string(s);
      std::string name = "s" + std::string(width - (int)d.size(), '0') + d;
      val_str2idx[v][name] = (dtype)s;
      val_idx2str[v].push_back(name);
    }
    dtype* col = colmajor.data() + v * num_of_datapoints;
    for (size_t r = 0; r < num_of_datapoints; ++r) {
      const dtype c = *(base + r * s0 + v * s1);
      if (c >= cards[v]) throw std::invalid_argument("code out of range for its cardinality");
      col[r] = c;
    }
  }
}

std::vector<std::vector<dtype>> DataframeWrapper::data_column_major() const {
  std::vector<std::vector<dtype>> out(num_of_vars);
  for (size_t v = 0; v < num_of_vars; ++v)
    out[v].assign(colmajor.begin() + v * num_of_datapoints, colmajor.begin() + (v + 1) * num_of_datapoints);
  return out;
}
std::vector<std::vector<dtype>> DataframeWrapper::data_row_major() const {
  std::vector<std::vector<dtype>> out(num_of_datapoints, std::vector<dtype>(num_of_vars));
  for (size_t v = 0; v < num_of_vars; ++v)
    for (size_t r = 0; r < num_of_datapoints; ++r) out[r][v] = colmajor[v * num_of_datapoints + r];
  return out;
}

DataframeWrapper::~DataframeWrapper() {
  dev_.reset();
  if (pinned_) obn_host_unpin(Engine::instance().ctx(), colmajor.data());
}

obn_dataset* DataframeWrapper::device() const {
  if (!dev_) {
    if (!pinned_ && !colmajor.empty())     // best effort: an unpinned upload is only slower
      pinned_ = obn_host_pin(Engine::instance().ctx(), colmajor.data(), colmajor.size()) == OBN_OK;
    std::vector<uint32_t> cards(num_of_values.begin(), num_of_values.end());
    dev_.reset(new DeviceDataset(colmajor.data(), num_of_vars, num_of_datapoints, num_of_datapoints, cards.data()));
  }
  return dev_->handle();
}

// contingency_table.h:34-56 -- counted on the GPU through the C ABI
ContingencyTable::ContingencyTable(const std::vector<size_t>& ids, const DataframeWrapper& d)
    : var_ids(ids), df(&d) {
  if (!std::is_sorted(var_ids.begin(), var_ids.end())) throw std::invalid_argument("var_ids must be sorted");
  radix_weights.resize(var_ids.size());
  cardinalities.resize(var_ids.size());
  size_t mult = 1;
  for (int i = (int)var_ids.size() - 1; i >= 0; --i) {
    if (var_ids[i] >= d.num_of_vars) throw std::out_of_range("variable id out of range");
    cardinalities[i] = d.num_of_values[var_ids[i]];
    radix_weights[i] = mult;
    mult *= cardinalities[i];
  }
  std::vector<uint32_t> v32(var_ids.begin(), var_ids.end());
  const uint64_t offs[2] = {0, (uint64_t)v32.size()};
  const uint64_t out_off[1] = {0};
  if (mult >= ((size_t)1 << 31)) throw std::runtime_error("dense table has >= 2^31 cells");
  std::vector<int32_t> dense(mult);
  check(obn_count_tables(Engine::instance().ctx(), d.device(), v32.data(), offs, 1, dense.data(), out_off));
  for (size_t key = 0; key < mult; ++key)
    if (dense[key]) counts.emplace_hint(counts.end(), key, (size_t)dense[key]);
}

}  // namespace obn
