#include "dataframe_wrapper.hpp"

#include <algorithm>
#include <cstring>
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

// Ranks of the distinct values of one column.  The reference inserts every cell's string into a
// std::set and later looks every cell up in a std::map (dataframe_wrapper.cpp:36-70: two string
// casts and two tree walks per cell); here a cell costs one probe of a small direct-mapped cache
// keyed by what identifies the value cheaply (the PyObject pointer, the integer itself), the
// distinct values get first-seen ids, and ONE 256-entry table turns ids into ranks at the end.
struct ColumnCoder {
  std::vector<std::string> seen;              // distinct strings in first-seen order (<= 256)
  int16_t slot[1024];                         // open addressing over `seen`: -1 = empty
  ColumnCoder() { std::fill(slot, slot + 1024, (int16_t)-1); }
  uint32_t id(const char* p, size_t len, DataframeWrapper& w, size_t v) {
    uint64_t h = 1469598103934665603ull;      // FNV-1a
    for (size_t i = 0; i < len; ++i) h = (h ^ (unsigned char)p[i]) * 1099511628211ull;
    size_t q = (size_t)(h >> 32) & 1023;
    for (; slot[q] >= 0; q = (q + 1) & 1023) {
      const std::string& s = seen[(size_t)slot[q]];
      if (s.size() == len && std::memcmp(s.data(), p, len) == 0) return (uint32_t)slot[q];
    }
    if (seen.size() == 256)
      throw std::invalid_argument("column '" + w.col_idx2str[v] + "' has more than 256 distinct values (uint8 codes)");
    slot[q] = (int16_t)seen.size();
    seen.emplace_back(p, len);
    return (uint32_t)slot[q];
  }
  // col holds first-seen ids; rewrite them as ranks among the sorted distinct strings
  void finish(DataframeWrapper& w, size_t v, dtype* col, size_t n) {
    std::set<std::string> uniq(seen.begin(), seen.end());
    finish_column(w, v, uniq);
    dtype rank[256];
    for (size_t i = 0; i < seen.size(); ++i) rank[i] = w.val_str2idx[v][seen[i]];
    for (size_t r = 0; r < n; ++r) col[r] = rank[col[r]];
  }
};

constexpr size_t kCacheSlots = 1024;          // direct-mapped; a miss only costs the slow path

template <typename Int>
void ingest_integers(DataframeWrapper& w, const py::array& arr) {
  const auto* base = static_cast<const unsigned char*>(arr.data());
  const ssize_t s0 = arr.strides(0), s1 = arr.strides(1);
  const size_t n = w.num_of_datapoints;
  std::vector<Int> key(kCacheSlots);
  std::vector<int32_t> val(kCacheSlots);
  for (size_t v = 0; v < w.num_of_vars; ++v) {
    ColumnCoder cc;
    std::fill(val.begin(), val.end(), -1);
    dtype* col = w.colmajor.data() + v * n;
    for (size_t r = 0; r < n; ++r) {
      const Int x = *reinterpret_cast<const Int*>(base + r * s0 + v * s1);
      const size_t h = (size_t)((uint64_t)x * 0x9E3779B97F4A7C15ull >> 54) & (kCacheSlots - 1);
      if (val[h] < 0 || key[h] != x) {
        key[h] = x;
        const std::string d = std::to_string(x);              // decimal string = .astype(str) of the reference's callers
        val[h] = (int32_t)cc.id(d.data(), d.size(), w, v);
      }
      col[r] = (dtype)val[h];
    }
    cc.finish(w, v, col, n);
  }
}

// Object cells (dataframe_wrapper.cpp:42,67).  A str is read through its cached UTF-8 form (no
// copy for ASCII); anything else goes through the same pybind11 string cast as before, so bytes
// still work and other objects raise the reference's error.  Equal strings are usually the SAME
// object (a frame built from categories, an interned literal, `table[codes]`), which the pointer
// cache turns into one pointer compare per cell; distinct objects with equal text still meet in
// the string table.
void ingest_objects(DataframeWrapper& w, const py::array& arr) {
  const auto* base = static_cast<const unsigned char*>(arr.data());
  const ssize_t s0 = arr.strides(0), s1 = arr.strides(1);
  const size_t n = w.num_of_datapoints;
  std::vector<PyObject*> key(kCacheSlots);
  std::vector<uint32_t> val(kCacheSlots);
  std::string tmp;
  for (size_t v = 0; v < w.num_of_vars; ++v) {
    ColumnCoder cc;
    std::fill(key.begin(), key.end(), nullptr);
    dtype* col = w.colmajor.data() + v * n;
    for (size_t r = 0; r < n; ++r) {
      PyObject* o = *reinterpret_cast<PyObject* const*>(base + r * s0 + v * s1);
      const size_t h = (size_t)(((uintptr_t)o >> 4) * 0x9E3779B97F4A7C15ull >> 54) & (kCacheSlots - 1);
      if (key[h] != o) {
        if (o && PyUnicode_Check(o)) {
          Py_ssize_t len = 0;
          const char* u = PyUnicode_AsUTF8AndSize(o, &len);
          if (!u) {
            PyErr_Clear();
            throw std::invalid_argument("Failed to cast value to string");
          }
          val[h] = cc.id(u, (size_t)len, w, v);
        } else {
          try {
            tmp = py::handle(o).cast<std::string>();
          } catch (const std::exception&) {
            throw std::invalid_argument("Failed to cast value to string");
          }
          val[h] = cc.id(tmp.data(), tmp.size(), w, v);
        }
        key[h] = o;                           // the array keeps the object alive for the whole scan
      }
      col[r] = (dtype)val[h];
    }
    cc.finish(w, v, col, n);
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
    ingest_objects(*this, arr);             // strides are honoured: C- and Fortran-ordered object arrays both work
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
