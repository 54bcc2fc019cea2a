// graph.hpp -- host-side graph containers with the reference's semantics.
//   PDAG              include/graph/pdag.h:16-37 (result type of every algorithm)
//   AdjMatGraph       the subset of PDAGwithAdjMat that PC uses
//                     (src/graph/pdag_with_adjmat.cpp:7-11,27-36,61-69,72-117,171-189)
//   is_d_separated    include/citest/graph_dsep.h:21-113 (OracleGraph CI test)
#pragma once
#include <algorithm>
#include <cstddef>
#include <cstdint>
#include <queue>
#include <set>
#include <stdexcept>
#include <vector>

namespace obn {

struct PDAG {
  std::size_t num_vars;
  std::vector<std::set<std::size_t>> parents;   // parents[v] = {u | u -> v}
  explicit PDAG(std::size_t n) : num_vars(n), parents(n) {}
  bool has_edge(std::size_t from, std::size_t to) const {
    return parents[to].find(from) != parents[to].end();
  }
  void add_edge(std::size_t from, std::size_t to) { parents[to].insert(from); }
  void remove_edge(std::size_t from, std::size_t to) { parents[to].erase(from); }
};

// Flat bit matrix: bit (u, v) set <=> arc u -> v present; undirected = both.
class AdjMatGraph {
 public:
  explicit AdjMatGraph(std::size_t n) : n_(n), words_((n + 63) / 64), bits_(n * ((n + 63) / 64), 0) {
    for (std::size_t u = 0; u < n; ++u)          // complete graph, no self loops
      for (std::size_t v = 0; v < n; ++v)
        if (u != v) set(u, v);
  }
  std::size_t num_vars() const { return n_; }
  bool has_directed_edge(std::size_t u, std::size_t v) const {
    check(u); check(v);
    return (bits_[u * words_ + (v >> 6)] >> (v & 63)) & 1u;
  }
  bool has_undirected_edge(std::size_t u, std::size_t v) const {
    return has_directed_edge(u, v) && has_directed_edge(v, u);
  }
  bool is_adjacent(std::size_t u, std::size_t v) const {
    return has_directed_edge(u, v) || has_directed_edge(v, u);
  }
  void remove_undirected_edge(std::size_t u, std::size_t v) {
    if (!has_undirected_edge(u, v)) throw std::invalid_argument("edge not found");
    clear(u, v);
    clear(v, u);
  }
  void orient_edge(std::size_t from, std::size_t to) {     // (from <-> to) => (from -> to)
    if (!has_undirected_edge(from, to)) return;
    clear(to, from);
  }
  // {w | v <-> w, w != excl}, ascending (pdag_with_adjmat.cpp:110-117)
  void undirected_neighbors_without(std::size_t v, std::size_t excl, std::vector<uint32_t>& out) const {
    out.clear();
    for (std::size_t j = 0; j < words_; ++j) {
      uint64_t w = bits_[v * words_ + j];
      while (w) {
        const std::size_t u = j * 64 + (std::size_t)__builtin_ctzll(w);
        w &= w - 1;
        if (u != excl && has_directed_edge(u, v)) out.push_back((uint32_t)u);
      }
    }
  }
  PDAG to_pdag() const {
    PDAG g(n_);
    for (std::size_t u = 0; u < n_; ++u)
      for (std::size_t v = 0; v < n_; ++v)
        if (has_directed_edge(u, v)) g.add_edge(u, v);
    return g;
  }

 private:
  void check(std::size_t i) const {
    if (i >= n_) throw std::out_of_range("PDAG index out of range");
  }
  void set(std::size_t u, std::size_t v) { bits_[u * words_ + (v >> 6)] |= (uint64_t)1 << (v & 63); }
  void clear(std::size_t u, std::size_t v) { bits_[u * words_ + (v >> 6)] &= ~((uint64_t)1 << (v & 63)); }
  std::size_t n_, words_;
  std::vector<uint64_t> bits_;
};

// d-separation by reachability in the moralised ancestral graph.
inline bool is_d_separated(const PDAG& g, std::size_t x, std::size_t y, const std::vector<std::size_t>& Z) {
  const std::size_t n = g.num_vars;
  if (x >= n || y >= n) throw std::out_of_range("x or y is out of range");
  std::vector<char> in_z(n, 0), anc(n, 0);
  for (auto z : Z) {
    if (z >= n) throw std::out_of_range("Z contains out-of-range index");
    if (z == x || z == y) throw std::invalid_argument("x or y is in Z");
    in_z[z] = 1;
  }
  std::vector<std::size_t> stack{x, y};
  stack.insert(stack.end(), Z.begin(), Z.end());
  while (!stack.empty()) {                                   // ancestral closure of {x,y} u Z
    const std::size_t v = stack.back();
    stack.pop_back();
    if (anc[v]) continue;
    anc[v] = 1;
    for (auto p : g.parents[v])
      if (!anc[p]) stack.push_back(p);
  }
  std::vector<std::vector<std::size_t>> adj(n);               // moral graph on the closure
  for (std::size_t v = 0; v < n; ++v) {
    if (!anc[v]) continue;
    std::vector<std::size_t> ps;
    for (auto p : g.parents[v])
      if (anc[p]) ps.push_back(p);
    for (std::size_t i = 0; i < ps.size(); ++i) {
      adj[ps[i]].push_back(v);
      adj[v].push_back(ps[i]);
      for (std::size_t j = i + 1; j < ps.size(); ++j) {
        adj[ps[i]].push_back(ps[j]);
        adj[ps[j]].push_back(ps[i]);
      }
    }
  }
  std::vector<char> vis(n, 0);
  std::queue<std::size_t> q;
  vis[x] = 1;
  q.push(x);
  while (!q.empty()) {
    const std::size_t u = q.front();
    q.pop();
    for (auto v : adj[u]) {
      if (in_z[v] || vis[v]) continue;
      if (v == y) return false;
      vis[v] = 1;
      q.push(v);
    }
  }
  return true;
}

}  // namespace obn
