// graph.hpp -- host-side graph containers with the reference's semantics.
//   PDAG              include/graph/pdag.h:16-37 (result type of every algorithm)
//   AdjMatGraph       the subset of PDAGwithAdjMat that PC uses
//                     (src/graph/pdag_with_adjmat.cpp:7-11,27-36,61-69,72-117,171-189)
//                     + Meek's rules as the reference states them (:191-279)
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
  // Arcs of an existing PDAG (u -> v for every u in parents[v]); used to post-process a result.
  explicit AdjMatGraph(const PDAG& g) : n_(g.num_vars), words_((g.num_vars + 63) / 64),
                                        bits_(g.num_vars * ((g.num_vars + 63) / 64), 0) {
    for (std::size_t v = 0; v < n_; ++v)
      for (auto u : g.parents[v]) {
        check(u);
        set(u, v);
      }
  }
  std::size_t num_vars() const { return n_; }
  bool has_directed_edge(std::size_t u, std::size_t v) const {
    check(u); check(v);
    return (bits_[u * words_ + (v >> 6)] >> (v & 63)) & 1u;
  }
  bool has_undirected_edge(std::size_t u, std::size_t v) const {
    return has_directed_edge(u, v) && has_directed_edge(v, u);
  }
  // the same without the range checks, for inner loops over ids that came out of this graph
  bool has_undirected_edge_unchecked(std::size_t u, std::size_t v) const { return get(u, v) && get(v, u); }
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
  // ---- Meek's rules (pdag_with_adjmat.cpp:191-279) -------------------------------------------
  // The reference's rule set is NOT the textbook one (rule 1 carries two extra guards, rule 3 has
  // no non-adjacency test and fires at most once per x and sweep, rule 4 orients every undirected
  // edge into b from the union of three neighbourhoods); the sweeps, their vertex order and the
  // in-place updates are reproduced as they are, because every orientation feeds the next test.
  // Lists are snapshots taken where the reference takes them (range-for over a returned vector).
  void successors(std::size_t v, std::vector<uint32_t>& out) const {          // {w | v -> w}, ascending (:72-86)
    out.clear();
    for (std::size_t j = 0; j < words_; ++j) {
      uint64_t w = bits_[v * words_ + j];
      while (w) {
        out.push_back((uint32_t)(j * 64 + (std::size_t)__builtin_ctzll(w)));
        w &= w - 1;
      }
    }
  }
  void undirected_neighbors(std::size_t v, std::vector<uint32_t>& out) const {  // (:103-109)
    undirected_neighbors_without(v, (std::size_t)-1, out);
  }
  void directed_parents(std::size_t v, std::vector<uint32_t>& out) const {     // u -> v and not v -> u (:192-197)
    out.clear();
    for (std::size_t u = 0; u < n_; ++u)
      if (get(u, v) && !get(v, u)) out.push_back((uint32_t)u);
  }
  void directed_children(std::size_t v, std::vector<uint32_t>& out) const {    // v -> u and not u -> v (:198-204)
    successors(v, out);
    out.erase(std::remove_if(out.begin(), out.end(), [&](uint32_t u) { return get(u, v); }), out.end());
  }
  // s -> ... -> t along arcs that have no reverse arc; s == t counts as a path (:120-137)
  bool has_directed_path(std::size_t s, std::size_t t) const {
    check(s); check(t);
    std::vector<char> seen(n_, 0);
    std::vector<uint32_t> stack{(uint32_t)s}, succ;
    while (!stack.empty()) {
      const std::size_t u = stack.back();
      stack.pop_back();
      if (u == t) return true;
      if (seen[u]) continue;
      seen[u] = 1;
      successors(u, succ);
      for (auto w : succ)
        if (!seen[w] && !get(w, u)) stack.push_back(w);
    }
    return false;
  }
  bool creates_unshielded_collider(std::size_t y, std::size_t z) const {       // (:209-214)
    for (std::size_t x = 0; x < n_; ++x)
      if (get(x, y) && !get(y, x) && !is_adjacent(x, z)) return true;
    return false;
  }
  void apply_meeks_rules(bool apply_r4 = false) {
    std::vector<uint32_t> a, b, cand, tmp;
    bool changed = true;
    while (changed) {
      changed = false;
      // rule 1 (:221-229): x -> y - z, x and z not adjacent  =>  y -> z
      for (std::size_t y = 0; y < n_; ++y) {
        directed_parents(y, a);
        for (auto x : a) {
          undirected_neighbors(y, b);
          for (auto z : b)
            if (!is_adjacent(x, z) && !creates_unshielded_collider(y, z) && !has_directed_path(z, y)) {
              orient_edge(y, z);
              changed = true;
            }
        }
      }
      // rule 2 (:231-238): x -> z -> y and x - y  =>  x -> y
      for (std::size_t z = 0; z < n_; ++z) {
        directed_parents(z, a);
        for (auto x : a) {
          directed_children(z, b);
          for (auto y : b)
            if (has_undirected_edge(x, y)) {
              orient_edge(x, y);
              changed = true;
            }
        }
      }
      // rule 3 (:240-258): the first (y, z, w) among x's undirected neighbours with y -> w <- z  =>  x -> w
      for (std::size_t x = 0; x < n_; ++x) {
        undirected_neighbors(x, a);
        if (a.size() < 3) continue;
        bool done = false;
        for (std::size_t iy = 0; iy < a.size() && !done; ++iy)
          for (std::size_t iz = 0; iz < a.size() && !done; ++iz) {
            if (iz == iy) continue;
            for (std::size_t iw = 0; iw < a.size(); ++iw) {
              if (iw == iy || iw == iz) continue;
              const std::size_t y = a[iy], z = a[iz], w = a[iw];
              if (get(y, w) && !get(w, y) && get(z, w) && !get(w, z)) {
                orient_edge(x, w);
                changed = true;
                done = true;
                break;
              }
            }
          }
      }
      // rule 4 (:260-277), off by default
      if (apply_r4)
        for (std::size_t c = 0; c < n_; ++c) {
          directed_children(c, a);
          for (auto bb : a) {
            directed_parents(c, b);
            for (auto d : b) {
              if (bb == d || is_adjacent(bb, d)) continue;
              undirected_neighbors(bb, cand);
              undirected_neighbors(d, tmp);
              cand.insert(cand.end(), tmp.begin(), tmp.end());
              for (std::size_t u = 0; u < n_; ++u)
                if (is_adjacent(c, u)) cand.push_back((uint32_t)u);
              std::sort(cand.begin(), cand.end());
              cand.erase(std::unique(cand.begin(), cand.end()), cand.end());
              for (auto v : cand)
                if (has_undirected_edge(v, bb)) {
                  orient_edge(v, bb);
                  changed = true;
                }
            }
          }
        }
    }
  }

  // Meek's rules as published (Meek 1995, rules 1-3; rule 4 only matters with background knowledge).
  // NOT a reference code path: the reference's rule 1 can never fire (its guards
  // `!is_adjacent(x, z)` and `!creates_unshielded_collider(y, z)` contradict each other for the
  // same x, pdag_with_adjmat.cpp:225-226), so apply_meeks_rules() above does not complete a
  // pattern to its CPDAG; this one does (tests: brute force over the Markov equivalence class).
  void complete_orientation(bool apply_r4 = false) {
    auto directed = [&](std::size_t u, std::size_t v) { return get(u, v) && !get(v, u); };
    auto undirected = [&](std::size_t u, std::size_t v) { return get(u, v) && get(v, u); };
    bool changed = true;
    while (changed) {
      changed = false;
      for (std::size_t b = 0; b < n_; ++b)
        for (std::size_t c = 0; c < n_; ++c) {
          if (b == c || !undirected(b, c)) continue;
          bool orient = false;                                // decide b -> c
          for (std::size_t a = 0; a < n_ && !orient; ++a) {
            if (a == b || a == c) continue;
            if (directed(a, b) && !is_adjacent(a, c)) orient = true;              // R1: a -> b - c
            else if (directed(b, a) && directed(a, c)) orient = true;             // R2: b -> a -> c, b - c
            else if (undirected(b, a) && directed(a, c))                          // R3: b - a -> c, b - d -> c, a and d apart
              for (std::size_t d = a + 1; d < n_ && !orient; ++d)
                if (d != b && d != c && undirected(b, d) && directed(d, c) && !is_adjacent(a, d)) orient = true;
          }
          if (!orient && apply_r4)                                                 // R4: b - d -> a -> c, b adjacent a, d and c apart
            for (std::size_t a = 0; a < n_ && !orient; ++a) {
              if (a == b || a == c || !directed(a, c) || !is_adjacent(b, a)) continue;
              for (std::size_t d = 0; d < n_ && !orient; ++d)
                if (d != a && d != b && d != c && undirected(b, d) && directed(d, a) && !is_adjacent(d, c)) orient = true;
            }
          if (orient) {
            clear(c, b);
            changed = true;
          }
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
  bool get(std::size_t u, std::size_t v) const { return (bits_[u * words_ + (v >> 6)] >> (v & 63)) & 1u; }
  void set(std::size_t u, std::size_t v) { bits_[u * words_ + (v >> 6)] |= (uint64_t)1 << (v & 63); }
  void clear(std::size_t u, std::size_t v) { bits_[u * words_ + (v >> 6)] &= ~((uint64_t)1 << (v & 63)); }
  std::size_t n_, words_;
  std::vector<uint64_t> bits_;
};

// Meek's rules on a finished PDAG (what pc() would return if pc.cpp:84 were enabled).
inline PDAG apply_meeks_rules(const PDAG& g, bool apply_r4 = false, bool textbook = false) {
  AdjMatGraph m(g);
  if (textbook) m.complete_orientation(apply_r4);
  else m.apply_meeks_rules(apply_r4);
  return m.to_pdag();
}

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
