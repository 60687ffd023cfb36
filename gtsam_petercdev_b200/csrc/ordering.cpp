// ordering.cpp -- a fill-reducing elimination order computed inside the product: automatic nested dissection on the variable
// adjacency graph (two variables are adjacent when a factor names both), separators taken from breadth-first level structures
// rooted at a pseudo-peripheral variable (George & Liu, "Computer Solution of Large Sparse Positive Definite Systems", ch. 8).
// The reference gets its orderings from the vendored CCOLAMD and METIS (inference/Ordering.cpp:43-125, 211-256); those libraries
// are not linked here, and callers that have them pass their permutation through gtb_set_ordering.  This routine is what
// gtb_set_ordering_builtin(2) and the ISAM2 layer use when nobody supplies one.  It is written for pose graphs (chains with
// loop closures), where level-structure separators are a few variables wide and the tree comes out balanced -- which is what
// the level-scheduled device elimination wants; for bundle adjustment use the Schur ordering (builtin 1).
// Deterministic: ties are broken by variable id.
#include <algorithm>
#include <cstdint>
#include <cstdlib>
#include <vector>

#include "plan.h"

namespace gtb {

namespace {

struct Graph {
  std::vector<int64_t> off;
  std::vector<int32_t> adj;   // ascending within a row, no self loops, no duplicates
};

Graph adjacency(const SymbolicInput& in) {
  const int32_t n = in.n_vars;
  Graph g;
  g.off.assign(n + 1, 0);
  auto each_edge = [&](auto&& fn) {
    for (int64_t f = 0; f < in.n_factors; ++f) {
      if (in.fac_nkeys[f] < 2) continue;
      const int32_t a = in.fac_var[2 * f], b = in.fac_var[2 * f + 1];
      if (a == b || a < 0 || b < 0) continue;
      fn(a, b);
      fn(b, a);
    }
  };
  each_edge([&](int32_t a, int32_t) { g.off[a + 1]++; });
  for (int32_t v = 0; v < n; ++v) g.off[v + 1] += g.off[v];
  std::vector<int32_t> raw(g.off[n]);
  std::vector<int64_t> fill(g.off.begin(), g.off.end() - 1);
  each_edge([&](int32_t a, int32_t b) { raw[fill[a]++] = b; });
  // rows ascending and without duplicates (two factors on the same pair of variables), compacted in place
  g.adj.reserve(raw.size());
  std::vector<int64_t> off(n + 1, 0);
  for (int32_t v = 0; v < n; ++v) {
    std::sort(raw.begin() + g.off[v], raw.begin() + g.off[v + 1]);
    for (int64_t k = g.off[v]; k < g.off[v + 1]; ++k)
      if (k == g.off[v] || raw[k] != raw[k - 1]) g.adj.push_back(raw[k]);
    off[v + 1] = (int64_t)g.adj.size();
  }
  g.off.swap(off);
  return g;
}

struct Dissector {
  const Graph& g;
  int leaf;
  std::vector<int32_t> label;   // which live subset a variable belongs to (0: already ordered)
  std::vector<int32_t> level;   // scratch: BFS level inside the current subset
  std::vector<int32_t> out;
  int32_t next_label = 1;

  Dissector(const Graph& graph, int32_t n, int leaf_size) : g(graph), leaf(leaf_size), label(n, 1), level(n, -1) { out.reserve(n); }

  int degree_in(int32_t v, int32_t L) const {
    int d = 0;
    for (int64_t k = g.off[v]; k < g.off[v + 1]; ++k) d += label[g.adj[k]] == L;
    return d;
  }

  // breadth-first level structure of the subset labelled L from root; returns the levels back to back
  void bfs(int32_t root, int32_t L, std::vector<int32_t>& order, std::vector<int64_t>& level_off) {
    order.clear();
    level_off.assign(1, 0);
    order.push_back(root);
    level[root] = 0;
    size_t head = 0;
    while (head < order.size()) {
      const size_t end = order.size();
      level_off.push_back((int64_t)end);
      for (; head < end; ++head) {
        const int32_t v = order[head];
        for (int64_t k = g.off[v]; k < g.off[v + 1]; ++k) {
          const int32_t w = g.adj[k];
          if (label[w] == L && level[w] < 0) {
            level[w] = level[v] + 1;
            order.push_back(w);
          }
        }
      }
    }
  }
  void clear_levels(const std::vector<int32_t>& order) {
    for (int32_t v : order) level[v] = -1;
  }

  void emit_by_degree(std::vector<int32_t>& nodes, int32_t L) {
    std::vector<std::pair<int, int32_t>> key;
    key.reserve(nodes.size());
    for (int32_t v : nodes) key.emplace_back(degree_in(v, L), v);
    std::sort(key.begin(), key.end());
    for (const auto& kv : key) {
      out.push_back(kv.second);
      label[kv.second] = 0;
    }
  }

  // `nodes` is one connected subset, all labelled L
  void dissect(std::vector<int32_t>& nodes, int32_t L) {
    if ((int)nodes.size() <= leaf) return emit_by_degree(nodes, L);
    // pseudo-peripheral root: from a variable of least degree to a least-degree variable of the last level of its level
    // structure, while that deepens the structure (each pass is one sweep over the subset's edges)
    int32_t root = nodes[0];
    int best = degree_in(root, L);
    for (int32_t v : nodes) {   // start from a variable of least degree (ties: smallest id)
      const int d = degree_in(v, L);
      if (d < best || (d == best && v < root)) { best = d; root = v; }
    }
    std::vector<int32_t> order;
    std::vector<int64_t> loff;
    int height = -1;
    for (int it = 0; it < 3; ++it) {
      bfs(root, L, order, loff);
      const int h = (int)loff.size() - 2;   // levels 0 .. h
      clear_levels(order);
      if (h <= height) break;
      height = h;
      int32_t cand = order[loff[h]];
      int cd = degree_in(cand, L);
      for (int64_t k = loff[h]; k < loff[h + 1]; ++k) {
        const int d = degree_in(order[k], L);
        if (d < cd || (d == cd && order[k] < cand)) { cd = d; cand = order[k]; }
      }
      if (cand == root) break;
      root = cand;
    }
    bfs(root, L, order, loff);
    const int h = (int)loff.size() - 2;
    if (h < 2) {   // no interior level: the subset is (nearly) complete
      clear_levels(order);
      return emit_by_degree(nodes, L);
    }
    // separator level: balance the two sides, prefer thin levels
    const int64_t total = (int64_t)order.size();
    int m = 1;
    int64_t best_cost = INT64_MAX;
    for (int l = 1; l <= h - 1; ++l) {
      const int64_t before = loff[l], width = loff[l + 1] - loff[l], after = total - loff[l + 1];
      const int64_t cost = std::llabs(before - after) + 4 * width;
      if (cost < best_cost) { best_cost = cost; m = l; }
    }
    // only the variables of level m that touch level m + 1 separate anything
    std::vector<int32_t> sep;
    for (int64_t k = loff[m]; k < loff[m + 1]; ++k) {
      const int32_t v = order[k];
      bool touches = false;
      for (int64_t q = g.off[v]; q < g.off[v + 1] && !touches; ++q) {
        const int32_t w = g.adj[q];
        touches = label[w] == L && level[w] == m + 1;
      }
      if (touches) sep.push_back(v);
    }
    clear_levels(order);
    for (int32_t v : sep) label[v] = -1;   // out of every subset until it is emitted
    // connected components of what is left
    // (seeded in the order of the level structure: deterministic, and the part before the separator comes first)
    std::vector<std::vector<int32_t>> comps;
    std::vector<int32_t> comp_label;
    for (int32_t s : order) {
      if (label[s] != L) continue;
      const int32_t nl = ++next_label;
      comps.emplace_back();
      std::vector<int32_t>& c = comps.back();
      c.push_back(s);
      label[s] = nl;
      for (size_t head = 0; head < c.size(); ++head) {
        const int32_t v = c[head];
        for (int64_t k = g.off[v]; k < g.off[v + 1]; ++k) {
          const int32_t w = g.adj[k];
          if (label[w] == L) { label[w] = nl; c.push_back(w); }
        }
      }
      comp_label.push_back(nl);
    }
    for (size_t i = 0; i < comps.size(); ++i) dissect(comps[i], comp_label[i]);
    std::sort(sep.begin(), sep.end());
    for (int32_t v : sep) {
      out.push_back(v);
      label[v] = 0;
    }
  }
};

}  // namespace

std::vector<int32_t> nested_dissection_order(const SymbolicInput& in, int leaf_size) {
  const int32_t n = in.n_vars;
  const Graph g = adjacency(in);
  Dissector d(g, n, leaf_size < 1 ? 1 : leaf_size);
  // connected components of the whole graph first
  for (int32_t s = 0; s < n; ++s) {
    if (d.label[s] != 1) continue;
    const int32_t nl = ++d.next_label;
    std::vector<int32_t> c(1, s);
    d.label[s] = nl;
    for (size_t head = 0; head < c.size(); ++head) {
      const int32_t v = c[head];
      for (int64_t k = g.off[v]; k < g.off[v + 1]; ++k) {
        const int32_t w = g.adj[k];
        if (d.label[w] == 1) { d.label[w] = nl; c.push_back(w); }
      }
    }
    d.dissect(c, nl);
  }
  return d.out;
}

}  // namespace gtb
