// symbolic.cpp -- host-side symbolic analysis: ordering -> elimination tree -> junction tree -> clique layout.
//
// A from-scratch restatement, over flat integer arrays, of what the reference does with shared_ptr trees on
// EVERY solve (inference/EliminateableFactorGraph-inst.h:131-137):
//   VariableIndex            inference/VariableIndex-inl.h:27-49      factor lists per variable, ascending
//   EliminationTree ctor     inference/EliminationTree-inst.h:78-156  parent = root of the tree holding the
//                                                                     previously eliminated key of a factor
//   JunctionTree ctor        inference/JunctionTree-inst.h:65-153     symbolic elimination per node + the
//                                                                     supernode rule  nrParents(me) + nrFrontals
//                                                                     (me so far) == nrParents(child)
//   Cluster::mergeChildren   inference/ClusterTree-inst.h:46-96       merged frontal keys appended reversed, then
//                                                                     the whole list reversed
//   EliminateSymbolic        symbolic/SymbolicFactor-inst.h:38-70     separator = sorted set difference
//   Scatter                  linear/Scatter.cpp:39-73                 front columns: frontals, sorted separator
// The output must be bit-identical to the reference's Bayes tree (tests/test_symbolic.py checks it against
// oracle dumps for COLAMD, METIS and Schur orderings).
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <stdexcept>

#include "plan.h"

namespace gtb {

namespace {
struct Cluster {
  std::vector<int32_t> frontals;  // variable ids, elimination order
  std::vector<int32_t> factors;   // graph indices, gather order
  std::vector<int32_t> children;  // elimination-tree node ids naming child clusters
};
}  // namespace

// Numbers the cliques of a cluster tree in DFS post-order and flattens it to the index arrays of a Plan.
static void flatten(const SymbolicInput& in, const std::vector<Cluster>& cl, const std::vector<std::vector<int32_t>>& sep,
                    const std::vector<int32_t>& eparent, Plan* P, std::vector<int32_t>* clique_of_out) {
  const int32_t n = in.n_vars;
  // ---- number the cliques in DFS post-order ----------------------------------------------------------
  std::vector<int32_t> node_of;       // clique id -> elimination-tree node id
  std::vector<int32_t> clique_of(n, -1);
  {
    struct Item { int32_t node; bool expanded; };
    std::vector<Item> st;
    std::vector<int32_t> roots;
    for (int32_t j = 0; j < n; ++j)
      if (eparent[j] == -1) roots.push_back(j);
    for (auto it = roots.rbegin(); it != roots.rend(); ++it) st.push_back({*it, false});
    while (!st.empty()) {
      Item& top = st.back();
      if (top.expanded) {
        clique_of[top.node] = (int32_t)node_of.size();
        node_of.push_back(top.node);
        st.pop_back();
        continue;
      }
      top.expanded = true;
      const int32_t nd = top.node;
      for (auto it = cl[nd].children.rbegin(); it != cl[nd].children.rend(); ++it) st.push_back({*it, false});
    }
  }
  const int32_t C = (int32_t)node_of.size();

  // ---- flatten ---------------------------------------------------------------------------------------
  Plan& p = *P;
  p = Plan();
  p.n_cliques = C;
  p.parent.assign(C, -1);
  p.level.assign(C, 0);
  p.n_frontal_vars.resize(C);
  p.n_vars.resize(C);
  p.vars_off.assign(C + 1, 0);
  p.F.resize(C);
  p.S.resize(C);
  p.ld.resize(C);
  p.front_off.resize(C);
  p.child_off.assign(C + 1, 0);
  p.fac_off.assign(C + 1, 0);
  p.pmap_off.resize(C);
  p.gidx_off.resize(C);
  p.cond_off.assign(C + 1, 0);

  int64_t arena = 0, gtot = 0, ptot = 0;
  for (int32_t c = 0; c < C; ++c) {
    const int32_t nd = node_of[c];
    const Cluster& k = cl[nd];
    p.n_frontal_vars[c] = (int32_t)k.frontals.size();
    p.n_vars[c] = (int32_t)(k.frontals.size() + sep[nd].size());
    p.vars_off[c + 1] = p.vars_off[c] + p.n_vars[c];
    p.vars.insert(p.vars.end(), k.frontals.begin(), k.frontals.end());
    p.vars.insert(p.vars.end(), sep[nd].begin(), sep[nd].end());
    int32_t F = 0, S = 0;
    for (int32_t v : k.frontals) F += in.var_dim[v];
    for (int32_t v : sep[nd]) S += in.var_dim[v];
    p.F[c] = F;
    p.S[c] = S;
    const int32_t dim = F + S + 1;
    const bool small = is_small_front(F, dim);
    p.ld[c] = small ? dim : ((dim + 3) & ~3);
    arena = small ? ((arena + 1) & ~(int64_t)1) : ((arena + 15) & ~(int64_t)15);
    p.front_off[c] = arena;
    arena += (int64_t)p.ld[c] * dim;
    p.child_off[c + 1] = p.child_off[c] + (int64_t)k.children.size();
    for (int32_t ch : k.children) {
      p.children.push_back(clique_of[ch]);
      p.parent[clique_of[ch]] = c;
      p.level[c] = std::max(p.level[c], p.level[clique_of[ch]] + 1);
    }
    p.fac_off[c + 1] = p.fac_off[c] + (int64_t)k.factors.size();
    p.gidx_off[c] = gtot;
    gtot += F + S;
    p.pmap_off[c] = ptot;
    ptot += S + 1;
    p.cond_off[c + 1] = p.cond_off[c] + (int64_t)F * dim;
    p.max_front = std::max(p.max_front, dim);
    const double f = F, s1 = S + 1;
    const double fl = f * f * f / 3.0 + f * f * s1 + f * s1 * (s1 + 1);
    p.flops += fl;
    if (small) p.n_small++; else { p.n_large++; p.flops_large += fl; }
    p.assembly_bytes += 8.0 * (0.5 * dim * (dim + 1) + 0.5 * s1 * (s1 + 1));
    p.bytes_backsolve += 8.0 * (0.5 * f * (f + 1) + f * s1 + f);
    if (small) {
      // own factor rows are added below once the slots are known; front written once, Schur block read by the parent
      p.bytes_small += 8.0 * 0.5 * dim * (dim + 1);
    } else {
      p.flops_syrk_schur += f * s1 * (s1 + 1);
      for (int32_t k0 = 0; k0 < F; k0 += 32) {
        const int32_t kb = std::min(32, F - k0);
        const double rb = k0 + kb, rows = F - rb;                       // rows [rb, F), columns [i, dim)
        p.flops_syrk_frontal += 2.0 * kb * (rows * (dim - rb) - 0.5 * rows * (rows - 1));
      }
    }
  }
  for (int32_t c = 0; c < C; ++c) {
    const int32_t par = p.parent[c];
    if (par >= 0 && is_small_front(p.F[par], p.F[par] + p.S[par] + 1)) {
      const double s1 = p.S[c] + 1;
      p.bytes_small += 8.0 * 0.5 * s1 * (s1 + 1);
    }
  }
  p.arena_doubles = arena;

  // columns, factor slots, global scalar indices, child -> parent maps
  p.gidx.resize(gtot);
  p.pmap.assign(ptot, -1);
  p.slots.resize(p.fac_off[C]);
  {
    std::vector<int32_t> col_of(n, -1), stamp(n, -1);
    for (int32_t c = 0; c < C; ++c) {
      const int32_t nd = node_of[c];
      int32_t col = 0;
      int64_t g = p.gidx_off[c];
      for (int64_t q = p.vars_off[c]; q < p.vars_off[c + 1]; ++q) {
        const int32_t v = p.vars[q];
        col_of[v] = col;
        stamp[v] = c;
        for (int32_t d = 0; d < in.var_dim[v]; ++d) p.gidx[g++] = (int32_t)(in.var_off[v] + d);
        col += in.var_dim[v];
      }
      const int32_t N = col;
      int64_t s = p.fac_off[c];
      for (int32_t gi : cl[nd].factors) {
        FactorSlot fs;
        fs.factor = gi;
        fs.pad = 0;
        const int32_t v0 = in.fac_var[2 * (int64_t)gi];
        if (stamp[v0] != c) throw std::runtime_error("internal: factor key not in clique");
        fs.col0 = col_of[v0];
        fs.col1 = -1;
        if (in.fac_nkeys[gi] == 2) {
          const int32_t v1 = in.fac_var[2 * (int64_t)gi + 1];
          if (stamp[v1] != c) throw std::runtime_error("internal: factor key not in clique");
          fs.col1 = col_of[v1];
        }
        p.slots[s++] = fs;
      }
      for (int64_t q = p.child_off[c]; q < p.child_off[c + 1]; ++q) {
        const int32_t ch = p.children[q];
        int64_t m = p.pmap_off[ch];
        const int64_t sb = p.vars_off[ch] + p.n_frontal_vars[ch];
        for (int64_t r = sb; r < p.vars_off[ch + 1]; ++r) {
          const int32_t v = p.vars[r];
          if (stamp[v] != c) throw std::runtime_error("internal: child separator key not in parent clique");
          for (int32_t d = 0; d < in.var_dim[v]; ++d) p.pmap[m++] = col_of[v] + d;
        }
        p.pmap[m] = N;  // rhs column
      }
    }
  }

  // ---- level schedule --------------------------------------------------------------------------------
  int32_t L = 0;
  for (int32_t c = 0; c < C; ++c) L = std::max(L, p.level[c] + 1);
  p.n_levels = L;
  p.level_off.assign(L + 1, 0);
  p.level_n_small.assign(L, 0);
  for (int32_t c = 0; c < C; ++c) p.level_off[p.level[c] + 1]++;
  for (int32_t l = 0; l < L; ++l) p.level_off[l + 1] += p.level_off[l];
  p.level_cliques.resize(C);
  {
    std::vector<int64_t> fill(p.level_off.begin(), p.level_off.end() - 1);
    for (int pass = 0; pass < 2; ++pass)
      for (int32_t c = 0; c < C; ++c) {
        const bool small = is_small_front(p.F[c], p.F[c] + p.S[c] + 1);
        if (small == (pass == 0)) {
          p.level_cliques[fill[p.level[c]]++] = c;
          if (small) p.level_n_small[p.level[c]]++;
        }
      }
  }
  if (clique_of_out) *clique_of_out = clique_of;
}

void build_plan(const SymbolicInput& in, int32_t small_max, Plan* P, Plan* N, double amalg_ratio, int32_t amalg_max_F,
                int32_t amalg_min_S) {
  const int32_t n = in.n_vars;
  const int64_t M = in.n_factors;
  static const bool times = getenv("GTB_PLAN_TIMES") != nullptr;
  auto t_prev = std::chrono::steady_clock::now();
  auto lap = [&](const char* what) {
    if (!times) return;
    const auto now = std::chrono::steady_clock::now();
    fprintf(stderr, "  build_plan: %-28s %.3f ms\n", what, std::chrono::duration<double, std::milli>(now - t_prev).count());
    t_prev = now;
  };
  if ((int64_t)in.order.size() != n)
    throw std::runtime_error("ordering must contain every variable exactly once (partial elimination is not supported)");
  std::vector<int32_t> pos_of(n, -1);
  for (int32_t j = 0; j < n; ++j) {
    int32_t v = in.order[j];
    if (v < 0 || v >= n || pos_of[v] != -1) throw std::runtime_error("ordering is not a permutation of the variables");
    pos_of[v] = j;
  }

  // ---- VariableIndex (CSR, factor indices ascending per variable) -------------------------------------
  std::vector<int64_t> vi_off(n + 1, 0);
  for (int64_t g = 0; g < M; ++g)
    for (int k = 0; k < in.fac_nkeys[g]; ++k) vi_off[in.fac_var[2 * g + k] + 1]++;
  for (int32_t v = 0; v < n; ++v) vi_off[v + 1] += vi_off[v];
  std::vector<int32_t> vi_fac(vi_off[n]);
  {
    std::vector<int64_t> fill(vi_off.begin(), vi_off.end() - 1);
    for (int64_t g = 0; g < M; ++g)
      for (int k = 0; k < in.fac_nkeys[g]; ++k) vi_fac[fill[in.fac_var[2 * g + k]]++] = (int32_t)g;
  }

  // ---- elimination tree -----------------------------------------------------------------------------
  std::vector<int32_t> eparent(n, -1), anc(n, -1);
  std::vector<int32_t> prev_col(M, -1);
  std::vector<std::vector<int32_t>> echildren(n), efactors(n);
  for (int32_t j = 0; j < n; ++j) {
    const int32_t v = in.order[j];
    if (vi_off[v] == vi_off[v + 1])
      throw std::runtime_error("EliminationTree: given ordering contains variables that are not involved in the factor graph");
    for (int64_t q = vi_off[v]; q < vi_off[v + 1]; ++q) {
      const int32_t i = vi_fac[q];
      if (prev_col[i] != -1) {
        // root of the tree containing the previously eliminated key of this factor (path-compressed walk;
        // the reference walks parents[] uncompressed, the root found is the same)
        int32_t r = prev_col[i];
        while (true) {
          const int32_t a = anc[r];
          if (a == -1 || a == j) break;
          anc[r] = j;
          r = a;
        }
        if (anc[r] == -1) {
          anc[r] = j;
          eparent[r] = j;
          echildren[j].push_back(r);
        }
      } else {
        efactors[j].push_back(i);
      }
      prev_col[i] = j;
    }
  }

  lap("variable index + etree");
  // ---- symbolic elimination: separator of every elimination-tree node ------------------------------
  // (ascending variable id.  The children's separators are sorted already: the union is built by linear merges onto the few
  // keys of the node's own factors instead of a sort of the whole set -- separators of a 10,000-pose graph hold hundreds of
  // variables and the ISAM2 layer plans at every update)
  std::vector<std::vector<int32_t>> sep(n);
  {
    std::vector<int32_t> mark(n, -1);
    std::vector<int32_t> own, merged;
    for (int32_t j = 0; j < n; ++j) {
      const int32_t v = in.order[j];
      std::vector<int32_t>& s = sep[j];
      mark[v] = j;
      own.clear();
      for (int32_t g : efactors[j])
        for (int k = 0; k < in.fac_nkeys[g]; ++k) {
          const int32_t u = in.fac_var[2 * (int64_t)g + k];
          if (mark[u] != j) { mark[u] = j; own.push_back(u); }
        }
      std::sort(own.begin(), own.end());
      s.swap(own);
      for (int32_t c : echildren[j]) {
        const std::vector<int32_t>& cs = sep[c];
        merged.clear();
        merged.reserve(s.size() + cs.size());
        size_t a = 0, b = 0;
        while (a < s.size() || b < cs.size()) {
          if (b == cs.size() || (a < s.size() && s[a] < cs[b])) {
            merged.push_back(s[a++]);
          } else if (a < s.size() && s[a] == cs[b]) {
            merged.push_back(s[a++]);
            ++b;
          } else {
            if (cs[b] != v) merged.push_back(cs[b]);
            ++b;
          }
        }
        s.swap(merged);
      }
      own.clear();
    }
  }

  lap("separators");
  // ---- junction tree: merge children into parents ----------------------------------------------------
  std::vector<Cluster> cl(n);
  std::vector<uint8_t> merged_away(n, 0);
  {
    std::vector<uint8_t> merge;
    for (int32_t j = 0; j < n; ++j) {
      Cluster& me = cl[j];
      me.frontals.assign(1, in.order[j]);
      me.factors = std::move(efactors[j]);
      me.children = std::move(echildren[j]);
      const size_t my_parents = sep[j].size();
      size_t my_frontals = 1;
      const size_t nc = me.children.size();
      merge.assign(nc, 0);
      bool any = false;
      for (size_t i = 0; i < nc; ++i) {
        const int32_t c = me.children[i];
        if (my_parents + my_frontals == sep[c].size()) {
          my_frontals += cl[c].frontals.size();
          merge[i] = 1;
          any = true;
        }
      }
      if (!any) continue;
      std::vector<int32_t> old = std::move(me.children);
      me.children.clear();
      for (size_t i = 0; i < nc; ++i) {
        const int32_t c = old[i];
        if (merge[i]) {
          Cluster& ch = cl[c];
          me.frontals.insert(me.frontals.end(), ch.frontals.rbegin(), ch.frontals.rend());
          me.factors.insert(me.factors.end(), ch.factors.begin(), ch.factors.end());
          me.children.insert(me.children.end(), ch.children.begin(), ch.children.end());
          merged_away[c] = 1;
          Cluster().frontals.swap(ch.frontals);
          Cluster().factors.swap(ch.factors);
          Cluster().children.swap(ch.children);
          std::vector<int32_t>().swap(sep[c]);
        } else {
          me.children.push_back(c);
        }
      }
      std::reverse(me.frontals.begin(), me.frontals.end());
    }
  }

  // ---- the reference's clique layout, bit-exact -------------------------------------------------------
  lap("supernodes");
  std::vector<int32_t> clique_exact;
  flatten(in, cl, sep, eparent, P, &clique_exact);
  lap("flatten (reported plan)");
  if (!N) return;

  // ---- relaxed amalgamation for the NUMERIC plan ------------------------------------------------------
  // The reference's supernode rule only merges a child whose separator is exactly the parent's variable set.  Trees of
  // large fronts are full of near misses: a parent that owns one or two variables its child does not see (every link of a
  // COLAMD-ordered bundle adjustment: 1,124 levels of ONE front with 9 frontal and ~1040 separator scalars; the lower
  // levels of a METIS tree).  Each link then costs a panel launch, a Schur launch and the hand-over of a ~1000 x 1000
  // Schur complement for a handful of pivots.  For the numeric factorisation such a child is merged into its parent
  // anyway: frontals = child's, then parent's; separator = parent's; the child's rows get structural zeros in the columns
  // it does not see (they stay exactly zero through the elimination).  A merge is taken while the padded area of the
  // merged front stays within `ratio` of the true area of its members and F stays below `max_F`.  The REPORTED layout (P)
  // is untouched; N records which reference cliques each numeric clique holds (orig_*), so that conditionals, the
  // failure rule of base/cholesky.cpp:146-158 (applied at every member's last pivot) and keys are reported per
  // reference clique.
  std::vector<std::vector<int32_t>> group(n);
  std::vector<double> area(n, 0.0);
  std::vector<int32_t> Fn(n, 0), Sn(n, 0);
  for (int32_t j = 0; j < n; ++j) {
    if (merged_away[j]) continue;
    group[j].assign(1, j);
    for (int32_t v : cl[j].frontals) Fn[j] += in.var_dim[v];
    for (int32_t v : sep[j]) Sn[j] += in.var_dim[v];
    area[j] = 0.5 * Fn[j] * (Fn[j] + 1.0) + (double)Fn[j] * (Sn[j] + 1.0);
  }
  bool any_amalgamated = false;
  if (amalg_ratio > 1.0) {
    for (int32_t j = 0; j < n; ++j) {
      if (merged_away[j] || is_small_front(Fn[j], Fn[j] + Sn[j] + 1)) continue;
      int32_t best = -1;
      size_t best_pos = 0;
      for (size_t i = 0; i < cl[j].children.size(); ++i) {
        const int32_t c = cl[j].children[i];
        if (is_small_front(Fn[c], Fn[c] + Sn[c] + 1)) continue;
        // below ~512 separator scalars the hand-over of the child's Schur complement (S^2 / 2 entries) is cheaper than the
        // zeros a merge pads into the parent's panel steps (measured: sphere2500-shape pose graphs get slower when merged)
        if (Sn[c] < amalg_min_S) continue;
        if (best < 0 || Fn[c] + Sn[c] > Fn[best] + Sn[best]) { best = c; best_pos = i; }
      }
      if (best < 0) continue;
      const int32_t c = best;
      const double Fm = (double)Fn[c] + Fn[j];
      const double merged_area = 0.5 * Fm * (Fm + 1.0) + Fm * (Sn[j] + 1.0);
      if (Fm > amalg_max_F || group[c].size() + group[j].size() > 255 || merged_area > amalg_ratio * (area[c] + area[j])) continue;
      Cluster& me = cl[j];
      Cluster& ch = cl[c];
      std::vector<int32_t> fr = std::move(ch.frontals);
      fr.insert(fr.end(), me.frontals.begin(), me.frontals.end());
      me.frontals = std::move(fr);
      me.factors.insert(me.factors.end(), ch.factors.begin(), ch.factors.end());
      std::vector<int32_t> kids(me.children.begin(), me.children.begin() + best_pos);
      kids.insert(kids.end(), ch.children.begin(), ch.children.end());
      kids.insert(kids.end(), me.children.begin() + best_pos + 1, me.children.end());
      me.children = std::move(kids);
      std::vector<int32_t> gr = std::move(group[c]);
      gr.insert(gr.end(), group[j].begin(), group[j].end());
      group[j] = std::move(gr);
      area[j] += area[c];
      Fn[j] += Fn[c];
      merged_away[c] = 1;
      any_amalgamated = true;
      Cluster().factors.swap(ch.factors);
      Cluster().children.swap(ch.children);
      std::vector<int32_t>().swap(sep[c]);
    }
  }
  lap("amalgamation");
  std::vector<int32_t> clique_num;
  if (any_amalgamated) {
    flatten(in, cl, sep, eparent, N, &clique_num);
  } else {   // nothing merged (every pose graph): the numeric plan IS the reported one
    *N = *P;
    clique_num = clique_exact;
  }
  lap("flatten (numeric plan)");
  // members of every numeric clique: reference clique ids in elimination order, with the frontal scalar where each ends
  Plan& q = *N;
  q.orig_off.assign(q.n_cliques + 1, 0);
  q.num_of.assign(P->n_cliques, -1);
  q.num_foff.assign(P->n_cliques, 0);
  std::vector<int32_t> node_of_num(q.n_cliques, -1);
  for (int32_t j = 0; j < n; ++j)
    if (!merged_away[j] && clique_num[j] >= 0) node_of_num[clique_num[j]] = j;
  for (int32_t c = 0; c < q.n_cliques; ++c) {
    const std::vector<int32_t>& g = group[node_of_num[c]];
    q.orig_off[c + 1] = q.orig_off[c] + (int64_t)g.size();
    int32_t end = 0;
    for (int32_t nd : g) {
      const int32_t ce = clique_exact[nd];
      q.num_of[ce] = c;
      q.num_foff[ce] = end;
      end += P->F[ce];
      q.orig_clique.push_back(ce);
      q.orig_end.push_back(end);
    }
    if (end != q.F[c]) throw std::runtime_error("internal: amalgamated clique does not add up");
  }
}

}  // namespace gtb
