// Multilevel k-way graph partitioner (host C++), the role METIS plays for the reference
// (experiments/OGB/preprocess.py:15-29 and experiments/GraphCast/data_utils/preprocess.py:14-30 call `metis.part_graph`;
// neither METIS nor its Python binding exist on this image).  Classic three phases:
//   1. coarsening by heavy-edge matching until the graph is small,
//   2. initial k-way partition of the coarsest graph by balanced region growing,
//   3. uncoarsening with greedy boundary refinement (move a vertex to the neighbouring part with the largest cut gain
//      when the balance constraint allows it).
// Input: undirected graph in CSR form (both directions present, no self loops needed), int64 row pointers, int32 columns.
// Output: part[v] in [0, k).  Deterministic for a given seed.  Single-threaded, O((V + E) log) per level.
#include <algorithm>
#include <cstdint>
#include <cstring>
#include <numeric>
#include <queue>
#include <random>
#include <unordered_map>
#include <vector>

namespace {

struct Graph {
  int64_t n = 0;
  std::vector<int64_t> ptr;   // n + 1
  std::vector<int32_t> adj;   // neighbours
  std::vector<int64_t> ew;    // edge weights (parallel to adj)
  std::vector<int64_t> vw;    // vertex weights
};

// ---- phase 1: heavy-edge matching -----------------------------------------------------------------------------------------
// cmap[v] = coarse vertex of v; returns the number of coarse vertices.
int64_t heavy_edge_matching(const Graph& g, std::mt19937_64& rng, int64_t max_vw, std::vector<int32_t>& cmap) {
  std::vector<int32_t> order(g.n);
  std::iota(order.begin(), order.end(), 0);
  std::shuffle(order.begin(), order.end(), rng);
  std::vector<int32_t> match(g.n, -1);
  for (int32_t v : order) {
    if (match[v] >= 0) continue;
    int32_t best = -1;
    int64_t best_w = -1;
    for (int64_t e = g.ptr[v]; e < g.ptr[v + 1]; ++e) {
      const int32_t u = g.adj[e];
      if (u == v || match[u] >= 0) continue;
      if (g.vw[v] + g.vw[u] > max_vw) continue;  // keep coarse vertices small enough to balance later
      if (g.ew[e] > best_w) { best_w = g.ew[e]; best = u; }
    }
    if (best >= 0) { match[v] = best; match[best] = v; } else { match[v] = v; }
  }
  cmap.assign(g.n, -1);
  int64_t nc = 0;
  for (int32_t v = 0; v < g.n; ++v) {
    if (cmap[v] >= 0) continue;
    cmap[v] = (int32_t)nc;
    if (match[v] != v) cmap[match[v]] = (int32_t)nc;
    ++nc;
  }
  return nc;
}

Graph contract(const Graph& g, const std::vector<int32_t>& cmap, int64_t nc) {
  Graph c;
  c.n = nc;
  c.vw.assign(nc, 0);
  for (int64_t v = 0; v < g.n; ++v) c.vw[cmap[v]] += g.vw[v];
  // members of every coarse vertex
  std::vector<int64_t> mptr(nc + 1, 0);
  for (int64_t v = 0; v < g.n; ++v) ++mptr[cmap[v] + 1];
  for (int64_t i = 0; i < nc; ++i) mptr[i + 1] += mptr[i];
  std::vector<int32_t> members(g.n);
  {
    std::vector<int64_t> fill(mptr.begin(), mptr.end() - 1);
    for (int64_t v = 0; v < g.n; ++v) members[fill[cmap[v]]++] = (int32_t)v;
  }
  c.ptr.assign(nc + 1, 0);
  c.adj.reserve(g.adj.size() / 2);
  c.ew.reserve(g.adj.size() / 2);
  std::vector<int64_t> slot(nc, -1);  // position of a coarse neighbour in the row being built
  for (int64_t cv = 0; cv < nc; ++cv) {
    const int64_t row_begin = (int64_t)c.adj.size();
    for (int64_t m = mptr[cv]; m < mptr[cv + 1]; ++m) {
      const int32_t v = members[m];
      for (int64_t e = g.ptr[v]; e < g.ptr[v + 1]; ++e) {
        const int32_t cu = cmap[g.adj[e]];
        if (cu == cv) continue;  // edge inside the coarse vertex disappears
        if (slot[cu] >= row_begin) {
          c.ew[slot[cu]] += g.ew[e];
        } else {
          slot[cu] = (int64_t)c.adj.size();
          c.adj.push_back(cu);
          c.ew.push_back(g.ew[e]);
        }
      }
    }
    c.ptr[cv + 1] = (int64_t)c.adj.size();
  }
  return c;
}

// ---- phase 2: balanced region growing on the coarsest graph --------------------------------------------------------------
void initial_partition(const Graph& g, int k, std::mt19937_64& rng, std::vector<int32_t>& part) {
  part.assign(g.n, -1);
  const int64_t total = std::accumulate(g.vw.begin(), g.vw.end(), (int64_t)0);
  const int64_t target = (total + k - 1) / k;
  std::vector<int32_t> order(g.n);
  std::iota(order.begin(), order.end(), 0);
  std::shuffle(order.begin(), order.end(), rng);
  size_t next_seed = 0;
  for (int p = 0; p < k; ++p) {
    int64_t w = 0;
    std::queue<int32_t> q;
    const int64_t want = (p == k - 1) ? total : target;  // the last part takes whatever is left
    while (w < want) {
      if (q.empty()) {  // (re)seed: the component ran dry or the part has not started yet
        while (next_seed < order.size() && part[order[next_seed]] >= 0) ++next_seed;
        if (next_seed >= order.size()) break;
        const int32_t s = order[next_seed];
        part[s] = p;
        w += g.vw[s];
        q.push(s);
        continue;
      }
      const int32_t v = q.front();
      q.pop();
      for (int64_t e = g.ptr[v]; e < g.ptr[v + 1] && w < want; ++e) {
        const int32_t u = g.adj[e];
        if (part[u] >= 0) continue;
        part[u] = p;
        w += g.vw[u];
        q.push(u);
      }
    }
  }
  for (int64_t v = 0; v < g.n; ++v)
    if (part[v] < 0) part[v] = k - 1;
}

// Alternative start: breadth-first order from a pseudo-peripheral vertex (the last vertex of a BFS from a random one), cut into
// k consecutive chunks of equal weight -- "slabs", the right shape for path- / ring-like graphs where blobs grown from random
// seeds leave 2k boundaries instead of k.
void initial_partition_slabs(const Graph& g, int k, std::mt19937_64& rng, std::vector<int32_t>& part) {
  std::vector<int32_t> order;
  order.reserve(g.n);
  std::vector<char> seen(g.n, 0);
  auto bfs = [&](int32_t start) {
    order.clear();
    std::fill(seen.begin(), seen.end(), 0);
    size_t head = 0;
    int64_t scan = 0;
    order.push_back(start);
    seen[start] = 1;
    while ((int64_t)order.size() < g.n) {
      if (head == order.size()) {  // next connected component
        while (scan < g.n && seen[scan]) ++scan;
        if (scan >= g.n) break;
        order.push_back((int32_t)scan);
        seen[scan] = 1;
      }
      const int32_t v = order[head++];
      for (int64_t e = g.ptr[v]; e < g.ptr[v + 1]; ++e) {
        const int32_t u = g.adj[e];
        if (!seen[u]) { seen[u] = 1; order.push_back(u); }
      }
    }
  };
  std::uniform_int_distribution<int64_t> pick(0, g.n - 1);
  bfs((int32_t)pick(rng));
  bfs(order.back());  // restart from the far end
  const int64_t total = std::accumulate(g.vw.begin(), g.vw.end(), (int64_t)0);
  part.assign(g.n, k - 1);
  int64_t acc = 0;
  for (int32_t v : order) {
    part[v] = (int32_t)std::min<int64_t>(k - 1, acc * k / std::max<int64_t>(total, 1));
    acc += g.vw[v];
  }
}

// ---- phase 3: greedy boundary refinement ---------------------------------------------------------------------------------
// One pass visits every vertex in random order and moves it to the adjacent part with the largest positive gain
// (external weight to that part - internal weight), or with zero gain when that evens out the two parts, provided the target
// part stays below max_pw.  Returns the number of moves.
int64_t refine_pass(const Graph& g, int k, int64_t max_pw, std::mt19937_64& rng, std::vector<int32_t>& part,
                    std::vector<int64_t>& pw) {
  std::vector<int32_t> order(g.n);
  std::iota(order.begin(), order.end(), 0);
  std::shuffle(order.begin(), order.end(), rng);
  std::vector<int64_t> conn(k, 0);
  std::vector<int32_t> touched;
  int64_t moves = 0;
  for (int32_t v : order) {
    const int32_t pv = part[v];
    touched.clear();
    bool boundary = false;
    for (int64_t e = g.ptr[v]; e < g.ptr[v + 1]; ++e) {
      const int32_t pu = part[g.adj[e]];
      if (conn[pu] == 0) touched.push_back(pu);
      conn[pu] += g.ew[e];
      boundary |= (pu != pv);
    }
    if (boundary) {
      const int64_t internal = conn[pv];
      int32_t best = -1;
      int64_t best_gain = 0;
      for (int32_t p : touched) {
        if (p == pv || pw[p] + g.vw[v] > max_pw) continue;
        const int64_t gain = conn[p] - internal;
        // strictly positive gains, or zero-gain moves that even out the two parts (they open up later positive moves)
        const bool evens_out = gain == 0 && best < 0 && pw[p] + 2 * g.vw[v] <= pw[pv];
        if (gain > best_gain || evens_out) {
          best = p;
          best_gain = gain;
        }
      }
      if (best >= 0) {
        part[v] = best;
        pw[pv] -= g.vw[v];
        pw[best] += g.vw[v];
        ++moves;
      }
    }
    for (int32_t p : touched) conn[p] = 0;
  }
  return moves;
}

// Fiduccia-Mattheyses style pass with hill climbing: repeatedly apply the best available boundary move (also when its gain is
// negative), lock the moved vertex, remember the best prefix of the move sequence and roll back everything after it.  This is
// what gets the greedy refinement above out of its local minima (a straight boundary often needs a few zero / negative moves
// before a whole bump can be shaved off).  Lazy priority queue: stale entries are recognised by a per-vertex stamp.
int64_t fm_pass(const Graph& g, int k, int64_t max_pw, std::vector<int32_t>& part, std::vector<int64_t>& pw) {
  struct Entry { int64_t gain; int32_t v, to; uint32_t stamp; };
  struct Less { bool operator()(const Entry& a, const Entry& b) const { return a.gain < b.gain; } };
  std::priority_queue<Entry, std::vector<Entry>, Less> pq;
  std::vector<uint32_t> stamp(g.n, 0);
  std::vector<char> locked(g.n, 0);
  std::vector<int64_t> conn(k, 0);
  std::vector<int32_t> touched;
  auto best_move = [&](int32_t v, int64_t& gain, int32_t& to) -> bool {
    const int32_t pv = part[v];
    touched.clear();
    for (int64_t e = g.ptr[v]; e < g.ptr[v + 1]; ++e) {
      const int32_t pu = part[g.adj[e]];
      if (conn[pu] == 0) touched.push_back(pu);
      conn[pu] += g.ew[e];
    }
    to = -1;
    gain = 0;
    for (int32_t p : touched) {
      if (p == pv || pw[p] + g.vw[v] > max_pw) continue;
      const int64_t gn = conn[p] - conn[pv];
      if (to < 0 || gn > gain) { to = p; gain = gn; }
    }
    for (int32_t p : touched) conn[p] = 0;
    return to >= 0;
  };
  for (int32_t v = 0; v < g.n; ++v) {
    int64_t gain;
    int32_t to;
    if (best_move(v, gain, to)) pq.push({gain, v, to, 0});
  }
  struct Move { int32_t v, from, to; };
  std::vector<Move> moves;
  int64_t cur = 0, best = 0;
  size_t best_len = 0;
  const size_t patience = 64 + (size_t)g.n / 200;
  while (!pq.empty() && moves.size() - best_len < patience) {
    const Entry e = pq.top();
    pq.pop();
    if (locked[e.v] || e.stamp != stamp[e.v]) continue;
    int64_t gain;
    int32_t to;
    if (!best_move(e.v, gain, to)) continue;
    if (gain != e.gain || to != e.to) {  // neighbourhood changed since the entry was made: re-queue with the fresh gain
      pq.push({gain, e.v, to, stamp[e.v]});
      continue;
    }
    const int32_t from = part[e.v];
    part[e.v] = to;
    pw[from] -= g.vw[e.v];
    pw[to] += g.vw[e.v];
    locked[e.v] = 1;
    moves.push_back({e.v, from, to});
    cur += gain;
    if (cur > best) { best = cur; best_len = moves.size(); }
    for (int64_t f = g.ptr[e.v]; f < g.ptr[e.v + 1]; ++f) {
      const int32_t u = g.adj[f];
      if (locked[u]) continue;
      ++stamp[u];
      int64_t gu;
      int32_t tu;
      if (best_move(u, gu, tu)) pq.push({gu, u, tu, stamp[u]});
    }
  }
  for (size_t i = moves.size(); i > best_len; --i) {  // roll back the tail after the best prefix
    const Move& m = moves[i - 1];
    part[m.v] = m.from;
    pw[m.to] -= g.vw[m.v];
    pw[m.from] += g.vw[m.v];
  }
  return best;
}

// Moves vertices out of overweight parts (cheapest cut increase first come, first served) until every part fits.
void rebalance(const Graph& g, int k, int64_t max_pw, std::vector<int32_t>& part, std::vector<int64_t>& pw) {
  std::vector<int64_t> conn(k, 0);
  for (int iter = 0; iter < 4; ++iter) {
    bool over = false;
    for (int p = 0; p < k; ++p) over |= pw[p] > max_pw;
    if (!over) return;
    for (int64_t v = 0; v < g.n; ++v) {
      const int32_t pv = part[v];
      if (pw[pv] <= max_pw) continue;
      std::fill(conn.begin(), conn.end(), 0);
      for (int64_t e = g.ptr[v]; e < g.ptr[v + 1]; ++e) conn[part[g.adj[e]]] += g.ew[e];
      int32_t best = -1;
      for (int p = 0; p < k; ++p) {
        if (p == pv || pw[p] + g.vw[v] > max_pw) continue;
        if (best < 0 || conn[p] > conn[best] || (conn[p] == conn[best] && pw[p] < pw[best])) best = p;
      }
      if (best >= 0) {
        part[v] = best;
        pw[pv] -= g.vw[v];
        pw[best] += g.vw[v];
      }
    }
  }
}

// ew[e] = 1 + |N(v) & N(u)| for every edge (v, u): edges inside a dense neighbourhood become heavy, long-range "small-world"
// shortcuts (no common neighbours) stay light.  Heavy-edge matching then contracts along the locality of the graph instead of
// gluing far-apart vertices together, which is what keeps unit-weight matching from finding good cuts on such graphs.
void triangle_weights(Graph& g) {
  std::vector<int32_t> stamp(g.n, -1);
  for (int32_t v = 0; v < g.n; ++v) {
    for (int64_t e = g.ptr[v]; e < g.ptr[v + 1]; ++e) stamp[g.adj[e]] = v;
    for (int64_t e = g.ptr[v]; e < g.ptr[v + 1]; ++e) {
      const int32_t u = g.adj[e];
      int64_t common = 0;
      const int64_t b = g.ptr[u], en = std::min<int64_t>(g.ptr[u + 1], b + 256);  // cap the scan of hub neighbours
      for (int64_t f = b; f < en; ++f) common += (stamp[g.adj[f]] == v);
      g.ew[e] = 1 + common;
    }
  }
}

int64_t edge_cut(const Graph& g, const std::vector<int32_t>& part) {
  int64_t cut = 0;
  for (int64_t v = 0; v < g.n; ++v)
    for (int64_t e = g.ptr[v]; e < g.ptr[v + 1]; ++e)
      if (part[v] != part[g.adj[e]]) cut += g.ew[e];
  return cut / 2;
}

}  // namespace

// n vertices; CSR of the undirected graph (rowptr int64 [n+1], col int32); optional vertex weights (null = 1);
// k parts; imbalance = allowed max part weight / ideal (e.g. 1.03); use_triangle_weights: coarsen along common-neighbour-weighted
// edges (see triangle_weights); part_out int32 [n].  Returns the number of cut edges (>= 0) or -1.
extern "C" __attribute__((visibility("default"))) int64_t dgb_partition_kway(int64_t n, const int64_t* rowptr,
                                                                              const int32_t* col, const int64_t* vweights,
                                                                              int k, double imbalance, int refine_iters,
                                                                              int use_triangle_weights, uint64_t seed,
                                                                              int32_t* part_out) {
  if (n < 0 || k < 1 || !rowptr || !part_out || (n > 0 && rowptr[n] > 0 && !col)) return -1;
  if (k == 1 || n == 0) {
    if (n > 0) std::memset(part_out, 0, sizeof(int32_t) * (size_t)n);
    return 0;
  }
  std::mt19937_64 rng(seed);
  std::vector<Graph> levels(1);
  {
    Graph& g = levels[0];
    g.n = n;
    g.ptr.assign(rowptr, rowptr + n + 1);
    g.adj.assign(col, col + rowptr[n]);
    g.ew.assign((size_t)rowptr[n], 1);
    if (vweights) g.vw.assign(vweights, vweights + n); else g.vw.assign(n, 1);
    if (use_triangle_weights) triangle_weights(g);
  }
  const int64_t total = std::accumulate(levels[0].vw.begin(), levels[0].vw.end(), (int64_t)0);
  const int64_t ideal = (total + k - 1) / k;
  const int64_t max_pw = std::max<int64_t>((int64_t)(imbalance * (double)ideal), ideal + 1);
  const int64_t coarse_target = std::max<int64_t>(64 * (int64_t)k, 2000);
  std::vector<std::vector<int32_t>> cmaps;
  while (levels.back().n > coarse_target && levels.size() < 40) {
    const Graph& g = levels.back();
    std::vector<int32_t> cmap;
    const int64_t nc = heavy_edge_matching(g, rng, std::max<int64_t>(ideal / 8, 2), cmap);
    if (nc > (int64_t)(0.95 * (double)g.n)) break;  // matching no longer shrinks the graph (star-like remainder)
    Graph c = contract(g, cmap, nc);
    cmaps.push_back(std::move(cmap));
    levels.push_back(std::move(c));
  }
  // several region-growing starts on the (small) coarsest graph, each refined there; keep the smallest cut
  std::vector<int32_t> part;
  {
    const Graph& g = levels.back();
    int64_t best_cut = -1;
    for (int trial = 0; trial < 8; ++trial) {
      std::vector<int32_t> cand;
      if (trial & 1) initial_partition_slabs(g, k, rng, cand); else initial_partition(g, k, rng, cand);
      std::vector<int64_t> pw(k, 0);
      for (int64_t v = 0; v < g.n; ++v) pw[cand[v]] += g.vw[v];
      rebalance(g, k, max_pw, cand, pw);
      for (int it = 0; it < refine_iters; ++it)
        if (refine_pass(g, k, max_pw, rng, cand, pw) == 0) break;
      const int64_t c = edge_cut(g, cand);
      if (best_cut < 0 || c < best_cut) { best_cut = c; part.swap(cand); }
    }
  }
  for (int lvl = (int)levels.size() - 1; lvl >= 0; --lvl) {
    const Graph& g = levels[lvl];
    std::vector<int64_t> pw(k, 0);
    for (int64_t v = 0; v < g.n; ++v) pw[part[v]] += g.vw[v];
    rebalance(g, k, max_pw, part, pw);
    for (int it = 0; it < refine_iters; ++it)
      if (refine_pass(g, k, max_pw, rng, part, pw) == 0) break;
    for (int it = 0; it < 4; ++it)
      if (fm_pass(g, k, max_pw, part, pw) <= 0) break;
    rebalance(g, k, max_pw, part, pw);
    if (lvl > 0) {  // project onto the finer graph
      const std::vector<int32_t>& cmap = cmaps[lvl - 1];
      std::vector<int32_t> fine(levels[lvl - 1].n);
      for (int64_t v = 0; v < levels[lvl - 1].n; ++v) fine[v] = part[cmap[v]];
      part.swap(fine);
    }
  }
  if (use_triangle_weights) {  // the real objective is the number of cut edges: finish with unit weights on the input graph
    Graph& g = levels[0];
    std::fill(g.ew.begin(), g.ew.end(), 1);
    std::vector<int64_t> pw(k, 0);
    for (int64_t v = 0; v < g.n; ++v) pw[part[v]] += g.vw[v];
    for (int it = 0; it < refine_iters; ++it)
      if (refine_pass(g, k, max_pw, rng, part, pw) == 0) break;
    for (int it = 0; it < 4; ++it)
      if (fm_pass(g, k, max_pw, part, pw) <= 0) break;
    rebalance(g, k, max_pw, part, pw);
  }
  std::memcpy(part_out, part.data(), sizeof(int32_t) * (size_t)n);
  return edge_cut(levels[0], part);
}
