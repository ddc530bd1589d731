// plan.cu -- conflict-free shared-memory placement for the quantise kernel's permuted gather (host code, index arithmetic only).
//
// The kernel stages R rows of a tile channel-major ("[channel][R rows]", one 2R-byte chunk per channel) so that the permuted
// gather of reorder.cu:115-118 is one shared load per channel for all R rows.  WHERE a channel's chunk sits in that buffer is
// free, and it decides the bank conflicts of both accesses that touch it:
//   * the transposing STORE: one wavefront = D = 64/R lanes, lane T of sub-step (k, j) holds channel
//     (256/R) k + 32 j + 8 (T % 4) + T / 4 (what ldmatrix.trans leaves in the lane);
//   * the permuted LOAD: one wavefront = the D consecutive 16-groups of a group block at step i, i.e. channels
//     reorder_index[16 (D gb + j) + i], j < D.
// Every channel is in exactly one store wavefront and one load wavefront, both kinds hold D channels, and a wavefront is
// conflict-free iff its D chunks fall into D different 2R-byte bank slots.  That is a proper D-edge-colouring of a D-regular
// bipartite multigraph (store wavefronts x load wavefronts, one edge per channel), which always exists (Koenig) and is
// found here by repeated Euler splits, O(KQ log D).  position(c) = D * (rank of c inside its colour class) + colour(c).
// A random reorder_index costs 2.3 - 3.5 wavefronts per gather instruction without this (measured in round 1: 42 % of all
// shared-memory wavefronts of the kernel were bank conflicts); with it every wavefront is one pass.
//
// The plan is a pure function of reorder_index; the kernel only needs it to be a bijection of [0, KQ) (it builds its gather
// offsets from plan[reorder_index[k]] itself), so a stale plan can cost speed but never correctness.
#include <stdint.h>

#include <vector>

#include "arcq.h"
#include "common.cuh"

namespace arcq {

// tile heights with a plan section, in section order; a section is usable iff KQ % (16 * 64 / R) == 0
static const int kPlanR[3] = {8, 4, 2};

bool plan_usable(int R, int KQ) {
  if (R != 8 && R != 4 && R != 2) return false;
  const int D = 64 / R;
  return KQ > 0 && KQ % (16 * D) == 0 && KQ <= 32768;
}
int plan_section(int R) { return R == 8 ? 0 : R == 4 ? 1 : 2; }

static inline int store_node(int c, int R) {  // store wavefront of channel c
  const int D = 64 / R, CW = 256 / R;
  const int k = c / CW, rem = c % CW, j = rem / 32, w = rem % 32;
  const int T = (w / 8) + 4 * (w % 8);          // w = 8 (T % 4) + T / 4  ->  T % 4 = w / 8, T / 4 = w % 8
  return (k * (CW / 32) + j) * (32 / D) + T / D;
}

// Euler split of a multigraph whose nodes all have even degree: edges are 2-coloured so that every node keeps exactly half
// of its edges in each part.  Bipartite => every closed trail has even length => alternating labels balance every node.
static void euler_split(const std::vector<int> &edges, const std::vector<int> &eu, const std::vector<int> &ev, int nleft, int nright,
                        std::vector<int> &part_a, std::vector<int> &part_b) {
  const int nn = nleft + nright;
  std::vector<int> head(nn + 1, 0);
  for (int e : edges) { ++head[eu[e] + 1]; ++head[nleft + ev[e] + 1]; }
  for (int i = 0; i < nn; ++i) head[i + 1] += head[i];
  std::vector<int> adj(head[nn]), fill(head.begin(), head.end() - 1);
  for (int e : edges) { adj[fill[eu[e]]++] = e; adj[fill[nleft + ev[e]]++] = e; }
  std::vector<int> cur(head.begin(), head.end() - 1);
  std::vector<char> used;
  int maxe = 0;
  for (int e : edges) if (e > maxe) maxe = e;
  used.assign(maxe + 1, 0);
  for (int start = 0; start < nleft; ++start) {
    for (;;) {
      // next unused edge at the start node
      while (cur[start] < head[start + 1] && used[adj[cur[start]]]) ++cur[start];
      if (cur[start] >= head[start + 1]) break;
      int node = start, label = 0;
      for (;;) {  // walk until stuck (which, with even degrees, happens back at `start`)
        while (cur[node] < head[node + 1] && used[adj[cur[node]]]) ++cur[node];
        if (cur[node] >= head[node + 1]) break;
        const int e = adj[cur[node]++];
        used[e] = 1;
        (label ? part_b : part_a).push_back(e);
        label ^= 1;
        node = (node < nleft) ? nleft + ev[e] : eu[e];
      }
    }
  }
}

static void colour(const std::vector<int> &edges, const std::vector<int> &eu, const std::vector<int> &ev, int nleft, int nright, int degree,
                   int base, std::vector<int> &col) {
  if (degree == 1) { for (int e : edges) col[e] = base; return; }
  std::vector<int> a, b;
  a.reserve(edges.size() / 2); b.reserve(edges.size() / 2);
  euler_split(edges, eu, ev, nleft, nright, a, b);
  colour(a, eu, ev, nleft, nright, degree / 2, base, col);
  colour(b, eu, ev, nleft, nright, degree / 2, base + degree / 2, col);
}

// worst number of chunks of one wavefront that share a bank slot, over all store and load wavefronts (1 = conflict-free)
static int max_conflict(const int16_t *idx, int KQ, int R, const uint16_t *pos) {
  const int D = 64 / R;
  int worst = 0;
  std::vector<int> cnt(D);
  const int nstore = KQ / D;
  std::vector<std::vector<int>> st(nstore);
  for (int c = 0; c < KQ; ++c) st[store_node(c, R)].push_back(c);
  for (auto &v : st) {
    if ((int)v.size() != D) return 1 << 20;
    std::fill(cnt.begin(), cnt.end(), 0);
    for (int c : v) { int n = ++cnt[pos[c] % D]; if (n > worst) worst = n; }
  }
  for (int gb = 0; gb < KQ / 16 / D; ++gb)
    for (int i = 0; i < 16; ++i) {
      std::fill(cnt.begin(), cnt.end(), 0);
      for (int j = 0; j < D; ++j) { int n = ++cnt[pos[(uint16_t)idx[16 * (gb * D + j) + i]] % D]; if (n > worst) worst = n; }
    }
  return worst;
}

}  // namespace arcq

using namespace arcq;

// per tile height (8, 4, 2) one section of 2 KQ entries: pos[KQ] (position of channel c) followed by tabt[KQ] (the same
// positions in the order the transposing lanes need them: tabt[(k 32 + T) J + j] = pos[(256/R) k + 32 j + 8 (T % 4) + T / 4],
// J = 8/R), so that the kernel fetches both with one bulk copy each instead of KQ dependent L2 lookups per CTA
extern "C" int64_t arcq_plan_entries(int KQ) { return KQ > 0 ? 6 * (int64_t)KQ : 0; }

extern "C" int arcq_plan_build(const int16_t *idx_host, int KQ, uint16_t *plan_host) {
  if (!idx_host || !plan_host || KQ <= 0 || KQ > 32768 || KQ % 16) { set_error("arcq_plan_build: bad argument (KQ=%d)", KQ); return ARCQ_EINVAL; }
  std::vector<char> seen(KQ, 0);
  for (int k = 0; k < KQ; ++k) {
    const int c = (uint16_t)idx_host[k];
    if (c >= KQ || seen[c]) { set_error("arcq_plan_build: reorder_index is not a permutation of [0, %d)", KQ); return ARCQ_EINVAL; }
    seen[c] = 1;
  }
  for (int s = 0; s < 3; ++s) {
    const int R = kPlanR[s], D = 64 / R;
    uint16_t *pos = plan_host + (int64_t)s * 2 * KQ, *tabt = pos + KQ;
    for (int c = 0; c < KQ; ++c) tabt[c] = (uint16_t)c;
    if (!plan_usable(R, KQ)) {  // identity placement: a bijection, never used by the launcher for this tile height
      for (int c = 0; c < KQ; ++c) pos[c] = (uint16_t)c;
      continue;
    }
    const int nnodes = KQ / D;
    std::vector<int> eu(KQ), ev(KQ), col(KQ, -1), edges(KQ);
    for (int c = 0; c < KQ; ++c) { eu[c] = store_node(c, R); edges[c] = c; }
    for (int k = 0; k < KQ; ++k) {
      const int g = k / 16, i = k % 16;
      ev[(uint16_t)idx_host[k]] = (g / D) * 16 + i;
    }
    colour(edges, eu, ev, nnodes, nnodes, D, 0, col);
    std::vector<int> rank(D, 0);
    for (int c = 0; c < KQ; ++c) {
      if (col[c] < 0 || col[c] >= D) { set_error("arcq_plan_build: colouring failed"); return ARCQ_EINVAL; }
      pos[c] = (uint16_t)(D * rank[col[c]]++ + col[c]);
    }
    if (max_conflict(idx_host, KQ, R, pos) != 1) { set_error("arcq_plan_build: plan for R=%d is not conflict-free", R); return ARCQ_EINVAL; }
    const int CW = 256 / R, J = CW / 32;
    for (int u = 0; u < KQ; ++u) {
      const int j = u % J, T = (u / J) & 31, k = u / (32 * J);
      tabt[u] = pos[CW * k + 32 * j + 8 * (T & 3) + (T >> 2)];
    }
  }
  return ARCQ_OK;
}

// worst bank-slot multiplicity of any store / load wavefront under `plan_host` for tile height R (1 = conflict-free; the
// default arithmetic placement of a random permutation gives 3 - 6); -1 if R has no usable section.  For tests and tools.
extern "C" int arcq_plan_max_conflict(const int16_t *idx_host, int KQ, int R, const uint16_t *plan_host) {
  if (!idx_host || !plan_host || !plan_usable(R, KQ)) return -1;
  const uint16_t *pos = plan_host + (int64_t)plan_section(R) * 2 * KQ;
  std::vector<char> seen(KQ, 0);
  for (int c = 0; c < KQ; ++c) { if (pos[c] >= KQ || seen[pos[c]]) return 1 << 20; seen[pos[c]] = 1; }
  return max_conflict(idx_host, KQ, R, pos);
}
