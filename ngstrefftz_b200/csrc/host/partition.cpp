#include "partition.hpp"

#include <algorithm>
#include <stdexcept>

namespace ngst {

int PartitionPlan::max_rows_per_level() const {
  int mx = 0;
  for (int l = 0; l < nlev; l++) {
    mx = std::max(mx, send_off[(size_t)(l + 1) * nranks] - send_off[(size_t)l * nranks]);
    mx = std::max(mx, recv_off[(size_t)(l + 1) * nranks] - recv_off[(size_t)l * nranks]);
  }
  return mx;
}

PartitionPlan make_partition_plan(const SimplexMesh& m, const TentSlab& slab, const std::vector<int>& vowner, int nranks,
                                  int rank) {
  if (nranks < 1 || rank < 0 || rank >= nranks) throw std::invalid_argument("partition: bad rank / nranks");
  if ((int)vowner.size() != m.nv) throw std::invalid_argument("partition: need one owner per vertex");
  for (int o : vowner)
    if (o < 0 || o >= nranks) throw std::invalid_argument("partition: vertex owner out of range");
  if (!slab.pitched) throw std::invalid_argument("partition: the slab has not been pitched");
  const int d1 = m.D + 1;
  PartitionPlan p;
  p.rank = rank;
  p.nranks = nranks;
  p.nlev = (int)slab.level_off.size() - 1;

  for (int t = 0; t < slab.ntents; t++)
    if (vowner[slab.vertex[t]] == rank) p.own_tents.push_back(t);

  p.glob2loc.assign(m.nel, -1);
  for (int e = 0; e < m.nel; e++) {
    const int* ev = m.elverts(e);
    bool local = false;
    int lowest = ev[0];
    for (int k = 0; k < d1; k++) {
      local = local || vowner[ev[k]] == rank;
      lowest = std::min(lowest, ev[k]);
    }
    if (local) {
      p.glob2loc[e] = (int)p.loc2glob.size();
      p.loc2glob.push_back(e);
      p.owned.push_back(vowner[lowest] == rank ? 1 : 0);
    }
  }

  p.final_level.assign(p.loc2glob.size(), -1);
  for (int t = 0; t < slab.ntents; t++)
    for (int j = slab.els.off[t]; j < slab.els.off[t + 1]; j++) {
      const int le = p.glob2loc[slab.els.val[j]];
      if (le >= 0) p.final_level[le] = std::max(p.final_level[le], slab.level[t]);
    }

  // exchange lists, level by level
  std::vector<std::vector<int>> snd(nranks), rcv(nranks);
  p.send_off.assign(1, 0);
  p.recv_off.assign(1, 0);
  for (int l = 0; l < p.nlev; l++) {
    for (auto& v : snd) v.clear();
    for (auto& v : rcv) v.clear();
    for (int i = slab.level_off[l]; i < slab.level_off[l + 1]; i++) {
      const int t = slab.level_tents[i];
      const int r = vowner[slab.vertex[t]];
      for (int j = slab.els.off[t]; j < slab.els.off[t + 1]; j++) {
        const int e = slab.els.val[j];
        const int* ev = m.elverts(e);
        for (int k = 0; k < d1; k++) {
          const int q = vowner[ev[k]];
          if (q == r) continue;
          bool seen = false;
          for (int kk = 0; kk < k; kk++) seen = seen || vowner[ev[kk]] == q;
          if (seen) continue;
          if (r == rank) snd[q].push_back(e);
          else if (q == rank) rcv[r].push_back(e);
        }
      }
    }
    for (int q = 0; q < nranks; q++) {
      std::sort(snd[q].begin(), snd[q].end());
      std::sort(rcv[q].begin(), rcv[q].end());
      for (int e : snd[q]) p.send_el.push_back(p.glob2loc[e]);
      for (int e : rcv[q]) p.recv_el.push_back(p.glob2loc[e]);
      p.send_off.push_back((int)p.send_el.size());
      p.recv_off.push_back((int)p.recv_el.size());
    }
  }
  // ---- cross-rank hazards and push lists
  {
    std::vector<int> towner(slab.ntents), tindex(slab.ntents), cnt(nranks, 0);
    for (int t = 0; t < slab.ntents; t++) {
      towner[t] = vowner[slab.vertex[t]];
      tindex[t] = cnt[towner[t]]++;
    }
    const int nown = (int)p.own_tents.size();
    p.p2p_indeg.assign(nown, 0);
    std::vector<std::vector<int>> lsucc(nown), rsucc(nown);
    std::vector<int> last(m.nel, -1), preds;
    for (int t = 0; t < slab.ntents; t++) {
      preds.clear();
      for (int j = slab.els.off[t]; j < slab.els.off[t + 1]; j++) {
        const int e = slab.els.val[j];
        if (last[e] >= 0) preds.push_back(last[e]);
        last[e] = t;
      }
      std::sort(preds.begin(), preds.end());
      preds.erase(std::unique(preds.begin(), preds.end()), preds.end());
      if (towner[t] == rank) p.p2p_indeg[tindex[t]] = (int)preds.size();
      for (int q : preds) {
        if (slab.level[q] >= slab.level[t]) throw std::invalid_argument("partition: tent levels do not respect the element hazards");
        if (towner[q] != rank) continue;
        (towner[t] == rank ? lsucc : rsucc)[tindex[q]].push_back(t);
      }
    }
    p.p2p_lsucc_off.assign(1, 0);
    p.p2p_rsucc_off.assign(1, 0);
    for (int i = 0; i < nown; i++) {
      for (int t : lsucc[i]) p.p2p_lsucc.push_back(tindex[t]);
      for (int t : rsucc[i]) { p.p2p_rsucc_peer.push_back(towner[t]); p.p2p_rsucc_tent.push_back(tindex[t]); }
      p.p2p_lsucc_off.push_back((int)p.p2p_lsucc.size());
      p.p2p_rsucc_off.push_back((int)p.p2p_rsucc_peer.size());
    }
    // local element numbering of every rank: ascending global element among the elements with a vertex it owns
    std::vector<int> nloc(nranks, 0);
    std::vector<int> holders_off(m.nel + 1, 0), holders, holder_loc;
    for (int e = 0; e < m.nel; e++) {
      const int* ev = m.elverts(e);
      for (int k = 0; k < d1; k++) {
        const int q = vowner[ev[k]];
        bool seen = false;
        for (int kk = 0; kk < k; kk++) seen = seen || vowner[ev[kk]] == q;
        if (seen) continue;
        holders.push_back(q);
        holder_loc.push_back(nloc[q]++);
      }
      holders_off[e + 1] = (int)holders.size();
    }
    p.p2p_push_off.assign(1, 0);
    for (int t : p.own_tents)
      for (int j = slab.els.off[t]; j < slab.els.off[t + 1]; j++) {
        const int e = slab.els.val[j];
        for (int h = holders_off[e]; h < holders_off[e + 1]; h++)
          if (holders[h] != rank) { p.p2p_push_peer.push_back(holders[h]); p.p2p_push_el.push_back(holder_loc[h]); }
        p.p2p_push_off.push_back((int)p.p2p_push_peer.size());
      }
  }
  return p;
}

TentSlab restrict_slab(const TentSlab& s, const std::vector<int>& own) {
  TentSlab r;
  r.mesh = s.mesh;
  r.method = s.method;
  r.cmax_el = s.cmax_el;
  r.dt = s.dt;
  r.pitched = s.pitched;
  r.nlayers = s.nlayers;
  r.ntents = (int)own.size();
  std::vector<int> g2l(s.ntents, -1);
  for (int i = 0; i < r.ntents; i++) g2l[own[i]] = i;
  auto sub = [&](const Csr& c, Csr& out) {
    out.off.assign(1, 0);
    out.val.clear();
    for (int t : own) {
      out.val.insert(out.val.end(), c.val.begin() + c.off[t], c.val.begin() + c.off[t + 1]);
      out.off.push_back((int)out.val.size());
    }
  };
  for (int t : own) {
    r.vertex.push_back(s.vertex[t]);
    r.level.push_back(s.level[t]);
    r.tbot.push_back(s.tbot[t]);
    r.ttop.push_back(s.ttop[t]);
    if (!s.maxslope.empty()) r.maxslope.push_back(s.maxslope[t]);
    r.nbtime.insert(r.nbtime.end(), s.nbtime.begin() + s.nbv.off[t], s.nbtime.begin() + s.nbv.off[t + 1]);
  }
  sub(s.nbv, r.nbv);
  sub(s.els, r.els);
  sub(s.facets, r.facets);
  r.dag.off.assign(1, 0);
  for (int t : own) {
    for (int j = s.dag.off[t]; j < s.dag.off[t + 1]; j++)
      if (g2l[s.dag.val[j]] >= 0) r.dag.val.push_back(g2l[s.dag.val[j]]);
    r.dag.off.push_back((int)r.dag.val.size());
  }
  const int nlev = (int)s.level_off.size() - 1;
  r.level_off.assign(1, 0);
  for (int l = 0; l < nlev; l++) {
    for (int i = s.level_off[l]; i < s.level_off[l + 1]; i++)
      if (g2l[s.level_tents[i]] >= 0) r.level_tents.push_back(g2l[s.level_tents[i]]);
    r.level_off.push_back((int)r.level_tents.size());
  }
  return r;
}

}  // namespace ngst
