#include "solver.hpp"
#include "sample.hpp"

#include <dlfcn.h>

#include <algorithm>
#include <cmath>
#include <cstring>
#include <stdexcept>

namespace ngst {

namespace {
struct CudaFailure : std::runtime_error {
  using std::runtime_error::runtime_error;
};
inline void cuda_check(cudaError_t e, const char* what) {
  if (e != cudaSuccess) throw CudaFailure(std::string("CUDA: ") + what + ": " + cudaGetErrorString(e));
}
#define CU(x) cuda_check((x), #x)
}  // namespace

template <typename T>
T* Solver::Upload(const std::vector<T>& v) {
  T* d = nullptr;
  const size_t bytes = std::max<size_t>(v.size(), 1) * sizeof(T);
  CU(cudaMalloc(&d, bytes));
  allocs_.push_back(d);
  if (!v.empty()) CU(cudaMemcpy(d, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice));
  return d;
}

Solver::Solver(const TentSlab& slab, int order, const std::vector<double>& c_el, const std::vector<double>& c_vertex,
               int device, const PartitionPlan* plan, const QtHostData* qt)
    : mesh_(*slab.mesh), slab_(slab), plan_(plan), order_(order), device_(device), c_el_(c_el), c_vertex_(c_vertex) {
  qt_ = qt != nullptr;
  if (qt_ && (mesh_.D > 2 || order < 2))
    throw std::invalid_argument("TWave: the quasi-Trefftz solver exists for D = 1, 2 (twavetents.cpp:1150-1151) and order >= 2");
  if (order < 1 || order > 12) throw std::invalid_argument("TWave: order must be in 1..12");
  if (slab.ntents == 0 && !plan) throw std::logic_error("TWave: the slab has no tents (call PitchTents first)");
  if ((int)c_el_.size() != mesh_.nel) throw std::invalid_argument("TWave: need one wavespeed per element");
  for (double c : c_el_)
    if (!(c > 0)) throw std::invalid_argument("TWave: wavespeed must be positive");
  if (c_vertex_.empty()) {
    c_vertex_.resize(mesh_.nv);
    for (int v = 0; v < mesh_.nv; v++) c_vertex_[v] = c_el_[mesh_.v2el.row(v)[0]];
  }
  const int D = mesh_.D, d1 = D + 1;
  if (plan_) loc2glob_ = plan_->loc2glob;
  else { loc2glob_.resize(mesh_.nel); for (int e = 0; e < mesh_.nel; e++) loc2glob_[e] = e; }
  nel_ = (int)loc2glob_.size();

  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0)
    throw CudaFailure(std::string("CUDA: no usable device (") + cudaGetErrorString(e) +
                      "); this library has no CPU fallback");
  if (device < 0 || device >= ndev) throw CudaFailure("CUDA: device ordinal out of range");
  CU(cudaSetDevice(device));
  CU(cudaStreamCreateWithFlags(&stream_, cudaStreamNonBlocking));
  CU(cudaEventCreate(&ev0_));
  CU(cudaEventCreate(&ev1_));
  CU(cudaEventCreateWithFlags(&ev_done_, cudaEventDisableTiming));

  rule_ = make_simplex_rule(D, 2 * order);
  basis_so_ = twave_basis(d1, order, 0);
  basis_fo_ = twave_basis(d1, order, 1);
  basis_ = basis_so_;
  staged_ = stage_tents(mesh_, slab_, c_el_, c_vertex_, rule_, qt_);

  const int Q = rule_.nip;
  std::vector<double> lam((size_t)Q * d1);
  for (int q = 0; q < Q; q++) {
    double s = 0;
    for (int d = 0; d < D; d++) { lam[(size_t)q * d1 + d] = rule_.xi[(size_t)q * D + d]; s += rule_.xi[(size_t)q * D + d]; }
    lam[(size_t)q * d1 + D] = 1.0 - s;
  }

  TentKernelArgs& a = args_;
  a.pts = Upload(mesh_.pts);
  {
    std::vector<int> el2v_loc((size_t)nel_ * d1);
    std::vector<double> c_loc(nel_);
    for (int e = 0; e < nel_; e++) {
      for (int k = 0; k < d1; k++) el2v_loc[(size_t)e * d1 + k] = mesh_.el2v[(size_t)loc2glob_[e] * d1 + k];
      c_loc[e] = c_el_[loc2glob_[e]];
    }
    a.el2v = Upload(el2v_loc);
    a.c_el = Upload(c_loc);
  }
  a.lam = Upload(lam);
  a.w = Upload(rule_.w);
  a.expo = Upload(basis_so_.expo);
  for (int f = 0; f < 2; f++) {
    const BasisCsr& b = f ? basis_fo_ : basis_so_;
    d_rowptr_[f] = Upload(b.rowptr);
    d_col_[f] = Upload(b.col);
    d_val_[f] = Upload(b.val);
  }
  a.npoly = basis_so_.npoly;
  a.order = order;
  a.Q = Q;
  a.center = Upload(staged_.center);
  a.scale = Upload(staged_.scale);
  a.ndom = Upload(staged_.ndom);
  a.te_off = Upload(staged_.te_off);
  a.tf_off = Upload(staged_.tf_off);
  {
    std::vector<int> te_el_loc(staged_.te_el);  // staged_ keeps global element ids; the device sees local ones
    if (plan_)
      for (int& e : te_el_loc) e = plan_->glob2loc[e];
    a.te_el = Upload(te_el_loc);
  }
  a.te_dom = Upload(staged_.te_dom);
  a.te_tbot = Upload(staged_.te_tbot);
  a.te_ttop = Upload(staged_.te_ttop);
  a.te_geo = Upload(staged_.te_geo);
  a.tf_kind = Upload(staged_.tf_kind);
  a.tf_dom = Upload(staged_.tf_dom);
  a.tf_c = Upload(staged_.tf_c);
  a.tf_verts = Upload(staged_.tf_verts);
  a.tf_n = Upload(staged_.tf_n);
  a.tf_area = Upload(staged_.tf_area);
  a.tf_data = Upload(staged_.tf_data);
  a.bnd_data = nullptr;
  a.max_ndom = staged_.max_ndom;
  // ---- kernel routing: single-wavespeed tents run on the fast DMMA kernel when a variant exists for (D, order)
  fast_ = !qt_ && tent_fast_available(D, order);
  if (fast_) {
    fstage_ = stage_fast(mesh_, slab_, staged_, c_el_, rule_, basis_so_, basis_fo_, fast_fe_stride(D), fast_ff_stride(D));
    fargs_.fe = Upload(fstage_.fe);
    fargs_.ff = Upload(fstage_.ff);
    fargs_.sqrtw = Upload(fstage_.sqrtw);
    {
      // top-face Gram: compact rule exact to degree 2p-2 (the degree of the integrand, tent_fast.cu Cfg::QT)
      const SimplexRule rt = make_compact_rule(D, 2 * order - 2);
      std::vector<double> lt((size_t)rt.nip * d1), sw(rt.nip);
      for (int q = 0; q < rt.nip; q++) {
        double s = 0;
        for (int d = 0; d < D; d++) { lt[(size_t)q * d1 + d] = rt.xi[(size_t)q * D + d]; s += rt.xi[(size_t)q * D + d]; }
        lt[(size_t)q * d1 + D] = 1.0 - s;
        sw[q] = std::sqrt(rt.w[q]);
      }
      fargs_.lam_top = Upload(lt);
      fargs_.sqrtw_top = Upload(sw);
      fargs_.Qt = rt.nip;
    }
    if (compact_rule_points(D, 2 * order) > 0) {
      // stabilisation Gram of the bottom faces (degree 2p): compact rule in a pass of its own (tent_fast.cu Cfg::QS)
      const SimplexRule rs = make_compact_rule(D, 2 * order);
      std::vector<double> ls((size_t)rs.nip * d1), sw(rs.nip);
      for (int q = 0; q < rs.nip; q++) {
        double s = 0;
        for (int d = 0; d < D; d++) { ls[(size_t)q * d1 + d] = rs.xi[(size_t)q * D + d]; s += rs.xi[(size_t)q * D + d]; }
        ls[(size_t)q * d1 + D] = 1.0 - s;
        sw[q] = std::sqrt(rs.w[q]);
      }
      fargs_.lam_stab = Upload(ls);
      fargs_.sqrtw_stab = Upload(sw);
      fargs_.w_stab = Upload(rs.w);
      fargs_.Qs = rs.nip;
    }
    fargs_.dmap = Upload(fstage_.dmap);
    fargs_.dfac = Upload(fstage_.dfac);
    for (int f = 0; f < 2; f++) {
      d_cptr_[f] = Upload(fstage_.cptr[f]);
      d_crow_[f] = Upload(fstage_.crow[f]);
      d_cval_[f] = Upload(fstage_.cval[f]);
    }
  }
  // [0] singular pivots, [1] dependency time-outs of the single-launch mode; copied into page-locked host words by
  // an asynchronous 8-byte copy at the end of every Propagate (no extra synchronisation)
  CU(cudaMalloc(&d_singular_, 2 * sizeof(int)));
  allocs_.push_back(d_singular_);
  CU(cudaMemset(d_singular_, 0, 2 * sizeof(int)));
  CU(cudaHostAlloc(&h_flags_, 2 * sizeof(int), cudaHostAllocDefault));
  h_flags_[0] = h_flags_[1] = 0;
  a.singular = d_singular_;
  a.stuck = d_singular_ + 1;

  if (qt_) {
    const int nv = mesh_.nv;
    const std::vector<int> eA = monomial_exponents(D, order - 2), eB = monomial_exponents(D, order - 1);
    const int nA = (int)eA.size() / D, nB = (int)eB.size() / D;
    if ((int)qt->aa.size() != nv * nA || (int)qt->bb.size() != nv * nB)
      throw std::invalid_argument("TWave: Taylor tables must hold nv x (#monomials of degree <= order-2 / order-1) values");
    auto find = [&](const std::vector<int>& ex, int dim, const int* e) {
      for (int j = 0; j < (int)ex.size() / dim; j++) {
        bool same = true;
        for (int k = 0; k < dim; k++) same = same && ex[(size_t)j * dim + k] == e[k];
        if (same) return j;
      }
      return -1;
    };
    const int P = order, P1 = P + 1;
    std::vector<int> idxA((size_t)(P - 1) * (P - 1), -1), idxB((size_t)P * P, -1), idx3((size_t)P1 * P1 * P1, -1);
    for (int nx = 0; nx < P; nx++)
      for (int ny = 0; ny < P; ny++) {
        if (D == 1 && ny > 0) continue;
        const int e[2] = {nx, ny};
        if (nx + ny <= P - 1) idxB[(size_t)nx * P + ny] = find(eB, D, e);
        if (nx < P - 1 && ny < P - 1 && nx + ny <= P - 2) idxA[(size_t)nx * (P - 1) + ny] = find(eA, D, e);
      }
    for (int x = 0; x <= P; x++)
      for (int y = 0; y <= P; y++)
        for (int t = 0; t <= P; t++) {
          if (x + y + t > P || (D == 1 && y > 0)) continue;
          const int e3[3] = {x, y, t}, e2[2] = {x, t};
          idx3[((size_t)x * P1 + y) * P1 + t] = find(basis_so_.expo, d1, D == 1 ? e2 : e3);
        }
    // wavespeed polynomial: non-zero terms only
    const std::vector<int> eC = monomial_exponents(D, qt->cdeg);
    if ((int)qt->ccoef.size() != (int)eC.size() / D) throw std::invalid_argument("TWave: wavespeed polynomial has the wrong number of coefficients");
    std::vector<double> cc;
    std::vector<int> ce;
    for (size_t j = 0; j < qt->ccoef.size(); j++)
      if (qt->ccoef[j] != 0.0) {
        cc.push_back(qt->ccoef[j]);
        for (int k = 0; k < D; k++) ce.push_back(eC[j * D + k]);
      }
    if (cc.empty()) throw std::invalid_argument("TWave: the wavespeed polynomial is zero");
    const SimplexRule vrule = make_simplex_rule(d1, 2 * order);
    std::vector<double> tsize(slab_.ntents);
    for (int t = 0; t < slab_.ntents; t++) tsize[t] = 1.0 / staged_.scale[t];
    QtArgs& q = qargs_;
    q.ccoef = Upload(cc);
    q.cexpo = Upload(ce);
    q.ncterm = (int)cc.size();
    q.aa = Upload(qt->aa);
    q.bb = Upload(qt->bb);
    q.nA = nA;
    q.nB = nB;
    q.idxA = Upload(idxA);
    q.idxB = Upload(idxB);
    q.idx3 = Upload(idx3);
    q.tent_vertex = Upload(slab_.vertex);
    q.tentsize = Upload(tsize);
    q.vxi = Upload(vrule.xi);
    q.vw = Upload(vrule.w);
    q.Qv = vrule.nip;
    SetWavespeedPoly(qt->cdeg, qt->ccoef.data());
    q.c_point = d_cpoint_;
  }
  if (qt_ && tent_qt_fast_available(D, order)) {
    // fast quasi-Trefftz kernel: per-slab records and routing flags computed on the device (tent_qt_fast.cu)
    double *fe = nullptr, *cone = nullptr, *ff = nullptr;
    int* ok = nullptr;
    CU(cudaMalloc(&fe, sizeof(double) * std::max<size_t>((size_t)staged_.nte * qt_fast_fe_stride(), 1)));
    allocs_.push_back(fe);
    CU(cudaMalloc(&cone, sizeof(double) * std::max<size_t>((size_t)staged_.nte * 2 * qt_fast_cone_stride(), 1)));
    allocs_.push_back(cone);
    CU(cudaMalloc(&ff, sizeof(double) * std::max<size_t>((size_t)staged_.ntf * qt_fast_ff_stride(), 1)));
    allocs_.push_back(ff);
    CU(cudaMalloc(&ok, sizeof(int) * std::max(slab_.ntents, 1)));
    allocs_.push_back(ok);
    CU(launch_qt_fast_stage(args_, qargs_, slab_.ntents, fe, cone, ff, ok, stream_));
    qt_ok_.assign(slab_.ntents, 0);
    CU(cudaMemcpyAsync(qt_ok_.data(), ok, sizeof(int) * slab_.ntents, cudaMemcpyDeviceToHost, stream_));
    CU(cudaStreamSynchronize(stream_));
    qfargs_.fe = fe;
    qfargs_.cone = cone;
    qfargs_.ff = ff;
    qtfast_ = true;
  }
  {
    const int nlev = (int)slab_.level_off.size() - 1;
    std::vector<int> gen, fast;
    gen_off_.assign(1, 0);
    fast_off_.assign(1, 0);
    fast_nhalo_.assign(nlev, 0);
    gen_nhalo_.assign(nlev, 0);
    std::vector<char> sent(nel_, 0);
    for (int l = 0; l < nlev; l++) {
      const size_t f0 = fast.size(), g0 = gen.size();
      // halo tents (multi-GPU): the tents of this level that write a row another rank holds. They run first, on their own
      // stream, so that the exchange of the level overlaps the interior tents (Propagate()).
      int s0 = 0, s1 = 0;
      if (plan_ && plan_->nranks > 1) {
        s0 = plan_->send_off[(size_t)l * plan_->nranks];
        s1 = plan_->send_off[(size_t)(l + 1) * plan_->nranks];
        for (int i = s0; i < s1; i++) sent[plan_->send_el[i]] = 1;
      }
      auto is_halo = [&](int t) {
        if (s1 == s0) return false;
        for (int te = staged_.te_off[t]; te < staged_.te_off[t + 1]; te++)
          if (sent[plan_->glob2loc[staged_.te_el[te]]]) return true;
        return false;
      };
      std::vector<int> fh, fi, gh, gi;
      for (int i = slab_.level_off[l]; i < slab_.level_off[l + 1]; i++) {
        const int t = slab_.level_tents[i];
        const bool h = is_halo(t);
        if ((fast_ && fstage_.ok[t]) || (qtfast_ && qt_ok_[t])) (h ? fh : fi).push_back(t);
        else (h ? gh : gi).push_back(t);
      }
      for (int i = s0; i < s1; i++) sent[plan_->send_el[i]] = 0;
      // most expensive tents first (the order inside a level is free: tents of a level share no element)
      auto by_cost = [&](int x, int y) {
        return staged_.te_off[x + 1] - staged_.te_off[x] > staged_.te_off[y + 1] - staged_.te_off[y];
      };
      std::stable_sort(fh.begin(), fh.end(), by_cost);
      std::stable_sort(fi.begin(), fi.end(), by_cost);
      fast.insert(fast.end(), fh.begin(), fh.end());
      fast.insert(fast.end(), fi.begin(), fi.end());
      // peer-to-peer exchange (one launch per slab): the halo tents of a level behind the first third of its interior tents.
      // A halo tent may have to wait for a tent of another rank; with all of them first, every CTA slot of the GPU holds
      // a waiting tent whenever the neighbour is a little behind (560 halo tents per level at config 4 on 8 GPUs, 444 slots).
      {
        const size_t k = fi.size() / 3;
        p2p_order_.insert(p2p_order_.end(), fi.begin(), fi.begin() + k);
        p2p_order_.insert(p2p_order_.end(), fh.begin(), fh.end());
        p2p_order_.insert(p2p_order_.end(), fi.begin() + k, fi.end());
      }
      gen.insert(gen.end(), gh.begin(), gh.end());
      gen.insert(gen.end(), gi.begin(), gi.end());
      fast_nhalo_[l] = (int)fh.size();
      gen_nhalo_[l] = (int)gh.size();
      (void)f0; (void)g0;
      gen_off_.push_back((int)gen.size());
      fast_off_.push_back((int)fast.size());
    }
    d_level_tents_ = Upload(gen);
    d_level_all_ = Upload(slab_.level_tents);
    d_level_fast_ = Upload(fast);
    if (plan_ && plan_->nranks > 1) d_p2p_tents_ = Upload(p2p_order_);
    // Dependency-driven scheduling (single GPU, every tent on the fast kernel): tent t must wait for the last earlier tent
    // (pitching order = the reference's serial order) that touched each of its elements -- exactly the read/write
    // hazards of the state rows. One launch per slab then replaces the per-level launches and their tails.
    if (fast_ && !qt_ && !plan_ && gen.empty() && slab_.ntents > 0) {
      const int nt = slab_.ntents;
      std::vector<int> last(mesh_.nel, -1), indeg(nt, 0), soff(nt + 1, 0);
      std::vector<std::vector<int>> pred(nt);
      for (int t = 0; t < nt; t++) {
        for (int te = staged_.te_off[t]; te < staged_.te_off[t + 1]; te++) {
          const int e = staged_.te_el[te];
          if (last[e] >= 0) pred[t].push_back(last[e]);
          last[e] = t;
        }
        std::sort(pred[t].begin(), pred[t].end());
        pred[t].erase(std::unique(pred[t].begin(), pred[t].end()), pred[t].end());
        indeg[t] = (int)pred[t].size();
        for (int p : pred[t]) {
          if (slab_.level[p] >= slab_.level[t]) throw std::logic_error("tent levels do not respect the element hazards");
          soff[p + 1]++;
        }
      }
      for (int t = 0; t < nt; t++) soff[t + 1] += soff[t];
      std::vector<int> succ(soff[nt]), fill(soff.begin(), soff.end() - 1);
      for (int t = 0; t < nt; t++)
        for (int p : pred[t]) succ[fill[p]++] = t;
      d_indeg_ = Upload(indeg);
      d_succ_off_ = Upload(soff);
      d_succ_ = Upload(succ);
      CU(cudaMalloc(&d_remaining_, sizeof(int) * nt));
      allocs_.push_back(d_remaining_);
      CU(cudaMalloc(&d_ticket_, sizeof(int)));
      allocs_.push_back(d_ticket_);
      dag_ = true;
    }
  }

  // ---- compact state (tent_fast.cu, CS = 1): which tent is the last one of the slab to contain each element
  compact_possible_ = fast_ && !qt_ && slab_.ntents > 0 && gen_off_.back() == 0;
  if (plan_) {
    el_final_level_ = plan_->final_level;  // of the GLOBAL slab: the last tent of a ghost element may run on another rank
  } else {
    el_final_level_.assign(nel_, -1);
    for (int t = 0; t < staged_.ntents; t++)
      for (int te = staged_.te_off[t]; te < staged_.te_off[t + 1]; te++) {
        int& l = el_final_level_[staged_.te_el[te]];
        l = std::max(l, slab_.level[t]);
      }
  }
  if (compact_possible_) d_te_final_ = const_cast<unsigned char*>(FinalFlags(-1));

  // element quadrature weights |detJ| * w (GetWeight, twavetents.cpp:793,818,843)
  {
    std::vector<double> elw((size_t)nel_ * Q);
    Quadrature(nullptr, elw.data());
    d_elw_ = Upload(elw);
  }
  CU(cudaMalloc(&d_scratch_, sizeof(double) * std::max(nel_, 1)));
  allocs_.push_back(d_scratch_);
}

Solver::~Solver() {
  cudaSetDevice(device_);
  if (stream_halo_) cudaStreamSynchronize(stream_halo_);
  if (stream_) cudaStreamSynchronize(stream_);
  for (void* p : allocs_) cudaFree(p);
  if (d_wavefront_) cudaFree(d_wavefront_);
  if (d_bnd_) cudaFree(d_bnd_);
  ReleaseMirror();
  ReleaseLoadBuffer();
  for (cudaStream_t st : {stream_in_, stream_out_, stream_proj_})
    if (st) { cudaStreamSynchronize(st); cudaStreamDestroy(st); }
  for (cudaEvent_t e : ev_chunk_) cudaEventDestroy(e);
  for (cudaEvent_t e : {ev_in_start_, ev_in_done_, ev_out_start_, ev_out_done_, ev_stage_, ev_snap_, ev_copied_, ev_used_})
    if (e) cudaEventDestroy(e);
  for (double* p : {d_stage_wf_, d_stage_cs_, d_stage_bnd_, d_snap_})
    if (p) cudaFree(p);
  if (stream_mirror_) cudaStreamDestroy(stream_mirror_);
  if (ev_level_) cudaEventDestroy(ev_level_);
  if (ev_mirror_) cudaEventDestroy(ev_mirror_);
  if (d_bnd_params_) cudaFree(d_bnd_params_);
  if (d_tmp_params_) cudaFree(d_tmp_params_);
  if (d_tmp_a_) cudaFree(d_tmp_a_);
  if (d_tmp_b_) cudaFree(d_tmp_b_);
  if (d_cstate_) cudaFree(d_cstate_);
  for (double* p : d_ainv_) if (p) cudaFree(p);
  for (void* b : ipc_opened_) cudaIpcCloseMemHandle(b);
  for (double* p : d_ginv_) if (p) cudaFree(p);
  if (h_flags_) cudaFreeHost(h_flags_);
  if (d_sendbuf_) cudaFree(d_sendbuf_);
  if (d_recvbuf_) cudaFree(d_recvbuf_);
  comm_.reset();
  if (ev0_) cudaEventDestroy(ev0_);
  if (ev1_) cudaEventDestroy(ev1_);
  if (ev_done_) cudaEventDestroy(ev_done_);
  if (ev_halo_) cudaEventDestroy(ev_halo_);
  if (ev_int_) cudaEventDestroy(ev_int_);
  if (stream_halo_) cudaStreamDestroy(stream_halo_);
  if (stream_) cudaStreamDestroy(stream_);
}

void Solver::SetWavespeedPoly(int deg, const double* coef) {
  CU(cudaSetDevice(device_));
  if (!coef) { d_cpoint_ = nullptr; return; }
  const int D = mesh_.D, Q = rule_.nip;
  const std::vector<int> ex = monomial_exponents(D, deg);
  const int nt = (int)ex.size() / D;
  std::vector<double> x((size_t)nel_ * Q * D), cp((size_t)nel_ * Q);
  Quadrature(x.data(), nullptr);
#pragma omp parallel for schedule(static)
  for (long long i = 0; i < (long long)nel_ * Q; i++) {
    double s = 0;
    for (int j = 0; j < nt; j++) {
      double m = coef[j];
      if (m == 0.0) continue;
      for (int d = 0; d < D; d++)
        for (int k = 0; k < ex[(size_t)j * D + d]; k++) m *= x[(size_t)i * D + d];
      s += m;
    }
    cp[i] = s;
  }
  d_cpoint_ = Upload(cp);
}

void Solver::Quadrature(double* x, double* weight) const {
  const int D = mesh_.D, d1 = D + 1, Q = rule_.nip;
#pragma omp parallel for schedule(static)
  for (int el = 0; el < nel_; el++) {
    const int* ev = mesh_.elverts(loc2glob_[el]);
    double J[9];
    for (int i = 0; i < D; i++)
      for (int d = 0; d < D; d++) J[d * D + i] = mesh_.point(ev[i])[d] - mesh_.point(ev[D])[d];
    double det;
    if (D == 1) det = J[0];
    else if (D == 2) det = J[0] * J[3] - J[1] * J[2];
    else det = J[0] * (J[4] * J[8] - J[5] * J[7]) - J[1] * (J[3] * J[8] - J[5] * J[6]) + J[2] * (J[3] * J[7] - J[4] * J[6]);
    det = std::fabs(det);
    for (int q = 0; q < Q; q++) {
      if (weight) weight[(size_t)el * Q + q] = det * rule_.w[q];
      if (x) {
        double lam[4], s = 0;
        for (int d = 0; d < D; d++) { lam[d] = rule_.xi[(size_t)q * D + d]; s += lam[d]; }
        lam[D] = 1.0 - s;
        for (int d = 0; d < D; d++) {
          double v = 0;
          for (int k = 0; k < d1; k++) v += lam[k] * mesh_.point(ev[k])[d];
          x[((size_t)el * Q + q) * D + d] = v;
        }
      }
    }
  }
}

// te_final flags (device) for a full slab (level_limit < 0, built once) or for the first level_limit levels
const unsigned char* Solver::FinalFlags(int level_limit) {
  if (level_limit < 0 && d_te_final_) return d_te_final_;
  if (level_limit >= 0 && level_limit == partial_levels_ && d_te_final_partial_) return d_te_final_partial_;
  std::vector<int> last(el_final_level_);
  if (level_limit >= 0) {
    // last tent among the levels < level_limit. Multi-GPU: a rank only sees its own tents, so the restriction is exact for
    // single-GPU solvers only (partial runs of partitioned solvers are refused in Propagate).
    std::fill(last.begin(), last.end(), -1);
    for (int t = 0; t < staged_.ntents; t++)
      if (slab_.level[t] < level_limit)
        for (int te = staged_.te_off[t]; te < staged_.te_off[t + 1]; te++) {
          const int el = plan_ ? plan_->glob2loc[staged_.te_el[te]] : staged_.te_el[te];
          last[el] = std::max(last[el], slab_.level[t]);
        }
  }
  std::vector<unsigned char> fin((size_t)std::max(staged_.nte, 1), 0);
  for (int t = 0; t < staged_.ntents; t++)
    for (int te = staged_.te_off[t]; te < staged_.te_off[t + 1]; te++) {
      const int el = plan_ ? plan_->glob2loc[staged_.te_el[te]] : staged_.te_el[te];
      fin[te] = last[el] == slab_.level[t];
    }
  if (level_limit < 0) return Upload(fin);
  if (!d_te_final_partial_) {
    CU(cudaMalloc(&d_te_final_partial_, fin.size()));
    allocs_.push_back(d_te_final_partial_);
  }
  CU(cudaMemcpyAsync(d_te_final_partial_, fin.data(), fin.size(), cudaMemcpyHostToDevice, stream_));
  CU(cudaStreamSynchronize(stream_));  // `fin` is a local
  partial_levels_ = level_limit;
  return d_te_final_partial_;
}

// compact state <- public state: every row through the degree-p polynomials onto the compact rule (exact for rows the
// tents have written; for sampled initial rows it is the projection under which the bottom-face integrals agree)
void Solver::ProjectCompact() {
  const int Q = rule_.nip;
  if (transfer_points_ != cs_points_) {
    const SimplexRule rs = make_compact_rule(mesh_.D, 2 * order_);
    if (rs.nip != cs_points_) throw std::logic_error("compact state: rule table and kernel disagree on the point count");
    d_transfer_ = Upload(rule_transfer_matrix(rule_, rs, order_));
    transfer_points_ = cs_points_;
  }
  // (sums over the points in ascending order, as the tents of a spatial sweep do for the rows they are the first to read:
  // both routes give the same bits)
  CU(launch_project_rows(nel_, Q, cs_points_, 0, ncomp_, ncomp_ * Q, d_transfer_, d_wavefront_, d_cstate_, stream_));
  cstate_valid_ = true;
}

void Solver::SetInitial(const double* wf, int ncomp) {
  const int D = mesh_.D;
  if (ncomp != D + 2 && ncomp != D + 1)
    throw std::invalid_argument("SetInitial: initial data must have D+2 (second order) or D+1 (first order) components");
  if (qt_ && ncomp == D + 1)  // checked before any member changes: a failed call leaves the solver as it was
    throw std::invalid_argument("SetInitial: the quasi-Trefftz solver takes second-order initial data (D+2 components)");
  CU(cudaSetDevice(device_));
  if (load_pending_) {  // a streamed input nobody consumed: its copies still write into the buffers
    CU(cudaStreamSynchronize(stream_in_));
    load_pending_ = false;
  }
  fosystem_ = (ncomp == D + 1) ? 1 : 0;  // twavetents.hpp:198-203
  ncomp_ = ncomp;
  basis_ = fosystem_ ? basis_fo_ : basis_so_;
  const size_t bytes = (size_t)nel_ * ncomp * rule_.nip * sizeof(double);
  if (wavefront_bytes_ != bytes) {
    ReleaseMirror();  // the mirror buffer and the scratch wavefronts of Energy / Error were sized for the previous row length
    if (d_tmp_a_) { CU(cudaFree(d_tmp_a_)); d_tmp_a_ = nullptr; }
    if (d_tmp_b_) { CU(cudaFree(d_tmp_b_)); d_tmp_b_ = nullptr; }
  }
  if (state_capacity_ > 0) {
    // peer-to-peer exchange: the buffers were allocated once for the largest system and are mapped by the other ranks
    if (bytes > state_capacity_) throw std::logic_error("SetInitial: state larger than the buffer shared with the other ranks");
    wavefront_bytes_ = bytes;
  } else {
    if (d_wavefront_ && wavefront_bytes_ != bytes) { CU(cudaFree(d_wavefront_)); d_wavefront_ = nullptr; }
    if (!d_wavefront_) { CU(cudaMalloc(&d_wavefront_, bytes)); wavefront_bytes_ = bytes; }
  }
  cs_points_ = compact_possible_ ? tent_fast_compact_points(D, order_, fosystem_) : 0;
  cstate_valid_ = false;
  if (cs_points_ > 0) {
    const size_t cb = (size_t)nel_ * ncomp * cs_points_ * sizeof(double);
    if (cstate_capacity_ > 0) {
      if (cb > cstate_capacity_) throw std::logic_error("SetInitial: compact state larger than the buffer shared with the other ranks");
      cstate_bytes_ = cb;
    } else {
      if (d_cstate_ && cstate_bytes_ != cb) { CU(cudaFree(d_cstate_)); d_cstate_ = nullptr; }
      if (!d_cstate_) { CU(cudaMalloc(&d_cstate_, cb)); cstate_bytes_ = cb; }
    }
  }
  if (wf) {
    CU(cudaMemcpyAsync(d_wavefront_, wf, bytes, cudaMemcpyHostToDevice, stream_));
    if (cs_points_ > 0 && compact_enabled_) ProjectCompact();
    CU(cudaStreamSynchronize(stream_));
  }
  have_initial_ = true;

  TentKernelArgs& a = args_;
  a.rowptr = d_rowptr_[fosystem_];
  a.col = d_col_[fosystem_];
  a.val = d_val_[fosystem_];
  a.N = basis_.nbasis;
  a.gnnz = 0;
  for (int i = 0; i + 1 < (int)basis_.rowptr.size(); i++) a.gnnz = std::max(a.gnnz, basis_.rowptr[i + 1] - basis_.rowptr[i]);
  a.Npad = (a.N + 3) / 4 * 4;
  a.ncomp = ncomp_;
  a.fosystem = fosystem_;
  a.wavefront = d_wavefront_;
  fargs_.cptr = d_cptr_[fosystem_];
  fargs_.crow = d_crow_[fosystem_];
  fargs_.cval = d_cval_[fosystem_];
  if (qt_) {
    a.rowptr = nullptr; a.col = nullptr; a.val = nullptr;
    const size_t budget = 150 * 1024;
    int qc = std::min(std::max(rule_.nip, qargs_.Qv), 16);
    for (; qc > 1; qc--) {
      a.Qc = qc;
      if (tent_qt_smem_bytes(D, a, qargs_) <= budget) break;
    }
    a.Qc = qc;
    smem_ = tent_qt_smem_bytes(D, a, qargs_);
    if (smem_ > 227 * 1024 || a.N > 64)
      throw std::invalid_argument("TWave: quasi-Trefftz local system too large for shared memory");
    return;
  }
  // chunk of quadrature points staged in shared memory at a time
  const size_t budget = 100 * 1024;
  int qc = std::min(rule_.nip, 32);
  for (; qc > 1; qc--) {
    a.Qc = qc;
    if (tent_generic_smem_bytes(D, a) <= budget) break;
  }
  a.Qc = qc;
  smem_ = tent_generic_smem_bytes(D, a);
  if (smem_ > 227 * 1024)
    throw std::invalid_argument("TWave: local system too large for shared memory (order / number of wavespeed regions per tent)");
  const int tiles = (a.Npad / 4) * (a.Npad / 4);
  threads_ = tiles > 128 ? 256 : (tiles > 32 ? 128 : 64);
}

namespace {
// stream memory operations of the driver API (no SM involved: they run on the front end of a stream)
struct StreamMemOps {
  using WriteFn = int (*)(cudaStream_t, unsigned long long, unsigned int, unsigned int);
  using WaitFn = int (*)(cudaStream_t, unsigned long long, unsigned int, unsigned int);
  WriteFn write = nullptr;
  WaitFn wait = nullptr;
  StreamMemOps() {
    void* lib = dlopen("libcuda.so.1", RTLD_NOW | RTLD_GLOBAL);
    if (!lib) return;
    write = reinterpret_cast<WriteFn>(dlsym(lib, "cuStreamWriteValue32_v2"));
    wait = reinterpret_cast<WaitFn>(dlsym(lib, "cuStreamWaitValue32_v2"));
    if (!write) write = reinterpret_cast<WriteFn>(dlsym(lib, "cuStreamWriteValue32"));
    if (!wait) wait = reinterpret_cast<WaitFn>(dlsym(lib, "cuStreamWaitValue32"));
  }
  bool ok() const { return write && wait; }
};
StreamMemOps& memops() {
  static StreamMemOps m;
  return m;
}
}  // namespace

void Solver::ReleaseLoadBuffer() {
  if (load_host_ && load_registered_) {
    cudaSetDevice(device_);
    if (stream_in_) cudaStreamSynchronize(stream_in_);
    cudaHostUnregister(const_cast<double*>(load_host_));
  }
  load_host_ = nullptr;
  load_registered_ = false;
}

void Solver::FinishLoad() {
  if (!load_pending_) return;
  CU(cudaStreamWaitEvent(stream_, ev_in_done_, 0));
  load_pending_ = false;
}

// Sweep tables (once per solver): chunks of consecutive elements, the ticket order, first-touch flags, per-chunk counts
void Solver::BuildSweep() {
  sweep_built_ = true;
  sweep_possible_ = false;
  if (!dag_ || !compact_possible_ || plan_ || !memops().ok()) return;
  const int nt = staged_.ntents, Q = rule_.nip;
  const size_t rowbytes = (size_t)(mesh_.D + 2) * Q * sizeof(double);
  sweep_chunk_elems_ = (int)std::max<size_t>(256, (16u << 20) / rowbytes);  // ~16 MB per chunk
  sweep_nchunks_ = (nel_ + sweep_chunk_elems_ - 1) / sweep_chunk_elems_;
  // Ticket order = estimated earliest start of every tent: it can start when the chunk it reads has landed (chunk c after
  // (c + 1) chunk copies at the measured PCIe rate) and when the last earlier tent of each of its elements is done (a tent
  // takes about 1.6 us per element when it applies a stored inverse). Predecessors have strictly smaller estimates, so
  // they draw earlier tickets (no deadlock), and CTAs pick up tents along the diagonal front of the sweep -- chunk after
  // chunk in space, level after level in time -- instead of finishing one chunk's 42 levels before touching the next.
  const double chunk_s = (double)sweep_chunk_elems_ * rowbytes / 53e9, el_s = 1.6e-6;
  std::vector<int> need(nt, 0);
  std::vector<double> est(nt, 0.0), done_of_last(nel_, 0.0);
  std::vector<unsigned char> first(std::max(staged_.nte, 1), 0), touched(nel_, 0);
  for (int t = 0; t < nt; t++) {  // pitching order: predecessors first
    int nd = 0;
    double st = 0.0;
    for (int te = staged_.te_off[t]; te < staged_.te_off[t + 1]; te++) {
      const int el = staged_.te_el[te];
      nd = std::max(nd, el / sweep_chunk_elems_);
      st = std::max(st, done_of_last[el]);
      if (!touched[el]) { touched[el] = 1; first[te] = 1; }
    }
    need[t] = nd;
    est[t] = std::max(st, (nd + 1) * chunk_s);
    const double fin = est[t] + el_s * (staged_.te_off[t + 1] - staged_.te_off[t]);
    for (int te = staged_.te_off[t]; te < staged_.te_off[t + 1]; te++) done_of_last[staged_.te_el[te]] = fin;
  }
  std::vector<int> order(nt);
  for (int t = 0; t < nt; t++) order[t] = t;
  std::stable_sort(order.begin(), order.end(), [&](int x, int y) { return est[x] < est[y]; });
  chunk_left_init_.assign(sweep_nchunks_, 0);
  std::vector<int> untouched;
  for (int el = 0; el < nel_; el++) {
    if (touched[el]) chunk_left_init_[el / sweep_chunk_elems_]++;
    else untouched.push_back(el);
  }
  sweep_untouched_ = (int)untouched.size();
  d_sweep_tents_ = Upload(order);
  d_tent_chunk_ = Upload(need);
  d_te_first_ = Upload(first);
  d_chunk_left_init_ = Upload(chunk_left_init_);
  d_untouched_ = Upload(untouched);
  CU(cudaMalloc(&d_chunk_loaded_, sizeof(int) * sweep_nchunks_));
  allocs_.push_back(d_chunk_loaded_);
  CU(cudaMalloc(&d_chunk_left_, sizeof(int) * sweep_nchunks_));
  allocs_.push_back(d_chunk_left_);
  EnsurePipeStreams();
  sweep_possible_ = true;
}

void Solver::SetInitialStreamed(const double* wf, int ncomp) {
  if (!wf) throw std::invalid_argument("SetInitialStreamed: null buffer");
  CU(cudaSetDevice(device_));
  if (!sweep_built_) BuildSweep();
  const bool can = sweep_possible_ && compact_enabled_ && !force_generic_ &&
                   tent_fast_stages_rows(mesh_.D, order_, ncomp == mesh_.D + 1 ? 1 : 0);
  if (!can) {  // no sweep for this solver: the plain, blocking path
    SetInitial(wf, ncomp);
    return;
  }
  SetInitial(nullptr, ncomp);  // validates, allocates, configures the kernels (waits for an input nobody consumed)
  const size_t rowlen = (size_t)ncomp_ * rule_.nip;
  const size_t bytes = (size_t)nel_ * rowlen * sizeof(double);
  if (wf != load_host_) {
    ReleaseLoadBuffer();
    cudaPointerAttributes attr{};
    if (cudaPointerGetAttributes(&attr, wf) != cudaSuccess) {
      cudaGetLastError();
      attr.type = cudaMemoryTypeUnregistered;
    }
    if (attr.type == cudaMemoryTypeUnregistered) {
      CU(cudaHostRegister(const_cast<double*>(wf), bytes, cudaHostRegisterDefault));  // asynchronous copies need page-locked memory
      load_registered_ = true;
    } else if (attr.type != cudaMemoryTypeHost) {
      throw std::invalid_argument("SetInitialStreamed: the buffer must be host memory");
    }
    load_host_ = wf;
  }
  if (transfer_points_ != cs_points_) {
    const SimplexRule rs = make_compact_rule(mesh_.D, 2 * order_);
    d_transfer_ = Upload(rule_transfer_matrix(rule_, rs, order_));
    transfer_points_ = cs_points_;
  }
  // flags down in the order of the solver's stream (behind whatever still uses the old state), then the chunk copies
  CU(cudaMemsetAsync(d_chunk_loaded_, 0, sizeof(int) * sweep_nchunks_, stream_));
  CU(cudaEventRecord(ev_in_start_, stream_));
  CU(cudaStreamWaitEvent(stream_in_, ev_in_start_, 0));
  for (int c = 0; c < sweep_nchunks_; c++) {
    const size_t e0 = (size_t)c * sweep_chunk_elems_, ne = std::min<size_t>(sweep_chunk_elems_, nel_ - e0);
    CU(cudaMemcpyAsync(d_wavefront_ + e0 * rowlen, wf + e0 * rowlen, ne * rowlen * sizeof(double), cudaMemcpyHostToDevice, stream_in_));
    if (memops().write(stream_in_, (unsigned long long)(d_chunk_loaded_ + c), 1u, 0u) != 0)
      throw CudaFailure("CUDA: cuStreamWriteValue32 failed");
  }
  CU(cudaEventRecord(ev_in_done_, stream_in_));
  cstate_valid_ = false;  // the tents of the sweep build the compact rows of everything they touch
  load_pending_ = true;
}

void Solver::EnsurePipeStreams() {
  if (!stream_in_) {
    int lo = 0, hi = 0;
    CU(cudaDeviceGetStreamPriorityRange(&lo, &hi));
    CU(cudaStreamCreateWithPriority(&stream_in_, cudaStreamNonBlocking, hi));
    CU(cudaStreamCreateWithPriority(&stream_out_, cudaStreamNonBlocking, hi));
    CU(cudaStreamCreateWithPriority(&stream_proj_, cudaStreamNonBlocking, hi));
    for (cudaEvent_t* e : {&ev_in_start_, &ev_in_done_, &ev_out_start_, &ev_out_done_})
      CU(cudaEventCreateWithFlags(e, cudaEventDisableTiming));
  }
  if (!ev_stage_)
    for (cudaEvent_t* e : {&ev_stage_, &ev_snap_, &ev_copied_, &ev_used_}) CU(cudaEventCreateWithFlags(e, cudaEventDisableTiming));
}

// The next step's input, fetched while the current step computes: H2D copies back to back on the input stream (copy
// engine); the transfer of each chunk to the compact state runs on a second stream behind that chunk's event, so the copy
// engine never waits for a kernel (r04m had both in one stream: every transfer kernel stalled the next chunk's copy).
void Solver::PrefetchInput(const double* wf, int ncomp, const double* bnd_values, long long nbnd) {
  const int D = mesh_.D, Q = rule_.nip;
  if (!wf) throw std::invalid_argument("PrefetchInput: null buffer");
  if (ncomp != D + 2 && ncomp != D + 1) throw std::invalid_argument("PrefetchInput: initial data must have D+2 or D+1 components");
  if (qt_ && ncomp == D + 1) throw std::invalid_argument("PrefetchInput: the quasi-Trefftz solver takes second-order initial data");
  if (bnd_values && nbnd != staged_.nbnd_values) throw std::invalid_argument("PrefetchInput: wrong number of boundary values");
  CU(cudaSetDevice(device_));
  EnsurePipeStreams();
  const size_t bytes = (size_t)nel_ * ncomp * Q * sizeof(double);
  if (stage_wf_bytes_ < bytes) {
    CU(cudaStreamSynchronize(stream_in_));
    CU(cudaStreamSynchronize(stream_proj_));
    if (d_stage_wf_) CU(cudaFree(d_stage_wf_));
    CU(cudaMalloc(&d_stage_wf_, bytes));
    stage_wf_bytes_ = bytes;
  }
  const int fo = ncomp == D + 1 ? 1 : 0;
  const int cs = (compact_possible_ && compact_enabled_) ? tent_fast_compact_points(D, order_, fo) : 0;
  const size_t cbytes = (size_t)nel_ * ncomp * cs * sizeof(double);
  if (cs > 0 && stage_cs_bytes_ < cbytes) {
    CU(cudaStreamSynchronize(stream_in_));
    CU(cudaStreamSynchronize(stream_proj_));
    if (d_stage_cs_) CU(cudaFree(d_stage_cs_));
    CU(cudaMalloc(&d_stage_cs_, cbytes));
    stage_cs_bytes_ = cbytes;
  }
  if (bnd_values && nbnd > 0 && !d_stage_bnd_) CU(cudaMalloc(&d_stage_bnd_, sizeof(double) * nbnd));
  if (cs > 0 && transfer_points_ != cs) {
    const SimplexRule rs = make_compact_rule(D, 2 * order_);
    d_transfer_ = Upload(rule_transfer_matrix(rule_, rs, order_));
    transfer_points_ = cs;
  }
  // the staging buffers are free once the previous prefetch has been taken over
  CU(cudaStreamWaitEvent(stream_in_, ev_used_, 0));
  if (bnd_values && nbnd > 0)
    CU(cudaMemcpyAsync(d_stage_bnd_, bnd_values, sizeof(double) * nbnd, cudaMemcpyHostToDevice, stream_in_));
  // in chunks, so that the transfer of a chunk runs behind the copy of the next one
  const size_t rowlen = (size_t)ncomp * Q;
  const int chunk = (int)std::max<size_t>(256, (16u << 20) / (rowlen * sizeof(double)));
  const int nchunks = (nel_ + chunk - 1) / chunk;
  while (cs > 0 && (int)ev_chunk_.size() < nchunks) {
    cudaEvent_t e;
    CU(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    ev_chunk_.push_back(e);
  }
  for (int c = 0, e0 = 0; e0 < nel_; e0 += chunk, c++) {
    const int ne = std::min(chunk, nel_ - e0);
    CU(cudaMemcpyAsync(d_stage_wf_ + (size_t)e0 * rowlen, wf + (size_t)e0 * rowlen, (size_t)ne * rowlen * sizeof(double),
                       cudaMemcpyHostToDevice, stream_in_));
    if (cs > 0) {
      CU(cudaEventRecord(ev_chunk_[c], stream_in_));
      CU(cudaStreamWaitEvent(stream_proj_, ev_chunk_[c], 0));
      CU(launch_project_rows(ne, Q, cs, 0, ncomp, ncomp * Q, d_transfer_, d_stage_wf_ + (size_t)e0 * rowlen,
                             d_stage_cs_ + (size_t)e0 * ncomp * cs, stream_proj_));
    }
  }
  if (cs > 0) {  // staged = every copy has landed (the boundary samples included) and every chunk is transferred
    CU(cudaEventRecord(ev_stage_, stream_in_));
    CU(cudaStreamWaitEvent(stream_proj_, ev_stage_, 0));
    CU(cudaEventRecord(ev_stage_, stream_proj_));
  } else {
    CU(cudaEventRecord(ev_stage_, stream_in_));
  }
  stage_ncomp_ = ncomp;
  stage_has_bnd_ = bnd_values && nbnd > 0;
  stage_valid_ = true;
}

// SetInitial (+ SetBoundaryValues) from the prefetched input: device-to-device copies on the solver's stream
void Solver::UsePrefetched() {
  if (!stage_valid_) throw std::logic_error("UsePrefetched: nothing has been prefetched");
  SetInitial(nullptr, stage_ncomp_);  // validates, allocates, configures
  const size_t bytes = (size_t)nel_ * ncomp_ * rule_.nip * sizeof(double);
  CU(cudaStreamWaitEvent(stream_, ev_stage_, 0));
  CU(cudaMemcpyAsync(d_wavefront_, d_stage_wf_, bytes, cudaMemcpyDeviceToDevice, stream_));
  if (cs_points_ > 0 && compact_enabled_) {
    CU(cudaMemcpyAsync(d_cstate_, d_stage_cs_, (size_t)nel_ * ncomp_ * cs_points_ * sizeof(double), cudaMemcpyDeviceToDevice, stream_));
    cstate_valid_ = true;
  }
  if (stage_has_bnd_) {
    bnd_family_ = -1;
    if (!d_bnd_) CU(cudaMalloc(&d_bnd_, sizeof(double) * staged_.nbnd_values));
    CU(cudaMemcpyAsync(d_bnd_, d_stage_bnd_, sizeof(double) * staged_.nbnd_values, cudaMemcpyDeviceToDevice, stream_));
    args_.bnd_data = d_bnd_;
  }
  CU(cudaEventRecord(ev_used_, stream_));
  stage_valid_ = false;
}

void Solver::GetWavefrontAsync(double* out) {
  if (!have_initial_) throw std::logic_error("GetWavefrontAsync: SetInitial has not been called");
  if (!out) throw std::invalid_argument("GetWavefrontAsync: null buffer");
  CU(cudaSetDevice(device_));
  FinishLoad();
  EnsurePipeStreams();
  const size_t bytes = (size_t)nel_ * ncomp_ * rule_.nip * sizeof(double);
  if (snap_bytes_ < bytes) {
    CU(cudaStreamSynchronize(stream_out_));
    if (d_snap_) CU(cudaFree(d_snap_));
    CU(cudaMalloc(&d_snap_, bytes));
    snap_bytes_ = bytes;
  }
  // snapshot on the device (the state may be overwritten by the next step at once), then the copy engine takes it out
  if (out_pending_) CU(cudaStreamWaitEvent(stream_, ev_copied_, 0));  // the previous snapshot has left
  CU(cudaMemcpyAsync(d_snap_, d_wavefront_, bytes, cudaMemcpyDeviceToDevice, stream_));
  CU(cudaEventRecord(ev_snap_, stream_));
  CU(cudaStreamWaitEvent(stream_out_, ev_snap_, 0));
  CU(cudaMemcpyAsync(out, d_snap_, bytes, cudaMemcpyDeviceToHost, stream_out_));
  CU(cudaEventRecord(ev_copied_, stream_out_));
  out_pending_ = true;
}

void Solver::WaitOutput() {
  if (!out_pending_) return;
  CU(cudaSetDevice(device_));
  CU(cudaEventSynchronize(ev_copied_));
  out_pending_ = false;
}

namespace {
void check_family(int D, int family, const double* params, int nparams) {
  if (!params) throw std::invalid_argument("analytic data: params is null");
  if (family == kStandingWave) {
    if (nparams != 1) throw std::invalid_argument("analytic data: the standing wave takes 1 parameter (c)");
  } else if (family == kPlaneGauss) {
    if (nparams != 3 + D) throw std::invalid_argument("analytic data: the plane pulse takes 3 + D parameters (c, a, s0, k)");
  } else if (family == kPolynomial) {
    if (nparams < 1 || params[0] < 0 || nparams != 1 + (int)params[0] * (D + 2))
      throw std::invalid_argument("analytic data: the polynomial takes 1 + nterms * (D + 2) parameters");
    for (int m = 0; m < (int)params[0]; m++)
      for (int d = 0; d <= D; d++) {
        const double e = params[1 + m * (D + 2) + 1 + d];
        if (e < 0 || e > 64 || e != std::floor(e)) throw std::invalid_argument("analytic data: exponents must be integers in [0, 64]");
      }
  } else {
    throw std::invalid_argument("analytic data: unknown family");
  }
}
}  // namespace

double* Solver::UploadParams(const double* params, int nparams, double** slot) {
  if (*slot) { CU(cudaFree(*slot)); *slot = nullptr; }
  CU(cudaMalloc(slot, sizeof(double) * nparams));
  CU(cudaMemcpyAsync(*slot, params, sizeof(double) * nparams, cudaMemcpyHostToDevice, stream_));
  CU(cudaStreamSynchronize(stream_));
  return *slot;
}

void Solver::SetInitialAnalytic(int family, const double* params, int nparams, int first_order) {
  const int D = mesh_.D;
  check_family(D, family, params, nparams);
  SetInitial(nullptr, first_order ? D + 1 : D + 2);  // allocates the state and configures the kernels
  double* d_p = UploadParams(params, nparams, &d_tmp_params_);
  CU(launch_sample_wavefront(D, family, d_p, nel_, rule_.nip, first_order ? 1 : 0, 0.0, args_.pts, args_.el2v, args_.lam,
                             d_wavefront_, stream_));
  if (cs_points_ > 0 && compact_enabled_) ProjectCompact();
  CU(cudaStreamSynchronize(stream_));
}

void Solver::MakeWavefrontAnalytic(int family, const double* params, int nparams, double time, int first_order,
                                   double* out) {
  const int D = mesh_.D;
  check_family(D, family, params, nparams);
  if (!out) throw std::invalid_argument("MakeWavefrontAnalytic: out is null");
  CU(cudaSetDevice(device_));
  const int ncomp = first_order ? D + 1 : D + 2;
  const size_t bytes = (size_t)nel_ * ncomp * rule_.nip * sizeof(double);
  double* d_out = nullptr;
  CU(cudaMalloc(&d_out, bytes));
  try {
    double* d_p = UploadParams(params, nparams, &d_tmp_params_);
    CU(launch_sample_wavefront(D, family, d_p, nel_, rule_.nip, first_order ? 1 : 0, time, args_.pts, args_.el2v,
                               args_.lam, d_out, stream_));
    CU(cudaMemcpyAsync(out, d_out, bytes, cudaMemcpyDeviceToHost, stream_));
    CU(cudaStreamSynchronize(stream_));
  } catch (...) {
    cudaFree(d_out);
    throw;
  }
  CU(cudaFree(d_out));
}

void Solver::SetBoundaryAnalytic(int family, const double* params, int nparams) {
  check_family(mesh_.D, family, params, nparams);
  CU(cudaSetDevice(device_));
  UploadParams(params, nparams, &d_bnd_params_);
  bnd_family_ = family;
  if (staged_.nbnd_values > 0 && !d_bnd_) CU(cudaMalloc(&d_bnd_, sizeof(double) * staged_.nbnd_values));
  args_.bnd_data = staged_.nbnd_values > 0 ? d_bnd_ : nullptr;
}

void Solver::SetBoundaryValues(const double* values, long long n) {
  CU(cudaSetDevice(device_));
  bnd_family_ = -1;  // host-sampled values replace an analytic datum
  if (values == nullptr) { args_.bnd_data = nullptr; return; }
  if (n != staged_.nbnd_values) throw std::invalid_argument("SetBoundaryValues: wrong number of values");
  if (n == 0) { args_.bnd_data = nullptr; return; }
  if (!d_bnd_) CU(cudaMalloc(&d_bnd_, sizeof(double) * n));
  CU(cudaMemcpyAsync(d_bnd_, values, sizeof(double) * n, cudaMemcpyHostToDevice, stream_));
  CU(cudaStreamSynchronize(stream_));
  args_.bnd_data = d_bnd_;
}

void Solver::Propagate(int level_limit) {
  if (!have_initial_) throw std::logic_error("Propagate: SetInitial has not been called");
  CU(cudaSetDevice(device_));
  const int nlev_all = (int)slab_.level_off.size() - 1;
  // level_limit >= 0 (diagnostics, ngst_twave_propagate_levels): only the tents of the first level_limit levels, one
  // launch per level, no advance of timeshift -- the state the reference's serial loop holds once these tents are done
  const bool partial = level_limit >= 0 && level_limit < nlev_all;
  const int nlev = partial ? level_limit : nlev_all;
  if (partial && plan_) throw std::logic_error("Propagate: partial slabs (propagate_levels) are for single-GPU solvers");
  // streamed input (SetInitialStreamed): the spatial sweep below consumes the chunks as they land; anything else waits
  // for the whole input
  const bool sweep = load_pending_ && sweep_possible_ && compact_enabled_ && cs_points_ > 0 && !qt_ && !force_generic_ && !partial &&
                     !comm_ && dag_mode_ != 0;
  if (load_pending_ && !sweep) FinishLoad();
  // compact state: the fast kernel reads / writes d_cstate_ and writes the public rows of final elements only
  const bool compact = UseCompact() && !qt_;
  FastArgs fa = fargs_;
  fa.cstate = nullptr;
  fa.te_final = nullptr;
  if (compact) {
    if (!cstate_valid_ && !load_pending_) ProjectCompact();
    fa.cstate = d_cstate_;
    fa.te_final = FinalFlags(partial ? level_limit : -1);
  }
  // factor cache: first full Propagate of this system stores the inverses (MODE 1), later ones apply them (MODE 2)
  fa.ainv = nullptr;
  fa.mode = 0;
  fa.p2p = nullptr;
  fa.tent_chunk = nullptr;
  fa.chunk_loaded = nullptr;
  fa.te_first = nullptr;
  fa.transfer = nullptr;
  fa.chunk_left = nullptr;
  fa.chunk_elems = 1;
  const int fo = fosystem_;
  if (compact && factor_cache_enabled_ && !cache_failed_ && (ainv_valid_[fo] || !partial)) {
    if (!d_ainv_[fo]) {
      const size_t bytes = sizeof(double) * tent_fast_inverse_doubles(mesh_.D, order_, fo) * (size_t)std::max(staged_.ntents, 1);
      if (cudaMalloc(&d_ainv_[fo], bytes) != cudaSuccess) {
        cudaGetLastError();  // not enough memory for the cache: keep assembling in every Propagate
        d_ainv_[fo] = nullptr;
        cache_failed_ = true;
      } else {
        ainv_bytes_[fo] = bytes;
      }
    }
    if (d_ainv_[fo]) {
      fa.ainv = d_ainv_[fo];
      fa.mode = ainv_valid_[fo] ? (apply_four_ctas_ ? 3 : 2) : 1;
    }
  }
  // factor cache of the generic kernel (small systems): inverses formed by the first full Propagate, applied afterwards
  args_.ginv = nullptr;
  args_.gmode = 0;
  args_.gwarp = 0;
  size_t gen_smem = smem_;
  if (GenericCachePossible() && (ginv_valid_[fosystem_] || !partial)) {
    const int fo = fosystem_;
    const size_t nmax = (size_t)args_.max_ndom * args_.Npad;
    if (!d_ginv_[fo]) {
      const size_t bytes = sizeof(double) * nmax * nmax * (size_t)std::max(staged_.ntents, 1);
      if (cudaMalloc(&d_ginv_[fo], bytes) != cudaSuccess) {
        cudaGetLastError();
        d_ginv_[fo] = nullptr;
        ginv_failed_ = true;
      } else {
        ginv_bytes_[fo] = bytes;
      }
    }
    if (d_ginv_[fo]) {
      args_.ginv = d_ginv_[fo];
      args_.gmode = ginv_valid_[fo] ? 2 : 1;
      args_.gwarp = args_.gmode == 2 ? GenericWarpApply() : 0;
      if (args_.gmode == 1) {
        TentKernelArgs probe = args_;
        gen_smem = tent_generic_smem_bytes(mesh_.D, probe);
        if (gen_smem > 200 * 1024) { args_.ginv = nullptr; args_.gmode = 0; gen_smem = smem_; }
      }
    }
  }
  int launches = 0;
  CU(cudaEventRecord(ev0_, stream_));
  if (bnd_family_ >= 0 && staged_.nbnd_values > 0) {
    // analytic boundary datum: sampled for this slab at t_local + timeshift (twavetents.cpp:373-374) ahead of the tents
    CU(launch_sample_boundary(mesh_.D, bnd_family_, d_bnd_params_, staged_.ntf, rule_.nip, timeshift_, args_.tf_kind,
                              args_.tf_verts, args_.tf_data, args_.lam, kNeumann, d_bnd_, stream_));
    launches++;
  }
  if (mirror_dev_ && !sweep) MirrorRows(-1, -1, launches);  // rows no tent touches are final from the start
  for (int l = 0; l < nlev && qt_; l++) {
    // quasi-Trefftz: fast monomial-space DMMA kernel where a variant exists (second-order form, causal top faces),
    // generic kernel otherwise
    const bool usefast = qtfast_ && !force_generic_ && !fosystem_;
    TentKernelArgs a = args_;
    if (usefast) {
      a.tents = d_level_fast_ + fast_off_[l];
      a.ntents = fast_off_[l + 1] - fast_off_[l];
      if (a.ntents > 0) {
        CU(launch_tent_qt_fast(order_, a, qargs_, qfargs_, stream_));
        launches++;
      }
      a.tents = d_level_tents_ + gen_off_[l];
      a.ntents = gen_off_[l + 1] - gen_off_[l];
    } else {
      a.tents = d_level_all_ + slab_.level_off[l];
      a.ntents = slab_.level_off[l + 1] - slab_.level_off[l];
    }
    if (a.ntents > 0) {
      CU(launch_tent_qt(mesh_.D, a, qargs_, smem_, stream_));
      launches++;
    }
    if (comm_) launches += ExchangeLevel(l, stream_);
    if (mirror_dev_) MirrorRows(l, l, launches);
  }
  for (int l = 0; l < nlev && force_generic_ && !qt_; l++) {
    TentKernelArgs a = args_;
    a.tents = d_level_all_ + slab_.level_off[l];
    a.ntents = slab_.level_off[l + 1] - slab_.level_off[l];
    if (a.ntents > 0) {
      CU(launch_tent_generic(mesh_.D, a, threads_, gen_smem, stream_));
      launches++;
    }
    if (comm_) launches += ExchangeLevel(l, stream_);
    if (mirror_dev_) MirrorRows(l, l, launches);
  }
  // Measured on B200 (profiles/r01o): with ~2700 tents per level (config 4) the single launch is 3 % slower than one launch
  // per level (waiting CTAs hold SM slots and poll), with small levels it removes the launch gaps and tails.
  const bool small_levels = slab_.ntents < 2 * 444 * nlev;
  // (r04d, B200: with stored inverses a tent is 6x cheaper and the single launch wins at config 4 as well, 11.74 vs 12.17 ms;
  // not with an output mirror, whose copies overlap the per-level launches)
  const bool cheap_tents = fa.mode >= 2 && !mirror_dev_;
  const bool dag = dag_ && (dag_mode_ == 2 || (dag_mode_ == 1 && (small_levels || cheap_tents))) && !force_generic_ && !comm_ && !partial &&
                   !sweep;
  if (sweep) {
    // ---- spatial sweep: ONE dependency-driven launch whose tents are ticketed in the order in which their dependency cones
    // are complete. A tent waits for its predecessors (hazard counters) and for the chunk of public rows it is the first
    // to read (flag written by the copy engine's stream behind the chunk's copy); the first tent to touch an element builds
    // its compact row on the fly. With an output mirror the last tent of an element counts its chunk down and the copy
    // engine takes every finished chunk out (stream wait on the counter) while the sweep goes on.
    const size_t rowlen = (size_t)ncomp_ * rule_.nip;
    CU(cudaMemcpyAsync(d_remaining_, d_indeg_, sizeof(int) * slab_.ntents, cudaMemcpyDeviceToDevice, stream_));
    CU(cudaMemsetAsync(d_ticket_, 0, sizeof(int), stream_));
    const bool out = mirror_host_ != nullptr;
    if (out) {
      CU(cudaMemcpyAsync(d_chunk_left_, d_chunk_left_init_, sizeof(int) * sweep_nchunks_, cudaMemcpyDeviceToDevice, stream_));
      CU(cudaEventRecord(ev_out_start_, stream_));
      CU(cudaStreamWaitEvent(stream_out_, ev_out_start_, 0));
    }
    TentKernelArgs a = args_;
    a.tents = d_sweep_tents_;
    a.ntents = slab_.ntents;
    a.ticket = d_ticket_;
    a.remaining = d_remaining_;
    a.succ_off = d_succ_off_;
    a.succ = d_succ_;
    fa.tent_chunk = d_tent_chunk_;
    fa.chunk_loaded = d_chunk_loaded_;
    fa.te_first = d_te_first_;
    fa.transfer = d_transfer_;
    fa.chunk_left = out ? d_chunk_left_ : nullptr;
    fa.chunk_elems = sweep_chunk_elems_;
    CU(launch_tent_fast(mesh_.D, order_, fosystem_, a, fa, stream_));
    launches++;
    if (out) {
      for (int c = 0; c < sweep_nchunks_; c++) {
        const size_t e0 = (size_t)c * sweep_chunk_elems_, ne = std::min<size_t>(sweep_chunk_elems_, nel_ - e0);
        if (chunk_left_init_[c] == 0) CU(cudaStreamWaitEvent(stream_out_, ev_in_done_, 0));  // nobody touches it: plain copy of the input
        else if (memops().wait(stream_out_, (unsigned long long)(d_chunk_left_ + c), 0u, 1u /* EQ */) != 0)
          throw CudaFailure("CUDA: cuStreamWaitValue32 failed");
        CU(cudaMemcpyAsync(mirror_host_ + e0 * rowlen, d_wavefront_ + e0 * rowlen, ne * rowlen * sizeof(double), cudaMemcpyDeviceToHost, stream_out_));
      }
      CU(cudaEventRecord(ev_out_done_, stream_out_));
      CU(cudaStreamWaitEvent(stream_, ev_out_done_, 0));
    }
    CU(cudaStreamWaitEvent(stream_, ev_in_done_, 0));  // (chunks no tent reads)
    load_pending_ = false;
    if (sweep_untouched_ > 0)  // elements no tent touches: their compact rows by the ordinary transfer
      CU(launch_project_rows(sweep_untouched_, rule_.nip, cs_points_, 0, ncomp_, ncomp_ * rule_.nip, d_transfer_, d_wavefront_, d_cstate_,
                             stream_, d_untouched_));
    cstate_valid_ = true;
  }
  if (dag) {
    CU(cudaMemcpyAsync(d_remaining_, d_indeg_, sizeof(int) * slab_.ntents, cudaMemcpyDeviceToDevice, stream_));
    CU(cudaMemsetAsync(d_ticket_, 0, sizeof(int), stream_));
    TentKernelArgs a = args_;
    a.tents = d_level_fast_;  // level after level, most expensive tents first inside a level
    a.ntents = fast_off_.back();
    a.ticket = d_ticket_;
    a.remaining = d_remaining_;
    a.succ_off = d_succ_off_;
    a.succ = d_succ_;
    CU(launch_tent_fast(mesh_.D, order_, fosystem_, a, fa, stream_));
    launches++;
    if (mirror_dev_) MirrorRows(0, nlev - 1, launches);  // single launch per slab: nothing to overlap with
  }
  // multi-GPU, peer-to-peer: ONE dependency-driven launch per slab on every rank. Tents wait on hazard counters that tents of
  // other ranks decrement with system-scope atomics after pushing their rows into this rank's buffers (tent_fast.cu).
  const bool p2p = comm_ && p2p_ready_ && p2p_enabled_ && compact && overlap_ && !force_generic_ && !qt_ && !partial;
  if (p2p) {
    TentKernelArgs a = args_;
    a.tents = p2p_halo_first_ ? d_level_fast_ : d_p2p_tents_;
    a.ntents = fast_off_.back();
    a.ticket = d_ticket_;
    a.remaining = d_remaining_;
    a.succ_off = d_p2p_lsucc_off_;
    a.succ = d_p2p_lsucc_;
    fa.p2p = d_p2p_;
    // nobody's tents push into a buffer that its owner is still filling (SetInitial, sampling, the compact transfer)
    comm_->Barrier(d_barrier_, stream_);
    CU(launch_tent_fast(mesh_.D, order_, fosystem_, a, fa, stream_));
    launches++;
    // arm the counters for the next slab, then wait for everybody: when the barrier completes on this stream every rank has
    // finished its tents (all pushes into this rank's buffers have landed) and nobody starts the next slab before all
    // counters are armed
    CU(cudaMemcpyAsync(d_remaining_, d_p2p_indeg_, sizeof(int) * slab_.ntents, cudaMemcpyDeviceToDevice, stream_));
    CU(cudaMemsetAsync(d_ticket_, 0, sizeof(int), stream_));
    comm_->Barrier(d_barrier_, stream_);
    if (mirror_dev_) MirrorRows(0, nlev - 1, launches);
  }
  const bool overlap = comm_ && !force_generic_ && !qt_ && overlap_ && !p2p;
  if (overlap && !stream_halo_) {
    int lo = 0, hi = 0;
    CU(cudaDeviceGetStreamPriorityRange(&lo, &hi));
    CU(cudaStreamCreateWithPriority(&stream_halo_, cudaStreamNonBlocking, hi));
    CU(cudaEventCreateWithFlags(&ev_halo_, cudaEventDisableTiming));
    CU(cudaEventCreateWithFlags(&ev_int_, cudaEventDisableTiming));
  }
  if (overlap) CU(cudaEventRecord(ev_int_, stream_));
  for (int l = 0; l < nlev && !force_generic_ && !qt_ && !dag && !p2p && !sweep; l++) {
    // level l: [halo tents | interior tents] of the fast and of the generic list. Single GPU: one launch per list.
    // Multi-GPU: halo tents + exchange on the high-priority stream, interior tents concurrently on the main stream;
    // level l+1 starts when both are done (tents of a level touch disjoint rows, received rows belong to no local tent
    // of the level).
    const int fh = overlap ? fast_nhalo_[l] : 0, gh = overlap ? gen_nhalo_[l] : 0;
    if (overlap) {
      CU(cudaStreamWaitEvent(stream_halo_, ev_int_, 0));
      if (l > 0) {
        CU(cudaStreamWaitEvent(stream_, ev_halo_, 0));
        // level l-1 is complete on both streams now (interior tents, halo tents, exchange): its final rows can go out
        if (mirror_dev_) MirrorRows(l - 1, l - 1, launches);
      }
      TentKernelArgs h = args_;
      h.tents = d_level_fast_ + fast_off_[l];
      h.ntents = fh;
      if (h.ntents > 0) { CU(launch_tent_fast(mesh_.D, order_, fosystem_, h, fa, stream_halo_)); launches++; }
      h.tents = d_level_tents_ + gen_off_[l];
      h.ntents = gh;
      if (h.ntents > 0) { CU(launch_tent_generic(mesh_.D, h, threads_, gen_smem, stream_halo_)); launches++; }
      launches += ExchangeLevel(l, stream_halo_);
      CU(cudaEventRecord(ev_halo_, stream_halo_));
    }
    TentKernelArgs a = args_;
    a.tents = d_level_fast_ + fast_off_[l] + fh;
    a.ntents = fast_off_[l + 1] - fast_off_[l] - fh;
    if (a.ntents > 0) {
      CU(launch_tent_fast(mesh_.D, order_, fosystem_, a, fa, stream_));
      launches++;
    }
    a.tents = d_level_tents_ + gen_off_[l] + gh;
    a.ntents = gen_off_[l + 1] - gen_off_[l] - gh;
    if (a.ntents > 0) {
      CU(launch_tent_generic(mesh_.D, a, threads_, gen_smem, stream_));
      launches++;
    }
    if (overlap) CU(cudaEventRecord(ev_int_, stream_));
    else if (comm_) launches += ExchangeLevel(l, stream_);
    if (mirror_dev_ && !overlap) MirrorRows(l, l, launches);
  }
  if (overlap) {
    CU(cudaStreamWaitEvent(stream_, ev_halo_, 0));
    if (mirror_dev_ && nlev > 0 && !force_generic_ && !qt_ && !dag && !p2p) MirrorRows(nlev - 1, nlev - 1, launches);
  }
  if (mirror_dev_) {  // the slab is done when the last rows have reached the host buffer
    CU(cudaEventRecord(ev_mirror_, stream_mirror_));
    CU(cudaStreamWaitEvent(stream_, ev_mirror_, 0));
  }
  CU(cudaEventRecord(ev1_, stream_));
  CU(cudaMemcpyAsync(h_flags_, d_singular_, 2 * sizeof(int), cudaMemcpyDeviceToHost, stream_));
  CU(cudaEventRecord(ev_done_, stream_));
  CU(cudaEventSynchronize(ev_done_));
  float ms = 0;
  CU(cudaEventElapsedTime(&ms, ev0_, ev1_));
  last_ms_ = ms;
  last_launches_ = launches;
  const int sing = h_flags_[0], stuck = h_flags_[1];
  if (sing || stuck) {
    CU(cudaMemsetAsync(d_singular_, 0, 2 * sizeof(int), stream_));
    if (stuck)
      throw std::logic_error("Propagate: " + std::to_string(stuck) + " tent(s) gave up waiting for their predecessors in "
                             "the dependency-driven launch (scheduling failure, not a numerical one); the state is invalid");
    throw std::domain_error("Propagate: singular local system in " + std::to_string(sing) + " pivot step(s)");
  }
  if (fa.mode == 1) ainv_valid_[fo] = true;  // (a singular system has thrown above: the inverses are then not trusted)
  if (args_.gmode == 1) ginv_valid_[fosystem_] = true;
  if (!compact) cstate_valid_ = false;   // the public rows moved on without the compact copy
  if (!partial) timeshift_ += slab_.dt;  // twavetents.cpp:173
}

// rows updated in level l that other ranks hold -> peers; rows peers updated that this rank holds <- peers
int Solver::ExchangeLevel(int l, cudaStream_t stream) {
  const PartitionPlan& p = *plan_;
  const int nr = p.nranks, rowlen = ncomp_ * rule_.nip;
  std::vector<int> sc(nr), rc(nr);
  int ns = 0, nrv = 0;
  for (int q = 0; q < nr; q++) {
    sc[q] = p.send_off[(size_t)l * nr + q + 1] - p.send_off[(size_t)l * nr + q];
    rc[q] = p.recv_off[(size_t)l * nr + q + 1] - p.recv_off[(size_t)l * nr + q];
    ns += sc[q];
    nrv += rc[q];
  }
  if (ns == 0 && nrv == 0) return 0;
  const size_t need = (size_t)p.max_rows_per_level() * rowlen * sizeof(double);
  if (halo_bytes_ < need) {
    if (d_sendbuf_) CU(cudaFree(d_sendbuf_));
    if (d_recvbuf_) CU(cudaFree(d_recvbuf_));
    CU(cudaMalloc(&d_sendbuf_, need));
    CU(cudaMalloc(&d_recvbuf_, need));
    halo_bytes_ = need;
  }
  const bool compact = UseCompact() && !qt_;
  int extra = 0;
  if (!compact) {
    comm_->Exchange(d_wavefront_, rowlen, d_send_el_ + p.send_off[(size_t)l * nr], sc.data(),
                    d_recv_el_ + p.recv_off[(size_t)l * nr], rc.data(), d_sendbuf_, d_recvbuf_, stream);
  } else {
    // compact rows of everything the level updated (what the peers' later tents read) ...
    comm_->Exchange(d_cstate_, ncomp_ * cs_points_, d_send_el_ + p.send_off[(size_t)l * nr], sc.data(),
                    d_recv_el_ + p.recv_off[(size_t)l * nr], rc.data(), d_sendbuf_, d_recvbuf_, stream);
    // ... and the public rows of the elements that became final in this level
    std::vector<int> sf(nr), rf(nr);
    int nsf = 0, nrf = 0;
    for (int q = 0; q < nr; q++) {
      sf[q] = sendf_off_[(size_t)l * nr + q + 1] - sendf_off_[(size_t)l * nr + q];
      rf[q] = recvf_off_[(size_t)l * nr + q + 1] - recvf_off_[(size_t)l * nr + q];
      nsf += sf[q];
      nrf += rf[q];
    }
    if (nsf > 0 || nrf > 0) {
      comm_->Exchange(d_wavefront_, rowlen, d_sendf_el_ + sendf_off_[(size_t)l * nr], sf.data(),
                      d_recvf_el_ + recvf_off_[(size_t)l * nr], rf.data(), d_sendbuf_, d_recvbuf_, stream);
      extra = (nsf > 0) + (nrf > 0);
    }
  }
  return (ns > 0) + (nrv > 0) + extra;
}

void Solver::ReleaseMirror() {
  if (mirror_host_) {
    cudaSetDevice(device_);
    if (stream_mirror_) cudaStreamSynchronize(stream_mirror_);
    if (mirror_registered_) cudaHostUnregister(mirror_host_);
  }
  mirror_host_ = nullptr;
  mirror_dev_ = nullptr;
  mirror_registered_ = false;
  mirror_bytes_ = 0;
}

void Solver::SetOutputMirror(double* host) {
  CU(cudaSetDevice(device_));
  ReleaseMirror();
  if (!host) return;
  if (!have_initial_) throw std::logic_error("SetOutputMirror: SetInitial has not been called (row length unknown)");
  const size_t bytes = (size_t)nel_ * ncomp_ * rule_.nip * sizeof(double);
  cudaPointerAttributes attr{};
  if (cudaPointerGetAttributes(&attr, host) != cudaSuccess) {
    cudaGetLastError();
    attr.type = cudaMemoryTypeUnregistered;
  }
  if (attr.type == cudaMemoryTypeUnregistered) {
    CU(cudaHostRegister(host, bytes, cudaHostRegisterMapped));  // pageable memory of the caller: page-lock and map it
    mirror_registered_ = true;
  } else if (attr.type != cudaMemoryTypeHost) {
    throw std::invalid_argument("SetOutputMirror: the buffer must be host memory");
  }
  mirror_host_ = host;
  mirror_bytes_ = bytes;
  void* dev = nullptr;
  cudaError_t e = cudaHostGetDevicePointer(&dev, host, 0);
  if (e != cudaSuccess) {
    ReleaseMirror();
    cuda_check(e, "cudaHostGetDevicePointer (is the buffer mapped, page-locked memory?)");
  }
  mirror_dev_ = static_cast<double*>(dev);
  if (!stream_mirror_) {
    int lo = 0, hi = 0;
    CU(cudaDeviceGetStreamPriorityRange(&lo, &hi));
    CU(cudaStreamCreateWithPriority(&stream_mirror_, cudaStreamNonBlocking, hi));
    CU(cudaEventCreateWithFlags(&ev_level_, cudaEventDisableTiming));
    CU(cudaEventCreateWithFlags(&ev_mirror_, cudaEventDisableTiming));
  }
  if (!d_final_rows_) {
    // level of the last tent that contains each element (the tents of one level touch disjoint elements); partitioned
    // solvers: of the global slab -- a ghost row is final once the exchange of that level has delivered it
    const int nlev = (int)slab_.level_off.size() - 1;
    const std::vector<int>& last = el_final_level_;
    final_off_.assign(nlev + 2, 0);
    for (int el = 0; el < nel_; el++) final_off_[last[el] + 2]++;
    for (int i = 0; i < nlev + 1; i++) final_off_[i + 1] += final_off_[i];
    std::vector<int> rows(nel_), pos(final_off_.begin(), final_off_.end() - 1);
    for (int el = 0; el < nel_; el++) rows[pos[last[el] + 1]++] = el;
    d_final_rows_ = Upload(rows);
  }
}

void Solver::MirrorRows(int first_level, int last_level, int& launches) {
  const int b = final_off_[first_level + 1], e = final_off_[last_level + 2];
  if (e <= b) return;
  CU(cudaEventRecord(ev_level_, stream_));
  CU(cudaStreamWaitEvent(stream_mirror_, ev_level_, 0));
  // a few CTAs keep PCIe busy (16 KB in flight each); each takes the place of one tent CTA (of 3 per SM) while it runs.
  // 24 measured best of 12 / 24 / 48 on B200 (profiles/r03c); ngst_twave_set_mirror_ctas changes it
  CU(launch_mirror_rows(d_final_rows_ + b, e - b, ncomp_ * rule_.nip, d_wavefront_, mirror_dev_, mirror_ctas_, stream_mirror_));
  launches++;
}

void Solver::GetWavefront(double* out) {
  if (!have_initial_) throw std::logic_error("GetWavefront: SetInitial has not been called");
  CU(cudaSetDevice(device_));
  FinishLoad();
  const size_t bytes = (size_t)nel_ * ncomp_ * rule_.nip * sizeof(double);
  CU(cudaMemcpyAsync(out, d_wavefront_, bytes, cudaMemcpyDeviceToHost, stream_));
  CU(cudaStreamSynchronize(stream_));
}

double Solver::ReduceElements(const double* a, const double* b, int mode) {
  CU(launch_element_reduce(mode, mesh_.D, nel_, rule_.nip, ncomp_, fosystem_, d_elw_, args_.c_el, d_cpoint_, a, b,
                           d_scratch_, stream_));
  std::vector<double> part(nel_);
  CU(cudaMemcpyAsync(part.data(), d_scratch_, sizeof(double) * nel_, cudaMemcpyDeviceToHost, stream_));
  CU(cudaStreamSynchronize(stream_));
  double s = 0;
  if (plan_) {
    for (int e = 0; e < nel_; e++)
      if (plan_->owned[e]) s += part[e];  // every element is counted by exactly one rank
    if (comm_) s = comm_->AllReduceSum(s, d_scalar_, stream_);
  } else {
    for (double v : part) s += v;  // fixed element order: deterministic
  }
  return s;
}

void Solver::GetWave(int* sizes, double* coef, double* psi_at_points, double* mono_coef, int* expo) {
  if (!have_initial_) throw std::logic_error("GetWave: SetInitial has not been called");
  CU(cudaSetDevice(device_));
  FinishLoad();
  const int D = mesh_.D, Q = rule_.nip;
  const int deg = fosystem_ ? order_ - 1 : order_;
  const int comps = fosystem_ ? D + 1 : 1;
  if (proj_deg_ != deg) {
    // orthonormal basis on the reference simplex: the monomials xi^alpha, |alpha| <= deg (first coordinate's exponent
    // outermost), orthonormalised in that order by modified Gram-Schmidt (twice) in the inner product of the order-2p
    // rule, which is exact for these products. NGSolve's own L2 dof numbering is not part of the reference sources.
    const std::vector<int> ex = monomial_exponents(D, deg);
    const int nd = (int)ex.size() / D;
    std::vector<double> V((size_t)Q * nd), R((size_t)nd * nd, 0.0);  // V[q*nd + k] -> orthonormal columns; mono = psi R
    for (int q = 0; q < Q; q++)
      for (int k = 0; k < nd; k++) {
        double m = 1.0;
        for (int d = 0; d < D; d++) m *= std::pow(rule_.xi[(size_t)q * D + d], ex[(size_t)k * D + d]);
        V[(size_t)q * nd + k] = m;
      }
    auto dot = [&](int a, int b) {
      double s = 0;
      for (int q = 0; q < Q; q++) s += rule_.w[q] * V[(size_t)q * nd + a] * V[(size_t)q * nd + b];
      return s;
    };
    for (int k = 0; k < nd; k++) {
      for (int pass = 0; pass < 2; pass++)
        for (int j = 0; j < k; j++) {
          const double r = dot(j, k);
          R[(size_t)j * nd + k] += r;
          for (int q = 0; q < Q; q++) V[(size_t)q * nd + k] -= r * V[(size_t)q * nd + j];
        }
      const double nrm = std::sqrt(dot(k, k));
      if (!(nrm > 0)) throw std::domain_error("GetWave: the quadrature rule does not resolve the polynomial space");
      R[(size_t)k * nd + k] = nrm;
      for (int q = 0; q < Q; q++) V[(size_t)q * nd + k] /= nrm;
    }
    // psi = mono R^-1: coefficient of monomial j in psi_k is (R^-1)[j][k]
    std::vector<double> Rinv((size_t)nd * nd, 0.0);
    for (int c = 0; c < nd; c++) {
      Rinv[(size_t)c * nd + c] = 1.0 / R[(size_t)c * nd + c];
      for (int r = c - 1; r >= 0; r--) {
        double s = 0;
        for (int j = r + 1; j <= c; j++) s += R[(size_t)r * nd + j] * Rinv[(size_t)j * nd + c];
        Rinv[(size_t)r * nd + c] = -s / R[(size_t)r * nd + r];
      }
    }
    proj_mono_.assign((size_t)nd * nd, 0.0);
    for (int k = 0; k < nd; k++)
      for (int j = 0; j < nd; j++) proj_mono_[(size_t)k * nd + j] = Rinv[(size_t)j * nd + k];
    proj_psi_ = V;
    proj_expo_ = ex;
    std::vector<double> P((size_t)nd * Q);
    for (int k = 0; k < nd; k++)
      for (int q = 0; q < Q; q++) P[(size_t)k * Q + q] = rule_.w[q] * V[(size_t)q * nd + k];
    d_proj_ = Upload(P);
    CU(cudaMalloc(&d_proj_out_, sizeof(double) * std::max<size_t>((size_t)nel_ * (D + 1) * nd, 1)));
    allocs_.push_back(d_proj_out_);
    proj_deg_ = deg;
    proj_ndof_ = nd;
  }
  const int nd = proj_ndof_;
  if (sizes) { sizes[0] = comps; sizes[1] = nd; sizes[2] = deg; }
  if (psi_at_points) std::copy(proj_psi_.begin(), proj_psi_.end(), psi_at_points);
  if (mono_coef) std::copy(proj_mono_.begin(), proj_mono_.end(), mono_coef);
  if (expo) std::copy(proj_expo_.begin(), proj_expo_.end(), expo);
  if (coef) {
    CU(launch_project_rows(nel_, Q, nd, 0, comps, ncomp_ * Q, d_proj_, d_wavefront_, d_proj_out_, stream_));
    CU(cudaMemcpyAsync(coef, d_proj_out_, sizeof(double) * (size_t)nel_ * comps * nd, cudaMemcpyDeviceToHost, stream_));
    CU(cudaStreamSynchronize(stream_));
  }
}

void Solver::InitComm(const void* id128) {
  if (!plan_) throw std::logic_error("InitComm: the solver was not created with a partition");
  CU(cudaSetDevice(device_));
  comm_ = std::make_unique<HaloComm>(plan_->rank, plan_->nranks, id128);
  if (!d_send_el_) {
    d_send_el_ = Upload(plan_->send_el);
    d_recv_el_ = Upload(plan_->recv_el);
    if (compact_possible_) {
      // the exchanged rows whose element becomes final in the level of the exchange (same order on both sides)
      auto filter = [&](const std::vector<int>& off, const std::vector<int>& el, std::vector<int>& foff, int** d_out) {
        std::vector<int> fel;
        foff.assign(1, 0);
        for (int l = 0; l < plan_->nlev; l++)
          for (int q = 0; q < plan_->nranks; q++) {
            for (int i = off[(size_t)l * plan_->nranks + q]; i < off[(size_t)l * plan_->nranks + q + 1]; i++)
              if (el_final_level_[el[i]] == l) fel.push_back(el[i]);
            foff.push_back((int)fel.size());
          }
        *d_out = Upload(fel);
      };
      filter(plan_->send_off, plan_->send_el, sendf_off_, &d_sendf_el_);
      filter(plan_->recv_off, plan_->recv_el, recvf_off_, &d_recvf_el_);
    }
    CU(cudaMalloc(&d_scalar_, sizeof(double)));
    allocs_.push_back(d_scalar_);
    InitP2P();
  }
}

namespace {
// offset of a device pointer inside its allocation: an IPC handle maps the allocation, not the pointer
size_t offset_in_allocation(const void* p) {
  using Fn = int (*)(unsigned long long*, size_t*, unsigned long long);
  static Fn fn = [] {
    void* lib = dlopen("libcuda.so.1", RTLD_NOW | RTLD_GLOBAL);
    return lib ? reinterpret_cast<Fn>(dlsym(lib, "cuMemGetAddressRange_v2")) : nullptr;
  }();
  unsigned long long base = 0;
  size_t size = 0;
  if (!fn || fn(&base, &size, (unsigned long long)p) != 0) return 0;
  return (size_t)((unsigned long long)p - base);
}
struct P2PMsg {
  int ok;
  int pad;
  cudaIpcMemHandle_t handle[3];
  unsigned long long offset[3];
};
}  // namespace

// Collective over the ranks of the partition (called from InitComm). Sets p2p_ready_ only if EVERY rank could set it up.
void Solver::InitP2P() {
  p2p_ready_ = false;
  const int nr = plan_->nranks, me = plan_->rank, D = mesh_.D, Q = rule_.nip;
  if (nr < 2) return;
  const int qs = tent_fast_compact_points(D, order_, 0);
  bool ok = compact_possible_ && qs > 0 && nr <= kMaxPeers && !have_initial_;
  P2PMsg mine{};
  const int nown = slab_.ntents;
  try {
    if (ok) {
      // state buffers once, for the second-order system (the larger one): their IPC mappings stay valid
      state_capacity_ = (size_t)nel_ * (D + 2) * Q * sizeof(double);
      cstate_capacity_ = (size_t)nel_ * (D + 2) * qs * sizeof(double);
      CU(cudaMalloc(&d_wavefront_, state_capacity_));
      CU(cudaMalloc(&d_cstate_, cstate_capacity_));
      CU(cudaMemset(d_wavefront_, 0, state_capacity_));
      CU(cudaMemset(d_cstate_, 0, cstate_capacity_));
      if (!d_remaining_) {
        CU(cudaMalloc(&d_remaining_, sizeof(int) * std::max(nown, 1)));
        allocs_.push_back(d_remaining_);
        CU(cudaMalloc(&d_ticket_, sizeof(int)));
        allocs_.push_back(d_ticket_);
      }
      void* ptrs[3] = {d_cstate_, d_wavefront_, d_remaining_};
      for (int k = 0; k < 3; k++) {
        CU(cudaIpcGetMemHandle(&mine.handle[k], ptrs[k]));
        mine.offset[k] = offset_in_allocation(ptrs[k]);
      }
    }
  } catch (const std::exception& e) {
    p2p_note_ = e.what();
    cudaGetLastError();
    ok = false;
  }
  mine.ok = ok ? 1 : 0;
  // agreement and handle exchange in one all-gather
  std::vector<P2PMsg> all(nr);
  {
    P2PMsg *d_in = nullptr, *d_out = nullptr;
    CU(cudaMalloc(&d_in, sizeof(P2PMsg)));
    CU(cudaMalloc(&d_out, sizeof(P2PMsg) * nr));
    CU(cudaMemcpyAsync(d_in, &mine, sizeof(P2PMsg), cudaMemcpyHostToDevice, stream_));
    comm_->AllGatherBytes(d_in, d_out, sizeof(P2PMsg), stream_);
    CU(cudaMemcpyAsync(all.data(), d_out, sizeof(P2PMsg) * nr, cudaMemcpyDeviceToHost, stream_));
    CU(cudaStreamSynchronize(stream_));
    CU(cudaFree(d_in));
    CU(cudaFree(d_out));
  }
  bool all_ok = true;
  for (const P2PMsg& m : all) all_ok = all_ok && m.ok;
  int failed = 0;
  if (all_ok) {
    for (int q = 0; q < nr && !failed; q++) {
      void* ptrs[3] = {d_cstate_, d_wavefront_, d_remaining_};
      if (q != me)
        for (int k = 0; k < 3; k++) {
          void* base = nullptr;
          if (cudaIpcOpenMemHandle(&base, all[q].handle[k], cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
            p2p_note_ = std::string("cudaIpcOpenMemHandle: ") + cudaGetErrorString(cudaGetLastError());
            failed = 1;
            break;
          }
          ipc_opened_.push_back(base);
          ptrs[k] = static_cast<char*>(base) + all[q].offset[k];
        }
      p2p_host_.cstate[q] = static_cast<double*>(ptrs[0]);
      p2p_host_.wavefront[q] = static_cast<double*>(ptrs[1]);
      p2p_host_.remaining[q] = static_cast<int*>(ptrs[2]);
    }
  }
  // second agreement: every rank could open every handle
  const double nfail = comm_->AllReduceSum(all_ok ? (double)failed : 1.0, d_scalar_, stream_);
  if (nfail != 0.0) {
    for (void* b : ipc_opened_) cudaIpcCloseMemHandle(b);
  for (double* p : d_ginv_) if (p) cudaFree(p);
    ipc_opened_.clear();
    if (p2p_note_.empty()) p2p_note_ = all_ok ? "another rank could not map the peer buffers" : "not every rank can use the compact state";
    // keep the fixed-capacity buffers (harmless); the per-level NCCL exchange is used
    return;
  }
  const PartitionPlan& p = *plan_;
  p2p_host_.push_off = Upload(p.p2p_push_off);
  p2p_host_.push_peer = Upload(p.p2p_push_peer);
  p2p_host_.push_el = Upload(p.p2p_push_el);
  p2p_host_.rsucc_off = Upload(p.p2p_rsucc_off);
  p2p_host_.rsucc_peer = Upload(p.p2p_rsucc_peer);
  p2p_host_.rsucc_tent = Upload(p.p2p_rsucc_tent);
  CU(cudaMalloc(&d_p2p_, sizeof(P2PArgs)));
  allocs_.push_back(d_p2p_);
  CU(cudaMemcpy(d_p2p_, &p2p_host_, sizeof(P2PArgs), cudaMemcpyHostToDevice));
  d_p2p_indeg_ = Upload(p.p2p_indeg);
  d_p2p_lsucc_off_ = Upload(p.p2p_lsucc_off);
  d_p2p_lsucc_ = Upload(p.p2p_lsucc);
  CU(cudaMalloc(&d_barrier_, sizeof(double)));
  allocs_.push_back(d_barrier_);
  CU(cudaMemset(d_barrier_, 0, sizeof(double)));
  // counters armed for the first slab, then nobody starts before everybody is armed
  CU(cudaMemcpyAsync(d_remaining_, d_p2p_indeg_, sizeof(int) * nown, cudaMemcpyDeviceToDevice, stream_));
  CU(cudaMemsetAsync(d_ticket_, 0, sizeof(int), stream_));
  comm_->Barrier(d_barrier_, stream_);
  CU(cudaStreamSynchronize(stream_));
  p2p_ready_ = true;
  p2p_note_ = "peer-to-peer exchange inside the tent kernel";
}

double Solver::Energy(const double* wf) {
  if (!have_initial_) throw std::logic_error("Energy: SetInitial has not been called");
  CU(cudaSetDevice(device_));
  FinishLoad();
  const size_t bytes = (size_t)nel_ * ncomp_ * rule_.nip * sizeof(double);
  const double* src = d_wavefront_;
  if (wf) {
    if (!d_tmp_a_) CU(cudaMalloc(&d_tmp_a_, bytes));
    CU(cudaMemcpyAsync(d_tmp_a_, wf, bytes, cudaMemcpyHostToDevice, stream_));
    src = d_tmp_a_;
  }
  return ReduceElements(src, nullptr, 0);
}

// the next Propagate applies stored inverses of the generic kernel with one warp per tent
int Solver::GenericWarpApply() const {
  if (!generic_warp_apply_ || !GenericCachePossible() || !ginv_valid_[fosystem_] || !d_ginv_[fosystem_]) return 0;
  // automatic: the orders the 1-D kernel has been run and compared bitwise at on a B200 (tests/test_gpu_generic_cache.py);
  // 3 = wherever it is compiled (orders 1..7)
  if (generic_warp_apply_ == 2)
    return (mesh_.D == 1 && order_ >= 4 && order_ <= 6 && tent_small1d_supported(args_)) ? 2 : 0;
  if (generic_warp_apply_ == 3) return (mesh_.D == 1 && tent_small1d_supported(args_)) ? 2 : 0;
  return tent_generic_warp_smem_bytes(mesh_.D, args_) <= 100 * 1024 ? 1 : 0;
}

void Solver::EnergyResidentNoAlloc(double* out) {
  if (!have_initial_) throw std::logic_error("Energy: SetInitial has not been called");
  if (plan_) throw std::logic_error("ngst_twave_energy_resident_noalloc: single-GPU solvers only (a partitioned solver sums over ranks)");
  if (nel_ > static_reduce_max_elements())
    throw std::invalid_argument("ngst_twave_energy_resident_noalloc: " + std::to_string(nel_) + " elements exceed the static scratch of " +
                                std::to_string(static_reduce_max_elements()) + " elements; use ngst_twave_energy");
  CU(cudaSetDevice(device_));
  FinishLoad();
  CU(launch_energy_static(mesh_.D, nel_, rule_.nip, ncomp_, fosystem_, d_elw_, args_.c_el, d_cpoint_, d_wavefront_, out, stream_));
}

double Solver::Error(const double* wf, const double* ref, bool l2only) {
  if (!have_initial_) throw std::logic_error("Error: SetInitial has not been called");
  CU(cudaSetDevice(device_));
  const size_t bytes = (size_t)nel_ * ncomp_ * rule_.nip * sizeof(double);
  if (!d_tmp_a_) CU(cudaMalloc(&d_tmp_a_, bytes));
  if (!d_tmp_b_) CU(cudaMalloc(&d_tmp_b_, bytes));
  CU(cudaMemcpyAsync(d_tmp_a_, wf, bytes, cudaMemcpyHostToDevice, stream_));
  CU(cudaMemcpyAsync(d_tmp_b_, ref, bytes, cudaMemcpyHostToDevice, stream_));
  return std::sqrt(ReduceElements(d_tmp_a_, d_tmp_b_, l2only ? 2 : 1));
}

// Error / L2Error (twavetents.cpp:766-822) of the RESIDENT state against analytic data sampled on the device at `time`:
// neither the state nor the exact solution crosses PCIe.
double Solver::ErrorAnalytic(int family, const double* params, int nparams, double time, bool l2only) {
  if (!have_initial_) throw std::logic_error("Error: SetInitial has not been called");
  check_family(mesh_.D, family, params, nparams);
  if (l2only && fosystem_) throw std::invalid_argument("L2Error: the first-order state has no u component");
  CU(cudaSetDevice(device_));
  FinishLoad();
  const size_t bytes = (size_t)nel_ * ncomp_ * rule_.nip * sizeof(double);
  if (!d_tmp_b_) CU(cudaMalloc(&d_tmp_b_, bytes));
  double* d_p = UploadParams(params, nparams, &d_tmp_params_);
  CU(launch_sample_wavefront(mesh_.D, family, d_p, nel_, rule_.nip, fosystem_, time, args_.pts, args_.el2v, args_.lam,
                             d_tmp_b_, stream_));
  return std::sqrt(ReduceElements(d_wavefront_, d_tmp_b_, l2only ? 2 : 1));
}

double Solver::MaxAdiam() const {
  double h = 0;
  for (double s : staged_.scale) h = std::max(h, 1.0 / s);
  return h;
}

}  // namespace ngst
