// Host-side owner of the device state of one TWave solver (TWaveTents<D>, src/twavetents.hpp:95-224).
#pragma once
#include <cuda_runtime.h>

#include <memory>
#include <string>
#include <vector>

#include "../host/basis.hpp"
#include "../host/mesh.hpp"
#include "../host/pitcher.hpp"
#include "../host/quadrature.hpp"
#include "../host/partition.hpp"
#include "../host/stage.hpp"
#include "halo.hpp"
#include "tent_fast.cuh"
#include "tent_kernel.cuh"
#include "tent_qt.cuh"

namespace ngst {

// Inputs of the quasi-Trefftz solver (QTWaveTents, src/twavetents.hpp:225-252) that NGSolve CoefficientFunctions provide in
// the reference: the wavespeed as a polynomial in x and the Taylor tables of 1/c^2 and B at the mesh vertices.
struct QtHostData {
  int cdeg = 0;
  std::vector<double> ccoef;  // monomial_exponents(D, cdeg) order
  std::vector<double> aa;     // nv * nA, nA = #monomials of degree <= order-2
  std::vector<double> bb;     // nv * nB, nB = #monomials of degree <= order-1
};

class Solver {
 public:
  // plan == nullptr: the whole slab on one GPU. Otherwise `slab` is the rank's restricted slab (restrict_slab) and
  // plan its partition plan; c_el / c_vertex stay GLOBAL arrays; every element-indexed array of the public interface
  // (wavefronts, quadrature) is then indexed by LOCAL element (plan->loc2glob).
  Solver(const TentSlab& slab, int order, const std::vector<double>& c_el, const std::vector<double>& c_vertex,
         int device, const PartitionPlan* plan = nullptr, const QtHostData* qt = nullptr);
  // wavespeed polynomial for Energy / Error with c evaluated at every quadrature point (twavetents.cpp:785-793, 835-843)
  void SetWavespeedPoly(int deg, const double* coef);
  bool is_qt() const { return qt_; }
  ~Solver();
  Solver(const Solver&) = delete;
  Solver& operator=(const Solver&) = delete;

  int D() const { return mesh_.D; }
  int order() const { return order_; }
  int nbasis() const { return basis_.nbasis; }
  int nip() const { return rule_.nip; }
  int ncomp() const { return ncomp_; }
  int fosystem() const { return fosystem_; }
  int nel() const { return nel_; }
  const PartitionPlan* plan() const { return plan_; }
  void InitComm(const void* id128);
  int ntents() const { return slab_.ntents; }
  const BasisCsr& basis() const { return basis_; }
  const StagedTents& staged() const { return staged_; }
  const SimplexRule& rule() const { return rule_; }

  void Quadrature(double* x, double* weight) const;
  void SetInitial(const double* wf, int ncomp);
  // Streamed input: as SetInitial, but returns at once: the copy engine delivers `wf` (page-locked by the call if it is
  // pageable; must stay unchanged until the next Propagate has returned) in chunks of consecutive elements while that
  // Propagate runs its tents in the order in which their dependency cones are complete (and, with an output mirror, copies
  // finished chunks out at the other end). Falls back to SetInitial where the sweep does not apply.
  void SetInitialStreamed(const double* wf, int ncomp);
  bool sweep_active() const { return load_pending_ && sweep_built_; }
  // Pipelining of independent steps (double buffering, copy engines): PrefetchInput copies the NEXT step's initial rows and
  // boundary samples into staging buffers on the input stream (and transfers them to the compact state there) while the
  // current step computes; UsePrefetched makes them the state with device-to-device copies. GetWavefrontAsync snapshots the
  // result on the device and copies it to `out` (page-locked) on the output stream; WaitOutput blocks until it has landed.
  void PrefetchInput(const double* wf, int ncomp, const double* bnd_values, long long nbnd);
  void UsePrefetched();
  void GetWavefrontAsync(double* out);
  void WaitOutput();
  void SetBoundaryValues(const double* values, long long n);
  // Analytic data sampled on the device (cuda/sample.hpp): SetInitial / MakeWavefront
  // (twavetents.cpp:734-764) and the boundary datum, re-sampled by every Propagate at t_local + timeshift
  // (twavetents.cpp:373-374), without a host evaluation or an H2D copy of samples.
  void SetInitialAnalytic(int family, const double* params, int nparams, int first_order);
  void MakeWavefrontAnalytic(int family, const double* params, int nparams, double time, int first_order, double* out);
  void SetBoundaryAnalytic(int family, const double* params, int nparams);
  double ErrorAnalytic(int family, const double* params, int nparams, double time, bool l2only);
  void Propagate(int level_limit = -1);
  void GetWavefront(double* out);
  // Output mirror (cuda/mirror.cu): `host` (nel x ncomp*nip doubles; page-locked here if it is not already) receives every
  // state row as soon as it is final during Propagate, overlapped with the remaining tents; when Propagate returns it
  // holds what GetWavefront would copy (partitioned solvers: their local rows). nullptr switches the mirror off.
  void SetOutputMirror(double* host);
  double Energy(const double* wf);
  // Energy of the resident state without cudaMalloc / cudaFree / host allocation (scratch: static device array in
  // reduce.cu, shared by all calls on a GPU). Single-GPU solvers with at most static_reduce_max_elements() elements.
  void EnergyResidentNoAlloc(double* out);
  double Error(const double* wf, const double* ref, bool l2only);
  double MaxAdiam() const;
  // GetWave (src/__init__.py:7-43): L2 projection of the state onto polynomials, element by element. Second-order state:
  // u onto degree p (1 component); first-order state: (grad u, d_t u) onto degree p-1 (D+1 components).
  // sizes[3] = {components, ndof per component, degree}. Null pointers are skipped. coef: nel x comps x ndof;
  // psi_at_points: nip x ndof values of the orthonormal reference basis at the rule points; mono_coef: ndof x ndof,
  // psi_k = sum_j mono_coef[k*ndof + j] xi^expo_j; expo: ndof x D.
  void GetWave(int* sizes, double* coef, double* psi_at_points, double* mono_coef, int* expo);

  double timeshift() const { return timeshift_; }
  void set_timeshift(double t) { timeshift_ = t; }
  const char* kernel_name() const { if (!qt_ && !(fast_ && !force_generic_) && GenericWarpApply()) return GenericWarpApply() == 2 ? "tent_kernel_small1d" : "tent_kernel_generic_warp"; return qt_ ? ((qtfast_ && !force_generic_ && !fosystem_) ? "tent_kernel_qt_fast" : "tent_kernel_qt") : ((fast_ && !force_generic_) ? "tent_kernel_fast" : "tent_kernel_generic"); }
  int ntents_fast() const { return (fast_ || qtfast_) ? fast_off_.back() : 0; }
  // 0 = automatic routing (fast kernel where possible), 1 = every tent on the generic DFMA kernel (A/B evidence)
  // 0: fast kernel where available (default), 1: generic kernel everywhere, 2: fast kernel with one launch per level
  // (no dependency-driven scheduling on one GPU, exchange not overlapped on several), 3: as 0 with the dependency-driven
  // single launch forced (default: only for slabs with small levels, where launches and level tails dominate)
  void set_kernel_mode(int mode) { force_generic_ = mode == 1; overlap_ = mode != 2; dag_mode_ = mode == 2 ? 0 : (mode == 3 ? 2 : 1); }
  void set_mirror_ctas(int n) { mirror_ctas_ = n; }
  // Compact state (tent_fast.cu, CS = 1; default on): slabs whose tents all run on the fast kernel keep an internal copy of
  // the state on the compact degree-2p rule that the tents read and write; off = the tents work on the public rows (A/B).
  void set_compact_state(bool on) { compact_enabled_ = on; }
  bool compact_state_active() const { return UseCompact(); }
  // Factor cache (tent_fast.cu MODE 1 / 2; default on, needs the compact state): the inverse of every tent matrix is formed
  // by the first Propagate and stored in HBM (NP x N doubles per tent); later Propagates of the same slab geometry only
  // build right-hand sides, apply the stored inverses and evaluate. off: assemble and solve in every Propagate.
  void set_factor_cache(bool on, bool four_ctas = false) { factor_cache_enabled_ = on; apply_four_ctas_ = four_ctas; }
  // generic kernel, stored inverses: one warp per tent (tent_kernel_generic<D, true>) instead of one CTA
  // (0 = CTA per tent, 1 = the CTA code on a warp (A/B), 2 = automatic, the default: the 1-D lanes-as-functions kernel
  // tent_kernel_small1d at D = 1, orders 4..6 -- the ones run on a B200 --, a CTA per tent otherwise, 3 = that kernel at
  // every order it is compiled for, 1..7)
  void set_generic_warp_apply(int mode) { generic_warp_apply_ = mode; }
  int GenericWarpApply() const;   // 0 = CTA, 1 = tent_kernel_generic<D, true>, 2 = tent_kernel_small1d
  // multi-GPU: exchange inside the tent kernel over peer memory (default where it could be set up) or per level with NCCL
  void set_peer_exchange(bool on, bool halo_first = false) { p2p_enabled_ = on; p2p_halo_first_ = halo_first; }
  bool peer_exchange_active() const { return p2p_ready_ && p2p_enabled_ && UseCompact() && !qt_; }
  const std::string& peer_exchange_note() const { return p2p_note_; }
  // 0 = not in use, 1 = the next Propagate assembles, inverts and stores, 2 = the next Propagate applies stored inverses
  int factor_cache_state() const {
    if (UseCompact() && factor_cache_enabled_ && !cache_failed_) return ainv_valid_[fosystem_] ? 2 : 1;
    if (GenericCachePossible()) return ginv_valid_[fosystem_] ? 2 : 1;
    return 0;
  }
  size_t factor_cache_bytes() const { return ainv_bytes_[0] + ainv_bytes_[1] + ginv_bytes_[0] + ginv_bytes_[1]; }
  // the generic kernel keeps inverses of systems up to 32 unknowns (D = 1 up to order 7 with two wavespeed regions), when
  // every tent of the solver runs on it
  bool GenericCachePossible() const {
    return factor_cache_enabled_ && !ginv_failed_ && !qt_ && have_initial_ && (force_generic_ || !fast_) &&
           args_.max_ndom * args_.Npad <= 32 && threads_ >= args_.max_ndom * args_.Npad;
  }
  double last_ms() const { return last_ms_; }
  int last_launches() const { return last_launches_; }

 private:
  void BuildBasis();
  template <typename T>
  T* Upload(const std::vector<T>& v);
  double ReduceElements(const double* d_a, const double* d_b, int mode);
  int ExchangeLevel(int level, cudaStream_t stream);
  size_t halo_bytes_ = 0;

  const SimplexMesh& mesh_;
  const TentSlab& slab_;
  const PartitionPlan* plan_ = nullptr;
  int nel_ = 0;                      // local elements (== mesh_.nel without a partition)
  std::vector<int> loc2glob_;        // local -> global element
  std::unique_ptr<HaloComm> comm_;
  int* d_send_el_ = nullptr;
  int* d_recv_el_ = nullptr;
  double* d_sendbuf_ = nullptr;
  double* d_recvbuf_ = nullptr;
  double* d_scalar_ = nullptr;
  int order_, device_;
  std::vector<double> c_el_, c_vertex_;
  SimplexRule rule_;
  BasisCsr basis_;
  StagedTents staged_;
  int ncomp_ = 0, fosystem_ = 0;
  bool have_initial_ = false;
  double timeshift_ = 0, last_ms_ = 0;
  int last_launches_ = 0;

  cudaStream_t stream_ = nullptr;
  cudaEvent_t ev0_ = nullptr, ev1_ = nullptr, ev_done_ = nullptr;
  // multi-GPU overlap: halo tents + exchange of a level run on a high-priority stream beside the interior tents
  cudaStream_t stream_halo_ = nullptr;
  cudaEvent_t ev_halo_ = nullptr, ev_int_ = nullptr;
  bool overlap_ = true;
  // single GPU: dependency-driven scheduling of the fast kernel (one launch per slab)
  bool dag_ = false;
  int dag_mode_ = 1;  // 0 never, 1 automatic (average level smaller than two waves of CTAs), 2 always
  int *d_indeg_ = nullptr, *d_remaining_ = nullptr, *d_ticket_ = nullptr, *d_succ_off_ = nullptr, *d_succ_ = nullptr;
  std::vector<int> fast_nhalo_, gen_nhalo_;  // per level: leading halo tents of the fast / generic list
  std::vector<void*> allocs_;
  TentKernelArgs args_{};
  int* d_level_tents_ = nullptr;   // per level: generic-kernel tents (material interfaces or no fast variant)
  int* d_level_fast_ = nullptr;    // per level: fast-kernel tents, most elements first
  std::vector<int> gen_off_, fast_off_;  // nlevels + 1 offsets into the two lists
  bool fast_ = false, force_generic_ = false, qt_ = false;
  QtArgs qargs_{};
  QtFastArgs qfargs_{};
  bool qtfast_ = false;            // quasi-Trefftz tents with qt_ok_ run on tent_kernel_qt_fast
  std::vector<int> qt_ok_;
  double* d_cpoint_ = nullptr;     // nel * nip wavespeed per quadrature point (QT; optional for reductions)
  int* d_level_all_ = nullptr;     // per level: all tents (slab order), for force_generic_
  FastStage fstage_;
  FastArgs fargs_{};
  int* d_cptr_[2] = {nullptr, nullptr};
  int* d_crow_[2] = {nullptr, nullptr};
  double* d_cval_[2] = {nullptr, nullptr};
  double* d_wavefront_ = nullptr;
  double* d_bnd_ = nullptr;
  double* mirror_host_ = nullptr;   // caller's buffer, its device alias, and whether this solver page-locked it
  double* mirror_dev_ = nullptr;
  bool mirror_registered_ = false;
  size_t mirror_bytes_ = 0;
  int* d_final_rows_ = nullptr;     // local elements ordered by the level of the last tent that contains them
  std::vector<int> final_off_;      // nlevels + 2 offsets: [0] rows no tent touches, [l + 1] rows final after level l
  cudaStream_t stream_mirror_ = nullptr;
  cudaEvent_t ev_level_ = nullptr, ev_mirror_ = nullptr;
  void ReleaseMirror();
  void MirrorRows(int first_level, int last_level, int& launches);  // rows final after levels first..last (-1 = untouched)
  int bnd_family_ = -1;            // >= 0: analytic boundary datum (AnalyticFamily), sampled on the device per slab
  double* d_bnd_params_ = nullptr;
  double* d_tmp_params_ = nullptr;
  double* UploadParams(const double* params, int nparams, double** slot);
  double* d_elw_ = nullptr;   // nel*nip weights |detJ| w
  double* d_scratch_ = nullptr;
  double* d_tmp_a_ = nullptr;
  double* d_tmp_b_ = nullptr;
  double* d_proj_ = nullptr;       // projection matrix of GetWave (ndof x nip), built on first use
  double* d_proj_out_ = nullptr;
  int proj_deg_ = -1, proj_ndof_ = 0;
  std::vector<double> proj_psi_, proj_mono_;
  std::vector<int> proj_expo_;
  // ---- spatial sweep: streamed input and output (SetInitialStreamed; single GPU, every tent on the fast kernel, compact
  // state). The copy engine delivers the public rows in chunks of consecutive elements and takes finished chunks out while
  // ONE dependency-driven launch runs the tents in the order in which their dependency cones are complete.
  const double* load_host_ = nullptr;   // caller's input buffer, page-locked by us if load_registered_
  bool load_registered_ = false;
  bool load_pending_ = false;           // chunk copies are enqueued on stream_in_; ev_in_done_ = all have landed
  bool sweep_built_ = false, sweep_possible_ = false;
  int sweep_nchunks_ = 0, sweep_chunk_elems_ = 0, sweep_untouched_ = 0;
  int* d_sweep_tents_ = nullptr;        // tents ordered by (last chunk their cone needs, level)
  int* d_tent_chunk_ = nullptr;         // per tent: last chunk its own elements lie in
  int* d_chunk_loaded_ = nullptr;       // per chunk: set by a stream memory operation behind the chunk's copy
  int* d_chunk_left_ = nullptr;         // per chunk: touched elements not yet final (armed from d_chunk_left_init_)
  int* d_chunk_left_init_ = nullptr;
  std::vector<int> chunk_left_init_;
  unsigned char* d_te_first_ = nullptr; // per tent-element: first tent of the slab to touch the element
  int* d_untouched_ = nullptr;          // elements no tent touches (their compact rows are transferred separately)
  cudaStream_t stream_in_ = nullptr, stream_out_ = nullptr;
  cudaStream_t stream_proj_ = nullptr;          // compact transfer of prefetched chunks: keeps kernels out of the copy stream
  std::vector<cudaEvent_t> ev_chunk_;           // chunk c of the prefetched input has landed
  cudaEvent_t ev_in_start_ = nullptr, ev_in_done_ = nullptr, ev_out_start_ = nullptr, ev_out_done_ = nullptr;
  double *d_stage_wf_ = nullptr, *d_stage_cs_ = nullptr, *d_stage_bnd_ = nullptr, *d_snap_ = nullptr;
  size_t stage_wf_bytes_ = 0, stage_cs_bytes_ = 0, snap_bytes_ = 0;
  int stage_ncomp_ = 0;
  bool stage_valid_ = false, stage_has_bnd_ = false, out_pending_ = false;
  cudaEvent_t ev_stage_ = nullptr, ev_snap_ = nullptr, ev_copied_ = nullptr, ev_used_ = nullptr;
  void EnsurePipeStreams();
  void BuildSweep();
  void FinishLoad();                    // the whole input has landed before the solver's stream goes on
  void ReleaseLoadBuffer();
  // ---- compact state
  bool compact_enabled_ = true;
  bool compact_possible_ = false;     // every tent of the slab runs on the fast kernel
  int cs_points_ = 0;                 // points of the compact rule for the current system, 0 = no compact variant
  double* d_cstate_ = nullptr;        // nel * ncomp * cs_points_
  size_t cstate_bytes_ = 0;
  bool cstate_valid_ = false;         // d_cstate_ corresponds to d_wavefront_
  double* d_transfer_ = nullptr;      // cs_points_ x nip: order-2p points -> compact points through degree-p polynomials
  int transfer_points_ = 0;
  std::vector<int> el_final_level_;   // per local element: level of the last tent (of the global slab) containing it
  unsigned char* d_te_final_ = nullptr;          // per staged tent-element: its tent is that last tent
  unsigned char* d_te_final_partial_ = nullptr;  // the same for a partial run (propagate_levels)
  int partial_levels_ = -1;
  // multi-GPU: per (level, peer) the exchanged rows whose element becomes final in that level (public rows follow)
  std::vector<int> sendf_off_, recvf_off_;
  int *d_sendf_el_ = nullptr, *d_recvf_el_ = nullptr;
  // ---- peer-to-peer exchange inside the tent kernel (multi-GPU): one dependency-driven launch per slab on every rank; tents
  // push the rows they write into the other holders' buffers (CUDA IPC mappings over NVLink) and release their successors
  // on other ranks with system-scope atomics. Falls back to the per-level NCCL exchange when it cannot be set up.
  bool p2p_enabled_ = true, p2p_ready_ = false, p2p_halo_first_ = false;
  std::vector<int> p2p_order_;   // ticket order of the peer-to-peer launch: level by level, halo tents behind a third of the interior ones
  int* d_p2p_tents_ = nullptr;
  P2PArgs p2p_host_{};
  P2PArgs* d_p2p_ = nullptr;
  std::vector<void*> ipc_opened_;
  int* d_p2p_indeg_ = nullptr;
  int *d_p2p_lsucc_off_ = nullptr, *d_p2p_lsucc_ = nullptr;
  double* d_barrier_ = nullptr;
  size_t state_capacity_ = 0, cstate_capacity_ = 0;  // p2p: buffers allocated once for the largest system (IPC handles stay valid)
  void InitP2P();
  std::string p2p_note_;
  // factor cache of the generic kernel (small systems only: 1-D, low orders, material-interface tents of such solvers)
  double* d_ginv_[2] = {nullptr, nullptr};  // [fosystem]
  size_t ginv_bytes_[2] = {0, 0};
  bool ginv_valid_[2] = {false, false}, ginv_failed_ = false;
  size_t smem_build_ = 0;                   // shared memory of the generic kernel while it forms the inverses
  bool factor_cache_enabled_ = true, cache_failed_ = false;
  int generic_warp_apply_ = 2;  // stored inverses of the generic kernel applied by one warp per tent (set_generic_warp_apply)
  bool apply_four_ctas_ = false;   // apply variant with 128 registers / 55 KB: four CTAs per SM, inverse read from L2 (A/B)
  double* d_ainv_[2] = {nullptr, nullptr};  // [fosystem]
  size_t ainv_bytes_[2] = {0, 0};
  bool ainv_valid_[2] = {false, false};
  bool UseCompact() const { return cs_points_ > 0 && compact_enabled_ && !force_generic_; }
  void ProjectCompact();
  const unsigned char* FinalFlags(int level_limit);
  int* d_singular_ = nullptr;      // device words: [0] singular pivot steps, [1] dependency time-outs
  int* h_flags_ = nullptr;         // their page-locked host copy, refreshed by every Propagate
  int mirror_ctas_ = 24;
  int* d_rowptr_[2] = {nullptr, nullptr};  // [fosystem]
  int* d_col_[2] = {nullptr, nullptr};
  double* d_val_[2] = {nullptr, nullptr};
  BasisCsr basis_so_, basis_fo_;
  size_t wavefront_bytes_ = 0;
  int threads_ = 256;
  size_t smem_ = 0;
};

// kernels (tent_generic.cu, reduce.cu)
size_t tent_generic_smem_bytes(int D, const TentKernelArgs& a);
bool tent_small1d_supported(const TentKernelArgs& a);   // sizes tent_kernel_small1d (gwarp 2) is compiled for
size_t tent_generic_warp_smem_bytes(int D, const TentKernelArgs& a);  // per CTA of the warp-per-tent variant (gmode 2, gwarp)
cudaError_t launch_tent_generic(int D, const TentKernelArgs& a, int threads, size_t smem, cudaStream_t stream);
cudaError_t launch_project_rows(int nel, int Q, int ndof, int comp0, int ncomp_used, int rowlen, const double* P,
                                const double* wf, double* out, cudaStream_t stream, const int* elems = nullptr);
cudaError_t launch_mirror_rows(const int* rows, int nrows, int rowlen, const double* wf, double* host_alias, int ctas,
                               cudaStream_t stream);
cudaError_t launch_element_reduce(int mode, int D, int nel, int Q, int ncomp, int fosystem, const double* elw,
                                  const double* c_el, const double* c_point, const double* a, const double* b,
                                  double* out, cudaStream_t stream);
// Energy with scratch and result slot in a file-scope device array (reduce.cu): no allocation; nel <= static_reduce_max_elements()
int static_reduce_max_elements();
cudaError_t launch_energy_static(int D, int nel, int Q, int ncomp, int fosystem, const double* elw, const double* c_el,
                                 const double* c_point, const double* a, double* host_out, cudaStream_t stream);

// fp64_peak.cu: which 0 = DFMA pipe, 1 = DMMA (mma.sync m8n8k4 f64); TFLOP/s on the current device
cudaError_t measure_fp64_peak(int which, int reps, double* tflops);

}  // namespace ngst
