/* C ABI of libngst_b200.so: the B200-native tent-pitched Trefftz-DG wave propagator.
 *
 * Every entry point is what a binding of the reference's pybind11 surface for this path would call
 * (the reference file:line each one replaces is cited per function; paths are relative to the
 * NGSTrefftz tree, tents.* under external_dependencies/ngstents/src/). Plain pointers and sizes only.
 * All functions return 0 on success and a negative code on failure; the message is available from
 * ngst_last_error() (thread-local). No C++ exception crosses this boundary.
 *
 * Ownership: the caller owns every array passed in (it is copied during the call) and every output
 * buffer. Handles own host tables and device buffers. A handle is not thread-safe.
 */
#ifndef NGST_B200_H
#define NGST_B200_H

#ifdef __cplusplus
extern "C" {
#endif

typedef struct ngst_mesh ngst_mesh;
typedef struct ngst_slab ngst_slab;
typedef struct ngst_twave ngst_twave;

enum {
  NGST_OK = 0,
  NGST_ERR_ARG = -1,     /* invalid argument / malformed input (ngcore::Exception in the reference) */
  NGST_ERR_STATE = -2,   /* call order: not pitched, no initial data, ... */
  NGST_ERR_CUDA = -3,    /* CUDA runtime failure, including "no device": there is no CPU fallback */
  NGST_ERR_NUMERIC = -4, /* singular local system met during Propagate */
  NGST_ERR_COMM = -5     /* multi-GPU exchange failure */
};

const char* ngst_last_error(void);
const char* ngst_version(void);

/* ---------------------------------------------------------------- spatial mesh
 * Replaces the MeshAccess queries of the path (src/twavetents.cpp:113,117,380,455,472,601,616,941-952;
 * tents.cpp:233-236,248,507,532,568). bnd_kind[i] = 1 marks a boundary element named "neumann"
 * (src/twavetents.cpp:380), 0 anything else (Dirichlet). */
int ngst_mesh_create(int D, int nv, const double* pts, int nel, const int* el2v, int nbnd, const int* bnd2v,
                     const int* bnd_kind, ngst_mesh** out);
/* structured stand-ins for netgen's SegMesh / unit_square / unit_cube meshes (test/tents.py:24,38,142) */
int ngst_mesh_structured(int D, const int* ncells, const double* lo, const double* hi, ngst_mesh** out);
int ngst_mesh_set_boundary_kind(ngst_mesh* m, const int* bnd_kind);
/* counts[8] = D, nv, nel, nbnd, nedges, nfacets, 0, 0 */
int ngst_mesh_counts(const ngst_mesh* m, int* counts);
/* flat tables; pass NULL to skip. Sizes: pts nv*D, el2v nel*(D+1), bnd2v nbnd*D, bnd_kind nbnd,
 * edge2v nedges*2, facet2v nfacets*D, el2facet nel*(D+1), facet2el nfacets*2, facet2bnd nfacets,
 * el2edge nel*{1,3,6} */
int ngst_mesh_tables(const ngst_mesh* m, double* pts, int* el2v, int* bnd2v, int* bnd_kind, int* edge2v,
                     int* facet2v, int* el2facet, int* facet2el, int* facet2bnd, int* el2edge);
void ngst_mesh_destroy(ngst_mesh* m);

/* ---------------------------------------------------------------- tent slab
 * TentSlab(mesh, method, heapsize) python_tents.cpp:13-29; method 2 = "edge", 1 = "vol". */
int ngst_slab_create(const ngst_mesh* m, int method, ngst_slab** out);
/* SetMaxWavespeed python_tents.cpp:35-43: one value per element (the CF evaluated at centroids,
 * tents.cpp:491-500), or n == 1 for a constant. */
int ngst_slab_set_max_wavespeed(ngst_slab* s, const double* c, int n);
/* PitchTents python_tents.cpp:44-65 -> tents.cpp:78-359. *ok = 0 when the slab could not be completed. */
int ngst_slab_pitch(ngst_slab* s, double dt, int local_ct, double global_ct, int* ok);
/* Level-synchronous pitcher (no reference counterpart; same front-advancing rules and pole heights, tents.cpp:796-813, but a
 * pick rule that depends on the state alone: the ready vertex of minimal (level, vertex number) instead of "first vertex of
 * minimal level in the ready list", tents.cpp:431-443, whose order is a history of the whole run). All ready vertices of the
 * minimal level are handled as one phase -- the lexicographically first maximal independent set of them is pitched, in
 * parallel -- so the cost is a few passes over the mesh per level instead of a serial loop over the tents, and the result does
 * not depend on how the work is split. The tents DIFFER from the reference's (on the tested meshes: fewer tents, fewer
 * levels); they are bit-identical to the CPU oracle's serial loop with the same rule. Use ngst_slab_pitch where the
 * reference's tent order is wanted. */
int ngst_slab_pitch_level_sync(ngst_slab* s, double dt, int local_ct, double global_ct, int* ok);
/* A slab from tents pitched elsewhere, e.g. by the reference's own ngstents (what TWave(order, tps, ...) is handed,
 * src/twavetents.cpp:1194-1228): the Tent fields of python_tents.cpp:119-128 -- vertex, level, tbot, ttop, nbv, nbtime,
 * els, internal_facets -- flattened with CSR offsets (*_off: ntents+1 entries), in the slab's tent order. They are taken
 * as they are (so tent ids, list order and times are the caller's, bit for bit); dependency table, slopes and level
 * sets are derived as after ngst_slab_pitch. level == NULL: replayed as the greedy loop assigns levels (tents.cpp:199-217);
 * facet_off == facets == NULL: internal facets from the mesh. NGST_ERR_ARG for lists that do not fit the mesh or whose
 * levels do not order the tents sharing an element. Replaces ngst_slab_pitch for this slab (solvers created before
 * must be re-created). */
int ngst_slab_import_tents(ngst_slab* s, int ntents, double dt, const int* vertex, const int* level, const double* tbot,
                           const double* ttop, const int* nb_off, const int* nbv, const double* nbtime, const int* el_off,
                           const int* els, const int* facet_off, const int* facets);
int ngst_slab_ntents(const ngst_slab* s);    /* GetNTents  tents.hpp:217 */
int ngst_slab_nlayers(const ngst_slab* s);   /* GetNLayers tents.hpp:218 */
double ngst_slab_height(const ngst_slab* s); /* GetSlabHeight tents.hpp:223 */
double ngst_slab_max_slope(const ngst_slab* s); /* MaxSlope tents.cpp:366-374 */
/* sizes[4] = total nbv entries, total els entries, total facet entries, total dag entries */
int ngst_slab_sizes(const ngst_slab* s, long long* sizes);
/* Flattened struct Tent (tents.hpp:38-75, python_tents.cpp:119-128) for all tents; NULL skips.
 * *_off arrays have ntents+1 entries. */
int ngst_slab_export(const ngst_slab* s, int* vertex, int* level, double* tbot, double* ttop, double* maxslope,
                     int* nb_off, int* nbv, double* nbtime, int* el_off, int* els, int* facet_off, int* facets,
                     int* dag_off, int* dag);
/* level sets: level_off has nlevels+1 entries, level_tents ntents entries (ascending tent id per level) */
int ngst_slab_levels(const ngst_slab* s, int* nlevels, int* level_off, int* level_tents);
void ngst_slab_destroy(ngst_slab* s);

/* ---------------------------------------------------------------- Trefftz tent solver
 * TWave(order, tps, wavespeedcf) src/twavetents.cpp:1194-1228 -> TWaveTents<D> (src/twavetents.hpp:168-186).
 * c_el: wavespeed per element (the CF at element centroids, twavetents.hpp:177-185); c_vertex: the CF at the
 * vertices (TentAdiam, twavetents.cpp:871-895) or NULL (then the first adjacent element's value).
 * device: CUDA device ordinal. Fails with NGST_ERR_CUDA when no device is usable. */
int ngst_twave_create(const ngst_slab* s, int order, const double* c_el, const double* c_vertex, int device,
                      ngst_twave** out);
/* Quasi-Trefftz solver: TWave(order, tps, wavespeedcf, BBcf) -> QTWaveTents<D>, D = 1, 2 (src/twavetents.cpp:1201-1216,
 * src/twavetents.hpp:225-252; Propagate src/twavetents.cpp:967-1144). NGSolve CoefficientFunctions are replaced by
 *  - the wavespeed as a polynomial c(x) of total degree cdeg in the D space coordinates, coefficients in
 *    first-coordinate-outermost monomial order (evaluated on the device wherever the reference evaluates the CF:
 *    src/twavetents.cpp:1008, 1071, 1099; 409-412);
 *  - Taylor tables at the mesh vertices, aa_taylor[v][j] = d^alpha_j (1/c^2)(x_v) / alpha_j!, |alpha| <= order-2, and
 *    bb_taylor[v][j] = d^alpha_j B(x_v) / alpha_j!, |alpha| <= order-1 (what ComputeDerivs + Evaluate produce in
 *    src/trefftzfespace.hpp:59-126, src/trefftzfespace.cpp:1076-1089), same monomial order.
 * c_el / c_vertex as in ngst_twave_create (used for MaxAdiam only). All other ngst_twave_* calls apply. */
int ngst_qtwave_create(const ngst_slab* s, int order, const double* c_el, const double* c_vertex, int device, int cdeg,
                       const double* ccoef, const double* aa_taylor, const double* bb_taylor, ngst_twave** out);
/* Energy / Error with the wavespeed evaluated at every quadrature point, as the reference does with its CF
 * (src/twavetents.cpp:785-793, 835-843), instead of one value per element; ccoef == NULL restores the element values. */
int ngst_twave_set_wavespeed_poly(ngst_twave* t, int cdeg, const double* ccoef);

/* ---- multi-GPU (one process per GPU; the reference has no distributed Propagate).
 * The spatial mesh is partitioned by vertex ownership: vertex_owner[v] in [0, nranks). A tent runs on the owner of its
 * vertex; an element is LOCAL to every rank owning one of its vertices. Every rank pitches the same global slab and
 * passes the same owner array. For a partitioned solver every element-indexed array of this interface (wavefronts,
 * quadrature) has one row per LOCAL element, in ascending global element order (loc2glob).
 * ngst_partition_plan (host only, no device): counts[5] = own tents, local elements, total send rows, total recv rows,
 * levels; arrays may be NULL. send_off/recv_off have levels*nranks+1 entries (index level*nranks+peer); *_el are local
 * element ids; owned[i] = 1 when this rank counts local element i in Energy/Error. */
int ngst_partition_plan(const ngst_slab* s, const int* vertex_owner, int nranks, int rank, int* counts, int* own_tents,
                        int* loc2glob, int* owned, int* send_off, int* send_el, int* recv_off, int* recv_el);
/* The cross-rank hazard graph and push lists of the peer-to-peer exchange (host only; tests): counts[5] = own tents, local
 * successor entries, remote successor entries, push entries, own tent-elements; indeg per own tent (predecessors on any
 * rank); successors on this rank (own-tent index) and on other ranks (rank, its own-tent index); per own tent-element the
 * other ranks holding the element with its local index there. Arrays may be NULL. */
int ngst_partition_p2p(const ngst_slab* s, const int* vertex_owner, int nranks, int rank, long long* counts, int* indeg,
                       int* lsucc_off, int* lsucc, int* rsucc_off, int* rsucc_peer, int* rsucc_tent, int* push_off,
                       int* push_peer, int* push_el);
int ngst_twave_create_partitioned(const ngst_slab* s, int order, const double* c_el, const double* c_vertex, int device,
                                  const int* vertex_owner, int nranks, int rank, ngst_twave** out);
/* the quasi-Trefftz solver (ngst_qtwave_create) on this rank's part of the partition; the Taylor tables stay global
 * (one row per mesh vertex) */
int ngst_qtwave_create_partitioned(const ngst_slab* s, int order, const double* c_el, const double* c_vertex, int device,
                                   int cdeg, const double* ccoef, const double* aa_taylor, const double* bb_taylor,
                                   const int* vertex_owner, int nranks, int rank, ngst_twave** out);
/* NCCL bootstrap: one rank creates the 128-byte id, the caller distributes it (MPI / torch.distributed broadcast), every
 * rank calls ngst_twave_comm_init. Afterwards Propagate exchanges the updated ghost rows after every tent level with
 * grouped ncclSend/ncclRecv, and Energy / Error / L2Error return the all-reduced global value. */
int ngst_comm_unique_id(void* out128);
int ngst_twave_comm_init(ngst_twave* t, const void* id128);
/* Peer-to-peer exchange (default where ngst_twave_comm_init could set it up on EVERY rank: compact state in use, CUDA IPC
 * mappings of the other ranks' state buffers over NVLink). Propagate is then ONE dependency-driven launch per slab on every
 * rank: a tent waits for its predecessors on any rank (hazard counters), pushes the rows it writes into the buffers of the
 * other ranks that hold the element and releases its successors elsewhere with system-scope atomics -- compute and exchange
 * in one kernel, no per-level synchronisation between the GPUs; one stream-ordered NCCL barrier per slab. on == 0 (or kernel
 * mode 2): halo tents + grouped ncclSend/ncclRecv per level. on == 2: as 1, but the halo tents of a level are ticketed before
 * all its interior tents instead of behind the first third of them (A/B). ngst_twave_peer_exchange returns 1 when the next
 * Propagate will use it and copies a short reason / description into note (may be NULL). */
int ngst_twave_set_peer_exchange(ngst_twave* t, int on);
int ngst_twave_peer_exchange(const ngst_twave* t, char* note, int note_len);
/* own_tents: info[7] global tent ids; loc2glob / owned: info[6] entries; NULL skips */
int ngst_twave_partition(const ngst_twave* t, int* own_tents, int* loc2glob, int* owned);
/* info[8] = D, order, nbasis (LocalDofs), nip, ncomp, fosystem, nel (local), ntents (of this rank) */
int ngst_twave_info(const ngst_twave* t, int* info);
/* Mapped quadrature points (nel*nip*D) and weights |detJ|*w (nel*nip) of every element: what
 * MakeWavefront (src/twavetents.cpp:734-764) samples a CoefficientFunction at. */
int ngst_twave_quadrature(const ngst_twave* t, double* x, double* weight);
/* SetInitial src/twavetents.hpp:195-204: wf is nel x (ncomp*nip), row = [comp][q]; ncomp == D+2 second-order
 * input (u, grad u, dt u), ncomp == D+1 first-order input (switches to fosystem, nbasis-1). Host pointer. */
int ngst_twave_set_initial(ngst_twave* t, const double* wf, int ncomp);
/* Streamed form of ngst_twave_set_initial (no reference counterpart): returns at once. The copy engine delivers `wf` in
 * chunks of consecutive elements (~16 MB) while the NEXT ngst_twave_propagate runs as a spatial sweep: one dependency-driven
 * launch whose tents are ticketed in the order in which their dependency cones are complete -- a tent waits for its
 * predecessors and for the chunk of rows it is the first to read, whose compact form it builds on the fly. With an output
 * mirror (ngst_twave_set_output_mirror) the same propagate copies every chunk out as soon as the last tent of its elements
 * is done, so that host-to-device copy, tents and device-to-host copy overlap along the sweep. `wf` must be page-locked
 * (pageable memory is page-locked by the call) and stay unchanged until that propagate has returned; any other call that
 * reads the state waits for the whole input first. Single GPU, every tent on the fast kernel with the compact state;
 * otherwise the call is ngst_twave_set_initial. The result is bitwise that of ngst_twave_set_initial + propagate. */
int ngst_twave_set_initial_streamed(ngst_twave* t, const double* wf, int ncomp);
/* Boundary datum (SetBoundaryCF src/twavetents.hpp:206-209, evaluated at src/twavetents.cpp:373-407):
 * number of sample points and their (x, t_local) coordinates, npoints*(D+1); the datum must be evaluated at
 * (x, t_local + timeshift). values: for each boundary face in the same order nip Dirichlet values (dt u) or
 * D*nip Neumann values laid out [component][q]. NULL values = homogeneous datum. */
long long ngst_twave_boundary_npoints(const ngst_twave* t);
long long ngst_twave_boundary_nvalues(const ngst_twave* t);
int ngst_twave_boundary_points(const ngst_twave* t, double* xt, int* kind_per_point);
int ngst_twave_set_boundary_values(ngst_twave* t, const double* values, long long n);
/* Analytic data sampled on the device instead of a host-evaluated CoefficientFunction:
 * SetInitial / MakeWavefront (src/twavetents.cpp:734-764) and the boundary datum (src/twavetents.cpp:373-374, 386,
 * 407-411), which every later Propagate re-samples at (x, t_local + timeshift) until ngst_twave_set_boundary_values
 * replaces it. family / params (doubles), data = (u, grad u, d_t u):
 *   0 standing wave  [c]                          u = prod_d sin(pi x_d) sin(pi sqrt(D) c t) / (sqrt(D) pi)
 *   1 plane pulse    [c, a, s0, k_0..k_{D-1}]     u = exp(-a (k.x - c t - s0)^2), |k| = 1
 *   2 polynomial     [nterms, (coef, e_0..e_{D-1}, e_t) per term]   u = sum coef x^e t^e_t
 * first_order != 0: rows hold (grad u, d_t u) only (switches the solver to the first-order system like
 * ngst_twave_set_initial with ncomp == D+1). out: host buffer, nel x (ncomp*nip). */
int ngst_twave_set_initial_analytic(ngst_twave* t, int family, const double* params, int nparams, int first_order);
int ngst_twave_make_wavefront_analytic(ngst_twave* t, int family, const double* params, int nparams, double time,
                                       int first_order, double* out);
int ngst_twave_set_boundary_analytic(ngst_twave* t, int family, const double* params, int nparams);
/* Error (l2only == 0) / L2Error (l2only != 0) src/twavetents.cpp:766-822 of the state RESIDENT on the device against
 * the analytic data sampled on the device at `time`: no wavefront crosses PCIe. */
int ngst_twave_error_analytic(ngst_twave* t, int family, const double* params, int nparams, double time, int l2only,
                              double* out);
/* Pipelining of independent steps over the copy engines (no reference counterpart; double buffering as in any input
 * pipeline). ngst_twave_prefetch_input returns at once: the NEXT step's initial rows (wf, nel x ncomp*nip) and boundary
 * samples (bnd_values, n as ngst_twave_set_boundary_values, or NULL) are copied host-to-device into staging buffers and
 * transferred to the compact state on an input stream while the current step computes; both host buffers must be page-locked
 * and stay unchanged until ngst_twave_use_prefetched has been called. ngst_twave_use_prefetched = ngst_twave_set_initial
 * (+ ngst_twave_set_boundary_values) from the staged data, device to device. ngst_twave_get_wavefront_async snapshots the
 * state on the device and copies it to `out` (page-locked) on an output stream while the next step's input arrives and its
 * tents run (PCIe is full duplex); ngst_twave_wait_output blocks until the last such copy has landed. */
int ngst_twave_prefetch_input(ngst_twave* t, const double* wf, int ncomp, const double* bnd_values, long long nbnd);
int ngst_twave_use_prefetched(ngst_twave* t);
int ngst_twave_get_wavefront_async(ngst_twave* t, double* out);
int ngst_twave_wait_output(ngst_twave* t);
/* Propagate src/twavetents.cpp:72-175: one slab, timeshift += slab height. Fails with NGST_ERR_STATE when the slab was
 * re-pitched after the solver was created (the device tables are staged once; the reference re-reads its slab on every
 * call): create a new solver for the new tents. */
int ngst_twave_propagate(ngst_twave* t);
/* Diagnostics (no reference counterpart): the tents of the first `nlevels` tent levels only, one launch per level,
 * timeshift unchanged -- the state the reference's loop over the tents (src/twavetents.cpp:90-171) holds once these
 * tents are done. Lets the parity tests compare slabs of the benchmark's size with the CPU oracle level by level. */
int ngst_twave_propagate_levels(ngst_twave* t, int nlevels);
double ngst_twave_timeshift(const ngst_twave* t);
/* Rewind / set the accumulated time (the reference's `timeshift` member, src/twavetents.hpp:106, is only ever
 * advanced by Propagate; re-running a slab from the same initial data needs it reset). */
int ngst_twave_set_timeshift(ngst_twave* t, double timeshift);
/* name of the CUDA kernel variant Propagate launches for this solver (diagnostics / bench.py) */
const char* ngst_twave_kernel_name(const ngst_twave* t);
/* Kernel routing (diagnostics; no reference counterpart): mode 0 = automatic (single-wavespeed tents on the DMMA
 * kernel tent_kernel_fast, tents with material interfaces on tent_kernel_generic), 1 = every tent on the generic
 * DFMA kernel, 2 = as 0 but strictly level by level: on one GPU one launch per level instead of the dependency-driven
 * single launch per slab, on several GPUs the halo exchange of a level runs after all tents of the level instead of
 * overlapping the interior tents; 3 = as 0 with the dependency-driven single launch forced (mode 0 uses it only for
 * slabs with small levels). Results are bitwise identical in modes 0, 2 and 3.
 * ngst_twave_ntents_fast: tents of the slab routed to the fast kernel in mode 0. */
int ngst_twave_set_kernel_mode(ngst_twave* t, int mode);
/* Compact state (no reference counterpart; default on). Where every tent of the slab runs on the fast kernel, the solver
 * keeps a second copy of the state at the points of the compact rule exact to degree 2p (46 instead of 125 points at
 * D = 3, p = 4), which the tents read and write; a row written by a tent is a polynomial of degree <= p on its element
 * (src/twavetents.cpp:541-590), so nothing is lost, and the bottom-face integrals of src/twavetents.cpp:230-267 agree for
 * sampled initial rows as well (ngst_rule_transfer). The public rows (ngst_twave_get_wavefront, Energy, Error, mirror) are
 * written by the last tent of the slab that contains the element. on == 0: the tents work on the public rows (A/B).
 * ngst_twave_compact_state: 1 when the next Propagate will use it. */
int ngst_twave_set_compact_state(ngst_twave* t, int on);
int ngst_twave_compact_state(const ngst_twave* t);
/* Factor cache (no reference counterpart; default on where the compact state is in use). The matrix of a tent depends on
 * the tent's geometry and the wavespeed only -- not on the state, the boundary datum or the time -- and a slab is
 * propagated many times (docs/notebooks/twavetents.ipynb: one TentSlab, a loop of Propagate calls). The reference
 * re-assembles and re-inverts it in every Propagate (src/twavetents.cpp:119-164, 64-70). Here the first Propagate forms
 * the inverse of every tent matrix (as the reference's Solve does) and keeps it in HBM (8 N (N rounded up to 8) bytes per
 * tent: 24.6 kB at D = 3, p = 4); every later Propagate builds the right-hand sides (src/twavetents.cpp:230-267, 373-433),
 * multiplies by the stored inverses and evaluates (src/twavetents.cpp:541-590). Results are those of inverse-then-multiply,
 * i.e. the reference's own arithmetic. on == 0: assemble and solve in every Propagate (the memory is kept until the solver
 * is destroyed). If the memory cannot be allocated the solver silently keeps assembling.
 * ngst_twave_factor_cache returns 0 = not in use, 1 = the next Propagate builds the cache, 2 = the next Propagate applies it;
 * *bytes (may be NULL) = device memory held by the cache.
 * Values of `on` above 1 select how stored inverses are applied and leave the cache switch alone: 2 = fast kernel compiled for four
 * CTAs per SM (A/B; measured slower); generic kernel (systems of up to 32 unknowns, e.g. all 1-D tents): 3 = automatic (default:
 * one warp per tent with lanes as basis functions, tent_kernel_small1d, for D = 1 and orders 4..6; one CTA per tent otherwise),
 * 4 = one CTA per tent, 5 = the CTA code executed by one warp per tent (A/B), 6 = tent_kernel_small1d at every order it is
 * compiled for (1..7; orders 1..3 and 7 have not been run on a GPU yet, hence not automatic). All variants give the same bits. */
int ngst_twave_set_factor_cache(ngst_twave* t, int on);
int ngst_twave_factor_cache(const ngst_twave* t, long long* bytes);
int ngst_twave_ntents_fast(const ngst_twave* t);
/* device time of the last Propagate in ms (CUDA events on the solver's stream) and kernel launches in it */
double ngst_twave_last_ms(const ngst_twave* t);
int ngst_twave_last_launches(const ngst_twave* t);
/* GetWavefront src/twavetents.hpp:193: copy of the nel x (ncomp*nip) state to a host buffer */
int ngst_twave_get_wavefront(ngst_twave* t, double* out);
/* Output mirror (no reference counterpart; the fused form of Propagate + GetWavefront src/twavetents.hpp:193): `host`
 * (nel x ncomp*nip doubles; page-locked and mapped by the call if it is plain pageable memory) receives every state row
 * during ngst_twave_propagate as soon as the last tent that contains its element has run, over PCIe while the remaining
 * tents are solved; when ngst_twave_propagate returns it holds exactly what ngst_twave_get_wavefront would copy. The
 * buffer must stay valid until the mirror is switched off (host == NULL), ngst_twave_set_initial changes the row length,
 * or the solver is destroyed. Call after ngst_twave_set_initial. Partitioned solvers mirror their local rows (owned and ghost),
 * a ghost row after the halo exchange of the level that completes it. */
int ngst_twave_set_output_mirror(ngst_twave* t, double* host);
/* number of copy CTAs the output mirror keeps in flight (default 24, the best of 12 / 24 / 48 measured on B200) */
int ngst_twave_set_mirror_ctas(ngst_twave* t, int ctas);
/* GetWave src/__init__.py:7-43 (per-element L2 projection of the quadrature-point state, IntegrationPointFunction
 * src/specialcoefficientfunction.cpp:48-172): second-order state -> u onto polynomials of degree p (1 component),
 * first-order state -> (grad u, d_t u) onto degree p-1 (D+1 components). The basis is orthonormal on the reference
 * simplex (monomials, first coordinate outermost, Gram-Schmidt in that order); NGSolve's L2 dof numbering is not part of
 * the reference sources. sizes[3] = {components, ndof, degree}; coef: nel x components x ndof; psi_at_points: nip x ndof;
 * mono_coef: ndof x ndof (psi_k = sum_j mono_coef[k*ndof+j] xi^expo_j); expo: ndof x D. Null outputs are skipped. */
int ngst_twave_get_wave(ngst_twave* t, int* sizes, double* coef, double* psi_at_points, double* mono_coef, int* expo);
/* Energy / Error / L2Error src/twavetents.cpp:766-847 on host wavefronts (nel x ncomp*nip) */
int ngst_twave_energy(ngst_twave* t, const double* wf, double* out);
/* Energy of the RESIDENT state (what ngst_twave_energy(t, NULL, out) returns, same summation order, same bits) for
 * callers that must not allocate: no counterpart in the reference. The
 * per-element partial sums and the result slot live in a file-scope `static __device__` array of the library
 * (ngst_twave_energy_noalloc_max_elements() = 262 144 elements, 2 MB held by every GPU that loads the library until
 * exit); the call does no cudaMalloc / cudaFree and no host allocation: two kernel launches, one 8-byte copy into *out
 * and a stream synchronisation on the solver's stream. Single-GPU solvers only; more elements than the maximum ->
 * NGST_ERR_ARG (use ngst_twave_energy). All calls on a GPU share the array -- it keeps the last call's contents, none of
 * which is read by the next call -- so calls that overlap on the GPU (two solvers from two host threads) race. */
int ngst_twave_energy_resident_noalloc(ngst_twave* t, double* out);
int ngst_twave_energy_noalloc_max_elements(void);
int ngst_twave_error(ngst_twave* t, const double* wf, const double* wf_ref, double* out);
int ngst_twave_l2error(ngst_twave* t, const double* wf, const double* wf_ref, double* out);
/* MaxAdiam src/twavetents.cpp:906-914 */
int ngst_twave_max_adiam(const ngst_twave* t, double* out);
/* CSR Trefftz basis (TWaveBasis<D+1>::Basis src/trefftzfespace.cpp:611-682); sizes[3] = nbasis, npoly, nnz */
int ngst_twave_basis(const ngst_twave* t, int* sizes, int* rowptr, int* col, double* val);
void ngst_twave_destroy(ngst_twave* t);

/* Measured FP64 roofline denominator of `device` in TFLOP/s: which 0 = DFMA pipe, 1 = DMMA (mma.sync m8n8k4 f64),
 * 2 = both interleaved in one instruction stream (4:1 flops; shows whether the two datapaths overlap).
 * No reference counterpart (the reference is CPU-only); used by bench.py for roofline.peak. */
int ngst_device_fp64_peak(int device, int which, double* tflops);

/* host-only helpers exposed for tests (no device needed) */
int ngst_basis(int D, int order, int fowave, int* sizes, int* rowptr, int* col, double* val, int* expo);
int ngst_rule(int dim, int order, int* nip, double* xi, double* w);
/* Install a rule table dumped from a real NGSolve install (IntegrationRule(eltyp, order) points and weights, without the
 * zero-weight SIMD pad points) in place of the restated rule of (dim, order): every solver
 * created afterwards samples, stores and exchanges its state at these points in this order, so that wavefront matrices
 * correspond column for column to the reference's (src/twavetents.cpp:77, 739). The table must have as many points as the
 * built-in rule (the kernels are sized for it), positive weights, and integrate all monomials of total degree <= order
 * exactly (checked to 1e-11 relative); otherwise NGST_ERR_ARG. nip == 0 removes the override. ngst_rule returns the rule in
 * force. */
int ngst_rule_override(int dim, int order, int nip, const double* xi, const double* w);
/* Point-value transfer from the state rule (ngst_rule(dim, order_from)) to the compact rule (ngst_compact_rule(dim,
 * degree_to)) through the polynomials of total degree <= poly_degree: sizes[2] = {rows = compact points, cols = state
 * points}, M row-major. With both rules exact to 2*poly_degree, sum_q w_q phi(x_q) g_q = sum_r v_r phi(y_r) (M g)_r for every
 * g and every polynomial phi of degree <= poly_degree, and M g = g at the compact points for polynomial g: the identity
 * behind integrating a tent's right-hand side (src/twavetents.cpp:230-267) on the compact rule. */
int ngst_rule_transfer(int dim, int order_from, int degree_to, int poly_degree, int* sizes, double* M);
/* compact fully symmetric rule exact to `degree` (top-face Gram of the fast kernel; no reference counterpart); falls back
 * to ngst_rule(dim, degree) where no table exists */
int ngst_compact_rule(int dim, int degree, int* nip, double* xi, double* w);

#ifdef __cplusplus
}
#endif
#endif
