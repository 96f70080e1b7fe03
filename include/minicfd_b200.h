/* minicfd_b200.h — C ABI of libminicfd_b200.so
 *
 * B200-native (sm_100a) implementation of miniCFD's per-Runge-Kutta-stage spatial residual
 * (volume flux divergence + face Riemann flux + lifting) and the explicit RK update that
 * consumes it, on Hexahedron8 meshes with nodal Lobatto expansions, FP64 throughout.
 *
 * The reference has no C ABI: its seam is a set of C++ templates.  Every
 * entry point below therefore cites the reference member function(s) it stands in for; the header
 * shim under include/mini/ keeps the reference's template names and forwards to these calls.
 * Paths are relative to the reference repository root (pvcStillinGradSchool/miniCFD @ e14f896).
 *
 * Conventions: plain pointers and sizes only; every call returns 0 on success or a negative
 * mcfd_status, never throws; mcfd_last_error() gives the message.  All calls for one context come
 * from one host thread (the reference is single-threaded per MPI rank).  Kernels run on the stream
 * given by mcfd_set_stream (default: the legacy default stream).  There is NO CPU fallback: every
 * compute entry point fails with MCFD_ERR_CUDA when no usable device is present.
 *
 * Column layout = the reference's: column[cell_id * kFields + node * K + component], node index
 * ijk = (i*Q + j)*Q + k (include/mini/mesh/part.hpp:755-759,784-787;
 * include/mini/polynomial/hexahedron.hpp:662-677; include/mini/basis/lagrange.hpp:239-253), K = 5,
 * kFields = 5*(P+1)^3.  On the device the state lives in coefficient-major tiles of 32 cells
 * (DESIGN.md section 3); the entry points convert.
 */
#ifndef MINICFD_B200_H_
#define MINICFD_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct mcfd_ctx mcfd_ctx;

typedef enum {
  MCFD_OK = 0,
  MCFD_ERR_ARG = -1,     /* invalid argument / unsupported configuration */
  MCFD_ERR_CUDA = -2,    /* CUDA runtime failure, including "no device" */
  MCFD_ERR_MESH = -3,    /* inconsistent mesh description */
  MCFD_ERR_STATE = -4    /* call sequence error (e.g. no mesh yet) */
} mcfd_status;

/* Spatial scheme: include/mini/spatial/fr/lobatto.hpp:36-434 (Huynh-lumped FR on Lobatto points)
 * or include/mini/spatial/dg/lobatto.hpp:32-303 (DGSEM). */
/* MCFD_FR_GENERAL: include/mini/spatial/fr/general.hpp:34-641, the Vincent correction function with the
 * given c_next spread over the whole node line (basis/vincent.hpp:18-102). */
/* MCFD_DG_GENERAL: include/mini/spatial/dg/general.hpp:20-200 on cells whose polynomial is the stand-alone modal
 * mini::polynomial::Projection<Scalar, 3, P, 5> (include/mini/polynomial/projection.hpp:67-176) with Gauss-Legendre rules
 * (demo/euler/sourceless.hpp:28-33, 64-66 under DGFEM).  A column then holds K x M modal coefficients per cell
 * (M = 10 at P = 2, 20 at P = 3), column[cell * 5 M + mode * 5 + component] (polynomial/expansion.hpp:83-90). */
typedef enum { MCFD_FR_LOBATTO = 0, MCFD_DG_LOBATTO = 1, MCFD_FR_GENERAL = 2, MCFD_DG_GENERAL = 3 } mcfd_scheme;

/* Riemann solver behind mini::riemann::rotated::Euler<...> (include/mini/riemann/rotated/euler.hpp:
 * 153-159): include/mini/riemann/euler/exact.hpp:44-46, hllc.hpp:30-44, ausm.hpp:30-33. */
typedef enum { MCFD_EXACT = 0, MCFD_HLLC = 1, MCFD_AUSM = 2 } mcfd_riemann;

/* Boundary kinds: the seven Set* methods of include/mini/spatial/fem.hpp:196-222; fluxes from
 * include/mini/riemann/rotated/euler.hpp:144-174, 181-236, 260-280, 296-299.  MCFD_BC_NO_SLIP_WALL is
 * a no-op for inviscid Riemann types (fem.hpp:382-383), as in the reference. */
typedef enum {
  MCFD_BC_INVISCID_WALL = 0,
  MCFD_BC_SUPERSONIC_OUTLET = 1,
  MCFD_BC_SUPERSONIC_INLET = 2,
  MCFD_BC_SUBSONIC_INLET = 3,
  MCFD_BC_SUBSONIC_OUTLET = 4,
  MCFD_BC_SMART = 5,
  MCFD_BC_NO_SLIP_WALL = 6
} mcfd_bc_kind;

/* Diffusive part of the Riemann type.  MCFD_NAVIER_STOKES_DDG = mini::riemann::ConvectionDiffusion<
 * rotated::Euler<..>, diffusive::Direct<diffusive::NavierStokes<Gas>>> as composed in
 * demo/laminar/shockless.hpp:43-57 (include/mini/riemann/concept.hpp:93-121, diffusive/navier_stokes.hpp,
 * diffusive/direct.hpp), on expansions with kLocal == false. */
/* MCFD_CELL_VISCOSITY_DDG = ConvectionDiffusion<rotated::Euler<..>, diffusive::Direct<diffusive::Isotropic<Scalar, 5>>>
 * (include/mini/riemann/diffusive/linear.hpp:103-121; demo/euler/sourceless.hpp:88-93 composes it for the VISCOSITY
 * build), generalised to one viscosity per component and cell, which is the flux EnergyBasedViscosity adds
 * (include/mini/spatial/viscosity.hpp:68-71, 114-129): F_k -= nu_k grad u_k. */
/* MCFD_NAVIER_STOKES_CELL_VISCOSITY_DDG = both: EnergyBasedViscosity over a diffusive base (demo/laminar/shocked.hpp:
 * 100-111; viscosity.hpp:76-129): the base's Navier-Stokes flux with the properties of mcfd_set_diffusion, then
 * F_k -= nu_k grad u_k with the per-cell viscosities (zero until set or updated); on no-slip walls the base's wall flux,
 * then the base's and the artificial normal fluxes of the cell's own gradient (viscosity.hpp:132-139). */
typedef enum { MCFD_INVISCID = 0, MCFD_NAVIER_STOKES_DDG = 1, MCFD_CELL_VISCOSITY_DDG = 2,
               MCFD_NAVIER_STOKES_CELL_VISCOSITY_DDG = 3 } mcfd_diffusion;

typedef struct {
  int32_t scheme;    /* mcfd_scheme */
  int32_t degree;    /* P: 2 or 3 (template parameter kDegrees, demo/euler/sourceless.hpp:27) */
  int32_t riemann;   /* mcfd_riemann */
  int32_t device;    /* CUDA device ordinal, -1 = current device */
  double fr_c_next;  /* MCFD_FR_GENERAL only: the `c_next` of fr::General(part, c_next), fr/general.hpp:218-219 */
  int32_t diffusion; /* mcfd_diffusion; the viscous path exists for MCFD_FR_LOBATTO */
  int32_t k_local;   /* 0 | 1: the kLocal flag of mini::polynomial::Hexahedron<Gx, Gy, Gz, K, kLocal>
                      * (include/mini/polynomial/hexahedron.hpp:41-56, 338-343).  With 1 a column holds det(J) u
                      * instead of u — the reference's laminar drivers use that for FR
                      * (demo/laminar/shockless.hpp:62-69) — gradients and Hessians are those of that interpolant
                      * (hexahedron.hpp:397-408, 526-562) and residual columns are d(det u)/dt (fr/general.hpp:251-253).
                      * FR schemes, Euler and Navier-Stokes; on the device and in the halo buffers the state stays u. */
} mcfd_config;

/* -- life cycle ------------------------------------------------------------------------------ */

/* Replaces: `Spatial spatial(&part)` (demo/euler/sourceless.cpp:120) as far as device resources
 * go.  gamma is fixed to 1.4 like `IdealGas<Scalar, 1.4>` in demo/euler/sourceless.hpp:21. */
int mcfd_create(const mcfd_config *config, mcfd_ctx **ctx_out);
int mcfd_destroy(mcfd_ctx *ctx);
const char *mcfd_last_error(const mcfd_ctx *ctx);   /* ctx may be NULL: last create() error */
/* Kernels and copies of this context run on `cuda_stream` (a cudaStream_t passed as void*). */
int mcfd_set_stream(mcfd_ctx *ctx, void *cuda_stream);

/* -- Part / Cell / Face storage (include/mini/mesh/part.hpp:97-291, 404-445, 932-1019, 1456-1549) */

/* Describe one part.  Cells [0,n_local) are local in the part's own id order (inner cells first,
 * then inter cells, part.hpp:746-767), cells [n_local, n_local+n_ghost) are ghosts (part.hpp:768-
 * 775).  cell_nodes: [n_cells][8] node ids in CGNS HEXA_8 order
 * (include/mini/coordinate/hexahedron.hpp:378-389).  Interior faces are found by shared node sets
 * (part.hpp:939-966); bface_nodes [n_bfaces][4] name the boundary quads, bface_tags their section.
 * periodic_pairs [n_pairs][2][4] optionally identifies quads (vertex m of the first with vertex m of
 * the second) — no counterpart in the reference, which has no periodic boundary. */
int mcfd_set_mesh(mcfd_ctx *ctx, int64_t n_nodes, const double *xyz,
                  int64_t n_local, int64_t n_ghost, const int64_t *cell_nodes,
                  int64_t n_bfaces, const int64_t *bface_nodes, const int32_t *bface_tags,
                  int64_t n_pairs, const int64_t *periodic_pairs);

/* Optional, before mcfd_set_mesh: the ids of the cells [0, n_local + n_ghost) in the whole (shuffled) mesh, i.e.
 * Cell::metis_id (part.hpp:190).  The reference makes the cell with the smaller such id the holder of a face between two
 * local cells (include/mini/mesh/shuffler.hpp:417-420 writes the pairs, part.hpp:932-966 reads them); without the ids the
 * part-local id decides.  The holder only matters where a face's tangent frame does: mcfd_limiter_weno_eigen. */
int mcfd_set_cell_global_ids(mcfd_ctx *ctx, int64_t n_cells, const int64_t *ids);

/* Connectivity as the reference numbers it.  kind: 0 local, 1 ghost, 2 boundary faces
 * (part.hpp:115).  Any output pointer may be NULL.  holder/sharer are cell ids (sharer = -1 on
 * boundaries), holder_face/sharer_face local face ids 0..5, orientation the vertex correspondence
 * code.  Returns the number of faces of that kind. */
int64_t mcfd_get_faces(const mcfd_ctx *ctx, int kind, int64_t *holder, int64_t *sharer,
                       int32_t *holder_face, int32_t *sharer_face, int32_t *orientation);
/* Cell::length() and Cell::volume() (part.hpp:230, 239-244) of the local cells. */
int mcfd_get_cell_geometry(const mcfd_ctx *ctx, double *length, double *volume);

/* Halo lists: Part::send_cell_ptrs_/recv_cell_ptrs_ (part.hpp:904-931, 1435-1438).  `cells` are
 * cell ids in pack order; the wire format per cell is the reference's kFields doubles
 * (include/mini/polynomial/expansion.hpp:83-90). */
int mcfd_set_halo(mcfd_ctx *ctx, int32_t n_send_parts, const int64_t *send_counts,
                  const int64_t *send_cells, int32_t n_recv_parts, const int64_t *recv_counts,
                  const int64_t *recv_cells);
/* Device buffers of the packed halo (all neighbours back to back, in list order). */
int mcfd_halo_buffers(mcfd_ctx *ctx, double **send_dev, int64_t *send_doubles,
                      double **recv_dev, int64_t *recv_doubles);
/* Part::ShareGhostCellCoeffs (part.hpp:1273-1278) up to, not including, MPI_Isend: pack the
 * current solution of the send cells into the send buffer. */
int mcfd_halo_pack(mcfd_ctx *ctx);
/* Part::UpdateGhostCellCoeffs (part.hpp:1309-1314) after MPI_Waitall: unpack the recv buffer into
 * the ghost cells of the current solution. */
int mcfd_halo_unpack(mcfd_ctx *ctx);

/* -- the same exchange between parts on the GPUs of one node ------------------------------------------------------
 * Part::ShareGhostCellCoeffs / UpdateGhostCellCoeffs (part.hpp:1242-1314) and the MPI_Allreduce(MIN) of the time step
 * (demo/euler/sourceless.cpp:194) for parts placed one per GPU of an NVLink / NVSwitch node, without a message library
 * and without the host in the loop (minicfd_b200/csrc/peer.cuh): the send cells' rows are packed into a send stage —
 * by the stage kernel itself where it can —, the copy engines take them to the neighbours' receive stages through the
 * peer mapping, an epoch word per neighbour tells when they have arrived, and the receiving stage kernel polls that
 * word itself.  Set-up, after mcfd_set_mesh and mcfd_set_halo on every rank:
 *   1. mcfd_peer_export allocates the stages, words and time-step slots and returns three buffers — the epoch words,
 *      the time-step slots, the receive stage, in this order — as plain device pointers (contexts driven by one
 *      process; different GPUs need cudaDeviceEnablePeerAccess) and / or CUDA IPC handles (one process per GPU);
 *   2. mcfd_peer_open turns a handle received from another process into a pointer valid on this context's device;
 *   3. mcfd_peer_connect, per neighbour i (the order of mcfd_set_halo): neighbour_flags[i] = its epoch words,
 *      slot_at_neighbour[i] = this part's index in THAT part's neighbour list, neighbour_recv[i] = its receive stage,
 *      neighbour_recv_items[i] = the number of cells in its recv lists (all its neighbours), offset_at_neighbour[i] =
 *      where this part's cells start among them; world_flags[r], world_dt_slots[r] = the epoch words and time-step
 *      slots of every rank r, this one included.  The two parts' lists must name the same cells in the same order.
 * mcfd_peer_push delivers the send cells of the state the next residual reads (the current solution, or the stage input
 * inside an open RK step); mcfd_peer_wait holds the stream until every neighbour's delivery of the same count has
 * arrived and unpacks it into the ghost rows.  Every rank must push equally often, and again after every
 * mcfd_set_solution*; push and wait alternate (the same state may be pushed again after its wait — the limiter's
 * unlimited, then limited coefficients: the receive stages keep a spare slot for that).  mcfd_peer_min_time_step replaces the value in dt_dev (NULL: the context's own,
 * mcfd_time_step_device) by the minimum over all ranks.  mcfd_peer_step is one whole GetTimeStep + MPI_Allreduce(MIN) +
 * RungeKutta<order>::Update (sourceless.cpp:191-196, rk.hpp:16-56) for the inviscid nodal schemes; the deliveries of
 * the stage results run on a stream of the context's own beside the next stage.  All counts live on the device, so
 * captured steps can be replayed as a CUDA graph: capture three steps (the buffers rotate with period three) after at
 * least one step has run, end the capture with mcfd_peer_join (the side stream rejoins the context's), and have every
 * rank replay equally often.  Nothing synchronises with the host; mcfd_peer_check does, and reports a wait that gave
 * up (about ten seconds of polling).  While a step is enqueued do not call anything that synchronises the device
 * implicitly (cudaMalloc, cudaFree, the blocking getters of this library) on a rank whose neighbours have not
 * enqueued theirs: the kernels that poll would wait for work that never gets enqueued. */
typedef struct { unsigned char bytes[64]; } mcfd_ipc_handle;   /* cudaIpcMemHandle_t */
int mcfd_peer_export(mcfd_ctx *ctx, int32_t world, mcfd_ipc_handle *handles /* [3] or NULL */, void **pointers /* [3] or NULL */);
int mcfd_peer_open(mcfd_ctx *ctx, const mcfd_ipc_handle *handle, void **pointer);
int mcfd_peer_connect(mcfd_ctx *ctx, int32_t rank, int32_t world, int32_t n_neighbours, void *const *neighbour_flags,
                      const int32_t *slot_at_neighbour, void *const *neighbour_recv, const int64_t *neighbour_recv_items,
                      const int64_t *offset_at_neighbour, void *const *world_flags, void *const *world_dt_slots);
int mcfd_peer_push(mcfd_ctx *ctx);
int mcfd_peer_wait(mcfd_ctx *ctx);
int mcfd_peer_min_time_step(mcfd_ctx *ctx, double *dt_dev);
int mcfd_peer_step(mcfd_ctx *ctx, int order, double t, double dt_guess);
/* The same for several contexts driven by one host thread (one per GPU, or several on one GPU): they advance in
 * lockstep, every delivery enqueued before any wait that depends on it.  With contexts of ONE process use this call,
 * not mcfd_peer_step on each in turn: a context's step polls for deliveries that the next context's call has yet to
 * enqueue. */
int mcfd_peer_step_group(mcfd_ctx *const *ctxs, int32_t n, int order, double t, double dt_guess);
int mcfd_peer_join(mcfd_ctx *ctx);
int mcfd_peer_check(mcfd_ctx *ctx);

/* -- boundary conditions (fem.hpp:71-74, 196-222) -------------------------------------------- */

int mcfd_set_boundary(mcfd_ctx *ctx, int32_t tag, int32_t bc_kind);
/* Number of boundary faces carrying `tag`, in the order of the boundary face table. */
int64_t mcfd_count_boundary_faces(const mcfd_ctx *ctx, int32_t tag);
/* Flux-point coordinates [face][Q*Q][3] at which the reference evaluates its
 * std::function<Value(Global const&, double)> (fr/lobatto.hpp:350, 373, 396, 419). */
int mcfd_get_boundary_points(const mcfd_ctx *ctx, int32_t tag, double *xyz);
/* The pre-sampled values of that function, [face][Q*Q][5], host pointer. */
int mcfd_set_boundary_values(mcfd_ctx *ctx, int32_t tag, const double *values);
/* Called with the stage time before every residual evaluation inside mcfd_rk_step, so that
 * time-dependent boundary functions can be re-sampled (fem.hpp:225-227 SetTime). */
typedef void (*mcfd_time_callback)(void *user, double t);
int mcfd_set_time_callback(mcfd_ctx *ctx, mcfd_time_callback fn, void *user);

/* -- diffusive part (MCFD_NAVIER_STOKES_DDG contexts) -------------------------------------------- */

/* Replaces the class-static setters `Riemann::Diffusion::SetBetaValues(beta_0, beta_1)`
 * (include/mini/riemann/diffusive/direct.hpp:80-83) and `Riemann::Diffusion::SetProperty(nu, prandtl)`
 * (include/mini/riemann/diffusive/navier_stokes.hpp:57-60) as demo/laminar/cavity.cpp:50-51 calls them;
 * per context here.  Defaults: beta = (2, 1/12), nu = 0.  With it the residual becomes the one of
 * fr::Lobatto on ConvectionDiffusion: viscous volume flux (fem.hpp:397-404, navier_stokes.hpp:150-175),
 * direct-DG common gradient and viscous face flux (fr/general.hpp:316-350, direct.hpp:45-78,
 * navier_stokes.hpp:193-202) from the nodal gradients and Hessians of both cells
 * (polynomial/hexahedron.hpp:479-524, 586-594), with the distance of fem.hpp:83-85; MCFD_BC_NO_SLIP_WALL
 * becomes fr/general.hpp:460-479 + navier_stokes.hpp:204-224, its boundary values being
 * (unused, wall velocity u v w, normal temperature gradient) as in demo/laminar/cavity.cpp:30-38. */
int mcfd_set_diffusion(mcfd_ctx *ctx, double beta_0, double beta_1, double nu, double prandtl);
/* MCFD_CELL_VISCOSITY_DDG contexts: mcfd_set_diffusion's nu is Isotropic::SetProperty(nu) (linear.hpp:118-120), one
 * value for every cell and component (prandtl unused).  mcfd_set_cell_viscosity replaces
 * EnergyBasedViscosity::properties_ (viscosity.hpp:57, 68-71, 564-566): nu [n_local + n_ghost][5], constant inside a
 * cell; a face uses the mean of its two cells (fr/general.hpp:331-335).  The ghost cells' values are the caller's to
 * exchange (viscosity.hpp:163-189; mcfd_viscosity_halo_* below).  The rule that CHOOSES the values (UpdateProperties,
 * viscosity.hpp:556-572, with the damping matrices of :299-384) is mcfd_update_cell_viscosity below. */
int mcfd_set_cell_viscosity(mcfd_ctx *ctx, int64_t n_cells, const double *nu);
int mcfd_get_cell_viscosity(mcfd_ctx *ctx, double *nu);   /* [n_local + n_ghost][5] */

/* EnergyBasedViscosity (include/mini/spatial/viscosity.hpp) + WithViscosity (with_viscosity.hpp:20-73) on
 * MCFD_CELL_VISCOSITY_DDG contexts (the reference keeps this state in class statics: one spatial object per process;
 * here it is per context).  On a part with ghost cells UpdateProperties reads whatever the ghost cells hold at the time
 * of the call — the reference calls it from SetSolutionColumn, i.e. before the new state has been exchanged.
 * mcfd_build_damping_matrices = EnergyBasedViscosity::Initialize / BuildDampingMatrices (viscosity.hpp:299-384,
 * 549-554): per cell the N x N matrix of the viscous operator (nu = 1, zero on the far side of every face) applied to
 * the unit nodal vectors, rows scaled by the quadrature weights; built from N probe launches of the trace and stage
 * kernels with the beta values of mcfd_set_diffusion.  From then on mcfd_time_step[_device] is WithViscosity::
 * GetTimeStep: CFL * min_dt_ of the last update (with_viscosity.hpp:67-71).
 * mcfd_update_cell_viscosity = EnergyBasedViscosity::UpdateProperties (viscosity.hpp:556-572) on the state the next
 * residual reads: face-jump integrals (IntegrateJumpOnFaces, :446-475), then per cell and component
 * nu = clamp(jump / (-u^T D u * h / lambda_max * time_scale)) or 0 below the reference value (GetViscosityValues,
 * :477-544).  WithViscosity calls it after every SetSolutionColumn (with_viscosity.hpp:52-55). */
/* EnergyBasedViscosity::ShareGhostCellProperties / UpdateGhostCellProperties (viscosity.hpp:163-189; called by
 * WithViscosity::GetResidualColumn before every residual, with_viscosity.hpp:45-50): 5 doubles per cell, the send cells'
 * viscosities packed into the send buffer / the receive buffer unpacked into the ghost cells, in the order of
 * mcfd_set_halo.  As with the state, moving the bytes between the parts is the host's job. */
int mcfd_viscosity_halo_buffers(mcfd_ctx *ctx, double **send_dev, int64_t *send_doubles, double **recv_dev,
                                int64_t *recv_doubles);
int mcfd_viscosity_halo_pack(mcfd_ctx *ctx);
int mcfd_viscosity_halo_unpack(mcfd_ctx *ctx);
int mcfd_build_damping_matrices(mcfd_ctx *ctx);
int mcfd_get_damping_matrices(mcfd_ctx *ctx, double *matrices);   /* [n_local][N][N], row-major */
int mcfd_update_cell_viscosity(mcfd_ctx *ctx, double time_scale);

/* -- modal polynomial (MCFD_DG_GENERAL contexts) --------------------------------------------------- */

/* Gauss-Legendre points of the local cells, [n_local][Q^3][3], in the order of integrator::Hexahedron
 * (include/mini/integrator/hexahedron.hpp:66-80): where FiniteElement::Approximate samples its function. */
int mcfd_get_quadrature_points(const mcfd_ctx *ctx, double *xyz);
/* det J at the nodes of the local cells, [n_local][N] (integrator/hexahedron.hpp GetJacobianDeterminant): what a
 * k_local host multiplies nodal values by to obtain column entries (Hexahedron::SetValue, hexahedron.hpp:344-349). */
int mcfd_get_jacobian_determinants(const mcfd_ctx *ctx, double *det);
/* Replaces Cell::Approximate -> Projection::Approximate (include/mini/polynomial/projection.hpp:28-43, 170-173): the L2
 * projection of a function given by its values at those points, values [n_local][Q^3][5], becomes the solution.
 * For MCFD_DG_GENERAL mcfd_get_boundary_points returns the faces' Gauss points, [face][Q*Q][3]. */
int mcfd_project(mcfd_ctx *ctx, const double *values);

/* -- temporal::System interface (include/mini/temporal/ode.hpp:17-48) -------------------------- */

int64_t mcfd_count_local_cells(const mcfd_ctx *ctx);
int64_t mcfd_cell_data_size(const mcfd_ctx *ctx);          /* Part::GetCellDataSize, part.hpp:793 */
/* FiniteElement::SetSolutionColumn / GetSolutionColumn (fem.hpp:228-247); host pointers. */
int mcfd_set_solution(mcfd_ctx *ctx, const double *column);
int mcfd_get_solution(mcfd_ctx *ctx, double *column);
/* Same with device pointers (column layout, kFields*n_local doubles). */
int mcfd_set_solution_device(mcfd_ctx *ctx, const double *column_dev);
int mcfd_get_solution_device(mcfd_ctx *ctx, double *column_dev);
/* FiniteElement::GetResidualColumn (fem.hpp:248-258; fr/general.hpp:249-268; dg/lobatto.hpp:143-
 * 158) at time t for the current solution; ghost cells must be current (mcfd_halo_unpack).
 * column: host pointer or NULL; column_dev: device pointer or NULL. */
int mcfd_residual(mcfd_ctx *ctx, double t, double *column, double *column_dev);

/* -- temporal::RungeKutta (include/mini/temporal/rk.hpp:16-56) --------------------------------- */

/* One full RungeKutta<order>::Update(system, t, dt) for a part without ghosts (or whose ghosts the
 * caller does not need to refresh), fused: each stage is one kernel that evaluates the residual and
 * writes the next stage state.  order = 1, 2 or 3. */
int mcfd_rk_step(mcfd_ctx *ctx, int order, double t, double dt);
/* The same, stage by stage, for parts with ghost cells:
 *   mcfd_rk_begin; for s in 0..order-1 { mcfd_halo_pack; <exchange>; mcfd_rk_stage_inner(s);
 *   mcfd_halo_unpack; mcfd_rk_stage_inter(s); }
 * stage_inner advances the tiles holding only inner cells (no ghost neighbour) and can overlap the
 * exchange, stage_inter the rest — the overlap of fem.hpp:249-256. */
int mcfd_rk_begin(mcfd_ctx *ctx, int order, double t, double dt);
int mcfd_rk_stage_inner(mcfd_ctx *ctx, int stage);
int mcfd_rk_stage_inter(mcfd_ctx *ctx, int stage);
/* Time of the residual evaluation of `stage` in the step opened by mcfd_rk_begin. */
double mcfd_rk_stage_time(const mcfd_ctx *ctx, int stage);

/* FiniteElement::GetTimeStep (fem.hpp:411-435) incl. the FR-Lobatto doubling
 * (fr/lobatto.hpp:220-222): the part-local value, before MPI_Allreduce(MIN)
 * (demo/euler/sourceless.cpp:191-194). */
int mcfd_time_step(mcfd_ctx *ctx, double dt_guess, int rk_order, double *dt_local);

/* Device-resident variants for time loops that must not wait for the host: the part-local time step is
 * left in device memory (dt_dev, or an internal scalar when NULL), may be min-reduced across ranks there
 * (ncclAllReduce MIN = the MPI_Allreduce of demo/euler/sourceless.cpp:194), and the RK stage kernels read
 * it from there.  Nothing synchronises the stream. */
/* The stage times t + dt and t + dt / 2 (rk.hpp:24-47) are not known to the host then.  A context whose boundary
 * values depend on time (mcfd_set_time_dependent(ctx, 1): its time callback re-samples them per stage) therefore
 * refuses mcfd_rk_begin_device / mcfd_rk_step_device / mcfd_peer_step for orders above one with MCFD_ERR_STATE;
 * read the step back with mcfd_time_step and use mcfd_rk_step.  Without that flag the callback is still called,
 * with the time of the step's beginning for every stage. */
int mcfd_set_time_dependent(mcfd_ctx *ctx, int flag);
int mcfd_time_step_device(mcfd_ctx *ctx, double dt_guess, int rk_order, double *dt_dev);
int mcfd_rk_begin_device(mcfd_ctx *ctx, int order, double t, const double *dt_dev);
int mcfd_rk_step_device(mcfd_ctx *ctx, int order, double t, const double *dt_dev);

/* The same time step as mcfd_time_step, for callers that may not allocate device memory while they run
 * (a host that gave the whole GPU to its own memory pool, a rank held to a fixed memory budget, a time
 * loop that must not meet the device-wide synchronisation hidden in cudaMalloc/cudaFree).  No counterpart
 * in the reference: its GetTimeStep (fem.hpp:417-435) runs on host memory it already owns.
 * It needs no context and allocates no device memory, pinned memory or stream (only a failing call touches the host
 * heap, for its error message): inputs and output are the caller's device memory —
 * column_dev: the solution column [n_cells][N][5] in the reference layout of expansions with kLocal == false, i.e.
 * the nodal values u themselves (a kLocal == true column holds det(J) u and would give a wrong wave speed: divide
 * first); length_dev: Cell::length() of
 * each cell (mcfd_get_cell_geometry); dt_dev: one double receiving the part-local time step.  The
 * reduction's scratch is a fixed 1 MiB array inside the library image, held by every GPU that loads
 * the library until the process ends; it limits a call to 1 048 576 cells and larger n_cells are refused
 * with MCFD_ERR_ARG (message via mcfd_last_error(NULL)).  The first call on each device looks up the
 * array's device address (a cudaGetSymbolAddress, no allocation) with that device current: make that
 * first call outside any stream capture.  All calls share the array: nothing in it has to be cleared
 * beforehand (every entry a call reads it has written first), but two calls that overlap on the same GPU
 * (different streams or threads) race on it — serialise them.  Asynchronous on cuda_stream. */
int mcfd_time_step_static(int32_t scheme, int32_t degree, int64_t n_cells, const double *column_dev,
                          const double *length_dev, double dt_guess, int32_t rk_order, double *dt_dev,
                          void *cuda_stream);

/* Frame output, the step after the path: what mini::mesh::vtk::Writer<Part>::Prepare (include/mini/mesh/vtk.hpp:421-494)
 * collects for WriteSolutions (:526-668).  Every local hexahedron becomes one VTK tri-quadratic hexahedron (type 29,
 * vtk.hpp:400-409) of 27 points in VTK's order (vtk.hpp:305-325); points [n_local][27][3] are
 * coordinate().LocalToGlobal(local), values [n_local][27][5] the current solution at those local coordinates:
 * modal == 0: Hexahedron::LocalToValue (polynomial/hexahedron.hpp:256-264; on kLocal contexts the interpolated
 * det(J) u divided by det J at the point); modal != 0: through the L2 projection on the cell's orthonormal basis, as
 * Extrapolation<Hexahedron> and Projection evaluate it (extrapolation.hpp:180-183, projection.hpp:128-132;
 * MCFD_DG_GENERAL contexts always).  Either pointer may be null (the points do not change between frames).  The
 * values are sampled by one kernel over the device-resident state; formatting the .vtu/.pvtu text is the host's job
 * (include/mini/mesh/vtk.hpp of this repository, minicfd_b200/vtk.py). */
int mcfd_sample_vtk(mcfd_ctx *ctx, int32_t modal, double *points, double *values);

/* The text of a frame's arrays, built on the GPU.  WriteSolutions prints every number through `vtu << v << ' '`
 * (vtk.hpp:600-602, 610-612, 622-624; the stream's default format = printf's "%g", six significant digits correctly
 * rounded) and the integer arrays through `vtu << i << ' '` (vtk.hpp:650-665); at 10 M numbers per frame that costs the
 * host far more than the time steps between two frames.  These calls return exactly those characters (no terminating
 * zero, *length bytes) into text[capacity]; 16 bytes per value always suffice; on MCFD_ERR_ARG for a short buffer
 * *length holds the size needed.  Values whose sixth digit the device cannot decide (a scaled fraction within 1e-6 of
 * one half, |v| outside 1e-294..1e+304) are formatted by the host's snprintf inside the call.
 * mcfd_vtk_component_text: component k of the samples of the last mcfd_sample_vtk (which may be called with
 * values == NULL to keep them on the device), point after point, cell after cell.
 * mcfd_format_doubles: n host values (extra point / cell data).  mcfd_format_integers: first + i * step, i < n
 * (connectivity: 0, 1; offsets: 27, 27; types: 29, 0). */
int mcfd_vtk_component_text(mcfd_ctx *ctx, int32_t component, char *text, int64_t capacity, int64_t *length);
int mcfd_format_doubles(mcfd_ctx *ctx, int64_t n, const double *values, char *text, int64_t capacity, int64_t *length);
int mcfd_format_integers(mcfd_ctx *ctx, int64_t n, int64_t first, int64_t step, char *text, int64_t capacity,
                         int64_t *length);

/* The step before the path: the deterministic index work of mini::mesh::Shuffler that turns a cell partition (METIS's
 * output, which stays with the host) into the part-contiguous numbering the solver reads.  Context-free: `device` is
 * the CUDA ordinal, all arrays are host arrays, results are bit-exact.
 * mcfd_get_node_parts = metis::GetNodeParts (include/mini/mesh/metis.hpp:314-335): a node belongs to the smallest part
 *   among the cells that use it; range[n_cells + 1] / nodes[] are the mesh's CSR lists (0-based), untouched nodes get
 *   n_parts.
 * mcfd_get_new_order = GetNewOrder (include/mini/mesh/shuffler.hpp:35-49): indices sorted by (part, index);
 *   new_to_old[n], old_to_new[n].
 * mcfd_shuffle_data = ShuffleData (shuffler.hpp:57-64) on rows of `width` doubles, in place.
 * mcfd_shuffle_connectivity = ShuffleConnectivity (shuffler.hpp:85-105): 1-based node ids renamed through
 *   old_to_new_for_nodes[n_nodes], then the cells' rows of `npe` ids reordered by new_to_old_for_cells, in place. */
int mcfd_get_node_parts(int32_t device, int64_t n_nodes, int64_t n_cells, const int64_t *range, const int64_t *nodes,
                        const int64_t *cell_parts, int64_t n_parts, int64_t *node_parts);
int mcfd_get_new_order(int32_t device, int64_t n, const int64_t *parts, int64_t *new_to_old, int64_t *old_to_new);
int mcfd_shuffle_data(int32_t device, int64_t n, int64_t width, const int64_t *new_to_old, double *data);
int mcfd_shuffle_connectivity(int32_t device, int64_t n_cells, int32_t npe, int64_t n_nodes,
                              const int64_t *old_to_new_for_nodes, const int64_t *new_to_old_for_cells,
                              int64_t *cell_nodes);

/* Initial-condition clean-up on the current solution of the local cells: kind 0 =
 * mini::limiter::average::Reconstruct (include/mini/limiter/average.hpp:11-28), kind 1 =
 * mini::limiter::majority::Reconstruct (include/mini/limiter/majority.hpp:14-42), as called by
 * demo/euler/sourceless.cpp:146-153.  Ghost cells are not touched: exchange afterwards. */
int mcfd_limiter_reconstruct(mcfd_ctx *ctx, int kind);

/* -- diagnostics --------------------------------------------------------------------------------- */

/* Number of kernels this context has launched so far (bench.py's gpu_launches). */
/* Replaces mini::limiter::Reconstruct(part, &limiter) with limiter = mini::limiter::weno::Lazy<Cell>(w0, eps)
 * (include/mini/limiter/reconstruct.hpp:21-54, include/mini/limiter/weno.hpp:97-195) on cells whose polynomial is
 * mini::polynomial::Extrapolation<Hexahedron> (include/mini/polynomial/extrapolation.hpp:20-188): the nodal
 * values are projected on the cell's orthonormal basis (include/mini/basis/linear.hpp:87-141,
 * include/mini/polynomial/projection.hpp:28-51), every neighbour's polynomial is re-projected on that basis and
 * shifted to the cell's mean, the candidates are blended with weights w / (beta + eps)^2 from the Taylor-derivative
 * smoothness indicators (weno.hpp:25-95), and the nodal values are re-sampled from the blend.  Acts on the state the
 * next residual would read (the current solution, or the newest stage inside an open RK step) — the reference's
 * WithLimiter does the same after every SetSolutionColumn (include/mini/spatial/with_limiter.hpp:57-68).
 * phase 0 = inner cells (they need no ghost data), phase 1 = inter cells (call after mcfd_halo_unpack of the
 * UNLIMITED state, exactly like reconstruct.hpp:46-51; they borrow the already limited inner cells), -1 = both. */
int mcfd_limiter_weno_lazy(mcfd_ctx *ctx, double w0, double eps, int phase);
/* The same with limiter = mini::limiter::weno::Eigen<Cell, Riemann>(w0, eps) (include/mini/limiter/weno.hpp:197-349):
 * only the cells flagged by Eigen::IsNotSmooth (weno.hpp:230-258) are rebuilt, once per adjacent face in the
 * characteristic variables of that face's frame — the Riemann solver of the face's first quadrature point, as
 * installed by InstallRiemannSolvers (demo/euler/rotorcraft.cpp:93-96) — with EigenMatrices of the cell mean
 * (include/mini/riemann/rotated/euler.hpp:330-338, include/mini/riemann/euler/eigen.hpp:27-56), and the per-face
 * results are averaged with the neighbours' volumes (weno.hpp:331-348). */
int mcfd_limiter_weno_eigen(mcfd_ctx *ctx, double w0, double eps, int phase);
/* mini::riemann::euler::EigenMatrices<Gas>(conservative, nu, mu, pi) (include/mini/riemann/euler/eigen.hpp:27-61) for n
 * states: conservative [n][5], frames [n][9] = (nu, mu, pi); left/right [n][5][5] row-major (L and R).  Test hook. */
int mcfd_eigen_matrices(mcfd_ctx *ctx, int64_t n, const double *conservative, const double *frames, double *left,
                        double *right);

int64_t mcfd_launch_count(const mcfd_ctx *ctx);
/* Which kernel a stage of this context launches, as a static string: "NodeStageKernel<affine>" (inviscid lumped Lobatto
 * schemes at p = 2 on a mesh whose local cells are all parallelepipeds up to rounding), "NodeStageKernel<general>" /
 * "LineStageKernel<affine|general>" (the same schemes otherwise, or as MCFD_STAGE_KERNEL=nodes|lines says),
 * "StageKernel", "DgmStageKernel".  Decided by mcfd_set_mesh; "?" before it.  No counterpart in the reference. */
const char *mcfd_stage_kernel(const mcfd_ctx *ctx);
/* Duration in ms of the stage kernels launched since the last call with reset != 0, measured with
 * CUDA events on the context's stream when timing is enabled (bench.py's roofline leg). */
int mcfd_enable_kernel_timing(mcfd_ctx *ctx, int enable);
int mcfd_kernel_time(mcfd_ctx *ctx, double *total_ms, int64_t *n_launches, int reset);
/* One rotated Riemann problem on the device (test hook): frame normal n[3], conservative states,
 * bc_kind < 0 for an interior face. */
int mcfd_riemann_flux(mcfd_ctx *ctx, int32_t bc_kind, int64_t n, const double *normals,
                      const double *u_left, const double *u_right, double *flux);

#ifdef __cplusplus
}
#endif
#endif  /* MINICFD_B200_H_ */
