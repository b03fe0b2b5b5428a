/* blitzdg_b200 -- C ABI of the B200-native sw2d nodal-DG RHS + LSERK4 hot path.
 *
 * The reference (WQCG/blitzdg) has no FFI seam around this path: the RHS is a
 * free function in a driver translation unit,
 *     void sw2d::computeRHS(fields, const numParams&, const physParams&,
 *                           const DGContext2D&, real_type t)
 *                                     (reference src/sw2d/SW2d.hpp:53, src/sw2d/main.cpp:279-484;
 *                                      simple variant src/sw2d-simple/SW2d.hpp:15, main.cpp:181-356)
 * fed by the DGContext2D accessor bag (reference include/DGContext2D.hpp:82-196).
 * This header is the seam a maintainer binds instead: plain
 * C, opaque handle, int status returns, no exceptions across the boundary.
 *
 * Ownership: the caller keeps every host pointer; the library copies at create /
 * set_* time and never retains host pointers.  One handle = one GPU (= one rank).
 * Calls on one handle are not thread-safe; distinct handles are independent.
 *
 * Host array layouts are the reference's native ones:
 *   volume arrays  [Np][K]  row-major, element index fastest  (blitz default storage)
 *   face arrays    [Nfaces*Nfp][K] row-major
 *   vmapM/vmapP    [Nfaces*Nfp*K], element-major then face then face node, values n + Np*k
 *   BC lists       face-node indices i in [0, Nfaces*Nfp*K)       (DGContext2D::bcmap())
 */
#ifndef BLITZDG_B200_H
#define BLITZDG_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct bdg_sw2d bdg_sw2d; /* opaque solver handle */

/* status codes */
#define BDG_OK 0
#define BDG_E_INVALID 1    /* bad argument / inconsistent context                         */
#define BDG_E_CUDA 2       /* CUDA runtime error (message in bdg_last_error)              */
#define BDG_E_UNSTABLE 3   /* the reference's "A numerical instability has occurred!"     */
#define BDG_E_NCCL 4       /* NCCL error                                                  */
#define BDG_E_UNSUPPORTED 5

/* Mirrors DGContext2D (reference include/DGContext2D.hpp:11-39). */
typedef struct {
  int32_t N, Np, Nfp, Nfaces, K;          /* order, nodes/element, nodes/face, faces/element, elements */
  const double *Dr, *Ds;                  /* [Np][Np]                                       */
  const double *Lift;                     /* [Np][Nfaces*Nfp]                               */
  const double *Filt;                     /* [Np][Np] or NULL                               */
  const double *rx, *sx, *ry, *sy;        /* [Np][K]                                        */
  const double *nx, *ny, *Fscale;         /* [Nfaces*Nfp][K]                                */
  const int32_t *vmapM, *vmapP;           /* [Nfaces*Nfp*K]                                 */
  const int32_t *bc_face_nodes[9];        /* BCmap lists by BCTag 1..8 (reference include/BCtypes.hpp:7-16) */
  int32_t bc_count[9];                    /* absent tag = count 0, never an error           */
} bdg_dgcontext2d;

/* physParams + numParams (reference src/sw2d/SW2d.hpp:15-30, src/sw2d/main.cpp:49-60,281-283). */
typedef struct {
  double g;            /* 9.81                                                               */
  double CD;           /* bottom drag, 2.5e-3                                                */
  double f;            /* Coriolis, 1.0070e-4                                                */
  double CFL;          /* 0.25 (sw2d) / 0.65 (sw2d-simple)                                   */
  int32_t lf_mode;     /* 0: per-face Lax-Friedrichs speed (sw2d-simple main.cpp:305-312)
                          1: global max over all face nodes (sw2d main.cpp:413-414)          */
  int32_t sources;     /* 1: star states + bathymetry/drag/Coriolis (sw2d main.cpp:355-368,460-483) */
  int32_t filter;      /* 1: RHS := Filt*RHS (sw2d-simple main.cpp:134-136)                  */
  int32_t reserved;
  double tide_amp;     /* 3.0                       (sw2d main.cpp:283)                      */
  double tide_period;  /* 3600*12.42                (sw2d main.cpp:281)                      */
  double tide_ramp;    /* 0.15/3600                 (sw2d main.cpp:352)                      */
} bdg_sw2d_params;

/* Halo description of one rank of an element-partitioned mesh (the reference
 * only computes MeshManager::partitionMesh, src/MeshManager.cpp:491-544, and never consumes it).
 * Local vmapP entries >= Np*K address ghost slots: ghost g lives at index Np*K + g.
 * Ghost slots are grouped by neighbour rank: recv_offset[i]..recv_offset[i+1].
 * send_nodes[send_offset[i]..send_offset[i+1]) are LOCAL volume node ids whose values the
 * neighbour nbr_rank[i] expects, in that neighbour's ghost order. */
typedef struct {
  int32_t rank, nranks;
  int32_t n_ghost;
  int32_t n_neighbors;
  const int32_t *nbr_rank;     /* [n_neighbors]   */
  const int32_t *send_offset;  /* [n_neighbors+1] */
  const int32_t *send_nodes;   /* [send_offset[n_neighbors]] */
  const int32_t *recv_offset;  /* [n_neighbors+1] */
  int32_t n_interior;          /* elements [0,n_interior) touch no ghost slot (0 = unknown) */
} bdg_halo;

/* --- lifetime ---------------------------------------------------------------------------- */
/* halo may be NULL (single GPU). device = CUDA ordinal. */
int bdg_sw2d_create(bdg_sw2d **out, const bdg_dgcontext2d *ctx, const bdg_sw2d_params *prm,
                    const bdg_halo *halo, int device);
void bdg_sw2d_destroy(bdg_sw2d *s);
const char *bdg_last_error(const bdg_sw2d *s); /* s may be NULL: error of the last failed create */

/* --- state (host arrays [Np][K]) ---------------------------------------------------------- */
/* fields H, Hx, Hy of sw2d::fields (reference src/sw2d/SW2d.hpp:32-49); NULL H => flat bottom unused */
int bdg_sw2d_set_bathymetry(bdg_sw2d *s, const double *H, const double *Hx, const double *Hy);
/* sets h, hu, hv and the time; zeroes the RK residuals (reference src/sw2d/main.cpp:117-119) */
int bdg_sw2d_set_state(bdg_sw2d *s, const double *h, const double *hu, const double *hv, double t);
int bdg_sw2d_get_state(bdg_sw2d *s, double *h, double *hu, double *hv, double *t);

/* Streaming variants for batches of independent states (ensemble members, parameter sweeps): the upload of the next
 * state and the download of the previous result use both directions of the PCIe link while the steps of the
 * current state run.  stage_state starts the upload (host arrays pinned and valid until the commit has executed);
 * commit_state is a stream-ordered set_state of the staged arrays; get_state_async snapshots the live state behind
 * everything enqueued so far and copies it out on a side stream; get_state_wait blocks until it has arrived.
 *     stage(0); for i: commit(); stage(i+1); step(...); get_state_async(out[i%2]); ...; get_state_wait();          */
int bdg_sw2d_stage_state(bdg_sw2d *s, const double *h, const double *hu, const double *hv, double t);
int bdg_sw2d_commit_state(bdg_sw2d *s);
int bdg_sw2d_get_state_async(bdg_sw2d *s, double *h, double *hu, double *hv);
int bdg_sw2d_get_state_wait(bdg_sw2d *s);

/* --- the hot path ------------------------------------------------------------------------- */
/* one evaluation of sw2d::computeRHS on the current state at the current time; host out [Np][K] */
int bdg_sw2d_rhs(bdg_sw2d *s, double *RHS1, double *RHS2, double *RHS3);
/* sw2d::computeTimeStep (reference src/sw2d/main.cpp:253-277); global over ranks when a comm is set */
int bdg_sw2d_timestep(bdg_sw2d *s, double *dt_out);
/* nsteps LSERK4 steps of 5 fused RHS+stage launches, all on device (reference
 * include/LSERK4.hpp:16-27 + src/advec1d/main.cpp:92-102).  dt <= 0: adaptive, dt recomputed on
 * device each step.  Asynchronous: returns after enqueueing; bdg_sw2d_sync waits. */
int bdg_sw2d_step(bdg_sw2d *s, int32_t nsteps, double dt);
int bdg_sw2d_sync(bdg_sw2d *s);
/* The rest of the src/sw2d driver's time loop.
 * set_sponge: spongeCoeff [Np][K] (reference src/sw2d/main.cpp:184-186, sw2d::buildSpongeCoeff :517-553); NULL = none.
 * step_ssprk2: nsteps steps of the loop at reference src/sw2d/main.cpp:192-244 -- dt = computeTimeStep (dt <= 0) or
 * the caller's, Heun's SSP-RK2 with both RHS evaluations at the step-start time (:213-232), the relaxation
 * hu /= 1 + sponge*hu*hu, hv likewise, after each stage (:222-224, :235-236), t += dt.  Runs on the two-launch
 * path (any element shape, any BC mix), all on device, asynchronous like bdg_sw2d_step. */
int bdg_sw2d_set_sponge(bdg_sw2d *s, const double *sponge);
int bdg_sw2d_step_ssprk2(bdg_sw2d *s, int32_t nsteps, double dt);
/* blow-up guard of the drivers (reference src/sw2d/main.cpp:238-240): max|h-H| > 1e8 or NaN
 * => BDG_E_UNSTABLE.  eta_max may be NULL. */
int bdg_sw2d_check(bdg_sw2d *s, double *eta_max);

/* positivityPreservingLimiter2D of the Python drivers (reference swhelpers/limiters.py:3-21) applied
 * to the live state in place: per element theta = min(1, mean(h)/(mean(h) - max(min(h),1e-3) + 1e-4)) and
 * q = mean(q) + theta*(q - mean(q)) for h, hu, hv (arithmetic node means).  Stream-ordered like bdg_sw2d_step. */
int bdg_sw2d_limit_positivity(bdg_sw2d *s);

/* --- tracer --------------------------------------------------------------------------------
 * The 4th conserved field hN of the reference's Python drivers (swhelpers/rhs.py:206-297, swhelpers/flux.py:18-20,
 * sw2d.py:30-33,129-140): F4 = hN u, G4 = hN v, the flow's per-face Lax-Friedrichs speed, hN+ = hN- on walls.
 * set_tracer: hN [Np][K] (copied; zeroes its RK residual); NULL removes the tracer.  While a tracer is set,
 * bdg_sw2d_step / bdg_sw2d_step_ssprk2 advance it with the flow (one extra launch per stage).  Needs lf_mode = 0
 * (the only form the reference has) and a single rank; otherwise BDG_E_UNSUPPORTED.
 * rhs_tracer: RHS4 of the current state, host out [Np][K] -- the parity entry point. */
int bdg_sw2d_set_tracer(bdg_sw2d *s, const double *hN);
int bdg_sw2d_get_tracer(bdg_sw2d *s, double *hN);
int bdg_sw2d_rhs_tracer(bdg_sw2d *s, double *RHS4);

/* --- over-integrated RHS on curved triangles -----------------------------------------------------
 * sw2dComputeRHS_curved of the reference's Python drivers (swhelpers/rhs.py:6-176, called by sw2d_curved.py:240-262): the
 * fluxes are evaluated at the cubature points of the elements and at the Gauss nodes of their faces, tested against the
 * nodal basis, and each element's mass matrix is inverted (V V^T / J on straight elements, the stored Cholesky factor on
 * the curved ones), then the nodal sources (Coriolis, drag, bed slope) are added.
 * bdg_curved2d carries what that function reads from its three contexts, in the reference's native layouts:
 *   cub_ctx   (include/CubatureContext2D.hpp, built by TriangleNodesProvisioner::buildCubatureVolumeMesh,
 *              src/TriangleNodesProvisioner.cpp:81-202):  V, Dr, Ds [Ncub][Np]; W, rx, sx, ry, sy [Ncub][K];
 *              MMChol [Np][Np][K] (upper factors; only the curved elements are read)
 *   gauss_ctx (include/GaussFaceContext2D.hpp, buildGaussFaceNodes, src/TriangleNodesProvisioner.cpp:205-381):
 *              Interp [3*NGauss][Np]; W, nx, ny [3*NGauss][K]; mapM, mapP [3*NGauss*K] (possibly made periodic by the
 *              caller, swhelpers/maps.py); BCmap[Wall] as a list of Gauss-node ids
 *   ctx.V [Np][Np], the nodal Jacobian J [Np][K] and the list of curved elements (the function's own arguments).
 * set_curved copies everything (NULL removes it).  rhs_curved evaluates the function once on the device state: h, hu, hv
 * of set_state, hN of set_tracer (zero if none), zx = -Hx, zy = -Hy of set_bathymetry (z = -H; zero if unset), g and f of
 * the parameters, CD = CD_nodal [Np][K] (sw2d_curved.py:166-177 builds a nodal drag coefficient) or, if NULL, the
 * parameters' scalar.  Results in host arrays [Np][K], any may be NULL.  Triangles, single rank. */
typedef struct {
  int32_t Ncub;                                      /* cubature points per element                          */
  int32_t NGauss;                                    /* Gauss nodes per FACE (gauss_ctx.NGauss)              */
  const double *cubV, *cubDr, *cubDs;                /* [Ncub][Np]                                           */
  const double *cubW, *cub_rx, *cub_sx, *cub_ry, *cub_sy; /* [Ncub][K]                                       */
  const double *MMChol;                              /* [Np][Np][K] or NULL when n_curved = 0                */
  const int32_t *curvedEls;                          /* [n_curved] element ids                               */
  int32_t n_curved;
  int32_t n_gmapW;
  const double *gInterp;                             /* [3*NGauss][Np]                                       */
  const double *gW, *gnx, *gny;                      /* [3*NGauss][K]                                        */
  const int32_t *gmapM, *gmapP;                      /* [3*NGauss*K], values g + 3*NGauss*k                  */
  const int32_t *gmapW;                              /* [n_gmapW] gauss_ctx.BCmap[Wall]                      */
  const double *V;                                   /* [Np][Np] ctx.V                                       */
  const double *J;                                   /* [Np][K]                                              */
} bdg_curved2d;
int bdg_sw2d_set_curved(bdg_sw2d *s, const bdg_curved2d *c);
int bdg_sw2d_rhs_curved(bdg_sw2d *s, const double *CD_nodal, double *RHS1, double *RHS2, double *RHS3, double *RHS4);
/* The curved driver's time loop on the device (sw2d_curved.py:240-262), nsteps times:
 *   R = Filt*RHS_curved(q);  q1 = q + dt/2 R;  R = Filt*RHS_curved(q1);  q += dt R;  t += dt
 * for h, hu, hv and the tracer (if set).  Filt = the context's Filt when the parameters' `filter` is 1 (the driver always
 * filters, sw2d_curved.py:244-247), else none.  dt > 0 (the driver's own dt formula, :236, is host arithmetic on u, v).
 * Only enqueues, like bdg_sw2d_step. */
int bdg_sw2d_step_curved(bdg_sw2d *s, int32_t nsteps, double dt, const double *CD_nodal);

/* --- output path ---------------------------------------------------------------------------
 * Every outputInterval steps the drivers derive eta = h - H, u = hu/h, v = hv/h (reference src/sw2d/main.cpp:199-210)
 * and VtkOutputter::writeFieldToFile (reference include/VtkOutputter.hpp:48-101) interpolates each to the equispaced
 * nodes of the sub-element split, resultField = IM * field (splitElements, src/TriangleNodesProvisioner.cpp:1154-1227,
 * src/QuadNodesProvisioner.cpp:721-800).  set_output: interp = that IM, [n_out][Np] row-major (copied).
 * output_begin: enqueues the derivation + the three interpolations of the live state and the copies of the results
 * into eta, u, v (host, [n_out][K] row-major, any may be NULL; pinned memory keeps the copies asynchronous) on a side
 * stream: later bdg_sw2d_step calls only wait for the interpolation kernel, not for the copies.
 * output_wait: blocks until the host arrays of the last output_begin are complete; it only waits on an event and may
 * be called from a second (writer) thread while the owning thread keeps stepping. */
int bdg_sw2d_set_output(bdg_sw2d *s, const double *interp, int32_t n_out);
int bdg_sw2d_output_begin(bdg_sw2d *s, double *eta, double *u, double *v);
int bdg_sw2d_output_wait(bdg_sw2d *s);

/* --- plumbing ----------------------------------------------------------------------------- */
/* kernel variant used by step(): 0 = auto, 1 = generic two-kernel path, 2 = fused DMMA path */
int bdg_sw2d_set_variant(bdg_sw2d *s, int32_t variant);
int bdg_sw2d_get_variant(const bdg_sw2d *s);
/* number of kernel launches enqueued by this handle since create */
int64_t bdg_sw2d_launch_count(const bdg_sw2d *s);
/* the CUDA stream (cudaStream_t) all work of this handle is enqueued on, for event timing */
void *bdg_sw2d_stream(bdg_sw2d *s);
/* algorithmic bytes of one RHS+stage over this handle's elements (B_alg) */
double bdg_sw2d_algorithmic_bytes_per_stage(const bdg_sw2d *s);

/* --- allocation-free entry point ---------------------------------------------------------------------------------
 * No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * The reduction of bdg_sw2d_timestep -- dt = CFL / ((N+1)^2/2 * max over face nodes of Fscale*(|u| + sqrt(g h))) -- on
 * the CALLER's device arrays, in the reference's native layouts: d_h, d_hu, d_hv [Np][K] row-major, d_Fscale
 * [Nfaces*Nfp][K] row-major, d_fmask [Nfaces*Nfp] int32 = local node of each face-node slot (Fmask flattened face by
 * face; vmapM of element 0).  Every d_* pointer is memory of `device`; d_dt_out receives one double.  The kernel is
 * enqueued on `stream` (a cudaStream_t, NULL = the default stream) and the call returns without synchronising.
 * It calls no cudaMalloc, cudaFree or allocating library routine: its scratch (one partial maximum per 4096 face nodes
 * and the ticket word electing the last thread block) is a static __device__ array of 512.25 KB that every GPU loading
 * the library holds until the process exits.  Hence the limit: Nfaces*Nfp*K <= BDG_NOALLOC_MAX_FACE_NODES, beyond it
 * BDG_E_UNSUPPORTED (bdg_sw2d_timestep has no such limit).  All calls on a device share that array, which keeps the
 * last call's contents (the ticket word is cleared on the stream before each launch): calls that overlap on the GPU --
 * different streams, or host threads without ordering -- race on it and must be serialised by the caller.
 * err (may be NULL): receives the message of a failed call, at most err_cap bytes including the terminator. */
#define BDG_NOALLOC_MAX_FACE_NODES (1 << 28)
int bdg_sw2d_timestep_noalloc(int32_t device, int32_t N, int32_t Np, int32_t Nfp, int32_t Nfaces, int32_t K, double g,
                              double CFL, const double *d_h, const double *d_hu, const double *d_hv,
                              const double *d_Fscale, const int32_t *d_fmask, double *d_dt_out, void *stream, char *err,
                              int32_t err_cap);

/* --- multi-GPU (one process per GPU) ------------------------------------------------------- */
#define BDG_NCCL_ID_BYTES 128
int bdg_nccl_unique_id(void *id_out /* BDG_NCCL_ID_BYTES */);
/* joins the communicator; collective over all ranks of the halo description */
int bdg_sw2d_comm_init(bdg_sw2d *s, const void *id /* BDG_NCCL_ID_BYTES */);
/* test hook: pack this rank's send buffer to host / load ghost values from host, so the
 * exchange can be emulated without NCCL (world_size-2 gloo tests on CPU boxes use the oracle) */
int bdg_sw2d_halo_pack_to_host(bdg_sw2d *s, double *send_buf /* 3*n_send */);
int bdg_sw2d_halo_set_ghosts(bdg_sw2d *s, const double *recv_buf /* 3*n_ghost */);

/* Introspection (host only, no GPU needed): the device element order bdg_sw2d_create would choose for this
 * connectivity -- perm_out[k_device] = caller's element.  The reference keeps the .msh order
 * (src/MeshManager.cpp:130-247); the device path regroups elements into compact 32-element patches
 * (blitzdg_b200/csrc/sw2d_order.cu) and hides the permutation behind set_state/get_state/rhs.
 * vmapP as in bdg_dgcontext2d; n_interior as in bdg_halo (K when there is no halo). */
int bdg_element_order(const int32_t *vmapP, int32_t Np, int32_t Nfp, int32_t Nfaces, int32_t K, int32_t n_interior,
                      int32_t *perm_out);

int bdg_version(void);

#ifdef __cplusplus
}
#endif
#endif /* BLITZDG_B200_H */
