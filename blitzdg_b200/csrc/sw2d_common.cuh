// Shared declarations of the sw2d device path: solver handle, device views, launch entry points.
//
// Device data layout: every field is stored ELEMENT-MAJOR, index m = n + Np*k --
// exactly the index space vmapM/vmapP address -- so the reference's fullToVector/vectorToFull
// sweeps (include/BlitzHelpers.hpp:215-297) disappear; conversion from the host's row-major
// [Np][K] happens once at upload.  Ghost values of off-rank neighbours sit behind the local
// nodes at m = Np*K + g.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/blitzdg_b200.h"

struct ncclComm;

// scalar slots kept on the device (no host round trips inside bdg_sw2d_step)
// S_HALO_TIMEOUT is sticky: set by k_wait_epoch when a neighbour's epoch never arrives, read (never cleared) by bdg_sw2d_check
// S_LAM0..2: global LF speed slots of the fused full-physics path (this stage's / the next stage's / the one being re-armed)
enum { S_T = 0, S_DT = 1, S_GMAX0 = 2, S_GMAX1 = 3, S_DTMAX = 4, S_ETAMAX = 5, S_NANFLAG = 6, S_HALO_TIMEOUT = 7, S_LAM0 = 8, S_LAM1 = 9, S_LAM2 = 10, S_COUNT = 16 };

// everything a kernel needs, passed by value
struct Sw2dView {
  int N, Np, Nfp, Nf, Nfn, K;
  int full;        // star states + sources (prm.sources)
  int lf_global;   // prm.lf_mode == 1
  int use_filter;
  long long stride;  // distance between the h, hu, hv planes of a state buffer (>= Np*K + n_ghost)
  double g, CD, fcor, tide_amp, tide_period, tide_ramp;
  const double *Dr, *Ds, *Lift, *Filt;   // row-major as uploaded
  const double *rx, *sx, *ry, *sy;       // [K][Np] element-major (per node)
  const double *nx, *ny, *Fscale;        // [K][Nfn] element-major (per face node)
  const double *aff;                     // affine triangles: [K][4+3*Nf] = rx,sx,ry,sy, then (nx,ny,Fscale) per face; else NULL
  const int32_t *vmapP;                  // [K][Nfn]
  const int32_t *fmask;                  // [Nfn] local node of face-node slot
  const uint8_t *bctag;                  // [K][Nfn] BCTag per face node (0 interior)
  const uint8_t *facetag;                // [K][Nf]  BCTag per face (fused path), 255 = mixed
  const double *H, *Hx, *Hy;             // H: [Np*K + n_ghost]; Hx,Hy: [Np*K]; NULL when unset
  const double *sponge;                  // [Np*K] sponge coefficient (src/sw2d/main.cpp:184-186); NULL = none
  double *scal;                          // S_COUNT device scalars
};

// Time-stepper forms of the volume kernels' epilogue; st = dt*RHS_i (modes 2, 3) or a*res_i + dt*RHS_i (mode 1),
// qin = the state the RHS was evaluated on, qold = the previous content of the destination buffer:
//   mode 1  LSERK4 stage            q  = qin + b*st                  (src/advec1d/main.cpp:98-101)
//   mode 2  SSP-RK2 first stage     q1 = qin + st                    (src/sw2d/main.cpp:217-219)
//   mode 3  SSP-RK2 second stage    q  = 0.5*(qold + qin + st)       (src/sw2d/main.cpp:230-232; qin = q1, qold = q^n)
// followed in modes 2, 3 by the sponge relaxation of the momenta, hu /= 1 + sigma*hu*hu (main.cpp:222-224,235-236).
#ifdef __CUDACC__
__device__ __forceinline__ double stepper_value(int mode, double b, double st, double qin, double qold, int field, double sigma) {
  if (mode == 1) return fma(b, st, qin);
  double val = mode == 2 ? qin + st : 0.5 * ((qold + qin) + st);
  if (field > 0 && sigma != 0.0) val = val / (1.0 + sigma * val * val);
  return val;
}
#endif

// peer-memory halo (sw2d_halo.cu): per neighbour i, the peer's two state buffers, plane stride, flag words
constexpr int kMaxNbr = 8;
struct PeerTable {
  int n;
  int rank[kMaxNbr];
  double *q[2][kMaxNbr];
  long long stride[kMaxNbr];
  unsigned long long *flags[kMaxNbr];  // the peer's flag array [nranks]; I write entry [my rank]
};

// Device copies of the cubature volume mesh and the Gauss face nodes (sw2d_curved.cu; bdg_curved2d of the header), every
// per-element table element-major in DEVICE element order.
struct CurvedOps {
  int Ncub, NG3;                              // cubature points per element, Gauss nodes per element (3 faces)
  const double *V, *Dr, *Ds;                  // [Ncub][Np]
  const double *cW, *crx, *csx, *cry, *csy;   // [K][Ncub]
  const double *gInterp;                      // [NG3][Np]
  const double *gW, *gnx, *gny;               // [K][NG3]
  const int32_t *mapM, *mapP;                 // [K][NG3], device numbering g + NG3*k
  const uint8_t *wall;                        // [K][NG3] 1 on Wall faces
  const double *J;                            // [K][Np]
  const double *mminv;                        // [Np][Np] V V^T
  const int32_t *chol_idx;                    // [K]: index into chol, -1 = straight element
  const double *chol;                         // [n_curved][Np][Np] upper Cholesky factors
  double *gq;                                 // [4][NG3*K] scratch: fields at the Gauss nodes
  double *CD;                                 // [K][Np] scratch: nodal drag coefficient of the current call
  double *out;                                // [4][Np*K] scratch: RHS in device order
};

struct bdg_sw2d {
  int device;
  cudaStream_t stream, comm_stream;
  cudaEvent_t ev_pack, ev_recv;
  bdg_sw2d_params prm;
  Sw2dView v;
  int n_ghost;
  size_t nT, nF;        // Np*K, Nfn*K
  int is_affine;
  int variant;          // 0 auto, 1 generic, 2 fused
  int cur;              // ping-pong index of the live state buffer
  double *q[2];         // [3][stride]
  double *res;          // [3][stride] (same plane stride as q: 128-B aligned planes for the bulk copies)
  double *rhs;          // [3][nT] scratch (bdg_sw2d_rhs) -- allocated on demand
  double *sflux;        // [3][nF] generic path: Fscale*dFlux
  double *d_Dr, *d_Ds, *d_Lift, *d_Filt, *d_rx, *d_sx, *d_ry, *d_sy, *d_nx, *d_ny, *d_Fscale, *d_aff;
  double *d_H, *d_Hx, *d_Hy, *d_sponge;
  double *qN[2], *resN;   // tracer hN (sw2d_tracer.cu): ping-pong pair + residual, [Np*K]; NULL = no tracer
  int curN;
  int32_t *d_vmapP, *d_fmask;
  int32_t *d_perm;      // device element k = caller's element d_perm[k] (sw2d_order.cu); NULL = identity
  uint8_t *d_bctag, *d_facetag;
  double *d_scal;
  double *d_Dq;         // quads with a tensor-product basis: 1-D differentiation matrix [quad_n][quad_n], else NULL
  int quad_n;
  double *d_vfrag;      // operator fragments of the generic DMMA volume kernel (sw2d_volume_dmma.cu), NULL if Np > 64
  void *fused_ops;      // operator fragments of the fused path (device), NULL if unsupported
  int fused_ok;
  int delta_dirty;      // fused full physics: the star-state table must be rebuilt from the bathymetry before the next launch
  int lam_valid;        // ... and scal[S_LAM0 + lam_slot] holds the global LF speed of the live state
  int lam_slot;
  double t;             // host mirror of the time (exact for fixed dt)
  int adaptive_used;
  // halo
  int rank, nranks, n_nbr, n_send, n_interior;
  int32_t *h_nbr, *h_send_off, *h_recv_off;
  int32_t *d_send_nodes;
  double *d_sendbuf;    // [n_nbr segments][3 (or 4)][count]
  ncclComm *comm;
  // peer-memory halo
  int peer_ok;                    // IPC mappings established: the fused per-face path exchanges halos by direct stores
  PeerTable peer;
  void *peer_base[3][kMaxNbr];    // opened IPC mappings (q0, q1, flags) to close at destroy
  unsigned long long *d_flags;    // [nranks] epochs published by my neighbours
  unsigned int *d_put_counter;    // CTAs of the current k_put launch that have finished their stores (last one publishes)
  int put_ctas;                   // grid of k_put
  long long *d_put_dst;           // [n_send] ghost slot index in the receiver's plane
  unsigned long long epoch;       // number of puts issued (identical on all ranks: same call sequence)
  int put_pending;                // a put on the communication stream has not been joined by the main stream yet
  int ghost_valid;                // the live buffer's ghost slots hold its neighbours' traces
  // output path (sw2d_output.cu): interpolation matrix [out_n][Np], result [3][out_n*K] in the caller's layout
  double *d_IM, *d_out;
  int out_n;
  cudaStream_t out_stream;
  cudaEvent_t ev_out_state, ev_out_kernel, ev_out_done;
  // streaming interface (capi.cu: stage_state / commit_state / get_state_async): raw row-major planes [3][Np*K]
  double *in_raw, *dl_raw;
  double *stage_buf;              // transposition staging of set_state / get_state / rhs / set_bathymetry (per handle)
  size_t stage_cap;
  cudaStream_t in_stream, dl_stream;
  cudaEvent_t ev_in_done, ev_in_free, ev_dl_conv, ev_dl_done;
  int staged, dl_pending;
  double staged_t;
  CurvedOps *curved;              // cubature / Gauss contexts of the over-integrated RHS (bdg_sw2d_set_curved), NULL = none
  int64_t launches;
  char err[512];
};

// ---- launchers (each returns cudaGetLastError()) ------------------------------------------------
// generic two-kernel path over elements [k0,k1): face fluxes into sflux, then volume+lift+update
cudaError_t launch_generic_speed(bdg_sw2d *s, const double *q, int slot);
cudaError_t launch_generic_face(bdg_sw2d *s, const double *q, int k0, int k1, int slot);
// mode 0: write RHS to out (3 planes of nT); mode 1: LSERK4 stage update of q/res in place; modes 2, 3: the two
// SSP-RK2 stages of the src/sw2d driver, state read from q, result stored to qdst (see stepper_value)
cudaError_t launch_generic_volume(bdg_sw2d *s, double *q, double *res, double *out, int mode, double a, double b, double *qdst);
// over-integrated RHS on curved triangles (sw2d_curved.cu): out [4][Np*K] in device order
cudaError_t launch_curved_rhs(bdg_sw2d *s, const double *q, const double *qN, const double *CDn, double *out);
cudaError_t launch_curved_update(bdg_sw2d *s, const double *Filt, const double *R, double c, const double *base, const double *baseN, double *dst,
                                 double *dstN);
cudaError_t launch_dt_reduce(bdg_sw2d *s, const double *q);
cudaError_t launch_dt_finalize(bdg_sw2d *s);
cudaError_t launch_advance_time(bdg_sw2d *s);
cudaError_t launch_check(bdg_sw2d *s, const double *q);
cudaError_t launch_pack(bdg_sw2d *s, const double *q, int with_H);
cudaError_t launch_positivity_limiter(bdg_sw2d *s, double *q);  // swhelpers/limiters.py:3-21 on the live state, in place
cudaError_t launch_output_fields(bdg_sw2d *s, const double *q, cudaStream_t stream);  // eta, u, v interpolated into s->d_out
cudaError_t launch_put(bdg_sw2d *s, const double *q, int buf, unsigned long long epoch, int side);
// tracer stage / RHS beside the flow's stage kernel (mode 0: RHS4 into rhs_out; 1 LSERK4; 2, 3 SSP-RK2 forms)
cudaError_t launch_tracer(bdg_sw2d *s, const double *q, const double *hN_in, double *hN_out, double *rhs_out, int mode, double a, double b);
cudaError_t launch_wait_epoch(bdg_sw2d *s, unsigned long long epoch);
// fused DMMA path (affine triangles): q_in -> q_out, res in place, elements [k0,k1)
int fused_supported(const bdg_sw2d *s);
cudaError_t fused_prepare(bdg_sw2d *s, const double *Dr, const double *Ds, const double *Lift, const int32_t *vmapP,
                          const uint8_t *facetag, const int32_t *fmask);  // host arrays of create; sets s->fused_ops when the path applies
struct FusedStage {   // stepper form and global-LF bookkeeping of one fused launch (NULL = LSERK4 stage, nothing carried)
  int mode;           // 1 LSERK4, 2 / 3 first / second SSP-RK2 stage (stepper_value above)
  int slot_next;      // full mode: scal slot collecting the global LF speed of the new state, -1 = none
  int slot_zero;      // ... and the slot to re-arm, -1 = none
  int last_stage;     // the next RHS evaluation is at t + dt
};
cudaError_t launch_fused_stage(bdg_sw2d *s, const double *q_in, double *q_out, double *res, int k0, int k1, double a,
                               double b, int write_rhs, double *rhs_out, int slot, const FusedStage *fs = nullptr);
cudaError_t fused_build_delta(bdg_sw2d *s);  // star-state table of the fused full-physics path from v.H (device)
void fused_release(bdg_sw2d *s);
// generic path, second launch on the fp64 tensor pipe (sw2d_volume_dmma.cu)
cudaError_t volume_dmma_prepare(bdg_sw2d *s, const double *Dr, const double *Ds, const double *Lift);
int volume_dmma_supported(const bdg_sw2d *s);
cudaError_t launch_volume_dmma(bdg_sw2d *s, double *q, double *res, double *out, int mode, double a, double b, double *qdst);
// device element order (sw2d_order.cu): perm[new] = old, empty = identity
#include <vector>
void build_element_order(const int32_t *vmapP, int Np, int Nfp, int Nf, int K, int n_interior, int tile, std::vector<int32_t> &perm);
