// C ABI of the sw2d device path (include/blitzdg_b200.h): upload once, keep all state on the GPU,
// enqueue 5 fused RHS+stage launches per LSERK4 step, exchange face-trace halos over NCCL.
#include <nccl.h>

#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <vector>

#include "sw2d_common.cuh"

namespace {

char g_create_err[512] = "";

const double rk4a[5] = {0.0, -567301805773.0 / 1357537059087.0, -2404267990393.0 / 2016746695238.0,
                        -3550918686646.0 / 2091501179385.0, -1275806237668.0 / 842570457699.0};
const double rk4b[5] = {1432997174477.0 / 9575080441755.0, 5161836677717.0 / 13612068292357.0,
                        1720146321549.0 / 2090206949498.0, 3134564353537.0 / 4481467310338.0,
                        2277821191437.0 / 14882151754819.0};  // reference include/LSERK4.hpp:17-27

int fail(bdg_sw2d *s, int code, const char *fmt, ...) {
  char *dst = s ? s->err : g_create_err;
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(dst, 512, fmt, ap);
  va_end(ap);
  return code;
}

#define CU(call)                                                                                   \
  do {                                                                                             \
    cudaError_t e_ = (call);                                                                       \
    if (e_ != cudaSuccess) return fail(s, BDG_E_CUDA, "%s -> %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
  } while (0)
#define NC(call)                                                                                   \
  do {                                                                                             \
    ncclResult_t r_ = (call);                                                                      \
    if (r_ != ncclSuccess) return fail(s, BDG_E_NCCL, "%s -> %s (%s:%d)", #call, ncclGetErrorString(r_), __FILE__, __LINE__); \
  } while (0)

// ---- layout conversion on the device -------------------------------------------------------------
// in: row-major [R][K] (k fastest, the reference's blitz storage) -> out: element-major [K][R]
// perm (optional): device element k holds the caller's element perm[k] (sw2d_order.cu)
__global__ void k_rowmajor_to_elem(const double *__restrict__ in, double *__restrict__ out, int R, long long K, int KB,
                                   const int32_t *__restrict__ perm) {
  extern __shared__ double tile[];  // [R][KB+1]
  const long long k0 = (long long)blockIdx.x * KB;
  const int nk = (int)min((long long)KB, K - k0);
  for (int x = threadIdx.x; x < R * KB; x += blockDim.x) {
    const int n = x / KB, tk = x - n * KB;
    if (tk < nk) tile[n * (KB + 1) + tk] = in[(long long)n * K + (perm ? (long long)perm[k0 + tk] : k0 + tk)];
  }
  __syncthreads();
  for (int x = threadIdx.x; x < nk * R; x += blockDim.x) {
    const int tk = x / R, n = x - tk * R;
    out[k0 * R + x] = tile[n * (KB + 1) + tk];
  }
}
__global__ void k_elem_to_rowmajor(const double *__restrict__ in, double *__restrict__ out, int R, long long K, int KB,
                                   const int32_t *__restrict__ perm) {
  extern __shared__ double tile[];
  const long long k0 = (long long)blockIdx.x * KB;
  const int nk = (int)min((long long)KB, K - k0);
  for (int x = threadIdx.x; x < nk * R; x += blockDim.x) {
    const int tk = x / R, n = x - tk * R;
    tile[n * (KB + 1) + tk] = in[k0 * R + x];
  }
  __syncthreads();
  for (int x = threadIdx.x; x < R * KB; x += blockDim.x) {
    const int n = x / KB, tk = x - n * KB;
    if (tk < nk) out[(long long)n * K + (perm ? (long long)perm[k0 + tk] : k0 + tk)] = tile[n * (KB + 1) + tk];
  }
}

int tile_kb(int R) {
  int kb = 5800 / R - 1;
  if (kb > 128) kb = 128;
  if (kb < 4) kb = 4;
  return kb;
}

// transposition staging buffer of one handle (grown on demand, freed at destroy): every use is enqueued on the handle's
// own stream, so two handles -- even on the same device, driven from different threads -- never share it
double *staging(bdg_sw2d *s, size_t n) {
  if (s->stage_cap < n) {
    if (s->stage_buf) {
      cudaStreamSynchronize(s->stream);  // earlier conversions may still read the old buffer
      cudaFree(s->stage_buf);
    }
    s->stage_buf = nullptr;
    s->stage_cap = 0;
    if (cudaMalloc(&s->stage_buf, n * sizeof(double)) != cudaSuccess) return nullptr;
    s->stage_cap = n;
  }
  return s->stage_buf;
}

// host row-major [R][K] -> device element-major [K][R]
int upload_transposed(bdg_sw2d *s, const double *host, double *dev, int R, long long K) {
  double *tmp = staging(s, (size_t)R * K);
  if (!tmp) return fail(s, BDG_E_CUDA, "staging allocation of %zu doubles failed", (size_t)R * K);
  CU(cudaMemcpyAsync(tmp, host, sizeof(double) * (size_t)R * K, cudaMemcpyHostToDevice, s->stream));
  const int KB = tile_kb(R);
  k_rowmajor_to_elem<<<(unsigned)((K + KB - 1) / KB), 256, sizeof(double) * R * (KB + 1), s->stream>>>(tmp, dev, R, K, KB, s->d_perm);
  s->launches++;
  CU(cudaGetLastError());
  return BDG_OK;
}
int download_transposed(bdg_sw2d *s, const double *dev, double *host, int R, long long K) {
  double *tmp = staging(s, (size_t)R * K);
  if (!tmp) return fail(s, BDG_E_CUDA, "staging allocation failed");
  const int KB = tile_kb(R);
  k_elem_to_rowmajor<<<(unsigned)((K + KB - 1) / KB), 256, sizeof(double) * R * (KB + 1), s->stream>>>(dev, tmp, R, K, KB, s->d_perm);
  s->launches++;
  CU(cudaGetLastError());
  CU(cudaMemcpyAsync(host, tmp, sizeof(double) * (size_t)R * K, cudaMemcpyDeviceToHost, s->stream));
  return BDG_OK;
}

template <typename T>
int dev_alloc(bdg_sw2d *s, T **p, size_t n) {
  CU(cudaMalloc((void **)p, sizeof(T) * (n ? n : 1)));
  return BDG_OK;
}
template <typename T>
int dev_upload(bdg_sw2d *s, T **p, const T *host, size_t n) {
  int rc = dev_alloc(s, p, n);
  if (rc) return rc;
  CU(cudaMemcpyAsync(*p, host, sizeof(T) * n, cudaMemcpyHostToDevice, s->stream));
  return BDG_OK;
}

}  // namespace

static void curved_free(bdg_sw2d *s);

// ====================================================================================== lifetime
extern "C" int bdg_version(void) { return 102; }

extern "C" int bdg_element_order(const int32_t *vmapP, int32_t Np, int32_t Nfp, int32_t Nfaces, int32_t K, int32_t n_interior,
                                 int32_t *perm_out) {
  if (!vmapP || !perm_out || Np <= 0 || Nfp <= 0 || K <= 0 || (Nfaces != 3 && Nfaces != 4)) return BDG_E_INVALID;
  std::vector<int32_t> perm;
  build_element_order(vmapP, Np, Nfp, Nfaces, K, n_interior, 32, perm);
  for (int32_t k = 0; k < K; ++k) perm_out[k] = perm.empty() ? k : perm[k];
  return BDG_OK;
}

extern "C" const char *bdg_last_error(const bdg_sw2d *s) { return s ? s->err : g_create_err; }

extern "C" void bdg_sw2d_destroy(bdg_sw2d *s) {
  if (!s) return;
  cudaSetDevice(s->device);
  cudaStreamSynchronize(s->stream);
  for (int k = 0; k < 3; ++k)
    for (int i = 0; i < kMaxNbr; ++i)
      if (s->peer_base[k][i]) cudaIpcCloseMemHandle(s->peer_base[k][i]);
  if (s->comm) ncclCommDestroy((ncclComm_t)s->comm);
  if (s->d_flags) cudaFree(s->d_flags);
  if (s->d_put_counter) cudaFree(s->d_put_counter);
  if (s->d_put_dst) cudaFree(s->d_put_dst);
  fused_release(s);
  curved_free(s);
  void *ptrs[] = {s->q[0], s->q[1], s->res, s->rhs, s->sflux, s->d_Dr, s->d_Ds, s->d_Lift, s->d_Filt, s->d_rx, s->d_sx,
                  s->d_ry, s->d_sy, s->d_nx, s->d_ny, s->d_Fscale, s->d_aff, s->d_H, s->d_Hx, s->d_Hy, s->d_vmapP,
                  s->d_fmask, s->d_bctag, s->d_facetag, s->d_scal, s->d_send_nodes, s->d_sendbuf, s->d_perm, s->d_vfrag, s->d_Dq, s->d_sponge, s->d_IM, s->d_out, s->in_raw, s->dl_raw, s->stage_buf, s->qN[0], s->qN[1], s->resN};
  for (void *p : ptrs)
    if (p) cudaFree(p);
  free(s->h_nbr); free(s->h_send_off); free(s->h_recv_off);
  for (cudaEvent_t ev : {s->ev_in_done, s->ev_in_free, s->ev_dl_conv, s->ev_dl_done})
    if (ev) cudaEventDestroy(ev);
  if (s->in_stream) cudaStreamDestroy(s->in_stream);
  if (s->dl_stream) cudaStreamDestroy(s->dl_stream);
  if (s->ev_out_state) cudaEventDestroy(s->ev_out_state);
  if (s->ev_out_kernel) cudaEventDestroy(s->ev_out_kernel);
  if (s->ev_out_done) cudaEventDestroy(s->ev_out_done);
  if (s->out_stream) cudaStreamDestroy(s->out_stream);
  if (s->ev_pack) cudaEventDestroy(s->ev_pack);
  if (s->ev_recv) cudaEventDestroy(s->ev_recv);
  if (s->comm_stream) cudaStreamDestroy(s->comm_stream);
  if (s->stream) cudaStreamDestroy(s->stream);
  delete s;
}

static int create_impl(bdg_sw2d *s, const bdg_dgcontext2d *c, const bdg_sw2d_params *prm, const bdg_halo *halo) {
  const int Np = c->Np, Nfp = c->Nfp, Nf = c->Nfaces, K = c->K, Nfn = Nf * Nfp;
  if (Np <= 0 || Nfp <= 0 || K <= 0 || (Nf != 3 && Nf != 4)) return fail(s, BDG_E_INVALID, "bad sizes Np=%d Nfp=%d Nfaces=%d K=%d", Np, Nfp, Nf, K);
  if (Np > 1024) return fail(s, BDG_E_UNSUPPORTED, "Np=%d > 1024", Np);
  if (!c->Dr || !c->Ds || !c->Lift || !c->rx || !c->sx || !c->ry || !c->sy || !c->nx || !c->ny || !c->Fscale || !c->vmapM || !c->vmapP)
    return fail(s, BDG_E_INVALID, "NULL array in bdg_dgcontext2d");
  if ((long long)Np * K + (halo ? halo->n_ghost : 0) >= (1LL << 31)) return fail(s, BDG_E_UNSUPPORTED, "Np*K exceeds int32 (reference index_type)");
  if (prm->filter && !c->Filt) return fail(s, BDG_E_INVALID, "filter requested but Filt is NULL");
  s->prm = *prm;
  s->n_ghost = halo ? halo->n_ghost : 0;
  s->nT = (size_t)Np * K;
  s->nF = (size_t)Nfn * K;
  const long long nTot = (long long)s->nT + s->n_ghost;
  const long long stride = (nTot + 15) / 16 * 16;

  // Fmask from vmapM of element 0, then check vmapM[k][sl] == Fmask[sl] + Np*k everywhere (reference
  // src/TriangleNodesProvisioner.cpp:930-936): the device path derives '-' traces from Fmask alone.
  std::vector<int32_t> fmask(Nfn);
  for (int sl = 0; sl < Nfn; ++sl) {
    fmask[sl] = c->vmapM[sl];
    if (fmask[sl] < 0 || fmask[sl] >= Np) return fail(s, BDG_E_INVALID, "vmapM[%d]=%d is not a local node of element 0", sl, fmask[sl]);
  }
  for (long long k = 0; k < K; ++k)
    for (int sl = 0; sl < Nfn; ++sl)
      if (c->vmapM[k * Nfn + sl] != fmask[sl] + Np * k) return fail(s, BDG_E_INVALID, "vmapM is not Fmask + Np*k at element %lld slot %d", k, sl);
  long long first_ghost_elem = K;  // first element with a '+' node in a ghost slot
  for (size_t i = 0; i < s->nF; ++i) {
    if (c->vmapP[i] < 0 || c->vmapP[i] >= nTot) return fail(s, BDG_E_INVALID, "vmapP[%zu]=%d out of range", i, c->vmapP[i]);
    if ((size_t)c->vmapP[i] >= s->nT && (long long)(i / Nfn) < first_ghost_elem) first_ghost_elem = (long long)(i / Nfn);
  }
  // The interior launch over [0,n_interior) overlaps the halo exchange on the promise that those elements read no ghost
  // slot: a caller-supplied split that breaks the promise would be a silent data race, so it is rejected here.
  if (halo && halo->n_interior > first_ghost_elem)
    return fail(s, BDG_E_INVALID, "halo->n_interior=%d but element %lld reads a ghost slot (interior elements must come first)", halo->n_interior,
                first_ghost_elem);

  // per-face-node BC tags from the BCmap lists (reference include/DGContext2D.hpp bcmap())
  std::vector<uint8_t> bctag(s->nF, 0), facetag((size_t)K * Nf, 0);
  for (int tag = 1; tag < 9; ++tag)
    for (int j = 0; j < c->bc_count[tag]; ++j) {
      const int32_t i = c->bc_face_nodes[tag][j];
      if (i < 0 || (size_t)i >= s->nF) return fail(s, BDG_E_INVALID, "BC list %d entry %d out of range", tag, i);
      bctag[i] = (uint8_t)tag;
    }
  bool face_uniform = true;
  for (size_t fc = 0; fc < (size_t)K * Nf; ++fc) {
    const uint8_t t0 = bctag[fc * Nfp];
    facetag[fc] = t0;
    for (int j = 1; j < Nfp; ++j)
      if (bctag[fc * Nfp + j] != t0) { facetag[fc] = 255; face_uniform = false; }
  }

  // Device element order (sw2d_order.cu): compact 32-element patches for the fused kernel's tiles.  From here on
  // every per-element table is built in device order; perm[k_dev] = caller's element.
  const int n_interior_in = (halo && halo->n_interior > 0 && halo->n_interior <= K) ? halo->n_interior : (halo && halo->n_neighbors > 0 ? 0 : K);
  std::vector<int32_t> perm, inv, vmapP_dev;
  build_element_order(c->vmapP, Np, Nfp, Nf, K, n_interior_in, 32, perm);
  const bool renum = !perm.empty();
  const int32_t *vmapP_h = c->vmapP;
  auto node_dev = [&](int32_t m) -> int32_t {  // volume node id: caller's numbering -> device numbering (ghost slots unchanged)
    if (!renum || (size_t)m >= s->nT) return m;
    return (int32_t)(m % Np + (long long)Np * inv[m / Np]);
  };
  if (renum) {
    inv.resize(K);
    for (int k = 0; k < K; ++k) inv[perm[k]] = k;
    vmapP_dev.resize(s->nF);
    std::vector<uint8_t> bt(s->nF), ft((size_t)K * Nf);
    for (long long k = 0; k < K; ++k) {
      const long long ko = perm[k];
      for (int sl = 0; sl < Nfn; ++sl) {
        vmapP_dev[k * Nfn + sl] = node_dev(c->vmapP[ko * Nfn + sl]);
        bt[k * Nfn + sl] = bctag[ko * Nfn + sl];
      }
      for (int f = 0; f < Nf; ++f) ft[k * Nf + f] = facetag[ko * Nf + f];
    }
    bctag.swap(bt);
    facetag.swap(ft);
    vmapP_h = vmapP_dev.data();
  }

  CU(cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking));
  if (renum) {
    int rcp = dev_upload(s, &s->d_perm, perm.data(), (size_t)K);
    if (rcp) return rcp;
  }
  {  // the halo exchange must win the race for SM slots against the interior-element kernel it overlaps with
    int lo = 0, hi = 0;
    CU(cudaDeviceGetStreamPriorityRange(&lo, &hi));
    CU(cudaStreamCreateWithPriority(&s->comm_stream, cudaStreamNonBlocking, hi));
  }
  CU(cudaEventCreateWithFlags(&s->ev_pack, cudaEventDisableTiming));
  CU(cudaEventCreateWithFlags(&s->ev_recv, cudaEventDisableTiming));

  int rc;
  for (int b = 0; b < 2; ++b) {
    if ((rc = dev_alloc(s, &s->q[b], (size_t)3 * stride))) return rc;
    CU(cudaMemsetAsync(s->q[b], 0, sizeof(double) * 3 * stride, s->stream));
  }
  if ((rc = dev_alloc(s, &s->res, (size_t)3 * stride))) return rc;
  CU(cudaMemsetAsync(s->res, 0, sizeof(double) * 3 * stride, s->stream));
  if ((rc = dev_alloc(s, &s->sflux, 3 * s->nF))) return rc;
  if ((rc = dev_alloc(s, &s->d_scal, (size_t)S_COUNT))) return rc;
  CU(cudaMemsetAsync(s->d_scal, 0, sizeof(double) * S_COUNT, s->stream));
  if ((rc = dev_upload(s, &s->d_Dr, c->Dr, (size_t)Np * Np))) return rc;
  if ((rc = dev_upload(s, &s->d_Ds, c->Ds, (size_t)Np * Np))) return rc;
  if ((rc = dev_upload(s, &s->d_Lift, c->Lift, (size_t)Np * Nfn))) return rc;
  if (c->Filt && (rc = dev_upload(s, &s->d_Filt, c->Filt, (size_t)Np * Np))) return rc;
  if ((rc = dev_upload(s, &s->d_vmapP, vmapP_h, s->nF))) return rc;
  if ((rc = dev_upload(s, &s->d_fmask, fmask.data(), (size_t)Nfn))) return rc;
  if ((rc = dev_upload(s, &s->d_bctag, bctag.data(), s->nF))) return rc;
  if ((rc = dev_upload(s, &s->d_facetag, facetag.data(), (size_t)K * Nf))) return rc;
  struct { double **dst; const double *src; int R; } geo[] = {
      {&s->d_rx, c->rx, Np}, {&s->d_sx, c->sx, Np}, {&s->d_ry, c->ry, Np}, {&s->d_sy, c->sy, Np},
      {&s->d_nx, c->nx, Nfn}, {&s->d_ny, c->ny, Nfn}, {&s->d_Fscale, c->Fscale, Nfn}};
  for (auto &gm : geo) {
    if ((rc = dev_alloc(s, gm.dst, (size_t)gm.R * K))) return rc;
    if ((rc = upload_transposed(s, gm.src, *gm.dst, gm.R, K))) return rc;
  }

  // affine detection: metrics constant per element, normals / Fscale constant per face (straight-sided
  // triangles, reference src/TriangleNodesProvisioner.cpp:807-821 stores them per node regardless)
  s->is_affine = 0;
  if (Nf == 3) {
    const int W = 4 + 3 * Nf;
    std::vector<double> aff((size_t)K * W, 0.0), lo((size_t)K), hi((size_t)K);
    bool ok = true;
    const double *vol[4] = {c->rx, c->sx, c->ry, c->sy};
    double scale = 0.0;
    for (int a = 0; a < 4 && ok; ++a) {
      for (long long k = 0; k < K; ++k) { lo[k] = hi[k] = vol[a][k]; aff[k * W + a] = 0.0; }
      for (int n = 0; n < Np; ++n) {
        const double *row = vol[a] + (size_t)n * K;
        for (long long k = 0; k < K; ++k) {
          const double x = row[k];
          aff[k * W + a] += x;
          if (x < lo[k]) lo[k] = x;
          if (x > hi[k]) hi[k] = x;
        }
      }
      for (long long k = 0; k < K; ++k) { aff[k * W + a] /= Np; scale = fmax(scale, fabs(aff[k * W + a])); }
      for (long long k = 0; k < K; ++k)
        if (hi[k] - lo[k] > 1e-9 * scale) { ok = false; break; }
    }
    const double *fac[3] = {c->nx, c->ny, c->Fscale};
    for (int a = 0; a < 3 && ok; ++a)
      for (int f = 0; f < Nf && ok; ++f) {
        for (long long k = 0; k < K; ++k) { lo[k] = hi[k] = fac[a][(size_t)(f * Nfp) * K + k]; }
        for (int j = 0; j < Nfp; ++j) {
          const double *row = fac[a] + (size_t)(f * Nfp + j) * K;
          for (long long k = 0; k < K; ++k) {
            const double x = row[k];
            aff[k * W + 4 + 3 * f + a] += x;
            if (x < lo[k]) lo[k] = x;
            if (x > hi[k]) hi[k] = x;
          }
        }
        for (long long k = 0; k < K; ++k) {
          aff[k * W + 4 + 3 * f + a] /= Nfp;
          if (hi[k] - lo[k] > 1e-9 * fmax(1.0, fabs(aff[k * W + 4 + 3 * f + a]))) { ok = false; break; }
        }
      }
    if (ok) {
      s->is_affine = 1;
      // The fused kernel wants the face triple as (nx, ny, 1) * Fscale/2: the 1/2 of the flux jump (main.cpp:416-421) and
      // the Fscale of the surface term (:451-458) are folded into the normal once, here, instead of per face node per stage.
      for (long long k = 0; k < K; ++k)
        for (int f = 0; f < Nf; ++f) {
          double *t3 = &aff[k * W + 4 + 3 * f];
          const double hfs = 0.5 * t3[2];
          t3[0] *= hfs; t3[1] *= hfs; t3[2] = hfs;
        }
      if (renum) {
        std::vector<double> a2(aff.size());
        for (long long k = 0; k < K; ++k) memcpy(&a2[k * W], &aff[(size_t)perm[k] * W], sizeof(double) * W);
        aff.swap(a2);
      }
      if ((rc = dev_upload(s, &s->d_aff, aff.data(), aff.size()))) return rc;
      CU(cudaStreamSynchronize(s->stream));
    }
  }

  // halo
  s->rank = halo ? halo->rank : 0;
  s->nranks = halo ? halo->nranks : 1;
  s->n_nbr = halo ? halo->n_neighbors : 0;
  s->n_interior = n_interior_in;
  if (s->n_nbr > 0) {
    const int nn = s->n_nbr;
    s->h_nbr = (int32_t *)malloc(sizeof(int32_t) * nn);
    s->h_send_off = (int32_t *)malloc(sizeof(int32_t) * (nn + 1));
    s->h_recv_off = (int32_t *)malloc(sizeof(int32_t) * (nn + 1));
    memcpy(s->h_nbr, halo->nbr_rank, sizeof(int32_t) * nn);
    memcpy(s->h_send_off, halo->send_offset, sizeof(int32_t) * (nn + 1));
    memcpy(s->h_recv_off, halo->recv_offset, sizeof(int32_t) * (nn + 1));
    s->n_send = s->h_send_off[nn];
    if (s->h_recv_off[nn] != s->n_ghost) return fail(s, BDG_E_INVALID, "recv_offset end %d != n_ghost %d", s->h_recv_off[nn], s->n_ghost);
    std::vector<int32_t> pk((size_t)2 * s->n_send + nn + 1);
    for (int i = 0; i < nn; ++i)
      for (int j = s->h_send_off[i]; j < s->h_send_off[i + 1]; ++j) {
        const int32_t m = halo->send_nodes[j];
        if (m < 0 || (size_t)m >= s->nT) return fail(s, BDG_E_INVALID, "send_nodes[%d]=%d out of range", j, m);
        pk[j] = node_dev(m);
        pk[s->n_send + j] = i;
      }
    memcpy(pk.data() + 2 * (size_t)s->n_send, s->h_send_off, sizeof(int32_t) * (nn + 1));
    if ((rc = dev_upload(s, &s->d_send_nodes, pk.data(), pk.size()))) return rc;
    if ((rc = dev_alloc(s, &s->d_sendbuf, (size_t)4 * s->n_send))) return rc;
    CU(cudaStreamSynchronize(s->stream));
  }

  Sw2dView &v = s->v;
  v.N = c->N; v.Np = Np; v.Nfp = Nfp; v.Nf = Nf; v.Nfn = Nfn; v.K = K;
  v.full = prm->sources != 0; v.lf_global = prm->lf_mode == 1; v.use_filter = prm->filter != 0;
  v.stride = stride;
  v.g = prm->g; v.CD = prm->CD; v.fcor = prm->f;
  v.tide_amp = prm->tide_amp; v.tide_period = prm->tide_period; v.tide_ramp = prm->tide_ramp;
  v.Dr = s->d_Dr; v.Ds = s->d_Ds; v.Lift = s->d_Lift; v.Filt = s->d_Filt;
  v.rx = s->d_rx; v.sx = s->d_sx; v.ry = s->d_ry; v.sy = s->d_sy;
  v.nx = s->d_nx; v.ny = s->d_ny; v.Fscale = s->d_Fscale; v.aff = s->d_aff;
  v.vmapP = s->d_vmapP; v.fmask = s->d_fmask; v.bctag = s->d_bctag; v.facetag = s->d_facetag;
  v.H = v.Hx = v.Hy = nullptr;
  v.scal = s->d_scal;

  {
    cudaError_t e = volume_dmma_prepare(s, c->Dr, c->Ds, c->Lift);
    if (e != cudaSuccess) return fail(s, BDG_E_CUDA, "volume_dmma_prepare: %s", cudaGetErrorString(e));
  }
  s->fused_ok = 0;
  if (s->is_affine && face_uniform && !v.use_filter && fused_supported(s)) {
    cudaError_t e = fused_prepare(s, c->Dr, c->Ds, c->Lift, vmapP_h, facetag.data(), fmask.data());
    if (e != cudaSuccess) return fail(s, BDG_E_CUDA, "fused_prepare: %s", cudaGetErrorString(e));
    s->fused_ok = s->fused_ops != nullptr;
  }
  CU(cudaStreamSynchronize(s->stream));
  return BDG_OK;
}

extern "C" int bdg_sw2d_create(bdg_sw2d **out, const bdg_dgcontext2d *ctx, const bdg_sw2d_params *prm, const bdg_halo *halo,
                               int device) {
  if (!out || !ctx || !prm) return fail(nullptr, BDG_E_INVALID, "NULL argument");
  *out = nullptr;
  cudaError_t e = cudaSetDevice(device);
  if (e != cudaSuccess) return fail(nullptr, BDG_E_CUDA, "cudaSetDevice(%d): %s", device, cudaGetErrorString(e));
  bdg_sw2d *s = new (std::nothrow) bdg_sw2d();
  if (!s) return fail(nullptr, BDG_E_INVALID, "out of host memory");
  memset(s, 0, sizeof(*s));
  s->device = device;
  int rc = create_impl(s, ctx, prm, halo);
  if (rc != BDG_OK) {
    snprintf(g_create_err, sizeof(g_create_err), "%s", s->err);
    bdg_sw2d_destroy(s);
    return rc;
  }
  *out = s;
  return BDG_OK;
}

// ========================================================================================= state
extern "C" int bdg_sw2d_set_bathymetry(bdg_sw2d *s, const double *H, const double *Hx, const double *Hy) {
  if (!s) return BDG_E_INVALID;
  CU(cudaSetDevice(s->device));
  const int Np = s->v.Np;
  const long long K = s->v.K;
  int rc;
  if (!H) {
    s->v.H = s->v.Hx = s->v.Hy = nullptr;
    s->lam_valid = 0;
    return BDG_OK;
  }
  if (!s->d_H) {
    if ((rc = dev_alloc(s, &s->d_H, s->nT + s->n_ghost))) return rc;
    if ((rc = dev_alloc(s, &s->d_Hx, s->nT))) return rc;
    if ((rc = dev_alloc(s, &s->d_Hy, s->nT))) return rc;
    CU(cudaMemsetAsync(s->d_H, 0, sizeof(double) * (s->nT + s->n_ghost), s->stream));
    CU(cudaMemsetAsync(s->d_Hx, 0, sizeof(double) * s->nT, s->stream));
    CU(cudaMemsetAsync(s->d_Hy, 0, sizeof(double) * s->nT, s->stream));
  }
  if ((rc = upload_transposed(s, H, s->d_H, Np, K))) return rc;
  if (Hx && (rc = upload_transposed(s, Hx, s->d_Hx, Np, K))) return rc;
  if (Hy && (rc = upload_transposed(s, Hy, s->d_Hy, Np, K))) return rc;
  s->v.H = s->d_H; s->v.Hx = s->d_Hx; s->v.Hy = s->d_Hy;
  s->delta_dirty = 1;  // the fused full-physics path rebuilds its star-state table before the next launch
  s->lam_valid = 0;
  CU(cudaStreamSynchronize(s->stream));
  return BDG_OK;
}

// spongeCoeff of the src/sw2d driver (main.cpp:184-186, built by buildSpongeCoeff :517-553); used by bdg_sw2d_step_ssprk2
extern "C" int bdg_sw2d_set_sponge(bdg_sw2d *s, const double *sigma) {
  if (!s) return BDG_E_INVALID;
  CU(cudaSetDevice(s->device));
  if (!sigma) {
    s->v.sponge = nullptr;
    return BDG_OK;
  }
  int rc;
  if (!s->d_sponge && (rc = dev_alloc(s, &s->d_sponge, s->nT))) return rc;
  if ((rc = upload_transposed(s, sigma, s->d_sponge, s->v.Np, s->v.K))) return rc;
  s->v.sponge = s->d_sponge;
  s->lam_valid = 0;
  CU(cudaStreamSynchronize(s->stream));
  return BDG_OK;
}

extern "C" int bdg_sw2d_set_state(bdg_sw2d *s, const double *h, const double *hu, const double *hv, double t) {
  if (!s || !h || !hu || !hv) return BDG_E_INVALID;
  CU(cudaSetDevice(s->device));
  if (s->put_pending) {
    CU(cudaStreamSynchronize(s->comm_stream));
    s->put_pending = 0;
  }
  double *q = s->q[s->cur];
  const double *src[3] = {h, hu, hv};
  for (int c = 0; c < 3; ++c) {
    int rc = upload_transposed(s, src[c], q + c * s->v.stride, s->v.Np, s->v.K);
    if (rc) return rc;
  }
  CU(cudaMemsetAsync(s->res, 0, sizeof(double) * 3 * s->v.stride, s->stream));  // reference src/sw2d/main.cpp:117-119
  s->t = t;
  s->ghost_valid = 0;
  s->lam_valid = 0;
  CU(cudaMemcpyAsync(s->d_scal + S_T, &s->t, sizeof(double), cudaMemcpyHostToDevice, s->stream));
  CU(cudaStreamSynchronize(s->stream));
  return BDG_OK;
}

extern "C" int bdg_sw2d_get_state(bdg_sw2d *s, double *h, double *hu, double *hv, double *t) {
  if (!s) return BDG_E_INVALID;
  CU(cudaSetDevice(s->device));
  const double *q = s->q[s->cur];
  double *dst[3] = {h, hu, hv};
  for (int c = 0; c < 3; ++c)
    if (dst[c]) {
      int rc = download_transposed(s, q + c * s->v.stride, dst[c], s->v.Np, s->v.K);
      if (rc) return rc;
      CU(cudaStreamSynchronize(s->stream));  // staging buffer is reused by the next plane
    }
  if (t) {
    CU(cudaMemcpyAsync(&s->t, s->d_scal + S_T, sizeof(double), cudaMemcpyDeviceToHost, s->stream));
    CU(cudaStreamSynchronize(s->stream));
    *t = s->t;
  }
  return BDG_OK;
}

// ---- streaming interface: batches of independent states (ensemble members, parameter sweeps, restarts) -----------
// The upload of the next state and the download of the previous result run on their own streams (both directions
// of the PCIe link at once) while the steps of the current state occupy the SMs:
//   stage_state(i+1) | commit_state(i), step(i) ... | get_state_async(i-1)
// The layout conversions run on the main stream (they are HBM-speed and short), only the raw copies are forked.
static int streaming_setup(bdg_sw2d *s) {
  if (s->in_stream) return BDG_OK;
  int rc;
  if ((rc = dev_alloc(s, &s->in_raw, 3 * s->nT))) return rc;
  if ((rc = dev_alloc(s, &s->dl_raw, 3 * s->nT))) return rc;
  CU(cudaStreamCreateWithFlags(&s->in_stream, cudaStreamNonBlocking));
  CU(cudaStreamCreateWithFlags(&s->dl_stream, cudaStreamNonBlocking));
  for (cudaEvent_t *ev : {&s->ev_in_done, &s->ev_in_free, &s->ev_dl_conv, &s->ev_dl_done}) CU(cudaEventCreateWithFlags(ev, cudaEventDisableTiming));
  return BDG_OK;
}

// Starts the upload of a state that bdg_sw2d_commit_state will make the live one.  The host arrays must stay valid
// (and should be pinned) until commit_state has been called and the handle synchronised or stepped.
extern "C" int bdg_sw2d_stage_state(bdg_sw2d *s, const double *h, const double *hu, const double *hv, double t) {
  if (!s || !h || !hu || !hv) return BDG_E_INVALID;
  CU(cudaSetDevice(s->device));
  int rc = streaming_setup(s);
  if (rc) return rc;
  if (s->staged) return fail(s, BDG_E_INVALID, "bdg_sw2d_stage_state: a staged state is waiting for bdg_sw2d_commit_state");
  CU(cudaStreamWaitEvent(s->in_stream, s->ev_in_free, 0));  // the previous commit has consumed in_raw (no-op the first time)
  const double *src[3] = {h, hu, hv};
  for (int c = 0; c < 3; ++c) CU(cudaMemcpyAsync(s->in_raw + c * s->nT, src[c], sizeof(double) * s->nT, cudaMemcpyHostToDevice, s->in_stream));
  CU(cudaEventRecord(s->ev_in_done, s->in_stream));
  s->staged = 1;
  s->staged_t = t;
  return BDG_OK;
}

// Stream-ordered bdg_sw2d_set_state of the staged arrays: everything enqueued later sees the new state.
extern "C" int bdg_sw2d_commit_state(bdg_sw2d *s) {
  if (!s) return BDG_E_INVALID;
  if (!s->staged) return fail(s, BDG_E_INVALID, "bdg_sw2d_commit_state without bdg_sw2d_stage_state");
  CU(cudaSetDevice(s->device));
  if (s->put_pending) {
    CU(cudaStreamSynchronize(s->comm_stream));
    s->put_pending = 0;
  }
  CU(cudaStreamWaitEvent(s->stream, s->ev_in_done, 0));
  double *q = s->q[s->cur];
  const int R = s->v.Np, KB = tile_kb(R);
  const long long K = s->v.K;
  for (int c = 0; c < 3; ++c) {
    k_rowmajor_to_elem<<<(unsigned)((K + KB - 1) / KB), 256, sizeof(double) * R * (KB + 1), s->stream>>>(s->in_raw + c * s->nT, q + c * s->v.stride, R, K, KB, s->d_perm);
    s->launches++;
  }
  CU(cudaGetLastError());
  CU(cudaEventRecord(s->ev_in_free, s->stream));
  CU(cudaMemsetAsync(s->res, 0, sizeof(double) * 3 * s->v.stride, s->stream));
  s->t = s->staged_t;
  s->staged = 0;
  s->ghost_valid = 0;
  s->lam_valid = 0;
  CU(cudaMemcpyAsync(s->d_scal + S_T, &s->t, sizeof(double), cudaMemcpyHostToDevice, s->stream));
  return BDG_OK;
}

// Snapshot of the live state behind everything enqueued so far; the copies into h, hu, hv ([Np][K], pinned) run on a
// side stream while later work proceeds.  bdg_sw2d_get_state_wait blocks until they have arrived.
extern "C" int bdg_sw2d_get_state_async(bdg_sw2d *s, double *h, double *hu, double *hv) {
  if (!s || !h || !hu || !hv) return BDG_E_INVALID;
  CU(cudaSetDevice(s->device));
  int rc = streaming_setup(s);
  if (rc) return rc;
  if (s->dl_pending) CU(cudaStreamWaitEvent(s->stream, s->ev_dl_done, 0));  // the previous snapshot has left dl_raw
  const double *q = s->q[s->cur];
  const int R = s->v.Np, KB = tile_kb(R);
  const long long K = s->v.K;
  for (int c = 0; c < 3; ++c) {
    k_elem_to_rowmajor<<<(unsigned)((K + KB - 1) / KB), 256, sizeof(double) * R * (KB + 1), s->stream>>>(q + c * s->v.stride, s->dl_raw + c * s->nT, R, K, KB, s->d_perm);
    s->launches++;
  }
  CU(cudaGetLastError());
  CU(cudaEventRecord(s->ev_dl_conv, s->stream));
  CU(cudaStreamWaitEvent(s->dl_stream, s->ev_dl_conv, 0));
  double *dst[3] = {h, hu, hv};
  for (int c = 0; c < 3; ++c) CU(cudaMemcpyAsync(dst[c], s->dl_raw + c * s->nT, sizeof(double) * s->nT, cudaMemcpyDeviceToHost, s->dl_stream));
  CU(cudaEventRecord(s->ev_dl_done, s->dl_stream));
  s->dl_pending = 1;
  return BDG_OK;
}

extern "C" int bdg_sw2d_get_state_wait(bdg_sw2d *s) {
  if (!s) return BDG_E_INVALID;
  if (!s->dl_pending) return BDG_OK;
  CU(cudaSetDevice(s->device));
  CU(cudaEventSynchronize(s->ev_dl_done));
  return BDG_OK;
}

// ========================================================================================== halo
// Per stage: pack the face traces the neighbours need (main stream) -> grouped ncclSend/ncclRecv on the
// comm stream straight into the ghost slots of the live state buffer; the caller overlaps the interior
// elements and waits on ev_recv before touching cut elements.
static int halo_begin(bdg_sw2d *s, double *q) {
  if (!s->comm || s->n_nbr == 0) return BDG_OK;
  const int np = s->v.full ? 4 : 3;
  cudaError_t e = launch_pack(s, q, np == 4);
  if (e != cudaSuccess) return fail(s, BDG_E_CUDA, "pack: %s", cudaGetErrorString(e));
  CU(cudaEventRecord(s->ev_pack, s->stream));
  CU(cudaStreamWaitEvent(s->comm_stream, s->ev_pack, 0));
  ncclComm_t comm = (ncclComm_t)s->comm;
  NC(ncclGroupStart());
  for (int i = 0; i < s->n_nbr; ++i) {
    const int so = s->h_send_off[i], sc = s->h_send_off[i + 1] - so;
    const int ro = s->h_recv_off[i], rcn = s->h_recv_off[i + 1] - ro;
    for (int c = 0; c < np; ++c) {
      if (sc > 0) NC(ncclSend(s->d_sendbuf + (size_t)np * so + (size_t)c * sc, sc, ncclDouble, s->h_nbr[i], comm, s->comm_stream));
      double *dst = (c < 3) ? q + c * s->v.stride + s->nT + ro : s->d_H + s->nT + ro;
      if (rcn > 0) NC(ncclRecv(dst, rcn, ncclDouble, s->h_nbr[i], comm, s->comm_stream));
    }
  }
  NC(ncclGroupEnd());
  CU(cudaEventRecord(s->ev_recv, s->comm_stream));
  return BDG_OK;
}
static int halo_wait(bdg_sw2d *s) {
  if (!s->comm || s->n_nbr == 0) return BDG_OK;
  CU(cudaStreamWaitEvent(s->stream, s->ev_recv, 0));
  return BDG_OK;
}
// max-allreduce of one device scalar across ranks (global LF speed, adaptive dt)
static int allreduce_max(bdg_sw2d *s, double *dptr) {
  if (!s->comm || s->nranks == 1) return BDG_OK;
  CU(cudaEventRecord(s->ev_pack, s->stream));
  CU(cudaStreamWaitEvent(s->comm_stream, s->ev_pack, 0));
  NC(ncclAllReduce(dptr, dptr, 1, ncclDouble, ncclMax, (ncclComm_t)s->comm, s->comm_stream));
  CU(cudaEventRecord(s->ev_recv, s->comm_stream));
  CU(cudaStreamWaitEvent(s->stream, s->ev_recv, 0));
  return BDG_OK;
}

extern "C" int bdg_nccl_unique_id(void *id_out) {
  static_assert(sizeof(ncclUniqueId) == BDG_NCCL_ID_BYTES, "ncclUniqueId size");
  ncclUniqueId id;
  if (ncclGetUniqueId(&id) != ncclSuccess) return BDG_E_NCCL;
  memcpy(id_out, &id, sizeof(id));
  return BDG_OK;
}

// Peer-memory halo setup (sw2d_halo.cu): all-gather CUDA IPC handles of the two state buffers and the flag
// array, plus each rank's ghost base per sender, over the communicator; open the neighbours' handles.
// Any failure leaves peer_ok = 0 and the NCCL send/recv path in charge.
struct PeerRecord {
  cudaIpcMemHandle_t q[2], flags;
  long long stride;
};
static int peer_setup(bdg_sw2d *s) {
  s->peer_ok = 0;
  const char *env = getenv("BDG_HALO");
  if (env && !strcmp(env, "nccl")) return BDG_OK;
  if (s->nranks < 2) return BDG_OK;
  // Only conditions that are IDENTICAL on every rank may return before the handshake below (BDG_HALO, nranks): a
  // rank-local one -- no fused path on this partition, more neighbours than the peer table holds, global-LF mode --
  // is published as stride = -1 instead, so that all ranks leave the collectives together and fall back to NCCL together.
  const bool local_ok = s->fused_ok && !s->v.lf_global && s->n_nbr <= kMaxNbr;
  ncclComm_t comm = (ncclComm_t)s->comm;
  const int P = s->nranks;
  int rc;
  if ((rc = dev_alloc(s, &s->d_flags, (size_t)P))) return rc;
  CU(cudaMemsetAsync(s->d_flags, 0, sizeof(unsigned long long) * P, s->stream));
  if ((rc = dev_alloc(s, &s->d_put_counter, 1))) return rc;
  CU(cudaMemsetAsync(s->d_put_counter, 0, sizeof(unsigned int), s->stream));
  // my record and my ghost bases: gbase[r] = first ghost slot (plane index) that sender r fills, or -1
  PeerRecord mine;
  memset(&mine, 0, sizeof(mine));
  bool ok = local_ok && cudaIpcGetMemHandle(&mine.q[0], s->q[0]) == cudaSuccess && cudaIpcGetMemHandle(&mine.q[1], s->q[1]) == cudaSuccess &&
            cudaIpcGetMemHandle(&mine.flags, s->d_flags) == cudaSuccess;
  if (!ok) cudaGetLastError();
  mine.stride = ok ? s->v.stride : -1;
  std::vector<long long> gbase(P, -1);
  for (int i = 0; i < s->n_nbr; ++i) gbase[s->h_nbr[i]] = (long long)s->nT + s->h_recv_off[i];
  std::vector<PeerRecord> all(P);
  std::vector<long long> gall((size_t)P * P);
  unsigned char *d_rec = nullptr;
  long long *d_g = nullptr;
  if ((rc = dev_alloc(s, &d_rec, sizeof(PeerRecord) * (P + 1)))) return rc;
  if ((rc = dev_alloc(s, &d_g, (size_t)P * (P + 1)))) return rc;
  CU(cudaMemcpyAsync(d_rec + sizeof(PeerRecord) * P, &mine, sizeof(mine), cudaMemcpyHostToDevice, s->stream));
  CU(cudaMemcpyAsync(d_g + (size_t)P * P, gbase.data(), sizeof(long long) * P, cudaMemcpyHostToDevice, s->stream));
  NC(ncclAllGather(d_rec + sizeof(PeerRecord) * P, d_rec, sizeof(PeerRecord), ncclChar, comm, s->stream));
  NC(ncclAllGather(d_g + (size_t)P * P, d_g, P, ncclInt64, comm, s->stream));
  CU(cudaMemcpyAsync(all.data(), d_rec, sizeof(PeerRecord) * P, cudaMemcpyDeviceToHost, s->stream));
  CU(cudaMemcpyAsync(gall.data(), d_g, sizeof(long long) * P * P, cudaMemcpyDeviceToHost, s->stream));
  CU(cudaStreamSynchronize(s->stream));
  cudaFree(d_rec);
  cudaFree(d_g);
  for (int r = 0; r < P; ++r)
    if (all[r].stride < 0) ok = false;  // every rank must take the same decision
  PeerTable pt;
  memset(&pt, 0, sizeof(pt));
  pt.n = ok ? s->n_nbr : 0;
  for (int i = 0; i < s->n_nbr && ok; ++i) {
    const int r = s->h_nbr[i];
    void *p0 = nullptr, *p1 = nullptr, *pf = nullptr;
    if (cudaIpcOpenMemHandle(&p0, all[r].q[0], cudaIpcMemLazyEnablePeerAccess) != cudaSuccess ||
        cudaIpcOpenMemHandle(&p1, all[r].q[1], cudaIpcMemLazyEnablePeerAccess) != cudaSuccess ||
        cudaIpcOpenMemHandle(&pf, all[r].flags, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
      cudaGetLastError();
      ok = false;
    }
    s->peer_base[0][i] = p0; s->peer_base[1][i] = p1; s->peer_base[2][i] = pf;
    pt.rank[i] = r;
    pt.q[0][i] = (double *)p0; pt.q[1][i] = (double *)p1;
    pt.stride[i] = all[r].stride;
    pt.flags[i] = (unsigned long long *)pf;
    if (gall[(size_t)r * P + s->rank] < 0) ok = false;
  }
  // the decision must be unanimous: a rank that could not map a peer forces everyone back to NCCL
  int *d_ok = nullptr;
  if ((rc = dev_alloc(s, &d_ok, 1))) return rc;
  int h_ok = ok ? 1 : 0;
  CU(cudaMemcpyAsync(d_ok, &h_ok, sizeof(int), cudaMemcpyHostToDevice, s->stream));
  NC(ncclAllReduce(d_ok, d_ok, 1, ncclInt, ncclMin, comm, s->stream));
  CU(cudaMemcpyAsync(&h_ok, d_ok, sizeof(int), cudaMemcpyDeviceToHost, s->stream));
  CU(cudaStreamSynchronize(s->stream));
  cudaFree(d_ok);
  if (!h_ok) return BDG_OK;
  std::vector<long long> dst((size_t)(s->n_send ? s->n_send : 1));
  for (int i = 0; i < s->n_nbr; ++i) {
    const long long base = gall[(size_t)s->h_nbr[i] * P + s->rank];
    for (int j = s->h_send_off[i]; j < s->h_send_off[i + 1]; ++j) dst[j] = base + (j - s->h_send_off[i]);
  }
  if ((rc = dev_upload(s, &s->d_put_dst, dst.data(), dst.size()))) return rc;
  CU(cudaStreamSynchronize(s->stream));
  s->peer = pt;
  s->put_ctas = s->n_send > 0 ? (s->n_send + 2047) / 2048 : 1;  // >= 8 entries per thread, at most one CTA per SM
  if (s->put_ctas > 148) s->put_ctas = 148;
  s->epoch = 0;
  s->ghost_valid = 0;
  s->peer_ok = 1;
  return BDG_OK;
}

extern "C" int bdg_sw2d_comm_init(bdg_sw2d *s, const void *idp) {
  if (!s || !idp) return BDG_E_INVALID;
  CU(cudaSetDevice(s->device));
  ncclUniqueId id;
  memcpy(&id, idp, sizeof(id));
  ncclComm_t comm;
  NC(ncclCommInitRank(&comm, s->nranks, id, s->rank));
  s->comm = (ncclComm *)comm;
  return peer_setup(s);
}

extern "C" int bdg_sw2d_halo_pack_to_host(bdg_sw2d *s, double *send_buf) {
  if (!s) return BDG_E_INVALID;
  if (s->n_send == 0) return BDG_OK;
  CU(cudaSetDevice(s->device));
  const int np = s->v.full ? 4 : 3;
  cudaError_t e = launch_pack(s, s->q[s->cur], np == 4);
  if (e != cudaSuccess) return fail(s, BDG_E_CUDA, "pack: %s", cudaGetErrorString(e));
  CU(cudaMemcpyAsync(send_buf, s->d_sendbuf, sizeof(double) * np * s->n_send, cudaMemcpyDeviceToHost, s->stream));
  CU(cudaStreamSynchronize(s->stream));
  return BDG_OK;
}

// recv_buf uses the send-buffer layout of the peers: per neighbour segment, np planes of cnt values
extern "C" int bdg_sw2d_halo_set_ghosts(bdg_sw2d *s, const double *recv_buf) {
  if (!s) return BDG_E_INVALID;
  CU(cudaSetDevice(s->device));
  const int np = s->v.full ? 4 : 3;
  double *q = s->q[s->cur];
  for (int i = 0; i < s->n_nbr; ++i) {
    const int ro = s->h_recv_off[i], cnt = s->h_recv_off[i + 1] - ro;
    for (int c = 0; c < np; ++c) {
      double *dst = (c < 3) ? q + c * s->v.stride + s->nT + ro : s->d_H + s->nT + ro;
      if (c == 3 && !s->d_H) continue;
      CU(cudaMemcpyAsync(dst, recv_buf + (size_t)np * ro + (size_t)c * cnt, sizeof(double) * cnt, cudaMemcpyHostToDevice, s->stream));
    }
  }
  CU(cudaStreamSynchronize(s->stream));
  return BDG_OK;
}

// ======================================================================================== tracer
extern "C" int bdg_sw2d_set_tracer(bdg_sw2d *s, const double *hN) {
  if (!s) return BDG_E_INVALID;
  CU(cudaSetDevice(s->device));
  if (!hN) {
    CU(cudaStreamSynchronize(s->stream));
    for (double **p : {&s->qN[0], &s->qN[1], &s->resN}) {
      if (*p) cudaFree(*p);
      *p = nullptr;
    }
    return BDG_OK;
  }
  if (s->v.lf_global) return fail(s, BDG_E_UNSUPPORTED, "tracer: needs the per-face Lax-Friedrichs speed (lf_mode = 0), the form of swhelpers/rhs.py");
  if (s->n_nbr > 0) return fail(s, BDG_E_UNSUPPORTED, "tracer: single rank only (no halo exchange of hN yet)");
  int rc;
  for (double **p : {&s->qN[0], &s->qN[1], &s->resN})
    if (!*p && (rc = dev_alloc(s, p, s->nT))) return rc;
  s->curN = 0;
  if ((rc = upload_transposed(s, hN, s->qN[0], s->v.Np, s->v.K))) return rc;
  CU(cudaMemsetAsync(s->resN, 0, sizeof(double) * s->nT, s->stream));
  CU(cudaMemsetAsync(s->qN[1], 0, sizeof(double) * s->nT, s->stream));
  CU(cudaStreamSynchronize(s->stream));
  return BDG_OK;
}

extern "C" int bdg_sw2d_get_tracer(bdg_sw2d *s, double *hN) {
  if (!s || !hN) return BDG_E_INVALID;
  if (!s->qN[0]) return fail(s, BDG_E_INVALID, "no tracer set");
  CU(cudaSetDevice(s->device));
  int rc = download_transposed(s, s->qN[s->curN], hN, s->v.Np, s->v.K);
  if (rc) return rc;
  CU(cudaStreamSynchronize(s->stream));
  return BDG_OK;
}

extern "C" int bdg_sw2d_rhs_tracer(bdg_sw2d *s, double *RHS4) {
  if (!s || !RHS4) return BDG_E_INVALID;
  if (!s->qN[0]) return fail(s, BDG_E_INVALID, "no tracer set");
  CU(cudaSetDevice(s->device));
  int rc;
  if (!s->rhs && (rc = dev_alloc(s, &s->rhs, 3 * s->nT))) return rc;
  cudaError_t e = launch_tracer(s, s->q[s->cur], s->qN[s->curN], nullptr, s->rhs, 0, 0.0, 0.0);
  if (e != cudaSuccess) return fail(s, BDG_E_CUDA, "tracer: %s", cudaGetErrorString(e));
  if ((rc = download_transposed(s, s->rhs, RHS4, s->v.Np, s->v.K))) return rc;
  CU(cudaStreamSynchronize(s->stream));
  return BDG_OK;
}

// ---- over-integrated RHS on curved triangles (sw2d_curved.cu; reference swhelpers/rhs.py:6-176) -------------------------
static void curved_free(bdg_sw2d *s) {
  CurvedOps *o = s->curved;
  if (!o) return;
  const void *ptrs[] = {o->V, o->Dr, o->Ds, o->cW, o->crx, o->csx, o->cry, o->csy, o->gInterp, o->gW, o->gnx, o->gny, o->mapM, o->mapP,
                        o->wall, o->J, o->mminv, o->chol_idx, o->chol, o->gq, o->CD, o->out};
  for (const void *p : ptrs)
    if (p) cudaFree(const_cast<void *>(p));
  delete o;
  s->curved = nullptr;
}

extern "C" int bdg_sw2d_set_curved(bdg_sw2d *s, const bdg_curved2d *c) {
  if (!s) return BDG_E_INVALID;
  CU(cudaSetDevice(s->device));
  CU(cudaStreamSynchronize(s->stream));
  curved_free(s);
  if (!c) return BDG_OK;
  const int Np = s->v.Np, K = s->v.K, Ncub = c->Ncub, NG3 = 3 * c->NGauss;
  if (s->v.Nf != 3) return fail(s, BDG_E_UNSUPPORTED, "the over-integrated RHS is built for triangles (the reference has no quad cubature context)");
  if (s->nranks > 1 || s->n_ghost > 0) return fail(s, BDG_E_UNSUPPORTED, "the over-integrated RHS runs on a single rank");
  if (Ncub <= 0 || c->NGauss <= 0 || Ncub > 1024 || NG3 > 768) return fail(s, BDG_E_INVALID, "bad sizes Ncub=%d NGauss=%d", Ncub, c->NGauss);
  if (!c->cubV || !c->cubDr || !c->cubDs || !c->cubW || !c->cub_rx || !c->cub_sx || !c->cub_ry || !c->cub_sy || !c->gInterp || !c->gW ||
      !c->gnx || !c->gny || !c->gmapM || !c->gmapP || !c->V || !c->J)
    return fail(s, BDG_E_INVALID, "NULL array in bdg_curved2d");
  if (c->n_curved < 0 || c->n_gmapW < 0 || (c->n_curved > 0 && (!c->MMChol || !c->curvedEls)) || (c->n_gmapW > 0 && !c->gmapW))
    return fail(s, BDG_E_INVALID, "curved element / wall lists: count without array");
  const long long nG = (long long)NG3 * K;
  if (nG >= (1LL << 31)) return fail(s, BDG_E_UNSUPPORTED, "3*NGauss*K exceeds int32 (the reference's index_type of gmapM / gmapP)");
  // every per-element table goes to DEVICE element order (sw2d_order.cu): perm[k_dev] = caller's element
  std::vector<int32_t> perm((size_t)K), inv((size_t)K);
  if (s->d_perm) {
    CU(cudaMemcpy(perm.data(), s->d_perm, sizeof(int32_t) * (size_t)K, cudaMemcpyDeviceToHost));
  } else {
    for (int k = 0; k < K; ++k) perm[k] = k;
  }
  for (int k = 0; k < K; ++k) inv[perm[k]] = k;
  std::vector<int32_t> mM((size_t)nG), mP((size_t)nG), cidx((size_t)K, -1);
  std::vector<uint8_t> wall((size_t)nG, 0);
  for (long long kd = 0; kd < K; ++kd)
    for (int g = 0; g < NG3; ++g) {
      const long long ic = g + (long long)NG3 * perm[kd];
      const long long a = c->gmapM[ic], b = c->gmapP[ic];
      if (a < 0 || a >= nG || b < 0 || b >= nG) return fail(s, BDG_E_INVALID, "gmapM/gmapP[%lld] out of range", ic);
      mM[g + NG3 * kd] = (int32_t)(a % NG3 + (long long)NG3 * inv[a / NG3]);
      mP[g + NG3 * kd] = (int32_t)(b % NG3 + (long long)NG3 * inv[b / NG3]);
    }
  for (int i = 0; i < c->n_gmapW; ++i) {
    const long long a = c->gmapW[i];
    if (a < 0 || a >= nG) return fail(s, BDG_E_INVALID, "gmapW[%d]=%lld out of range", i, a);
    wall[a % NG3 + (long long)NG3 * inv[a / NG3]] = 1;
  }
  std::vector<double> chol((size_t)(c->n_curved > 0 ? c->n_curved : 1) * Np * Np, 0.0), mminv((size_t)Np * Np);
  for (int ci = 0; ci < c->n_curved; ++ci) {
    const int kc = c->curvedEls[ci];
    if (kc < 0 || kc >= K) return fail(s, BDG_E_INVALID, "curvedEls[%d]=%d out of range", ci, kc);
    cidx[inv[kc]] = ci;
    for (int i = 0; i < Np * Np; ++i) chol[(size_t)ci * Np * Np + i] = c->MMChol[(size_t)i * K + kc];  // [Np][Np][K], k fastest
  }
  for (int i = 0; i < Np; ++i)
    for (int j = 0; j < Np; ++j) {
      double a = 0.0;
      for (int m = 0; m < Np; ++m) a += c->V[i * Np + m] * c->V[j * Np + m];  // V V^T (rhs.py:152-154)
      mminv[i * Np + j] = a;
    }
  CurvedOps *o = new (std::nothrow) CurvedOps();
  if (!o) return fail(s, BDG_E_INVALID, "out of host memory");
  memset(o, 0, sizeof(*o));
  s->curved = o;
  o->Ncub = Ncub;
  o->NG3 = NG3;
  int rc = BDG_OK;
  auto up = [&](const double **dst, const double *host, size_t n) {
    double *d = nullptr;
    if (!rc) rc = dev_upload(s, &d, host, n);
    *dst = d;
  };
  auto upT = [&](const double **dst, const double *host, int R) {  // [R][K] row-major -> [K][R] in device element order
    double *d = nullptr;
    if (!rc) rc = dev_alloc(s, &d, (size_t)R * K);
    *dst = d;
    if (!rc) rc = upload_transposed(s, host, d, R, K);
  };
  up(&o->V, c->cubV, (size_t)Ncub * Np); up(&o->Dr, c->cubDr, (size_t)Ncub * Np); up(&o->Ds, c->cubDs, (size_t)Ncub * Np);
  up(&o->gInterp, c->gInterp, (size_t)NG3 * Np); up(&o->mminv, mminv.data(), mminv.size()); up(&o->chol, chol.data(), chol.size());
  upT(&o->cW, c->cubW, Ncub); upT(&o->crx, c->cub_rx, Ncub); upT(&o->csx, c->cub_sx, Ncub); upT(&o->cry, c->cub_ry, Ncub); upT(&o->csy, c->cub_sy, Ncub);
  upT(&o->gW, c->gW, NG3); upT(&o->gnx, c->gnx, NG3); upT(&o->gny, c->gny, NG3); upT(&o->J, c->J, Np);
  {
    int32_t *d = nullptr;
    if (!rc) rc = dev_upload(s, &d, mM.data(), mM.size());
    o->mapM = d; d = nullptr;
    if (!rc) rc = dev_upload(s, &d, mP.data(), mP.size());
    o->mapP = d; d = nullptr;
    if (!rc) rc = dev_upload(s, &d, cidx.data(), cidx.size());
    o->chol_idx = d;
    uint8_t *w = nullptr;
    if (!rc) rc = dev_upload(s, &w, wall.data(), wall.size());
    o->wall = w;
  }
  if (!rc) rc = dev_alloc(s, &o->gq, (size_t)4 * nG);
  if (!rc) rc = dev_alloc(s, &o->CD, s->nT);
  if (!rc) rc = dev_alloc(s, &o->out, 4 * s->nT);
  cudaStreamSynchronize(s->stream);  // the host vectors above go out of scope
  if (rc) curved_free(s);
  return rc;
}

extern "C" int bdg_sw2d_rhs_curved(bdg_sw2d *s, const double *CD_nodal, double *RHS1, double *RHS2, double *RHS3, double *RHS4) {
  if (!s) return BDG_E_INVALID;
  if (!s->curved) return fail(s, BDG_E_INVALID, "bdg_sw2d_set_curved first");
  CU(cudaSetDevice(s->device));
  int rc;
  const double *CDn = nullptr;
  if (CD_nodal) {
    if ((rc = upload_transposed(s, CD_nodal, s->curved->CD, s->v.Np, s->v.K))) return rc;
    CDn = s->curved->CD;
  }
  cudaError_t e = launch_curved_rhs(s, s->q[s->cur], s->qN[0] ? s->qN[s->curN] : nullptr, CDn, s->curved->out);
  if (e != cudaSuccess) return fail(s, BDG_E_CUDA, "curved rhs: %s", cudaGetErrorString(e));
  double *dst[4] = {RHS1, RHS2, RHS3, RHS4};
  for (int i = 0; i < 4; ++i)
    if (dst[i]) {
      if ((rc = download_transposed(s, s->curved->out + (size_t)i * s->nT, dst[i], s->v.Np, s->v.K))) return rc;
      CU(cudaStreamSynchronize(s->stream));  // one staging buffer
    }
  CU(cudaStreamSynchronize(s->stream));
  return BDG_OK;
}

// The time loop of the curved driver (sw2d_curved.py:240-262), device resident: per step
//   R = Filt * RHS_curved(q);  q1 = q + dt/2 R;  R = Filt * RHS_curved(q1);  q += dt R;  t += dt
// for h, hu, hv and (if set) the tracer.  Filt = the context's filter when prm.filter = 1 (the driver always filters), else none.
extern "C" int bdg_sw2d_step_curved(bdg_sw2d *s, int32_t nsteps, double dt, const double *CD_nodal) {
  if (!s || nsteps < 0 || !(dt > 0.0)) return BDG_E_INVALID;
  if (!s->curved) return fail(s, BDG_E_INVALID, "bdg_sw2d_set_curved first");
  CU(cudaSetDevice(s->device));
  int rc;
  const double *CDn = nullptr;
  if (CD_nodal) {
    if ((rc = upload_transposed(s, CD_nodal, s->curved->CD, s->v.Np, s->v.K))) return rc;
    CDn = s->curved->CD;
  }
  const double *Filt = (s->prm.filter && s->d_Filt) ? s->d_Filt : nullptr;
  cudaError_t e;
  for (int st = 0; st < nsteps; ++st) {
    double *q = s->q[s->cur], *q1 = s->q[s->cur ^ 1];
    double *qN = s->qN[0] ? s->qN[s->curN] : nullptr, *qN1 = s->qN[0] ? s->qN[s->curN ^ 1] : nullptr;
    if ((e = launch_curved_rhs(s, q, qN, CDn, s->curved->out)) != cudaSuccess ||
        (e = launch_curved_update(s, Filt, s->curved->out, 0.5 * dt, q, qN, q1, qN1)) != cudaSuccess ||
        (e = launch_curved_rhs(s, q1, qN1, CDn, s->curved->out)) != cudaSuccess ||
        (e = launch_curved_update(s, Filt, s->curved->out, dt, q, qN, q, qN)) != cudaSuccess)
      return fail(s, BDG_E_CUDA, "curved step: %s", cudaGetErrorString(e));
    s->t += dt;
  }
  s->lam_valid = 0;
  s->ghost_valid = 0;
  if (nsteps > 0) CU(cudaMemcpyAsync(s->d_scal + S_T, &s->t, sizeof(double), cudaMemcpyHostToDevice, s->stream));
  return BDG_OK;
}

// tracer stage beside a flow stage: reads the PRE-stage flow state q[cur] (the flow kernels write the other buffer or,
// on the in-place LSERK4 two-launch path, run after this launch on the same stream)
static int enqueue_tracer_stage(bdg_sw2d *s, double a, double b, int mode) {
  if (!s->qN[0]) return BDG_OK;
  cudaError_t e = launch_tracer(s, s->q[s->cur], s->qN[s->curN], s->qN[s->curN ^ 1], nullptr, mode, a, b);
  if (e != cudaSuccess) return fail(s, BDG_E_CUDA, "tracer: %s", cudaGetErrorString(e));
  s->curN ^= 1;
  return BDG_OK;
}

// ====================================================================================== hot path
static int use_fused(const bdg_sw2d *s) {
  if (s->variant == 1) return 0;
  // variant 0 (auto) = the faster path as measured on B200 (profiles/r02_fused_orders.md): the fused single-launch kernel
  // for N <= 5 and N = 8; at N = 6, 7 (one CTA of four compute warps per SM) the two-launch DMMA path is ahead (66 vs 58 and
  // 64 vs 58 G DSU/s), so auto keeps it there; variant 2 still selects the fused kernel explicitly.
  if (s->variant == 0 && (s->v.Np == 28 || s->v.Np == 36)) return 0;
  return s->fused_ok;
}

// One RHS evaluation of the live state.  write_rhs: store RHS planes into rhs_out (bdg_sw2d_rhs);
// otherwise apply LSERK4 stage (a,b).  slot alternates the global-LF scalar between S_GMAX0/1.
// rk2 = 2, 3: first / second SSP-RK2 stage of the src/sw2d driver instead (the state is read from q[cur], the result
// goes to q[cur^1], which in the second stage still holds q^n, and the buffers swap).  last: the next RHS evaluation
// happens at t + dt (last stage of a time step).
static int enqueue_stage(bdg_sw2d *s, int write_rhs, double *rhs_out, double a, double b, int slot, int rk2 = 0, int last = 0) {
  if (s->v.full && !s->v.H) return fail(s, BDG_E_INVALID, "sources=1 needs bdg_sw2d_set_bathymetry first");
  double *q = s->q[s->cur];
  const int K = s->v.K;
  const bool halo = s->comm && s->n_nbr > 0;
  // interior/cut split rounded down to a whole 32-element tile of the fused kernel: computing a few interior
  // elements with the cut launch is harmless and keeps both launches free of partially active tiles
  const bool fused = use_fused(s);
  const int ksplit = halo ? (fused ? s->n_interior / 32 * 32 : s->n_interior) : K;
  FusedStage fs;
  fs.mode = rk2 ? rk2 : 1; fs.slot_next = -1; fs.slot_zero = -1; fs.last_stage = last;
  int rc;
  cudaError_t e;
  if (halo && s->peer_ok && fused) {
    // Peer-memory halo (sw2d_halo.cu): ghosts of the live buffer were stored by the neighbours during the
    // previous stage.  Cut elements first, then publish their new traces into the neighbours' next buffer,
    // then the interior elements -- the NVLink transfer hides behind the interior launch.
    double *qo = s->q[s->cur ^ 1];
    if (s->put_pending) {  // join a put that is still in flight before its source buffer can be written again
      CU(cudaStreamWaitEvent(s->stream, s->ev_recv, 0));
      s->put_pending = 0;
    }
    if (!s->ghost_valid) {  // first evaluation after set_state: exchange the traces of the live buffer itself
      if ((e = launch_put(s, q, s->cur, ++s->epoch, 0)) != cudaSuccess) return fail(s, BDG_E_CUDA, "put: %s", cudaGetErrorString(e));
      s->ghost_valid = 1;
    }
    if (write_rhs) {  // bdg_sw2d_rhs: one evaluation, nothing to overlap
      if ((e = launch_wait_epoch(s, s->epoch)) != cudaSuccess) return fail(s, BDG_E_CUDA, "wait: %s", cudaGetErrorString(e));
      if ((e = launch_fused_stage(s, q, qo, s->res, 0, K, a, b, write_rhs, rhs_out, slot, &fs)) != cudaSuccess)
        return fail(s, BDG_E_CUDA, "fused(rhs): %s", cudaGetErrorString(e));
      return BDG_OK;
    }
    // Fork: the high-priority side stream runs  wait-for-epoch -> cut elements -> put (publishes epoch+1)  while the
    // main stream runs the interior elements; the two launches touch disjoint elements of q[cur^1] / res and both
    // only read q[cur].  Join: the main stream waits for the side stream, so everything enqueued after this stage
    // (the next stage, get_state, the dt reduction ...) sees the complete buffer.
    cudaStream_t main_stream = s->stream;
    CU(cudaEventRecord(s->ev_pack, main_stream));
    CU(cudaStreamWaitEvent(s->comm_stream, s->ev_pack, 0));
    s->stream = s->comm_stream;
    e = launch_wait_epoch(s, s->epoch);
    if (e == cudaSuccess) e = launch_fused_stage(s, q, qo, s->res, ksplit, K, a, b, 0, nullptr, slot, &fs);
    if (e == cudaSuccess) e = launch_put(s, qo, s->cur ^ 1, ++s->epoch, 0);
    s->stream = main_stream;
    if (e != cudaSuccess) return fail(s, BDG_E_CUDA, "fused(cut)/put: %s", cudaGetErrorString(e));
    CU(cudaEventRecord(s->ev_recv, s->comm_stream));
    if ((e = launch_fused_stage(s, q, qo, s->res, 0, ksplit, a, b, 0, nullptr, slot, &fs)) != cudaSuccess)
      return fail(s, BDG_E_CUDA, "fused(interior): %s", cudaGetErrorString(e));
    CU(cudaStreamWaitEvent(main_stream, s->ev_recv, 0));
    if (!write_rhs) s->cur ^= 1;
    return BDG_OK;
  }
  if (s->put_pending) {
    CU(cudaStreamWaitEvent(s->stream, s->ev_recv, 0));
    s->put_pending = 0;
  }
  s->ghost_valid = 0;  // this path refreshes the ghosts itself (NCCL) and may update q in place
  if ((rc = halo_begin(s, q))) return rc;
  if (fused && s->v.full) {
    // Fused full physics (sw2d_fused.cu, FULL): ONE launch per stage.  The global LF speed of the state is carried over
    // from the previous stage's epilogue (three rotating slots: this stage's, the next one's, the one being re-armed);
    // only the first evaluation after a state change, and bdg_sw2d_rhs, run the reduction launch k_speed.
    if ((rc = halo_wait(s))) return rc;  // the star states of cut faces and the speed reduction need every ghost
    if (s->delta_dirty) {
      if ((e = fused_build_delta(s)) != cudaSuccess) return fail(s, BDG_E_CUDA, "star-state table: %s", cudaGetErrorString(e));
      s->delta_dirty = 0;
    }
    if (!s->lam_valid || write_rhs) {
      CU(cudaMemsetAsync(s->d_scal + S_LAM0, 0, 3 * sizeof(double), s->stream));
      s->lam_slot = 0;
      if ((e = launch_generic_speed(s, q, S_LAM0)) != cudaSuccess) return fail(s, BDG_E_CUDA, "speed: %s", cudaGetErrorString(e));
      if ((rc = allreduce_max(s, s->d_scal + S_LAM0))) return rc;
    }
    const int A = S_LAM0 + s->lam_slot, B = S_LAM0 + (s->lam_slot + 1) % 3, C3 = S_LAM0 + (s->lam_slot + 2) % 3;
    if (!write_rhs) { fs.slot_next = B; fs.slot_zero = C3; }
    if ((e = launch_fused_stage(s, q, s->q[s->cur ^ 1], s->res, 0, K, a, b, write_rhs, rhs_out, A, &fs)) != cudaSuccess)
      return fail(s, BDG_E_CUDA, "fused(full): %s", cudaGetErrorString(e));
    if (write_rhs) {
      s->lam_valid = 0;
    } else {
      if ((rc = allreduce_max(s, s->d_scal + B))) return rc;
      s->lam_slot = (s->lam_slot + 1) % 3;
      s->lam_valid = 1;
      s->cur ^= 1;
    }
    return BDG_OK;
  }
  if (s->v.lf_global) {  // needs every trace, ghosts included: no overlap in this mode
    if ((rc = halo_wait(s))) return rc;
    if ((e = launch_generic_speed(s, q, slot)) != cudaSuccess) return fail(s, BDG_E_CUDA, "speed: %s", cudaGetErrorString(e));
    if ((rc = allreduce_max(s, s->d_scal + slot))) return rc;
  }
  if (fused) {
    double *qo = s->q[s->cur ^ 1];
    if ((e = launch_fused_stage(s, q, qo, s->res, 0, ksplit, a, b, write_rhs, rhs_out, slot, &fs)) != cudaSuccess)
      return fail(s, BDG_E_CUDA, "fused(interior): %s", cudaGetErrorString(e));
    if (ksplit < K) {
      if ((rc = halo_wait(s))) return rc;
      if ((e = launch_fused_stage(s, q, qo, s->res, ksplit, K, a, b, write_rhs, rhs_out, slot, &fs)) != cudaSuccess)
        return fail(s, BDG_E_CUDA, "fused(cut): %s", cudaGetErrorString(e));
    } else if (halo && !s->v.lf_global) {
      if ((rc = halo_wait(s))) return rc;
    }
    if (!write_rhs) s->cur ^= 1;
  } else {
    if ((e = launch_generic_face(s, q, 0, ksplit, slot)) != cudaSuccess) return fail(s, BDG_E_CUDA, "face: %s", cudaGetErrorString(e));
    if (halo && !s->v.lf_global && (rc = halo_wait(s))) return rc;
    if ((e = launch_generic_face(s, q, ksplit, K, slot)) != cudaSuccess) return fail(s, BDG_E_CUDA, "face: %s", cudaGetErrorString(e));
    if ((e = launch_generic_volume(s, q, s->res, rhs_out, write_rhs ? 0 : (rk2 ? rk2 : 1), a, b, s->q[s->cur ^ 1])) != cudaSuccess)
      return fail(s, BDG_E_CUDA, "volume: %s", cudaGetErrorString(e));
    if (rk2 && !write_rhs) s->cur ^= 1;
  }
  return BDG_OK;
}

extern "C" int bdg_sw2d_rhs(bdg_sw2d *s, double *RHS1, double *RHS2, double *RHS3) {
  if (!s) return BDG_E_INVALID;
  CU(cudaSetDevice(s->device));
  int rc;
  if (!s->rhs && (rc = dev_alloc(s, &s->rhs, 3 * s->nT))) return rc;
  if (s->v.lf_global) CU(cudaMemsetAsync(s->d_scal + S_GMAX0, 0, 2 * sizeof(double), s->stream));
  if ((rc = enqueue_stage(s, 1, s->rhs, 0.0, 0.0, S_GMAX0))) return rc;
  double *dst[3] = {RHS1, RHS2, RHS3};
  for (int c = 0; c < 3; ++c)
    if (dst[c]) {
      if ((rc = download_transposed(s, s->rhs + c * s->nT, dst[c], s->v.Np, s->v.K))) return rc;
      CU(cudaStreamSynchronize(s->stream));
    }
  CU(cudaStreamSynchronize(s->stream));
  return BDG_OK;
}

static int enqueue_dt(bdg_sw2d *s) {
  cudaError_t e;
  int rc;
  if ((e = launch_dt_reduce(s, s->q[s->cur])) != cudaSuccess) return fail(s, BDG_E_CUDA, "dt_reduce: %s", cudaGetErrorString(e));
  if ((rc = allreduce_max(s, s->d_scal + S_DTMAX))) return rc;
  if ((e = launch_dt_finalize(s)) != cudaSuccess) return fail(s, BDG_E_CUDA, "dt_finalize: %s", cudaGetErrorString(e));
  return BDG_OK;
}

extern "C" int bdg_sw2d_timestep(bdg_sw2d *s, double *dt_out) {
  if (!s || !dt_out) return BDG_E_INVALID;
  CU(cudaSetDevice(s->device));
  CU(cudaMemsetAsync(s->d_scal + S_DTMAX, 0, sizeof(double), s->stream));
  int rc = enqueue_dt(s);
  if (rc) return rc;
  CU(cudaMemcpyAsync(dt_out, s->d_scal + S_DT, sizeof(double), cudaMemcpyDeviceToHost, s->stream));
  CU(cudaStreamSynchronize(s->stream));
  return BDG_OK;
}

extern "C" int bdg_sw2d_step(bdg_sw2d *s, int32_t nsteps, double dt) {
  if (!s || nsteps < 0) return BDG_E_INVALID;
  CU(cudaSetDevice(s->device));
  const bool adaptive = !(dt > 0.0);
  const bool device_time = adaptive || (s->v.full && s->prm.tide_amp != 0.0);
  int rc;
  cudaError_t e;
  if (!adaptive) CU(cudaMemcpyAsync(s->d_scal + S_DT, &dt, sizeof(double), cudaMemcpyHostToDevice, s->stream));
  if (s->v.lf_global) CU(cudaMemsetAsync(s->d_scal + S_GMAX0, 0, 2 * sizeof(double), s->stream));
  if (adaptive) {
    CU(cudaMemsetAsync(s->d_scal + S_DTMAX, 0, sizeof(double), s->stream));
    s->adaptive_used = 1;
  }
  int slot = S_GMAX0;
  for (int st = 0; st < nsteps; ++st) {
    if (adaptive && (rc = enqueue_dt(s))) return rc;
    for (int stage = 0; stage < 5; ++stage) {  // reference src/advec1d/main.cpp:92-102
      if ((rc = enqueue_tracer_stage(s, rk4a[stage], rk4b[stage], 1))) return rc;
      if ((rc = enqueue_stage(s, 0, nullptr, rk4a[stage], rk4b[stage], slot, 0, stage == 4))) return rc;
      slot ^= 1;
    }
    if (device_time) {
      if ((e = launch_advance_time(s)) != cudaSuccess) return fail(s, BDG_E_CUDA, "advance: %s", cudaGetErrorString(e));
    } else {
      s->t += dt;
    }
  }
  if (!device_time && nsteps > 0) CU(cudaMemcpyAsync(s->d_scal + S_T, &s->t, sizeof(double), cudaMemcpyHostToDevice, s->stream));
  return BDG_OK;
}

// The time loop of the src/sw2d driver (reference src/sw2d/main.cpp:192-244): per step dt = computeTimeStep (dt <= 0)
// or the caller's, two RHS evaluations at the same time t in Heun's SSP-RK2 form, the sponge relaxation of hu, hv
// after each of them, t += dt.  Four launches per step (face + volume per stage), nothing leaves the device.
extern "C" int bdg_sw2d_step_ssprk2(bdg_sw2d *s, int32_t nsteps, double dt) {
  if (!s || nsteps < 0) return BDG_E_INVALID;
  CU(cudaSetDevice(s->device));
  const bool adaptive = !(dt > 0.0);
  const bool device_time = adaptive || (s->v.full && s->prm.tide_amp != 0.0);
  int rc;
  cudaError_t e;
  if (!adaptive) CU(cudaMemcpyAsync(s->d_scal + S_DT, &dt, sizeof(double), cudaMemcpyHostToDevice, s->stream));
  if (s->v.lf_global) CU(cudaMemsetAsync(s->d_scal + S_GMAX0, 0, 2 * sizeof(double), s->stream));
  if (adaptive) {
    CU(cudaMemsetAsync(s->d_scal + S_DTMAX, 0, sizeof(double), s->stream));
    s->adaptive_used = 1;
  }
  int slot = S_GMAX0;
  for (int st = 0; st < nsteps; ++st) {
    if (adaptive && (rc = enqueue_dt(s))) return rc;
    for (int stage = 0; stage < 2; ++stage) {
      if ((rc = enqueue_tracer_stage(s, 0.0, 1.0, 2 + stage))) return rc;
      if ((rc = enqueue_stage(s, 0, nullptr, 0.0, 1.0, slot, 2 + stage, stage == 1))) return rc;
      slot ^= 1;
    }
    if (device_time) {
      if ((e = launch_advance_time(s)) != cudaSuccess) return fail(s, BDG_E_CUDA, "advance: %s", cudaGetErrorString(e));
    } else {
      s->t += dt;
    }
  }
  if (!device_time && nsteps > 0) CU(cudaMemcpyAsync(s->d_scal + S_T, &s->t, sizeof(double), cudaMemcpyHostToDevice, s->stream));
  return BDG_OK;
}

extern "C" int bdg_sw2d_sync(bdg_sw2d *s) {
  if (!s) return BDG_E_INVALID;
  CU(cudaSetDevice(s->device));
  CU(cudaStreamSynchronize(s->stream));
  CU(cudaStreamSynchronize(s->comm_stream));
  return BDG_OK;
}

extern "C" int bdg_sw2d_check(bdg_sw2d *s, double *eta_max) {
  if (!s) return BDG_E_INVALID;
  CU(cudaSetDevice(s->device));
  static_assert(S_NANFLAG == S_ETAMAX + 1 && S_HALO_TIMEOUT == S_ETAMAX + 2, "scalar slots");
  CU(cudaMemsetAsync(s->d_scal + S_ETAMAX, 0, 2 * sizeof(double), s->stream));  // S_HALO_TIMEOUT stays: it is sticky
  cudaError_t e = launch_check(s, s->q[s->cur]);
  if (e != cudaSuccess) return fail(s, BDG_E_CUDA, "check: %s", cudaGetErrorString(e));
  double r[3];
  CU(cudaMemcpyAsync(r, s->d_scal + S_ETAMAX, 3 * sizeof(double), cudaMemcpyDeviceToHost, s->stream));
  CU(cudaStreamSynchronize(s->stream));
  if (eta_max) *eta_max = r[1] != 0.0 ? NAN : r[0];
  if (r[2] != 0.0)  // a cut launch ran on stale ghost traces: every result since then is invalid
    return fail(s, BDG_E_NCCL, "halo exchange timed out: a neighbour rank never published its face traces (peer-memory halo)");
  if (r[1] != 0.0 || r[0] > 1e8) return fail(s, BDG_E_UNSTABLE, "A numerical instability has occurred!");  // reference main.cpp:238-240
  return BDG_OK;
}

// positivityPreservingLimiter2D of the Python drivers (reference swhelpers/limiters.py:3-21) on the live state
extern "C" int bdg_sw2d_limit_positivity(bdg_sw2d *s) {
  if (!s) return BDG_E_INVALID;
  if (s->v.Np > 64) return fail(s, BDG_E_UNSUPPORTED, "positivity limiter: Np = %d > 64", s->v.Np);
  CU(cudaSetDevice(s->device));
  if (s->put_pending) {
    CU(cudaStreamWaitEvent(s->stream, s->ev_recv, 0));
    s->put_pending = 0;
  }
  cudaError_t e = launch_positivity_limiter(s, s->q[s->cur]);
  if (e != cudaSuccess) return fail(s, BDG_E_CUDA, "limiter: %s", cudaGetErrorString(e));
  s->ghost_valid = 0;  // neighbours hold traces of the unlimited state
  s->lam_valid = 0;
  return BDG_OK;
}

// ======================================================================================== output
// SURVEY.md 8f rank 2 (sw2d_output.cu).  interp [n_out][Np] row-major: IM of splitElements (or any other interpolation).
extern "C" int bdg_sw2d_set_output(bdg_sw2d *s, const double *interp, int32_t n_out) {
  if (!s || !interp || n_out < 1 || n_out > 4096) return BDG_E_INVALID;
  CU(cudaSetDevice(s->device));
  if (s->out_stream) CU(cudaStreamSynchronize(s->out_stream));
  if (s->d_IM) cudaFree(s->d_IM);
  if (s->d_out) cudaFree(s->d_out);
  s->d_IM = s->d_out = nullptr;
  s->out_n = 0;
  int rc;
  if ((rc = dev_upload(s, &s->d_IM, interp, (size_t)n_out * s->v.Np))) return rc;
  if ((rc = dev_alloc(s, &s->d_out, 3 * (size_t)n_out * s->v.K))) return rc;
  if (!s->out_stream) {
    CU(cudaStreamCreateWithFlags(&s->out_stream, cudaStreamNonBlocking));
    CU(cudaEventCreateWithFlags(&s->ev_out_state, cudaEventDisableTiming));
    CU(cudaEventCreateWithFlags(&s->ev_out_kernel, cudaEventDisableTiming));
    CU(cudaEventCreateWithFlags(&s->ev_out_done, cudaEventDisableTiming));
  }
  CU(cudaStreamSynchronize(s->stream));
  s->out_n = n_out;
  return BDG_OK;
}

// Enqueues, behind everything already enqueued on this handle: eta = h - H, u = hu/h, v = hv/h of the live state,
// each interpolated with interp, then their copies into the host arrays ([n_out][K] row-major; pinned memory makes
// the copies asynchronous).  Only the interpolation kernel holds up later steps; the copies overlap them.
extern "C" int bdg_sw2d_output_begin(bdg_sw2d *s, double *eta, double *u, double *v) {
  if (!s) return BDG_E_INVALID;
  if (!s->out_n) return fail(s, BDG_E_INVALID, "bdg_sw2d_output_begin needs bdg_sw2d_set_output first");
  CU(cudaSetDevice(s->device));
  CU(cudaEventRecord(s->ev_out_state, s->stream));
  CU(cudaStreamWaitEvent(s->out_stream, s->ev_out_state, 0));
  cudaError_t e = launch_output_fields(s, s->q[s->cur], s->out_stream);
  if (e != cudaSuccess) return fail(s, BDG_E_CUDA, "output: %s", cudaGetErrorString(e));
  CU(cudaEventRecord(s->ev_out_kernel, s->out_stream));
  CU(cudaStreamWaitEvent(s->stream, s->ev_out_kernel, 0));  // the state buffer may be overwritten from here on
  const size_t plane = (size_t)s->out_n * s->v.K;
  double *dst[3] = {eta, u, v};
  for (int c = 0; c < 3; ++c)
    if (dst[c]) CU(cudaMemcpyAsync(dst[c], s->d_out + c * plane, sizeof(double) * plane, cudaMemcpyDeviceToHost, s->out_stream));
  CU(cudaEventRecord(s->ev_out_done, s->out_stream));
  return BDG_OK;
}

extern "C" int bdg_sw2d_output_wait(bdg_sw2d *s) {
  if (!s) return BDG_E_INVALID;
  if (!s->out_n) return BDG_OK;
  CU(cudaSetDevice(s->device));
  CU(cudaEventSynchronize(s->ev_out_done));
  return BDG_OK;
}

// ====================================================================================== plumbing
extern "C" int bdg_sw2d_set_variant(bdg_sw2d *s, int32_t variant) {
  if (!s || variant < 0 || variant > 2) return BDG_E_INVALID;
  if (variant == 2 && !s->fused_ok) return fail(s, BDG_E_UNSUPPORTED, "fused path unavailable for this context (needs affine triangles, uniform face tags, no filter, supported order)");
  s->variant = variant;
  s->ghost_valid = 0;
  s->lam_valid = 0;
  return BDG_OK;
}
extern "C" int bdg_sw2d_get_variant(const bdg_sw2d *s) { return s ? (use_fused(s) ? 2 : 1) : -1; }
extern "C" int64_t bdg_sw2d_launch_count(const bdg_sw2d *s) { return s ? s->launches : 0; }
extern "C" void *bdg_sw2d_stream(bdg_sw2d *s) { return s ? (void *)s->stream : nullptr; }

// B_alg of SURVEY.md §8d: q r+w and res r+w (96*Np), vmapP (4*Nfn), geometry, BC tag per face,
// plus the static bathymetry reads in full mode.
extern "C" double bdg_sw2d_algorithmic_bytes_per_stage(const bdg_sw2d *s) {
  if (!s) return 0.0;
  const double Np = s->v.Np, Nfn = s->v.Nfn, Nf = s->v.Nf;
  double per = 96.0 * Np + 4.0 * Nfn + 4.0 * Nf;
  per += s->is_affine ? (4 * 8 + Nf * 3 * 8) : (32.0 * Np + 24.0 * Nfn);
  if (s->v.full) per += 16.0 * Np + 16.0 * Nfn;
  return per * (double)s->v.K;
}
