// Face-trace halo exchange over NVLink peer memory, the fused path's alternative to
// ncclSend/ncclRecv: the sender's kernel stores the traces its neighbours need DIRECTLY into their ghost
// slots (peer pointers from CUDA IPC handles exchanged once at comm_init) and then publishes an epoch
// number in the receiver's flag word; the receiver's stream runs a one-warp kernel that waits for the epoch
// before the cut-element launch.  No SM-resident communication kernel competes with the persistent
// stage kernel, nothing is staged, and the transfer hides behind the interior-element launch.
//
// Ordering per stage (capi.cu enqueue_stage): wait(epoch) -> cut elements (read ghosts of buffer b, write
// buffer b^1) -> put(buffer b^1 traces into the peers' buffer b^1 ghosts, epoch+1) -> interior elements.
// A peer can only overwrite my buffer-b ghosts after it has seen my epoch+1, which I publish after my
// cut launch -- the only reader of those ghosts -- has finished.
#include "sw2d_common.cuh"

namespace {

__device__ __forceinline__ void st_release_sys(unsigned long long *p, unsigned long long v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long *p) {
  unsigned long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}

// Grid-stride over the send list: entry j = trace node send_nodes[j] of fields h,hu,hv -> ghost slot put_dst[j] of
// the neighbour seg_of[j].  Every CTA fences its own stores at system scope and counts itself done; the LAST CTA to
// finish publishes the epoch to every neighbour (and re-arms the counter), so a METIS cut of 50 k+ nodes is spread
// over many SMs instead of being serialised through one block, while the flag protocol stays a single release store
// per neighbour.
__global__ void __launch_bounds__(256) k_put(PeerTable pt, const double *__restrict__ q, long long stride,
                                             const int32_t *__restrict__ send_nodes, const int32_t *__restrict__ seg_of,
                                             const long long *__restrict__ put_dst, int n_send, int buf, int my_rank,
                                             unsigned long long epoch, unsigned int *counter) {
  for (int j = blockIdx.x * blockDim.x + threadIdx.x; j < n_send; j += gridDim.x * blockDim.x) {
    const int sg = seg_of[j];
    const long long m = send_nodes[j], d = put_dst[j], ps = pt.stride[sg];
    double *dst = pt.q[buf][sg];
    dst[d] = q[m];
    dst[ps + d] = q[stride + m];
    dst[2 * ps + d] = q[2 * stride + m];
  }
  __threadfence_system();  // my stores are visible system-wide before my arrival below can be observed
  __syncthreads();
  __shared__ int last;
  if (threadIdx.x == 0) {
    last = atomicAdd(counter, 1u) == gridDim.x - 1;
    if (last) *counter = 0u;  // the next put on this stream starts from zero (launches of one handle never overlap)
  }
  __syncthreads();
  if (last) {
    __threadfence_system();  // acquire side of the counter: every other CTA's fenced stores precede the flags
    if (threadIdx.x < pt.n) st_release_sys(pt.flags[threadIdx.x] + my_rank, epoch);
  }
}

// lanes 0..n-1 wait until neighbour nbr[lane] has published `epoch`; ~60 s budget, then the sticky S_HALO_TIMEOUT
// flag is raised (bdg_sw2d_check returns BDG_E_NCCL from then on) instead of hanging the GPU
__global__ void k_wait_epoch(const unsigned long long *flags, PeerTable pt, unsigned long long epoch, double *scal) {
  const int i = threadIdx.x;
  if (i >= pt.n) return;
  const unsigned long long *f = flags + pt.rank[i];
  for (long long spin = 0; spin < (1LL << 28); ++spin) {
    if (ld_acquire_sys(f) >= epoch) return;
    __nanosleep(200);
  }
  scal[S_HALO_TIMEOUT] = 1.0;
}

}  // namespace

// side != 0: on the high-priority communication stream, ordered after everything enqueued on the main stream so
// far; the main stream does not wait for it (the next stage does, see capi.cu)
cudaError_t launch_put(bdg_sw2d *s, const double *q, int buf, unsigned long long epoch, int side) {
  if (s->n_send == 0) return cudaSuccess;
  cudaStream_t st = s->stream;
  if (side) {
    cudaError_t e;
    if ((e = cudaEventRecord(s->ev_pack, s->stream)) != cudaSuccess) return e;
    if ((e = cudaStreamWaitEvent(s->comm_stream, s->ev_pack, 0)) != cudaSuccess) return e;
    st = s->comm_stream;
  }
  const int32_t *nodes = s->d_send_nodes, *seg = s->d_send_nodes + s->n_send;
  k_put<<<s->put_ctas, 256, 0, st>>>(s->peer, q, s->v.stride, nodes, seg, s->d_put_dst, s->n_send, buf, s->rank, epoch, s->d_put_counter);
  s->launches++;
  if (side) {
    cudaError_t e = cudaEventRecord(s->ev_recv, s->comm_stream);
    if (e != cudaSuccess) return e;
    s->put_pending = 1;
  }
  return cudaGetLastError();
}

cudaError_t launch_wait_epoch(bdg_sw2d *s, unsigned long long epoch) {
  if (s->n_nbr == 0) return cudaSuccess;
  k_wait_epoch<<<1, 32, 0, s->stream>>>(s->d_flags, s->peer, epoch, s->d_scal);
  s->launches++;
  return cudaGetLastError();
}
