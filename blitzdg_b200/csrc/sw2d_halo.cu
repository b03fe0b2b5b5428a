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

// ONE CTA: thread t handles send entries t, t+1024, ...: trace node send_nodes[j] of fields h,hu,hv -> ghost slot
// put_dst[j] of the neighbour seg_of[j]; then a single system-scope fence and the epoch is published to every
// neighbour.  (A halo is a few thousand nodes: one CTA keeps the fence + flag protocol trivial and short.)
__global__ void __launch_bounds__(1024) k_put(PeerTable pt, const double *__restrict__ q, long long stride,
                                              const int32_t *__restrict__ send_nodes, const int32_t *__restrict__ seg_of,
                                              const long long *__restrict__ put_dst, int n_send, int buf, int my_rank,
                                              unsigned long long epoch) {
  for (int j = threadIdx.x; j < n_send; j += blockDim.x) {
    const int sg = seg_of[j];
    const long long m = send_nodes[j], d = put_dst[j], ps = pt.stride[sg];
    double *dst = pt.q[buf][sg];
    dst[d] = q[m];
    dst[ps + d] = q[stride + m];
    dst[2 * ps + d] = q[2 * stride + m];
  }
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x < pt.n) st_release_sys(pt.flags[threadIdx.x] + my_rank, epoch);
}

// lanes 0..n-1 wait until neighbour nbr[lane] has published `epoch`; ~60 s budget, then the NaN flag is raised
// (bdg_sw2d_check reports it) instead of hanging the GPU
__global__ void k_wait_epoch(const unsigned long long *flags, PeerTable pt, unsigned long long epoch, double *scal) {
  const int i = threadIdx.x;
  if (i >= pt.n) return;
  const unsigned long long *f = flags + pt.rank[i];
  for (long long spin = 0; spin < (1LL << 28); ++spin) {
    if (ld_acquire_sys(f) >= epoch) return;
    __nanosleep(200);
  }
  scal[S_NANFLAG] = 1.0;
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
  k_put<<<1, 1024, 0, st>>>(s->peer, q, s->v.stride, nodes, seg, s->d_put_dst, s->n_send, buf, s->rank, epoch);
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
