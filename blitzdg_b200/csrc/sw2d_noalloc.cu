// bdg_sw2d_timestep_noalloc (include/blitzdg_b200.h): the time-step reduction of bdg_sw2d_timestep for callers that
// cannot allocate -- code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc.
//
// No counterpart in the reference.  The handle-based bdg_sw2d_timestep
// (capi.cu, k_dt_reduce in sw2d_generic.cu) reduces into the handle's cudaMalloc'ed scalar block; this variant does the
// same arithmetic on the CALLER's device arrays and keeps its scratch -- one partial maximum per thread block and the
// ticket word that elects the last block -- in a file-scope static __device__ array.  Nothing here calls cudaMalloc,
// cudaFree or any allocating library routine; the existing code paths are untouched.
//
// The array is shared by every call on a device and keeps the last call's contents: the ticket word is cleared on the
// stream before each launch, every partial that is read is written first by the same launch.  Calls that overlap on the
// GPU (two streams, two host threads) race on it: serialise them.
#include <cstdarg>
#include <cstdio>
#include <mutex>

#include "../../include/blitzdg_b200.h"

namespace {

constexpr int kThreads = 256;
constexpr int kItems = 16;                      // face nodes per thread
constexpr int kChunk = kThreads * kItems;       // face nodes per block: 4096
constexpr int kMaxBlocks = BDG_NOALLOC_MAX_FACE_NODES / kChunk;  // 65536 partial maxima
constexpr int kHeader = 32;                     // 256 B in front of the partials; word 0 = ticket counter
constexpr int kMaxDevices = 64;

static_assert((long long)kMaxBlocks * kChunk == BDG_NOALLOC_MAX_FACE_NODES, "limit must be a whole number of chunks");

// 512.25 KB, held by every GPU that loads the library until the process exits
static __device__ __align__(256) double g_noalloc_scratch[kHeader + kMaxBlocks];

// device address of g_noalloc_scratch per device: in host code the symbol's name is a host-side stand-in, only
// cudaGetSymbolAddress (with that device current) gives the pointer a kernel may dereference
double *g_scratch_ptr[kMaxDevices];
std::mutex g_scratch_mutex;

__device__ __forceinline__ double max_pos(double a, double b) {  // both >= 0 or NaN: the integer order of k_dt_reduce
  return __double_as_longlong(a) > __double_as_longlong(b) ? a : b;
}

__device__ __forceinline__ double block_max_pos(double v, double *sm) {
  for (int o = 16; o > 0; o >>= 1) v = max_pos(v, __shfl_xor_sync(0xffffffffu, v, o));
  if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = v;
  __syncthreads();
  if (threadIdx.x < 32) {
    v = threadIdx.x < kThreads / 32 ? sm[threadIdx.x] : 0.0;
    for (int o = 16; o > 0; o >>= 1) v = max_pos(v, __shfl_xor_sync(0xffffffffu, v, o));
  }
  __syncthreads();
  return v;  // valid in warp 0
}

// max over face nodes of Fscale*(|u| + sqrt(g h)) (reference src/sw2d/main.cpp:259-276, as k_dt_reduce), inputs in the
// reference's native layouts: h, hu, hv [Np][K] and Fscale [Nfn][K] row-major, fmask [Nfn] = local node of a face slot.
// Face node i = sl*K + k: consecutive threads read consecutive k, every load is coalesced.
__global__ void __launch_bounds__(kThreads) k_dt_noalloc(const double *__restrict__ h, const double *__restrict__ hu,
                                                         const double *__restrict__ hv, const double *__restrict__ Fscale,
                                                         const int32_t *__restrict__ fmask, int Np, long long K, long long nF,
                                                         double g, double CFL, double nn, double *scratch, double *__restrict__ dt_out) {
  __shared__ double sm[kThreads / 32];
  __shared__ bool last;
  double *partial = scratch + kHeader;
  unsigned long long *ticket = reinterpret_cast<unsigned long long *>(scratch);
  const long long base = (long long)blockIdx.x * kChunk;
  double mx = 0.0;
#pragma unroll 4
  for (int it = 0; it < kItems; ++it) {
    const long long i = base + it * kThreads + threadIdx.x;
    if (i < nF) {
      const long long sl = i / K, k = i - sl * K;
      int n = fmask[sl];
      n = n < 0 ? 0 : (n >= Np ? Np - 1 : n);  // a corrupt Fmask must not fault
      const long long m = n * K + k;
      const double hh = h[m], u = hu[m] / hh, w = hv[m] / hh;
      const double val = Fscale[i] * (sqrt(u * u + w * w) + sqrt(g * hh));
      mx = max_pos(val, mx);
    }
  }
  mx = block_max_pos(mx, sm);
  if (threadIdx.x == 0) {
    partial[blockIdx.x] = mx;
    __threadfence();
    last = atomicAdd(ticket, 1ULL) == gridDim.x - 1;
  }
  __syncthreads();
  if (!last) return;
  __threadfence();
  mx = 0.0;
  for (int b = threadIdx.x; b < (int)gridDim.x; b += kThreads) mx = max_pos(__ldcg(partial + b), mx);
  mx = block_max_pos(mx, sm);
  if (threadIdx.x == 0) {
    *dt_out = CFL / (nn * mx);  // nn = (N+1)^2 * 0.5, the operation order of k_dt_finalize
    *ticket = 0ULL;        // leave the word as the next call needs it
  }
}

int fail(char *err, int cap, int code, const char *fmt, ...) {
  if (err && cap > 0) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(err, (size_t)cap, fmt, ap);
    va_end(ap);
  }
  return code;
}

// once per device, with that device current
int scratch_for(int device, double **out, char *err, int cap) {
  std::lock_guard<std::mutex> lock(g_scratch_mutex);
  if (!g_scratch_ptr[device]) {
    double *p = nullptr;
    cudaError_t e = cudaGetSymbolAddress((void **)&p, g_noalloc_scratch);
    if (e != cudaSuccess || !p)
      return fail(err, cap, BDG_E_CUDA, "bdg_sw2d_timestep_noalloc: cudaGetSymbolAddress on device %d: %s", device, cudaGetErrorString(e));
    g_scratch_ptr[device] = p;
  }
  *out = g_scratch_ptr[device];
  return BDG_OK;
}

bool on_device(const void *p, int device) {
  cudaPointerAttributes a;
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
    cudaGetLastError();
    return false;
  }
  return a.type == cudaMemoryTypeManaged || (a.type == cudaMemoryTypeDevice && a.device == device);
}

}  // namespace

extern "C" int bdg_sw2d_timestep_noalloc(int32_t device, int32_t N, int32_t Np, int32_t Nfp, int32_t Nfaces, int32_t K, double g,
                                         double CFL, const double *d_h, const double *d_hu, const double *d_hv,
                                         const double *d_Fscale, const int32_t *d_fmask, double *d_dt_out, void *stream,
                                         char *err, int32_t err_cap) {
  if (err && err_cap > 0) err[0] = 0;
  if (device < 0 || device >= kMaxDevices) return fail(err, err_cap, BDG_E_INVALID, "bdg_sw2d_timestep_noalloc: device %d outside 0..%d", device, kMaxDevices - 1);
  if (N < 1 || Np <= 0 || Nfp <= 0 || K <= 0 || (Nfaces != 3 && Nfaces != 4))
    return fail(err, err_cap, BDG_E_INVALID, "bdg_sw2d_timestep_noalloc: bad sizes N=%d Np=%d Nfp=%d Nfaces=%d K=%d", N, Np, Nfp, Nfaces, K);
  if (!d_h || !d_hu || !d_hv || !d_Fscale || !d_fmask || !d_dt_out) return fail(err, err_cap, BDG_E_INVALID, "bdg_sw2d_timestep_noalloc: NULL array");
  const long long nF = (long long)Nfaces * Nfp * K;
  if (nF > (long long)BDG_NOALLOC_MAX_FACE_NODES)
    return fail(err, err_cap, BDG_E_UNSUPPORTED,
                "bdg_sw2d_timestep_noalloc: Nfaces*Nfp*K = %lld face nodes exceed BDG_NOALLOC_MAX_FACE_NODES = %lld (the static scratch "
                "holds one partial maximum per %d face nodes); use bdg_sw2d_timestep",
                nF, (long long)BDG_NOALLOC_MAX_FACE_NODES, kChunk);
  cudaError_t e = cudaSetDevice(device);
  if (e != cudaSuccess) return fail(err, err_cap, BDG_E_CUDA, "bdg_sw2d_timestep_noalloc: cudaSetDevice(%d): %s", device, cudaGetErrorString(e));
  const void *ptrs[6] = {d_h, d_hu, d_hv, d_Fscale, d_fmask, d_dt_out};
  const char *names[6] = {"d_h", "d_hu", "d_hv", "d_Fscale", "d_fmask", "d_dt_out"};
  for (int i = 0; i < 6; ++i)
    if (!on_device(ptrs[i], device)) return fail(err, err_cap, BDG_E_INVALID, "bdg_sw2d_timestep_noalloc: %s is not memory of device %d", names[i], device);
  double *scratch = nullptr;
  int rc = scratch_for(device, &scratch, err, err_cap);
  if (rc) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  e = cudaMemsetAsync(scratch, 0, sizeof(double) * kHeader, st);  // the ticket word must start at zero
  if (e != cudaSuccess) return fail(err, err_cap, BDG_E_CUDA, "bdg_sw2d_timestep_noalloc: clearing the scratch header: %s", cudaGetErrorString(e));
  const int blocks = (int)((nF + kChunk - 1) / kChunk);
  const double nn = (N + 1) * (N + 1) * 0.5;
  k_dt_noalloc<<<blocks, kThreads, 0, st>>>(d_h, d_hu, d_hv, d_Fscale, d_fmask, Np, (long long)K, nF, g, CFL, nn, scratch, d_dt_out);
  e = cudaGetLastError();
  if (e != cudaSuccess) return fail(err, err_cap, BDG_E_CUDA, "bdg_sw2d_timestep_noalloc: launch: %s", cudaGetErrorString(e));
  return BDG_OK;
}
