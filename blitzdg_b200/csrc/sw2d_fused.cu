// Fused sw2d stage kernel for straight-sided (affine) triangles: ONE launch per LSERK4 stage does
//   trace gather (vmapP) + BC + Lax-Friedrichs flux            (reference src/sw2d-simple/main.cpp:246-319, src/sw2d/main.cpp:327-421)
//   volume term + LIFT*(Fscale*flux)                            (:332-355 / :434-458)
//   res = a*res + dt*RHS ; q += b*res                           (src/advec1d/main.cpp:98,101)
// reading q/res once and writing them once (B_alg of SURVEY.md §8d).
//
// Tiles.  A CTA of 4 warps works on tiles of 32 consecutive elements.  At create time every tile gets
//   * a LOCAL trace map: for each face node the index of its '+' node inside the tile's node array, either one
//     of the tile's own 32*Np nodes or one of the tile's GHOST nodes (the distinct '+' nodes that live in other
//     tiles or in off-rank halo slots), packed with the face's BC tag in one 32-bit word, and
//   * the ghost source list (global node ids).
// A tile in shared memory is therefore a closed problem: h,hu,hv of own + ghost nodes, res, 13 geometry doubles
// per element, the local map.  Own nodes / res / geometry / map arrive through 1-D TMA bulk copies
// (cp.async.bulk + mbarrier), ghost nodes through 8-byte cp.async gathers from L2 that signal the same mbarrier;
// results leave through TMA bulk stores.  The persistent kernel double-buffers tiles so the next tile's copies
// and gathers fly while the current one is computed.
//
// Arithmetic.  Because the element is affine,
//   RHS_i = -( Dr (rx F_i + ry G_i) + Ds (sx F_i + sy G_i) ) + LIFT (Fscale dF_i)
// is one small GEMM  C[Np x cols] = A[Np x Kp] * B[Kp x cols]  with the constant operator
//   A = [ -Dr | -Ds | LIFT ]   (each block padded to a multiple of 4 columns, rows to a multiple of 8)
// run on the fp64 tensor pipe (mma.sync.m8n8k4.f64 = DMMA) with A fragments in registers (N<=4) or in shared
// memory in fragment order.  Phase 1 evaluates nodal fluxes and wave speeds for own AND ghost nodes with a flat,
// branch-free thread->node mapping; phase 2 builds the LIFT-part B fragments directly in fragment layout
//   lane = 4*c + t  ->  element E(c) of the tile, face slots j = 4*kb + t
// so the 4 lanes of a quad hold all face nodes of one element and the per-face max of the wave speed
// (sw2d-simple main.cpp:305-312) is two xor-shuffles; phase 3 builds the volume B fragments, issues the DMMAs and
// applies the stage update in place in shared memory.
//
// Deviations from the reference's operation order (inside the parity tolerance, SURVEY.md §7 hard part 1):
// FMA contraction, x/h evaluated as x*(1/h) with a Newton-refined reciprocal, chain-rule factors applied
// before the contraction (exact for affine elements up to round-off).
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <type_traits>
#include <vector>

#include "sw2d_common.cuh"

namespace {

#ifndef BDG_MINBLOCKS
#define BDG_MINBLOCKS 3
#endif
constexpr int kWarps = 4;
constexpr int kThreads = kWarps * 32;
constexpr int kTile = kWarps * 8;  // elements per CTA tile
constexpr int kAW = 13;            // doubles per element in v.aff: rx,sx,ry,sy,(nx*c,ny*c,c)x3 with c = Fscale/2 (folded at create)
constexpr int kGI = 5;             // ghost indices a thread can carry: Gmax <= kGI*kThreads

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t *bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
  uint32_t ok;
  do {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
  } while (!ok);
}
// 1-D TMA bulk copy global -> shared, completion counted in bytes on an mbarrier
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, uint32_t bytes, uint64_t *bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void bulk_s2g(void *dst, const void *src, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(smem_u32(src)), "r"(bytes) : "memory");
}
// 8-byte asynchronous gather global -> shared (LDGSTS); completion through cp_async_arrive on an mbarrier
__device__ __forceinline__ void cp_async8(void *dst, const void *src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(smem_u32(dst)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async16(void *dst, const void *src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"(smem_u32(dst)), "l"(src) : "memory");
}
__device__ __forceinline__ void prefetch_l2(const void *src, uint32_t bytes) {
  asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(src), "r"(bytes) : "memory");
}
__device__ __forceinline__ void cp_async_arrive(uint64_t *bar) {
  asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}

__device__ __forceinline__ void dmma884(double &d0, double &d1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(d0), "+d"(d1) : "d"(a), "d"(b));
}

// barrier of the kThreads compute threads (the persistent kernel has a fifth, copy-only warp that never joins it)
template <int NT>
__device__ __forceinline__ void compute_sync() { asm volatile("bar.sync 1, %0;" ::"n"(NT) : "memory"); }

// max of two NON-NEGATIVE doubles (wave speeds) through their bit patterns: integer compares instead of DSETP keep
// the shared fp64 pipe for the arithmetic; a NaN compares largest and so reaches the blow-up guard
__device__ __forceinline__ double dmax(double a, double b) { return __double_as_longlong(a) > __double_as_longlong(b) ? a : b; }

// max(x, 0), min(x, 0) and "x > 0" through the sign / bit pattern (integer pipe instead of DSETP / DMNMX on the shared fp64 pipe).
// A NaN produced by the arithmetic here is the canonical positive one and passes through pos_part / is_pos unchanged.
__device__ __forceinline__ double pos_part(double x) { return __double2hiint(x) < 0 ? 0.0 : x; }
__device__ __forceinline__ double neg_part(double x) { return __double2hiint(x) < 0 ? x : 0.0; }
__device__ __forceinline__ bool is_pos(double x) { return __double_as_longlong(x) > 0; }

// 1/x: MUFU seed (20+ bits) + two Newton steps; no denormal / inf handling (h > 0 here)
__device__ __forceinline__ double frcp(double x) {
  double r;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
  double e = fma(-x, r, 1.0);
  r = fma(r, e, r);
  e = fma(-x, r, 1.0);
  return fma(r, e, r);
}
// sqrt(x), x > 0: coupled Newton on (sqrt, 1/(2 sqrt)) from the MUFU rsqrt seed, one final correction
__device__ __forceinline__ double fsqrt(double x) {
  double r;
  asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
  double hh = 0.5 * r, s = x * r;
  double e = fma(-hh, s, 0.5);
  s = fma(s, e, s);
  e = fma(-s, s, x);  // seed error d -> ~1.5 d^2 after the coupled step -> O(d^3) after this correction (d <= 2^-20):
  return fma(e, hh, s);  // the correction term is already ~d^2 small, so the unrefined 1/(2 sqrt) (relative error d) serves
}

#ifdef BDG_PHASEPROF
// phase profiler (debug builds only): lane 0 of every warp accumulates clock64() deltas per code section
#define PROF_DECL long long pacc[12] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0}; long long plast = clock64();
#define PROF_T(i) do { long long now_ = clock64(); pacc[i] += now_ - plast; plast = now_; } while (0)
#define PROF_ARGS , long long *pacc, long long &plast
#define PROF_PASS , pacc, plast
#else
#define PROF_DECL
#define PROF_T(i) do { } while (0)
#define PROF_ARGS
#define PROF_PASS
#endif

template <int NP, int NFP>
struct Cfg {
  static constexpr int NFN = 3 * NFP;
  static constexpr int NPB = (NP + 3) / 4;    // k-blocks of the Dr part (and of the Ds part)
  static constexpr int NFB = (NFN + 3) / 4;   // k-blocks of the LIFT part
  static constexpr int KB = 2 * NPB + NFB;
  static constexpr int MB = (NP + 7) / 8;     // row blocks
#ifndef BDG_AREGS_MAX
#define BDG_AREGS_MAX 36
#endif
  static constexpr bool A_REGS = (MB * KB <= BDG_AREGS_MAX);
  // Face blocks: k-block kb < 3 of the LIFT part holds (up to 4) slots of face kb only, the remaining slots sit in
  // the k-blocks >= 3 -- the face of a position is then a compile-time constant except in those last blocks.
  static constexpr bool FACE_BLOCKS = NFB >= 3 && (NFP <= 4 || 3 * (NFP - 4) <= 4 * (NFB - 3));
  static constexpr int TNP = kTile * NP;      // own nodes per tile
  static constexpr uint32_t txQ = sizeof(double) * TNP;  // one field plane of a tile
  static constexpr uint32_t txMeta = sizeof(uint32_t) * kTile * NFN + sizeof(double) * kTile * kAW;
  static constexpr size_t sizeA = A_REGS ? 0 : sizeof(double) * MB * KB * 32;
  static constexpr uint32_t txDl = sizeof(double) * kTile * NFN;  // full mode: one static double per face slot of the tile
  // per tile buffer: Q[3][PS], Aff[kTile][13], Lm[kTile][NFN] (+ Dl[kTile][NFN] in full mode); CTA-wide: R[3][TNP],
  // Fl[4][PS] (full mode: Fl[4][TNP] -- nodal fluxes and 1/h of the OWN nodes only, the traces are re-evaluated from their
  // star states), A, barriers
  __host__ __device__ static constexpr size_t buf_bytes(int G, bool full) {
    return sizeof(double) * 3 * (TNP + G) + sizeof(double) * kTile * kAW + sizeof(uint32_t) * kTile * NFN + (full ? txDl : 0);
  }
  static constexpr size_t total(int nbuf, int G, bool full = false) {
    return nbuf * buf_bytes(G, full) + sizeof(double) * 3 * TNP + sizeof(double) * 4 * (TNP + (full ? 0 : G)) + sizeA + 64 + 128 +
           (nbuf > 1 ? nbuf * sizeof(int32_t) * G : 0);
  }
  // Ghost slots of the persistent kernel's tile buffers: a COMPILE-TIME constant (every plane offset folds into the
  // LDS/STS immediates), the largest multiple of 8 that keeps `ctas` CTAs per SM (228 KB per SM, 1 KB reserved per
  // CTA); tiles with more ghosts run in the edge kernel.  Phase-1 work follows a tile's own ghost count, not this.
  static constexpr int gfit(int ctas, bool full) {
    const long budget = 233472 / ctas - 1024, fixed = (long)total(2, 0, full), per8 = (long)(total(2, 8, full) - total(2, 0, full));
    const long g = budget > fixed ? (budget - fixed) / per8 * 8 : 0;
    return (int)(g > 600 ? 600 : g);
  }
  static constexpr int gfit_best(bool full) {
    return gfit(BDG_MINBLOCKS, full) >= 96 ? gfit(BDG_MINBLOCKS, full) : (gfit(2, full) >= 96 ? gfit(2, full) : gfit(1, full));
  }
  // CTAs per SM the persistent kernel is built for: the first of (BDG_MINBLOCKS, 2, 1) whose buffers hold >= 96 ghosts.
  // N <= 4: 3; N = 5: 2; N >= 6 (operator in shared memory, 0.9-1.4 K nodes per tile): 1.
  static constexpr int ctas_best(bool full) { return gfit(BDG_MINBLOCKS, full) >= 96 ? BDG_MINBLOCKS : (gfit(2, full) >= 96 ? 2 : 1); }
  static constexpr int GFIT = gfit_best(false);
  static constexpr int GFIT_FULL = gfit_best(true);
  // register split of the persistent kernel's two warpgroups (setmaxnreg): compute / copy
#ifndef BDG_REGS_COMPUTE
#define BDG_REGS_COMPUTE 136
#endif
  static constexpr int regs_compute(bool full) { return ctas_best(full) >= 3 ? BDG_REGS_COMPUTE : (ctas_best(full) == 2 ? 208 : 0); }
  static constexpr int regs_copy(bool full) { return ctas_best(full) >= 3 ? 24 : (ctas_best(full) == 2 ? 40 : 0); }
  static constexpr int CTAS = ctas_best(false), CTAS_FULL = ctas_best(true);
  // Compute warps per tile.  With ONE CTA per SM (N >= 6) four warps leave a single warp per scheduler, which cannot hide
  // the A-fragment loads and the DMMA latency (measured: DMMA pipe 47 % active at N = 8); there TWO warps share each group
  // of 8 elements and split the row blocks of the contraction (WH = 2): both build the same B fragments (pointwise work
  // that is small against 560 DMMAs per group), each owns half of the accumulators and of the epilogue rows.
  // Measured (profiles/r02_fused_orders.md): N = 8 62.9 -> 67.0 G DSU/s; at N = 6, 7 (4-5 row blocks) the duplicated phase 2 costs
  // more than the split saves (58 -> 51 G), so those keep four warps.
  static constexpr int WH = (CTAS == 1 && MB >= 6) ? 2 : 1;
  static constexpr int NCW = 4 * WH, NCT = 32 * NCW;  // compute warps / threads
  static constexpr int NTWS = NCT + 128;              // persistent kernel: + the copy warpgroup
  static constexpr int RBH = (MB + WH - 1) / WH;      // row blocks per warp
  static constexpr int REGS_COMPUTE = regs_compute(false), REGS_COMPUTE_FULL = regs_compute(true);
  static constexpr int REGS_COPY = regs_copy(false), REGS_COPY_FULL = regs_copy(true);
  // Residual in C-FRAGMENT order in HBM (simple physics, LSERK4): the residual is private to these kernels, so a tile's block
  // is stored as [field][row block][4 warps x 32 lanes][2 columns] -- every compute thread reads and writes exactly the
  // accumulator positions it owns with 16-byte accesses (a warp instruction covers 512 contiguous bytes), and the residual
  // never passes through shared memory: no R planes, no bulk load / store of them, no barrier between the copy warp and the
  // epilogue.  Rows NP..8*MB-1 are padding (+1/15 of the residual's traffic at N = 4), so only orders with little padding use it.
  // Measured at N = 4 (profiles/r02_rfrag_ab.md): the epilogue's waits on the R planes were the largest stall site of the
  // kernel (21 % of the samples); 119.7 -> 123.5 G DSU/s on the same box.  -DBDG_NO_RES_FRAG restores the staged residual.
#ifndef BDG_NO_RES_FRAG
  static constexpr bool RFRAG = (NP == 15);
#else
  static constexpr bool RFRAG = false;
#endif
  static constexpr int RBLK = 3 * MB * 256;  // doubles per tile of the fragment-ordered residual
};

struct StageArgs {
  const double *q_in;
  double *q_out;
  double *res;
  double *rhs_out;        // write_rhs mode: [3][nT]
  const double *Afrag;    // [MB][KB][32] operator in DMMA A-fragment order
  const uint32_t *lmap;   // [ntiles_all*kTile][NFN]: local '+' node (low 16 bits) | face tag << 16
  const int32_t *gflat;   // ghost node sources (global node ids) of all tiles back to back
  const long long *goff;  // [ntiles_all + 1]: tile T owns gflat[goff[T] .. goff[T+1])
  const int4 *rec;        // persistent kernel: regular tiles {tile id, ghost count, goff low / high 32 bits}
  const int32_t *biglist; // edge kernel in list mode: tile ids with more than Gcap ghosts
  int G;                  // ghost slots per tile in shared memory (Gcap: persistent kernel, Gmax: edge kernel), multiple of 8
  int list_first, use_list;  // first list position of this launch (rec / biglist); edge kernel: 0 = tiles tile0 + blockIdx.x
  int k0, k1;             // active element range
  int tile0, ntiles;      // tiles [tile0, tile0 + ntiles) of this launch
  int write_rhs;
  int slot;               // scal slot of the global LF speed
  const int32_t *slotinfo;  // [4*NFB]: face-slot arrangement: local '-' node | face << 16 | slot << 20 per B-fragment position
  unsigned int *sched;    // persistent kernel: {next list position to claim, CTAs finished} of this stream's launches
  long long *prof;        // BDG_PHASEPROF builds: [kWarps][12] cycle sums
  double a, b;            // LSERK4 stage coefficients
  // full physics / stepper forms
  const double *dl;       // full mode: [ntiles_all*kTile][NFN] static double per face slot: H+ - H- (Out faces: H-), see k_build_delta
  const double *r_src;    // what the R planes are loaded from: res (LSERK4, a != 0), q^n (second SSP-RK2 stage), NULL (nothing)
  int mode;               // stepper_value(): 1 LSERK4 stage, 2 / 3 first / second SSP-RK2 stage (sw2d_common.cuh)
  int store_res;          // R planes are written back to res (LSERK4)
  int slot_next;          // full mode: scal slot that collects the global LF speed of the NEW state (< 0: none)
  int slot_zero;          // ... and the slot to re-arm for the stage after the next (< 0: none)
  int last_stage;         // the next RHS evaluation happens at t + dt (tidal forcing of the Out faces in the speed above)
};

template <int NP, int NFP, bool FULL = false>
struct Smem {
  using C = Cfg<NP, NFP>;
  unsigned char *base;
  int nbuf, G;
  __device__ int PS() const { return C::TNP + G; }
  __device__ size_t bufb() const { return C::buf_bytes(G, FULL); }
  __device__ double *Q(int b) const { return reinterpret_cast<double *>(base + b * bufb()); }
  __device__ double *Aff(int b) const { return Q(b) + 3 * PS(); }
  __device__ uint32_t *Lm(int b) const { return reinterpret_cast<uint32_t *>(Aff(b) + kTile * kAW); }
  __device__ double *Dl(int b) const { return reinterpret_cast<double *>(Lm(b) + kTile * C::NFN); }  // FULL only (8-byte aligned: kTile*NFN is even)
  __device__ double *R() const { return reinterpret_cast<double *>(base + nbuf * bufb()); }
  __device__ double *Fl() const { return R() + 3 * C::TNP; }
  __device__ double *A() const { return Fl() + 4 * (FULL ? C::TNP : PS()); }
  __device__ uint64_t *Bar() const { return reinterpret_cast<uint64_t *>(reinterpret_cast<unsigned char *>(A()) + C::sizeA); }
  __device__ int4 *Ring() const { return reinterpret_cast<int4 *>(Bar() + 8); }  // 8 tile records (persistent kernel)
  __device__ int32_t *Ids(int b) const { return reinterpret_cast<int32_t *>(Ring() + 8) + b * G; }  // ghost sources of the tile after buffer b's
};

// Phase 1 for U nodes of one thread at once (j, j + kThreads, ...): the U dependency chains (reciprocal, two
// square roots) are independent straight-line code, so the scheduler interleaves them.  Indices behind the
// tile's last node are clamped: such lanes recompute (and re-store) the last node's values.
template <int U, bool FULL, int NT>
__device__ __forceinline__ void phase1_nodes(const double *__restrict__ sQ, double *__restrict__ sFl, int PS, int FPS, int j0, int last, double g,
                                             double half_g) {
  int j[U];
  double h[U], hu[U], hv[U];
#pragma unroll
  for (int u = 0; u < U; ++u) {
    j[u] = min(j0 + u * NT, last);
    h[u] = sQ[j[u]];
    hu[u] = sQ[PS + j[u]];
    hv[u] = sQ[2 * PS + j[u]];
  }
#pragma unroll
  for (int u = 0; u < U; ++u) {
    const double r = frcp(h[u]), uu = hu[u] * r, vv = hv[u] * r, gh2 = (half_g * h[u]) * h[u];
    sFl[j[u]] = fma(hu[u], uu, gh2);             // F2
    sFl[FPS + j[u]] = hu[u] * vv;                // G2 = F3
    sFl[2 * FPS + j[u]] = fma(hv[u], vv, gh2);   // G3
    if (FULL)
      sFl[3 * FPS + j[u]] = r;                   // 1/h for the source terms of the epilogue
    else
      sFl[3 * FPS + j[u]] = fsqrt(fma(uu, uu, fma(vv, vv, 1e-300))) + fsqrt(g * h[u]);
  }
}

// Unit normal of ONE wall face node as the reference stores it (nx, ny [K][Nfn]), straight from global memory.  Everywhere else
// the kernel uses the 13 geometry doubles per element (element / face means of the reference's per-node copies); on a
// wall the reflection hu+ = hu- - 2 nx (hu- nx + hv- ny) turns a normal that is off by the reference's own setup noise
// (~1e-13) into a spurious jump of the tangential momentum that Fscale*LIFT ~ 1e4 / h_cell amplifies: with the mean normal
// the result moves by 5e-10 (config 2) to 4e-8 (config 3) of max|RHS| relative to the reference's, with the face node's own
// normal by < 1e-11 (measured with the extended-precision arbiter).  Walls are rare: a warp that holds one
// takes this path.
__device__ __forceinline__ void wall_normal(const Sw2dView &v, long long k, int slot, int NFN, double &ux, double &uy) {
  const long long gi = min(k, (long long)v.K - 1) * NFN + slot;
  ux = __ldg(v.nx + gi);
  uy = __ldg(v.ny + gi);
}

// tidal free surface of the Out faces at time t (main.cpp:352): amp cos(om t) (tanh(ramp (t - T)) + 1) / 2
__device__ __forceinline__ double tide_level(const Sw2dView &v, double t) {
  const double om = 2.0 * 3.1415926535897932 / v.tide_period;
  return v.tide_amp * cos(om * t) * 0.5 * (tanh(v.tide_ramp * (t - v.tide_period)) + 1);
}

// The arithmetic of one tile that sits in shared memory (A1-A13 of SURVEY.md Appendix A).  sQ: [3][PS] own +
// ghost nodes, sR: [3][TNP].  On return the tile's q and res planes hold the updated values (stage mode) -- or,
// in write_rhs mode, rhs_out has been written and the tile is untouched.  Contains two compute_sync() and an
// optional wait on the res barrier before phase 3; the caller makes the tile visible first.
// The operator fragments (Areg / sA) arrive pre-multiplied by dt (by 1 in write_rhs mode), so the accumulators
// hold dt*RHS and the new res is a*res + acc.  All three phases are branch-free on their hot path: padding slots are
// clamped to a valid neighbour (their operator columns are zero), boundary faces are a rare warp-level
// variant, and every thread keeps several independent dependency chains in flight.
//
// FULL = the physics of src/sw2d/main.cpp:279-484: star states from ONE static double per face slot (sDl: H+ - H-; on Out
// faces, where that is 0, H- for the tidal hP), every trace flux re-evaluated from its star state, global LF speed from
// scal[p.slot], bed-slope / drag / Coriolis sources in the epilogue (Hx, Hy straight from global memory: a thread's
// C fragment covers 4 nodes), and the global LF speed of the NEW state collected for the next stage (lam_next).
template <int NP, int NFP, bool FULL, bool RFK = false>
__device__ __forceinline__ void compute_tile(const Sw2dView &v, const StageArgs &p, double *__restrict__ sQ, double *__restrict__ sR,
                                             const double *__restrict__ sAff, const uint32_t *__restrict__ sLm, const double *__restrict__ sDl,
                                             double *__restrict__ sFl, const double *__restrict__ sA, const double *Areg, const int *fm, int PS,
                                             int nnodes, long long kt0, bool need_res, uint64_t *bar_res, uint32_t res_parity, double dtv,
                                             double lamg, double tide, double tide_next, double &lam_next, double *__restrict__ gres,
                                             bool guard PROF_ARGS) {
  using C = Cfg<NP, NFP>;
  constexpr bool RF = RFK;  // gres: this tile's residual block in C-fragment order; guard: store active elements only
  static_assert(!(RFK && FULL) && (!RFK || C::RFRAG), "the fragment-ordered residual serves the LSERK4 stages of the simple physics");
#ifdef BDG_RES_FRAG_PAD_IO
  constexpr bool kPadIO = true;   // study: move the padding rows of the fragment-ordered residual as well
#else
  constexpr bool kPadIO = false;
#endif
  constexpr int NFN = C::NFN, NPB = C::NPB, NFB = C::NFB, KB = C::KB, MB = C::MB, TNP = C::TNP, NT = C::NCT, RBH = C::RBH;
  static_assert(!(C::A_REGS && C::WH > 1) && !(FULL && C::WH > 1), "the split-row-block layout is built for the big orders");
  const int FPS = FULL ? TNP : PS;  // stride of the flux planes
  const int lane = threadIdx.x & 31, wfull = threadIdx.x >> 5;
  const int w = wfull & 3;             // element group of my warp: elements w + 4*c'
  const int rb0 = C::WH > 1 ? (wfull >> 2) * RBH : 0;  // first row block of my warp (WH = 2: the upper warps take the second half)
  const int c = lane >> 2, t = lane & 3;
  const int e = w + 4 * (c < 4 ? c : (c ^ 1));  // my element (B-fragment column c); E(c) keeps B and C accesses conflict-free
  const int eo = e * NP;
  const double g = v.g, half_g = 0.5 * v.g;
  const double *af = sAff + e * kAW;

  // ---- phase 1: nodal fluxes (+ wave speed / 1/h) at own (and, simple mode, ghost) nodes (A6, A7), flat mapping ----
  {
    const int nn = FULL ? TNP : nnodes;
    const int wbase = wfull * 32, last = nn - 1;
    int i = 0;
    for (; wbase + (i + 2) * NT < nn; i += 3) phase1_nodes<3, FULL, NT>(sQ, sFl, PS, FPS, threadIdx.x + i * NT, last, g, half_g);
    if (wbase + (i + 1) * NT < nn)
      phase1_nodes<2, FULL, NT>(sQ, sFl, PS, FPS, threadIdx.x + i * NT, last, g, half_g);
    else if (wbase + i * NT < nn)
      phase1_nodes<1, FULL, NT>(sQ, sFl, PS, FPS, threadIdx.x + i * NT, last, g, half_g);
  }
  PROF_T(2);
  compute_sync<NT>();
  PROF_T(3);

  // ---- phase 2: face slots of my element: '+' trace, BC, speed, flux jump (A1-A8) ---------------------
  // fm[kb] = local '-' node | face << 16 | slot << 20 of the face slot at my position 4*kb+t (arrange_slots).  Pass A finds the per-face maximum wave speed
  // (two loads per slot), pass B evaluates the flux jump straight into the LIFT-part B fragments, so only those
  // 3*NFB doubles stay live.  Wall faces (main.cpp:340-345: hP = hM, normal momentum reflected) are rare: a warp
  // that holds one runs the second variant of each pass.
  auto wall_trace = [&](double nx, double ny, double hM, double huM, double hvM, double &huP, double &hvP, double &F2P, double &G2P,
                        double &G3P, double &spdP) {
    const double d = huM * nx + hvM * ny;
    huP = fma(-2.0 * nx, d, huM);
    hvP = fma(-2.0 * ny, d, hvM);
    const double rr = frcp(hM), u = huP * rr, vv = hvP * rr, gh2 = (half_g * hM) * hM;
    F2P = fma(huP, u, gh2); G2P = huP * vv; G3P = fma(hvP, vv, gh2);
    spdP = fsqrt(fma(u, u, fma(vv, vv, 1e-300))) + fsqrt(g * hM);
  };
  uint32_t word[NFB];
  uint32_t bcmask = 0;  // positions whose face carries a BC tag (simple mode: Wall only)
#pragma unroll
  for (int kb = 0; kb < NFB; ++kb) {
    word[kb] = sLm[e * NFN + (fm[kb] >> 20)];
    const uint32_t tag = word[kb] >> 16;
    bcmask |= ((FULL ? (tag == 3u || tag == 2u) : tag == 3u) ? 1u : 0u) << kb;
  }
  const bool any_wall = __any_sync(0xffffffffu, bcmask != 0u);
  double sfl[NFB][3];  // Fscale * dFlux_i at my slots = LIFT-part B fragments
  if constexpr (FULL) {
    // lamg: global LF speed of this state (main.cpp:413-414), from the previous stage's epilogue or k_speed; read ONCE per
    // launch by the caller (a load per tile would put an L2 round trip into every phase 2)
    const double qnan = __longlong_as_double(0x7ff8000000000000LL);
    auto pass_f = [&](auto with_bc) {
#pragma unroll
      for (int kb = 0; kb < NFB; ++kb) {
        const int f = (C::FACE_BLOCKS && kb < 3) ? kb : (fm[kb] >> 16) & 15;
        const int m = eo + (fm[kb] & 0xffff), r = word[kb] & 0xffffu;
        double nx = af[4 + 3 * f], ny = af[5 + 3 * f];
        const double hfs = af[6 + 3 * f];  // (nx, ny, 1) * Fscale/2
        double hM = sQ[m], huM = sQ[PS + m], hvM = sQ[2 * PS + m];
        double hP = sQ[r], huP = sQ[PS + r], hvP = sQ[2 * PS + r];
        double dlt = sDl[e * NFN + (fm[kb] >> 20)];
        if (decltype(with_bc)::value && ((bcmask >> kb) & 1u)) {
          if ((word[kb] >> 16) == 3u) {  // Wall (main.cpp:340-345), with the reference's OWN normal of this face node (see wall_normal)
            double ux, uy;
            wall_normal(v, kt0 + e, fm[kb] >> 20, NFN, ux, uy);
            const double d = huM * ux + hvM * uy;
            hP = hM;
            huP = fma(-2.0 * ux, d, huM);
            hvP = fma(-2.0 * uy, d, hvM);
            nx = ux * hfs; ny = uy * hfs;
          } else {  // Out (main.cpp:348-353): the slot carries H-
            hP = dlt + tide;
            huP = huM; hvP = hvM;
          }
          dlt = 0.0;
        }
        // star states (main.cpp:355-368): h-* = max(0, h- + min(H+ - H-, 0)), h+* = max(0, h+ - max(H+ - H-, 0)); the momenta
        // keep their values (hu* = h* (hu / h*) after h = h*), a dry trace poisons them like the reference's 0/0
        const double hMs = pos_part(hM + neg_part(dlt)), hPs = pos_part(hP - pos_part(dlt));
        if (!is_pos(hMs)) huM = hvM = qnan;
        if (!is_pos(hPs)) huP = hvP = qnan;
        const double rM = frcp(hMs), rP = frcp(hPs);
        const double uM = huM * rM, vM = hvM * rM, uP = huP * rP, vP = hvP * rP;
        const double pM = (half_g * hMs) * hMs, pP = (half_g * hPs) * hPs;
        const double dF2 = fma(huM, uM, pM) - fma(huP, uP, pP), dG2 = huM * vM - huP * vP, dG3 = fma(hvM, vM, pM) - fma(hvP, vP, pP);
        const double l = hfs * lamg;
        const double dh = hMs - hPs, dhu = huM - huP, dhv = hvM - hvP;
        sfl[kb][0] = fma(-l, dh, fma(dhu, nx, dhv * ny));
        sfl[kb][1] = fma(-l, dhu, fma(dF2, nx, dG2 * ny));
        sfl[kb][2] = fma(-l, dhv, fma(dG2, nx, dG3 * ny));
      }
    };
    if (any_wall) pass_f(std::true_type{}); else pass_f(std::false_type{});
  } else {
  const uint32_t wallmask = bcmask;
  double lam0 = 0.0, lam1 = 0.0, lam2 = 0.0;  // per-face max speed (scalars: a runtime face index must not spill)
  if (v.lf_global) {
    lam0 = lam1 = lam2 = lamg;
  } else {
    auto pass_a = [&](auto with_wall) {
#pragma unroll
      for (int kb = 0; kb < NFB; ++kb) {
        const bool fixed_face = C::FACE_BLOCKS && kb < 3;
        const int f = fixed_face ? kb : (fm[kb] >> 16) & 15;
        const int m = eo + (fm[kb] & 0xffff), r = word[kb] & 0xffffu;
        double spdP = sFl[3 * PS + r];
        const double spdM = sFl[3 * PS + m];
        if (decltype(with_wall)::value && ((wallmask >> kb) & 1u)) {
          double a0, a1, a2, a3, a4;
          const double ih = frcp(af[6 + 3 * f]);  // unit normal back from the folded one (walls are rare)
          wall_trace(af[4 + 3 * f] * ih, af[5 + 3 * f] * ih, sQ[m], sQ[PS + m], sQ[2 * PS + m], a0, a1, a2, a3, a4, spdP);
        }
        const double sp = dmax(spdM, spdP);
        if (fixed_face) {  // the only position of this thread on face kb so far
          if (kb == 0) lam0 = sp;
          if (kb == 1) lam1 = sp;
          if (kb == 2) lam2 = sp;
        } else {
          lam0 = f == 0 ? dmax(lam0, sp) : lam0;
          lam1 = f == 1 ? dmax(lam1, sp) : lam1;
          lam2 = f == 2 ? dmax(lam2, sp) : lam2;
        }
      }
    };
    if (any_wall) pass_a(std::true_type{}); else pass_a(std::false_type{});
    // per-face max over the quad (the 4 lanes of a quad hold all face nodes of element e)
    lam0 = dmax(lam0, __shfl_xor_sync(0xffffffffu, lam0, 1));
    lam1 = dmax(lam1, __shfl_xor_sync(0xffffffffu, lam1, 1));
    lam2 = dmax(lam2, __shfl_xor_sync(0xffffffffu, lam2, 1));
    lam0 = dmax(lam0, __shfl_xor_sync(0xffffffffu, lam0, 2));
    lam1 = dmax(lam1, __shfl_xor_sync(0xffffffffu, lam1, 2));
    lam2 = dmax(lam2, __shfl_xor_sync(0xffffffffu, lam2, 2));
  }
  auto pass_b = [&](auto with_wall) {
#pragma unroll
  for (int kb = 0; kb < NFB; ++kb) {
    const int f = (C::FACE_BLOCKS && kb < 3) ? kb : (fm[kb] >> 16) & 15;
    const int m = eo + (fm[kb] & 0xffff), r = word[kb] & 0xffffu;
    double nx = af[4 + 3 * f], ny = af[5 + 3 * f];
    const double hfs = af[6 + 3 * f];  // (nx, ny, 1) * Fscale/2
    const double l = hfs * (f == 0 ? lam0 : (f == 1 ? lam1 : lam2));
    const double hM = sQ[m], huM = sQ[PS + m], hvM = sQ[2 * PS + m];
    const double F2M = sFl[m], G2M = sFl[PS + m], G3M = sFl[2 * PS + m];
    double hP = sQ[r], huP = sQ[PS + r], hvP = sQ[2 * PS + r];
    double F2P = sFl[r], G2P = sFl[PS + r], G3P = sFl[2 * PS + r];
    if (decltype(with_wall)::value && ((wallmask >> kb) & 1u)) {
      double spd_unused, ux, uy;
      hP = hM;
      wall_normal(v, kt0 + e, fm[kb] >> 20, NFN, ux, uy);
      wall_trace(ux, uy, hM, huM, hvM, huP, hvP, F2P, G2P, G3P, spd_unused);
      nx = ux * hfs; ny = uy * hfs;
    }
    const double dh = hM - hP, dhu = huM - huP, dhv = hvM - hvP, dG2 = G2M - G2P;
    sfl[kb][0] = fma(-l, dh, fma(dhu, nx, dhv * ny));
    sfl[kb][1] = fma(-l, dhu, fma(F2M - F2P, nx, dG2 * ny));
    sfl[kb][2] = fma(-l, dhv, fma(dG2, nx, (G3M - G3P) * ny));
  }
  };
  if (any_wall) pass_b(std::true_type{}); else pass_b(std::false_type{});
  }
  PROF_T(4);
  compute_sync<NT>();  // all '+' lookups into sQ/sFl are done: the tile may now be updated in place
  PROF_T(5);
  // ---- phase 3: B fragments -> DMMA (3 fields x RBH row blocks of independent accumulator chains) -> stage update
  // My row blocks are rb0 .. rb0+RBH-1 (WH = 1: all of them, rb0 = 0 at compile time).  A second-half warp whose last
  // row block lies behind MB (odd MB) computes it on the clamped operator block and drops the result.
  const double rx = af[0], sx = af[1], ry = af[2], sy = af[3];
  const int gr = lane >> 2;  // C fragment row within a row block
  const int c0 = 2 * t, c1 = 2 * t + 1;
  const int e0 = w + 4 * (c0 < 4 ? c0 : (c0 ^ 1));  // C fragment column 2t   -> element E(2t)
  const int e1 = w + 4 * (c1 < 4 ? c1 : (c1 ^ 1));  // C fragment column 2t+1 -> element E(2t+1)
  // full mode: static Hx, Hy (and the sponge coefficient) of my 4 C-fragment nodes, fetched now, used after the contraction
  double sHx[RBH][2], sHy[RBH][2], sSg[RBH][2];
  if constexpr (FULL) {
#pragma unroll
    for (int r = 0; r < RBH; ++r)
#pragma unroll
      for (int cc = 0; cc < 2; ++cc) {
        const int n = min(8 * (rb0 + r) + gr, NP - 1);
        const long long gm = min(kt0 + (cc ? e1 : e0), (long long)v.K - 1) * NP + n;
        sHx[r][cc] = __ldg(v.Hx + gm);
        sHy[r][cc] = __ldg(v.Hy + gm);
      }
  }
  const bool sponge = p.mode != 1 && v.sponge != nullptr;
  if (sponge) {
#pragma unroll
    for (int r = 0; r < RBH; ++r)
#pragma unroll
      for (int cc = 0; cc < 2; ++cc) {
        const int n = min(8 * (rb0 + r) + gr, NP - 1);
        sSg[r][cc] = __ldg(v.sponge + min(kt0 + (cc ? e1 : e0), (long long)v.K - 1) * NP + n);
      }
  }
  // Full physics: phase 2 needs most of the register file, so operator fragments that live across the tile loop get
  // rematerialised by the compiler one at a time right at their DMMA (a global load + the dt scaling in front of every
  // DMMA: 14 % of the stall samples).  Here they are re-read explicitly instead -- 24 independent L1 hits issued together at
  // the start of the contraction, when the phase-2 temporaries are dead -- unscaled; dt is applied to the accumulators.
#ifdef BDG_A_RELOAD_SIMPLE
  constexpr bool A_RELOAD = C::A_REGS;
#else
  constexpr bool A_RELOAD = FULL && C::A_REGS;
#endif
  double Aloc[A_RELOAD ? MB * KB : 1];
  if constexpr (A_RELOAD) {
#pragma unroll
    for (int i = 0; i < MB * KB; ++i) asm volatile("ld.global.nc.f64 %0, [%1];" : "=d"(Aloc[i]) : "l"(p.Afrag + i * 32 + lane));
  }
  double acc[3][RBH][2];
#pragma unroll
  for (int r = 0; r < RBH; ++r)
#pragma unroll
    for (int i = 0; i < 3; ++i) acc[i][r][0] = acc[i][r][1] = 0.0;
  auto vol_part = [&] {
  #pragma unroll
    for (int kb = 0; kb < NPB; ++kb) {
      const int n = eo + min(4 * kb + t, NP - 1);  // padding rows of B meet zero columns of A
      const double hu = sQ[PS + n], hv = sQ[2 * PS + n], F2 = sFl[n], G2 = sFl[FPS + n], G3 = sFl[2 * FPS + n];
      double ba[3], bb[3];
      ba[0] = fma(rx, hu, ry * hv); bb[0] = fma(sx, hu, sy * hv);
      ba[1] = fma(rx, F2, ry * G2); bb[1] = fma(sx, F2, sy * G2);
      ba[2] = fma(rx, G2, ry * G3); bb[2] = fma(sx, G2, sy * G3);
  #pragma unroll
      for (int r = 0; r < RBH; ++r) {
        const int rb = min(rb0 + r, MB - 1);
        const double a1 = C::A_REGS ? (A_RELOAD ? Aloc[r * KB + kb] : Areg[r * KB + kb]) : sA[(rb * KB + kb) * 32 + lane];
        const double a2 = C::A_REGS ? (A_RELOAD ? Aloc[r * KB + NPB + kb] : Areg[r * KB + NPB + kb]) : sA[(rb * KB + NPB + kb) * 32 + lane];
  #pragma unroll
        for (int i = 0; i < 3; ++i) dmma884(acc[i][r][0], acc[i][r][1], a1, ba[i]);
  #pragma unroll
        for (int i = 0; i < 3; ++i) dmma884(acc[i][r][0], acc[i][r][1], a2, bb[i]);
      }
    }
  };
  auto lift_part = [&] {
  #pragma unroll
    for (int kb = 0; kb < NFB; ++kb) {
  #pragma unroll
      for (int r = 0; r < RBH; ++r) {
        const int rb = min(rb0 + r, MB - 1);
        const double a3 = C::A_REGS ? (A_RELOAD ? Aloc[r * KB + 2 * NPB + kb] : Areg[r * KB + 2 * NPB + kb]) : sA[(rb * KB + 2 * NPB + kb) * 32 + lane];
  #pragma unroll
        for (int i = 0; i < 3; ++i) dmma884(acc[i][r][0], acc[i][r][1], a3, sfl[kb][i]);
      }
    }
  };
  double rres[RF ? 3 : 1][RF ? RBH : 1][2], qold[RF ? 3 : 1][RF ? RBH : 1][2];
  const int tslot = (wfull & 3) * 32 + lane;
  if constexpr (RF) {
    // LIFT part first: it consumes the 3*NFB flux-jump registers; then my residual values are requested (16-byte loads, L2
    // hits: the copy warp prefetched the block) and arrive behind the 2*NPB*MB*3 DMMAs of the volume part
    lift_part();
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
      for (int r = 0; r < RBH; ++r) {
        double2 x = make_double2(0.0, 0.0);  // first stage (a = 0): nothing to read, a*0 + acc below
        // lanes of a padding row (n >= NP) neither read nor write their slots: those 64-byte pieces never move
        if (need_res && (kPadIO || 8 * (rb0 + r) + gr < NP)) x = __ldcg(reinterpret_cast<const double2 *>(gres) + (i * MB + rb0 + r) * 128 + tslot);
        rres[i][r][0] = x.x; rres[i][r][1] = x.y;
      }
    vol_part();
    // the old values of MY accumulator positions (nobody else writes them): requested while the last DMMAs drain, ahead of the
    // warp barrier that orders the other lanes' B-fragment reads before my in-place update
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
      for (int r = 0; r < RBH; ++r) {
        const int n = min(8 * (rb0 + r) + gr, NP - 1);
        qold[i][r][0] = sQ[i * PS + e0 * NP + n];
        qold[i][r][1] = sQ[i * PS + e1 * NP + n];
      }
  } else {
    vol_part();
    lift_part();
  }
  // res is first needed here, a whole DMMA chain after the tile's other inputs: its load (which has to wait for the
  // previous tile's store to drain from the same buffer) hides behind the contraction
  if (!RF && bar_res) mbar_wait(bar_res, res_parity);
  PROF_T(6);
  // every B-fragment read of an element precedes its in-place update below: the readers are the lanes of this warp
  // (WH = 1) or of the two warps that share the element group (WH = 2: a CTA-wide barrier of the compute warps)
  if constexpr (C::WH > 1) compute_sync<NT>(); else __syncwarp();
  if constexpr (A_RELOAD) {  // the operator was not pre-scaled: acc = RHS -> dt * RHS
#pragma unroll
    for (int r = 0; r < RBH; ++r)
#pragma unroll
      for (int i = 0; i < 3; ++i) { acc[i][r][0] *= dtv; acc[i][r][1] *= dtv; }
  }
  // sources (main.cpp:460-483) at my C-fragment nodes: g h Hx - CD u |u| + f hv  and  g h Hy - CD v |u| - f hu, times dt
  if constexpr (FULL) {
#pragma unroll
    for (int r = 0; r < RBH; ++r)
#pragma unroll
      for (int cc = 0; cc < 2; ++cc) {
        const int o = (cc ? e1 : e0) * NP + min(8 * (rb0 + r) + gr, NP - 1);
        const double h = sQ[o], hu = sQ[PS + o], hv = sQ[2 * PS + o], rr = sFl[3 * FPS + o];
        const double u = hu * rr, vv = hv * rr, nu = fsqrt(fma(u, u, fma(vv, vv, 1e-300))), gh = g * h;
        acc[1][r][cc] = fma(dtv, fma(v.fcor, hv, fma(-v.CD * u, nu, gh * sHx[r][cc])), acc[1][r][cc]);
        acc[2][r][cc] = fma(dtv, fma(-v.fcor, hu, fma(-v.CD * vv, nu, gh * sHy[r][cc])), acc[2][r][cc]);
      }
  }
  // epilogue: rows n = 8*rb + gr of elements e0, e1; acc = a*res + dt*rhs (stage mode) or rhs (write_rhs mode)
  if (p.write_rhs) {
#pragma unroll
    for (int r = 0; r < RBH; ++r) {
      const int n = 8 * (rb0 + r) + gr;
#pragma unroll
      for (int cc = 0; cc < 2; ++cc) {
        const long long k = kt0 + (cc ? e1 : e0);
        if (n < NP && k >= p.k0 && k < p.k1) {
#pragma unroll
          for (int i = 0; i < 3; ++i) p.rhs_out[(long long)i * NP * v.K + k * NP + n] = acc[i][r][cc];
        }
      }
    }
    return;
  }
  if constexpr (RF) {  // LSERK4 stage with the residual in registers: res = a*res + dt*rhs straight back to HBM, q += b*res in the tile
    const bool act0 = !guard || (kt0 + e0 >= p.k0 && kt0 + e0 < p.k1), act1 = !guard || (kt0 + e1 >= p.k0 && kt0 + e1 < p.k1);
#pragma unroll
    for (int r = 0; r < RBH; ++r) {
      const int n = 8 * (rb0 + r) + gr;
      const bool real = n < NP;  // padding rows carry zeros
#pragma unroll
      for (int i = 0; i < 3; ++i) {
        double rs[2];
#pragma unroll
        for (int cc = 0; cc < 2; ++cc) {
          const int o = (cc ? e1 : e0) * NP + n;
          rs[cc] = fma(p.a, rres[i][r][cc], acc[i][r][cc]);
          if (real) sQ[i * PS + o] = fma(p.b, rs[cc], qold[i][r][cc]);
          else rs[cc] = 0.0;
        }
        double *dst = gres + ((i * MB + rb0 + r) * 128 + tslot) * 2;
        if (!real && !kPadIO) continue;
        if (!guard) {
          __stcg(reinterpret_cast<double2 *>(dst), make_double2(rs[0], rs[1]));
        } else {
          if (act0) dst[0] = rs[0];
          if (act1) dst[1] = rs[1];
        }
      }
    }
    return;
  }
#pragma unroll
  for (int r = 0; r < RBH; ++r) {
    const int n = 8 * (rb0 + r) + gr;
    if (n < NP) {
#pragma unroll
      for (int i = 0; i < 3; ++i) {
#pragma unroll
        for (int cc = 0; cc < 2; ++cc) {
          const int o = (cc ? e1 : e0) * NP + n;
          if (p.mode == 1) {  // LSERK4 stage (src/advec1d/main.cpp:98,101)
            const double rs = need_res ? fma(p.a, sR[i * TNP + o], acc[i][r][cc]) : acc[i][r][cc];
            sR[i * TNP + o] = rs;
            sQ[i * PS + o] = fma(p.b, rs, sQ[i * PS + o]);
          } else {  // SSP-RK2 stages of the src/sw2d driver (main.cpp:217-219, 230-232) + sponge relaxation (:222-224, :235-236)
            double val = p.mode == 2 ? sQ[i * PS + o] + acc[i][r][cc] : 0.5 * ((sR[i * TNP + o] + sQ[i * PS + o]) + acc[i][r][cc]);
            if (sponge && i > 0 && sSg[r][cc] != 0.0) val = val * frcp(fma(sSg[r][cc] * val, val, 1.0));
            sQ[i * PS + o] = val;
          }
        }
      }
    }
  }
  // full mode: global LF speed of the NEW state at my face slots (main.cpp:397-414), for the next stage.  An interior
  // face's '+' side is the '-' side of the neighbour (same star state: its H+ - H- is minus mine), which that
  // neighbour -- another tile, or another rank before the all-reduce -- evaluates; BC faces take both sides here.
  if constexpr (FULL) {
    if (p.slot_next >= 0) {
      __syncwarp();  // the new values of my element were written by other lanes of this warp
      const double qnan = __longlong_as_double(0x7ff8000000000000LL);
#pragma unroll
      for (int kb = 0; kb < NFB; ++kb) {
        const int f = (C::FACE_BLOCKS && kb < 3) ? kb : (fm[kb] >> 16) & 15;
        const int m = eo + (fm[kb] & 0xffff);
        const uint32_t tag = word[kb] >> 16;
        double dlt = sDl[e * NFN + (fm[kb] >> 20)];
        const double hM = sQ[m];
        double huM = sQ[PS + m], hvM = sQ[2 * PS + m];
        const double Hm = dlt;
        if (tag == 3u || tag == 2u) dlt = 0.0;
        const double hMs = pos_part(hM + neg_part(dlt));
        if (!is_pos(hMs)) huM = hvM = qnan;
        const double rM = frcp(hMs), uM = huM * rM, vM = hvM * rM;
        double sp = fsqrt(fma(uM, uM, fma(vM, vM, 1e-300))) + fsqrt(g * hMs);
        if (tag == 3u) {  // wall: reflected momentum over the same depth
          const double ih = frcp(af[6 + 3 * f]), ux = af[4 + 3 * f] * ih, uy = af[5 + 3 * f] * ih, d = huM * ux + hvM * uy;
          const double uP = fma(-2.0 * ux, d, huM) * rM, vP = fma(-2.0 * uy, d, hvM) * rM;
          sp = dmax(sp, fsqrt(fma(uP, uP, fma(vP, vP, 1e-300))) + fsqrt(g * hMs));
        } else if (tag == 2u) {  // Out: tidal depth of the next evaluation, momentum copied
          const double hPs = pos_part(Hm + tide_next);
          const double rP = frcp(hPs), uP = (is_pos(hPs) ? huM : qnan) * rP, vP = (is_pos(hPs) ? hvM : qnan) * rP;
          sp = dmax(sp, fsqrt(fma(uP, uP, fma(vP, vP, 1e-300))) + fsqrt(g * hPs));
        }
        lam_next = dmax(lam_next, sp);
      }
    }
  }
}

// end of a launch's work in one warp: fold the collected speed into scal[slot_next] (one atomic per warp and launch)
__device__ __forceinline__ void publish_speed(const Sw2dView &v, const StageArgs &p, double lam_next) {
  if (p.slot_next < 0) return;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) lam_next = dmax(lam_next, __shfl_xor_sync(0xffffffffu, lam_next, o));
  if ((threadIdx.x & 31) == 0) atomicMax(reinterpret_cast<unsigned long long *>(v.scal + p.slot_next), (unsigned long long)__double_as_longlong(lam_next));
}

// per-thread constants: operator fragments scaled by `scale` (dt, or 1 in write_rhs mode), and for each of my
// B-fragment positions 4*kb+t the face slot that sits there: '-' node offset | face << 16 | slot << 20
template <int NP, int NFP, bool FULL = false>
__device__ __forceinline__ void load_constants(const StageArgs &p, const Sw2dView &v, double *Areg, double *sA, int *fm, double scale) {
  using C = Cfg<NP, NFP>;
  const int tid = threadIdx.x, lane = tid & 31;
#ifdef BDG_A_RELOAD_SIMPLE
  constexpr bool kReload = true;
#else
  constexpr bool kReload = FULL;
#endif
  if (C::A_REGS && kReload) {
    // full physics: the operator fragments are re-read (L1) at the start of every contraction, see compute_tile
  } else if (C::A_REGS) {
#pragma unroll
    for (int i = 0; i < C::MB * C::KB; ++i) Areg[i] = scale * __ldg(p.Afrag + i * 32 + lane);
  } else {
    for (int x = tid; x < C::MB * C::KB * 32; x += C::NCT) sA[x] = scale * __ldg(p.Afrag + x);
  }
#pragma unroll
  for (int kb = 0; kb < C::NFB; ++kb) fm[kb] = __ldg(p.slotinfo + 4 * kb + (lane & 3));
}


// ---- edge kernel: one tile per CTA, any tile (ragged mesh end, partially active range, write_rhs, tiles with
// more ghost nodes than the persistent kernel's buffers hold) --------------------------------------------------
template <int NP, int NFP, bool FULL, bool RFK = false>
__global__ void __launch_bounds__(Cfg<NP, NFP>::NCT, (FULL || Cfg<NP, NFP>::CTAS < 3) ? (Cfg<NP, NFP>::CTAS < 2 ? 1 : 2) : 3) k_fused_edge(Sw2dView v, StageArgs p) {
  using C = Cfg<NP, NFP>;
  constexpr int NFN = C::NFN, TNP = C::TNP, kThreads = C::NCT;
  constexpr bool RF = RFK;  // residual in C-fragment order, read and written by compute_tile itself
  extern __shared__ __align__(128) unsigned char smem[];
  const Smem<NP, NFP, FULL> S{smem, 1, p.G};
  const int PS = S.PS();
  double *sQ = S.Q(0), *sR = S.R(), *sAff = S.Aff(0), *sDl = S.Dl(0);
  uint32_t *sLm = S.Lm(0);

  const int tid = threadIdx.x;
  const long long tile = p.use_list ? p.biglist[p.list_first + blockIdx.x] : p.tile0 + blockIdx.x;
  const long long kt0 = tile * kTile;
  const int nload = (int)min((long long)kTile, (long long)v.K - kt0);
  const bool all_active = nload == kTile && kt0 >= p.k0 && kt0 + kTile <= p.k1;
  const bool need_res = !p.write_rhs && p.r_src != nullptr;
  const long long stride = v.stride;
  const long long base = kt0 * NP;
  const long long g0 = p.goff[tile];
  const int gc = (int)(p.goff[tile + 1] - g0);
  if (FULL && blockIdx.x == 0 && tid == 0 && p.slot_zero >= 0) v.scal[p.slot_zero] = 0.0;

  // plain loads; neutral fill (h = 1) behind the mesh end; lmap is padded to whole tiles at create
  for (int x = tid; x < 3 * TNP; x += kThreads) {
    const int f = x / TNP, r = x - f * TNP;
    const bool in = r < nload * NP;
    sQ[f * PS + r] = in ? p.q_in[f * stride + base + r] : (f == 0 ? 1.0 : 0.0);
    if (!RF) sR[x] = (in && need_res) ? p.r_src[f * stride + base + r] : 0.0;
  }
  for (int x = tid; x < gc; x += kThreads) {
    const long long m = p.gflat[g0 + x];
    for (int f = 0; f < 3; ++f) sQ[f * PS + TNP + x] = p.q_in[f * stride + m];
  }
  for (int x = tid; x < kTile * NFN; x += kThreads) sLm[x] = p.lmap[kt0 * NFN + x];
  if (FULL)
    for (int x = tid; x < kTile * NFN; x += kThreads) sDl[x] = p.dl[kt0 * NFN + x];  // padded to whole tiles like lmap
  for (int x = tid; x < kTile * kAW; x += kThreads) sAff[x] = x < nload * kAW ? v.aff[kt0 * kAW + x] : 0.0;
  double Areg[C::A_REGS ? C::MB * C::KB : 1];
  int fm[C::NFB];
  const double dtv = p.write_rhs ? 1.0 : v.scal[S_DT];
  load_constants<NP, NFP, FULL>(p, v, Areg, S.A(), fm, dtv);
  double tide = 0.0, tide_next = 0.0, lam_next = 0.0;
  const double lamg = (FULL || v.lf_global) ? v.scal[p.slot] : 0.0;
  if (FULL) {
    const double t = v.scal[S_T];
    tide = tide_level(v, t);
    tide_next = p.last_stage ? tide_level(v, t + v.scal[S_DT]) : tide;
  }
  __syncthreads();

  PROF_DECL
  compute_tile<NP, NFP, FULL, RFK>(v, p, sQ, sR, sAff, sLm, sDl, S.Fl(), S.A(), Areg, fm, PS, TNP + gc, kt0, need_res, nullptr, 0, dtv, lamg, tide,
                              tide_next, lam_next, RF ? p.res + tile * C::RBLK : nullptr, !all_active PROF_PASS);
  if (p.write_rhs) return;
  if (FULL) {
    // elements outside the active range / behind the mesh end must not contribute (their values are not stored)
    const int lane = tid & 31, w = (tid >> 5) & 3, c = lane >> 2;
    const long long k = kt0 + w + 4 * (c < 4 ? c : (c ^ 1));
    if (!(k >= p.k0 && k < p.k1)) lam_next = 0.0;
    publish_speed(v, p, lam_next);
  }

  if (all_active) {
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    if (tid == 0) {
      for (int f = 0; f < 3; ++f) {
        bulk_s2g(p.q_out + f * stride + base, sQ + f * PS, C::txQ);
        if (p.store_res && !RF) bulk_s2g(p.res + f * stride + base, sR + f * TNP, C::txQ);
      }
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
    }
  } else {
    __syncthreads();
    for (int x = tid; x < 3 * TNP; x += kThreads) {
      const int f = x / TNP, r = x - f * TNP;
      const long long k = kt0 + r / NP;
      if (k >= p.k0 && k < p.k1) {
        p.q_out[f * stride + base + r] = sQ[f * PS + r];
        if (p.store_res && !RF) p.res[f * stride + base + r] = sR[x];
      }
    }
  }
}

// ---- persistent kernel: full, fully active tiles with at most Gcap ghosts, addressed through the record list
// rec[] = {tile id, ghost count, ghost offset}.  Warp-specialised: warps 0..3 only compute (compute_tile); warp 4
// owns every copy -- the bulk loads of a tile's own nodes / geometry / trace map (which also carry record j+3
// into a ring and the ghost source list of tile j+1 into an id buffer), the 8-byte cp.async ghost gathers, the res
// loads, the bulk stores of the results and their drain -- so no compute warp ever waits on the copy engine.
// Two tile buffers; mbarriers: full[b] (tile in buffer b landed: 1 expect_tx arrival + 32 cp.async arrivals),
// resfull (res planes landed, or -- in a stage with a = 0 -- merely "R is free again"), done[b] (the compute warps
// have finished the tile in buffer b).  Copy warp, step it: wait done(it-1) -> store tile it-1 -> gather the ghosts
// of tile it+1 (ids came with tile it) -> drain -> res of tile it -> bulk load of tile it+1 into the drained buffer.
// 2 warpgroups: warps 0..3 compute, warp 4 copies, warps 5..7 only give their registers back (setmaxnreg works per
// warpgroup): launched at 80 registers/thread (3 CTAs/SM), the compute warpgroup grows to 136, the other shrinks to 24.

__device__ __forceinline__ void mbar_arrive(uint64_t *bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}

template <int NP, int NFP, bool FULL, bool RFK = false>
__global__ void __launch_bounds__(Cfg<NP, NFP>::NTWS, FULL ? Cfg<NP, NFP>::CTAS_FULL : Cfg<NP, NFP>::CTAS) k_fused_persistent(Sw2dView v, StageArgs p) {
  using C = Cfg<NP, NFP>;
  constexpr int TNP = C::TNP;
  constexpr bool RF = RFK;  // residual in C-fragment order: the compute threads move it, the copy warp only prefetches it into L2
  constexpr int GF = FULL ? C::GFIT_FULL : C::GFIT;
  extern __shared__ __align__(128) unsigned char smem[];
  const Smem<NP, NFP, FULL> S{smem, 2, GF};
  constexpr int PS = TNP + GF;
  uint64_t *bar = S.Bar();  // [0],[1]: full[b]; [2]: resfull; [3],[4]: done[b]
  int4 *ring = S.Ring();
  const int tid = threadIdx.x, w = tid >> 5, lane = tid & 31;
  const bool need_res = p.r_src != nullptr;
  const long long stride = v.stride;
  if (FULL && blockIdx.x == 0 && tid == 0 && p.slot_zero >= 0) v.scal[p.slot_zero] = 0.0;
  // Tiles are claimed dynamically (atomic counter p.sched[0], list positions 0..ntiles-1): a CTA that starts late
  // -- its SM slot was still held by the cut-element launch of a partitioned run -- or sits on a slower SM simply
  // claims fewer.  The record of local tile j+3 is claimed and fetched together with tile j; a failed claim leaves an
  // end marker (x < 0) in the ring instead.
  const int4 *recs = p.rec + p.list_first;
  const int4 kEnd = make_int4(-1, 0, 0, 0);

  if (tid == 0) {
    mbar_init(&bar[0], 1 + 32);
    mbar_init(&bar[1], 1 + 32);
    mbar_init(&bar[2], 1);
    mbar_init(&bar[3], 1);
    mbar_init(&bar[4], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (tid == 0) {
    const unsigned base = atomicAdd(p.sched, 3u);
    for (unsigned i = 0; i < 3; ++i) ring[i] = base + i < (unsigned)p.ntiles ? __ldg(recs + base + i) : kEnd;
  }
  __syncthreads();

  if (w >= C::NCW) {  // ------------------------------------------------------------------ copy warp
    constexpr int kRegsCopy = FULL ? C::REGS_COPY_FULL : C::REGS_COPY;
    if constexpr (kRegsCopy > 0) asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(kRegsCopy));
    if (w > C::NCW) return;
    auto gather = [&](int b, int gcount, const int32_t *ids) {
      double *dst = S.Q(b) + TNP;
      for (int x = lane; x < gcount; x += 32) {
        const double *src = p.q_in + ids[x];
        cp_async8(dst + x, src);
        cp_async8(dst + PS + x, src + stride);
        cp_async8(dst + 2 * PS + x, src + 2 * stride);
      }
      cp_async_arrive(&bar[b]);
    };
    bool more = true;  // lane 0: a claim may still succeed
    auto finish = [&] {  // lane 0, last action of the CTA: the last CTA of the launch re-arms the scheduler
      __threadfence();
      if (atomicAdd(p.sched + 1, 1u) == gridDim.x - 1) {
        p.sched[0] = 0u;
        p.sched[1] = 0u;
      }
    };
    // lane 0: bulk loads of local tile j into buffer b (rj = its record, rj1 = the next one); claims record j+3
    auto tile_load = [&](int b, int j, const int4 &rj, const int4 &rj1) {
      const long long kt0 = (long long)rj.x * kTile;
      unsigned pos3 = 0;
      bool has_rec = false;
      if (more) {
        pos3 = atomicAdd(p.sched, 1u);
        has_rec = more = pos3 < (unsigned)p.ntiles;
      }
      if (!has_rec) ring[(j + 3) & 7] = kEnd;  // ordered before the arrive below, which the readers acquire
      const bool has_ids = rj1.x >= 0 && rj1.y > 0;
      const uint32_t idb = has_ids ? sizeof(int32_t) * rj1.y : 0;
      mbar_expect_tx(&bar[b], 3 * C::txQ + C::txMeta + (FULL ? C::txDl : 0u) + (has_rec ? 16u : 0u) + idb);
      for (int f = 0; f < 3; ++f) bulk_g2s(S.Q(b) + f * PS, p.q_in + f * stride + kt0 * NP, C::txQ, &bar[b]);
      bulk_g2s(S.Lm(b), p.lmap + kt0 * C::NFN, sizeof(uint32_t) * kTile * C::NFN, &bar[b]);
      if (FULL) bulk_g2s(S.Dl(b), p.dl + kt0 * C::NFN, C::txDl, &bar[b]);
      bulk_g2s(S.Aff(b), v.aff + kt0 * kAW, sizeof(double) * kTile * kAW, &bar[b]);
      if (has_rec) bulk_g2s(&ring[(j + 3) & 7], recs + pos3, 16, &bar[b]);
      if (has_ids) bulk_g2s(S.Ids(b), p.gflat + (((long long)rj1.w << 32) | (unsigned)rj1.z), idb, &bar[b]);
#ifndef BDG_NO_RES_PREFETCH
      if (RF && need_res) prefetch_l2(p.res + (long long)rj.x * C::RBLK, sizeof(double) * C::RBLK);
#endif
    };
    auto res_ready = [&](int tile) {  // lane 0: R is free; fill it (a != 0) or just say so
      if (RF) return;
      if (need_res) {
        mbar_expect_tx(&bar[2], 3 * C::txQ);
        for (int f = 0; f < 3; ++f) bulk_g2s(S.R() + f * TNP, p.r_src + f * stride + (long long)tile * kTile * NP, C::txQ, &bar[2]);
      } else {
        mbar_arrive(&bar[2]);
      }
    };
    auto store_tile = [&](int b, int tile) {  // lane 0
      const long long base = (long long)tile * kTile * NP;
      for (int f = 0; f < 3; ++f) {
        bulk_s2g(p.q_out + f * stride + base, S.Q(b) + f * PS, C::txQ);
        if (p.store_res && !RF) bulk_s2g(p.res + f * stride + base, S.R() + f * TNP, C::txQ);
      }
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    };
    {
      const int4 r0 = ring[0], r1 = ring[1];
      if (r0.x < 0) {  // every tile was claimed before this CTA started
        if (lane == 0) finish();
        return;
      }
      if (lane == 0) {
        tile_load(0, 0, r0, r1);
        res_ready(r0.x);
      }
      gather(0, r0.y, p.gflat + (((long long)r0.w << 32) | (unsigned)r0.z));  // first tile: ids straight from global memory
    }
    int prev_tile = 0, it = 0;
    for (;; ++it) {
      const int b = it & 1, pb = b ^ 1;
      __syncwarp();
      const int4 rc = ring[it & 7];  // came with tile it-3 (or from the prologue)
      if (rc.x < 0) break;
      if (it > 0) {
        mbar_wait(&bar[3 + pb], ((it - 1) >> 1) & 1);
        if (lane == 0) store_tile(pb, prev_tile);
      }
      const int4 rn = ring[(it + 1) & 7];  // came with tile it-2
      int4 rn2 = kEnd;
      if (rn.x >= 0) {
        mbar_wait(&bar[b], (it >> 1) & 1);  // tile it has landed: ids of tile it+1, records up to it+3
        rn2 = ring[(it + 2) & 7];
        gather(pb, rn.y, S.Ids(b));
      }
      if (lane == 0) {
        if (it > 0) {
          asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
          res_ready(rc.x);
        }
        if (rn.x >= 0) tile_load(pb, it + 1, rn, rn2);
      }
      prev_tile = rc.x;
    }
    mbar_wait(&bar[3 + ((it - 1) & 1)], ((it - 1) >> 1) & 1);  // it = number of tiles this CTA processed (>= 1)
    if (lane == 0) {
      store_tile((it - 1) & 1, prev_tile);
      asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
      finish();
    }
    return;
  }

  // ---------------------------------------------------------------------------------------- compute warps
  constexpr int kRegsCompute = FULL ? C::REGS_COMPUTE_FULL : C::REGS_COMPUTE;
  if constexpr (kRegsCompute > 0) asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(kRegsCompute));
  double Areg[C::A_REGS ? C::MB * C::KB : 1];
  int fm[C::NFB];
  const double dtv = v.scal[S_DT];
  load_constants<NP, NFP, FULL>(p, v, Areg, S.A(), fm, dtv);
  double tide = 0.0, tide_next = 0.0, lam_next = 0.0;
  const double lamg = (FULL || v.lf_global) ? v.scal[p.slot] : 0.0;
  if (FULL) {
    const double t = v.scal[S_T];
    tide = tide_level(v, t);
    tide_next = p.last_stage ? tide_level(v, t + dtv) : tide;
  }
  compute_sync<C::NCT>();
  PROF_DECL
  for (int it = 0;; ++it) {
    const int b = it & 1;
    const int4 rc = ring[it & 7];  // came with tile it-3 (or from the prologue); x < 0: no more tiles for this CTA
    if (rc.x < 0) break;
    PROF_T(0);
    mbar_wait(&bar[b], (it >> 1) & 1);
    PROF_T(1);
    compute_tile<NP, NFP, FULL, RFK>(v, p, S.Q(b), S.R(), S.Aff(b), S.Lm(b), S.Dl(b), S.Fl(), S.A(), Areg, fm, PS, TNP + rc.y, (long long)rc.x * kTile,
                                need_res, RF ? nullptr : &bar[2], it & 1, dtv, lamg, tide, tide_next, lam_next,
                                RF ? p.res + (long long)rc.x * C::RBLK : nullptr, false PROF_PASS);
    PROF_T(7);
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    compute_sync<C::NCT>();
    if (tid == 0) mbar_arrive(&bar[3 + b]);
    PROF_T(8);
  }
  if (FULL) publish_speed(v, p, lam_next);
#ifdef BDG_PHASEPROF
  if ((tid & 31) == 0 && p.prof)
    for (int i = 0; i < 12; ++i) atomicAdd((unsigned long long *)p.prof + (tid >> 5) * 12 + i, (unsigned long long)pacc[i]);
#endif
}

struct FusedOps {
  double *Afrag;
  int32_t *slotinfo;
  unsigned int *sched;  // [2 streams][2]
  uint32_t *lmap;
  int32_t *gflat;
  long long *goff;
  int4 *rec;
  int32_t *biglist;
  std::vector<int> reg_tiles, big_tiles;  // host copies (sorted tile ids) to cut launch ranges
  // the full-physics instantiation has its own (larger) ghost capacity, hence its own split into regular and big tiles
  int4 *rec_full = nullptr;
  int32_t *biglist_full = nullptr;
  std::vector<int> reg_tiles_full, big_tiles_full;
  int Gcap, Gmax;                         // ghost slots of the persistent / edge kernel buffers (multiples of 8)
  int Np, Nfp;
  double *res_frag = nullptr;  // residual in C-fragment order, [ntiles][Cfg::RBLK] (orders with Cfg::RFRAG, simple physics)
  double *dl = nullptr;     // full mode: static double per face slot, [ntiles*kTile][Nfn] (k_build_delta)
  long long n_slots = 0;    // ntiles*kTile*Nfn
  long long *prof = nullptr;
};

int g_sm_count[16] = {};

template <int NP, int NFP, bool FULL, bool RFK = false>
cudaError_t launch_t(bdg_sw2d *s, StageArgs a, const FusedOps *ops) {
  using C = Cfg<NP, NFP>;
  static int ctas_per_sm[16] = {};
  static int attr_key[16] = {};
  const int dev = s->device & 15;
  const int Gcap = FULL ? C::GFIT_FULL : C::GFIT;
  size_t smem_edge = C::total(1, ops->Gmax, FULL), smem_pers = C::total(2, Gcap, FULL);
#ifdef BDG_PHASEPROF
  if (getenv("BDG_EXTRA_SMEM")) smem_pers += atoi(getenv("BDG_EXTRA_SMEM"));  // occupancy experiments: 1 or 2 CTAs per SM
#endif
  cudaError_t e;
  if (attr_key[dev] != ops->Gmax * 4096 + Gcap + 1) {
    if ((e = cudaFuncSetAttribute(k_fused_edge<NP, NFP, FULL, RFK>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_edge)) != cudaSuccess) return e;
    if ((e = cudaFuncSetAttribute(k_fused_persistent<NP, NFP, FULL, RFK>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_pers)) != cudaSuccess) return e;
    cudaFuncSetAttribute(k_fused_edge<NP, NFP, FULL, RFK>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    cudaFuncSetAttribute(k_fused_persistent<NP, NFP, FULL, RFK>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    if ((e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&ctas_per_sm[dev], k_fused_persistent<NP, NFP, FULL, RFK>, C::NTWS, smem_pers)) != cudaSuccess) return e;
    if ((e = cudaDeviceGetAttribute(&g_sm_count[dev], cudaDevAttrMultiProcessorCount, s->device)) != cudaSuccess) return e;
    if (ctas_per_sm[dev] < 1) return cudaErrorLaunchOutOfResources;
    attr_key[dev] = ops->Gmax * 4096 + Gcap + 1;
  }
  // tiles [t0,t1) cover [k0,k1); the fully active ones [p0,p1) (k1 <= K, so they are also full) go to the
  // persistent kernel when they are regular (<= Gcap ghosts); ragged / partially active head and tail tiles and
  // the big tiles inside [p0,p1) go to the edge kernel
  const std::vector<int> &reg_tiles = FULL ? ops->reg_tiles_full : ops->reg_tiles;
  const std::vector<int> &big_tiles = FULL ? ops->big_tiles_full : ops->big_tiles;
  const int t0 = a.k0 / kTile, t1 = (a.k1 - 1) / kTile + 1;
  int p0 = (a.k0 + kTile - 1) / kTile, p1 = a.k1 / kTile;
  if (a.write_rhs || p1 <= p0) p0 = p1 = t1;
  auto edge = [&](int first, int count, int use_list) {
    if (count <= 0) return;
    StageArgs h = a;
    h.G = ops->Gmax;
    h.use_list = use_list;
    h.tile0 = use_list ? 0 : first;
    h.list_first = use_list ? first : 0;
    h.ntiles = count;
    k_fused_edge<NP, NFP, FULL, RFK><<<count, C::NCT, smem_edge, s->stream>>>(s->v, h);
    s->launches++;
  };
  edge(t0, p0 - t0, 0);
  if (p1 > p0) {
    const int jr0 = (int)(std::lower_bound(reg_tiles.begin(), reg_tiles.end(), p0) - reg_tiles.begin());
    const int jr1 = (int)(std::lower_bound(reg_tiles.begin(), reg_tiles.end(), p1) - reg_tiles.begin());
    if (jr1 > jr0) {
      StageArgs m = a;
      m.G = Gcap;
      m.rec = FULL ? ops->rec_full : ops->rec;
      m.list_first = jr0; m.ntiles = jr1 - jr0; m.use_list = 1; m.tile0 = 0;
      int grid = g_sm_count[dev] * ctas_per_sm[dev];
      // while a halo exchange is in flight leave a few SMs one CTA short, so the NCCL send/recv kernel finds room
      if (s->comm && !s->peer_ok && s->n_nbr > 0 && a.k0 == 0 && a.k1 < s->v.K) grid -= 16;
      if (grid > m.ntiles) grid = m.ntiles;
      k_fused_persistent<NP, NFP, FULL, RFK><<<grid, C::NTWS, smem_pers, s->stream>>>(s->v, m);
      s->launches++;
    }
    const int jb0 = (int)(std::lower_bound(big_tiles.begin(), big_tiles.end(), p0) - big_tiles.begin());
    const int jb1 = (int)(std::lower_bound(big_tiles.begin(), big_tiles.end(), p1) - big_tiles.begin());
    StageArgs hb = a;
    (void)hb;
    if (jb1 > jb0) {
      StageArgs h = a;
      h.G = ops->Gmax;
      h.biglist = FULL ? ops->biglist_full : ops->biglist;
      h.use_list = 1; h.tile0 = 0; h.list_first = jb0; h.ntiles = jb1 - jb0;
      k_fused_edge<NP, NFP, FULL, RFK><<<jb1 - jb0, C::NCT, smem_edge, s->stream>>>(s->v, h);
      s->launches++;
    }
  }
  edge(p1, t1 - p1, 0);
  return cudaGetLastError();
}

template <int NP, int NFP>
void build_afrag(const double *Dr, const double *Ds, const double *Lift, const std::vector<int> &pos2slot, std::vector<double> &out) {
  using C = Cfg<NP, NFP>;
  constexpr int NPB = C::NPB, KB = C::KB, MB = C::MB, NFN = C::NFN;
  out.assign((size_t)MB * KB * 32, 0.0);
  for (int rb = 0; rb < MB; ++rb)
    for (int kb = 0; kb < KB; ++kb)
      for (int lane = 0; lane < 32; ++lane) {
        const int row = 8 * rb + lane / 4;  // A fragment of m8n8k4: row = lane/4, col = lane%4
        double val = 0.0;
        if (row < NP) {
          if (kb < NPB) {
            const int j = 4 * kb + lane % 4;
            if (j < NP) val = -Dr[row * NP + j];
          } else if (kb < 2 * NPB) {
            const int j = 4 * (kb - NPB) + lane % 4;
            if (j < NP) val = -Ds[row * NP + j];
          } else {
            const int sl = pos2slot[4 * (kb - 2 * NPB) + lane % 4];  // -1: padding position (its B entry duplicates a neighbour)
            if (sl >= 0) val = Lift[row * NFN + sl];
          }
        }
        out[((size_t)rb * KB + kb) * 32 + lane] = val;
      }
}

// Face-slot arrangement.  The LIFT part of the contraction may take the Nfn face slots in any order (a column
// permutation of LIFT); the order decides which shared-memory banks the 16 lanes of a half warp hit when they
// read the '-' traces (element stride Np, 4 elements x 4 positions per k-block) and the '+' traces (wherever the
// mesh puts the neighbours).  A short local search over position swaps minimises the wavefronts of those loads,
// the '+' side counted on a sample of the mesh's own tiles.  pos2slot[j] = slot at B-fragment position j, -1 = pad.
void arrange_slots(int Np, int Nfp, int Nfn, int NFB, bool face_blocks, const int32_t *fmask, const std::vector<uint32_t> &lmap,
                   int ntiles_full, std::vector<int> &pos2slot) {
  const int NPOS = 4 * NFB;
  pos2slot.assign(NPOS, -1);
  if (face_blocks) {  // Cfg::FACE_BLOCKS: k-block kb < 3 holds slots of face kb only, the rest go to the k-blocks >= 3
    int spill = 12;
    for (int f = 0; f < 3; ++f)
      for (int j = 0; j < Nfp; ++j) pos2slot[j < 4 ? 4 * f + j : spill++] = f * Nfp + j;
  } else {
    for (int j = 0; j < Nfn; ++j) pos2slot[j] = j;
  }
  auto allowed = [&](int pos) {  // the face-block rule at one position
    return !face_blocks || pos >= 12 || pos2slot[pos] < 0 || pos2slot[pos] / Nfp == pos / 4;
  };
  std::vector<int> samples;
  for (int i = 0; i < 32 && ntiles_full > 0; ++i) samples.push_back((int)((long long)i * ntiles_full / 32));
  samples.erase(std::unique(samples.begin(), samples.end()), samples.end());
  auto mult = [](const int *addr, int n) {  // wavefronts of one 16-lane 8-byte access: max distinct addresses per bank
    int worst = 1;
    for (int b = 0; b < 16; ++b) {
      int seen[16], ns = 0;
      for (int i = 0; i < n; ++i) {
        if ((addr[i] & 15) != b) continue;
        bool dup = false;
        for (int q = 0; q < ns; ++q) dup = dup || seen[q] == addr[i];
        if (!dup) seen[ns++] = addr[i];
      }
      worst = std::max(worst, ns);
    }
    return worst;
  };
  auto group_cost = [&](const int *grp) {  // grp[4]: slots at positions 4*kb .. 4*kb+3
    int addr[16], n = 0;
    long long cost = 0;
    for (int c = 0; c < 4; ++c)
      for (int t = 0; t < 4; ++t)
        if (grp[t] >= 0) addr[n++] = 4 * c * Np + fmask[grp[t]];
    cost += (long long)mult(addr, n) * 8 * (long long)samples.size();
    for (int T : samples)
      for (int w = 0; w < 4; ++w)
        for (int half = 0; half < 2; ++half) {
          n = 0;
          for (int cc = 0; cc < 4; ++cc) {
            const int c = 4 * half + cc, e = w + 4 * (c < 4 ? c : (c ^ 1));
            for (int t = 0; t < 4; ++t)
              if (grp[t] >= 0) addr[n++] = (int)(lmap[((size_t)T * kTile + e) * Nfn + grp[t]] & 0xffffu);
          }
          cost += mult(addr, n);
        }
    return cost;
  };
  std::vector<long long> gc(NFB);
  for (int kb = 0; kb < NFB; ++kb) gc[kb] = group_cost(&pos2slot[4 * kb]);
  uint64_t rng = 0x9E3779B97F4A7C15ull;
  auto rnd = [&](int n) { rng = rng * 6364136223846793005ull + 1442695040888963407ull; return (int)((rng >> 33) % (uint64_t)n); };
  for (int iter = 0; iter < 20000 && NFB > 1; ++iter) {
    const int a = rnd(NPOS), b = rnd(NPOS);
    const int ka = a / 4, kb = b / 4;
    if (pos2slot[a] == pos2slot[b]) continue;
    std::swap(pos2slot[a], pos2slot[b]);
    bool ok = allowed(a) && allowed(b);  // and every k-block keeps at least one real slot (pads duplicate it)
    for (int k : {ka, kb}) ok = ok && (pos2slot[4 * k] >= 0 || pos2slot[4 * k + 1] >= 0 || pos2slot[4 * k + 2] >= 0 || pos2slot[4 * k + 3] >= 0);
    const long long na = ok ? group_cost(&pos2slot[4 * ka]) : 0, nb = ok ? (ka == kb ? 0 : group_cost(&pos2slot[4 * kb])) : 0;
    if (ok && na + nb <= gc[ka] + (ka == kb ? 0 : gc[kb])) {
      gc[ka] = na;
      if (ka != kb) gc[kb] = nb;
    } else {
      std::swap(pos2slot[a], pos2slot[b]);
    }
  }
}

// simple physics: N = 1..8; full physics (a second instantiation of every kernel): N = 1..5
#define BDG_FUSED_ORDERS_FULL(X) X(3, 2) X(6, 3) X(10, 4) X(15, 5) X(21, 6)
#define BDG_FUSED_ORDERS(X) BDG_FUSED_ORDERS_FULL(X) X(28, 7) X(36, 8) X(45, 9)

}  // namespace

// Full physics (star states + sources) is fused in its reference form, i.e. with the GLOBAL Lax-Friedrichs speed of
// src/sw2d/main.cpp:413-414; sources with a per-face speed (a combination only the Python drivers use) stay on the
// two-launch path.
int fused_supported(const bdg_sw2d *s) {
  if (!s->is_affine || s->v.Nf != 3 || s->v.use_filter) return 0;
  if (s->v.full && (!s->v.lf_global || s->v.Np > 21)) return 0;
#define X(np, nfp) \
  if (s->v.Np == np && s->v.Nfp == nfp) return 1;
  BDG_FUSED_ORDERS(X)
#undef X
  return 0;
}

// Builds the operator fragments and the per-tile local trace maps / ghost lists.  vmapP and facetag are the
// host arrays of create (vmapP entries >= Np*K address halo slots and are always ghosts).  Sets fused_ops on
// success; leaves it NULL (generic path) when a tile needs more ghost slots than the kernel can carry.
cudaError_t fused_prepare(bdg_sw2d *s, const double *Dr, const double *Ds, const double *Lift, const int32_t *vmapP,
                          const uint8_t *facetag, const int32_t *fmask) {
  const int Np = s->v.Np, Nfp = s->v.Nfp, Nfn = s->v.Nfn, K = s->v.K;
  if (!fused_supported(s)) return cudaErrorNotSupported;

  const int ntiles = (K + kTile - 1) / kTile;
  const int tnp = kTile * Np;
  const long long nT = (long long)Np * K;  // first halo slot
  std::vector<uint32_t> lmap((size_t)ntiles * kTile * Nfn);
  std::vector<int32_t> gflat;           // ghosts of all tiles back to back
  std::vector<size_t> goff(ntiles + 1, 0);
  int Gmax = 0;
  std::vector<int32_t> ext;
  for (int T = 0; T < ntiles; ++T) {
    const long long kt0 = (long long)T * kTile, base = kt0 * Np;
    ext.clear();
    for (int el = 0; el < kTile && kt0 + el < K; ++el)
      for (int sl = 0; sl < Nfn; ++sl) {
        const int32_t mP = vmapP[(kt0 + el) * Nfn + sl];
        const long long r = (long long)mP - base;
        // (halo slots Np*K + g are ALWAYS external: in the ragged last tile they would otherwise alias the padding nodes)
        if (r < 0 || r >= tnp || mP >= nT) ext.push_back(mP);
      }
    std::sort(ext.begin(), ext.end());
    ext.erase(std::unique(ext.begin(), ext.end()), ext.end());
    gflat.insert(gflat.end(), ext.begin(), ext.end());
    while (gflat.size() % 4) gflat.push_back(ext.empty() ? 0 : ext.back());  // 16-byte granules: the lists travel by bulk copy
    goff[T + 1] = gflat.size();
    Gmax = std::max(Gmax, (int)(goff[T + 1] - goff[T]));
    for (int el = 0; el < kTile; ++el)
      for (int sl = 0; sl < Nfn; ++sl) {
        uint32_t word;
        if (kt0 + el >= K) {
          word = (uint32_t)(el * Np);  // padding element behind the mesh end: any valid own node
        } else {
          const int32_t mP = vmapP[(kt0 + el) * Nfn + sl];
          const long long r = (long long)mP - base;
          uint32_t loc;
          if (r >= 0 && r < tnp && mP < nT)
            loc = (uint32_t)r;
          else
            loc = (uint32_t)(tnp + (std::lower_bound(ext.begin(), ext.end(), mP) - ext.begin()));
          word = loc | ((uint32_t)facetag[(kt0 + el) * 3 + sl / Nfp] << 16);
        }
        lmap[((size_t)kt0 + el) * Nfn + sl] = word;
      }
  }
  // face-slot arrangement, operator fragments in that arrangement, per-position slot table
  const int NFB = (Nfn + 3) / 4;
  std::vector<int> pos2slot;
  bool face_blocks = false;
#define X(np, nfp) \
  if (Np == np && Nfp == nfp) face_blocks = Cfg<np, nfp>::FACE_BLOCKS;
  BDG_FUSED_ORDERS(X)
#undef X
  arrange_slots(Np, Nfp, Nfn, NFB, face_blocks, fmask, lmap, K / kTile, pos2slot);
  std::vector<double> frag;
#define X(np, nfp) \
  if (Np == np && Nfp == nfp) build_afrag<np, nfp>(Dr, Ds, Lift, pos2slot, frag);
  BDG_FUSED_ORDERS(X)
#undef X
  std::vector<int32_t> slotinfo(4 * NFB);
  for (int j = 0; j < 4 * NFB; ++j) {
    int sl = pos2slot[j];
    for (int t = 0; sl < 0 && t < 4; ++t) sl = pos2slot[j / 4 * 4 + t];  // pad: duplicate a real slot of the same k-block
    slotinfo[j] = fmask[sl] | ((sl / Nfp) << 16) | (sl << 20);
  }
  if (getenv("BDG_VERBOSE")) {
    fprintf(stderr, "[blitzdg_b200] face-slot arrangement:");
    for (int j = 0; j < 4 * NFB; ++j) fprintf(stderr, " %d", pos2slot[j]);
    fprintf(stderr, "\n");
  }

  // Buffer sizing: the persistent kernel's tile buffers hold Gcap ghosts, chosen so that (almost) every tile
  // fits -- a few outliers (e.g. tiles whose elements are not neighbours of each other after a partition's
  // interior-first renumbering) must not dictate the shared memory and phase-1 work of all tiles.
  std::vector<int> cnt(ntiles);
  for (int T = 0; T < ntiles; ++T) cnt[T] = (int)(goff[T + 1] - goff[T]);
  std::vector<int> sorted = cnt;
  std::sort(sorted.begin(), sorted.end());
  int cap = 0;  // the persistent kernel's compile-time ghost capacity
#define X(np, nfp) \
  if (Np == np && Nfp == nfp) cap = Cfg<np, nfp>::GFIT;
  BDG_FUSED_ORDERS(X)
#undef X
  const int Gcap = cap, Gmax8 = std::max(8, (Gmax + 7) / 8 * 8);
  if (getenv("BDG_VERBOSE"))
    fprintf(stderr, "[blitzdg_b200] fused tiles %d, ghosts per tile: mean %.1f median %d cap %d max %d\n", ntiles,
            (double)gflat.size() / ntiles, sorted[ntiles / 2], Gcap, Gmax);
  if (Gmax8 > kGI * kThreads || tnp + Gmax8 > 65535) return cudaSuccess;  // not representable: stay on the generic path
  std::vector<int4> rec, rec_full;
  FusedOps *ops = new FusedOps();
  int cap_full = 0;
#define X(np, nfp) \
  if (Np == np && Nfp == nfp) cap_full = Cfg<np, nfp>::GFIT_FULL;
  BDG_FUSED_ORDERS_FULL(X)
#undef X
  for (int T = 0; T < ntiles; ++T) {
    const int4 r = make_int4(T, cnt[T], (int)(goff[T] & 0xffffffffLL), (int)(goff[T] >> 32));
    if (cnt[T] <= Gcap) {
      ops->reg_tiles.push_back(T);
      rec.push_back(r);
    } else {
      ops->big_tiles.push_back(T);
    }
    if (cnt[T] <= cap_full) {
      ops->reg_tiles_full.push_back(T);
      rec_full.push_back(r);
    } else {
      ops->big_tiles_full.push_back(T);
    }
  }
  std::vector<int32_t> bigf(ops->big_tiles_full.begin(), ops->big_tiles_full.end());
  if (bigf.empty()) bigf.push_back(0);
  if (rec_full.empty()) rec_full.push_back(make_int4(0, 0, 0, 0));
  std::vector<long long> goff64(goff.begin(), goff.end());
  if (gflat.empty()) gflat.push_back(0);
  std::vector<int32_t> big(ops->big_tiles.begin(), ops->big_tiles.end());
  if (big.empty()) big.push_back(0);
  if (rec.empty()) rec.push_back(make_int4(0, 0, 0, 0));

  ops->Np = Np; ops->Nfp = Nfp; ops->Gcap = Gcap; ops->Gmax = Gmax8;
  ops->slotinfo = nullptr;
  ops->sched = nullptr;
  cudaError_t e;
  if ((e = cudaMalloc((void **)&ops->sched, 4 * sizeof(unsigned int))) != cudaSuccess || (e = cudaMemset(ops->sched, 0, 4 * sizeof(unsigned int))) != cudaSuccess) {
    delete ops;
    return e;
  }
  ops->Afrag = nullptr; ops->lmap = nullptr; ops->gflat = nullptr; ops->goff = nullptr; ops->rec = nullptr; ops->biglist = nullptr;
  auto up = [&](auto **dst, const auto &vec) -> cudaError_t {
    using T = typename std::remove_reference<decltype(vec[0])>::type;
    cudaError_t r = cudaMalloc((void **)dst, sizeof(T) * vec.size());
    if (r != cudaSuccess) return r;
    return cudaMemcpy((void *)*dst, vec.data(), sizeof(T) * vec.size(), cudaMemcpyHostToDevice);
  };
  if ((e = up(&ops->slotinfo, slotinfo)) != cudaSuccess || (e = up(&ops->Afrag, frag)) != cudaSuccess || (e = up(&ops->lmap, lmap)) != cudaSuccess ||
      (e = up(&ops->gflat, gflat)) != cudaSuccess || (e = up(&ops->goff, goff64)) != cudaSuccess ||
      (e = up(&ops->rec, rec)) != cudaSuccess || (e = up(&ops->biglist, big)) != cudaSuccess ||
      (s->v.full && ((e = up(&ops->rec_full, rec_full)) != cudaSuccess || (e = up(&ops->biglist_full, bigf)) != cudaSuccess ||
                     (e = cudaMalloc((void **)&ops->dl, sizeof(double) * (size_t)ntiles * kTile * Nfn)) != cudaSuccess ||
                     (e = cudaMemset(ops->dl, 0, sizeof(double) * (size_t)ntiles * kTile * Nfn)) != cudaSuccess))) {
    s->fused_ops = ops;
    fused_release(s);
    return e;
  }
  size_t rblk = 0;
#define X(np, nfp) \
  if (Np == np && Nfp == nfp && Cfg<np, nfp>::RFRAG) rblk = Cfg<np, nfp>::RBLK;
  BDG_FUSED_ORDERS(X)
#undef X
  if (rblk && !s->v.full) {
    if ((e = cudaMalloc((void **)&ops->res_frag, sizeof(double) * rblk * (size_t)ntiles)) != cudaSuccess ||
        (e = cudaMemset(ops->res_frag, 0, sizeof(double) * rblk * (size_t)ntiles)) != cudaSuccess) {
      s->fused_ops = ops;
      fused_release(s);
      return e;
    }
  }
#ifdef BDG_PHASEPROF
  cudaMalloc((void **)&ops->prof, sizeof(long long) * 8 * 12);
  cudaMemset(ops->prof, 0, sizeof(long long) * 8 * 12);
#endif
  ops->n_slots = (long long)ntiles * kTile * Nfn;
  s->fused_ops = ops;
  return cudaSuccess;
}

namespace {
// static double per face slot for the star states (main.cpp:355-368): H+ - H- (exact for nearly equal depths); on Out
// faces, where '+' = '-', H- itself, which the tidal hP = H- + tide needs (main.cpp:352).  Padding slots stay 0.
__global__ void k_build_delta(Sw2dView v, double *__restrict__ dl) {
  const long long nF = (long long)v.Nfn * v.K;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < nF; i += (long long)gridDim.x * blockDim.x) {
    const long long k = i / v.Nfn;
    const int sl = (int)(i - k * v.Nfn);
    const double HM = v.H[k * v.Np + v.fmask[sl]], HP = v.H[v.vmapP[i]];
    const int tag = v.facetag[k * v.Nf + sl / v.Nfp];
    dl[i] = tag == 2 ? HM : (tag == 3 ? 0.0 : HP - HM);
  }
}
}  // namespace

// (re)builds the star-state table from the bathymetry on the device (ghost depths included): after set_bathymetry and,
// on a partitioned mesh, after the neighbours' depths have arrived
cudaError_t fused_build_delta(bdg_sw2d *s) {
  FusedOps *ops = static_cast<FusedOps *>(s->fused_ops);
  if (!ops || !ops->dl || !s->v.H) return cudaSuccess;
  k_build_delta<<<1184, 256, 0, s->stream>>>(s->v, ops->dl);
  s->launches++;
  return cudaGetLastError();
}

void fused_release(bdg_sw2d *s) {
  if (!s->fused_ops) return;
  FusedOps *ops = static_cast<FusedOps *>(s->fused_ops);
#ifdef BDG_PHASEPROF
  if (ops->prof) {
    long long h[8 * 12];
    cudaMemcpy(h, ops->prof, sizeof(h), cudaMemcpyDeviceToHost);
    static const char *nm[12] = {"top+gather", "wait tile", "phase1", "sync1", "phase2", "sync2", "wait res", "phase3", "fence+sync3", "store+reissue", "", ""};
    for (int w = 0; w < kWarps; ++w) {
      long long tot = 0;
      for (int i = 0; i < 10; ++i) tot += h[w * 12 + i];
      fprintf(stderr, "[PHASEPROF warp %d]", w);
      for (int i = 0; i < 10; ++i) fprintf(stderr, " %s %.1f%%", nm[i], 100.0 * h[w * 12 + i] / (tot ? tot : 1));
      fprintf(stderr, " | total cycles %lld\n", tot);
    }
    cudaFree(ops->prof);
  }
#endif
  cudaFree(ops->sched); cudaFree(ops->slotinfo); cudaFree(ops->Afrag); cudaFree(ops->lmap); cudaFree(ops->gflat); cudaFree(ops->goff); cudaFree(ops->rec); cudaFree(ops->biglist); cudaFree(ops->rec_full); cudaFree(ops->biglist_full); cudaFree(ops->dl); cudaFree(ops->res_frag);
  delete ops;
  s->fused_ops = nullptr;
}

cudaError_t launch_fused_stage(bdg_sw2d *s, const double *q_in, double *q_out, double *res, int k0, int k1, double a, double b,
                               int write_rhs, double *rhs_out, int slot, const FusedStage *fs) {
  if (k1 <= k0) return cudaSuccess;
  FusedOps *ops = static_cast<FusedOps *>(s->fused_ops);
  if (!ops) return cudaErrorNotSupported;
  StageArgs args;
  args.q_in = q_in; args.q_out = q_out; args.res = res; args.rhs_out = rhs_out; args.Afrag = ops->Afrag;
  args.lmap = ops->lmap; args.gflat = ops->gflat; args.goff = ops->goff; args.rec = ops->rec; args.biglist = ops->biglist;
  args.G = ops->Gmax; args.list_first = 0; args.use_list = 0;
  args.k0 = k0; args.k1 = k1; args.tile0 = 0; args.ntiles = 0;
  args.write_rhs = write_rhs; args.slot = slot; args.a = a; args.b = b; args.prof = ops->prof; args.slotinfo = ops->slotinfo;
  args.sched = ops->sched + (s->stream == s->comm_stream ? 2 : 0);  // launches on the two streams may overlap
  args.dl = ops->dl;
  args.mode = fs ? fs->mode : 1;
  args.slot_next = fs ? fs->slot_next : -1;
  args.slot_zero = fs ? fs->slot_zero : -1;
  args.last_stage = fs ? fs->last_stage : 0;
  // R planes: the residual (LSERK4, unless a = 0 wipes it), q^n still sitting in the destination buffer (second SSP-RK2
  // stage), nothing (first SSP-RK2 stage, RHS evaluation)
  // LSERK4 stages of an order with Cfg::RFRAG run the kernels that keep the residual in C-fragment order (private to them:
  // ops->res_frag); the SSP-RK2 stages (R planes = q^n, no residual) and every other order run the staged instantiation
  const bool rf = ops->res_frag && !s->v.full && args.mode == 1;
  if (rf) args.res = res = ops->res_frag;
  args.r_src = write_rhs ? nullptr : (args.mode == 1 ? (a != 0.0 ? res : nullptr) : (args.mode == 3 ? q_out : nullptr));
  args.store_res = !write_rhs && args.mode == 1;
  if (s->v.full) {
#define X(np, nfp) \
  if (ops->Np == np && ops->Nfp == nfp) return launch_t<np, nfp, true>(s, args, ops);
    BDG_FUSED_ORDERS_FULL(X)
#undef X
    return cudaErrorNotSupported;
  }
#define X(np, nfp)                                                                  \
  if (ops->Np == np && ops->Nfp == nfp) {                                           \
    if constexpr (Cfg<np, nfp>::RFRAG) {                                            \
      if (rf) return launch_t<np, nfp, false, true>(s, args, ops);                  \
    }                                                                               \
    return launch_t<np, nfp, false, false>(s, args, ops);                           \
  }
  BDG_FUSED_ORDERS(X)
#undef X
  return cudaErrorNotSupported;
}
