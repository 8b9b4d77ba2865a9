// Fused compute + collective kernels over NVLink peer memory (one process per GPU, buffers exchanged through
// torch.distributed._symmetric_memory; no NCCL call on these paths).
//
// Reference: the reduction / broadcast points R1-R4 of SURVEY §2.3 - helper.py:193-257 (FedAvg accumulate + apply),
// helper.py:295-418 (Weiszfeld weighted averages), helper.py:534-572 (FoolsGold gather + weighted reduce), and the
// global->client copy image_train.py:26-32.  The reference does them with Python loops on one device.
//
// Protocol.  Every rank launches the same kernel on its own GPU.  `peers[q]` are the device addresses of rank q's
// symmetric staging buffer, global-model replica and signal pad, mapped into this process.  A kernel has up to three
// phases separated by a cross-GPU barrier:
//   phase 1  local K-way reduce of the client rows into this rank's staging vector (plain HBM traffic);
//   barrier  (grid-wide arrival counter -> last block publishes this rank's ticket in every peer's signal pad with
//            st.release.sys; every block then spins with ld.acquire.sys until all peers' tickets arrived);
//   phase 2  rank r owns slice r of the vector: it loads that slice from every peer's staging buffer (P2P loads),
//            applies the update (FedAvg scale + DP noise / plain sum) and stores the result into EVERY peer's replica
//            (P2P stores) - reduce-scatter + update + all-gather in one launch, tile by tile;
//   barrier  so that nobody reads a replica before all slices landed.
// With NVLS (multicast mappings of the same buffers) phase 2 is one multimem.ld_reduce (sum formed inside the NVSwitch)
// and one multimem.st (broadcast by the switch) per 16 bytes instead of `world` peer loads and `world` peer stores.
// Tickets grow monotonically per call, so signal pads never need to be cleared.  All blocks of a launch must be
// co-resident (grid <= 148): they wait on one another.
#include "common.cuh"
#include <curand_kernel.h>

using namespace dba;

namespace {

constexpr int kMaxRanks = 8;

struct PeerSet {
  float* stage[kMaxRanks];        // [D] partial sums of each rank
  float* replica[kMaxRanks];      // [D] global model (or any broadcast target) of each rank
  unsigned int* signal[kMaxRanks];  // [2 * kMaxRanks] tickets: slot phase*kMaxRanks + sender
  int rank, world;
  // NVLS: multicast addresses of the staging / replica buffers (null = peer loads / stores).  A multimem.ld_reduce on
  // stage_mc returns the sum over ALL ranks' staging vectors, reduced inside the NVSwitch; a multimem.st on replica_mc
  // lands in every rank's replica.
  float* stage_mc;
  float* replica_mc;
};

__device__ __forceinline__ float4 mc_ld_reduce4(const float* p) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void mc_st4(float* p, const float4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};"
               ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

__device__ __forceinline__ void st_release_sys(unsigned int* p, unsigned int v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned int ld_acquire_sys(const unsigned int* p) {
  unsigned int v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

// Cross-GPU barrier usable inside a kernel whose blocks are all resident.  `counter` is a per-GPU arrival counter
// (monotonic), `ticket` identifies (call, phase).
__device__ void grid_and_peer_barrier(const PeerSet& ps, unsigned int* counter, unsigned int arrivals_target,
                                      int phase, unsigned int ticket) {
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence_system();
    const unsigned int prev = atomicAdd(counter, 1u);
    if (prev + 1 == arrivals_target) {        // last block of this GPU: publish to every peer (incl. self)
      for (int q = 0; q < ps.world; ++q) st_release_sys(ps.signal[q] + phase * kMaxRanks + ps.rank, ticket);
    }
    for (int q = 0; q < ps.world; ++q) {
      const unsigned int* flag = ps.signal[ps.rank] + phase * kMaxRanks + q;
      while ((int)(ld_acquire_sys(flag) - ticket) < 0) { __nanosleep(64); }
    }
    __threadfence_system();
  }
  __syncthreads();
}

__device__ __forceinline__ float4 ld4(const float* p) { return *reinterpret_cast<const float4*>(p); }
__device__ __forceinline__ void st4(float* p, const float4 v) { *reinterpret_cast<float4*>(p) = v; }
__device__ __forceinline__ float4 add4(float4 a, const float4 b) { a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w; return a; }

// Slice [lo, hi) of the vector owned by `rank` (4-element aligned so the slices can be moved with 16-byte accesses).
__device__ __forceinline__ void owned_slice(long long D, int rank, int world, long long& lo, long long& hi) {
  long long per = (D + world - 1) / world;
  per = (per + 3) & ~3LL;
  lo = per * rank;
  hi = lo + per;
  if (lo > D) lo = D;
  if (hi > D) hi = D;
}

// Phase 2 worker shared by both kernels: out = f(own value, sum over all peers' staging vectors) for the owned slice,
// written to every peer's replica.  Remote loads have ~2-3 us latency, so every thread keeps world x 2 independent
// 16-byte loads in flight before it touches the data.
template <typename F>
__device__ __forceinline__ void reduce_scatter_update_gather(const PeerSet& ps, long long D, F f) {
  long long lo, hi;
  owned_slice(D, ps.rank, ps.world, lo, hi);
  const float* mine = ps.replica[ps.rank];
  const long long nvec = (hi - lo) >> 2;
  const long long stride = (long long)gridDim.x * blockDim.x;
  long long v = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (ps.stage_mc != nullptr) {
    // NVLS path: one switch-reduced load and one multicast store per 16 bytes, whatever the number of ranks
    for (; v + stride < nvec; v += 2 * stride) {
      const long long i0 = lo + (v << 2), i1 = lo + ((v + stride) << 2);
      const float4 a0 = mc_ld_reduce4(ps.stage_mc + i0), a1 = mc_ld_reduce4(ps.stage_mc + i1);
      const float4 g0 = ld4(mine + i0), g1 = ld4(mine + i1);
      float4 o0, o1;
      o0.x = f(g0.x, a0.x, i0); o0.y = f(g0.y, a0.y, i0 + 1); o0.z = f(g0.z, a0.z, i0 + 2); o0.w = f(g0.w, a0.w, i0 + 3);
      o1.x = f(g1.x, a1.x, i1); o1.y = f(g1.y, a1.y, i1 + 1); o1.z = f(g1.z, a1.z, i1 + 2); o1.w = f(g1.w, a1.w, i1 + 3);
      mc_st4(ps.replica_mc + i0, o0);
      mc_st4(ps.replica_mc + i1, o1);
    }
    for (; v < nvec; v += stride) {
      const long long i0 = lo + (v << 2);
      const float4 a0 = mc_ld_reduce4(ps.stage_mc + i0);
      const float4 g0 = ld4(mine + i0);
      float4 o0;
      o0.x = f(g0.x, a0.x, i0); o0.y = f(g0.y, a0.y, i0 + 1); o0.z = f(g0.z, a0.z, i0 + 2); o0.w = f(g0.w, a0.w, i0 + 3);
      mc_st4(ps.replica_mc + i0, o0);
    }
  }
  for (; v + stride < nvec; v += 2 * stride) {
    const long long i0 = lo + (v << 2), i1 = lo + ((v + stride) << 2);
    float4 a0 = make_float4(0.f, 0.f, 0.f, 0.f), a1 = a0;
    float4 t0[kMaxRanks], t1[kMaxRanks];
#pragma unroll
    for (int q = 0; q < kMaxRanks; ++q)
      if (q < ps.world) { t0[q] = ld4(ps.stage[q] + i0); t1[q] = ld4(ps.stage[q] + i1); }
    const float4 g0 = ld4(mine + i0), g1 = ld4(mine + i1);
#pragma unroll
    for (int q = 0; q < kMaxRanks; ++q)
      if (q < ps.world) { a0 = add4(a0, t0[q]); a1 = add4(a1, t1[q]); }
    float4 o0, o1;
    o0.x = f(g0.x, a0.x, i0); o0.y = f(g0.y, a0.y, i0 + 1); o0.z = f(g0.z, a0.z, i0 + 2); o0.w = f(g0.w, a0.w, i0 + 3);
    o1.x = f(g1.x, a1.x, i1); o1.y = f(g1.y, a1.y, i1 + 1); o1.z = f(g1.z, a1.z, i1 + 2); o1.w = f(g1.w, a1.w, i1 + 3);
#pragma unroll
    for (int q = 0; q < kMaxRanks; ++q)
      if (q < ps.world) { st4(ps.replica[q] + i0, o0); st4(ps.replica[q] + i1, o1); }
  }
  for (; v < nvec; v += stride) {
    const long long i0 = lo + (v << 2);
    float4 a0 = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int q = 0; q < ps.world; ++q) a0 = add4(a0, ld4(ps.stage[q] + i0));
    const float4 g0 = ld4(mine + i0);
    float4 o0;
    o0.x = f(g0.x, a0.x, i0); o0.y = f(g0.y, a0.y, i0 + 1); o0.z = f(g0.z, a0.z, i0 + 2); o0.w = f(g0.w, a0.w, i0 + 3);
    for (int q = 0; q < ps.world; ++q) st4(ps.replica[q] + i0, o0);
  }
  // scalar tail of the slice (only the last rank's slice can have one)
  const long long t0s = lo + (nvec << 2);
  for (long long i = t0s + blockIdx.x * (long long)blockDim.x + threadIdx.x; i < hi; i += stride) {
    float acc = 0.f;
    for (int q = 0; q < ps.world; ++q) acc += ps.stage[q][i];
    const float out = f(mine[i], acc, i);
    for (int q = 0; q < ps.world; ++q) ps.replica[q][i] = out;
  }
}

// F1: FedAvg fused with delta extraction, cross-GPU reduce, global update (+DP noise) and broadcast of the new model
// into every replica.
__global__ void __launch_bounds__(256) fedavg_p2p_kernel(const float* __restrict__ W, long long ldw, int K, long long D,
                                                         float coef, float sigma, unsigned long long seed,
                                                         PeerSet ps, unsigned int* counter, unsigned int call_id,
                                                         int total_clients) {
  float* stage = ps.stage[ps.rank];
  // phase 1: local sum of client states (the -n*G term is applied once, in phase 2, with the global client count)
  const long long nvec = D >> 2;
  const bool vec_ok = (ldw & 3) == 0 && (((uintptr_t)W & 15) == 0);
  if (vec_ok) {
    for (long long v = blockIdx.x * (long long)blockDim.x + threadIdx.x; v < nvec; v += (long long)gridDim.x * blockDim.x) {
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
      for (int k = 0; k < K; ++k) acc = add4(acc, ld4(W + k * ldw + (v << 2)));
      st4(stage + (v << 2), acc);
    }
  }
  for (long long i = (vec_ok ? (nvec << 2) : 0) + blockIdx.x * (long long)blockDim.x + threadIdx.x; i < D;
       i += (long long)gridDim.x * blockDim.x) {
    float acc = 0.f;
    for (int k = 0; k < K; ++k) acc += W[k * ldw + i];
    stage[i] = acc;
  }
  grid_and_peer_barrier(ps, counter, (2 * call_id + 1) * gridDim.x, 0, 2 * call_id + 1);
  // phase 2: this rank reduces + updates its slice and scatters it to all replicas
  const float n = (float)total_clients;
  reduce_scatter_update_gather(ps, D, [=](float g, float acc, long long i) {
    float out = g + coef * (acc - n * g);
    if (sigma > 0.f) {
      curandStatePhilox4_32_10_t st;
      curand_init(seed, (unsigned long long)i, 0, &st);
      out += sigma * curand_normal(&st);
    }
    return out;
  });
  grid_and_peer_barrier(ps, counter, (2 * call_id + 2) * gridDim.x, 1, 2 * call_id + 2);
}

// Generic fused all-reduce (sum) of the ranks' staging vectors into every replica:  replica[q][i] = beta*replica[q][i]
// + alpha * sum_r stage[r][i].  Used per Weiszfeld iteration (median) and for FoolsGold's weighted gradient reduce
// fused with the server SGD step (alpha = -lr*eta/n, beta = 1 - lr*decay).
__global__ void __launch_bounds__(256) allreduce_axpby_p2p_kernel(long long D, float alpha, float beta, PeerSet ps,
                                                                  unsigned int* counter, unsigned int call_id) {
  grid_and_peer_barrier(ps, counter, (2 * call_id + 1) * gridDim.x, 0, 2 * call_id + 1);
  reduce_scatter_update_gather(ps, D, [=](float g, float acc, long long) {
    return (beta != 0.f ? beta * g : 0.f) + alpha * acc;
  });
  grid_and_peer_barrier(ps, counter, (2 * call_id + 2) * gridDim.x, 1, 2 * call_id + 2);
}

// F3 (gather part): Gram matrix over feature rows that live on DIFFERENT GPUs - rows are read through peer pointers
// inside the dot-product kernel, so the all-gather is never materialised.
__global__ void __launch_bounds__(256) gram_ptrs_kernel(const double* const* __restrict__ rows, int n, long long L,
                                                        double* __restrict__ gram) {
  __shared__ double red[32];
  const int i = blockIdx.x, j = blockIdx.y;
  if (j < i) return;
  const double* a = rows[i];
  const double* b = rows[j];
  double acc = 0.0;
  for (long long t = threadIdx.x; t < L; t += blockDim.x) acc += a[t] * b[t];
  acc = block_sum(acc, red);
  if (threadIdx.x == 0) {
    gram[(long long)i * n + j] = acc;
    gram[(long long)j * n + i] = acc;
  }
}

static int fill_peers(PeerSet& ps, const void* const* stage, const void* const* replica, const void* const* signal,
                      int rank, int world) {
  if (world < 1 || world > kMaxRanks || rank < 0 || rank >= world) return 1;
  ps.rank = rank;
  ps.world = world;
  for (int q = 0; q < world; ++q) {
    ps.stage[q] = (float*)stage[q];
    ps.replica[q] = (float*)replica[q];
    ps.signal[q] = (unsigned int*)signal[q];
  }
  ps.stage_mc = nullptr;
  ps.replica_mc = nullptr;
  return 0;
}

}  // namespace

extern "C" {

// stage / replica / signal: host arrays of `world` device pointers (peer-mapped symmetric memory).  signal pads need
// 2*8 uint32 each, zero-initialised once.  counter: one zero-initialised uint32 in local device memory per GPU.
// call_id: 0, 1, 2, ... identical on all ranks, strictly increasing per (signal pad, counter) pair.
int dba_fedavg_p2p(const float* W, long long ldw, int K, long long D, float coef, float sigma, unsigned long long seed,
                   const void* const* stage, const void* const* replica, const void* const* signal, int rank,
                   int world, int total_clients, unsigned int* counter, unsigned int call_id, void* stream) {
  PeerSet ps;
  DBA_REQUIRE(fill_peers(ps, stage, replica, signal, rank, world) == 0, "dba_fedavg_p2p: bad rank/world %d/%d", rank,
              world);
  DBA_REQUIRE(K >= 0 && D > 0, "dba_fedavg_p2p: K=%d D=%lld", K, D);
  fedavg_p2p_kernel<<<kNumSMs, 256, 0, (cudaStream_t)stream>>>(W, ldw, K, D, coef, sigma, seed, ps, counter, call_id,
                                                              total_clients);
  DBA_COUNT_LAUNCH(1);
  DBA_CHECK_LAUNCH();
  return DBA_OK;
}

// NVLS variants: stage_mc / replica_mc are the multicast addresses of the same symmetric buffers (both non-null and
// 16-byte aligned); phase 2 then uses multimem.ld_reduce / multimem.st (reduction inside the NVSwitch).
int dba_fedavg_nvls(const float* W, long long ldw, int K, long long D, float coef, float sigma, unsigned long long seed,
                    const void* const* stage, const void* const* replica, const void* const* signal, void* stage_mc,
                    void* replica_mc, int rank, int world, int total_clients, unsigned int* counter,
                    unsigned int call_id, void* stream) {
  PeerSet ps;
  DBA_REQUIRE(fill_peers(ps, stage, replica, signal, rank, world) == 0, "dba_fedavg_nvls: bad rank/world %d/%d", rank,
              world);
  DBA_REQUIRE(K >= 0 && D > 0 && stage_mc && replica_mc, "dba_fedavg_nvls: K=%d D=%lld", K, D);
  ps.stage_mc = (float*)stage_mc; ps.replica_mc = (float*)replica_mc;
  fedavg_p2p_kernel<<<kNumSMs, 256, 0, (cudaStream_t)stream>>>(W, ldw, K, D, coef, sigma, seed, ps, counter, call_id,
                                                              total_clients);
  DBA_COUNT_LAUNCH(1);
  DBA_CHECK_LAUNCH();
  return DBA_OK;
}

int dba_allreduce_axpby_nvls(long long D, float alpha, float beta, const void* const* stage,
                             const void* const* replica, const void* const* signal, void* stage_mc, void* replica_mc,
                             int rank, int world, unsigned int* counter, unsigned int call_id, void* stream) {
  PeerSet ps;
  DBA_REQUIRE(fill_peers(ps, stage, replica, signal, rank, world) == 0, "dba_allreduce_axpby_nvls: bad rank/world");
  DBA_REQUIRE(D > 0 && stage_mc && replica_mc, "dba_allreduce_axpby_nvls: D=%lld", D);
  ps.stage_mc = (float*)stage_mc; ps.replica_mc = (float*)replica_mc;
  allreduce_axpby_p2p_kernel<<<kNumSMs, 256, 0, (cudaStream_t)stream>>>(D, alpha, beta, ps, counter, call_id);
  DBA_COUNT_LAUNCH(1);
  DBA_CHECK_LAUNCH();
  return DBA_OK;
}

int dba_allreduce_axpby_p2p(long long D, float alpha, float beta, const void* const* stage,
                            const void* const* replica, const void* const* signal, int rank, int world,
                            unsigned int* counter, unsigned int call_id, void* stream) {
  PeerSet ps;
  DBA_REQUIRE(fill_peers(ps, stage, replica, signal, rank, world) == 0, "dba_allreduce_axpby_p2p: bad rank/world");
  DBA_REQUIRE(D > 0, "dba_allreduce_axpby_p2p: D=%lld", D);
  allreduce_axpby_p2p_kernel<<<kNumSMs, 256, 0, (cudaStream_t)stream>>>(D, alpha, beta, ps, counter, call_id);
  DBA_COUNT_LAUNCH(1);
  DBA_CHECK_LAUNCH();
  return DBA_OK;
}

// rows: DEVICE array of n pointers to fp64 feature rows (local or peer-mapped).
int dba_gram_ptrs(const double* const* rows, int n, long long L, double* gram, void* stream) {
  DBA_REQUIRE(n > 0 && n <= 1024 && L > 0, "dba_gram_ptrs: n=%d L=%lld", n, L);
  dim3 g(n, n);
  gram_ptrs_kernel<<<g, 256, 0, (cudaStream_t)stream>>>(rows, n, L, gram);
  DBA_COUNT_LAUNCH(1);
  DBA_CHECK_LAUNCH();
  return DBA_OK;
}

}  // extern "C"
