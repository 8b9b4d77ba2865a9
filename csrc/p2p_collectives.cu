// Fused compute + collective kernels over NVLink peer memory (one process per GPU, buffers exchanged through
// torch.distributed._symmetric_memory; no NCCL call on these paths).
//
// Reference: the reduction / broadcast points R1-R4 of SURVEY §2.3 - helper.py:193-257 (FedAvg accumulate + apply),
// helper.py:295-418 (RFA / Weiszfeld), helper.py:534-572 (FoolsGold gather + weighted reduce), and the global->client
// copy image_train.py:26-32.  The reference does them with Python loops on one device.
//
// Protocol.  Every rank launches the same kernel on its own GPU (cooperative launch: all blocks are guaranteed
// co-resident, they wait on one another).  `stage[q]`, `replica[q]`, `aux[q]`, `signal[q]`, `mail[q]` are the device
// addresses of rank q's symmetric buffers mapped into this process.  A kernel is a sequence of phases separated by a
// cross-GPU barrier:
//   local phase   e.g. the K-way (weighted) reduce of this rank's client rows into its staging vector (HBM traffic);
//   barrier       per-phase arrival counter on this GPU; the last block publishes (optionally after writing a small
//                 scalar payload into every peer's mailbox) this rank's ticket in every peer's signal pad with
//                 st.release.sys; then the lanes of one warp per block poll the peers' tickets in parallel with
//                 ld.acquire.sys;
//   owner phase   rank r owns slice r of the vector: it loads that slice from every peer's staging buffer (P2P loads),
//                 applies the update (FedAvg scale + DP noise / axpby / Weiszfeld normalisation) and stores the result
//                 into EVERY peer's target buffer (P2P stores) - reduce-scatter + update + all-gather, tile by tile;
//   barrier       so that nobody reads a target before all slices landed.
// With NVLS (multicast mappings of the same buffers) the owner phase is one multimem.ld_reduce (sum formed inside the
// NVSwitch) and one multimem.st (broadcast by the switch) per 16 bytes instead of `world` peer loads / stores.
// Tickets grow monotonically per call, so signal pads never need clearing.  A wait that exceeds ~4 s sets the error
// flag (a peer died or never launched) and falls through instead of hanging the GPU; the host checks the flag.
#include "common.cuh"
#include <curand_kernel.h>

using namespace dba;

namespace {

constexpr int kMaxRanks = 8;
constexpr int kMaxPhases = 8;
constexpr int kCollThreads = 512;      // one block per SM: 16 warps keep enough 16-byte loads in flight per SM
constexpr long long kSpinBudget = 8000000000LL;      // clock64 ticks (~4 s): far beyond any legitimate wait

struct PeerSet {
  float* stage[kMaxRanks];          // [D] partial sums of each rank
  float* replica[kMaxRanks];        // [D] target of the owner phase on each rank (global model or auxiliary vector)
  unsigned int* signal[kMaxRanks];  // [kMaxPhases * kMaxRanks] tickets: slot phase*kMaxRanks + sender
  double* mail[kMaxRanks];          // [kMaxPhases * kMaxRanks * 2] scalar payloads travelling with a ticket
  unsigned int* counter;            // [kMaxPhases] arrival counters of THIS GPU (reset by the last arriver)
  int* err;                         // [1] error flag of THIS GPU (1 = barrier timeout)
  int rank, world;
  // NVLS: multicast addresses of the staging / target buffers (null = peer loads / stores).  A multimem.ld_reduce on
  // stage_mc returns the sum over ALL ranks' staging vectors, reduced inside the NVSwitch; a multimem.st on replica_mc
  // lands in every rank's target.
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

// Cross-GPU barrier inside a cooperative kernel.  `ticket` identifies the call (same on all ranks, strictly increasing
// per phase slot).  pay0 / pay1: scalars this rank sends to every peer with the ticket (read them after the barrier
// with mail_read); pass has_payload = false when there is nothing to send.
__device__ void grid_and_peer_barrier(const PeerSet& ps, int phase, unsigned int ticket, bool has_payload = false,
                                      double pay0 = 0.0, double pay1 = 0.0) {
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence_system();
    const unsigned int prev = atomicAdd(ps.counter + phase, 1u);
    if (prev + 1 == gridDim.x) {              // last block of this GPU: publish to every peer (incl. self)
      ps.counter[phase] = 0;                  // everybody has arrived: the counter is free for the next call
      for (int q = 0; q < ps.world; ++q) {
        if (has_payload) {
          double* m = ps.mail[q] + ((long long)phase * kMaxRanks + ps.rank) * 2;
          m[0] = pay0;
          m[1] = pay1;
        }
        st_release_sys(ps.signal[q] + phase * kMaxRanks + ps.rank, ticket);
      }
    }
  }
  if (threadIdx.x < 32) {                     // lane q polls peer q's ticket: one round trip for all peers
    const int q = threadIdx.x;
    if (q < ps.world) {
      const unsigned int* flag = ps.signal[ps.rank] + phase * kMaxRanks + q;
      const long long t0 = clock64();
      while ((int)(ld_acquire_sys(flag) - ticket) < 0) {
        __nanosleep(32);
        if (clock64() - t0 > kSpinBudget) { *ps.err = 1; break; }
      }
    }
    __syncwarp();
    __threadfence_system();
  }
  __syncthreads();
}
__device__ __forceinline__ double mail_read(const PeerSet& ps, int phase, int sender, int which) {
  return ps.mail[ps.rank][((long long)phase * kMaxRanks + sender) * 2 + which];
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

// Owner phase shared by all kernels: out = f(own value, sum over all peers' staging vectors, index) for the owned
// slice, written to every peer's target.  Remote loads have ~2-3 us latency, so every thread keeps world x 2
// independent 16-byte loads in flight (4 switch-reduced loads on the NVLS path) before it touches the data.
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
    for (; v + 3 * stride < nvec; v += 4 * stride) {
      float4 a[4], g[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) a[u] = mc_ld_reduce4(ps.stage_mc + lo + ((v + u * stride) << 2));
#pragma unroll
      for (int u = 0; u < 4; ++u) g[u] = ld4(mine + lo + ((v + u * stride) << 2));
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const long long i = lo + ((v + u * stride) << 2);
        float4 o;
        o.x = f(g[u].x, a[u].x, i); o.y = f(g[u].y, a[u].y, i + 1); o.z = f(g[u].z, a[u].z, i + 2);
        o.w = f(g[u].w, a[u].w, i + 3);
        mc_st4(ps.replica_mc + i, o);
      }
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

// Local phase shared by the kernels:  stage[i] = sum_k c_k * (W[k][i] - (sub ? G[i] : 0)),  c_k = coef_s[k] (shared
// memory) or 1.  Subtracting the replica per row (instead of n*G after the sum) keeps the small client deltas from
// being cancelled against the large weights - the same arithmetic as the single-GPU kernel and the reference.
__device__ __forceinline__ void local_rows_reduce(const float* __restrict__ W, long long ldw, int K, long long D,
                                                  const float* __restrict__ G, const float* coef_s, float* stage) {
  const long long nvec = D >> 2;
  const bool vec_ok = (ldw & 3) == 0 && (((uintptr_t)W & 15) == 0) && (!G || ((uintptr_t)G & 15) == 0);
  if (vec_ok) {
    for (long long v = blockIdx.x * (long long)blockDim.x + threadIdx.x; v < nvec; v += (long long)gridDim.x * blockDim.x) {
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
      const float4 g = G ? ld4(G + (v << 2)) : acc;
      for (int k = 0; k < K; ++k) {
        const float4 w = ld4(W + k * ldw + (v << 2));
        const float c = coef_s ? coef_s[k] : 1.f;
        acc.x += c * (w.x - g.x); acc.y += c * (w.y - g.y); acc.z += c * (w.z - g.z); acc.w += c * (w.w - g.w);
      }
      st4(stage + (v << 2), acc);
    }
  }
  for (long long i = (vec_ok ? (nvec << 2) : 0) + blockIdx.x * (long long)blockDim.x + threadIdx.x; i < D;
       i += (long long)gridDim.x * blockDim.x) {
    float acc = 0.f;
    const float g = G ? G[i] : 0.f;
    for (int k = 0; k < K; ++k) acc += (coef_s ? coef_s[k] : 1.f) * (W[k * ldw + i] - g);
    stage[i] = acc;
  }
}

// F1: FedAvg fused with delta extraction, cross-GPU reduce, global update (+DP noise) and broadcast of the new model
// into every replica:  G += coef * sum over ALL ranks' rows of (W[k] - G).
__global__ void __launch_bounds__(kCollThreads) fedavg_p2p_kernel(const float* __restrict__ W, long long ldw, int K, long long D,
                                                         float coef, float sigma, unsigned long long seed,
                                                         PeerSet ps, unsigned int ticket) {
  local_rows_reduce(W, ldw, K, D, ps.replica[ps.rank], nullptr, ps.stage[ps.rank]);
  grid_and_peer_barrier(ps, 0, ticket);
  reduce_scatter_update_gather(ps, D, [=](float g, float acc, long long i) {
    float out = g + coef * acc;
    if (sigma > 0.f) {
      curandStatePhilox4_32_10_t st;
      curand_init(seed, (unsigned long long)i, 0, &st);
      out += sigma * curand_normal(&st);
    }
    return out;
  });
  grid_and_peer_barrier(ps, 1, ticket);
}

// Fused weighted all-reduce:  target[i] = beta*target[i] + alpha * sum over all ranks of (sum_k wv[k] * S[k][i]).
// K > 0: the local weighted K-way reduce of the rows S is phase 1 of the same launch (FoolsGold's weighted gradient
// reduce + server SGD step: alpha = -lr*eta/n, beta = 1 - lr*decay, helper.py:280-290,561-570).  K == 0: the staging
// vector was filled by the caller (plain all-reduce of a vector).
__global__ void __launch_bounds__(kCollThreads) axpby_p2p_kernel(const float* __restrict__ S, long long lds, int K,
                                                        const double* __restrict__ wv, long long D, float alpha,
                                                        float beta, PeerSet ps, unsigned int ticket) {
  extern __shared__ float coef_s[];
  if (K > 0) {
    for (int k = threadIdx.x; k < K; k += blockDim.x) coef_s[k] = wv ? (float)wv[k] : 1.f;
    __syncthreads();
    local_rows_reduce(S, lds, K, D, nullptr, coef_s, ps.stage[ps.rank]);
  }
  grid_and_peer_barrier(ps, 0, ticket);
  reduce_scatter_update_gather(ps, D, [=](float g, float acc, long long) {
    return (beta != 0.f ? beta * g : 0.f) + alpha * acc;
  });
  grid_and_peer_barrier(ps, 1, ticket);
}

// F2: ONE Weiszfeld iteration of the geometric median over the client deltas (W[k] - G) of ALL ranks
// (helper.py:295-373) in one launch per rank: local weights from the local distances -> local weighted partial sum
// -> barrier carrying (local weight sum, local objective part) -> every rank forms the global weight sum / objective
// and runs the convergence test on identical numbers -> owner phase: median slice = sum_r stage_r / weight sum,
// stored into every rank's median buffer -> barrier -> local squared distances to the new median (accumulated for the
// next launch).  Launch `it` tests the objective of iteration it-1 (deferred by one launch so that it needs no
// barrier of its own); launch maxiter+1 only tests and publishes the per-client results.
struct WzArgs {
  const float* W;              // [K][ldw] local client states
  long long ldw, D;
  const float* G;              // [D] global model (local replica)
  const float* alphas;         // [K] n_c / sum over ALL clients of n
  double* dist2;               // [2][kmax] squared distances of the local clients (parity buffers)
  float* wts;                  // [K] normalised weights of the previous launch's iteration
  float* wv;                   // [K] reported weights (Q8 semantics, see flat_ops.cu)
  double* obj;                 // [2] objective by launch parity
  int* state;                  // [4]: done, iterations, oracle calls, parity slot of the last objective
  float* const* table;         // [world] symmetric result tables [2][n_total]: wv and distance of every client
  int K, kmax, it, maxiter, slot0, n_total;
  float eps;
  double ftol;
};
__global__ void __launch_bounds__(kCollThreads) weiszfeld_p2p_kernel(WzArgs a, PeerSet ps, unsigned int ticket) {
  extern __shared__ float w_s[];             // [K] un-normalised weights of this launch
  __shared__ double red[32];
  __shared__ double s_tot[2];
  if (a.state[0]) return;                    // converged in an earlier launch: identical decision on every rank
  const bool first = a.it == 0, last = a.it == a.maxiter + 1;
  const int par = a.it & 1;
  // ---- local weights and scalars (every block computes them: K <= 1024 values)
  double wsum = 0.0, opart = 0.0;
  for (int c = threadIdx.x; c < a.K; c += blockDim.x) {
    const double d = first ? 0.0 : sqrt(a.dist2[(long long)par * a.kmax + c]);
    const float w = first ? a.alphas[c] : a.alphas[c] / fmaxf(a.eps, (float)d);
    w_s[c] = w;
    wsum += (double)w;
    opart += (double)a.alphas[c] * d;
  }
  wsum = block_sum(wsum, red, true);
  opart = block_sum(opart, red, true);
  if (blockIdx.x == 0)                       // the other parity buffer collects this launch's distances
    for (int c = threadIdx.x; c < a.K; c += blockDim.x) a.dist2[(long long)(par ^ 1) * a.kmax + c] = 0.0;
  __syncthreads();
  if (!last) local_rows_reduce(a.W, a.ldw, a.K, a.D, a.G, w_s, ps.stage[ps.rank]);
  grid_and_peer_barrier(ps, 2, ticket, true, wsum, opart);
  // ---- global scalars, summed in rank order on every rank
  if (threadIdx.x == 0) {
    double tw = 0.0, ob = 0.0;
    for (int q = 0; q < ps.world; ++q) { tw += mail_read(ps, 2, q, 0); ob += mail_read(ps, 2, q, 1); }
    s_tot[0] = tw;
    s_tot[1] = ob;
  }
  __syncthreads();
  const double tot_w = s_tot[0], obj = s_tot[1];
  bool conv = false;
  if (!first) {                              // test the iteration computed by the previous launch (t = it - 1)
    const bool first_t = a.it == 1;
    const double prev = a.obj[par ^ 1];
    conv = !first_t && fabs(prev - obj) < a.ftol * obj;
    const int iters_before = a.state[1];
    if (blockIdx.x == 0) {
      if (first_t || !conv || iters_before == 0)
        for (int c = threadIdx.x; c < a.K; c += blockDim.x) a.wv[c] = a.wts[c];
    }
    __syncthreads();
  }
  const bool finish = conv || last;
  if (!finish) {
    const float inv = (float)(1.0 / tot_w);
    if (blockIdx.x == 0)
      for (int c = threadIdx.x; c < a.K; c += blockDim.x) a.wts[c] = w_s[c] * inv;
    // ---- owner phase: normalised median slice into every rank's median buffer
    reduce_scatter_update_gather(ps, a.D, [=](float, float acc, long long) { return acc * inv; });
    grid_and_peer_barrier(ps, 3, ticket);
    // ---- local squared distances to the new median (double accumulation, one atomic per block and client)
    const float* med = ps.replica[ps.rank];
    for (int c = 0; c < a.K; ++c) {
      const float* x = a.W + c * a.ldw;
      double acc = 0.0;
      for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < a.D;
           i += (long long)gridDim.x * blockDim.x) {
        const float d = x[i] - a.G[i] - med[i];
        acc += (double)d * (double)d;
      }
      acc = block_sum(acc, red);
      if (threadIdx.x == 0) atomicAdd(&a.dist2[(long long)(par ^ 1) * a.kmax + c], acc);
    }
  } else {
    // ---- publish (wv, distance) of the local clients into every rank's result table
    if (blockIdx.x == 0) {
      for (int c = threadIdx.x; c < a.K; c += blockDim.x) {
        const float d = (float)sqrt(a.dist2[(long long)par * a.kmax + c]);
        const float w = a.wv[c];
        for (int q = 0; q < ps.world; ++q) {
          a.table[q][a.slot0 + c] = w;
          a.table[q][a.n_total + a.slot0 + c] = d;
        }
      }
    }
    grid_and_peer_barrier(ps, 4, ticket);
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    a.obj[par] = obj;
    a.state[3] = par;                        // where the last objective lives
    if (!first) {
      a.state[2] += 1;
      if (a.it > 1) a.state[1] += 1;
      if (finish) a.state[0] = 1;
    }
  }
}

// Standalone cross-GPU barrier (stream-ordered): everything the peers enqueued before their matching call is
// complete and visible when this kernel ends - replaces cudaDeviceSynchronize + an NCCL barrier.
__global__ void __launch_bounds__(64) peer_barrier_kernel(PeerSet ps, unsigned int ticket) {
  grid_and_peer_barrier(ps, 5, ticket);
}

// out[j][:] = *rows[j] for j < n: rows may live on other GPUs (peer-mapped pointers): the gather is a set of NVLink
// loads into THIS GPU's memory (evaluation exchange of client models, simulator._run_eval_jobs_balanced).
__global__ void __launch_bounds__(256) gather_rows_ptrs_kernel(const float* const* __restrict__ rows, int n, long long D,
                                                               float* __restrict__ out, long long ldo) {
  const int j = blockIdx.y;
  const float* src = rows[j];
  float* dst = out + (long long)j * ldo;
  const long long nvec = D >> 2;
  const bool vec_ok = (((uintptr_t)src | (uintptr_t)dst) & 15) == 0;
  if (vec_ok) {
    for (long long v = blockIdx.x * (long long)blockDim.x + threadIdx.x; v < nvec; v += (long long)gridDim.x * blockDim.x)
      st4(dst + (v << 2), ld4(src + (v << 2)));
  }
  for (long long i = (vec_ok ? (nvec << 2) : 0) + blockIdx.x * (long long)blockDim.x + threadIdx.x; i < D;
       i += (long long)gridDim.x * blockDim.x)
    dst[i] = src[i];
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

// ---------------------------------------------------------------------------------------------- host side
// One communicator context per process (one process per GPU).
struct PeerCtx {
  bool ready = false;
  PeerSet ps{};
  float* aux[kMaxRanks];
  float* table[kMaxRanks];
  float** table_dev = nullptr;     // device copy of the table pointers
  float* stage_mc = nullptr;
  float* replica_mc = nullptr;
  float* aux_mc = nullptr;
  unsigned int ticket = 0;
  int grid = 0;
  int table_cap = 0;
};
PeerCtx g_ctx;

// Cooperative launch: the driver refuses the launch when the grid cannot be co-resident (instead of a deadlock of
// blocks waiting for peers that were never scheduled).
template <typename... Args>
cudaError_t coop_launch(void (*kernel)(Args...), int grid, int block, size_t smem, cudaStream_t st, Args... args) {
  void* argv[] = {(void*)&args...};
  return cudaLaunchCooperativeKernel((const void*)kernel, dim3(grid), dim3(block), argv, smem, st);
}

PeerSet peers_for(int target, bool nvls) {
  PeerSet ps = g_ctx.ps;
  if (target == 1)
    for (int q = 0; q < ps.world; ++q) ps.replica[q] = g_ctx.aux[q];
  ps.stage_mc = nvls ? g_ctx.stage_mc : nullptr;
  ps.replica_mc = nvls ? (target == 1 ? g_ctx.aux_mc : g_ctx.replica_mc) : nullptr;
  return ps;
}

}  // namespace

extern "C" {

// Registers the symmetric buffers of this process' communicator.  stage / replica / aux: [D] floats per rank;
// signal: 64 uint32 per rank, mail: 128 doubles per rank, table: 2*table_cap floats per rank - all zero-initialised,
// all peer-mapped (host arrays of `world` device pointers).  *_mc: multicast addresses (nullable).  counters: 8
// zero-initialised uint32, err: 1 zero-initialised int, table_dev: room for `world` pointers - local device memory.
int dba_peer_init(int rank, int world, const void* const* stage, const void* const* replica, const void* const* aux,
                  const void* const* signal, const void* const* mail, const void* const* table, int table_cap,
                  void* stage_mc, void* replica_mc, void* aux_mc, unsigned int* counters, int* err, void* table_dev) {
  DBA_REQUIRE(world >= 1 && world <= kMaxRanks && rank >= 0 && rank < world, "dba_peer_init: bad rank/world %d/%d", rank,
              world);
  PeerCtx& c = g_ctx;
  c = PeerCtx();
  c.ps.rank = rank;
  c.ps.world = world;
  for (int q = 0; q < world; ++q) {
    c.ps.stage[q] = (float*)stage[q];
    c.ps.replica[q] = (float*)replica[q];
    c.aux[q] = (float*)aux[q];
    c.ps.signal[q] = (unsigned int*)signal[q];
    c.ps.mail[q] = (double*)mail[q];
    c.table[q] = (float*)table[q];
  }
  c.ps.counter = counters;
  c.ps.err = err;
  c.stage_mc = (float*)stage_mc; c.replica_mc = (float*)replica_mc; c.aux_mc = (float*)aux_mc;
  c.table_cap = table_cap;
  c.table_dev = (float**)table_dev;
  DBA_CHECK_CUDA(cudaMemcpy(c.table_dev, c.table, sizeof(float*) * world, cudaMemcpyHostToDevice));
  // one block per SM, as many as can be co-resident (cooperative launch requirement)
  int dev = 0, sms = 0, coop = 0, occ = 0;
  DBA_CHECK_CUDA(cudaGetDevice(&dev));
  DBA_CHECK_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  DBA_CHECK_CUDA(cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev));
  DBA_REQUIRE(coop != 0, "dba_peer_init: device %d does not support cooperative launches", dev);
  DBA_CHECK_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, weiszfeld_p2p_kernel, kCollThreads, 4096));
  DBA_REQUIRE(occ >= 1, "dba_peer_init: the collective kernels do not fit on an SM");
  c.grid = sms;
  c.ready = true;
  return DBA_OK;
}

/* 0 = fine; 1 = a cross-GPU barrier timed out on this GPU (a peer died or never launched the matching kernel). */
int dba_peer_error(void) {
  if (!g_ctx.ready) return 0;
  int e = 0;
  if (cudaMemcpy(&e, g_ctx.ps.err, sizeof(int), cudaMemcpyDeviceToHost) != cudaSuccess) return 2;
  return e;
}

// G (the replica) += coef * sum over ALL ranks' client rows (W[k] - G)  (+ sigma*N(0,1)), new G in every replica.
int dba_peer_fedavg(const float* W, long long ldw, int K, long long D, float coef, float sigma,
                    unsigned long long seed, int nvls, void* stream) {
  DBA_REQUIRE(g_ctx.ready, "dba_peer_fedavg: dba_peer_init has not been called");
  DBA_REQUIRE(K >= 0 && D > 0, "dba_peer_fedavg: K=%d D=%lld", K, D);
  PeerSet ps = peers_for(0, nvls != 0);
  DBA_CHECK_CUDA(coop_launch(fedavg_p2p_kernel, g_ctx.grid, kCollThreads, 0, (cudaStream_t)stream, W, ldw, K, D, coef, sigma, seed,
                             ps, ++g_ctx.ticket));
  DBA_COUNT_LAUNCH(1);
  return DBA_OK;
}

// target (0 = replica, 1 = aux)[:D] = beta*target + alpha * sum over all ranks of (K > 0 ? sum_k wv[k]*S[k] : stage).
int dba_peer_axpby(const float* S, long long lds, int K, const double* wv, long long D, float alpha, float beta,
                   int target, int nvls, void* stream) {
  DBA_REQUIRE(g_ctx.ready, "dba_peer_axpby: dba_peer_init has not been called");
  DBA_REQUIRE(D > 0 && K >= 0 && K <= 4096, "dba_peer_axpby: D=%lld K=%d", D, K);
  PeerSet ps = peers_for(target, nvls != 0);
  DBA_CHECK_CUDA(coop_launch(axpby_p2p_kernel, g_ctx.grid, kCollThreads, (size_t)K * sizeof(float), (cudaStream_t)stream, S, lds, K,
                             wv, D, alpha, beta, ps, ++g_ctx.ticket));
  DBA_COUNT_LAUNCH(1);
  return DBA_OK;
}

// Whole RFA loop over the deltas (W[k] - G) of all ranks: maxiter + 2 launches per rank, no host synchronisation,
// no NCCL.  Results: median in every rank's aux buffer; wv / distances of ALL clients (global slot order; this rank's
// K clients start at slot0) in this rank's result table [2][n_total]; state = (done, iterations, oracle calls);
// obj[(iterations ...)] see weiszfeld_p2p_kernel.  dist2: 2*kmax doubles, wts / wv: K floats, obj: 2 doubles,
// state: 4 ints - local scratch owned by the caller, zeroed here.
int dba_peer_weiszfeld(const float* W, long long ldw, int K, long long D, const float* G, const float* alphas,
                       int maxiter, float eps, double ftol, int slot0, int n_total, double* dist2, int kmax, float* wts,
                       float* wv, double* obj, int* state, void* stream) {
  DBA_REQUIRE(g_ctx.ready, "dba_peer_weiszfeld: dba_peer_init has not been called");
  DBA_REQUIRE(K >= 0 && K <= kmax && kmax <= 1024 && D > 0 && n_total <= g_ctx.table_cap && slot0 + K <= n_total,
              "dba_peer_weiszfeld: K=%d kmax=%d D=%lld n_total=%d slot0=%d", K, kmax, D, n_total, slot0);
  cudaStream_t st = (cudaStream_t)stream;
  DBA_CHECK_CUDA(cudaMemsetAsync(state, 0, 4 * sizeof(int), st));
  DBA_CHECK_CUDA(cudaMemsetAsync(obj, 0, 2 * sizeof(double), st));
  DBA_CHECK_CUDA(cudaMemsetAsync(dist2, 0, 2 * (size_t)kmax * sizeof(double), st));
  WzArgs a{W, ldw, D, G, alphas, dist2, wts, wv, obj, state, g_ctx.table_dev, K, kmax, 0, maxiter, slot0, n_total, eps,
           ftol};
  PeerSet ps = peers_for(1, false);         // the median lives in the auxiliary vector
  for (int it = 0; it <= maxiter + 1; ++it) {
    a.it = it;
    DBA_CHECK_CUDA(coop_launch(weiszfeld_p2p_kernel, g_ctx.grid, kCollThreads, (size_t)(K > 0 ? K : 1) * sizeof(float), st, a, ps,
                               ++g_ctx.ticket));
    DBA_COUNT_LAUNCH(1);
  }
  return DBA_OK;
}

// Stream-ordered barrier across all ranks (every rank must call it the same number of times).
int dba_peer_barrier(void* stream) {
  DBA_REQUIRE(g_ctx.ready, "dba_peer_barrier: dba_peer_init has not been called");
  PeerSet ps = peers_for(0, false);
  DBA_CHECK_CUDA(coop_launch(peer_barrier_kernel, 1, 64, 0, (cudaStream_t)stream, ps, ++g_ctx.ticket));
  DBA_COUNT_LAUNCH(1);
  return DBA_OK;
}

// rows: DEVICE array of n pointers to float rows of length D (local or peer-mapped) -> out[n][ldo] in local memory.
int dba_gather_rows_ptrs(const float* const* rows, int n, long long D, float* out, long long ldo, void* stream) {
  DBA_REQUIRE(n > 0 && n <= 65535 && D > 0 && ldo >= D, "dba_gather_rows_ptrs: n=%d D=%lld ldo=%lld", n, D, ldo);
  int gx = (int)((D / 4 + 255) / 256);
  if (gx > 64) gx = 64;
  if (gx < 1) gx = 1;
  gather_rows_ptrs_kernel<<<dim3(gx, n), 256, 0, (cudaStream_t)stream>>>(rows, n, D, out, ldo);
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
