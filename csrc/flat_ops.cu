// Flat-slab kernels: everything the reference does tensor-by-tensor in Python loops over a state_dict
// (SGD step, FoolsGold gradient accumulation, poison scaling, delta extraction, norms, FedAvg, Weiszfeld, FoolsGold)
// is ONE launch over [K, D] slabs here.  All buffers are fp32, 16-byte aligned rows (ld % 4 == 0) are vectorised.
//
// Reference semantics: SGD image_train.py:33-35,102; grad-sum image_train.py:94-100; scaling image_train.py:166-171;
// norms helper.py:60-71,376-381; FedAvg helper.py:240-257; RFA helper.py:295-418; FoolsGold helper.py:259-293,527-607.
#include "common.cuh"
#include <curand_kernel.h>
#include <stdarg.h>

static thread_local char g_err[512] = "";        // per host thread: dba_last_error() reports the caller's own failure
unsigned long long g_dba_launches = 0;
int g_dba_act_mode = 0;
int g_dba_pdl = 1;
extern "C" void dba_set_pdl(int on) { g_dba_pdl = on ? 1 : 0; }
extern "C" int dba_get_pdl() { return g_dba_pdl; }
// Activation element mode of the network kernels: 0 = bf16 (kind::f16 MMA), 1 = fp32 activations with kind::tf32 MMA.
extern "C" int dba_set_act_dtype(int mode) {
  if (mode != 0 && mode != 1) {
    dba_set_error("dba_set_act_dtype: mode %d (0 = bf16, 1 = tf32)", mode);
    return DBA_ERR_ARG;
  }
  g_dba_act_mode = mode;
  return DBA_OK;
}
extern "C" int dba_get_act_dtype() { return g_dba_act_mode; }
extern "C" const char* dba_last_error() { return g_err; }
void dba_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
extern "C" unsigned long long dba_launch_count() { return g_dba_launches; }
extern "C" void dba_reset_launch_count() { g_dba_launches = 0; }
// Kernels launched by replaying a captured CUDA graph are accounted for by the caller (n per replay).
extern "C" void dba_add_launch_count(unsigned long long n) { DBA_COUNT_LAUNCH(n); }

using namespace dba;

// ------------------------------------------------------------------------------------------ lock-step staging
// One launch per training step instead of six indexed copies: row r of the step's slab receives the schedule entries
// (sample indices, flip flags, valid count, active flag, log slot, learning rate) of client rows[r] (or r) at `step`.
struct StageArgs {
  const long long* idx;        // [K][T][B]
  const unsigned char* flip;   // [K][T][B] or null
  const int* nvalid;           // [K][T]
  const unsigned char* active; // [K][T]
  const long long* slot;       // [K][T]
  const float* lr;             // [K][T]
  const long long* rows;       // [n] or null
  long long* st_idx;           // [n][B]
  unsigned char* st_flip;      // [n][B] or null
  int* st_nvalid;
  unsigned char* st_active;
  long long* st_slot;
  float* st_lr;
  int T, B, step;
};
__global__ void __launch_bounds__(128) stage_step_kernel(const StageArgs a) {
  const int r = blockIdx.x;
  const long long k = a.rows ? a.rows[r] : r;
  const long long e = k * a.T + a.step;
  for (int b = threadIdx.x; b < a.B; b += blockDim.x) {
    a.st_idx[(long long)r * a.B + b] = a.idx[e * a.B + b];
    if (a.st_flip) a.st_flip[(long long)r * a.B + b] = a.flip[e * a.B + b];
  }
  if (threadIdx.x == 0) {
    a.st_nvalid[r] = a.nvalid[e];
    a.st_active[r] = a.active[e];
    a.st_slot[r] = a.slot[e];
    a.st_lr[r] = a.lr[e];
  }
}
extern "C" int dba_stage_step(const long long* idx, const unsigned char* flip, const int* nvalid,
                              const unsigned char* active, const long long* slot, const float* lr,
                              const long long* rows, int n, int T, int B, int step, long long* st_idx,
                              unsigned char* st_flip, int* st_nvalid, unsigned char* st_active, long long* st_slot,
                              float* st_lr, void* stream) {
  DBA_REQUIRE(n > 0 && T > 0 && B > 0 && step >= 0 && step < T, "dba_stage_step: n=%d T=%d B=%d step=%d", n, T, B, step);
  DBA_REQUIRE((flip == nullptr) == (st_flip == nullptr), "dba_stage_step: flip / st_flip must be given together");
  StageArgs a{idx, flip, nvalid, active, slot, lr, rows, st_idx, st_flip, st_nvalid, st_active, st_slot, st_lr, T, B, step};
  stage_step_kernel<<<n, 128, 0, (cudaStream_t)stream>>>(a);
  DBA_COUNT_LAUNCH(1);
  DBA_CHECK_LAUNCH();
  return DBA_OK;
}

static inline int grid_for(long long work_items, int block, int max_waves = 8) {
  long long g = (work_items + block - 1) / block;
  long long cap = (long long)kNumSMs * max_waves;
  if (g > cap) g = cap;
  if (g < 1) g = 1;
  return (int)g;
}

// ------------------------------------------------------------------------------------------------------ SGD
// w -= lr_k * v,  v = mu*v + (g + wd*w);  optional s += g (FoolsGold feature: raw gradient before weight decay).
// Rows with active[k]==0 are untouched.  One launch for all K clients of the rank.
__global__ void __launch_bounds__(256) sgd_step_kernel(float* __restrict__ W, long long ldw, float* __restrict__ V,
                                                       const float* __restrict__ Gr, float* __restrict__ S,
                                                       long long ldv, const float* __restrict__ lr,
                                                       const unsigned char* __restrict__ active, int K, long long Dp,
                                                       float mu, float wd, int vec_ok) {
  const long long nvec = vec_ok ? (Dp >> 2) : 0;
  const long long total = (long long)K * nvec;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    int k = (int)(i / nvec);
    if (active && !active[k]) continue;
    long long j = (i - (long long)k * nvec) << 2;
    float4 w = *reinterpret_cast<float4*>(W + k * ldw + j);
    float4 v = *reinterpret_cast<float4*>(V + k * ldv + j);
    float4 g = *reinterpret_cast<const float4*>(Gr + k * ldv + j);
    if (S) {
      float4 s = *reinterpret_cast<float4*>(S + k * ldv + j);
      s.x += g.x; s.y += g.y; s.z += g.z; s.w += g.w;
      *reinterpret_cast<float4*>(S + k * ldv + j) = s;
    }
    const float l = lr[k];
    v.x = mu * v.x + (g.x + wd * w.x); v.y = mu * v.y + (g.y + wd * w.y);
    v.z = mu * v.z + (g.z + wd * w.z); v.w = mu * v.w + (g.w + wd * w.w);
    w.x -= l * v.x; w.y -= l * v.y; w.z -= l * v.z; w.w -= l * v.w;
    *reinterpret_cast<float4*>(V + k * ldv + j) = v;
    *reinterpret_cast<float4*>(W + k * ldw + j) = w;
  }
  // scalar tail of every row (everything when rows are not 16-byte aligned)
  const long long tail0 = nvec << 2;
  const long long ntail = Dp - tail0;
  for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < (long long)K * ntail;
       t += (long long)gridDim.x * blockDim.x) {
    int k = (int)(t / ntail);
    long long j = tail0 + (t - (long long)k * ntail);
    if (active && !active[k]) continue;
    float w = W[k * ldw + j], g = Gr[k * ldv + j];
    if (S) S[k * ldv + j] += g;
    float v = mu * V[k * ldv + j] + (g + wd * w);
    V[k * ldv + j] = v;
    W[k * ldw + j] = w - lr[k] * v;
  }
}

// W [K rows, stride ldw] holds the parameters in its first Dp columns; V, Gr, S are [K rows, stride ldv].
extern "C" int dba_sgd_step(float* W, long long ldw, float* V, const float* Gr, float* S, long long ldv,
                            const float* lr, const unsigned char* active, int K, long long Dp, float mu, float wd,
                            void* stream) {
  DBA_REQUIRE(K > 0 && Dp > 0 && ldw >= Dp && ldv >= Dp, "dba_sgd_step: K=%d Dp=%lld ldw=%lld ldv=%lld", K, Dp, ldw,
              ldv);
  const int vec_ok = ((uintptr_t)W & 15) == 0 && ((uintptr_t)V & 15) == 0 && ((uintptr_t)Gr & 15) == 0 &&
                     (!S || ((uintptr_t)S & 15) == 0) && (K == 1 || ((ldw & 3) == 0 && (ldv & 3) == 0));
  sgd_step_kernel<<<grid_for((long long)K * (Dp >> 2) + 1, 256), 256, 0, (cudaStream_t)stream>>>(
      W, ldw, V, Gr, S, ldv, lr, active, K, Dp, mu, wd, vec_ok);
  DBA_COUNT_LAUNCH(1);
  DBA_CHECK_LAUNCH();
  return DBA_OK;
}

// ------------------------------------------------------------------- distance-regularised poison loss (K11, W14)
// Reference: loss = alpha*CE + (1-alpha)*||w - w_anchor||_2 for poisoning clients (image_train.py:85-90,
// helper.py:111-123).  After the backward pass produced the CE gradient in Gr, this turns it into the gradient of
// the combined loss,  Gr = alpha*Gr + (1-alpha)*(w - a)/||w - a||  (0 at w == a, torch.norm's subgradient), and adds
// the distance to the per-(client, internal epoch) trace so the logged loss can be rebuilt as alpha*CE + (1-alpha)*dist.
__global__ void __launch_bounds__(256) dist_grad_kernel(float* __restrict__ Gr, long long ldv,
                                                        const float* __restrict__ W, long long ldw,
                                                        const float* __restrict__ A, long long lda,
                                                        const double* __restrict__ dist,
                                                        const long long* __restrict__ poison_n,
                                                        const unsigned char* __restrict__ active,
                                                        const long long* __restrict__ slot, float* __restrict__ mf_dist,
                                                        int ld_m, float alpha, int K, long long Dp) {
  const long long total = (long long)K * Dp;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int k = (int)(i / Dp);
    if (poison_n[k] <= 0 || (active && !active[k])) continue;
    const long long j = i - (long long)k * Dp;
    const float d = (float)dist[k];
    const float inv = d > 0.f ? 1.f / d : 0.f;
    Gr[k * ldv + j] = alpha * Gr[k * ldv + j] + (1.f - alpha) * (W[k * ldw + j] - A[k * lda + j]) * inv;
    if (j == 0) mf_dist[k * ld_m + (int)slot[k]] += (float)dist[k];
  }
}
extern "C" int dba_dist_grad(float* Gr, long long ldv, const float* W, long long ldw, const float* A, long long lda,
                             const double* dist, const long long* poison_n, const unsigned char* active,
                             const long long* slot, float* mf_dist, int ld_m, float alpha, int K, long long Dp,
                             void* stream) {
  DBA_REQUIRE(K > 0 && Dp > 0, "dba_dist_grad: K=%d Dp=%lld", K, Dp);
  dist_grad_kernel<<<grid_for((long long)K * Dp, 256), 256, 0, (cudaStream_t)stream>>>(
      Gr, ldv, W, ldw, A, lda, dist, poison_n, active, slot, mf_dist, ld_m, alpha, K, Dp);
  DBA_COUNT_LAUNCH(1);
  DBA_CHECK_LAUNCH();
  return DBA_OK;
}

// ------------------------------------------------------------------------------------------- poison scaling
// W[k] = A[k] + (W[k] - A[k]) * s[k]   over the whole state row (params + BN buffers).
__global__ void __launch_bounds__(256) scale_rows_kernel(float* __restrict__ W, long long ldw,
                                                         const float* __restrict__ A, long long lda,
                                                         const float* __restrict__ s, int K, long long D) {
  const long long total = (long long)K * D;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    int k = (int)(i / D);
    long long j = i - (long long)k * D;
    float a = A[k * lda + j];
    W[k * ldw + j] = a + (W[k * ldw + j] - a) * s[k];
  }
}
extern "C" int dba_scale_rows(float* W, long long ldw, const float* A, long long lda, const float* s, int K,
                              long long D, void* stream) {
  DBA_REQUIRE(K > 0 && D > 0, "dba_scale_rows: K=%d D=%lld", K, D);
  scale_rows_kernel<<<grid_for((long long)K * D, 256), 256, 0, (cudaStream_t)stream>>>(W, ldw, A, lda, s, K, D);
  DBA_COUNT_LAUNCH(1);
  DBA_CHECK_LAUNCH();
  return DBA_OK;
}

// ---------------------------------------------------------------------------------- row-wise squared distances
// out[k] = sum_i (X[k,i] - Y[k*ldy + i] - Z[i])^2   (ldy == 0 broadcasts one Y row; Y / Z may be null = 0).
// Deterministic two-stage reduction: stage 1 writes per-(row, block) partial sums into `scratch`
// (K * DBA_SQDIST_BLOCKS doubles), stage 2 folds them in a fixed order.  No atomics, no zero-init needed.
#define DBA_SQDIST_BLOCKS 128
__global__ void __launch_bounds__(256) sqdist_stage1(const float* __restrict__ X, long long ldx,
                                                     const float* __restrict__ Y, long long ldy,
                                                     const float* __restrict__ Z, long long n,
                                                     double* __restrict__ scratch) {
  __shared__ double red[32];
  const int k = blockIdx.y;
  const float* x = X + k * ldx;
  const float* y = Y ? Y + k * ldy : nullptr;
  double acc = 0.0;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n;
       i += (long long)gridDim.x * blockDim.x) {
    float d = x[i];
    if (y) d -= y[i];
    if (Z) d -= Z[i];
    acc += (double)d * (double)d;
  }
  acc = block_sum(acc, red);
  if (threadIdx.x == 0) scratch[(long long)k * gridDim.x + blockIdx.x] = acc;
}
__global__ void sqdist_stage2(const double* __restrict__ scratch, int nblk, double* __restrict__ out, int take_sqrt) {
  const int k = blockIdx.x;
  double acc = 0.0;
  for (int i = threadIdx.x; i < nblk; i += 32) acc += scratch[(long long)k * nblk + i];
  acc = warp_sum(acc);
  if (threadIdx.x == 0) out[k] = take_sqrt ? sqrt(acc) : acc;
}
extern "C" long long dba_rows_sqdist_scratch_elems(int K) { return (long long)K * DBA_SQDIST_BLOCKS; }

extern "C" int dba_rows_sqdist(const float* X, long long ldx, const float* Y, long long ldy, const float* Z, int K,
                               long long n, double* out, int take_sqrt, double* scratch, long long scratch_elems,
                               void* stream) {
  DBA_REQUIRE(K > 0 && n > 0, "dba_rows_sqdist: K=%d n=%lld", K, n);
  DBA_REQUIRE(scratch && scratch_elems >= dba_rows_sqdist_scratch_elems(K),
              "dba_rows_sqdist: scratch needs %lld doubles, got %lld", dba_rows_sqdist_scratch_elems(K), scratch_elems);
  dim3 g(DBA_SQDIST_BLOCKS, K);
  sqdist_stage1<<<g, 256, 0, (cudaStream_t)stream>>>(X, ldx, Y, ldy, Z, n, scratch);
  sqdist_stage2<<<K, 32, 0, (cudaStream_t)stream>>>(scratch, DBA_SQDIST_BLOCKS, out, take_sqrt);
  DBA_COUNT_LAUNCH(2);
  DBA_CHECK_LAUNCH();
  return DBA_OK;
}

// --------------------------------------------------------------------------- row distances without a scratch argument
// Same computation as dba_rows_sqdist, for callers that have no buffer to hand in as scratch and must not create
// one: a process that has given all device memory to its own pool (so cudaMalloc fails or fragments it), or a side
// thread - say a monitor sampling update norms next to a running training stream - for which cudaMalloc / cudaFree
// are off limits because they synchronise the whole device.  The partial sums live in a fixed array inside the
// library image instead; every GPU that loads the library carries these 64 KiB until the process exits.
//
// Limits and rules for the caller:
//   * at most DBA_NOALLOC_MAX_ROWS (= 64) rows per call; more is refused with an error, nothing is launched;
//   * the first call on each device looks the array's device address up (cudaGetSymbolAddress with that device
//     current) and remembers it - make that first call during start-up, before any stream capture begins;
//   * all calls share the one array, which still holds the previous call's partial sums.  Nothing has to start at
//     zero: stage 1 overwrites every slot stage 2 reads.  But two calls whose kernels can overlap on the GPU
//     (different streams, no ordering between them) race on it - order them yourself;
//   * no cudaMalloc, cudaFree or other allocating routine is called, here or in anything this function calls;
//   * the device that owns X / out / the stream must be the CURRENT device of the calling thread (the array's address
//     is looked up and remembered per current device);
//   * host-side state is safe to share between threads: the per-device address cache is written without a lock, but
//     every writer stores the same value (idempotent), the launch counter is atomic and the error text is per thread.
#define DBA_NOALLOC_MAX_ROWS 64
static __device__ __align__(256) double g_noalloc_sqdist_scratch[DBA_NOALLOC_MAX_ROWS * DBA_SQDIST_BLOCKS];
#define DBA_NOALLOC_MAX_DEVICES 64
static double* g_noalloc_scratch_addr[DBA_NOALLOC_MAX_DEVICES] = {nullptr};

extern "C" int dba_rows_sqdist_noalloc_max_rows() { return DBA_NOALLOC_MAX_ROWS; }

extern "C" int dba_rows_sqdist_noalloc(const float* X, long long ldx, const float* Y, long long ldy, const float* Z,
                                       int K, long long n, double* out, int take_sqrt, void* stream) {
  DBA_REQUIRE(K > 0 && n > 0, "dba_rows_sqdist_noalloc: K=%d n=%lld", K, n);
  if (K > DBA_NOALLOC_MAX_ROWS) {
    dba_set_error("dba_rows_sqdist_noalloc: %d rows exceed the built-in scratch (max %d rows per call); split the "
                  "call or use dba_rows_sqdist with your own scratch buffer", K, DBA_NOALLOC_MAX_ROWS);
    return DBA_ERR_LIMIT;
  }
  int dev = -1;
  DBA_CHECK_CUDA(cudaGetDevice(&dev));
  DBA_REQUIRE(dev >= 0 && dev < DBA_NOALLOC_MAX_DEVICES, "dba_rows_sqdist_noalloc: device index %d not supported", dev);
  if (g_noalloc_scratch_addr[dev] == nullptr) {
    // In host code the array's name is only a handle; the device address of this device's copy has to be asked for.
    double* p = nullptr;
    DBA_CHECK_CUDA(cudaGetSymbolAddress((void**)&p, g_noalloc_sqdist_scratch));
    DBA_REQUIRE(p != nullptr, "dba_rows_sqdist_noalloc: no device address for the built-in scratch");
    g_noalloc_scratch_addr[dev] = p;
  }
  double* scratch = g_noalloc_scratch_addr[dev];
  dim3 g(DBA_SQDIST_BLOCKS, K);
  sqdist_stage1<<<g, 256, 0, (cudaStream_t)stream>>>(X, ldx, Y, ldy, Z, n, scratch);
  sqdist_stage2<<<K, 32, 0, (cudaStream_t)stream>>>(scratch, DBA_SQDIST_BLOCKS, out, take_sqrt);
  DBA_COUNT_LAUNCH(2);
  DBA_CHECK_LAUNCH();
  return DBA_OK;
}

// --------------------------------------------------------------------------------------------------- FedAvg
// Single-GPU fused form of  delta extraction (K13) + accumulate (R2) + apply (K15) + DP noise (K18) + next round's
// global->client broadcast (R1):   G[i] += coef * sum_k (W[k,i] - G[i]) + sigma * N(0,1);   W[k,i] = G[i] if bcast.
__global__ void __launch_bounds__(256) fedavg_kernel(float* __restrict__ G, float* __restrict__ W, long long ldw, int K,
                                                     long long D, float coef, float sigma, unsigned long long seed,
                                                     int bcast) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < D;
       i += (long long)gridDim.x * blockDim.x) {
    const float g = G[i];
    float acc = 0.f;
    for (int k = 0; k < K; ++k) acc += W[k * ldw + i] - g;
    float out = g + coef * acc;
    if (sigma > 0.f) {
      curandStatePhilox4_32_10_t st;
      curand_init(seed, (unsigned long long)i, 0, &st);
      out += sigma * curand_normal(&st);
    }
    G[i] = out;
    if (bcast)
      for (int k = 0; k < K; ++k) W[k * ldw + i] = out;
  }
}
extern "C" int dba_fedavg_update(float* G, float* W, long long ldw, int K, long long D, float coef, float sigma,
                                 unsigned long long seed, int bcast, void* stream) {
  DBA_REQUIRE(K >= 0 && D > 0, "dba_fedavg_update: K=%d D=%lld", K, D);
  fedavg_kernel<<<grid_for(D, 256), 256, 0, (cudaStream_t)stream>>>(G, W, ldw, K, D, coef, sigma, seed, bcast);
  DBA_COUNT_LAUNCH(1);
  DBA_CHECK_LAUNCH();
  return DBA_OK;
}

// ------------------------------------------------------------------------------------------------ Weiszfeld
// state[0] = done flag, state[1] = iterations run, state[2] = oracle calls.  obj[0] = current objective.
// One iteration = wavg kernel (new median from the distances of the previous one) + distance kernels.
// Every block recomputes the n weights from dist[] in shared memory (n <= 1024), so there is no extra launch and
// no host round trip; once the device-side convergence flag is set the remaining launches fall through.
__global__ void __launch_bounds__(256) weiszfeld_wavg_kernel(const float* __restrict__ W, long long ldw,
                                                             const float* __restrict__ G, int n, long long D,
                                                             const float* __restrict__ alphas,
                                                             const double* __restrict__ dist, float eps, int first,
                                                             float* __restrict__ median, float* __restrict__ wts_out,
                                                             const int* __restrict__ state) {
  extern __shared__ float w_s[];
  __shared__ float red[32];
  if (state[0]) return;
  float part = 0.f;
  for (int c = threadIdx.x; c < n; c += blockDim.x) {
    float w = first ? alphas[c] : alphas[c] / fmaxf(eps, (float)dist[c]);
    w_s[c] = w;
    part += w;
  }
  float tot = block_sum(part, red, true);
  for (int c = threadIdx.x; c < n; c += blockDim.x) {
    w_s[c] /= tot;
    if (blockIdx.x == 0 && wts_out) wts_out[c] = w_s[c];
  }
  __syncthreads();
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < D;
       i += (long long)gridDim.x * blockDim.x) {
    const float g = G[i];
    float acc = 0.f;
    for (int c = 0; c < n; ++c) acc += w_s[c] * (W[c * ldw + i] - g);
    median[i] = acc;
  }
}

// Guarded stage 1 of the distance reduction:  partial sums of || (W[c] - G) - median ||^2.
__global__ void __launch_bounds__(256) weiszfeld_dist_stage1(const float* __restrict__ W, long long ldw,
                                                             const float* __restrict__ G,
                                                             const float* __restrict__ median, long long D,
                                                             double* __restrict__ scratch,
                                                             const int* __restrict__ state) {
  __shared__ double red[32];
  if (state[0]) return;
  const int c = blockIdx.y;
  const float* x = W + c * ldw;
  double acc = 0.0;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < D;
       i += (long long)gridDim.x * blockDim.x) {
    float d = x[i] - G[i] - median[i];
    acc += (double)d * (double)d;
  }
  acc = block_sum(acc, red);
  if (threadIdx.x == 0) scratch[(long long)c * gridDim.x + blockIdx.x] = acc;
}

// Stage 2 + objective + convergence test + bookkeeping (one thread block; warp w folds client rows w, w+8, ...).
__global__ void __launch_bounds__(256) weiszfeld_finish_kernel(const float* __restrict__ alphas,
                                                               const double* __restrict__ scratch, int nblk,
                                                               double* __restrict__ dist, int n, double ftol,
                                                               double* __restrict__ obj,
                                                               const float* __restrict__ wts_cur,
                                                               float* __restrict__ wv, int* __restrict__ state,
                                                               int first) {
  __shared__ double red[32];
  if (state[0]) return;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
  for (int c = warp; c < n; c += nwarp) {
    double acc = 0.0;
    for (int i = lane; i < nblk; i += 32) acc += scratch[(long long)c * nblk + i];
    acc = warp_sum(acc);
    if (lane == 0) dist[c] = sqrt(acc);
  }
  __syncthreads();
  double part = 0.0;
  for (int c = threadIdx.x; c < n; c += blockDim.x) part += (double)alphas[c] * dist[c];
  const double o = block_sum(part, red, true);
  const double prev = obj[0];
  const bool conv = !first && (fabs(prev - o) < ftol * o);
  const int iters_before = state[1];
  __syncthreads();
  // helper.py:348-352: wv is only refreshed by iterations that did not converge (Q8).  The reference crashes when
  // the very first iteration converges (wv is None); here that case reports that iteration's weights.
  if (first || !conv || iters_before == 0)
    for (int c = threadIdx.x; c < n; c += blockDim.x) wv[c] = wts_cur[c];
  if (threadIdx.x == 0) {
    obj[0] = o;
    state[2] += 1;
    if (!first) state[1] += 1;
    if (conv) state[0] = 1;
  }
}

// Runs the whole RFA loop on the device: 3*(maxiter+1) fixed launches, no host synchronisation; once the
// device-side convergence flag is set the remaining launches fall through.
// Buffers: median[D], dist[n] (double), wts[n], wv[n], obj[1] (double), state[4] (int; [0]=done, [1]=iterations,
// [2]=oracle calls), scratch >= dba_rows_sqdist_scratch_elems(n) doubles.
extern "C" int dba_weiszfeld(const float* W, long long ldw, const float* G, int n, long long D, const float* alphas,
                             int maxiter, float eps, double ftol, float* median, double* dist, float* wts, float* wv,
                             double* obj, int* state, double* scratch, long long scratch_elems, void* stream) {
  DBA_REQUIRE(n > 0 && n <= 1024 && D > 0, "dba_weiszfeld: n=%d D=%lld", n, D);
  DBA_REQUIRE(scratch && scratch_elems >= dba_rows_sqdist_scratch_elems(n), "dba_weiszfeld: scratch too small");
  cudaStream_t st = (cudaStream_t)stream;
  DBA_CHECK_CUDA(cudaMemsetAsync(state, 0, 4 * sizeof(int), st));
  DBA_CHECK_CUDA(cudaMemsetAsync(obj, 0, sizeof(double), st));
  const int grid = grid_for(D, 256);
  dim3 g1(DBA_SQDIST_BLOCKS, n);
  for (int it = 0; it <= maxiter; ++it) {
    const int first = it == 0;
    weiszfeld_wavg_kernel<<<grid, 256, n * sizeof(float), st>>>(W, ldw, G, n, D, alphas, dist, eps, first, median, wts,
                                                               state);
    weiszfeld_dist_stage1<<<g1, 256, 0, st>>>(W, ldw, G, median, D, scratch, state);
    weiszfeld_finish_kernel<<<1, 256, 0, st>>>(alphas, scratch, DBA_SQDIST_BLOCKS, dist, n, ftol, obj, wts, wv, state,
                                               first);
    DBA_COUNT_LAUNCH(3);
  }
  DBA_CHECK_LAUNCH();
  return DBA_OK;
}

// ------------------------------------------------------------------------------------------------ FoolsGold
// Gram matrix of the n history rows in fp64 (the reference uses float64 numpy + sklearn): gram[i,j] = <f_i, f_j>.
__global__ void __launch_bounds__(256) gram_kernel(const double* __restrict__ F, long long ldf, int n, long long L,
                                                   double* __restrict__ gram) {
  __shared__ double red[32];
  const int i = blockIdx.x, j = blockIdx.y;
  if (j < i) return;
  const double* a = F + i * ldf;
  const double* b = F + j * ldf;
  double acc = 0.0;
  for (long long t = threadIdx.x; t < L; t += blockDim.x) acc += a[t] * b[t];
  acc = block_sum(acc, red);
  if (threadIdx.x == 0) {
    gram[(long long)i * n + j] = acc;
    gram[(long long)j * n + i] = acc;
  }
}
// cosine similarity - I, pardoning, 1 - max, clip, rescale, logit (helper.py:574-607).  One block, n <= 1024.
__global__ void foolsgold_finish_kernel(const double* __restrict__ gram, int n, double* __restrict__ cs,
                                        double* __restrict__ wv, double* __restrict__ alpha) {
  extern __shared__ double sm[];
  double* nrm = sm;        // [n]
  double* maxcs = sm + n;  // [n]
  __shared__ double red[32];
  const int t = threadIdx.x;
  for (int i = t; i < n; i += blockDim.x) {
    double d = sqrt(gram[(long long)i * n + i]);
    nrm[i] = d == 0.0 ? 1.0 : d;
  }
  __syncthreads();
  for (long long e = t; e < (long long)n * n; e += blockDim.x) {
    int i = (int)(e / n), j = (int)(e % n);
    cs[e] = gram[e] / (nrm[i] * nrm[j]) - (i == j ? 1.0 : 0.0);
  }
  __syncthreads();
  for (int i = t; i < n; i += blockDim.x) {
    double m = -1e300;
    for (int j = 0; j < n; ++j) m = fmax(m, cs[(long long)i * n + j]);
    maxcs[i] = m;
  }
  __syncthreads();
  for (int i = t; i < n; i += blockDim.x) {
    double m = -1e300;
    for (int j = 0; j < n; ++j) {
      double v = cs[(long long)i * n + j];
      if (i != j && maxcs[i] < maxcs[j]) v = v * maxcs[i] / maxcs[j];
      m = fmax(m, v);
    }
    alpha[i] = m;
    double w = 1.0 - m;
    w = fmin(fmax(w, 0.0), 1.0);
    wv[i] = w;
  }
  __syncthreads();
  double mx = 0.0;
  for (int i = t; i < n; i += blockDim.x) mx = fmax(mx, wv[i]);
  // block max via the sum helper's smem: do a simple shared reduction
  __syncthreads();
  red[0] = 0.0;
  __syncthreads();
  for (int o = 16; o > 0; o >>= 1) mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  __shared__ double wmax[32];
  if ((t & 31) == 0) wmax[t >> 5] = mx;
  __syncthreads();
  if (t == 0) {
    double m = 0.0;
    for (int w = 0; w < (blockDim.x + 31) / 32; ++w) m = fmax(m, wmax[w]);
    red[0] = m;
  }
  __syncthreads();
  mx = red[0];
  for (int i = t; i < n; i += blockDim.x) {
    double w = wv[i];
    if (mx != 0.0) w = w / mx;
    if (w == 1.0) w = 0.99;
    w = log(w / (1.0 - w)) + 0.5;
    if (isnan(w)) w = 0.0;
    if (w > 1.0) w = 1.0;   // also +inf
    if (w < 0.0) w = 0.0;   // also -inf
    wv[i] = w;
  }
}
// history[id[c]] += feat[c] (fp64 accumulate of the fp32 last-layer gradient sums) and gather the rows to look at.
__global__ void fg_history_kernel(double* __restrict__ H, long long ldh, const int* __restrict__ ids,
                                  const float* __restrict__ feat, long long ldf, int n, long long L, int use_memory,
                                  double* __restrict__ out, long long ldo) {
  const int c = blockIdx.y;
  double* h = H + (long long)ids[c] * ldh;
  for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < L; t += (long long)gridDim.x * blockDim.x) {
    double f = (double)feat[c * ldf + t];
    double nv = h[t] + f;
    h[t] = nv;
    out[c * ldo + t] = use_memory ? nv : f;
  }
}
// Same, but every feature row is given by its own pointer: rows that live in another GPU's (peer-mapped) symmetric
// buffer are pulled over NVLink by the loads of this kernel - the all-gather of FoolsGold's client features is fused
// into the history update and never materialised (F3 of SURVEY §2.4).
__global__ void fg_history_ptrs_kernel(double* __restrict__ H, long long ldh, const int* __restrict__ ids,
                                       const float* const* __restrict__ feat_rows, int n, long long L, int use_memory,
                                       double* __restrict__ out, long long ldo) {
  const int c = blockIdx.y;
  double* h = H + (long long)ids[c] * ldh;
  const float* f_row = feat_rows[c];
  for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < L; t += (long long)gridDim.x * blockDim.x) {
    double f = (double)f_row[t];
    double nv = h[t] + f;
    h[t] = nv;
    out[c * ldo + t] = use_memory ? nv : f;
  }
}
extern "C" int dba_foolsgold_weights_ptrs(double* history, long long ldh, const int* ids, const float* const* feat_rows,
                                          int n, long long L, int use_memory, double* work_rows, double* gram,
                                          double* cs, double* wv, double* alpha, void* stream) {
  DBA_REQUIRE(n > 0 && n <= 1024 && L > 0, "dba_foolsgold_weights_ptrs: n=%d L=%lld", n, L);
  cudaStream_t st = (cudaStream_t)stream;
  dim3 gh(grid_for(L, 256, 2), n);
  fg_history_ptrs_kernel<<<gh, 256, 0, st>>>(history, ldh, ids, feat_rows, n, L, use_memory, work_rows, L);
  dim3 gg(n, n);
  gram_kernel<<<gg, 256, 0, st>>>(work_rows, L, n, L, gram);
  foolsgold_finish_kernel<<<1, 256, 2 * n * sizeof(double), st>>>(gram, n, cs, wv, alpha);
  DBA_COUNT_LAUNCH(3);
  DBA_CHECK_LAUNCH();
  return DBA_OK;
}

// Client ids must be distinct within one call (they are: a client is selected at most once per round).
extern "C" int dba_foolsgold_weights(double* history, long long ldh, const int* ids, const float* feat, long long ldf,
                                     int n, long long L, int use_memory, double* work_rows, double* gram, double* cs,
                                     double* wv, double* alpha, void* stream) {
  DBA_REQUIRE(n > 0 && n <= 1024 && L > 0, "dba_foolsgold_weights: n=%d L=%lld", n, L);
  cudaStream_t st = (cudaStream_t)stream;
  dim3 gh(grid_for(L, 256, 2), n);
  fg_history_kernel<<<gh, 256, 0, st>>>(history, ldh, ids, feat, ldf, n, L, use_memory, work_rows, L);
  dim3 gg(n, n);
  gram_kernel<<<gg, 256, 0, st>>>(work_rows, L, n, L, gram);
  foolsgold_finish_kernel<<<1, 256, 2 * n * sizeof(double), st>>>(gram, n, cs, wv, alpha);
  DBA_COUNT_LAUNCH(3);
  DBA_CHECK_LAUNCH();
  return DBA_OK;
}

// G[i] -= lr * (eta/n_total * sum_c wv[c]*S[c,i] + decay*G[i])   (weighted gradient reduce + the server SGD step)
__global__ void __launch_bounds__(256) foolsgold_apply_kernel(float* __restrict__ G, const float* __restrict__ S,
                                                              long long lds, const double* __restrict__ wv, int K,
                                                              long long Dp, float scale, float lr, float decay) {
  extern __shared__ float wv_s[];
  for (int c = threadIdx.x; c < K; c += blockDim.x) wv_s[c] = (float)wv[c];
  __syncthreads();
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < Dp;
       i += (long long)gridDim.x * blockDim.x) {
    float acc = 0.f;
    for (int c = 0; c < K; ++c) acc += wv_s[c] * S[c * lds + i];
    const float g = G[i];
    G[i] = g - lr * (scale * acc + decay * g);
  }
}
extern "C" int dba_foolsgold_apply(float* G, const float* S, long long lds, const double* wv, int K, long long Dp,
                                   float eta_over_n, float lr, float decay, void* stream) {
  DBA_REQUIRE(K >= 0 && Dp > 0, "dba_foolsgold_apply: K=%d Dp=%lld", K, Dp);
  foolsgold_apply_kernel<<<grid_for(Dp, 256), 256, (K + 1) * sizeof(float), (cudaStream_t)stream>>>(
      G, S, lds, wv, K, Dp, eta_over_n, lr, decay);
  DBA_COUNT_LAUNCH(1);
  DBA_CHECK_LAUNCH();
  return DBA_OK;
}
