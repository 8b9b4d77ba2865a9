// LOAN MLP (91 -> 46 -> 23 -> 9, Linear -> Dropout(0.5) -> ReLU twice, then Linear): the whole local training of a
// client in ONE persistent thread block, and batched evaluation of M models.
//
// Reference: models/loan_model.py:10-27 trained by loan_train.py:11-261 (one client after another, ~30 kernel
// launches and two host syncs per batch: the NaN check in LoanNet.forward and the accuracy counter) and evaluated by
// test.py:7-239.  The model is 5 529 parameters = 22 KB, so weights, momentum and gradient of a client live in shared
// memory for the entire round: one CTA per client runs all T optimiser steps (gather of the batch rows by index,
// feature-trigger stamping + label swap on the first `poison_n` samples, forward with counter-based dropout fused into
// the linear epilogue, softmax cross-entropy + argmax counters, backward, SGD with momentum / weight decay and optional
// FoolsGold gradient sum) without returning to the host.  K clients = K CTAs in one launch.
// K6 / K7 / K8 / K9 / K12 of SURVEY §2.4 for the LOAN family (CUDA-core kernel: N < 16 rules out tensor cores).
#include "common.cuh"

using namespace dba;

namespace {

constexpr int F0 = 91, F1 = 46, F2 = 23, F3 = 9;
constexpr int OFF_W1 = 0, OFF_B1 = OFF_W1 + F1 * F0, OFF_W2 = OFF_B1 + F1, OFF_B2 = OFF_W2 + F2 * F1,
              OFF_W3 = OFF_B2 + F2, OFF_B3 = OFF_W3 + F3 * F2, NPARAM = OFF_B3 + F3;   // 5529
static_assert(NPARAM == 5529, "LoanNet parameter count");
constexpr int MAXB = 64;
constexpr int THREADS = 256;

__device__ __forceinline__ unsigned int hash32(unsigned int a, unsigned int b, unsigned int c, unsigned int d) {
  unsigned int h = a * 0x9E3779B1u ^ (b + 0x7F4A7C15u) * 0x85EBCA77u ^ (c + 0x165667B1u) * 0xC2B2AE3Du ^
                   (d + 0x27D4EB2Fu) * 0x2545F491u;
  h ^= h >> 16; h *= 0x7FEB352Du; h ^= h >> 15; h *= 0x846CA68Bu; h ^= h >> 16;
  return h;
}

struct LoanTrainArgs {
  float* W;                 // [K][ldw] states (parameters in the first 5529 columns), updated in place
  float* gsum;              // [K][ldg] FoolsGold raw-gradient sums (+=) or null
  const float* X;           // [N][91] training features
  const long long* Y;       // [N]
  const long long* idx;     // [K][T][B] row indices, -1 = padding
  const float* lr;          // [K][T]
  const unsigned char* active;  // [K][T]
  const long long* epoch_of;    // [K][T]
  const long long* poison_n;    // [K]
  const long long* trig_row;    // [K]
  const unsigned char* trig_mask;  // [R][91]
  const float* trig_val;           // [R][91]
  float* loss_sum;          // [K][E]  sum of batch-mean losses (Q6)
  int* correct;             // [K][E]
  int* count;               // [K][E]
  int* pcount;              // [K][E]
  unsigned char* nan_flag;  // [K]
  long long ldw, ldg;
  int K, T, B, E, label_swap;
  float mu, wd;
  unsigned int seed;
  int dropout;               // 1 = Dropout(0.5) as in the reference, 0 = off (tests)
};

__global__ void __launch_bounds__(THREADS) loan_train_kernel(const LoanTrainArgs a) {
  extern __shared__ float sm[];
  float* w = sm;                       // [5529] (+pad)
  float* v = w + 5536;                 // momentum
  float* g = v + 5536;                 // gradient
  float* x = g + 5536;                 // [B][92]
  float* h1 = x + MAXB * 92;           // [B][46] post-activation
  float* h2 = h1 + MAXB * F1;          // [B][23]
  float* d3 = h2 + MAXB * F2;          // [B][9]  dlogits
  float* d2 = d3 + MAXB * F3;          // [B][23]
  float* d1 = d2 + MAXB * F2;          // [B][46]
  __shared__ int ylab[MAXB];
  __shared__ float red[32];
  __shared__ int s_cnt[4];

  const int k = blockIdx.x, tid = threadIdx.x;
  float* Wk = a.W + (long long)k * a.ldw;
  for (int i = tid; i < NPARAM; i += THREADS) { w[i] = Wk[i]; v[i] = 0.f; }
  __syncthreads();
  const int pn = (int)a.poison_n[k];
  const float dscale = a.dropout ? 2.f : 1.f;
  const unsigned char* tmask = a.trig_mask + a.trig_row[k] * F0;
  const float* tval = a.trig_val + a.trig_row[k] * F0;

  for (int t = 0; t < a.T; ++t) {
    if (!a.active[(long long)k * a.T + t]) continue;       // uniform per block
    const long long* ids = a.idx + ((long long)k * a.T + t) * a.B;
    // ---- gather + trigger
    for (int i = tid; i < a.B * F0; i += THREADS) {
      const int b = i / F0, f = i - b * F0;
      const long long id = ids[b];
      float val = 0.f;
      if (id >= 0) {
        val = a.X[id * F0 + f];
        if (b < pn && tmask[f]) val = tval[f];
      }
      x[b * 92 + f] = val;
    }
    if (tid < a.B) {
      const long long id = ids[tid];
      ylab[tid] = id < 0 ? -1 : (tid < pn ? a.label_swap : (int)a.Y[id]);
    }
    if (tid == 0) { s_cnt[0] = 0; s_cnt[1] = 0; s_cnt[2] = 0; }
    for (int i = tid; i < NPARAM; i += THREADS) g[i] = 0.f;
    __syncthreads();
    if (tid < a.B && ylab[tid] >= 0) { atomicAdd(&s_cnt[0], 1); if (tid < pn) atomicAdd(&s_cnt[2], 1); }
    __syncthreads();
    const int nvalid = s_cnt[0];
    const float inv_n = 1.f / (float)max(nvalid, 1);
    // ---- layer 1: h1 = relu(drop(W1 x + b1)); dropout keeps with p = 0.5 and scales by 2 (mask stored in sign of h1:
    //      dropped or negative pre-activations give exactly 0, which is also what their gradient needs)
    for (int i = tid; i < a.B * F1; i += THREADS) {
      const int b = i / F1, j = i - b * F1;
      float acc = w[OFF_B1 + j];
      const float* wr = w + OFF_W1 + j * F0;
      const float* xr = x + b * 92;
#pragma unroll 7
      for (int f = 0; f < F0; ++f) acc += wr[f] * xr[f];
      const bool keep = !a.dropout || (hash32(a.seed, (unsigned)k * 65537u + (unsigned)t, (unsigned)b, (unsigned)j) & 1u);
      h1[i] = (keep && acc > 0.f) ? dscale * acc : 0.f;
    }
    __syncthreads();
    for (int i = tid; i < a.B * F2; i += THREADS) {
      const int b = i / F2, j = i - b * F2;
      float acc = w[OFF_B2 + j];
      const float* wr = w + OFF_W2 + j * F1;
      const float* hr = h1 + b * F1;
#pragma unroll 2
      for (int f = 0; f < F1; ++f) acc += wr[f] * hr[f];
      const bool keep = !a.dropout || (hash32(a.seed ^ 0xA511E9B3u, (unsigned)k * 65537u + (unsigned)t, (unsigned)b, (unsigned)j) & 1u);
      h2[i] = (keep && acc > 0.f) ? dscale * acc : 0.f;
    }
    __syncthreads();
    // ---- logits, CE, metrics, dlogits (one thread per sample)
    float my_loss = 0.f;
    int my_ok = 0;
    bool bad = false;
    if (tid < a.B) {
      const int b = tid, lab = ylab[b];
      float lg[F3];
      float mx = -INFINITY;
      int arg = 0;
      for (int j = 0; j < F3; ++j) {
        float acc = w[OFF_B3 + j];
        for (int f = 0; f < F2; ++f) acc += w[OFF_W3 + j * F2 + f] * h2[b * F2 + f];
        lg[j] = acc;
        if (acc > mx) { mx = acc; arg = j; }
      }
      float se = 0.f;
      for (int j = 0; j < F3; ++j) se += expf(lg[j] - mx);
      const float lse = mx + logf(se);
      if (lab >= 0) {
        my_loss = (lse - lg[lab]) * inv_n;
        my_ok = arg == lab;
        bad = !isfinite(lse);
        for (int j = 0; j < F3; ++j) d3[b * F3 + j] = (expf(lg[j] - lse) - (j == lab ? 1.f : 0.f)) * inv_n;
      } else {
        for (int j = 0; j < F3; ++j) d3[b * F3 + j] = 0.f;
      }
    }
    my_loss = block_sum(my_loss, red, false);
    if (tid < a.B && my_ok) atomicAdd(&s_cnt[1], 1);
    if (bad) a.nan_flag[k] = 1;
    __syncthreads();
    if (tid == 0) {
      const int e = (int)a.epoch_of[(long long)k * a.T + t];
      a.loss_sum[k * a.E + e] += my_loss;
      a.correct[k * a.E + e] += s_cnt[1];
      a.count[k * a.E + e] += nvalid;
      a.pcount[k * a.E + e] += s_cnt[2];
    }
    // ---- backward: d2 = (W3^T d3) * [h2 > 0] * 2 ; d1 likewise
    for (int i = tid; i < a.B * F2; i += THREADS) {
      const int b = i / F2, f = i - b * F2;
      float acc = 0.f;
      for (int j = 0; j < F3; ++j) acc += w[OFF_W3 + j * F2 + f] * d3[b * F3 + j];
      d2[i] = h2[i] > 0.f ? dscale * acc : 0.f;
    }
    __syncthreads();
    for (int i = tid; i < a.B * F1; i += THREADS) {
      const int b = i / F1, f = i - b * F1;
      float acc = 0.f;
      for (int j = 0; j < F2; ++j) acc += w[OFF_W2 + j * F1 + f] * d2[b * F2 + j];
      d1[i] = h1[i] > 0.f ? dscale * acc : 0.f;
    }
    __syncthreads();
    // ---- weight gradients: one thread per parameter, reduction over the batch
    for (int i = tid; i < NPARAM; i += THREADS) {
      float acc = 0.f;
      if (i < OFF_B1) {
        const int j = i / F0, f = i - j * F0;
        for (int b = 0; b < a.B; ++b) acc += d1[b * F1 + j] * x[b * 92 + f];
      } else if (i < OFF_W2) {
        const int j = i - OFF_B1;
        for (int b = 0; b < a.B; ++b) acc += d1[b * F1 + j];
      } else if (i < OFF_B2) {
        const int r = i - OFF_W2, j = r / F1, f = r - j * F1;
        for (int b = 0; b < a.B; ++b) acc += d2[b * F2 + j] * h1[b * F1 + f];
      } else if (i < OFF_W3) {
        const int j = i - OFF_B2;
        for (int b = 0; b < a.B; ++b) acc += d2[b * F2 + j];
      } else if (i < OFF_B3) {
        const int r = i - OFF_W3, j = r / F2, f = r - j * F2;
        for (int b = 0; b < a.B; ++b) acc += d3[b * F3 + j] * h2[b * F2 + f];
      } else {
        const int j = i - OFF_B3;
        for (int b = 0; b < a.B; ++b) acc += d3[b * F3 + j];
      }
      g[i] = acc;
    }
    __syncthreads();
    // ---- SGD (momentum, weight decay) + FoolsGold raw-gradient sum
    const float lr = a.lr[(long long)k * a.T + t];
    for (int i = tid; i < NPARAM; i += THREADS) {
      const float gi = g[i];
      if (a.gsum) a.gsum[(long long)k * a.ldg + i] += gi;
      const float vi = a.mu * v[i] + (gi + a.wd * w[i]);
      v[i] = vi;
      w[i] -= lr * vi;
    }
    __syncthreads();
  }
  for (int i = tid; i < NPARAM; i += THREADS) Wk[i] = w[i];
}

// Evaluation: M models x rows; grid (row tiles, M); each thread one row, weights of the model in shared memory.
struct LoanEvalArgs {
  const float* W; long long ldw;
  const float* X; const long long* Y; const long long* idx; long long n;
  const unsigned char* trig_mask; const float* trig_val;   // one row each (or null = clean)
  double* loss; long long* correct;                        // [M]
  int label_swap;
};
__global__ void __launch_bounds__(256) loan_eval_kernel(const LoanEvalArgs a) {
  __shared__ float w[NPARAM];
  __shared__ double redd[32];
  __shared__ int s_ok;
  const int m = blockIdx.y;
  for (int i = threadIdx.x; i < NPARAM; i += blockDim.x) w[i] = a.W[(long long)m * a.ldw + i];
  if (threadIdx.x == 0) s_ok = 0;
  __syncthreads();
  double my_loss = 0.0;
  int ok = 0;
  for (long long r = blockIdx.x * (long long)blockDim.x + threadIdx.x; r < a.n; r += (long long)gridDim.x * blockDim.x) {
    const long long id = a.idx[r];
    float xv[F0];
    for (int f = 0; f < F0; ++f) {
      float val = a.X[id * F0 + f];
      if (a.trig_mask && a.trig_mask[f]) val = a.trig_val[f];
      xv[f] = val;
    }
    const int lab = a.trig_mask ? a.label_swap : (int)a.Y[id];
    float h1v[F1], h2v[F2];
    for (int j = 0; j < F1; ++j) {
      float acc = w[OFF_B1 + j];
      for (int f = 0; f < F0; ++f) acc += w[OFF_W1 + j * F0 + f] * xv[f];
      h1v[j] = relu_nan(acc);
    }
    for (int j = 0; j < F2; ++j) {
      float acc = w[OFF_B2 + j];
      for (int f = 0; f < F1; ++f) acc += w[OFF_W2 + j * F1 + f] * h1v[f];
      h2v[j] = relu_nan(acc);
    }
    float lg[F3], mx = -INFINITY;
    int arg = 0;
    for (int j = 0; j < F3; ++j) {
      float acc = w[OFF_B3 + j];
      for (int f = 0; f < F2; ++f) acc += w[OFF_W3 + j * F2 + f] * h2v[f];
      lg[j] = acc;
      if (acc > mx) { mx = acc; arg = j; }
    }
    float se = 0.f;
    for (int j = 0; j < F3; ++j) se += expf(lg[j] - mx);
    my_loss += (double)(mx + logf(se) - lg[lab]);
    ok += arg == lab;
  }
  my_loss = block_sum(my_loss, redd, false);
  if (ok) atomicAdd(&s_ok, ok);
  __syncthreads();
  if (threadIdx.x == 0) {
    atomicAdd(&a.loss[m], my_loss);
    atomicAdd((unsigned long long*)&a.correct[m], (unsigned long long)s_ok);
  }
}

}  // namespace

extern "C" {

int dba_loan_train(float* W, long long ldw, float* gsum, long long ldg, const float* X, const long long* Y,
                   const long long* idx, const float* lr, const unsigned char* active, const long long* epoch_of,
                   const long long* poison_n, const long long* trig_row, const unsigned char* trig_mask,
                   const float* trig_val, float* loss_sum, int* correct, int* count, int* pcount,
                   unsigned char* nan_flag, int K, int T, int B, int E, int label_swap, float mu, float wd,
                   unsigned int seed, int dropout, void* stream) {
  DBA_REQUIRE(K > 0 && T >= 0 && B > 0 && B <= MAXB, "dba_loan_train: K=%d T=%d B=%d (B <= %d)", K, T, B, MAXB);
  if (T == 0) return DBA_OK;
  LoanTrainArgs a{W, gsum, X, Y, idx, lr, active, epoch_of, poison_n, trig_row, trig_mask, trig_val, loss_sum, correct,
                  count, pcount, nan_flag, ldw, ldg, K, T, B, E, label_swap, mu, wd, seed, dropout};
  const size_t smem = (3 * 5536 + MAXB * (92 + F1 + F2 + F3 + F2 + F1)) * sizeof(float);
  static bool attr_set[64] = {false};
  int dev = 0;
  DBA_CHECK_CUDA(cudaGetDevice(&dev));
  if (dev < 64 && !attr_set[dev]) {
    DBA_CHECK_CUDA(cudaFuncSetAttribute(loan_train_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    attr_set[dev] = true;
  }
  loan_train_kernel<<<K, THREADS, smem, (cudaStream_t)stream>>>(a);
  DBA_COUNT_LAUNCH(1);
  DBA_CHECK_LAUNCH();
  return DBA_OK;
}

// loss[m] / correct[m] are accumulated (+=): zero them first.  trig_mask/trig_val: one 91-entry row or null.
int dba_loan_eval(const float* W, long long ldw, int M, const float* X, const long long* Y, const long long* idx,
                  long long n, const unsigned char* trig_mask, const float* trig_val, int label_swap, double* loss,
                  long long* correct, void* stream) {
  DBA_REQUIRE(M > 0 && n >= 0, "dba_loan_eval: M=%d n=%lld", M, n);
  if (n == 0) return DBA_OK;
  LoanEvalArgs a{W, ldw, X, Y, idx, n, trig_mask, trig_val, loss, correct, label_swap};
  long long gx = (n + 255) / 256;
  if (gx > 4 * kNumSMs) gx = 4 * kNumSMs;
  dim3 g((unsigned)gx, M);
  loan_eval_kernel<<<g, 256, 0, (cudaStream_t)stream>>>(a);
  DBA_COUNT_LAUNCH(1);
  DBA_CHECK_LAUNCH();
  return DBA_OK;
}

}  // extern "C"
