// Elementwise kernels of the BatchNorm-free MnistNet path (reference models/MnistNet.py:17-30: relu(conv) -> 2x2
// max-pool, twice, flatten, relu(fc1), fc2) on planar activations (bf16 x 8 or fp32 x 4 per 16-byte chunk) for K
// client replicas at once.
//
//   maxpool2_fwd            2x2 stride-2 max-pool of the [off, off + 2*Ho) window of the input grid (a 'valid'
//                           convolution leaves its output on the centre of the input grid); writes either a planar
//                           (B, Ho, Wo) tensor or the flattened form: channel n = pos * C + c of a (B, 1, 1) tensor, the
//                           operand of the following linear layer run as a 1x1 tensor-core convolution.
//   maxpool2_bwd_relu_bias  routes the pooled gradient to the arg-max (first maximum in scan order, as torch does),
//                           multiplies by the ReLU mask of the pooled activation, writes dz on the input grid (zero
//                           outside the window) and adds the per-channel sum of dz (the convolution's bias gradient)
//                           into the flat gradient slab.
//   relu_bwd_bias           dz = g * (a > 0) on any grid, plus bias gradient (fc1).
#include "common.cuh"

using namespace dba;

namespace {

// All tensors are addressed in 16-byte rows of V channels (V = 8: bf16 mode, V = 4: fp32 / tf32 mode, see common.cuh).
// In tf32 mode every tensor a weight-gradient kernel reads (pooled activations, dz) also gets its bf16 hi/lo split copy
// (store_split).
struct PoolArgs2 {
  const uint4* x;              // [K][C/V][Ri] pooled activation (post-ReLU), grid (B, Hi, Wi)
  uint4* y;                    // fwd out: planar (B, Ho, Wo) [K][C/V][Ro]  or flattened [K][Ho*Wo*C/V][R1]
  const uint4* g;              // bwd in: gradient of y (same form as y)
  uint4* dz;                   // bwd out: [K][C/V][Ri]
  uint2* y_split;              // tf32 mode, optional: split copy of y
  uint2* dz_split;             // tf32 mode, optional: split copy of dz
  float* G;                    // [K][ldg] gradient slab (bias gradient at b_off) or null
  long long Ri, Ro, ldg;
  int b_off, c_real;
  int K, C, B, Hi, Wi, off, Ho, Wo, flatten;
};

// row (16-byte units, relative to client k's block) and channel-chunk plane of output pixel (b, i, j), chunk cv
template <int V>
__device__ __forceinline__ void out_pos(const PoolArgs2& a, int cv, int b, int i, int j, long long& plane, long long& row) {
  if (a.flatten) {      // (B, 1, 1) grid: Wp = 2, S = 4, lead = 3; channel chunk = pos * C/V + cv
    plane = (long long)(i * a.Wo + j) * (a.C / V) + cv;
    row = 3 + 4LL * b;
  } else {
    const int Wpo = a.Wo + 1, So = (a.Ho + 1) * Wpo, leado = Wpo + 1;
    plane = cv;
    row = leado + (long long)b * So + i * Wpo + j;
  }
}
template <int V> __device__ __forceinline__ long long out_planes(const PoolArgs2& a) {
  return a.flatten ? (long long)a.Ho * a.Wo * (a.C / V) : (long long)(a.C / V);
}

template <int V>
__global__ void __launch_bounds__(256) maxpool2_fwd_kernel(const PoolArgs2 a) {
  const int cv = blockIdx.y, k = blockIdx.z;
  const int Wpi = a.Wi + 1, Si = (a.Hi + 1) * Wpi, leadi = Wpi + 1;
  const int total = a.B * a.Ho * a.Wo;
  const uint4* xp = a.x + ((long long)k * (a.C / V) + cv) * a.Ri;
  for (int q = blockIdx.x * blockDim.x + threadIdx.x; q < total; q += gridDim.x * blockDim.x) {
    const int j = q % a.Wo, i = (q / a.Wo) % a.Ho, b = q / (a.Wo * a.Ho);
    const long long r0 = (long long)leadi + (long long)b * Si + (a.off + 2 * i) * Wpi + (a.off + 2 * j);
    float m[V], v[V];
    chunk_unpack<V>(xp[r0], m);
    chunk_unpack<V>(xp[r0 + 1], v);
#pragma unroll
    for (int e = 0; e < V; ++e) m[e] = fmaxf(m[e], v[e]);
    chunk_unpack<V>(xp[r0 + Wpi], v);
#pragma unroll
    for (int e = 0; e < V; ++e) m[e] = fmaxf(m[e], v[e]);
    chunk_unpack<V>(xp[r0 + Wpi + 1], v);
#pragma unroll
    for (int e = 0; e < V; ++e) m[e] = fmaxf(m[e], v[e]);
    long long plane, row;
    out_pos<V>(a, cv, b, i, j, plane, row);
    a.y[((long long)k * out_planes<V>(a) + plane) * a.Ro + row] = chunk_pack<V>(m);   // maxima of operand-grid values
    if constexpr (V == 4) {
      if (a.y_split) store_split(a.y_split, k, out_planes<V>(a) >> 1, a.Ro, (int)plane, row, m);
    }
  }
}

// block-level sum of V per-thread channel values -> one atomic per channel per block
template <int V>
__device__ __forceinline__ void block_bias_add(float (&s)[V], float* dst, int c0, int c_real) {
  __shared__ float red[8][V];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll
  for (int e = 0; e < V; ++e) s[e] = warp_sum(s[e]);
  if (lane == 0) {
#pragma unroll
    for (int e = 0; e < V; ++e) red[warp][e] = s[e];
  }
  __syncthreads();
  if (threadIdx.x < V) {
    float t = 0.f;
    for (int w = 0; w < (int)(blockDim.x >> 5); ++w) t += red[w][threadIdx.x];
    if (c0 + (int)threadIdx.x < c_real && t != 0.f) atomicAdd(dst + c0 + threadIdx.x, t);
  }
}

template <int V>
__global__ void __launch_bounds__(256) maxpool2_bwd_relu_bias_kernel(const PoolArgs2 a) {
  const int cv = blockIdx.y, k = blockIdx.z;
  const int Wpi = a.Wi + 1, Si = (a.Hi + 1) * Wpi, leadi = Wpi + 1;
  const int total = a.B * a.Hi * a.Wi;
  const long long pbase = ((long long)k * (a.C / V) + cv) * a.Ri;
  const uint4* xp = a.x + pbase;
  uint4* dp = a.dz + pbase;
  float bsum[V];
#pragma unroll
  for (int e = 0; e < V; ++e) bsum[e] = 0.f;
  for (int q = blockIdx.x * blockDim.x + threadIdx.x; q < total; q += gridDim.x * blockDim.x) {
    const int w = q % a.Wi, h = (q / a.Wi) % a.Hi, b = q / (a.Wi * a.Hi);
    const long long r = (long long)leadi + (long long)b * Si + h * Wpi + w;
    float d[V];
#pragma unroll
    for (int e = 0; e < V; ++e) d[e] = 0.f;
    const int hh = h - a.off, ww = w - a.off;
    if (hh >= 0 && ww >= 0 && hh < 2 * a.Ho && ww < 2 * a.Wo) {
      const int i = hh >> 1, j = ww >> 1, di = hh & 1, dj = ww & 1;
      float g[V], me[V], o[V];
      long long plane, row;
      out_pos<V>(a, cv, b, i, j, plane, row);
      chunk_unpack<V>(a.g[((long long)k * out_planes<V>(a) + plane) * a.Ro + row], g);
      chunk_unpack<V>(xp[r], me);
      bool win[V];
#pragma unroll
      for (int e = 0; e < V; ++e) win[e] = me[e] > 0.f;            // ReLU mask: a zero never carries gradient
      const long long r0 = r - di * Wpi - dj;
#pragma unroll
      for (int t = 0; t < 4; ++t) {
        if (t == di * 2 + dj) continue;
        chunk_unpack<V>(xp[r0 + (t >> 1) * Wpi + (t & 1)], o);
        const bool before = t < di * 2 + dj;
#pragma unroll
        for (int e = 0; e < V; ++e) win[e] = win[e] && (before ? o[e] < me[e] : o[e] <= me[e]);
      }
#pragma unroll
      for (int e = 0; e < V; ++e) { d[e] = win[e] ? g[e] : 0.f; bsum[e] += d[e]; }
    }
    dp[r] = chunk_pack_operand<V>(d);                // dz is the input-gradient convolution's operand
    if constexpr (V == 4) {
      if (a.dz_split) store_split(a.dz_split, k, a.C >> 3, a.Ri, cv, r, d);
    }
  }
  if (a.G) block_bias_add<V>(bsum, a.G + (long long)k * a.ldg + a.b_off, cv * V, a.c_real);
}

struct ReluArgs {
  const uint4* g;
  const uint4* act;
  uint4* dz;
  uint2* dz_split;
  float* G;
  long long R, ldg;
  int b_off, c_real, K, C, B, H, W;
};
template <int V>
__global__ void __launch_bounds__(256) relu_bwd_bias_kernel(const ReluArgs a) {
  const int cv = blockIdx.y, k = blockIdx.z;
  const int Wp = a.W + 1, S = (a.H + 1) * Wp, lead = Wp + 1;
  const int total = a.B * a.H * a.W;
  const long long pbase = ((long long)k * (a.C / V) + cv) * a.R;
  float bsum[V];
#pragma unroll
  for (int e = 0; e < V; ++e) bsum[e] = 0.f;
  for (int q = blockIdx.x * blockDim.x + threadIdx.x; q < total; q += gridDim.x * blockDim.x) {
    const int w = q % a.W, h = (q / a.W) % a.H, b = q / (a.W * a.H);
    const long long row = (long long)lead + (long long)b * S + h * Wp + w;
    float g[V], x[V];
    chunk_unpack<V>(a.g[pbase + row], g);
    chunk_unpack<V>(a.act[pbase + row], x);
#pragma unroll
    for (int e = 0; e < V; ++e) { g[e] = x[e] > 0.f ? g[e] : 0.f; bsum[e] += g[e]; }
    a.dz[pbase + row] = chunk_pack_operand<V>(g);
    if constexpr (V == 4) {
      if (a.dz_split) store_split(a.dz_split, k, a.C >> 3, a.R, cv, row, g);
    }
  }
  if (a.G) block_bias_add<V>(bsum, a.G + (long long)k * a.ldg + a.b_off, cv * V, a.c_real);
}

inline int row_blocks(int total) {
  int rb = (total + 1023) / 1024;
  return rb < 1 ? 1 : (rb > 64 ? 64 : rb);
}
inline int act_v() { return g_dba_act_mode == 1 ? 4 : 8; }

}  // namespace

extern "C" {

// x: [K][C/V][Ri] 16-byte rows on the (B, Hi, Wi) grid, Ri = dba_conv_rows(B, Hi, Wi); pooled window starts at
// (off, off) and covers 2*Ho x 2*Wo pixels.  flatten = 0: y is [K][C/V][Ro] on (B, Ho, Wo), Ro = dba_conv_rows(B, Ho,
// Wo); flatten = 1: y is [K][Ho*Wo*C/V][R1] on (B, 1, 1) with channel n = (i*Wo + j)*C + c, Ro = dba_conv_rows(B, 1, 1).
// y_split (nullable, tf32 mode only): bf16 hi/lo split copy of y for the weight-gradient kernel.
int dba_maxpool2_fwd_split(const void* x, void* y, void* y_split, int K, int C, int B, int Hi, int Wi, int off, int Ho,
                           int Wo, int flatten, long long Ri, long long Ro, void* stream) {
  DBA_REQUIRE(K > 0 && C % 8 == 0 && off >= 0 && off + 2 * Ho <= Hi && off + 2 * Wo <= Wi,
              "dba_maxpool2_fwd: K=%d C=%d window off=%d %dx%d on %dx%d", K, C, off, Ho, Wo, Hi, Wi);
  DBA_REQUIRE(!y_split || g_dba_act_mode == 1, "dba_maxpool2_fwd: split copies exist in tf32 mode only");
  PoolArgs2 a{};
  a.x = (const uint4*)x; a.y = (uint4*)y; a.y_split = (uint2*)y_split; a.Ri = Ri; a.Ro = Ro;
  a.K = K; a.C = C; a.B = B; a.Hi = Hi; a.Wi = Wi; a.off = off; a.Ho = Ho; a.Wo = Wo; a.flatten = flatten;
  const dim3 grid(row_blocks(B * Ho * Wo), C / act_v(), K);
  if (act_v() == 4) maxpool2_fwd_kernel<4><<<grid, 256, 0, (cudaStream_t)stream>>>(a);
  else maxpool2_fwd_kernel<8><<<grid, 256, 0, (cudaStream_t)stream>>>(a);
  DBA_COUNT_LAUNCH(1);
  DBA_CHECK_LAUNCH();
  return DBA_OK;
}
int dba_maxpool2_fwd(const void* x, void* y, int K, int C, int B, int Hi, int Wi, int off, int Ho, int Wo, int flatten,
                     long long Ri, long long Ro, void* stream) {
  return dba_maxpool2_fwd_split(x, y, nullptr, K, C, B, Hi, Wi, off, Ho, Wo, flatten, Ri, Ro, stream);
}

// g has the form of dba_maxpool2_fwd's y; dz gets the gradient with respect to the PRE-ReLU convolution output on the
// input grid (zero outside the window); G[k*ldg + b_off + c] += sum of dz over pixels and samples for c < c_real.
// dz_split (nullable, tf32 mode only): split copy of dz.
int dba_maxpool2_bwd_relu_bias_split(const void* g, const void* x, void* dz, void* dz_split, float* G, long long ldg,
                                     int b_off, int c_real, int K, int C, int B, int Hi, int Wi, int off, int Ho, int Wo,
                                     int flatten, long long Ri, long long Ro, void* stream) {
  DBA_REQUIRE(K > 0 && C % 8 == 0 && off >= 0 && off + 2 * Ho <= Hi && off + 2 * Wo <= Wi,
              "dba_maxpool2_bwd: K=%d C=%d window off=%d %dx%d on %dx%d", K, C, off, Ho, Wo, Hi, Wi);
  DBA_REQUIRE(!dz_split || g_dba_act_mode == 1, "dba_maxpool2_bwd: split copies exist in tf32 mode only");
  PoolArgs2 a{};
  a.x = (const uint4*)x; a.g = (const uint4*)g; a.dz = (uint4*)dz; a.dz_split = (uint2*)dz_split; a.G = G; a.ldg = ldg;
  a.b_off = b_off; a.c_real = c_real; a.Ri = Ri; a.Ro = Ro;
  a.K = K; a.C = C; a.B = B; a.Hi = Hi; a.Wi = Wi; a.off = off; a.Ho = Ho; a.Wo = Wo; a.flatten = flatten;
  const dim3 grid(row_blocks(B * Hi * Wi), C / act_v(), K);
  if (act_v() == 4) maxpool2_bwd_relu_bias_kernel<4><<<grid, 256, 0, (cudaStream_t)stream>>>(a);
  else maxpool2_bwd_relu_bias_kernel<8><<<grid, 256, 0, (cudaStream_t)stream>>>(a);
  DBA_COUNT_LAUNCH(1);
  DBA_CHECK_LAUNCH();
  return DBA_OK;
}
int dba_maxpool2_bwd_relu_bias(const void* g, const void* x, void* dz, float* G, long long ldg, int b_off, int c_real,
                               int K, int C, int B, int Hi, int Wi, int off, int Ho, int Wo, int flatten, long long Ri,
                               long long Ro, void* stream) {
  return dba_maxpool2_bwd_relu_bias_split(g, x, dz, nullptr, G, ldg, b_off, c_real, K, C, B, Hi, Wi, off, Ho, Wo, flatten,
                                          Ri, Ro, stream);
}

// dz = g * (act > 0) on the (B, H, W) grid, G[k*ldg + b_off + c] += sum dz (c < c_real).
int dba_relu_bwd_bias_split(const void* g, const void* act, void* dz, void* dz_split, float* G, long long ldg, int b_off,
                            int c_real, int K, int C, int B, int H, int W, long long R, void* stream) {
  DBA_REQUIRE(K > 0 && C % 8 == 0 && B > 0, "dba_relu_bwd_bias: K=%d C=%d B=%d", K, C, B);
  DBA_REQUIRE(!dz_split || g_dba_act_mode == 1, "dba_relu_bwd_bias: split copies exist in tf32 mode only");
  ReluArgs a{(const uint4*)g, (const uint4*)act, (uint4*)dz, (uint2*)dz_split, G, R, ldg, b_off, c_real, K, C, B, H, W};
  const dim3 grid(row_blocks(B * H * W), C / act_v(), K);
  if (act_v() == 4) relu_bwd_bias_kernel<4><<<grid, 256, 0, (cudaStream_t)stream>>>(a);
  else relu_bwd_bias_kernel<8><<<grid, 256, 0, (cudaStream_t)stream>>>(a);
  DBA_COUNT_LAUNCH(1);
  DBA_CHECK_LAUNCH();
  return DBA_OK;
}
int dba_relu_bwd_bias(const void* g, const void* act, void* dz, float* G, long long ldg, int b_off, int c_real, int K,
                      int C, int B, int H, int W, long long R, void* stream) {
  return dba_relu_bwd_bias_split(g, act, dz, nullptr, G, ldg, b_off, c_real, K, C, B, H, W, R, stream);
}

}  // extern "C"
