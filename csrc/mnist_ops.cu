// Elementwise kernels of the BatchNorm-free MnistNet path (reference models/MnistNet.py:17-30: relu(conv) -> 2x2
// max-pool, twice, flatten, relu(fc1), fc2) on planar-8 activations for K client replicas at once.
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

__device__ __forceinline__ void unpack8(const uint4 u, float (&f)[8]) {
  f[0] = bf16lo(u.x); f[1] = bf16hi(u.x); f[2] = bf16lo(u.y); f[3] = bf16hi(u.y);
  f[4] = bf16lo(u.z); f[5] = bf16hi(u.z); f[6] = bf16lo(u.w); f[7] = bf16hi(u.w);
}
__device__ __forceinline__ uint4 pack8(const float (&f)[8]) {
  uint4 p;
  p.x = pack_bf16x2(f[0], f[1]); p.y = pack_bf16x2(f[2], f[3]);
  p.z = pack_bf16x2(f[4], f[5]); p.w = pack_bf16x2(f[6], f[7]);
  return p;
}

struct PoolArgs2 {
  const __nv_bfloat16* x;      // [K][C/8][Ri][8] pooled activation (post-ReLU), grid (B, Hi, Wi)
  __nv_bfloat16* y;            // fwd out: planar (B, Ho, Wo) [K][C/8][Ro][8]  or flattened [K][Ho*Wo*C/8][R1][8]
  const __nv_bfloat16* g;      // bwd in: gradient of y (same form as y)
  __nv_bfloat16* dz;           // bwd out: [K][C/8][Ri][8]
  float* G;                    // [K][ldg] gradient slab (bias gradient at b_off) or null
  long long Ri, Ro, ldg;
  int b_off, c_real;
  int K, C, B, Hi, Wi, off, Ho, Wo, flatten;
};

__device__ __forceinline__ long long out_index(const PoolArgs2& a, int k, int c8, int b, int i, int j) {
  if (a.flatten) {      // (B, 1, 1) grid: Wp = 2, S = 4, lead = 3; channel chunk = pos * C/8 + c8
    const int pos = i * a.Wo + j;
    const long long plane = (long long)k * (a.Ho * a.Wo * (a.C >> 3)) + pos * (a.C >> 3) + c8;
    return (plane * a.Ro + 3 + 4LL * b) * 8;
  }
  const int Wpo = a.Wo + 1, So = (a.Ho + 1) * Wpo, leado = Wpo + 1;
  return ((((long long)k * (a.C >> 3) + c8) * a.Ro) + leado + (long long)b * So + i * Wpo + j) * 8;
}

__global__ void __launch_bounds__(256) maxpool2_fwd_kernel(const PoolArgs2 a) {
  const int c8 = blockIdx.y, k = blockIdx.z;
  const int Wpi = a.Wi + 1, Si = (a.Hi + 1) * Wpi, leadi = Wpi + 1;
  const int total = a.B * a.Ho * a.Wo;
  const __nv_bfloat16* xp = a.x + (((long long)k * (a.C >> 3) + c8) * a.Ri) * 8;
  for (int q = blockIdx.x * blockDim.x + threadIdx.x; q < total; q += gridDim.x * blockDim.x) {
    const int j = q % a.Wo, i = (q / a.Wo) % a.Ho, b = q / (a.Wo * a.Ho);
    const long long r0 = (long long)leadi + (long long)b * Si + (a.off + 2 * i) * Wpi + (a.off + 2 * j);
    float m[8], v[8];
    unpack8(*reinterpret_cast<const uint4*>(xp + r0 * 8), m);
    unpack8(*reinterpret_cast<const uint4*>(xp + (r0 + 1) * 8), v);
#pragma unroll
    for (int e = 0; e < 8; ++e) m[e] = fmaxf(m[e], v[e]);
    unpack8(*reinterpret_cast<const uint4*>(xp + (r0 + Wpi) * 8), v);
#pragma unroll
    for (int e = 0; e < 8; ++e) m[e] = fmaxf(m[e], v[e]);
    unpack8(*reinterpret_cast<const uint4*>(xp + (r0 + Wpi + 1) * 8), v);
#pragma unroll
    for (int e = 0; e < 8; ++e) m[e] = fmaxf(m[e], v[e]);
    *reinterpret_cast<uint4*>(a.y + out_index(a, k, c8, b, i, j)) = pack8(m);
  }
}

// block-level sum of 8 per-thread channel values -> one atomic per channel per block
__device__ __forceinline__ void block_bias_add(float (&s)[8], float* dst, int c0, int c_real) {
  __shared__ float red[8][8];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll
  for (int e = 0; e < 8; ++e) s[e] = warp_sum(s[e]);
  if (lane == 0) {
#pragma unroll
    for (int e = 0; e < 8; ++e) red[warp][e] = s[e];
  }
  __syncthreads();
  if (threadIdx.x < 8) {
    float t = 0.f;
    for (int w = 0; w < (int)(blockDim.x >> 5); ++w) t += red[w][threadIdx.x];
    if (c0 + (int)threadIdx.x < c_real && t != 0.f) atomicAdd(dst + c0 + threadIdx.x, t);
  }
}

__global__ void __launch_bounds__(256) maxpool2_bwd_relu_bias_kernel(const PoolArgs2 a) {
  const int c8 = blockIdx.y, k = blockIdx.z;
  const int Wpi = a.Wi + 1, Si = (a.Hi + 1) * Wpi, leadi = Wpi + 1;
  const int total = a.B * a.Hi * a.Wi;
  const long long pbase = (((long long)k * (a.C >> 3) + c8) * a.Ri) * 8;
  const __nv_bfloat16* xp = a.x + pbase;
  __nv_bfloat16* dp = a.dz + pbase;
  float bsum[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) bsum[e] = 0.f;
  for (int q = blockIdx.x * blockDim.x + threadIdx.x; q < total; q += gridDim.x * blockDim.x) {
    const int w = q % a.Wi, h = (q / a.Wi) % a.Hi, b = q / (a.Wi * a.Hi);
    const long long r = (long long)leadi + (long long)b * Si + h * Wpi + w;
    float d[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) d[e] = 0.f;
    const int hh = h - a.off, ww = w - a.off;
    if (hh >= 0 && ww >= 0 && hh < 2 * a.Ho && ww < 2 * a.Wo) {
      const int i = hh >> 1, j = ww >> 1, di = hh & 1, dj = ww & 1;
      float g[8], me[8], o[8];
      unpack8(*reinterpret_cast<const uint4*>(a.g + out_index(a, k, c8, b, i, j)), g);
      unpack8(*reinterpret_cast<const uint4*>(xp + r * 8), me);
      bool win[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) win[e] = me[e] > 0.f;            // ReLU mask: a zero never carries gradient
      const long long r0 = r - di * Wpi - dj;
#pragma unroll
      for (int t = 0; t < 4; ++t) {
        if (t == di * 2 + dj) continue;
        unpack8(*reinterpret_cast<const uint4*>(xp + (r0 + (t >> 1) * Wpi + (t & 1)) * 8), o);
        const bool before = t < di * 2 + dj;
#pragma unroll
        for (int e = 0; e < 8; ++e) win[e] = win[e] && (before ? o[e] < me[e] : o[e] <= me[e]);
      }
#pragma unroll
      for (int e = 0; e < 8; ++e) { d[e] = win[e] ? g[e] : 0.f; bsum[e] += d[e]; }
    }
    *reinterpret_cast<uint4*>(dp + r * 8) = pack8(d);
  }
  if (a.G) block_bias_add(bsum, a.G + (long long)k * a.ldg + a.b_off, c8 * 8, a.c_real);
}

struct ReluArgs {
  const __nv_bfloat16* g;
  const __nv_bfloat16* act;
  __nv_bfloat16* dz;
  float* G;
  long long R, ldg;
  int b_off, c_real, K, C, B, H, W;
};
__global__ void __launch_bounds__(256) relu_bwd_bias_kernel(const ReluArgs a) {
  const int c8 = blockIdx.y, k = blockIdx.z;
  const int Wp = a.W + 1, S = (a.H + 1) * Wp, lead = Wp + 1;
  const int total = a.B * a.H * a.W;
  const long long pbase = (((long long)k * (a.C >> 3) + c8) * a.R) * 8;
  float bsum[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) bsum[e] = 0.f;
  for (int q = blockIdx.x * blockDim.x + threadIdx.x; q < total; q += gridDim.x * blockDim.x) {
    const int w = q % a.W, h = (q / a.W) % a.H, b = q / (a.W * a.H);
    const long long off = pbase + ((long long)lead + (long long)b * S + h * Wp + w) * 8;
    float g[8], x[8];
    unpack8(*reinterpret_cast<const uint4*>(a.g + off), g);
    unpack8(*reinterpret_cast<const uint4*>(a.act + off), x);
#pragma unroll
    for (int e = 0; e < 8; ++e) { g[e] = x[e] > 0.f ? g[e] : 0.f; bsum[e] += g[e]; }
    *reinterpret_cast<uint4*>(a.dz + off) = pack8(g);
  }
  if (a.G) block_bias_add(bsum, a.G + (long long)k * a.ldg + a.b_off, c8 * 8, a.c_real);
}

inline int row_blocks(int total) {
  int rb = (total + 1023) / 1024;
  return rb < 1 ? 1 : (rb > 64 ? 64 : rb);
}

}  // namespace

extern "C" {

// x: [K][C/8][Ri][8] on the (B, Hi, Wi) grid, Ri = dba_conv_rows(B, Hi, Wi); pooled window starts at (off, off) and
// covers 2*Ho x 2*Wo pixels.  flatten = 0: y is [K][C/8][Ro][8] on (B, Ho, Wo), Ro = dba_conv_rows(B, Ho, Wo);
// flatten = 1: y is [K][Ho*Wo*C/8][R1][8] on (B, 1, 1) with channel n = (i*Wo + j)*C + c, Ro = dba_conv_rows(B, 1, 1).
int dba_maxpool2_fwd(const void* x, void* y, int K, int C, int B, int Hi, int Wi, int off, int Ho, int Wo, int flatten,
                     long long Ri, long long Ro, void* stream) {
  DBA_REQUIRE(g_dba_act_mode == 0, "MnistNet pooling kernels run in bf16 mode only (dba_set_act_dtype(0))");
  DBA_REQUIRE(K > 0 && C % 8 == 0 && off >= 0 && off + 2 * Ho <= Hi && off + 2 * Wo <= Wi,
              "dba_maxpool2_fwd: K=%d C=%d window off=%d %dx%d on %dx%d", K, C, off, Ho, Wo, Hi, Wi);
  PoolArgs2 a{};
  a.x = (const __nv_bfloat16*)x; a.y = (__nv_bfloat16*)y; a.Ri = Ri; a.Ro = Ro;
  a.K = K; a.C = C; a.B = B; a.Hi = Hi; a.Wi = Wi; a.off = off; a.Ho = Ho; a.Wo = Wo; a.flatten = flatten;
  maxpool2_fwd_kernel<<<dim3(row_blocks(B * Ho * Wo), C / 8, K), 256, 0, (cudaStream_t)stream>>>(a);
  DBA_COUNT_LAUNCH(1);
  DBA_CHECK_LAUNCH();
  return DBA_OK;
}

// g has the form of dba_maxpool2_fwd's y; dz gets the gradient with respect to the PRE-ReLU convolution output on the
// input grid (zero outside the window); G[k*ldg + b_off + c] += sum of dz over pixels and samples for c < c_real.
int dba_maxpool2_bwd_relu_bias(const void* g, const void* x, void* dz, float* G, long long ldg, int b_off, int c_real,
                               int K, int C, int B, int Hi, int Wi, int off, int Ho, int Wo, int flatten, long long Ri,
                               long long Ro, void* stream) {
  DBA_REQUIRE(g_dba_act_mode == 0, "MnistNet pooling kernels run in bf16 mode only (dba_set_act_dtype(0))");
  DBA_REQUIRE(K > 0 && C % 8 == 0 && off >= 0 && off + 2 * Ho <= Hi && off + 2 * Wo <= Wi,
              "dba_maxpool2_bwd: K=%d C=%d window off=%d %dx%d on %dx%d", K, C, off, Ho, Wo, Hi, Wi);
  PoolArgs2 a{};
  a.x = (const __nv_bfloat16*)x; a.g = (const __nv_bfloat16*)g; a.dz = (__nv_bfloat16*)dz; a.G = G; a.ldg = ldg;
  a.b_off = b_off; a.c_real = c_real; a.Ri = Ri; a.Ro = Ro;
  a.K = K; a.C = C; a.B = B; a.Hi = Hi; a.Wi = Wi; a.off = off; a.Ho = Ho; a.Wo = Wo; a.flatten = flatten;
  maxpool2_bwd_relu_bias_kernel<<<dim3(row_blocks(B * Hi * Wi), C / 8, K), 256, 0, (cudaStream_t)stream>>>(a);
  DBA_COUNT_LAUNCH(1);
  DBA_CHECK_LAUNCH();
  return DBA_OK;
}

// dz = g * (act > 0) on the (B, H, W) grid, G[k*ldg + b_off + c] += sum dz (c < c_real).
int dba_relu_bwd_bias(const void* g, const void* act, void* dz, float* G, long long ldg, int b_off, int c_real, int K,
                      int C, int B, int H, int W, long long R, void* stream) {
  DBA_REQUIRE(g_dba_act_mode == 0, "MnistNet pooling kernels run in bf16 mode only (dba_set_act_dtype(0))");
  DBA_REQUIRE(K > 0 && C % 8 == 0 && B > 0, "dba_relu_bwd_bias: K=%d C=%d B=%d", K, C, B);
  ReluArgs a{(const __nv_bfloat16*)g, (const __nv_bfloat16*)act, (__nv_bfloat16*)dz, G, R, ldg, b_off, c_real, K, C, B,
             H, W};
  relu_bwd_bias_kernel<<<dim3(row_blocks(B * H * W), C / 8, K), 256, 0, (cudaStream_t)stream>>>(a);
  DBA_COUNT_LAUNCH(1);
  DBA_CHECK_LAUNCH();
  return DBA_OK;
}

}  // extern "C"
