// Training-mode BatchNorm on planar-8 activations for K client replicas at once (K3/K4 in SURVEY §2.4; reference
// nn.BatchNorm2d + F.relu + residual add in models/resnet_cifar.py:20-35).
//
//   forward : the per-channel sum / sum-of-squares come out of the convolution epilogue (conv_tc.cu); bn_apply turns
//             them into mean / rstd (saved for backward), updates running_mean / running_var (momentum 0.1, unbiased
//             variance, only for clients active in this step), and writes relu(y*scale + shift (+ residual)).
//   backward: bn_bwd_reduce -> per-channel sums of g' and g'*xhat (g' = upstream gradient masked by the ReLU);
//             bn_bwd_apply  -> dy = gamma*rstd*(g' - mean(g') - xhat*mean(g'*xhat)), d_gamma, d_beta; can scatter dy
//             onto the even positions of the 2x finer grid (input-gradient / weight-gradient of stride-2 convolutions
//             are computed at full resolution) and can emit g' for the identity-shortcut path.
// Ragged batches: samples b >= nvalid[k] are padding; they are excluded from the statistics (count = nvalid*H*W),
// produce zero activations and zero gradients.  Halo rows of every buffer are never written.
#include "common.cuh"

using namespace dba;

namespace {

struct BnArgs {
  // planar buffers in 16-byte rows of V channels (V = 8: bf16 mode, V = 4: fp32 / tf32 mode)
  const uint4* y;              // raw conv output      [K][C/V][R]
  const uint4* res;            // residual (fwd) or null
  uint4* out;                  // fwd: activation;  bwd: dy
  const uint4* g;              // bwd: upstream gradient
  const uint4* act;            // bwd: post-ReLU output of this BN (mask) or null
  uint4* gmask;                // bwd: optional copy of g' (same geometry as g)
  uint4* out2;                 // fwd: optional second copy of the activation in space-to-depth form (4C channels on the
                               // (H/2, W/2) grid, R2 rows per plane) for the next stage's stride-2 convolutions
  uint2* out_split;            // tf32 mode, optional: bf16 hi/lo split copy of `out` ([K][2][C/8][R or Rup][8]) for wgrad
  uint2* out2_split;           // tf32 mode, optional: split copy of out2 ([K][2][4C/8][R2][8])
  float* stats;                // fwd: [K][ld_stats] (sum, sumsq) pairs;  bwd: (sum g', sum g'*xhat) pairs
  float* saved;                // [K][ld_saved] (mean, rstd) pairs
  float* Wflat;                // [K][ldw] state rows (gamma, beta, running stats)
  float* Gflat;                // [K][ldg] gradient rows (bwd) or null
  const int* nvalid;           // [K]
  const unsigned char* active; // [K] or null
  long long ld_stats, ld_saved, ldw, ldg, R, Rup, R2;
  int stats_off, saved_off, g_off, b_off, rm_off, rv_off;
  int K, C, B, H, W, relu, up;
  float momentum, eps;
};


// grid (row blocks, C/V, K), 256 threads; each thread walks rows of one V-channel plane, kRowsInFlight rows per iteration
// with all loads issued before the first use.  Small grids (one or two clients per GPU) are latency-bound and take 4
// (K = 1 step: 1.43 -> 1.37 ms); big grids lose occupancy with 4 (K = 10: 5.62 -> 5.95 ms) and take 1.
template <int V, int kRowsInFlight>
__global__ void __launch_bounds__(256) bn_apply_kernel(const BnArgs a) {
  __shared__ float s_scale[V], s_shift[V];
  pdl_trigger();
  pdl_wait();
  const int chunk = blockIdx.y, k = blockIdx.z;
  const int nv = a.nvalid[k];
  const float n = (float)nv * (float)(a.H * a.W);
  if (threadIdx.x < V) {
    const int c = chunk * V + threadIdx.x;
    float* w = a.Wflat + (long long)k * a.ldw;
    const float s1 = a.stats[(long long)k * a.ld_stats + a.stats_off + 2 * c];
    const float s2 = a.stats[(long long)k * a.ld_stats + a.stats_off + 2 * c + 1];
    const float mean = n > 0.f ? s1 / n : 0.f;
    const float var = n > 0.f ? fmaxf(s2 / n - mean * mean, 0.f) : 0.f;
    const float rstd = 1.f / sqrtf(var + a.eps);
    const float sc = w[a.g_off + c] * rstd;
    s_scale[threadIdx.x] = sc;
    s_shift[threadIdx.x] = w[a.b_off + c] - mean * sc;
    if (blockIdx.x == 0) {
      a.saved[(long long)k * a.ld_saved + a.saved_off + 2 * c] = mean;
      a.saved[(long long)k * a.ld_saved + a.saved_off + 2 * c + 1] = rstd;
      if (nv > 0 && (!a.active || a.active[k])) {
        const float unb = n > 1.f ? var * n / (n - 1.f) : var;
        w[a.rm_off + c] = (1.f - a.momentum) * w[a.rm_off + c] + a.momentum * mean;
        w[a.rv_off + c] = (1.f - a.momentum) * w[a.rv_off + c] + a.momentum * unb;
      }
    }
  }
  __syncthreads();
  const int Wp = a.W + 1, S = (a.H + 1) * Wp, lead = Wp + 1;
  const long long plane = ((long long)k * (a.C / V) + chunk) * a.R;
  const int total = a.B * S;
  const int stride = gridDim.x * blockDim.x;
  for (int qb = blockIdx.x * blockDim.x + threadIdx.x; qb < total; qb += stride * kRowsInFlight) {
    uint4 ry[kRowsInFlight], rr[kRowsInFlight];
    bool ok[kRowsInFlight];
    int bb[kRowsInFlight], hh[kRowsInFlight], ww[kRowsInFlight];
#pragma unroll
    for (int u = 0; u < kRowsInFlight; ++u) {
      const int q = qb + u * stride;
      const int b = q / S, rem = q - b * S, h = rem / Wp, w = rem - h * Wp;
      bb[u] = b; hh[u] = h; ww[u] = w;
      ok[u] = q < total && h < a.H && w < a.W;
      if (ok[u] && b < nv) {
        ry[u] = a.y[plane + lead + q];
        if (a.res) rr[u] = a.res[plane + lead + q];
      }
    }
#pragma unroll
    for (int u = 0; u < kRowsInFlight; ++u) {
      if (!ok[u]) continue;
      const int q = qb + u * stride, b = bb[u], h = hh[u], w = ww[u];
      const long long off = plane + lead + q;
      float o[V];
      if (b < nv) {
        float v[V];
        chunk_unpack<V>(ry[u], v);
#pragma unroll
        for (int j = 0; j < V; ++j) o[j] = v[j] * s_scale[j] + s_shift[j];
        if (a.res) {
          float r[V];
          chunk_unpack<V>(rr[u], r);
#pragma unroll
          for (int j = 0; j < V; ++j) o[j] += r[j];
        }
        if (a.relu) {
#pragma unroll
          for (int j = 0; j < V; ++j) o[j] = relu_nan(o[j]);
        }
      } else {
#pragma unroll
        for (int j = 0; j < V; ++j) o[j] = 0.f;
      }
      const uint4 pk = chunk_pack_operand<V>(o);         // the activation is the next convolution's operand
      a.out[off] = pk;
      if constexpr (V == 4) {
        if (a.out_split) store_split(a.out_split, k, a.C >> 3, a.R, chunk, lead + q, o);
      }
      if (a.out2) {
        const int W2p = (a.W >> 1) + 1, S2 = ((a.H >> 1) + 1) * W2p, lead2 = W2p + 1;
        const int par = ((h & 1) << 1) | (w & 1);
        const long long plane2 = ((long long)k * 4 + par) * (a.C / V) + chunk;
        const long long row2 = lead2 + (long long)b * S2 + (h >> 1) * W2p + (w >> 1);
        a.out2[plane2 * a.R2 + row2] = pk;
        if constexpr (V == 4) {
          if (a.out2_split) store_split(a.out2_split, k, (4 * a.C) >> 3, a.R2, par * (a.C >> 2) + chunk, row2, o);
        }
      }
    }
  }
}

// tf32 mode with split copies: ONE thread handles the two 4-channel chunks that make up an 8-channel plane of the bf16
// hi/lo split tensor, so the split rows leave as full 16-byte stores (a thread of bn_apply_kernel<4> writes 8 bytes at a
// 16-byte stride: half-filled sectors, the kernel stayed at 48 % of the HBM copy bandwidth while the other BatchNorm
// kernels reach 67-70 %).  grid (row blocks, C/8, K).
__device__ __forceinline__ void split_pair_store(uint4* base, long long row16, long long lo_off16, const float (&o)[8]) {
  uint4 ph, pl;
  ph.x = pack_bf16x2(o[0], o[1]); ph.y = pack_bf16x2(o[2], o[3]); ph.z = pack_bf16x2(o[4], o[5]); ph.w = pack_bf16x2(o[6], o[7]);
  pl.x = pack_bf16x2(o[0] - bf16lo(ph.x), o[1] - bf16hi(ph.x));
  pl.y = pack_bf16x2(o[2] - bf16lo(ph.y), o[3] - bf16hi(ph.y));
  pl.z = pack_bf16x2(o[4] - bf16lo(ph.z), o[5] - bf16hi(ph.z));
  pl.w = pack_bf16x2(o[6] - bf16lo(ph.w), o[7] - bf16hi(ph.w));
  base[row16] = ph;
  base[row16 + lo_off16] = pl;
}

template <int kRowsInFlight>
__global__ void __launch_bounds__(256) bn_apply_pair_kernel(const BnArgs a) {
  __shared__ float s_scale[8], s_shift[8];
  pdl_trigger();
  pdl_wait();
  const int p8 = blockIdx.y, k = blockIdx.z;
  const int nv = a.nvalid[k];
  const float n = (float)nv * (float)(a.H * a.W);
  if (threadIdx.x < 8) {
    const int c = p8 * 8 + threadIdx.x;
    float* w = a.Wflat + (long long)k * a.ldw;
    const float s1 = a.stats[(long long)k * a.ld_stats + a.stats_off + 2 * c];
    const float s2 = a.stats[(long long)k * a.ld_stats + a.stats_off + 2 * c + 1];
    const float mean = n > 0.f ? s1 / n : 0.f;
    const float var = n > 0.f ? fmaxf(s2 / n - mean * mean, 0.f) : 0.f;
    const float rstd = 1.f / sqrtf(var + a.eps);
    const float sc = w[a.g_off + c] * rstd;
    s_scale[threadIdx.x] = sc;
    s_shift[threadIdx.x] = w[a.b_off + c] - mean * sc;
    if (blockIdx.x == 0) {
      a.saved[(long long)k * a.ld_saved + a.saved_off + 2 * c] = mean;
      a.saved[(long long)k * a.ld_saved + a.saved_off + 2 * c + 1] = rstd;
      if (nv > 0 && (!a.active || a.active[k])) {
        const float unb = n > 1.f ? var * n / (n - 1.f) : var;
        w[a.rm_off + c] = (1.f - a.momentum) * w[a.rm_off + c] + a.momentum * mean;
        w[a.rv_off + c] = (1.f - a.momentum) * w[a.rv_off + c] + a.momentum * unb;
      }
    }
  }
  __syncthreads();
  const int Wp = a.W + 1, S = (a.H + 1) * Wp, lead = Wp + 1;
  const int C4 = a.C >> 2, C8 = a.C >> 3;
  const long long plane0 = ((long long)k * C4 + 2 * p8) * a.R, plane1 = plane0 + a.R;
  uint4* split = reinterpret_cast<uint4*>(a.out_split);
  uint4* split2 = reinterpret_cast<uint4*>(a.out2_split);
  const long long sp_base = ((long long)k * 2 * C8 + p8) * a.R, sp_lo = (long long)C8 * a.R;
  const int total = a.B * S;
  const int stride = gridDim.x * blockDim.x;
  for (int qb = blockIdx.x * blockDim.x + threadIdx.x; qb < total; qb += stride * kRowsInFlight) {
    uint4 ry0[kRowsInFlight], ry1[kRowsInFlight], rr0[kRowsInFlight], rr1[kRowsInFlight];
    bool ok[kRowsInFlight];
    int bb[kRowsInFlight], hh[kRowsInFlight], ww[kRowsInFlight];
#pragma unroll
    for (int u = 0; u < kRowsInFlight; ++u) {
      const int q = qb + u * stride;
      const int b = q / S, rem = q - b * S, h = rem / Wp, w = rem - h * Wp;
      bb[u] = b; hh[u] = h; ww[u] = w;
      ok[u] = q < total && h < a.H && w < a.W;
      if (ok[u] && b < nv) {
        ry0[u] = a.y[plane0 + lead + q];
        ry1[u] = a.y[plane1 + lead + q];
        if (a.res) { rr0[u] = a.res[plane0 + lead + q]; rr1[u] = a.res[plane1 + lead + q]; }
      }
    }
#pragma unroll
    for (int u = 0; u < kRowsInFlight; ++u) {
      if (!ok[u]) continue;
      const int q = qb + u * stride, b = bb[u], h = hh[u], w = ww[u];
      float o[8];
      if (b < nv) {
        float v0[4], v1[4];
        chunk_unpack<4>(ry0[u], v0);
        chunk_unpack<4>(ry1[u], v1);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          o[j] = v0[j] * s_scale[j] + s_shift[j];
          o[4 + j] = v1[j] * s_scale[4 + j] + s_shift[4 + j];
        }
        if (a.res) {
          float r0[4], r1[4];
          chunk_unpack<4>(rr0[u], r0);
          chunk_unpack<4>(rr1[u], r1);
#pragma unroll
          for (int j = 0; j < 4; ++j) { o[j] += r0[j]; o[4 + j] += r1[j]; }
        }
        if (a.relu) {
#pragma unroll
          for (int j = 0; j < 8; ++j) o[j] = relu_nan(o[j]);
        }
      } else {
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = 0.f;
      }
      const float o0[4] = {o[0], o[1], o[2], o[3]}, o1[4] = {o[4], o[5], o[6], o[7]};
      const uint4 pk0 = chunk_pack_operand<4>(o0), pk1 = chunk_pack_operand<4>(o1);
      a.out[plane0 + lead + q] = pk0;
      a.out[plane1 + lead + q] = pk1;
      if (split) split_pair_store(split, sp_base + lead + q, sp_lo, o);
      if (a.out2) {
        const int W2p = (a.W >> 1) + 1, S2 = ((a.H >> 1) + 1) * W2p, lead2 = W2p + 1;
        const int par = ((h & 1) << 1) | (w & 1);
        const long long p2 = (((long long)k * 4 + par) * C4 + 2 * p8) * a.R2;
        const long long row2 = lead2 + (long long)b * S2 + (h >> 1) * W2p + (w >> 1);
        a.out2[p2 + row2] = pk0;
        a.out2[p2 + a.R2 + row2] = pk1;
        if (split2) {       // [K][2][4C/8][R2][8]: plane8 of the space-to-depth tensor = par * C/8 + p8
          const long long s2 = ((long long)k * 2 * (4 * C8) + (long long)par * C8 + p8) * a.R2;
          split_pair_store(split2, s2 + row2, (long long)(4 * C8) * a.R2, o);
        }
      }
    }
  }
}

template <int V, int kRowsInFlight>
__global__ void __launch_bounds__(256) bn_bwd_reduce_kernel(const BnArgs a) {
  __shared__ float red[2][8][V];
  pdl_trigger();
  pdl_wait();
  const int chunk = blockIdx.y, k = blockIdx.z;
  const int nv = a.nvalid[k];
  float mean[V], rstd[V];
#pragma unroll
  for (int j = 0; j < V; ++j) {
    mean[j] = a.saved[(long long)k * a.ld_saved + a.saved_off + 2 * (chunk * V + j)];
    rstd[j] = a.saved[(long long)k * a.ld_saved + a.saved_off + 2 * (chunk * V + j) + 1];
  }
  const int Wp = a.W + 1, S = (a.H + 1) * Wp, lead = Wp + 1;
  const long long plane = ((long long)k * (a.C / V) + chunk) * a.R;
  const int total = min(a.B, nv) * S;
  float sg[V], sgx[V];
#pragma unroll
  for (int j = 0; j < V; ++j) sg[j] = sgx[j] = 0.f;
  const int stride = gridDim.x * blockDim.x;
  for (int qb = blockIdx.x * blockDim.x + threadIdx.x; qb < total; qb += stride * kRowsInFlight) {
    uint4 rg[kRowsInFlight], ry[kRowsInFlight], ro[kRowsInFlight];
    bool ok[kRowsInFlight];
#pragma unroll
    for (int u = 0; u < kRowsInFlight; ++u) {
      const int q = qb + u * stride;
      const int b = q / S, rem = q - b * S, h = rem / Wp, w = rem - h * Wp;
      ok[u] = q < total && h < a.H && w < a.W;
      if (ok[u]) {
        const long long off = plane + lead + q;
        rg[u] = a.g[off];
        ry[u] = a.y[off];
        if (a.act) ro[u] = a.act[off];
      }
    }
#pragma unroll
    for (int u = 0; u < kRowsInFlight; ++u) {
      if (!ok[u]) continue;
      float g[V], y[V];
      chunk_unpack<V>(rg[u], g);
      chunk_unpack<V>(ry[u], y);
      if (a.act) {
        float o[V];
        chunk_unpack<V>(ro[u], o);
#pragma unroll
        for (int j = 0; j < V; ++j) g[j] = o[j] > 0.f ? g[j] : 0.f;
      }
#pragma unroll
      for (int j = 0; j < V; ++j) {
        sg[j] += g[j];
        sgx[j] += g[j] * (y[j] - mean[j]) * rstd[j];
      }
    }
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int j = 0; j < V; ++j) {
    sg[j] = warp_sum(sg[j]);
    sgx[j] = warp_sum(sgx[j]);
  }
  if (lane == 0) {
#pragma unroll
    for (int j = 0; j < V; ++j) { red[0][warp][j] = sg[j]; red[1][warp][j] = sgx[j]; }
  }
  __syncthreads();
  if (threadIdx.x < 2 * V) {
    const int which = threadIdx.x / V, j = threadIdx.x % V;
    float t = 0.f;
    for (int w = 0; w < 8; ++w) t += red[which][w][j];
    if (t != 0.f) atomicAdd(&a.stats[(long long)k * a.ld_stats + a.stats_off + 2 * (chunk * V + j) + which], t);
  }
}

template <int V, int kRowsInFlight>
__global__ void __launch_bounds__(256) bn_bwd_apply_kernel(const BnArgs a) {
  __shared__ float s_sc[V], s_mg[V], s_mgx[V], s_mean[V], s_rstd[V];
  pdl_trigger();
  pdl_wait();
  const int chunk = blockIdx.y, k = blockIdx.z;
  const int nv = a.nvalid[k];
  const float n = (float)nv * (float)(a.H * a.W);
  if (threadIdx.x < V) {
    const int c = chunk * V + threadIdx.x;
    const float sg = a.stats[(long long)k * a.ld_stats + a.stats_off + 2 * c];
    const float sgx = a.stats[(long long)k * a.ld_stats + a.stats_off + 2 * c + 1];
    const float mean = a.saved[(long long)k * a.ld_saved + a.saved_off + 2 * c];
    const float rstd = a.saved[(long long)k * a.ld_saved + a.saved_off + 2 * c + 1];
    const float gamma = a.Wflat[(long long)k * a.ldw + a.g_off + c];
    s_sc[threadIdx.x] = gamma * rstd;
    s_mg[threadIdx.x] = n > 0.f ? sg / n : 0.f;
    s_mgx[threadIdx.x] = n > 0.f ? sgx / n : 0.f;
    s_mean[threadIdx.x] = mean;
    s_rstd[threadIdx.x] = rstd;
    if (blockIdx.x == 0 && a.Gflat) {
      a.Gflat[(long long)k * a.ldg + a.g_off + c] = sgx;   // d gamma
      a.Gflat[(long long)k * a.ldg + a.b_off + c] = sg;    // d beta
    }
  }
  __syncthreads();
  const int Wp = a.W + 1, S = (a.H + 1) * Wp, lead = Wp + 1;
  const int Wpu = 2 * a.W + 1, Su = (2 * a.H + 1) * Wpu, leadu = Wpu + 1;
  const long long plane = ((long long)k * (a.C / V) + chunk) * a.R;
  const long long plane_up = ((long long)k * (a.C / V) + chunk) * a.Rup;
  const int total = a.B * S;
  const int stride = gridDim.x * blockDim.x;
  for (int qb = blockIdx.x * blockDim.x + threadIdx.x; qb < total; qb += stride * kRowsInFlight) {
    uint4 rg[kRowsInFlight], ry[kRowsInFlight], ro[kRowsInFlight];
    bool ok[kRowsInFlight];
    int bb[kRowsInFlight], hh[kRowsInFlight], ww[kRowsInFlight];
#pragma unroll
    for (int u = 0; u < kRowsInFlight; ++u) {
      const int q = qb + u * stride;
      const int b = q / S, rem = q - b * S, h = rem / Wp, w = rem - h * Wp;
      bb[u] = b; hh[u] = h; ww[u] = w;
      ok[u] = q < total && h < a.H && w < a.W;
      if (ok[u] && b < nv) {
        const long long off = plane + lead + q;
        rg[u] = a.g[off];
        ry[u] = a.y[off];
        if (a.act) ro[u] = a.act[off];
      }
    }
#pragma unroll
    for (int u = 0; u < kRowsInFlight; ++u) {
      if (!ok[u]) continue;
      const int q = qb + u * stride, b = bb[u], h = hh[u], w = ww[u];
      const long long off = plane + lead + q;
      float d[V], gm[V];
      if (b < nv) {
        float g[V], y[V];
        chunk_unpack<V>(rg[u], g);
        chunk_unpack<V>(ry[u], y);
        if (a.act) {
          float o[V];
          chunk_unpack<V>(ro[u], o);
#pragma unroll
          for (int j = 0; j < V; ++j) g[j] = o[j] > 0.f ? g[j] : 0.f;
        }
#pragma unroll
        for (int j = 0; j < V; ++j) {
          const float xh = (y[j] - s_mean[j]) * s_rstd[j];
          gm[j] = g[j];
          d[j] = s_sc[j] * (g[j] - s_mg[j] - xh * s_mgx[j]);
        }
      } else {
#pragma unroll
        for (int j = 0; j < V; ++j) d[j] = gm[j] = 0.f;
      }
      if (a.gmask) a.gmask[off] = chunk_pack<V>(gm);
      long long doff = off, drow = lead + q;
      if (a.up) {
        drow = leadu + (long long)b * Su + (2 * h) * Wpu + 2 * w;
        doff = plane_up + drow;
      }
      a.out[doff] = chunk_pack_operand<V>(d);            // dy is the input-gradient convolution's operand
      if constexpr (V == 4) {
        if (a.out_split) store_split(a.out_split, k, a.C >> 3, a.up ? a.Rup : a.R, chunk, drow, d);
      }
    }
  }
}

// bn_bwd_apply on 8-channel planes (tf32 mode with a split copy of dy): see bn_apply_pair_kernel.
template <int kRowsInFlight>
__global__ void __launch_bounds__(256) bn_bwd_apply_pair_kernel(const BnArgs a) {
  __shared__ float s_sc[8], s_mg[8], s_mgx[8], s_mean[8], s_rstd[8];
  pdl_trigger();
  pdl_wait();
  const int p8 = blockIdx.y, k = blockIdx.z;
  const int nv = a.nvalid[k];
  const float n = (float)nv * (float)(a.H * a.W);
  if (threadIdx.x < 8) {
    const int c = p8 * 8 + threadIdx.x;
    const float sg = a.stats[(long long)k * a.ld_stats + a.stats_off + 2 * c];
    const float sgx = a.stats[(long long)k * a.ld_stats + a.stats_off + 2 * c + 1];
    const float mean = a.saved[(long long)k * a.ld_saved + a.saved_off + 2 * c];
    const float rstd = a.saved[(long long)k * a.ld_saved + a.saved_off + 2 * c + 1];
    const float gamma = a.Wflat[(long long)k * a.ldw + a.g_off + c];
    s_sc[threadIdx.x] = gamma * rstd;
    s_mg[threadIdx.x] = n > 0.f ? sg / n : 0.f;
    s_mgx[threadIdx.x] = n > 0.f ? sgx / n : 0.f;
    s_mean[threadIdx.x] = mean;
    s_rstd[threadIdx.x] = rstd;
    if (blockIdx.x == 0 && a.Gflat) {
      a.Gflat[(long long)k * a.ldg + a.g_off + c] = sgx;   // d gamma
      a.Gflat[(long long)k * a.ldg + a.b_off + c] = sg;    // d beta
    }
  }
  __syncthreads();
  const int Wp = a.W + 1, S = (a.H + 1) * Wp, lead = Wp + 1;
  const int Wpu = 2 * a.W + 1, Su = (2 * a.H + 1) * Wpu, leadu = Wpu + 1;
  const int C4 = a.C >> 2, C8 = a.C >> 3;
  const long long plane0 = ((long long)k * C4 + 2 * p8) * a.R, plane1 = plane0 + a.R;
  const long long Rd = a.up ? a.Rup : a.R;                       // rows of a dy plane
  const long long dplane0 = ((long long)k * C4 + 2 * p8) * Rd, dplane1 = dplane0 + Rd;
  uint4* split = reinterpret_cast<uint4*>(a.out_split);
  const long long sp_base = ((long long)k * 2 * C8 + p8) * Rd, sp_lo = (long long)C8 * Rd;
  const int total = a.B * S;
  const int stride = gridDim.x * blockDim.x;
  for (int qb = blockIdx.x * blockDim.x + threadIdx.x; qb < total; qb += stride * kRowsInFlight) {
    uint4 rg0[kRowsInFlight], rg1[kRowsInFlight], ry0[kRowsInFlight], ry1[kRowsInFlight], ro0[kRowsInFlight],
        ro1[kRowsInFlight];
    bool ok[kRowsInFlight];
    int bb[kRowsInFlight], hh[kRowsInFlight], ww[kRowsInFlight];
#pragma unroll
    for (int u = 0; u < kRowsInFlight; ++u) {
      const int q = qb + u * stride;
      const int b = q / S, rem = q - b * S, h = rem / Wp, w = rem - h * Wp;
      bb[u] = b; hh[u] = h; ww[u] = w;
      ok[u] = q < total && h < a.H && w < a.W;
      if (ok[u] && b < nv) {
        const long long o0 = plane0 + lead + q, o1 = plane1 + lead + q;
        rg0[u] = a.g[o0]; rg1[u] = a.g[o1];
        ry0[u] = a.y[o0]; ry1[u] = a.y[o1];
        if (a.act) { ro0[u] = a.act[o0]; ro1[u] = a.act[o1]; }
      }
    }
#pragma unroll
    for (int u = 0; u < kRowsInFlight; ++u) {
      if (!ok[u]) continue;
      const int q = qb + u * stride, b = bb[u], h = hh[u], w = ww[u];
      float d[8], gm[8];
      if (b < nv) {
        float g[8], y[8];
        {
          float t0[4], t1[4];
          chunk_unpack<4>(rg0[u], t0); chunk_unpack<4>(rg1[u], t1);
#pragma unroll
          for (int j = 0; j < 4; ++j) { g[j] = t0[j]; g[4 + j] = t1[j]; }
          chunk_unpack<4>(ry0[u], t0); chunk_unpack<4>(ry1[u], t1);
#pragma unroll
          for (int j = 0; j < 4; ++j) { y[j] = t0[j]; y[4 + j] = t1[j]; }
          if (a.act) {
            chunk_unpack<4>(ro0[u], t0); chunk_unpack<4>(ro1[u], t1);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              g[j] = t0[j] > 0.f ? g[j] : 0.f;
              g[4 + j] = t1[j] > 0.f ? g[4 + j] : 0.f;
            }
          }
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float xh = (y[j] - s_mean[j]) * s_rstd[j];
          gm[j] = g[j];
          d[j] = s_sc[j] * (g[j] - s_mg[j] - xh * s_mgx[j]);
        }
      } else {
#pragma unroll
        for (int j = 0; j < 8; ++j) d[j] = gm[j] = 0.f;
      }
      if (a.gmask) {
        const float m0[4] = {gm[0], gm[1], gm[2], gm[3]}, m1[4] = {gm[4], gm[5], gm[6], gm[7]};
        a.gmask[plane0 + lead + q] = chunk_pack<4>(m0);
        a.gmask[plane1 + lead + q] = chunk_pack<4>(m1);
      }
      long long drow = lead + q;
      if (a.up) drow = leadu + (long long)b * Su + (2 * h) * Wpu + 2 * w;
      const float d0[4] = {d[0], d[1], d[2], d[3]}, d1[4] = {d[4], d[5], d[6], d[7]};
      a.out[dplane0 + drow] = chunk_pack_operand<4>(d0);          // dy is the input-gradient convolution's operand
      a.out[dplane1 + drow] = chunk_pack_operand<4>(d1);
      split_pair_store(split, sp_base + drow, sp_lo, d);
    }
  }
}

// ------------------------------------------------------------------------------------------------ fused backward
// Reduce + apply in ONE launch: the row blocks of a (client, channel-plane) pair form a thread-block cluster; every CTA
// reduces its slice, the partial sums are exchanged through distributed shared memory (fixed summation order, no
// atomics, no zeroed scratch), and every CTA applies the result to the same slice, which it re-reads from L2.
// Saves one launch + one grid drain per BatchNorm layer of a training step and makes d_gamma / d_beta deterministic.
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_arrive() { asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory"); }
__device__ __forceinline__ void cluster_wait() { asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory"); }
__device__ __forceinline__ float dsmem_ld(const float* local, uint32_t rank) {
  uint32_t addr = (uint32_t)__cvta_generic_to_shared(local), raddr;
  float v;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(raddr) : "r"(addr), "r"(rank));
  asm volatile("ld.shared::cluster.f32 %0, [%1];" : "=f"(v) : "r"(raddr) : "memory");
  return v;
}

constexpr int kBnU = 4;        // rows in flight per thread (independent 16-byte loads)

// grid (cluster size X, C/V, K), cluster (X, 1, 1), 256 threads.
template <int V>
__global__ void __launch_bounds__(256) bn_bwd_fused_kernel(const BnArgs a) {
  __shared__ float red[2][8][V];
  __shared__ float part[2][V];
  __shared__ float s_sc[V], s_mg[V], s_mgx[V], s_mean[V], s_rstd[V];
  pdl_trigger();
  pdl_wait();
  const int chunk = blockIdx.y, k = blockIdx.z;
  const int nv = a.nvalid[k];
  const float n = (float)nv * (float)(a.H * a.W);
  float mean[V], rstd[V];
#pragma unroll
  for (int j = 0; j < V; ++j) {
    mean[j] = a.saved[(long long)k * a.ld_saved + a.saved_off + 2 * (chunk * V + j)];
    rstd[j] = a.saved[(long long)k * a.ld_saved + a.saved_off + 2 * (chunk * V + j) + 1];
  }
  const int Wp = a.W + 1, S = (a.H + 1) * Wp, lead = Wp + 1;
  const long long plane = ((long long)k * (a.C / V) + chunk) * a.R;
  const int stride = gridDim.x * blockDim.x;
  const int q0 = blockIdx.x * blockDim.x + threadIdx.x;
  // ---------------------------------------------------------------- pass 1: sums of g' and g' * xhat over the slice
  {
    const int total = min(a.B, nv) * S;
    float sg[V], sgx[V];
#pragma unroll
    for (int j = 0; j < V; ++j) sg[j] = sgx[j] = 0.f;
    for (int qb = q0; qb < total; qb += stride * kBnU) {
      uint4 rg[kBnU], ry[kBnU], ro[kBnU];
      bool ok[kBnU];
#pragma unroll
      for (int u = 0; u < kBnU; ++u) {
        const int q = qb + u * stride;
        const int b = q / S, rem = q - b * S, h = rem / Wp, w = rem - h * Wp;
        ok[u] = q < total && h < a.H && w < a.W;
        if (ok[u]) {
          const long long off = plane + lead + q;
          rg[u] = a.g[off];
          ry[u] = a.y[off];
          if (a.act) ro[u] = a.act[off];
        }
      }
#pragma unroll
      for (int u = 0; u < kBnU; ++u) {
        if (!ok[u]) continue;
        float g[V], y[V];
        chunk_unpack<V>(rg[u], g);
        chunk_unpack<V>(ry[u], y);
        if (a.act) {
          float o[V];
          chunk_unpack<V>(ro[u], o);
#pragma unroll
          for (int j = 0; j < V; ++j) g[j] = o[j] > 0.f ? g[j] : 0.f;
        }
#pragma unroll
        for (int j = 0; j < V; ++j) {
          sg[j] += g[j];
          sgx[j] += g[j] * (y[j] - mean[j]) * rstd[j];
        }
      }
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int j = 0; j < V; ++j) {
      sg[j] = warp_sum(sg[j]);
      sgx[j] = warp_sum(sgx[j]);
    }
    if (lane == 0) {
#pragma unroll
      for (int j = 0; j < V; ++j) { red[0][warp][j] = sg[j]; red[1][warp][j] = sgx[j]; }
    }
    __syncthreads();
    if (threadIdx.x < 2 * V) {
      const int which = threadIdx.x / V, j = threadIdx.x % V;
      float t = 0.f;
      for (int w = 0; w < 8; ++w) t += red[which][w][j];
      part[which][j] = t;
    }
  }
  cluster_arrive();
  cluster_wait();                               // every CTA's partial sums are in its shared memory
  if (threadIdx.x < V) {
    const int j = threadIdx.x, c = chunk * V + j;
    float sg = 0.f, sgx = 0.f;
    for (uint32_t r = 0; r < gridDim.x; ++r) {   // rank order: the same sum in every CTA, run to run
      sg += dsmem_ld(&part[0][j], r);
      sgx += dsmem_ld(&part[1][j], r);
    }
    const float gamma = a.Wflat[(long long)k * a.ldw + a.g_off + c];
    const float mean_c = a.saved[(long long)k * a.ld_saved + a.saved_off + 2 * c];
    const float rstd_c = a.saved[(long long)k * a.ld_saved + a.saved_off + 2 * c + 1];
    s_sc[j] = gamma * rstd_c;
    s_mg[j] = n > 0.f ? sg / n : 0.f;
    s_mgx[j] = n > 0.f ? sgx / n : 0.f;
    s_mean[j] = mean_c;
    s_rstd[j] = rstd_c;
    if (cluster_ctarank() == 0 && a.Gflat) {
      a.Gflat[(long long)k * a.ldg + a.g_off + c] = sgx;   // d gamma
      a.Gflat[(long long)k * a.ldg + a.b_off + c] = sg;    // d beta
    }
  }
  cluster_arrive();                             // done reading the peers' shared memory (waited for before exit)
  __syncthreads();
  // ---------------------------------------------------------------- pass 2: dy (and g') for the same slice
  {
    const int Wpu = 2 * a.W + 1, Su = (2 * a.H + 1) * Wpu, leadu = Wpu + 1;
    const long long plane_up = ((long long)k * (a.C / V) + chunk) * a.Rup;
    const int total = a.B * S;
    for (int qb = q0; qb < total; qb += stride * kBnU) {
      uint4 rg[kBnU], ry[kBnU], ro[kBnU];
      bool ok[kBnU];
      int bb[kBnU], hh[kBnU], ww[kBnU];
#pragma unroll
      for (int u = 0; u < kBnU; ++u) {
        const int q = qb + u * stride;
        const int b = q / S, rem = q - b * S, h = rem / Wp, w = rem - h * Wp;
        bb[u] = b; hh[u] = h; ww[u] = w;
        ok[u] = q < total && h < a.H && w < a.W;
        if (ok[u] && b < nv) {
          const long long off = plane + lead + q;
          rg[u] = a.g[off];
          ry[u] = a.y[off];
          if (a.act) ro[u] = a.act[off];
        }
      }
#pragma unroll
      for (int u = 0; u < kBnU; ++u) {
        if (!ok[u]) continue;
        const int q = qb + u * stride;
        const long long off = plane + lead + q;
        float d[V], gm[V];
        if (bb[u] < nv) {
          float g[V], y[V];
          chunk_unpack<V>(rg[u], g);
          chunk_unpack<V>(ry[u], y);
          if (a.act) {
            float o[V];
            chunk_unpack<V>(ro[u], o);
#pragma unroll
            for (int j = 0; j < V; ++j) g[j] = o[j] > 0.f ? g[j] : 0.f;
          }
#pragma unroll
          for (int j = 0; j < V; ++j) {
            const float xh = (y[j] - s_mean[j]) * s_rstd[j];
            gm[j] = g[j];
            d[j] = s_sc[j] * (g[j] - s_mg[j] - xh * s_mgx[j]);
          }
        } else {
#pragma unroll
          for (int j = 0; j < V; ++j) d[j] = gm[j] = 0.f;
        }
        if (a.gmask) a.gmask[off] = chunk_pack<V>(gm);
        long long doff = off, drow = lead + q;
        if (a.up) {
          drow = leadu + (long long)bb[u] * Su + (2 * hh[u]) * Wpu + 2 * ww[u];
          doff = plane_up + drow;
        }
        a.out[doff] = chunk_pack_operand<V>(d);          // dy is the input-gradient convolution's operand
        if constexpr (V == 4) {
          if (a.out_split) store_split(a.out_split, k, a.C >> 3, a.up ? a.Rup : a.R, chunk, drow, d);
        }
      }
    }
  }
  cluster_wait();                               // no CTA leaves while a peer may still read its partial sums
}

// ------------------------------------------------------------------------------------------------ max pooling
// MaxPool2d(3, stride 2, pad 1) on planar activations (Tiny-ImageNet stem, models/resnet_tinyimagenet.py:147).
// Inputs are post-ReLU (>= 0) and the halo rows / columns of the layout are zero, so the window can be read without
// bounds checks: a zero never beats a real maximum.  Backward is gather-style (no atomics): an input pixel receives
// the gradient of every output window in which it is the first maximum in scan order (PyTorch's tie rule).
struct PoolArgs {
  const uint4* x;             // [K][C/V][Rin]   H x W grid
  uint4* y;                   // [K][C/V][Rout]  H/2 x W/2 grid (fwd out)
  const uint4* gy;            // bwd: gradient wrt y
  uint4* gx;                  // bwd: gradient wrt x
  uint2* y_split;             // fwd, tf32 mode, optional: bf16 hi/lo split copy of y for wgrad
  long long Rin, Rout;
  int K, C, B, H, W;
};
template <int V>
__device__ __forceinline__ uint4 chunk_max(const uint4 a, const uint4 b) {
  float x[V], y[V];
  chunk_unpack<V>(a, x); chunk_unpack<V>(b, y);
#pragma unroll
  for (int j = 0; j < V; ++j) x[j] = fmaxf(x[j], y[j]);
  return chunk_pack<V>(x);
}
template <int V>
__global__ void __launch_bounds__(256) maxpool_fwd_kernel(const PoolArgs a) {
  const int chunk = blockIdx.y, k = blockIdx.z;
  const int Ho = a.H >> 1, Wo = a.W >> 1;
  const int Wp = a.W + 1, S = (a.H + 1) * Wp, lead = Wp + 1;
  const int Wpo = Wo + 1, So = (Ho + 1) * Wpo, leado = Wpo + 1;
  const long long pin = ((long long)k * (a.C / V) + chunk) * a.Rin, pout = ((long long)k * (a.C / V) + chunk) * a.Rout;
  const int total = a.B * Ho * Wo;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int wo = i % Wo, ho = (i / Wo) % Ho, b = i / (Wo * Ho);
    const long long c = pin + lead + (long long)b * S + (2 * ho) * Wp + 2 * wo;      // window centre (h=2ho, w=2wo)
    uint4 m = a.x[c - Wp - 1];
#pragma unroll
    for (int dh = -1; dh <= 1; ++dh)
#pragma unroll
      for (int dw = -1; dw <= 1; ++dw)
        if (dh != -1 || dw != -1) m = chunk_max<V>(m, a.x[c + dh * Wp + dw]);
    a.y[pout + leado + (long long)b * So + ho * Wpo + wo] = m;
    if constexpr (V == 4) {
      if (a.y_split) {
        float f[4];
        chunk_unpack<4>(m, f);
        store_split(a.y_split, k, a.C >> 3, a.Rout, chunk, leado + (long long)b * So + ho * Wpo + wo, f);
      }
    }
  }
}
template <int V>
__global__ void __launch_bounds__(256) maxpool_bwd_kernel(const PoolArgs a) {
  const int chunk = blockIdx.y, k = blockIdx.z;
  const int Ho = a.H >> 1, Wo = a.W >> 1;
  const int Wp = a.W + 1, S = (a.H + 1) * Wp, lead = Wp + 1;
  const int Wpo = Wo + 1, So = (Ho + 1) * Wpo, leado = Wpo + 1;
  const long long pin = ((long long)k * (a.C / V) + chunk) * a.Rin, pout = ((long long)k * (a.C / V) + chunk) * a.Rout;
  const int total = a.B * a.H * a.W;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int w = i % a.W, h = (i / a.W) % a.H, b = i / (a.W * a.H);
    const long long me = pin + lead + (long long)b * S + h * Wp + w;
    float xv[V], acc[V];
    chunk_unpack<V>(a.x[me], xv);
#pragma unroll
    for (int j = 0; j < V; ++j) acc[j] = 0.f;
    // output windows (ho, wo) with |2ho - h| <= 1 and |2wo - w| <= 1
    for (int ho = (h >> 1); ho <= ((h + 1) >> 1) && ho < Ho; ++ho)
      for (int wo = (w >> 1); wo <= ((w + 1) >> 1) && wo < Wo; ++wo) {
        const long long c = pin + lead + (long long)b * S + (2 * ho) * Wp + 2 * wo;
        float g[V];
        chunk_unpack<V>(a.gy[pout + leado + (long long)b * So + ho * Wpo + wo], g);
        // first maximum in scan order per channel; positions before `me` in the scan must be strictly smaller,
        // positions after it smaller or equal
        bool first[V];
#pragma unroll
        for (int j = 0; j < V; ++j) first[j] = true;
        const int my_dh = h - 2 * ho, my_dw = w - 2 * wo;
        for (int dh = -1; dh <= 1; ++dh)
          for (int dw = -1; dw <= 1; ++dw) {
            if (dh == my_dh && dw == my_dw) continue;
            const int hh = 2 * ho + dh, ww = 2 * wo + dw;
            if (hh < 0 || hh >= a.H || ww < 0 || ww >= a.W) continue;       // padding never wins (-inf in PyTorch)
            float o[V];
            chunk_unpack<V>(a.x[c + dh * Wp + dw], o);
            const bool before = dh < my_dh || (dh == my_dh && dw < my_dw);
#pragma unroll
            for (int j = 0; j < V; ++j) first[j] = first[j] && (before ? o[j] < xv[j] : o[j] <= xv[j]);
          }
#pragma unroll
        for (int j = 0; j < V; ++j) acc[j] += first[j] ? g[j] : 0.f;
      }
    a.gx[me] = chunk_pack<V>(acc);
  }
}

int g_bn_pair = 1;     // tf32 mode: bn_apply on 8-channel planes (A/B switch dba_set_bn_pair)
int g_bn_fused = 0;   // measured slower than reduce + apply (K = 1: 1.65 vs 1.59 ms / step, K = 10: 6.36 vs 5.96)

template <typename... KArgs, typename... Args>
inline cudaError_t launch_cluster_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, int cluster_x, cudaStream_t st,
                                      Args... args) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = 0;
  cfg.stream = st;
  cudaLaunchAttribute at[2];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = cluster_x;
  at[0].val.clusterDim.y = 1;
  at[0].val.clusterDim.z = 1;
  at[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[1].val.programmaticStreamSerializationAllowed = g_dba_pdl ? 1 : 0;
  cfg.attrs = at;
  cfg.numAttrs = 2;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

inline int act_v() { return g_dba_act_mode == 1 ? 4 : 8; }
inline dim3 bn_grid(int B, int H, int W, int C, int K) {
  const int total = B * (H + 1) * (W + 1);
  int rb = (total + 1023) / 1024;
  if (rb < 1) rb = 1;
  if (rb > 64) rb = 64;
  return dim3(rb, C / act_v(), K);     // (fewer, longer row blocks for big grids were measured: K = 10 step 5.62 -> 5.86 ms)
}
inline bool bn_small(const dim3& g) { return (long long)g.x * g.y * g.z <= 1100; }

}  // namespace

extern "C" {

// out = relu?( (y - mean)*rstd*gamma + beta (+ res) ); also saves (mean, rstd) and updates the running statistics.
// stats: pairs (sum, sumsq) at stats[k*ld_stats + stats_off + 2c]; saved likewise.  Offsets g/b/rm/rv index Wflat rows.
int dba_bn_apply_split(const void* y, const void* res, void* out, void* out_s2d, long long R_s2d, void* out_split,
                       void* out_s2d_split, float* stats, long long ld_stats, int stats_off, float* saved,
                       long long ld_saved, int saved_off, float* Wflat, long long ldw, int g_off, int b_off, int rm_off,
                       int rv_off, const int* nvalid, const unsigned char* active, int K, int C, int B, int H, int W,
                       long long R, int relu, float momentum, float eps, void* stream);
int dba_bn_apply_ex(const void* y, const void* res, void* out, void* out_s2d, long long R_s2d, float* stats,
                    long long ld_stats, int stats_off, float* saved, long long ld_saved, int saved_off, float* Wflat,
                    long long ldw, int g_off, int b_off, int rm_off, int rv_off, const int* nvalid,
                    const unsigned char* active, int K, int C, int B, int H, int W, long long R, int relu,
                    float momentum, float eps, void* stream) {
  return dba_bn_apply_split(y, res, out, out_s2d, R_s2d, nullptr, nullptr, stats, ld_stats, stats_off, saved, ld_saved,
                            saved_off, Wflat, ldw, g_off, b_off, rm_off, rv_off, nvalid, active, K, C, B, H, W, R, relu,
                            momentum, eps, stream);
}
int dba_bn_apply(const void* y, const void* res, void* out, float* stats, long long ld_stats, int stats_off,
                 float* saved, long long ld_saved, int saved_off, float* Wflat, long long ldw, int g_off, int b_off,
                 int rm_off, int rv_off, const int* nvalid, const unsigned char* active, int K, int C, int B, int H,
                 int W, long long R, int relu, float momentum, float eps, void* stream) {
  return dba_bn_apply_ex(y, res, out, nullptr, 0, stats, ld_stats, stats_off, saved, ld_saved, saved_off, Wflat, ldw,
                         g_off, b_off, rm_off, rv_off, nvalid, active, K, C, B, H, W, R, relu, momentum, eps, stream);
}
// out_s2d (nullable): the activation is ALSO written in space-to-depth form ([K][4C/8][R_s2d][8] on the (H/2, W/2)
// grid, parity plane p = 2*(h&1) + (w&1)), the operand form of the next stage's stride-2 convolutions.
// out_split / out_s2d_split (nullable, tf32 mode only): bf16 hi/lo split copies of out / out_s2d, the weight-gradient
// kernel's operand form ([K][2][C/8][R][8], see store_split).
int dba_bn_apply_split(const void* y, const void* res, void* out, void* out_s2d, long long R_s2d, void* out_split,
                       void* out_s2d_split, float* stats, long long ld_stats, int stats_off, float* saved,
                       long long ld_saved, int saved_off, float* Wflat, long long ldw, int g_off, int b_off, int rm_off,
                       int rv_off, const int* nvalid, const unsigned char* active, int K, int C, int B, int H, int W,
                       long long R, int relu, float momentum, float eps, void* stream) {
  DBA_REQUIRE(K > 0 && C % 8 == 0 && B > 0, "dba_bn_apply: K=%d C=%d B=%d", K, C, B);
  DBA_REQUIRE((!out_split && !out_s2d_split) || g_dba_act_mode == 1, "dba_bn_apply: split copies exist in tf32 mode only");
  DBA_REQUIRE(!out_s2d || (H % 2 == 0 && W % 2 == 0), "dba_bn_apply: space-to-depth output needs even H, W");
  BnArgs a{};
  a.out2 = (uint4*)out_s2d; a.R2 = R_s2d;
  a.out_split = (uint2*)out_split; a.out2_split = (uint2*)out_s2d_split;
  a.y = (const uint4*)y; a.res = (const uint4*)res; a.out = (uint4*)out;
  a.stats = stats; a.saved = saved; a.Wflat = Wflat; a.nvalid = nvalid; a.active = active;
  a.ld_stats = ld_stats; a.ld_saved = ld_saved; a.ldw = ldw; a.R = R;
  a.stats_off = stats_off; a.saved_off = saved_off; a.g_off = g_off; a.b_off = b_off; a.rm_off = rm_off; a.rv_off = rv_off;
  a.K = K; a.C = C; a.B = B; a.H = H; a.W = W; a.relu = relu; a.momentum = momentum; a.eps = eps;
  const dim3 grid = bn_grid(B, H, W, C, K);
  const cudaStream_t st = (cudaStream_t)stream;
  if (act_v() == 4 && g_bn_pair && a.out_split && (a.out2 == nullptr || a.out2_split)) {
    // 8-channel planes per thread: full 16-byte stores of the split copies
    dim3 gp(grid.x, C / 8, K);
    if (bn_small(gp)) launch_pdl(bn_apply_pair_kernel<4>, gp, 256, 0, st, a);
    else launch_pdl(bn_apply_pair_kernel<1>, gp, 256, 0, st, a);
  } else if (act_v() == 4) {
    if (bn_small(grid)) launch_pdl(bn_apply_kernel<4, 4>, grid, 256, 0, st, a);
    else launch_pdl(bn_apply_kernel<4, 1>, grid, 256, 0, st, a);
  } else {
    if (bn_small(grid)) launch_pdl(bn_apply_kernel<8, 4>, grid, 256, 0, st, a);
    else launch_pdl(bn_apply_kernel<8, 1>, grid, 256, 0, st, a);
  }
  DBA_COUNT_LAUNCH(1);
  DBA_CHECK_LAUNCH();
  return DBA_OK;
}

// Backward of the above.  g: upstream gradient (geometry of y); act: this BN's post-ReLU output (null = no ReLU);
// dy: output, written on the (2H, 2W) grid at even positions when up != 0 (Rup = rows of that grid);
// gmask: optional copy of the ReLU-masked upstream gradient; sums: scratch pairs [K][ld_sums] zeroed by the caller;
// Gflat: d_gamma / d_beta are stored at g_off / b_off of the gradient rows.
int dba_bn_backward_split(const void* g, const void* y, const void* act, void* dy, void* dy_split, void* gmask,
                          float* sums, long long ld_sums, int sums_off, float* saved, long long ld_saved, int saved_off,
                          float* Wflat, long long ldw, float* Gflat, long long ldg, int g_off, int b_off,
                          const int* nvalid, int K, int C, int B, int H, int W, long long R, int up, long long Rup,
                          void* stream);
int dba_bn_backward(const void* g, const void* y, const void* act, void* dy, void* gmask, float* sums,
                    long long ld_sums, int sums_off, float* saved, long long ld_saved, int saved_off, float* Wflat,
                    long long ldw, float* Gflat, long long ldg, int g_off, int b_off, const int* nvalid, int K, int C,
                    int B, int H, int W, long long R, int up, long long Rup, void* stream) {
  return dba_bn_backward_split(g, y, act, dy, nullptr, gmask, sums, ld_sums, sums_off, saved, ld_saved, saved_off, Wflat,
                               ldw, Gflat, ldg, g_off, b_off, nvalid, K, C, B, H, W, R, up, Rup, stream);
}
// dy_split (nullable, tf32 mode only): bf16 hi/lo split copy of dy for the weight-gradient kernel.
// sums_ready != 0: the sums were accumulated by the convolution that produced g (dba_conv_tc_dgrad_bn): apply only.
int dba_bn_backward_ex(const void* g, const void* y, const void* act, void* dy, void* dy_split, void* gmask,
                       float* sums, long long ld_sums, int sums_off, float* saved, long long ld_saved, int saved_off,
                       float* Wflat, long long ldw, float* Gflat, long long ldg, int g_off, int b_off,
                       const int* nvalid, int K, int C, int B, int H, int W, long long R, int up, long long Rup,
                       int sums_ready, void* stream) {
  DBA_REQUIRE(K > 0 && C % 8 == 0 && B > 0, "dba_bn_backward: K=%d C=%d B=%d", K, C, B);
  DBA_REQUIRE(!dy_split || g_dba_act_mode == 1, "dba_bn_backward: split copies exist in tf32 mode only");
  BnArgs a{};
  a.g = (const uint4*)g; a.y = (const uint4*)y; a.act = (const uint4*)act;
  a.out = (uint4*)dy; a.gmask = (uint4*)gmask; a.stats = sums; a.out_split = (uint2*)dy_split; a.saved = saved; a.Wflat = Wflat;
  a.Gflat = Gflat; a.nvalid = nvalid;
  a.ld_stats = ld_sums; a.ld_saved = ld_saved; a.ldw = ldw; a.ldg = ldg; a.R = R; a.Rup = Rup;
  a.stats_off = sums_off; a.saved_off = saved_off; a.g_off = g_off; a.b_off = b_off;
  a.K = K; a.C = C; a.B = B; a.H = H; a.W = W; a.up = up;
  dim3 grid = bn_grid(B, H, W, C, K);
  if (g_bn_fused && !sums_ready) {
    // one cluster per (client, channel plane): as many row blocks as keep ~148 SMs busy, at most 8 (portable size)
    const long long pairs = (long long)(C / act_v()) * K;
    int X = (int)((2 * kNumSMs + pairs - 1) / pairs);
    if (X > (int)grid.x) X = (int)grid.x;
    X = X >= 8 ? 8 : X >= 4 ? 4 : X >= 2 ? 2 : 1;
    grid.x = X;
    if (act_v() == 4) launch_cluster_pdl(bn_bwd_fused_kernel<4>, grid, 256, X, (cudaStream_t)stream, a);
    else launch_cluster_pdl(bn_bwd_fused_kernel<8>, grid, 256, X, (cudaStream_t)stream, a);
    DBA_COUNT_LAUNCH(1);
    DBA_CHECK_LAUNCH();
    return DBA_OK;
  }
  const cudaStream_t st = (cudaStream_t)stream;
  const bool small = bn_small(grid);
  if (act_v() == 4) {
    if (!sums_ready) {
      if (small) launch_pdl(bn_bwd_reduce_kernel<4, 4>, grid, 256, 0, st, a);
      else launch_pdl(bn_bwd_reduce_kernel<4, 1>, grid, 256, 0, st, a);
    }
    if (g_bn_pair && a.out_split) {       // 8-channel planes per thread: 16-byte stores of the split copy of dy
      const dim3 gp(grid.x, C / 8, K);
      if (bn_small(gp)) launch_pdl(bn_bwd_apply_pair_kernel<2>, gp, 256, 0, st, a);   // <4> needs 176 registers
      else launch_pdl(bn_bwd_apply_pair_kernel<1>, gp, 256, 0, st, a);
    } else if (small) {
      launch_pdl(bn_bwd_apply_kernel<4, 4>, grid, 256, 0, st, a);
    } else {
      launch_pdl(bn_bwd_apply_kernel<4, 1>, grid, 256, 0, st, a);
    }
  } else {
    if (!sums_ready) {
      if (small) launch_pdl(bn_bwd_reduce_kernel<8, 4>, grid, 256, 0, st, a);
      else launch_pdl(bn_bwd_reduce_kernel<8, 1>, grid, 256, 0, st, a);
    }
    if (small) launch_pdl(bn_bwd_apply_kernel<8, 4>, grid, 256, 0, st, a);
    else launch_pdl(bn_bwd_apply_kernel<8, 1>, grid, 256, 0, st, a);
  }
  DBA_COUNT_LAUNCH(sums_ready ? 1 : 2);
  DBA_CHECK_LAUNCH();
  return DBA_OK;
}
int dba_bn_backward_split(const void* g, const void* y, const void* act, void* dy, void* dy_split, void* gmask,
                          float* sums, long long ld_sums, int sums_off, float* saved, long long ld_saved, int saved_off,
                          float* Wflat, long long ldw, float* Gflat, long long ldg, int g_off, int b_off,
                          const int* nvalid, int K, int C, int B, int H, int W, long long R, int up, long long Rup,
                          void* stream) {
  return dba_bn_backward_ex(g, y, act, dy, dy_split, gmask, sums, ld_sums, sums_off, saved, ld_saved, saved_off, Wflat, ldw,
                            Gflat, ldg, g_off, b_off, nvalid, K, C, B, H, W, R, up, Rup, 0, stream);
}

// 1: BatchNorm backward = one cluster launch per layer; 0 (default): the two-kernel form (reduce with atomics into the
// caller-zeroed `sums` scratch, then apply).  Both produce the same dy; the fused form needs no zeroed scratch and is
// deterministic, but clusters of <= 8 CTAs walk long row slices twice and lose to 64 row blocks x 2 launches.
void dba_set_bn_pair(int on) { g_bn_pair = on ? 1 : 0; }
void dba_set_bn_fused(int on) { g_bn_fused = on ? 1 : 0; }
int dba_get_bn_fused() { return g_bn_fused; }

// y = MaxPool2d(3, 2, 1)(x); x on the B x H x W grid, y on B x H/2 x W/2.  Rin / Rout = dba_conv_rows of the two grids.
int dba_maxpool_fwd_split(const void* x, void* y, void* y_split, int K, int C, int B, int H, int W, long long Rin,
                          long long Rout, void* stream);
int dba_maxpool_fwd(const void* x, void* y, int K, int C, int B, int H, int W, long long Rin, long long Rout,
                    void* stream) {
  return dba_maxpool_fwd_split(x, y, nullptr, K, C, B, H, W, Rin, Rout, stream);
}
int dba_maxpool_fwd_split(const void* x, void* y, void* y_split, int K, int C, int B, int H, int W, long long Rin,
                          long long Rout, void* stream) {
  DBA_REQUIRE(K > 0 && C % 8 == 0 && H % 2 == 0 && W % 2 == 0, "dba_maxpool_fwd: K=%d C=%d H=%d W=%d", K, C, H, W);
  DBA_REQUIRE(!y_split || g_dba_act_mode == 1, "dba_maxpool_fwd: split copies exist in tf32 mode only");
  PoolArgs a{(const uint4*)x, (uint4*)y, nullptr, nullptr, (uint2*)y_split, Rin, Rout, K, C, B, H, W};
  int rb = (B * (H / 2) * (W / 2) + 1023) / 1024;
  if (rb > 64) rb = 64;
  const dim3 grid(rb < 1 ? 1 : rb, C / act_v(), K);
  if (act_v() == 4) maxpool_fwd_kernel<4><<<grid, 256, 0, (cudaStream_t)stream>>>(a);
  else maxpool_fwd_kernel<8><<<grid, 256, 0, (cudaStream_t)stream>>>(a);
  DBA_COUNT_LAUNCH(1);
  DBA_CHECK_LAUNCH();
  return DBA_OK;
}
// gx = d loss / d x given gy = d loss / d y (x is the forward input).
int dba_maxpool_bwd(const void* x, const void* gy, void* gx, int K, int C, int B, int H, int W, long long Rin,
                    long long Rout, void* stream) {
  DBA_REQUIRE(K > 0 && C % 8 == 0 && H % 2 == 0 && W % 2 == 0, "dba_maxpool_bwd: K=%d C=%d H=%d W=%d", K, C, H, W);
  PoolArgs a{(const uint4*)x, nullptr, (const uint4*)gy, (uint4*)gx, nullptr, Rin, Rout, K, C, B, H, W};
  int rb = (B * H * W + 1023) / 1024;
  if (rb > 64) rb = 64;
  const dim3 grid(rb < 1 ? 1 : rb, C / act_v(), K);
  if (act_v() == 4) maxpool_bwd_kernel<4><<<grid, 256, 0, (cudaStream_t)stream>>>(a);
  else maxpool_bwd_kernel<8><<<grid, 256, 0, (cudaStream_t)stream>>>(a);
  DBA_COUNT_LAUNCH(1);
  DBA_CHECK_LAUNCH();
  return DBA_OK;
}

}  // extern "C"
