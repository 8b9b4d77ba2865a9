// Network plumbing kernels around the tensor-core convolution: weight packing, BatchNorm folding for evaluation,
// the fused input stage (gather + /255 + trigger stamping + flip + stem im2col) and the fused classifier head
// (avg-pool + linear + cross-entropy + metrics, forward and backward).
//
// Reference ops: image batches + triggers image_helper.py:289-350 (K12); F.avg_pool2d / Linear / cross_entropy /
// argmax accuracy resnet_cifar.py:94-96, image_train.py:85,104-105, test.py:20-24 (K5, K6, K8).
#include "common.cuh"

using namespace dba;

// ------------------------------------------------------------------------------------------- weight packing
// Table entry per convolution (all ints): [0] src offset in the flat parameter vector (weights stored
// [Cout][taps][Cin]), [1] Cout, [2] Cin, [3] taps, [4] dst offset (elements) of the forward pack
// [taps'][CinP/8][CoutP][8], [5] dst offset of the transposed pack [taps][CoutP/8][CinP][8] with flipped taps (-1 =
// none), [6] CinP, [7] CoutP, [8] mode: 1 = fold, treat (taps, Cin) as one K axis of length taps*Cin padded to CinP
// (stem); 2 = linear layer over a flattened [Cin][taps] feature map (torch's x.view(-1, C*h*w): source index
// co*(Cin*taps) + ci*taps + tap), packed as a 1x1 convolution over channels n = tap*CinP + ci - forward
// [1][taps*CinP/8][CoutP][8] (which IS the [taps][CinP/8][CoutP][8] layout), transposed [1][CoutP/8][taps*CinP][8].
// 3 = 3x3 stride-2 convolution whose input gradient is computed in depth-to-space form (dba_conv_tc_ex out mode 3):
// besides the forward pack, the transposed slot receives [4 taps (a,b)][CoutP/8][4*CinP][8] with
// Wd[(a,b)][co][(2*ph+pw)*CinP + ci] = W[dh(ph,a)][dw(pw,b)][co][ci], dh(0,0) = 1, dh(1,0) = 2, dh(1,1) = 0, none for (0,1).
#define PACK_FIELDS 9
template <int V>
__global__ void __launch_bounds__(256) pack_weights_kernel(const float* __restrict__ W, long long ldw,
                                                           const int* __restrict__ table,
                                                           uint4* __restrict__ dst, long long dst_client_stride) {
  // dst is addressed in 16-byte chunks of V values (V = 8: bf16, V = 4: fp32); table offsets are in elements
  pdl_trigger();
  pdl_wait();
  const int* e = table + blockIdx.y * PACK_FIELDS;
  const int k = blockIdx.z;
  const int Cout = e[1], Cin = e[2], taps = e[3], CinP = e[6], CoutP = e[7], fold = e[8] == 1, chw = e[8] == 2;
  const float* src = W + (long long)k * ldw + e[0];
  uint4* d = dst + (long long)k * dst_client_stride;
  const int ptaps = fold ? 1 : taps;             // taps as seen by the kernel
  const int kreal = fold ? taps * Cin : Cin;     // real reduction length per packed tap
  const int cinV = CinP / V, coutV = CoutP / V;
  // forward pack.  Plain [Cout][taps][Cin] sources go through a 32 x 32 shared-memory tile: rows of 32 consecutive input
  // channels are read coalesced (one 128-byte row per warp load), 16-byte chunks are stored with the output channel
  // running fastest (512 contiguous bytes per warp store).  A thread-per-chunk gather (kept for the fc1 'chw' order)
  // read 16 bytes per thread at a 9 KB stride: 190 us per step for ten clients.
  if (!chw) {
    __shared__ float tile[32][33];
    const int tco = (CoutP + 31) / 32, tci = (CinP + 31) / 32;
    const int ntiles = ptaps * tco * tci;
    const int lane = threadIdx.x & 31, wrp = threadIdx.x >> 5;
    for (int tl = blockIdx.x; tl < ntiles; tl += gridDim.x) {
      const int ci0 = (tl % tci) * 32, co0 = ((tl / tci) % tco) * 32, tap = tl / (tci * tco);
      for (int r = wrp; r < 32; r += 8) {
        const int co = co0 + r, ci = ci0 + lane;
        tile[r][lane] = (co < Cout && ci < kreal) ? src[((long long)co * ptaps + tap) * kreal + ci] : 0.f;
      }
      __syncthreads();
      const int ch = threadIdx.x >> 5;                  // chunk of V input channels inside the tile
      if (ch < 32 / V && ci0 + ch * V < CinP && co0 + lane < CoutP) {
        float v[V];
#pragma unroll
        for (int j = 0; j < V; ++j) v[j] = tile[lane][ch * V + j];
        d[e[4] / V + ((long long)tap * cinV + (ci0 / V + ch)) * CoutP + co0 + lane] = chunk_pack_operand<V>(v);
      }
      __syncthreads();
    }
  } else {
    const long long nf = (long long)ptaps * cinV * CoutP;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < nf; i += (long long)gridDim.x * blockDim.x) {
      const int co = (int)(i % CoutP);
      const int cv = (int)((i / CoutP) % cinV);
      const int tap = (int)(i / ((long long)cinV * CoutP));
      float v[V];
#pragma unroll
      for (int j = 0; j < V; ++j) {
        const int ci = cv * V + j;
        v[j] = (co < Cout && ci < kreal) ? src[(long long)co * (Cin * taps) + ci * taps + tap] : 0.f;
      }
      d[e[4] / V + ((long long)tap * cinV + cv) * CoutP + co] = chunk_pack_operand<V>(v);
    }
  }
  if (e[5] < 0 || fold) return;
  if (e[8] == 3) {
    const long long nd = 4LL * coutV * 4 * CinP;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < nd; i += (long long)gridDim.x * blockDim.x) {
      const int n = (int)(i % (4 * CinP));                 // output channel of the gradient conv: parity * CinP + ci
      const int ov = (int)((i / (4 * CinP)) % coutV);
      const int t2 = (int)(i / ((long long)4 * CinP * coutV));      // 2x2 tap (a, b)
      const int par = n / CinP, ci = n - par * CinP;
      const int ph = par >> 1, pw = par & 1, ta = t2 >> 1, tb = t2 & 1;
      const int dh = ph == 0 ? (ta == 0 ? 1 : -1) : (ta == 0 ? 2 : 0);
      const int dw = pw == 0 ? (tb == 0 ? 1 : -1) : (tb == 0 ? 2 : 0);
      float v[V];
#pragma unroll
      for (int j = 0; j < V; ++j) {
        const int co = ov * V + j;
        v[j] = (dh >= 0 && dw >= 0 && co < Cout && ci < Cin) ? src[((long long)co * taps + dh * 3 + dw) * Cin + ci] : 0.f;
      }
      d[e[5] / V + ((long long)t2 * coutV + ov) * (4 * CinP) + n] = chunk_pack_operand<V>(v);
    }
    return;
  }
  // transposed pack for the input-gradient convolution: out channels become the reduction axis, taps are flipped
  const long long nt = (long long)taps * coutV * CinP;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < nt; i += (long long)gridDim.x * blockDim.x) {
    const int ci = (int)(i % CinP);
    const int ov = (int)((i / CinP) % coutV);
    const int tap = (int)(i / ((long long)CinP * coutV));
    float v[V];
#pragma unroll
    for (int j = 0; j < V; ++j) {
      const int co = ov * V + j;
      v[j] = (co < Cout && ci < Cin)
                 ? (chw ? src[(long long)co * (Cin * taps) + ci * taps + tap] : src[((long long)co * taps + tap) * Cin + ci])
                 : 0.f;
    }
    const uint4 pk = chunk_pack_operand<V>(v);
    if (chw) {
      d[e[5] / V + ((long long)ov * taps + tap) * CinP + ci] = pk;
    } else {
      const int ftap = taps - 1 - tap;
      d[e[5] / V + ((long long)ftap * coutV + ov) * CinP + ci] = pk;
    }
  }
}
extern "C" int dba_pack_weights(const float* W, long long ldw, const int* table, int n_layers, int K, void* dst,
                                long long dst_client_stride, void* stream) {
  DBA_REQUIRE(n_layers > 0 && K > 0, "dba_pack_weights: n_layers=%d K=%d", n_layers, K);
  dim3 g(K >= 8 ? 48 : 128, n_layers, K);     // few clients: the widest layers (576 tiles / 147k chunks) need the blocks
  // dst_client_stride counts elements (bf16 or fp32) of the caller's pack tensor
  if (g_dba_act_mode == 1) {
    DBA_REQUIRE(dst_client_stride % 4 == 0, "dba_pack_weights: client stride %lld %% 4", dst_client_stride);
    launch_pdl(pack_weights_kernel<4>, g, 256, 0, (cudaStream_t)stream, W, ldw, table, (uint4*)dst, dst_client_stride / 4);
  } else {
    DBA_REQUIRE(dst_client_stride % 8 == 0, "dba_pack_weights: client stride %lld %% 8", dst_client_stride);
    launch_pdl(pack_weights_kernel<8>, g, 256, 0, (cudaStream_t)stream, W, ldw, table, (uint4*)dst, dst_client_stride / 8);
  }
  DBA_COUNT_LAUNCH(1);
  DBA_CHECK_LAUNCH();
  return DBA_OK;
}

// ------------------------------------------------------------------------------- BatchNorm folding (evaluation)
// Table entry per BN layer: [0] gamma offset, [1] beta offset, [2] running-mean offset, [3] running-var offset (all
// into the flat state row), [4] C, [5] dst offset into scale/shift rows.  scale = g/sqrt(rv+eps), shift = b - rm*scale.
#define FOLD_FIELDS 6
__global__ void bn_fold_kernel(const float* __restrict__ W, long long ldw, const int* __restrict__ table,
                               float* __restrict__ scale, float* __restrict__ shift, long long ld_out, float eps) {
  const int* e = table + blockIdx.y * FOLD_FIELDS;
  const int k = blockIdx.z;
  const float* w = W + (long long)k * ldw;
  for (int c = blockIdx.x * blockDim.x + threadIdx.x; c < e[4]; c += gridDim.x * blockDim.x) {
    const float s = w[e[0] + c] * (1.f / sqrtf(w[e[3] + c] + eps));
    scale[(long long)k * ld_out + e[5] + c] = s;
    shift[(long long)k * ld_out + e[5] + c] = w[e[1] + c] - w[e[2] + c] * s;
  }
}
extern "C" int dba_bn_fold(const float* W, long long ldw, const int* table, int n_layers, int K, float* scale,
                           float* shift, long long ld_out, float eps, void* stream) {
  DBA_REQUIRE(n_layers > 0 && K > 0, "dba_bn_fold: n_layers=%d K=%d", n_layers, K);
  dim3 g(2, n_layers, K);
  bn_fold_kernel<<<g, 256, 0, (cudaStream_t)stream>>>(W, ldw, table, scale, shift, ld_out, eps);
  DBA_COUNT_LAUNCH(1);
  DBA_CHECK_LAUNCH();
  return DBA_OK;
}

// ------------------------------------------------------------------------------------------------ input stage
// Builds the stem convolution's A operand straight from the uint8 dataset: for every (client, sample, output pixel)
// the ks x ks x C input patch (zero padded, /255, trigger pixels forced to 1.0 for poisoned samples, optional
// horizontal flip of the source image) is written as KP = roundup(ks*ks*C, 16) bf16 values in planar-8 layout, so the
// stem becomes a 1x1 convolution on the tensor cores.  Also emits the (possibly swapped) labels.
// idx [K][B] dataset indices (-1 = padding sample -> zero rows, label -1).  Samples b < poison_n[k] are poisoned with
// trigger mask row trig_row[k].  With shared != 0 there is one input set (K = 1 here) used by all models.
struct InputArgs {
  const unsigned char* data;     // [N][Hi][Wi][C]
  const long long* labels;       // [N]
  const long long* idx;          // [K][ld_idx] (+ t*B offset applied by the caller through idx pointer)
  long long ld_idx;
  const unsigned char* trig;     // [T][Hi][Wi]
  const long long* trig_row;     // [K]
  const long long* poison_n;     // [K]
  const unsigned char* flip;     // [K][ld_idx] or null
  uint4* out;                    // [K][KP/V][R] 16-byte rows of V values (V = 8: bf16, 4: fp32)
  uint2* out_split;              // tf32 mode, optional: bf16 hi/lo split copy ([K][2][KP/8][R][8]) for the stem's wgrad
  int* labels_out;               // [K][B]
  int K, B, Hi, Wi, C, ks, stride, pad, Ho, Wo, KP, label_swap;
  long long R;
};
// grid: x = 256-pixel blocks of one image, y = sample, z = client * chunks + chunk.  Everything that depends on
// (client, sample, chunk) - the sample id, poison / flip flags, the (kh, kw, c) of the chunk's 8 im2col columns - is
// block-uniform, so the per-thread work is one 32-bit division, 8 byte loads and one 16-byte store (the first version
// decomposed a flat 64-bit index per thread and spent its time in integer division: 0.5 TB/s of stores).
template <int V>
__global__ void __launch_bounds__(256) input_stage_kernel(const InputArgs a) {
  pdl_trigger();
  pdl_wait();
  const int Wp = a.Wo + 1, S = (a.Ho + 1) * Wp, lead = Wp + 1;
  const int nchunk = a.KP / V;
  const int b = blockIdx.y;
  const int k = blockIdx.z / nchunk, chunk = blockIdx.z - k * nchunk;
  __shared__ int s_dh[V], s_dw[V], s_c[V];
  if (threadIdx.x < V) {
    const int e = chunk * V + threadIdx.x;
    int c = -1, dh = 0, dw = 0;
    if (e < a.ks * a.ks * a.C) {
      const int tap = e / a.C;
      c = e - tap * a.C;
      const int kh = tap / a.ks;
      dh = kh - a.pad;
      dw = tap - kh * a.ks - a.pad;
    }
    s_dh[threadIdx.x] = dh; s_dw[threadIdx.x] = dw; s_c[threadIdx.x] = c;
  }
  __syncthreads();
  const long long id = a.idx[(long long)k * a.ld_idx + b];
  const bool poisoned = id >= 0 && b < a.poison_n[k];
  const bool flip = a.flip && a.flip[(long long)k * a.ld_idx + b];
  const unsigned char* img = a.data + (id < 0 ? 0 : id) * (long long)a.Hi * a.Wi * a.C;
  const unsigned char* tm = a.trig + a.trig_row[k] * (long long)a.Hi * a.Wi;
  uint4* dst = a.out + (((long long)k * nchunk + chunk) * a.R + lead + (long long)b * S);
  const int npix = a.Ho * a.Wo;
  for (int pix = blockIdx.x * blockDim.x + threadIdx.x; pix < npix; pix += gridDim.x * blockDim.x) {
    const int ho = pix / a.Wo, wo = pix - ho * a.Wo;
    float v[V];
#pragma unroll
    for (int j = 0; j < V; ++j) {
      float x = 0.f;
      const int c = s_c[j];
      if (id >= 0 && c >= 0) {
        const int hi = ho * a.stride + s_dh[j], wi = wo * a.stride + s_dw[j];
        if (hi >= 0 && hi < a.Hi && wi >= 0 && wi < a.Wi) {
          if (poisoned && tm[hi * a.Wi + wi]) x = 1.f;
          else x = (float)img[(hi * a.Wi + (flip ? a.Wi - 1 - wi : wi)) * a.C + c] * (1.f / 255.f);
        }
      }
      v[j] = x;
    }
    dst[ho * Wp + wo] = chunk_pack_operand<V>(v);
    if constexpr (V == 4) {
      if (a.out_split) store_split(a.out_split, k, a.KP >> 3, a.R, chunk, lead + (long long)b * S + ho * Wp + wo, v);
    }
  }
  if (a.labels_out && chunk == 0 && blockIdx.x == 0 && threadIdx.x == 0)
    a.labels_out[k * a.B + b] = id < 0 ? -1 : (poisoned ? a.label_swap : (int)a.labels[id]);
}
extern "C" int dba_input_stage_image_split(const void* data, const long long* labels, const long long* idx,
                                           long long ld_idx, const void* trig, const long long* trig_row,
                                           const long long* poison_n, const void* flip, void* out, void* out_split,
                                           int* labels_out, int K, int B, int Hi, int Wi, int C, int ks, int stride,
                                           int pad, int Ho, int Wo, int KP, int label_swap, long long R, void* stream);
extern "C" int dba_input_stage_image(const void* data, const long long* labels, const long long* idx, long long ld_idx,
                                     const void* trig, const long long* trig_row, const long long* poison_n,
                                     const void* flip, void* out, int* labels_out, int K, int B, int Hi, int Wi, int C,
                                     int ks, int stride, int pad, int Ho, int Wo, int KP, int label_swap, long long R,
                                     void* stream) {
  return dba_input_stage_image_split(data, labels, idx, ld_idx, trig, trig_row, poison_n, flip, out, nullptr, labels_out,
                                     K, B, Hi, Wi, C, ks, stride, pad, Ho, Wo, KP, label_swap, R, stream);
}
// out_split (nullable, tf32 mode only): bf16 hi/lo split copy of the im2col'd input, the stem weight gradient's operand.
extern "C" int dba_input_stage_image_split(const void* data, const long long* labels, const long long* idx,
                                           long long ld_idx, const void* trig, const long long* trig_row,
                                           const long long* poison_n, const void* flip, void* out, void* out_split,
                                           int* labels_out, int K, int B, int Hi, int Wi, int C, int ks, int stride,
                                           int pad, int Ho, int Wo, int KP, int label_swap, long long R, void* stream) {
  DBA_REQUIRE(!out_split || g_dba_act_mode == 1, "dba_input_stage_image: split copies exist in tf32 mode only");
  DBA_REQUIRE(K > 0 && B > 0 && KP % 16 == 0 && KP >= ks * ks * C, "dba_input_stage_image: K=%d B=%d KP=%d", K, B, KP);
  InputArgs a{(const unsigned char*)data, labels, idx, ld_idx, (const unsigned char*)trig, trig_row, poison_n,
              (const unsigned char*)flip, (uint4*)out, (uint2*)out_split, labels_out, K, B, Hi, Wi, C, ks, stride, pad,
              Ho, Wo, KP, label_swap, R};
  const int V = g_dba_act_mode == 1 ? 4 : 8;
  DBA_REQUIRE(B <= 65535 && (long long)K * (KP / V) <= 65535, "dba_input_stage_image: grid too large (B=%d K=%d)", B, K);
  int gx = (Ho * Wo + 255) / 256;
  if (gx > 64) gx = 64;
  if (V == 4) launch_pdl(input_stage_kernel<4>, dim3(gx, B, K * (KP / 4)), 256, 0, (cudaStream_t)stream, a);
  else launch_pdl(input_stage_kernel<8>, dim3(gx, B, K * (KP / 8)), 256, 0, (cudaStream_t)stream, a);
  DBA_COUNT_LAUNCH(1);
  DBA_CHECK_LAUNCH();
  return DBA_OK;
}

// ------------------------------------------------------------------------------------------- classifier head
// One warp per (client, sample): global average pool over the HxW interior rows of the planar activation, linear
// layer (fp32 master weights straight from the flat slab), softmax cross-entropy, argmax -> device-side metric
// accumulators (no host sync per batch).  In training mode also the backward pass of all three: gradient of the
// activation (planar bf16, scaled by 1/(n_valid*H*W)), and fp32 atomics into the flat gradient slab for the linear
// weight / bias.  Metric slots: metrics_f[k][slot] += loss, metrics_i[k][slot][0] += correct, [1] += count.
struct HeadArgs {
  const uint4* act;          // [K][Cp/V][R] 16-byte rows of V channels (V = 8: bf16, 4: fp32)
  const float* Wflat;        // [K][ldw]
  float* Gflat;              // [K][ldg] or null (evaluation)
  uint4* dact;               // same layout, or null
  const int* labels;         // [Kl][B]   (Kl = 1 when labels are shared by all models)
  const int* nvalid;         // [K] or null (=> every sample with label >= 0 counts, loss is a SUM)
  const long long* slot;     // [K] metric column (internal epoch) or null -> column 0
  float* metrics_f;          // [K][ld_m]
  int* metrics_i;            // [K][ld_m][2]
  unsigned char* nan_flag;   // [K] or null
  long long ldw, ldg, R;
  int w_off, b_off, K, B, C, H, W, ncls, ld_m, shared_labels;
  int Cp;                    // channels of the activation buffers (C rounded up to the planar layout's padding)
};
// Block = kHeadWarps samples of ONE client (grid.x = blocks per client, grid.y = client), so the gradient of the
// linear layer is first summed in shared memory and leaves the block as one atomic per weight (the first version
// issued ncls*C global atomics per sample on the same addresses: 68 us for 64 samples).
constexpr int kHeadWarps = 4;       // 16 blocks per client at B = 64: the head is latency-bound, not atomics-bound
template <int MAXC, int MAXCLS, bool SMEM_GRAD, int V>
__global__ void __launch_bounds__(kHeadWarps * 32) head_kernel(const HeadArgs a) {
  extern __shared__ float s_gw[];               // [ncls][C] + [ncls]  (SMEM_GRAD && training only)
  pdl_trigger();
  pdl_wait();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int k = blockIdx.y, b = blockIdx.x * kHeadWarps + warp;
  const bool train = a.Gflat != nullptr;
  if (SMEM_GRAD && train) {
    for (int i = threadIdx.x; i < a.ncls * a.C + a.ncls; i += blockDim.x) s_gw[i] = 0.f;
    __syncthreads();
  }
  const int label = b < a.B ? a.labels[(a.shared_labels ? 0 : k) * a.B + b] : -1;
  const int Wp = a.W + 1, S = (a.H + 1) * Wp, lead = Wp + 1;
  const long long plane = a.R;                   // rows per channel-chunk plane
  const uint4* act = a.act + (long long)k * (a.Cp / V) * plane;
  constexpr int PER = MAXC / 32;
  const float inv_hw = 1.f / (float)(a.H * a.W);
  if (b < a.B && label < 0 && a.dact) {        // padding sample: contributes nothing, its gradient rows must be zero
    uint4* d = a.dact + (long long)k * (a.Cp / V) * plane;
    for (int i = 0; i < PER; ++i) {
      const int c = lane + 32 * i;
      if (c < a.C)
        for (int h = 0; h < a.H; ++h)
          for (int w = 0; w < a.W; ++w)
            elem_store<V>(d, (long long)(c / V) * plane + (long long)lead + (long long)b * S + h * Wp + w, c % V, 0.f);
    }
  }
  if (label >= 0) {
    float pooled[PER];
#pragma unroll
    for (int i = 0; i < PER; ++i) {
      const int c = lane + 32 * i;
      float sum = 0.f;
      if (c < a.C) {
        const uint4* p = act + (long long)(c / V) * plane;
        for (int h = 0; h < a.H; ++h)
          for (int w = 0; w < a.W; ++w)
            sum += elem_load<V>(p, (long long)lead + (long long)b * S + h * Wp + w, c % V);
      }
      pooled[i] = sum * inv_hw;
    }
    const float* Wl = a.Wflat + (long long)k * a.ldw + a.w_off;
    const float* bl = a.Wflat + (long long)k * a.ldw + a.b_off;
    float logit[MAXCLS];           // every lane ends up holding all logits
    float mx = -INFINITY;
    int arg = 0;
    for (int j = 0; j < a.ncls; ++j) {
      float part = 0.f;
#pragma unroll
      for (int i = 0; i < PER; ++i) {
        const int c = lane + 32 * i;
        if (c < a.C) part += Wl[(long long)j * a.C + c] * pooled[i];
      }
      part = warp_sum(part) + bl[j];
      logit[j] = part;
      if (part > mx) { mx = part; arg = j; }
    }
    float se = 0.f;
    for (int j = 0; j < a.ncls; ++j) se += expf(logit[j] - mx);
    const float lse = mx + logf(se);
    const float loss = lse - logit[label];
    const int slot = a.slot ? (int)a.slot[k] : 0;
    const bool finite = isfinite(lse);
    if (lane == 0) {
      const float scale = a.nvalid ? 1.f / (float)max(a.nvalid[k], 1) : 1.f;   // training logs the batch-MEAN loss (Q6)
      atomicAdd(&a.metrics_f[(long long)k * a.ld_m + slot], loss * scale);
      atomicAdd(&a.metrics_i[((long long)k * a.ld_m + slot) * 2], arg == label ? 1 : 0);
      atomicAdd(&a.metrics_i[((long long)k * a.ld_m + slot) * 2 + 1], 1);
      if (a.nan_flag && !finite) a.nan_flag[k] = 1;
    }
    if (train) {
      // ---- backward
      const float inv_n = 1.f / (float)max(a.nvalid ? a.nvalid[k] : 1, 1);
      float* Gw = a.Gflat + (long long)k * a.ldg + a.w_off;
      float* Gb = a.Gflat + (long long)k * a.ldg + a.b_off;
      float dpooled[PER];
#pragma unroll
      for (int i = 0; i < PER; ++i) dpooled[i] = 0.f;
      for (int j = 0; j < a.ncls; ++j) {
        const float dl = (expf(logit[j] - lse) - (j == label ? 1.f : 0.f)) * inv_n;
#pragma unroll
        for (int i = 0; i < PER; ++i) {
          const int c = lane + 32 * i;
          if (c < a.C) {
            dpooled[i] += dl * Wl[(long long)j * a.C + c];
            if (SMEM_GRAD) atomicAdd(&s_gw[j * a.C + c], dl * pooled[i]);
            else atomicAdd(&Gw[(long long)j * a.C + c], dl * pooled[i]);
          }
        }
        if (lane == 0) {
          if (SMEM_GRAD) atomicAdd(&s_gw[a.ncls * a.C + j], dl);
          else atomicAdd(&Gb[j], dl);
        }
      }
      if (a.dact) {
        uint4* d = a.dact + (long long)k * (a.Cp / V) * plane;
#pragma unroll
        for (int i = 0; i < PER; ++i) {
          const int c = lane + 32 * i;
          if (c < a.C) {
            const float g = dpooled[i] * inv_hw;
            for (int h = 0; h < a.H; ++h)
              for (int w = 0; w < a.W; ++w)
                elem_store<V>(d, (long long)(c / V) * plane + (long long)lead + (long long)b * S + h * Wp + w, c % V, g);
          }
        }
      }
    }
  }
  if (SMEM_GRAD && train) {
    __syncthreads();
    float* Gw = a.Gflat + (long long)k * a.ldg + a.w_off;
    float* Gb = a.Gflat + (long long)k * a.ldg + a.b_off;
    for (int i = threadIdx.x; i < a.ncls * a.C; i += blockDim.x)
      if (s_gw[i] != 0.f) atomicAdd(&Gw[i], s_gw[i]);
    for (int i = threadIdx.x; i < a.ncls; i += blockDim.x)
      if (s_gw[a.ncls * a.C + i] != 0.f) atomicAdd(&Gb[i], s_gw[a.ncls * a.C + i]);
  }
}
extern "C" int dba_head_ex(const void* act, const float* Wflat, long long ldw, int w_off, int b_off, float* Gflat,
                           long long ldg, void* dact, const int* labels, int shared_labels, const int* nvalid,
                           const long long* slot, float* metrics_f, int* metrics_i, int ld_m, unsigned char* nan_flag,
                           int K, int B, int C, int c_pad, int H, int W, int ncls, long long R, void* stream);
extern "C" int dba_head(const void* act, const float* Wflat, long long ldw, int w_off, int b_off, float* Gflat,
                        long long ldg, void* dact, const int* labels, int shared_labels, const int* nvalid,
                        const long long* slot, float* metrics_f, int* metrics_i, int ld_m, unsigned char* nan_flag,
                        int K, int B, int C, int H, int W, int ncls, long long R, void* stream) {
  return dba_head_ex(act, Wflat, ldw, w_off, b_off, Gflat, ldg, dact, labels, shared_labels, nvalid, slot, metrics_f,
                     metrics_i, ld_m, nan_flag, K, B, C, (C + 7) / 8 * 8, H, W, ncls, R, stream);
}
// c_pad: channel count of the planar act / dact buffers (>= C, multiple of 8) when the layer is padded (500 -> 512).
extern "C" int dba_head_ex(const void* act, const float* Wflat, long long ldw, int w_off, int b_off, float* Gflat,
                           long long ldg, void* dact, const int* labels, int shared_labels, const int* nvalid,
                           const long long* slot, float* metrics_f, int* metrics_i, int ld_m, unsigned char* nan_flag,
                           int K, int B, int C, int c_pad, int H, int W, int ncls, long long R, void* stream) {
  DBA_REQUIRE(K > 0 && B > 0 && C <= 512 && ncls <= 200, "dba_head: K=%d B=%d C=%d ncls=%d", K, B, C, ncls);
  HeadArgs a{(const uint4*)act, Wflat, Gflat, (uint4*)dact, labels, nvalid, slot, metrics_f, metrics_i,
             nan_flag, ldw, ldg, R, w_off, b_off, K, B, C, H, W, ncls, ld_m, shared_labels, c_pad};
  DBA_REQUIRE(c_pad >= C && c_pad % 8 == 0, "dba_head: c_pad=%d for C=%d", c_pad, C);
  DBA_REQUIRE(K <= 65535, "dba_head: K=%d too large for grid.y", K);
  const dim3 grid((B + kHeadWarps - 1) / kHeadWarps, K);
  const size_t gsm = (size_t)(ncls * C + ncls) * sizeof(float);
  const bool smem_grad = Gflat != nullptr && gsm <= 40 * 1024;
  const cudaStream_t st = (cudaStream_t)stream;
  const int nt = kHeadWarps * 32;
  if (g_dba_act_mode == 1) {
    if (C <= 256 && ncls <= 16) {
      if (smem_grad) launch_pdl(head_kernel<256, 16, true, 4>, grid, nt, gsm, st, a);
      else launch_pdl(head_kernel<256, 16, false, 4>, grid, nt, 0, st, a);
    } else {
      if (smem_grad) launch_pdl(head_kernel<512, 200, true, 4>, grid, nt, gsm, st, a);
      else launch_pdl(head_kernel<512, 200, false, 4>, grid, nt, 0, st, a);
    }
  } else {
    if (C <= 256 && ncls <= 16) {
      if (smem_grad) launch_pdl(head_kernel<256, 16, true, 8>, grid, nt, gsm, st, a);
      else launch_pdl(head_kernel<256, 16, false, 8>, grid, nt, 0, st, a);
    } else {
      if (smem_grad) launch_pdl(head_kernel<512, 200, true, 8>, grid, nt, gsm, st, a);
      else launch_pdl(head_kernel<512, 200, false, 8>, grid, nt, 0, st, a);
    }
  }
  DBA_COUNT_LAUNCH(1);
  DBA_CHECK_LAUNCH();
  return DBA_OK;
}
