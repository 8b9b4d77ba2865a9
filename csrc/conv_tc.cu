// Batched-client implicit-GEMM convolution on 5th-gen tensor cores (tcgen05.mma, accumulators in TMEM), sm_100a.
//
// Reference op: every nn.Conv2d forward / input-gradient of the four model families (K1/K2 in SURVEY §2.4;
// models/resnet_cifar.py:20-35,89, models/resnet_tinyimagenet.py:64-72,206), which the reference runs one client at
// a time through cuDNN.  Here ONE launch convolves the activations of all K client replicas (or of M models under
// evaluation) with their K different weight sets.
//
// Data layout ("planar-8, shared-halo"):  an activation tensor is  X[k][c8][row][8]  (bf16): channel chunks of 8
// are the slow index, and for one (client, chunk) plane all pixels of all B images form ONE linear array of 16-byte
// rows:   row(b,h,w) = lead + b*S + h*Wp + w,   Wp = W+1, S = (H+1)*Wp, lead = Wp+1.
// Column W of every line and line H of every image are zero and never written: they are the right/left and
// bottom/top halo of the neighbouring pixels at once.  Consequences:
//   * a 3x3 tap (dh,dw) of an output row r is simply input row r + dh*Wp + dw  -> the A operand of every tap is the
//     SAME shared-memory block, addressed through a UMMA descriptor whose start address is shifted by whole rows
//     (no im2col copy, no 9x re-read: one halo'd block per K-slice is loaded once with 1-D bulk async copies);
//   * [c8][row][8] is exactly the SWIZZLE_NONE canonical K-major UMMA layout (8x16B core matrices, SBO = 128 B,
//     LBO = plane pitch), and at the same time the canonical MN-major layout needed by the weight-gradient GEMM.
// Rows that are halo positions produce garbage accumulators that the epilogue simply does not store (6-36 % of
// the MMA work depending on image size) - the price for having no gather at all.
//
// Kernel structure (192 threads, 1 CTA/SM, persistent over work items (client, pass, n-block)):
//   warp 0   : producer - cp.async.bulk (UBLKCP) of A blocks [kc][T*128+halo][8] and weight blocks [kc][N][8]
//   warp 1   : one elected thread issues tcgen05.mma (M=128, N<=256, K=16) for T row tiles x taps x k-steps,
//              weight-stationary inside a pass (a weight block is read once per T*128 rows), tcgen05.commit to free
//              smem stages and to publish accumulators
//   warps 2-5: epilogue - tcgen05.ld -> fused scale/shift (folded BatchNorm) + residual + ReLU, or raw output +
//              per-channel sum / sum-of-squares for training-mode BatchNorm -> bf16 planar store (coalesced 16 B
//              per lane), optional stride-2 subsampling of the output grid.
// Accumulator sets in TMEM are double-buffered when 2*T*N <= 512 columns so the epilogue overlaps the next pass.
#include "common.cuh"

using namespace dba;

namespace {

constexpr int kTileM = 128;
constexpr int kThreads = 192;

struct ConvArgs {
  const __nv_bfloat16* X;      // [Kx][Cin/8][R][8]
  const __nv_bfloat16* Wt;     // [K][taps][Cin/8][Cout][8]
  __nv_bfloat16* Y;            // [K][Cout/8][Ro][8]
  const __nv_bfloat16* Res;    // optional, layout of Y
  const float* scale;          // optional [K][Cout]
  const float* shift;          // optional [K][Cout]
  float* stats;                // optional [K][Cout][2]  (sum, sum of squares over stored outputs)
  long long x_client_stride;   // elements; 0 = all clients read the same input (shared evaluation batch)
  long long x_plane_stride;    // elements between channel-chunk planes (= R*8)
  long long y_client_stride, y_plane_stride;
  long long w_client_stride;   // elements between the clients' packed weight sets
  long long ss_stride;         // floats between the clients' scale / shift rows
  long long stats_stride;      // floats between the clients' statistics rows
  int K, Cin, Cout, taps;      // taps = 1 (1x1) or 9 (3x3, pad 1)
  int H, W, Wp, S, lead, B;    // input (= compute) geometry
  int T;                       // row tiles per pass
  int NB;                      // output channels per n-block (<= 256), Cout % NB == 0
  int KB;                      // input channels per k-block (multiple of 16, <= 64)
  int relu;
  int sub;                     // 1 = same grid, 2 = keep even (h, w) only (stride-2 convolution)
  int Ho, Wo, Wpo, So, leado;  // output geometry (equals input geometry when sub == 1)
  int SA, SB, NS;              // pipeline depths: A stages, B stages, accumulator sets
  int rowsA;                   // T*128 + halo rows
  int halo;                    // Wp + 1 for 3x3, 0 for 1x1
  int num_pass;                // ceil(B*S / (T*128))
  int lbo_sbo_swap;            // debug: swap descriptor LBO/SBO meaning
};

struct Pipe {
  int it = 0;
  __device__ __forceinline__ int stage(int n) const { return it % n; }
  __device__ __forceinline__ uint32_t phase(int n) const { return (uint32_t)((it / n) & 1); }
};

__device__ __forceinline__ void warp_transpose_reduce32(float (&v)[32], int lane) {
  // After the call v[0] of lane j holds the sum over all 32 lanes of the original v[j].
#pragma unroll
  for (int s = 16; s >= 1; s >>= 1) {
#pragma unroll
    for (int i = 0; i < s; ++i) {
      const bool hi = (lane & s) != 0;
      const float send = hi ? v[i] : v[i + s];
      const float keep = hi ? v[i + s] : v[i];
      v[i] = keep + __shfl_xor_sync(0xffffffffu, send, s);
    }
  }
}

__global__ void __launch_bounds__(kThreads, 1) conv_tc_kernel(const ConvArgs a) {
  extern __shared__ __align__(128) uint8_t smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int kc = a.KB >> 3;                      // 16-byte channel chunks per k-block
  const int nkb = a.Cin / a.KB;
  const int nnb = a.Cout / a.NB;
  const uint32_t pitchA = (uint32_t)a.rowsA * 16u;
  const uint32_t bytesA = (uint32_t)kc * pitchA;
  const uint32_t pitchB = (uint32_t)a.NB * 16u;
  const uint32_t bytesB = (uint32_t)kc * pitchB;

  uint8_t* smA = smem;
  uint8_t* smB = smA + (size_t)a.SA * bytesA;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smB + (size_t)a.SB * bytesB);
  uint64_t* a_full = bars;
  uint64_t* a_empty = a_full + a.SA;
  uint64_t* b_full = a_empty + a.SA;
  uint64_t* b_empty = b_full + a.SB;
  uint64_t* acc_full = b_empty + a.SB;
  uint64_t* acc_empty = acc_full + a.NS;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + a.NS);
  float* stat_s = reinterpret_cast<float*>(tmem_slot + 4);   // [NB][2] when stats requested

  if (threadIdx.x == 0) {
    for (int i = 0; i < a.SA; ++i) { mbar_init(&a_full[i], 1); mbar_init(&a_empty[i], 1); }
    for (int i = 0; i < a.SB; ++i) { mbar_init(&b_full[i], 1); mbar_init(&b_empty[i], 1); }
    for (int i = 0; i < a.NS; ++i) { mbar_init(&acc_full[i], 1); mbar_init(&acc_empty[i], 4); }
    mbar_fence_init();
  }
  if (warp == 1) {
    tmem_alloc(tmem_slot, 512);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  const int items = a.K * a.num_pass * nnb;
  const int accCols = a.T * a.NB;               // columns of one accumulator set

  if (warp == 0) {
    // ================================================================= producer
    if (elect_one()) {
      Pipe pa, pb;
      for (int item = blockIdx.x; item < items; item += gridDim.x) {
        const int nb = item % nnb;
        const int pass = (item / nnb) % a.num_pass;
        const int k = item / (nnb * a.num_pass);
        const long long row0 = (long long)a.lead + (long long)pass * a.T * kTileM - a.halo;
        const __nv_bfloat16* xk = a.X + (long long)k * a.x_client_stride;
        const __nv_bfloat16* wk = a.Wt + (long long)k * a.w_client_stride;
        for (int kb = 0; kb < nkb; ++kb) {
          const int sa = pa.stage(a.SA);
          mbar_wait(&a_empty[sa], pa.phase(a.SA) ^ 1u);
          mbar_arrive_expect_tx(&a_full[sa], bytesA);
          for (int c = 0; c < kc; ++c) {
            const __nv_bfloat16* src = xk + (long long)(kb * kc + c) * a.x_plane_stride + row0 * 8;
            bulk_g2s(smA + (size_t)sa * bytesA + (size_t)c * pitchA, src, pitchA, &a_full[sa]);
          }
          ++pa.it;
          for (int tap = 0; tap < a.taps; ++tap) {
            const int sb = pb.stage(a.SB);
            mbar_wait(&b_empty[sb], pb.phase(a.SB) ^ 1u);
            mbar_arrive_expect_tx(&b_full[sb], bytesB);
            // weights [tap][Cin/8][Cout][8]: rows nb*NB .. +NB of chunks kb*kc .. +kc
            for (int c = 0; c < kc; ++c) {
              const __nv_bfloat16* src =
                  wk + (((long long)tap * (a.Cin >> 3) + (kb * kc + c)) * a.Cout + (long long)nb * a.NB) * 8;
              bulk_g2s(smB + (size_t)sb * bytesB + (size_t)c * pitchB, src, pitchB, &b_full[sb]);
            }
            ++pb.it;
          }
        }
      }
    }
  } else if (warp == 1) {
    // =============================================================== MMA issuer
    if (elect_one()) {
      Pipe pa, pb, pacc;
      const uint32_t idesc = umma_idesc_bf16(kTileM, a.NB, 0, 0);
      for (int item = blockIdx.x; item < items; item += gridDim.x) {
        const int as = pacc.stage(a.NS);
        mbar_wait(&acc_empty[as], pacc.phase(a.NS) ^ 1u);
        tc_fence_after();
        for (int kb = 0; kb < nkb; ++kb) {
          const int sa = pa.stage(a.SA);
          mbar_wait(&a_full[sa], pa.phase(a.SA));
          tc_fence_after();
          const uint32_t a_base = smem_u32(smA + (size_t)sa * bytesA);
          for (int tap = 0; tap < a.taps; ++tap) {
            const int sb = pb.stage(a.SB);
            mbar_wait(&b_full[sb], pb.phase(a.SB));
            tc_fence_after();
            const uint32_t b_base = smem_u32(smB + (size_t)sb * bytesB);
            int shift = 0;
            if (a.taps == 9) shift = (tap / 3 - 1) * a.Wp + (tap % 3 - 1);
            for (int t = 0; t < a.T; ++t) {
              const uint32_t d_tmem = tmem_base + (uint32_t)(as * accCols + t * a.NB);
              const uint32_t a_row = (uint32_t)(t * kTileM + a.halo + shift) * 16u;
              for (int kk = 0; kk < (kc >> 1); ++kk) {
                const uint32_t a_addr = a_base + (uint32_t)(2 * kk) * pitchA + a_row;
                const uint32_t b_addr = b_base + (uint32_t)(2 * kk) * pitchB;
                uint64_t da, db;
                if (!a.lbo_sbo_swap) {
                  da = umma_desc_noswz(a_addr, pitchA, 128);
                  db = umma_desc_noswz(b_addr, pitchB, 128);
                } else {
                  da = umma_desc_noswz(a_addr, 128, pitchA);
                  db = umma_desc_noswz(b_addr, 128, pitchB);
                }
                umma_f16(d_tmem, da, db, idesc, (kb | tap | kk) != 0 ? 1u : 0u);
              }
            }
            umma_commit(&b_empty[sb]);
            ++pb.it;
          }
          umma_commit(&a_empty[sa]);
          ++pa.it;
        }
        umma_commit(&acc_full[as]);
        ++pacc.it;
      }
    }
  } else {
    // ================================================================= epilogue
    const int q = warp & 3;                       // TMEM lane quadrant this warp may access
    Pipe pacc;
    const int totalRows = a.B * a.S;
    for (int item = blockIdx.x; item < items; item += gridDim.x) {
      const int nb = item % nnb;
      const int pass = (item / nnb) % a.num_pass;
      const int k = item / (nnb * a.num_pass);
      const int as = pacc.stage(a.NS);
      if (a.stats) {
        for (int i = threadIdx.x - 64; i < 2 * a.NB; i += 128) stat_s[i] = 0.f;
        asm volatile("bar.sync 1, 128;" ::: "memory");
      }
      mbar_wait(&acc_full[as], pacc.phase(a.NS));
      tc_fence_after();
      const float* sc = a.scale ? a.scale + (long long)k * a.ss_stride + nb * a.NB : nullptr;
      const float* sh = a.shift ? a.shift + (long long)k * a.ss_stride + nb * a.NB : nullptr;
      __nv_bfloat16* yk = a.Y + (long long)k * a.y_client_stride;
      const __nv_bfloat16* rk = a.Res ? a.Res + (long long)k * a.y_client_stride : nullptr;
      for (int t = 0; t < a.T; ++t) {
        const int qrow = pass * a.T * kTileM + t * kTileM + q * 32 + lane;   // row index relative to `lead`
        bool store = qrow < totalRows;
        long long orow = 0;
        {
          const int b = qrow / a.S;
          const int rem = qrow - b * a.S;
          const int h = rem / a.Wp, w = rem - h * a.Wp;
          store = store && h < a.H && w < a.W;
          if (a.sub == 2) {
            store = store && !((h | w) & 1);
            orow = (long long)a.leado + (long long)b * a.So + (h >> 1) * a.Wpo + (w >> 1);
          } else {
            orow = (long long)a.lead + qrow;
          }
        }
        const uint32_t t_addr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(as * accCols + t * a.NB);
        for (int c0 = 0; c0 < a.NB; c0 += 32) {
          float sv[32];
#pragma unroll
          for (int h2 = 0; h2 < 2; ++h2) {
            uint32_t v[16];
            tmem_ld_32x32b_x16(t_addr + (uint32_t)(c0 + 16 * h2), v);
            tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 16; ++j) sv[16 * h2 + j] = __uint_as_float(v[j]);
          }
          // NB is a multiple of 16; a trailing half-group of 16 columns is handled by masking below
          const int ncol = min(32, a.NB - c0);
#pragma unroll
          for (int g = 0; g < 4; ++g) {            // 4 chunks of 8 channels
            if (g * 8 >= ncol) break;
            const int ch = c0 + g * 8;
            float o[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              float x = sv[g * 8 + j];
              if (sc) x = x * __ldg(sc + ch + j) + __ldg(sh + ch + j);
              o[j] = x;
            }
            const long long off = (long long)((nb * a.NB + ch) >> 3) * a.y_plane_stride + orow * 8;
            if (store) {
              if (rk) {
                const uint4 r = *reinterpret_cast<const uint4*>(rk + off);
                o[0] += bf16lo(r.x); o[1] += bf16hi(r.x); o[2] += bf16lo(r.y); o[3] += bf16hi(r.y);
                o[4] += bf16lo(r.z); o[5] += bf16hi(r.z); o[6] += bf16lo(r.w); o[7] += bf16hi(r.w);
              }
              if (a.relu) {
#pragma unroll
                for (int j = 0; j < 8; ++j) o[j] = fmaxf(o[j], 0.f);
              }
              uint4 pk;
              pk.x = pack_bf16x2(o[0], o[1]); pk.y = pack_bf16x2(o[2], o[3]);
              pk.z = pack_bf16x2(o[4], o[5]); pk.w = pack_bf16x2(o[6], o[7]);
              *reinterpret_cast<uint4*>(yk + off) = pk;
            }
#pragma unroll
            for (int j = 0; j < 8; ++j) sv[g * 8 + j] = store ? o[j] : 0.f;
          }
          if (a.stats) {
            float s1[32], s2[32];
#pragma unroll
            for (int j = 0; j < 32; ++j) {
              // statistics of the value as stored (bf16-rounded), so BatchNorm normalises exactly what it reads
              const float r = (j < ncol) ? __bfloat162float(__float2bfloat16(sv[j])) : 0.f;
              s1[j] = r;
              s2[j] = r * r;
            }
            warp_transpose_reduce32(s1, lane);
            warp_transpose_reduce32(s2, lane);
            if (lane < ncol) {
              atomicAdd(&stat_s[2 * (c0 + lane)], s1[0]);
              atomicAdd(&stat_s[2 * (c0 + lane) + 1], s2[0]);
            }
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&acc_empty[as]);
      ++pacc.it;
      if (a.stats) {
        asm volatile("bar.sync 1, 128;" ::: "memory");
        float* gs = a.stats + (long long)k * a.stats_stride + (long long)nb * a.NB * 2;
        for (int i = threadIdx.x - 64; i < 2 * a.NB; i += 128) {
          const float v = stat_s[i];
          if (v != 0.f) atomicAdd(gs + i, v);
        }
        asm volatile("bar.sync 1, 128;" ::: "memory");
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

}  // namespace

extern "C" {

// Geometry helpers shared with the Python side (dba_mod_b200/ops/layout.py mirrors these).
long long dba_conv_rows(int B, int H, int W) {
  const int Wp = W + 1, S = (H + 1) * Wp, lead = Wp + 1;
  long long body = (long long)B * S;
  body = (body + 511) / 512 * 512;            // any T*128 <= 512 rows per pass may overrun into the tail
  return lead + body + Wp + 2;
}

// Y = epilogue(conv(X, Wt)).  See ConvArgs for layouts.  X: [Kx][Cin/8][R][8] with R = dba_conv_rows(B,H,W);
// Wt: [K][taps][Cin/8][Cout][8]; Y (and Res): [K][Cout/8][Ro][8], Ro = dba_conv_rows(B,Ho,Wo), Ho = H/sub.
// scale/shift (nullable together): per (client, out-channel) affine = folded BatchNorm.  stats (nullable): [K][Cout][2]
// accumulated (+=) with sum / sum-of-squares of the stored outputs.  x_shared != 0: every client reads client 0's X.
int dba_conv_tc_strided(const void* X, const void* Wt, long long w_client_stride, void* Y, const void* Res,
                        const float* scale, const float* shift, long long ss_stride, float* stats,
                        long long stats_stride, int K, int B, int H, int W, int Cin, int Cout, int ksize, int sub,
                        int relu, int x_shared, void* stream) {
  const int debug_swap = 0;
  DBA_REQUIRE(K > 0 && B > 0 && H > 0 && W > 0, "dba_conv_tc: bad sizes K=%d B=%d H=%d W=%d", K, B, H, W);
  DBA_REQUIRE(ksize == 1 || ksize == 3, "dba_conv_tc: ksize must be 1 or 3, got %d", ksize);
  DBA_REQUIRE(Cin % 16 == 0 && Cout % 16 == 0, "dba_conv_tc: Cin (%d) and Cout (%d) must be multiples of 16", Cin, Cout);
  DBA_REQUIRE(sub == 1 || (sub == 2 && H % 2 == 0 && W % 2 == 0), "dba_conv_tc: sub=%d needs even H, W", sub);
  DBA_REQUIRE((scale == nullptr) == (shift == nullptr), "dba_conv_tc: scale and shift go together");
  ConvArgs a{};
  a.X = (const __nv_bfloat16*)X; a.Wt = (const __nv_bfloat16*)Wt; a.Y = (__nv_bfloat16*)Y;
  a.Res = (const __nv_bfloat16*)Res; a.scale = scale; a.shift = shift; a.stats = stats;
  a.K = K; a.Cin = Cin; a.Cout = Cout; a.taps = ksize * ksize;
  a.H = H; a.W = W; a.Wp = W + 1; a.S = (H + 1) * a.Wp; a.lead = a.Wp + 1; a.B = B;
  a.relu = relu; a.sub = sub;
  a.Ho = H / sub; a.Wo = W / sub; a.Wpo = a.Wo + 1; a.So = (a.Ho + 1) * a.Wpo; a.leado = a.Wpo + 1;
  const long long R = dba_conv_rows(B, H, W), Ro = dba_conv_rows(B, a.Ho, a.Wo);
  a.x_plane_stride = R * 8;
  a.x_client_stride = x_shared ? 0 : (long long)(Cin / 8) * R * 8;
  a.y_plane_stride = Ro * 8;
  a.y_client_stride = (long long)(Cout / 8) * Ro * 8;
  a.w_client_stride = w_client_stride;
  a.ss_stride = ss_stride;
  a.stats_stride = stats_stride;
  a.NB = Cout <= 256 ? Cout : 256;
  DBA_REQUIRE(Cout % a.NB == 0, "dba_conv_tc: Cout=%d not a multiple of the n-block %d", Cout, a.NB);
  a.KB = Cin >= 64 ? 64 : Cin;
  DBA_REQUIRE(Cin % a.KB == 0, "dba_conv_tc: Cin=%d not a multiple of the k-block %d", Cin, a.KB);
  a.halo = ksize == 3 ? a.Wp + 1 : 0;
  a.lbo_sbo_swap = debug_swap;
  // rows per pass: as many 128-row tiles as fit in TMEM (512 columns) and in shared memory
  const int kc = a.KB / 8;
  const long long body = (long long)B * a.S;
  DBA_REQUIRE(a.NB % 32 == 0, "dba_conv_tc: output channels per block (%d) must be a multiple of 32", a.NB);
  int T = 512 / a.NB;                          // T stays a power of two so that T*128 divides the 512-row tail slack
  if (T > 4) T = 4;
  while (T > 1 && (long long)(T / 2) * kTileM >= body) T /= 2;
  const size_t budget = 220 * 1024;
  size_t need = 0;
  for (;; T /= 2) {
    a.T = T;
    a.rowsA = T * kTileM + 2 * a.halo;
    const size_t bytesA = (size_t)kc * a.rowsA * 16, bytesB = (size_t)kc * a.NB * 16;
    a.SA = 2;
    a.SB = 4;
    a.NS = (2 * T * a.NB <= 512) ? 2 : 1;
    need = a.SA * bytesA + a.SB * bytesB + 1024 + (stats ? 2 * a.NB * sizeof(float) : 0);
    if (need > budget) { a.SB = 2; need = a.SA * bytesA + a.SB * bytesB + 1024 + (stats ? 2 * a.NB * sizeof(float) : 0); }
    if (need <= budget || T == 1) break;
  }
  DBA_REQUIRE(need <= budget, "dba_conv_tc: shared memory need %zu exceeds budget", need);
  a.num_pass = (int)((body + (long long)a.T * kTileM - 1) / ((long long)a.T * kTileM));
  const int items = K * a.num_pass * (Cout / a.NB);
  static bool attr_set[64] = {false};
  int dev = 0;
  DBA_CHECK_CUDA(cudaGetDevice(&dev));
  if (dev < 64 && !attr_set[dev]) {
    DBA_CHECK_CUDA(cudaFuncSetAttribute(conv_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)budget));
    attr_set[dev] = true;
  }
  if (need < 120 * 1024) need = 120 * 1024;    // one CTA per SM: each CTA allocates all 512 TMEM columns
  const int grid = items < kNumSMs ? items : kNumSMs;
  conv_tc_kernel<<<grid, kThreads, need, (cudaStream_t)stream>>>(a);
  DBA_COUNT_LAUNCH(1);
  DBA_CHECK_LAUNCH();
  return DBA_OK;
}

// Dense-stride convenience form: Wt [K][taps][Cin/8][Cout][8], scale/shift [K][Cout], stats [K][Cout][2].
int dba_conv_tc(const void* X, const void* Wt, void* Y, const void* Res, const float* scale, const float* shift,
                float* stats, int K, int B, int H, int W, int Cin, int Cout, int ksize, int sub, int relu,
                int x_shared, int reserved, void* stream) {
  (void)reserved;
  return dba_conv_tc_strided(X, Wt, (long long)ksize * ksize * Cin * Cout, Y, Res, scale, shift, Cout, stats,
                             2LL * Cout, K, B, H, W, Cin, Cout, ksize, sub, relu, x_shared, stream);
}

}  // extern "C"
