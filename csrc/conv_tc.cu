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
// Work items (client, n-block, pass) are dealt to CTAs in contiguous ranges, so a CTA stays with one weight set for
// as long as possible: when all taps of that weight set fit in shared memory ("resident" mode, every layer with
// Cin*Cout <= 64*64) they are fetched ONCE per (client, n-block) with a single bulk copy and the per-tap
// producer/consumer round trips disappear; larger layers stream one weight block per (tap, k-block).
#include <cstdlib>
#include "common.cuh"

using namespace dba;


namespace {

constexpr int kTileM = 128;
constexpr int kEpiWarps = 8;                  // two warps per TMEM lane quadrant
constexpr int kMaxSB = 16;      // weight-stream pipeline depth limit (barrier slots)
// Producer warps.  Measured (tools/gpu_bulk_probe.cu): a warp has ONE cp.async.bulk instruction in flight - the next one
// issues ~500 cycles + ~55 cycles per active lane later, however many pipeline stages are free - so a single producer
// warp feeds 8 KB weight stages at ~10 B/clk (1 KB rows: 930 cycles per stage against ~300 cycles of MMAs).  Stages are
// therefore dealt to kProd warps by stage index (stage s belongs to warp s % kProd: one owner per mbarrier, so nobody
// can run a whole phase ahead of a barrier it waits on).
constexpr int kProd = 3;
constexpr int kMmaWarp = kProd;
constexpr int kEpiWarp0 = kProd + 1;
constexpr int kEpiThread0 = 32 * kEpiWarp0;
constexpr int kThreads = kEpiThread0 + 32 * kEpiWarps;
static_assert(kEpiWarp0 % 4 == 0, "epilogue warp e must sit on TMEM lane quadrant e & 3");

struct ConvArgs {
  // All tensors are addressed in 16-byte chunks ("rows") holding V channels: V = 8 (bf16) or 4 (fp32 / tf32 mode).
  const uint4* X;              // [Kx][Cin/V][R]
  const uint4* Wt;             // [K][taps][Cin/V][Cout]
  uint4* Y;                    // [K][Cout/V][Ro]
  const uint4* Res;            // optional, layout of Y
  const float* scale;          // optional [K][Cout]
  const float* shift;          // optional [K][Cout]
  float* stats;                // optional [K][Cout][2]  (sum, sum of squares over stored outputs)
  long long x_client_stride;   // rows; 0 = all clients read the same input (shared evaluation batch)
  long long x_plane_stride;    // rows between channel-chunk planes (= R)
  long long y_client_stride, y_plane_stride;
  long long w_client_stride;   // rows between the clients' packed weight sets (0 = one model for all)
  long long ss_stride;         // floats between the clients' scale / shift rows
  long long stats_stride;      // floats between the clients' statistics rows
  int K, Cin, Cout, taps;      // taps = 1 (1x1) or 9 (3x3, pad 1)
  int H, W, Wp, S, lead, B;    // input (= compute) geometry
  int T;                       // row tiles per pass
  int NB;                      // output channels per n-block (<= 256), Cout % NB == 0
  int KB;                      // input channels per k-block: 2, 4 or 8 chunks (bf16: 16 / 32 / 64, tf32: 8 / 16 / 32)
  int relu;
  int sub;                     // 1 = same grid, 2 = keep even (h, w) only (stride-2 convolution)
  int Ho, Wo, Wpo, So, leado;  // output geometry (equals input geometry when sub == 1)
  int SA, SB, NS;              // pipeline depths: A stages, B stages, accumulator sets
  int rowsA;                   // T*128 + halo rows
  int halo;                    // Wp + 1 for 3x3, 0 for 1x1
  int num_pass;                // ceil(B*S / (T*128))
  int resident;                // 1 = whole weight set of a (client, n-block) kept in shared memory
  unsigned int wbytes;         // bytes of that weight set
  unsigned long long magicS, magicWp;   // ceil(2^40 / S), ceil(2^40 / Wp): exact division of row indices < 2^26
  long long* dbg;              // optional [grid][8] cycle counters (profiling aid), null in production
  int dbg_shift;               // profiling aid: 0 = real tap shifts, 1 = all zero, 2 = multiples of 8 rows (wrong results)
  // Stride-2 convolutions without the 4x full-grid waste: X is the space-to-depth form of the real input (parity
  // planes p = 2*(h&1) + (w&1), each Cin channels, on the OUTPUT grid); tap (dh, dw) of the 3x3 kernel reads parity
  // plane (dh != 1, dw != 1) shifted by (dh == 0 ? -1 : 0, dw == 0 ? -1 : 0); the 1x1 kernel reads plane 0 unshifted.
  int ksz;                     // 1, 3 or 5: taps = ksz*ksz, centred (pad = ksz/2)
  int vlo, vhi;                // outputs are stored for vlo <= h, w < vhi only (a 'valid' convolution is the centre
                               // crop of the 'same' one: 5x5 pad 0 on 12x12 -> vlo = 2, vhi = 10)
  int in_s2d;
  // 1: store Y (and only Y - Res stays in the plain layout) as the space-to-depth form of the (H, W) output.
  // 2: store Y - and read Res - at the even positions (2h, 2w) of the (2H, 2W) grid: the input gradient of a 1x1
  //    stride-2 convolution computed on its own (half-resolution) grid and added in place into the full-grid gradient.
  // 3: depth-to-space: the Cout = 4*C output channels are four parity groups; group p = 2*ph + pw of pixel (h, w) is
  //    stored as the C channels of pixel (2h + ph, 2w + pw) of the plain (2H, 2W) tensor.  With 2x2 taps and the
  //    re-packed weights of pack mode 3 this is the input gradient of a 3x3 stride-2 convolution from its compact dy.
  int out_s2d;
  int cgrp;                    // out mode 3: channels per parity group (C)
  long long res_client_stride, res_plane_stride;
  // Input-gradient convolutions whose output g feeds a BatchNorm backward: the epilogue also accumulates that layer's
  // reduction sums (sum g', sum g' * xhat; g' = g masked by the layer's ReLU) into `stats`, so bn_bwd_reduce is not
  // launched.  bn_y / bn_act (nullable: no ReLU) have Y's geometry; bn_saved = (mean, rstd) pairs per channel.
  const uint4* bn_y;
  const uint4* bn_act;
  const float* bn_saved;
  long long bn_saved_stride;
};

struct Pipe {
  // ring position kept incrementally: `it % n` and `(it / n) & 1` with a run-time n are ~50 instructions each, which is
  // what bounded the single MMA-issuing thread on layers with few MMAs per stage
  int st = 0;
  uint32_t ph = 0;
  __device__ __forceinline__ int stage(int) const { return st; }
  __device__ __forceinline__ uint32_t phase(int) const { return ph; }
  __device__ __forceinline__ void advance(int n) {
    if (++st == n) { st = 0; ph ^= 1u; }
  }
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

#define DBG_T0() long long t0__ = a.dbg ? clock64() : 0
#define DBG_ADD(slot) do { if (a.dbg) { long long t1__ = clock64(); dbgacc[slot] += t1__ - t0__; t0__ = t1__; } } while (0)

template <int KC2, bool TF32>
__global__ void __launch_bounds__(kThreads, 1) conv_tc_kernel(const ConvArgs a) {
  constexpr int V = TF32 ? 4 : 8;                // channels per 16-byte chunk
  constexpr int NG = 32 / V;                     // chunks per 32-channel epilogue group
  long long dbgacc[4] = {0, 0, 0, 0};
  const long long dbg_start = a.dbg ? clock64() : 0;
  extern __shared__ __align__(128) uint8_t smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int kc = a.KB / V;                       // 16-byte channel chunks per k-block
  const int nkb = a.Cin / a.KB;
  const int nnb = a.Cout / a.NB;
  const int planes = a.Cin / V;
  const uint32_t pitchA = (uint32_t)a.rowsA * 16u;
  const int npar = (a.in_s2d && a.taps == 9) ? 4 : 1;      // parity planes staged per k-block
  const uint32_t bytesA = (uint32_t)(npar * kc) * pitchA;
  const uint32_t pitchB = (uint32_t)a.NB * 16u;
  const uint32_t bytesB = (uint32_t)kc * pitchB;
  const size_t bRegion = a.resident ? (size_t)a.wbytes : (size_t)a.SB * bytesB;

  uint8_t* smA = smem;
  uint8_t* smB = smA + (size_t)a.SA * bytesA;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smB + bRegion);
  uint64_t* a_full = bars;
  uint64_t* a_empty = a_full + a.SA;
  uint64_t* b_full = a_empty + a.SA;            // streaming: SB stages; resident: [0] = weights landed
  uint64_t* b_empty = b_full + kMaxSB;             //                        resident: [0] = weights no longer needed
  uint64_t* acc_full = b_empty + kMaxSB;
  uint64_t* acc_empty = acc_full + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + 2);
  float* ss_s = reinterpret_cast<float*>(tmem_slot + 4);     // [2][NB] scale / shift of the current weight group
  float* mr_s = ss_s + 2 * a.NB;                             // [2][NB] mean / rstd (BatchNorm-backward sums mode)
  int* tap_shift = reinterpret_cast<int*>(ss_s + 4 * a.NB);  // [taps] A-view offset of every tap (16-byte units)

  if (threadIdx.x >= kEpiThread0 && threadIdx.x < kEpiThread0 + a.taps) {      // per-tap view offsets, computed once
    const int tap = threadIdx.x - kEpiThread0;
    int shift = 0;
    if (a.taps > 1) {
      const int th = tap / a.ksz;
      shift = (th - (a.ksz >> 1)) * a.Wp + (tap - th * a.ksz - (a.ksz >> 1));
    }
    if (a.ksz == 2) shift = (tap >> 1) * a.Wp + (tap & 1);        // 2x2 taps reaching down / right (out mode 3)
    if (npar == 4) {
      const int dh = tap / 3, dw = tap - dh * 3;
      shift = (dh == 0 ? -a.Wp : 0) + (dw == 0 ? -1 : 0)
              + (int)((uint32_t)((dh != 1) * 2 + (dw != 1)) * (uint32_t)kc * (pitchA >> 4));
    }
    if (a.dbg_shift == 1) shift = 0;
    if (a.dbg_shift == 2) shift = (tap / 3 - 1) * 24 + (tap % 3 - 1) * 8;
    tap_shift[tap] = shift;
  }
  if (a.stats) {                                // per-CTA partial statistics (the training epilogue reuses ss_s)
    for (int i = threadIdx.x; i < 2 * a.NB; i += kThreads) ss_s[i] = 0.f;
  }
  if (threadIdx.x == 0) {
    for (int i = 0; i < a.SA; ++i) { mbar_init(&a_full[i], 1); mbar_init(&a_empty[i], 1); }
    for (int i = 0; i < kMaxSB; ++i) { mbar_init(&b_full[i], 1); mbar_init(&b_empty[i], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&acc_full[i], 1); mbar_init(&acc_empty[i], kEpiWarps); }
    mbar_fence_init();
  }
  pdl_trigger();                                // the next kernel's blocks may be scheduled from now on
  if (warp == kMmaWarp) {
    tmem_alloc(tmem_slot, 512);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();                                   // everything above overlapped the previous kernel's tail

  // contiguous range of work items for this CTA; item = (k * nnb + nb) * num_pass + pass
  const long long items = (long long)a.K * nnb * a.num_pass;
  const int item_lo = (int)(items * blockIdx.x / gridDim.x);
  const int item_hi = (int)(items * (blockIdx.x + 1) / gridDim.x);
  const int accCols = a.T * a.NB;               // columns of one accumulator set
  if (warp < kProd) {
    // ================================================================= producers
    // Every producer warp walks the same sequence of load events and keeps the same pipeline positions, but issues
    // only the stages it owns (A stage s: warp (kProd - 1 - s) % kProd, B stage s: warp s % kProd; the resident weight
    // set: warp 0).  Within a stage the lanes issue the bulk copies in parallel, lane 0 waits for the free stage and
    // posts the byte count.
    {
      Pipe pa, pb;
      int cur_group = -1, wloads = 0;
      for (int item = item_lo; item < item_hi; ++item) {
        const int group = item / a.num_pass;
        const int pass = item - group * a.num_pass;
        const int nb = group % nnb, k = group / nnb;
        const uint4* xk = a.X + (long long)k * a.x_client_stride;
        const uint4* wk = a.Wt + (long long)k * a.w_client_stride;
        if (a.resident && group != cur_group) {
          cur_group = group;
          if (warp == 0) {
            if (lane == 0) {
              mbar_wait(&b_empty[0], (uint32_t)(wloads & 1) ^ 1u);
              mbar_arrive_expect_tx(&b_full[0], a.wbytes);
            }
            __syncwarp();
            if (nnb == 1) {
              if (lane == 0) bulk_g2s(smB, wk, a.wbytes, &b_full[0]);            // [tap][plane][Cout][8] is contiguous
            } else {
              for (int tp = lane; tp < a.taps * planes; tp += 32)
                bulk_g2s(smB + (size_t)tp * pitchB, wk + ((long long)tp * a.Cout + (long long)nb * a.NB), pitchB,
                         &b_full[0]);
            }
          }
          ++wloads;
        }
        const long long row0 = (long long)a.lead + (long long)pass * a.T * kTileM - a.halo;
        for (int kb = 0; kb < nkb; ++kb) {
          const int sa = pa.stage(a.SA);
          if (warp == (kProd - 1 - sa % kProd)) {
            DBG_T0();
            if (lane == 0) {
              mbar_wait(&a_empty[sa], pa.phase(a.SA) ^ 1u);
              mbar_arrive_expect_tx(&a_full[sa], bytesA);
            }
            __syncwarp();
            DBG_ADD(0);
            for (int i = lane; i < npar * kc; i += 32) {
              const int pp = i / kc, c = i - pp * kc;
              const uint4* src = xk + (long long)(pp * planes + kb * kc + c) * a.x_plane_stride + row0;
              bulk_g2s(smA + (size_t)sa * bytesA + (size_t)i * pitchA, src, pitchA, &a_full[sa]);
            }
          }
          pa.advance(a.SA);
          if (a.resident) continue;
          for (int tap = 0; tap < a.taps; ++tap) {
            const int sb = pb.stage(a.SB);
            if (warp == sb % kProd) {
              if (lane == 0) {
                mbar_wait(&b_empty[sb], pb.phase(a.SB) ^ 1u);
                mbar_arrive_expect_tx(&b_full[sb], bytesB);
              }
              __syncwarp();
              const uint4* wsrc = wk + (((long long)tap * planes + kb * kc) * a.Cout + (long long)nb * a.NB);
              if (nnb == 1) {
                if (lane == 0) bulk_g2s(smB + (size_t)sb * bytesB, wsrc, bytesB, &b_full[sb]);   // kc planes contiguous
              } else if (lane < kc) {
                bulk_g2s(smB + (size_t)sb * bytesB + (size_t)lane * pitchB, wsrc + (long long)lane * a.Cout, pitchB,
                         &b_full[sb]);
              }
            }
            pb.advance(a.SB);
          }
        }
      }
    }
  } else if (warp == kMmaWarp) {
    // =============================================================== MMA issuer
    if (elect_one()) {
      Pipe pa, pb, pacc;
      int cur_group = -1, wuses = 0;
      const uint32_t idesc = TF32 ? umma_idesc_tf32(kTileM, a.NB, 0, 0) : umma_idesc_bf16(kTileM, a.NB, 0, 0);
      const uint32_t kstepA = (2u * pitchA) >> 4, kstepB = (2u * pitchB) >> 4;
      const uint64_t dBres = umma_desc_noswz(smem_u32(smB), pitchB, 128);      // + stage / tap offset in 16-byte units
      for (int item = item_lo; item < item_hi; ++item) {
        const int group = item / a.num_pass;
        if (a.resident && group != cur_group) {
          cur_group = group;
          mbar_wait(&b_full[0], (uint32_t)(wuses & 1));
          tc_fence_after();
          ++wuses;
        }
        const int as = pacc.stage(a.NS);
        DBG_T0();
        mbar_wait(&acc_empty[as], pacc.phase(a.NS) ^ 1u);
        DBG_ADD(0);
        tc_fence_after();
        for (int kb = 0; kb < nkb; ++kb) {
          const int sa = pa.stage(a.SA);
          mbar_wait(&a_full[sa], pa.phase(a.SA));
          DBG_ADD(1);
          tc_fence_after();
          // Descriptors differ from one MMA to the next only in their 14-bit start-address field: build the stage's
          // base descriptors once and add row / k-step offsets (in 16-byte units) - the single issuing thread must
          // spend only a handful of instructions per tcgen05.mma, or small-N layers become issue-bound.
          const uint64_t dA0 = umma_desc_noswz(smem_u32(smA + (size_t)sa * bytesA), pitchA, 128) + (uint64_t)a.halo;
          for (int tap = 0; tap < a.taps; ++tap) {
            uint64_t dB0;
            int sb = 0;
            if (a.resident) {
              dB0 = dBres + (uint64_t)((uint32_t)(tap * planes + kb * kc) * (pitchB >> 4));
            } else {
              sb = pb.stage(a.SB);
              mbar_wait(&b_full[sb], pb.phase(a.SB));
              tc_fence_after();
              dB0 = dBres + (uint64_t)((uint32_t)sb * (bytesB >> 4));
            }
            const uint32_t first = (kb | tap) != 0 ? 1u : 0u;
            uint32_t d_tmem = tmem_base + (uint32_t)(as * accCols);
            uint64_t da = dA0 + (uint64_t)(int64_t)tap_shift[tap];    // halo + shift >= 0: no borrow out of the field
            for (int t = 0; t < a.T; ++t, d_tmem += (uint32_t)a.NB, da += (uint64_t)kTileM) {
              umma<TF32>(d_tmem, da, dB0, idesc, first);
              if (KC2 >= 2) umma<TF32>(d_tmem, da + kstepA, dB0 + kstepB, idesc, 1u);
              if (KC2 >= 4) {
                umma<TF32>(d_tmem, da + 2 * kstepA, dB0 + 2 * kstepB, idesc, 1u);
                umma<TF32>(d_tmem, da + 3 * kstepA, dB0 + 3 * kstepB, idesc, 1u);
              }
            }
            if (!a.resident) {
              umma_commit(&b_empty[sb]);
              pb.advance(a.SB);
            }
          }
          umma_commit(&a_empty[sa]);
          pa.advance(a.SA);
          DBG_ADD(2);
        }
        umma_commit(&acc_full[as]);
        pacc.advance(a.NS);
        if (a.resident && (item + 1 == item_hi || (item + 1) / a.num_pass != group)) umma_commit(&b_empty[0]);
      }
    }
  } else {
    // ================================================================= epilogue
    // 8 warps: warps e and e+4 share TMEM lane quadrant (e & 3) and split the pass' tiles (or, for single-tile
    // passes, its 32-channel groups) between them - with one warp per scheduler the epilogue is latency-bound.
    const int ew = warp - kEpiWarp0;
    const int q = warp & 3;                       // TMEM lane quadrant this warp may access
    const int half = ew >> 2;                     // 0 / 1
    Pipe pacc;
    const int totalRows = a.B * a.S;
    int cur_group = -1;
    const bool affine = a.shift != nullptr;       // scale may be null: bias-only epilogue (networks without BatchNorm)
    const int t_step = a.T >= 2 ? 2 : 1, t_first = a.T >= 2 ? half : 0;
    const int c_step = a.T >= 2 ? 32 : 64, c_first = a.T >= 2 ? 0 : 32 * half;
    // Training epilogue: the 8 warps add their per-pass channel sums into ss_s (shared-memory atomics) and the CTA
    // sends ONE global atomic per channel when its (client, n-block) group ends.  Per-warp global atomics put ~2000
    // same-address updates per channel on the L2 for a 64x32x32 layer: 17 of that convolution's 31 us (measured, K = 1).
    auto flush_stats = [&](int grp) {
      const int nb_ = grp % nnb, k_ = grp / nnb;
      float* g = a.stats + (long long)k_ * a.stats_stride + (long long)nb_ * a.NB * 2;
      asm volatile("bar.sync 1, 256;" ::: "memory");
      for (int i = threadIdx.x - kEpiThread0; i < 2 * a.NB; i += 32 * kEpiWarps) {
        const float v = ss_s[i];
        if (v != 0.f) { atomicAdd(g + i, v); ss_s[i] = 0.f; }
      }
      asm volatile("bar.sync 1, 256;" ::: "memory");
    };
    for (int item = item_lo; item < item_hi; ++item) {
      const int group = item / a.num_pass;
      const int pass = item - group * a.num_pass;
      const int nb = group % nnb, k = group / nnb;
      const int as = pacc.stage(a.NS);
      if (group != cur_group) {
        if (a.stats && cur_group >= 0) flush_stats(cur_group);
        cur_group = group;
        if (a.bn_y) {     // mean / rstd of the consuming BatchNorm layer for this (client, n-block)
          const float* sv_ = a.bn_saved + (long long)k * a.bn_saved_stride + 2 * nb * a.NB;
          asm volatile("bar.sync 1, 256;" ::: "memory");
          for (int i = threadIdx.x - kEpiThread0; i < a.NB; i += 32 * kEpiWarps) {
            mr_s[i] = sv_[2 * i];
            mr_s[a.NB + i] = sv_[2 * i + 1];
          }
          asm volatile("bar.sync 1, 256;" ::: "memory");
        }
        if (affine) {     // per-channel scale / shift of this (client, n-block) -> shared memory
          const float* gsc = a.scale ? a.scale + (long long)k * a.ss_stride + nb * a.NB : nullptr;
          const float* gsh = a.shift + (long long)k * a.ss_stride + nb * a.NB;
          asm volatile("bar.sync 1, 256;" ::: "memory");
          for (int i = threadIdx.x - kEpiThread0; i < a.NB; i += 32 * kEpiWarps) {
            ss_s[i] = gsc ? gsc[i] : 1.f;
            ss_s[a.NB + i] = gsh[i];
          }
          asm volatile("bar.sync 1, 256;" ::: "memory");
        }
      }
      // which of this thread's rows (one per tile) are real output pixels, and where they go
      bool store[2];
      long long orow[2], rrow[2];                 // output row / residual row (differ only for out_s2d)
      long long d2s_row[2];                       // out mode 3: row of pixel (2h, 2w) on the (2H, 2W) grid
#pragma unroll
      for (int i = 0; i < 2; ++i) {
        const int t = t_first + i * t_step;
        store[i] = false;
        orow[i] = rrow[i] = d2s_row[i] = 0;
        if (t < a.T) {
          const unsigned int qrow = (unsigned int)(pass * a.T * kTileM + t * kTileM + q * 32 + lane);
          const unsigned int b = (unsigned int)(((unsigned long long)qrow * a.magicS) >> 40);
          const unsigned int rem = qrow - b * (unsigned int)a.S;
          const unsigned int h = (unsigned int)(((unsigned long long)rem * a.magicWp) >> 40);
          const unsigned int w = rem - h * (unsigned int)a.Wp;
          bool st = (int)qrow < totalRows && (int)h < a.vhi && (int)w < a.vhi && (int)h >= a.vlo && (int)w >= a.vlo &&
                    (int)h < a.H && (int)w < a.W;
          if (a.sub == 2) {
            st = st && !((h | w) & 1u);
            orow[i] = rrow[i] = (long long)a.leado + (long long)b * a.So + (h >> 1) * a.Wpo + (w >> 1);
          } else if (a.out_s2d >= 2) {
            orow[i] = rrow[i] = d2s_row[i] = (long long)a.leado + (long long)b * a.So + (2 * h) * a.Wpo + 2 * w;
          } else if (a.out_s2d) {
            rrow[i] = (long long)a.lead + qrow;
            orow[i] = (long long)(((h & 1u) << 1) | (w & 1u)) * a.y_plane_stride * (a.Cout / V)
                      + (long long)a.leado + (long long)b * a.So + (h >> 1) * a.Wpo + (w >> 1);
          } else {
            orow[i] = rrow[i] = (long long)a.lead + qrow;
          }
          store[i] = st;
        }
      }
      DBG_T0();
      mbar_wait(&acc_full[as], pacc.phase(a.NS));
      DBG_ADD(0);
      tc_fence_after();
      uint4* yk = a.Y + (long long)k * a.y_client_stride;
      const uint4* rk = a.Res ? a.Res + (long long)k * a.res_client_stride : nullptr;
      const uint4* bnyk = a.bn_y ? a.bn_y + (long long)k * a.y_client_stride : nullptr;
      const uint4* bnak = a.bn_act ? a.bn_act + (long long)k * a.y_client_stride : nullptr;
      const bool gs = a.stats != nullptr;
      for (int c0 = c_first; c0 < a.NB; c0 += c_step) {
        // u1/u2 hold scale/shift (evaluation epilogue) or the running sum / sum of squares (training epilogue);
        // the two modes are mutually exclusive (checked on the host), sharing the registers avoids spills
        float u1[32], u2[32];
        if (affine) {
#pragma unroll
          for (int j = 0; j < 32; ++j) { u1[j] = ss_s[c0 + j]; u2[j] = ss_s[a.NB + c0 + j]; }
        } else {
#pragma unroll
          for (int j = 0; j < 32; ++j) u1[j] = u2[j] = 0.f;
        }
#pragma unroll
        for (int i = 0; i < 2; ++i) {
          const int t = t_first + i * t_step;
          if (t >= a.T) break;
          const uint32_t t_addr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(as * accCols + t * a.NB + c0);
          float sv[32];
          {
            uint32_t v[16], v2[16];
            tmem_ld_32x32b_x16(t_addr, v);
            tmem_ld_32x32b_x16(t_addr + 16u, v2);
            tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 16; ++j) { sv[j] = __uint_as_float(v[j]); sv[16 + j] = __uint_as_float(v2[j]); }
          }
          if (affine) {
#pragma unroll
            for (int j = 0; j < 32; ++j) sv[j] = fmaf(sv[j], u1[j], u2[j]);
          }
          if (!store[i]) continue;
          long long obase = (long long)((nb * a.NB + c0) / V) * a.y_plane_stride + orow[i];
          if (a.out_s2d == 3) {
            const int cb = nb * a.NB + c0, par = cb / a.cgrp, cc = cb - par * a.cgrp;
            obase = (long long)(cc / V) * a.y_plane_stride + (d2s_row[i] + (par >> 1) * a.Wpo + (par & 1));
          }
          if (rk) {
            const long long rbase = (long long)((nb * a.NB + c0) / V) * a.res_plane_stride + rrow[i];
#pragma unroll
            for (int g0 = 0; g0 < NG; g0 += 4) {       // four 16-byte loads in flight
              uint4 r[4];
#pragma unroll
              for (int g = 0; g < 4; ++g) r[g] = rk[rbase + (g0 + g) * a.res_plane_stride];
#pragma unroll
              for (int g = 0; g < 4; ++g) {
                float f[V];
                chunk_unpack<V>(r[g], f);
#pragma unroll
                for (int j = 0; j < V; ++j) sv[(g0 + g) * V + j] += f[j];
              }
            }
          }
          if (a.relu) {
#pragma unroll
            for (int j = 0; j < 32; ++j) sv[j] = relu_nan(sv[j]);
          }
#pragma unroll
          for (int g = 0; g < NG; ++g) {           // 32 channels -> NG coalesced 16-byte stores
            float f[V];
#pragma unroll
            for (int j = 0; j < V; ++j) f[j] = sv[g * V + j];
            // affine epilogue (evaluation, bias-only networks): the output is the next layer's operand -> tf32 grid
            const uint4 pk = affine ? chunk_pack_operand<V>(f) : chunk_pack<V>(f);
            yk[obase + g * a.y_plane_stride] = pk;
            if (gs) {
              // statistics of the value as stored (bf16-rounded in bf16 mode), so BatchNorm normalises exactly what
              // it reads
              chunk_unpack<V>(pk, f);
              if (bnyk) {       // sums of the BatchNorm backward that consumes this gradient
                float yv[V];
                chunk_unpack<V>(bnyk[obase + g * a.y_plane_stride], yv);
                if (bnak) {
                  float av[V];
                  chunk_unpack<V>(bnak[obase + g * a.y_plane_stride], av);
#pragma unroll
                  for (int j = 0; j < V; ++j) f[j] = av[j] > 0.f ? f[j] : 0.f;
                }
#pragma unroll
                for (int j = 0; j < V; ++j) {
                  const int c = c0 + g * V + j;
                  u1[g * V + j] += f[j];
                  u2[g * V + j] += f[j] * (yv[j] - mr_s[c]) * mr_s[a.NB + c];
                }
              } else {
#pragma unroll
                for (int j = 0; j < V; ++j) { u1[g * V + j] += f[j]; u2[g * V + j] += f[j] * f[j]; }
              }
            }
          }
        }
        if (gs) {       // one cross-lane reduction per 32 channels per pass, then one shared atomic per channel per warp
          warp_transpose_reduce32(u1, lane);
          warp_transpose_reduce32(u2, lane);
          if (u1[0] != 0.f || u2[0] != 0.f) {
            atomicAdd(&ss_s[2 * (c0 + lane)], u1[0]);
            atomicAdd(&ss_s[2 * (c0 + lane) + 1], u2[0]);
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&acc_empty[as]);
      pacc.advance(a.NS);
      DBG_ADD(1);
    }
    if (a.stats && cur_group >= 0) flush_stats(cur_group);
  }
  if (a.dbg && lane == 0 && (warp == kProd - 1 || warp == kMmaWarp || warp == kEpiWarp0)) {
    // [cta][role 0..2][4]; slot 3 = the role's whole lifetime (producer = the warp that owns A stage 0)
    const int role = warp == kProd - 1 ? 0 : warp == kMmaWarp ? 1 : 2;
    dbgacc[3] = clock64() - dbg_start;
    for (int i = 0; i < 4; ++i) a.dbg[((long long)blockIdx.x * 3 + role) * 4 + i] = dbgacc[i];
  }

  tc_fence_before();
  __syncthreads();
  if (warp == kMmaWarp) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

}  // namespace

static long long* g_conv_dbg = nullptr;
static int g_conv_dbg_shift = 0;

extern "C" {

// Profiling aid: when set, every conv launch writes per-CTA cycle counters [grid][3 roles][4] (producer: wait a_empty;
// MMA: wait acc_empty, wait a_full, issue; epilogue: wait acc_full, work).  Pass null to switch it off.
void dba_conv_set_debug(long long* buf) { g_conv_dbg = buf; }
void dba_conv_set_debug_shift(int mode) { g_conv_dbg_shift = mode; }

// Geometry helpers shared with the Python side (dba_mod_b200/ops/layout.py mirrors these).
long long dba_conv_rows(int B, int H, int W) {
  const int Wp = W + 1, S = (H + 1) * Wp, lead = Wp + 1;
  long long body = (long long)B * S;
  body = (body + 511) / 512 * 512;            // any T*128 <= 512 rows per pass may overrun into the tail
  return lead + body + Wp + 2;
}

// Y = epilogue(conv(X, Wt)).  See ConvArgs for layouts.  X: [Kx][Cin/8][R][8] with R = dba_conv_rows(B,H,W);
// Wt: per client [taps][Cin/8][Cout][8] (client stride w_client_stride elements, 0 = shared model); Y (and Res):
// [K][Cout/8][Ro][8], Ro = dba_conv_rows(B,Ho,Wo), Ho = H/sub.  scale/shift (nullable together): per (client,
// out-channel) affine = folded BatchNorm (row stride ss_stride).  stats (nullable): rows of (sum, sumsq) pairs,
// accumulated (+=).  x_shared != 0: every client reads client 0's X.
// in_s2d / out_s2d: space-to-depth forms for stride-2 convolutions (see ConvArgs).  With in_s2d, (H, W) is the
// OUTPUT grid of the stride-2 convolution (the grid X' lives on), X' has 4*Cin channels and sub must be 1.  With
// out_s2d, Y gets 4*Cout channels on the (H/2, W/2) grid; Res (if any) stays a plain (H, W) tensor.
struct ConvBnBwd {              // host-side description of the BatchNorm-backward sums mode (see ConvArgs::bn_y)
  const void* y;
  const void* act;
  const float* saved;
  long long saved_stride;
};

static int conv_launch(const void* X, const void* Wt, long long w_client_stride, void* Y, const void* Res,
                       const float* scale, const float* shift, long long ss_stride, float* stats,
                       long long stats_stride, int K, int B, int H, int W, int Cin, int Cout, int ksize, int sub,
                       int relu, int x_shared, int in_s2d, int out_s2d, int vlo, int vhi, void* stream,
                       const ConvBnBwd* bb = nullptr) {
  DBA_REQUIRE(K > 0 && B > 0 && H > 0 && W > 0, "dba_conv_tc: bad sizes K=%d B=%d H=%d W=%d", K, B, H, W);
  DBA_REQUIRE(ksize == 1 || ksize == 3 || ksize == 5 || (ksize == 2 && out_s2d == 3),
              "dba_conv_tc: ksize must be 1, 3 or 5 (2 with out mode 3), got %d", ksize);
  const bool tf32 = g_dba_act_mode == 1;
  const int V = tf32 ? 4 : 8;
  DBA_REQUIRE(Cin % (2 * V) == 0 && Cout % 32 == 0, "dba_conv_tc: Cin (%d) %% %d and Cout (%d) %% 32 must be 0", Cin,
              2 * V, Cout);
  DBA_REQUIRE(sub == 1 || (sub == 2 && H % 2 == 0 && W % 2 == 0), "dba_conv_tc: sub=%d needs even H, W", sub);
  DBA_REQUIRE(scale == nullptr || shift != nullptr, "dba_conv_tc: scale needs shift (shift alone = bias epilogue)");
  DBA_REQUIRE(!(shift && stats), "dba_conv_tc: the affine epilogue and the statistics epilogue are exclusive");
  DBA_REQUIRE(!(in_s2d || out_s2d) || sub == 1, "dba_conv_tc: space-to-depth forms replace sub=2, pass sub=1");
  DBA_REQUIRE(out_s2d != 1 || (H % 2 == 0 && W % 2 == 0), "dba_conv_tc: out_s2d needs even H, W");
  DBA_REQUIRE(out_s2d >= 0 && out_s2d <= 3, "dba_conv_tc: out mode %d", out_s2d);
  DBA_REQUIRE(out_s2d != 3 || (Cout % 4 == 0 && (Cout / 4) % 32 == 0 && Res == nullptr && stats == nullptr),
              "dba_conv_tc: out mode 3 needs Cout = 4*C, C %% 32 == 0, no residual / statistics (Cout=%d)", Cout);
  ConvArgs a{};
  a.in_s2d = in_s2d ? 1 : 0; a.out_s2d = out_s2d;
  a.X = (const uint4*)X; a.Wt = (const uint4*)Wt; a.Y = (uint4*)Y;
  a.Res = (const uint4*)Res; a.scale = scale; a.shift = shift; a.stats = stats;
  a.K = K; a.Cin = Cin; a.Cout = Cout; a.taps = ksize * ksize;
  a.H = H; a.W = W; a.Wp = W + 1; a.S = (H + 1) * a.Wp; a.lead = a.Wp + 1; a.B = B;
  a.relu = relu; a.sub = sub;
  const int osub = out_s2d == 1 ? 2 : sub;
  a.Ho = H / osub; a.Wo = W / osub;
  if (out_s2d >= 2) { a.Ho = 2 * H; a.Wo = 2 * W; }
  a.cgrp = out_s2d == 3 ? Cout / 4 : Cout;
  a.Wpo = a.Wo + 1; a.So = (a.Ho + 1) * a.Wpo; a.leado = a.Wpo + 1;
  const long long R = dba_conv_rows(B, H, W), Ro = dba_conv_rows(B, a.Ho, a.Wo);
  a.x_plane_stride = R;
  a.x_client_stride = x_shared ? 0 : (long long)(in_s2d ? 4 : 1) * (Cin / V) * R;
  a.y_plane_stride = Ro;
  a.y_client_stride = (long long)(out_s2d == 1 ? 4 : 1) * ((out_s2d == 3 ? Cout / 4 : Cout) / V) * Ro;
  a.res_plane_stride = out_s2d == 1 ? R : a.y_plane_stride;
  a.res_client_stride = out_s2d == 1 ? (long long)(Cout / V) * R : a.y_client_stride;
  DBA_REQUIRE(w_client_stride % V == 0, "dba_conv_tc: weight client stride %lld not a multiple of %d", w_client_stride, V);
  a.w_client_stride = w_client_stride / V;          // the caller counts elements
  a.ss_stride = ss_stride;
  a.stats_stride = stats_stride;
  a.KB = Cin >= 8 * V ? 8 * V : Cin;                      // at most 128 bytes of K per row and stage
  if (in_s2d && ksize == 3 && a.KB > 4 * V) a.KB = 4 * V;   // four parity planes per k-block share the stage
  DBA_REQUIRE(Cin % a.KB == 0 && (a.KB == 2 * V || a.KB == 4 * V || a.KB == 8 * V), "dba_conv_tc: unsupported Cin=%d",
              Cin);
  a.halo = ksize == 2 ? a.Wp + 1 : (ksize / 2) * (a.Wp + 1);
  a.ksz = ksize;
  a.vlo = vlo; a.vhi = vhi > 0 ? vhi : (H > W ? H : W);
  if (bb) {
    DBA_REQUIRE(stats && bb->y && bb->saved && sub == 1 && !in_s2d && !out_s2d && !shift && !relu,
                "dba_conv_tc: the BatchNorm-backward sums mode is for plain stride-1 input-gradient convolutions");
    a.bn_y = (const uint4*)bb->y; a.bn_act = (const uint4*)bb->act; a.bn_saved = bb->saved;
    a.bn_saved_stride = bb->saved_stride;
  }
  const int kc = a.KB / V;
  const long long body = (long long)B * a.S;
  // n-block and rows per pass: as many 128-row tiles as fit in TMEM (512 columns); when that leaves SMs idle, first
  // split wide output-channel blocks (halves the weight traffic per CTA), then shorten the passes.
  int NB = Cout <= 256 ? Cout : 256;
  DBA_REQUIRE(Cout % NB == 0, "dba_conv_tc: Cout=%d not a multiple of the n-block %d", Cout, NB);
  // T stays a power of two so that T*128 divides the 512-row tail slack; two accumulator sets (epilogue of pass p
  // overlaps the MMAs of pass p+1) are preferred over longer passes except for NB = 256, where the weight traffic
  // per row would double.
  int T = NB >= 256 ? 2 : 256 / NB;
  if (T > 4) T = 4;
  while (T > 1 && (long long)(T / 2) * kTileM >= body) T /= 2;
  auto n_items = [&](int T_, int NB_) {
    return (long long)K * (Cout / NB_) * ((body + (long long)T_ * kTileM - 1) / ((long long)T_ * kTileM));
  };
  while (n_items(T, NB) < kNumSMs) {
    if (NB >= 128) NB /= 2;
    else if (T > 1) T /= 2;
    else break;
  }
  const size_t budget = 220 * 1024;
  size_t need = 0;
  const int npar = (in_s2d && ksize == 3) ? 4 : 1;
  for (;; T /= 2) {
    a.T = T; a.NB = NB;
    a.rowsA = T * kTileM + (npar == 4 ? 1 : 2) * a.halo;      // parity-plane taps only reach up / left
    const size_t bytesA = (size_t)npar * kc * a.rowsA * 16, bytesB = (size_t)kc * NB * 16;
    const size_t misc = 1024 + 4 * NB * sizeof(float) + 128;
    a.NS = (2 * T * NB <= 512) ? 2 : 1;
    a.wbytes = (unsigned int)((size_t)a.taps * (Cin / V) * NB * 16);
    a.resident = a.wbytes <= 100 * 1024 && 2 * bytesA + a.wbytes + misc <= budget;
    size_t breg;
    if (a.resident) {
      a.SB = 1;
      breg = a.wbytes;
    } else {
      // Streamed weights: as many stages as fit (<= kMaxSB).  A CTA that owns few row tiles is bound by the latency of
      // its own weight stream (K = 1, 256 channels: 295 KB per CTA at 32 KB in flight = 11 B/clk); deeper is faster.
      const int sb_cap = kMaxSB;
      long long sb = budget > 2 * bytesA + misc ? (long long)((budget - 2 * bytesA - misc) / bytesB) : 0;
      a.SB = sb > sb_cap ? sb_cap : (int)sb;
      if (a.SB >= kProd) a.SB -= a.SB % kProd;          // stages are dealt to the producer warps round-robin
      if (a.SB < 2) a.SB = 2;
      breg = a.SB * bytesB;
    }
    // Two A stages.  Measured on B200: a third / fourth stage does not help even the bandwidth-hungry
    // space-to-depth layers (26.97 vs 26.92 ms per 40k evaluations) and costs the MMA-bound layers 8% (more bulk-copy
    // writes under the tensor core's operand reads); shortening the pass to make room for them costs 6% (halo rows).
    a.SA = 2;
    need = a.SA * bytesA + breg + misc;
    if (need <= budget || T == 1) break;
  }
  DBA_REQUIRE(need <= budget, "dba_conv_tc: shared memory need %zu exceeds budget", need);
  a.dbg = g_conv_dbg;
  a.dbg_shift = g_conv_dbg_shift;
  a.magicS = ((1ULL << 40) + (unsigned long long)a.S - 1) / (unsigned long long)a.S;
  a.magicWp = ((1ULL << 40) + (unsigned long long)a.Wp - 1) / (unsigned long long)a.Wp;
  a.num_pass = (int)((body + (long long)a.T * kTileM - 1) / ((long long)a.T * kTileM));
  const long long items = (long long)K * a.num_pass * (Cout / a.NB);
  static bool attr_set[64] = {false};
  int dev = 0;
  DBA_CHECK_CUDA(cudaGetDevice(&dev));
  if (dev < 64 && !attr_set[dev]) {
    DBA_CHECK_CUDA(cudaFuncSetAttribute(conv_tc_kernel<1, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)budget));
    DBA_CHECK_CUDA(cudaFuncSetAttribute(conv_tc_kernel<2, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)budget));
    DBA_CHECK_CUDA(cudaFuncSetAttribute(conv_tc_kernel<4, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)budget));
    DBA_CHECK_CUDA(cudaFuncSetAttribute(conv_tc_kernel<1, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)budget));
    DBA_CHECK_CUDA(cudaFuncSetAttribute(conv_tc_kernel<2, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)budget));
    DBA_CHECK_CUDA(cudaFuncSetAttribute(conv_tc_kernel<4, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)budget));
    attr_set[dev] = true;
  }
  if (need < 120 * 1024) need = 120 * 1024;    // one CTA per SM: each CTA allocates all 512 TMEM columns
  const int grid = items < kNumSMs ? (int)items : kNumSMs;
  const int kc2 = kc / 2;                      // MMAs (32 bytes of K each) per tap and k-block
  const cudaStream_t st = (cudaStream_t)stream;
  if (tf32) {
    if (kc2 == 4) launch_pdl(conv_tc_kernel<4, true>, grid, kThreads, need, st, a);
    else if (kc2 == 2) launch_pdl(conv_tc_kernel<2, true>, grid, kThreads, need, st, a);
    else launch_pdl(conv_tc_kernel<1, true>, grid, kThreads, need, st, a);
  } else {
    if (kc2 == 4) launch_pdl(conv_tc_kernel<4, false>, grid, kThreads, need, st, a);
    else if (kc2 == 2) launch_pdl(conv_tc_kernel<2, false>, grid, kThreads, need, st, a);
    else launch_pdl(conv_tc_kernel<1, false>, grid, kThreads, need, st, a);
  }
  DBA_COUNT_LAUNCH(1);
  DBA_CHECK_LAUNCH();
  return DBA_OK;
}

int dba_conv_tc_ex(const void* X, const void* Wt, long long w_client_stride, void* Y, const void* Res,
                   const float* scale, const float* shift, long long ss_stride, float* stats,
                   long long stats_stride, int K, int B, int H, int W, int Cin, int Cout, int ksize, int sub,
                   int relu, int x_shared, int in_s2d, int out_s2d, void* stream) {
  return conv_launch(X, Wt, w_client_stride, Y, Res, scale, shift, ss_stride, stats, stats_stride, K, B, H, W, Cin, Cout,
                     ksize, sub, relu, x_shared, in_s2d, out_s2d, 0, 0, stream);
}

// 'Valid' (unpadded) convolution as the centre crop of the padded one: only outputs with vlo <= h, w < vhi are stored
// (at their position on the input grid); everything else as dba_conv_tc_ex.  With ksize = 5 the operand block reaches
// 2*(W+1)+2 rows beyond a pass on either side: X needs (W+2) zero rows of slack before its first and after its last
// plane (see NativeMnist._zbuf).
int dba_conv_tc_valid(const void* X, const void* Wt, long long w_client_stride, void* Y, const void* Res,
                      const float* scale, const float* shift, long long ss_stride, float* stats,
                      long long stats_stride, int K, int B, int H, int W, int Cin, int Cout, int ksize, int relu,
                      int x_shared, int vlo, int vhi, void* stream) {
  return conv_launch(X, Wt, w_client_stride, Y, Res, scale, shift, ss_stride, stats, stats_stride, K, B, H, W, Cin, Cout,
                     ksize, 1, relu, x_shared, 0, 0, vlo, vhi, stream);
}

int dba_conv_tc_strided(const void* X, const void* Wt, long long w_client_stride, void* Y, const void* Res,
                        const float* scale, const float* shift, long long ss_stride, float* stats,
                        long long stats_stride, int K, int B, int H, int W, int Cin, int Cout, int ksize, int sub,
                        int relu, int x_shared, void* stream) {
  return dba_conv_tc_ex(X, Wt, w_client_stride, Y, Res, scale, shift, ss_stride, stats, stats_stride, K, B, H, W, Cin,
                        Cout, ksize, sub, relu, x_shared, 0, 0, stream);
}

// Input-gradient convolution (same operands as dba_conv_tc_strided without affine / ReLU / sub-sampling) that ALSO
// accumulates the reduction sums of the BatchNorm backward consuming its output g = conv(dY, Wt) (+ Res):
//   sums[k][2c] += sum g',  sums[k][2c + 1] += sum g' * (bn_y - mean_c) * rstd_c,  g' = bn_act > 0 ? g : 0 (bn_act may be
// null: no ReLU).  bn_y / bn_act have the output's geometry; saved = (mean, rstd) pairs of that layer, row stride
// ld_saved; sums must be zero on entry.  dba_bn_backward_split(..., sums_ready = 1) then skips its reduce kernel.
int dba_conv_tc_dgrad_bn(const void* dY, const void* Wt, long long w_client_stride, void* G, const void* Res,
                         const void* bn_y, const void* bn_act, const float* saved, long long ld_saved, float* sums,
                         long long ld_sums, int K, int B, int H, int W, int Cin, int Cout, int ksize, void* stream) {
  const ConvBnBwd bb{bn_y, bn_act, saved, ld_saved};
  return conv_launch(dY, Wt, w_client_stride, G, Res, nullptr, nullptr, 0, sums, ld_sums, K, B, H, W, Cin, Cout, ksize, 1,
                     0, 0, 0, 0, 0, 0, stream, &bb);
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
