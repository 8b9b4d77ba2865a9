// Convolution weight gradient on tcgen05 tensor cores for K client replicas (K2 "wgrad" in SURVEY §2.4; the autograd
// of nn.Conv2d that the reference gets from cuDNN, image_train.py:91,209).
//
//   dW[k][co][tap][ci] += sum_rows  dY[k][co][row] * X[k][ci][row + shift(tap)]
//
// Both operands are consumed in the SAME planar-8 / shared-halo layout the forward pass uses: a [c8][row][8] block
// is the SWIZZLE_NONE canonical *MN-major* UMMA operand (8 consecutive channels per 16 bytes, reduction rows at a
// 16-byte pitch), so dY is the A operand (M = output channels), X the B operand (N = input channels), the reduction
// axis is the pixel-row axis, and a 3x3 tap is once more just a row-shifted descriptor into the same X block.
// Per stage a CTA loads 128 rows of dY (all M planes) and 128+halo rows of X (all N planes) with bulk async copies and
// issues  taps_in_group x 8  MMAs (M=128, N=Cin, K=16) into taps_in_group accumulators living side by side in TMEM.
// Work item = (client, 128-wide block of Cout, tap group, slice of the row range); partial results are added into
// the fp32 gradient slab with vector reductions (red.global.add.v4.f32), rows of Cin being contiguous thanks to the
// [Cout][tap][Cin] internal weight order.
//
// tf32 mode (fp32 activations): tcgen05 accepts MN-major 32-bit operands only in the 128B-swizzle / 32-byte-atom
// shared-memory layout, and TMA loads with CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B faulted on this driver (DESIGN.md
// section 6b), so the weight gradient keeps bf16 MN-major operands but loses no precision against TF32: its operands
// are the "split" copies the producing kernels store next to every fp32 tensor, x = hi + lo with hi = bf16(x),
// lo = bf16(x - hi) (16 mantissa bits, TF32 has 10), and every k-step issues three MMAs into the same accumulator:
// hi*hi + hi*lo + lo*hi (the dropped lo*lo term is below 2^-17 relative).
#include "common.cuh"

using namespace dba;

namespace {

// Two producer warps (a warp has one bulk-copy instruction in flight): stage s is loaded by warp
// s % kProdW.  Warps: kProdW producers, one MMA issuer, four epilogue warps (one per TMEM lane quadrant, warp & 3).
constexpr int kProdW = 2;
constexpr int kMmaWarpW = kProdW;
constexpr int kThreadsW = 32 * (kProdW + 1 + 4);

struct WgArgs {
  const __nv_bfloat16* dY;     // [K][Cout/8][R][8]  (zero on halo rows and on padding samples)
  const __nv_bfloat16* X;      // [K][Cin/8][R][8]
                               // split mode: [K][2][C/8][R][8], the lo set dy_lo / x_lo elements after the hi set
  long long dy_lo, x_lo;
  float* G;                    // [K][ldg] gradient rows; this layer's weights at w_off, layout [Cout][taps][cin_real]
  long long dy_plane, dy_client, x_plane, x_client, ldg;
  int w_off;
  int K, Cout, Cin, cin_real, cout_real, taps, Wp, lead, halo;
  int body_chunks;             // number of 128-row chunks covering the body rows
  int nsplit;                  // row-range slices
  int tpg;                     // taps per group
  int ngroups;                 // tap groups
  int nmb;                     // Cout blocks of 128
  int N;                       // input channels per accumulator block (<= 256)
  int nnblk;                   // Cin / N
  int stages;
  int ksz;                     // 1, 3 or 5 (taps = ksz*ksz, centred)
  // Destination index of (co, tap, ci): co*g_co + tap*g_tap + ci*g_ci.  Convolutions: [Cout][taps][Cin] ->
  // (taps*cin_real, cin_real, 1).  A linear layer over a flattened C x h x w map, run as a 1x1 convolution over
  // channels n = pos*cpad + c: ci -> (pos, c) with c < cgrp_real, destination (in_features, 1 per pos, npos per c).
  int cgrp_pad, cgrp_real, npos;
  // s2d: X is the space-to-depth form (4*Cin channels, parity-major, on dY's grid) of a 3x3 stride-2 convolution's input;
  // tap (dh, dw) reads parity plane (dh != 1, dw != 1) shifted by (dh == 0 ? -1 : 0, dw == 0 ? -1 : 0).
  int s2d;
  // vs = 3 ("vertical stack", 3x3 convolutions with Cout <= 32): the M = 128 rows of the A operand hold THREE copies of
  // the 32 output channels, copy dh loaded from dY shifted by -(dh - 1) image rows, so one MMA per horizontal tap dw
  // (a +-1 row shift of the X view) yields the gradients of the three taps (dh, dw) in TMEM lanes 32*dh + co:
  // 3 MMAs per k-step instead of 9.  (M cannot be smaller than 128 without halving the tensor rate, and the A fetch of
  // 4 KB per MMA is what bounds the narrow layers - so the idle 96 rows are given work.)
  int vs;
};

__device__ __forceinline__ void red_add_v4(float* addr, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

template <bool SPLIT>
__global__ void __launch_bounds__(kThreadsW, 1) wgrad_tc_kernel(const WgArgs a) {
  extern __shared__ __align__(128) uint8_t smem[];
  constexpr int kRowsK = SPLIT ? 64 : 128;                  // reduction rows per pipeline stage
  constexpr int NSET = SPLIT ? 2 : 1;                       // operand sets per stage (hi, lo)
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int mplanes = 16;                                   // M = 128 -> 16 planes of 8 output channels
  const int nplanes = a.N >> 3;
  const uint32_t pitchA = kRowsK * 16u;
  const uint32_t setA = mplanes * pitchA;                   // one operand set of A
  const uint32_t bytesA = NSET * setA;
  const uint32_t rowsB = kRowsK + 2 * a.halo;
  const uint32_t pitchB = rowsB * 16u;
  const int npar = a.s2d ? 4 : 1;
  const uint32_t setB = (uint32_t)(npar * nplanes) * pitchB;
  const uint32_t bytesB = NSET * setB;
  uint8_t* smA = smem;
  uint8_t* smB = smA + (size_t)a.stages * bytesA;
  uint64_t* full = reinterpret_cast<uint64_t*>(smB + (size_t)a.stages * bytesB);
  uint64_t* empty = full + a.stages;
  uint64_t* acc_full = empty + a.stages;
  uint64_t* acc_empty = acc_full + 1;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + 1);
  int* tap_shift = reinterpret_cast<int*>(tmem_slot + 4);    // [taps] row shift of every tap, computed once
  if (threadIdx.x >= 96 && threadIdx.x < 96 + a.taps) {
    const int tap = threadIdx.x - 96;
    int shift = 0;
    if (a.taps > 1) {
      const int th = tap / a.ksz;
      shift = (th - (a.ksz >> 1)) * a.Wp + (tap - th * a.ksz - (a.ksz >> 1));
    }
    if (a.s2d) {
      const int dh = tap / 3, dw = tap - dh * 3;
      shift = (dh == 0 ? -a.Wp : 0) + (dw == 0 ? -1 : 0)
              + (int)((uint32_t)((dh != 1) * 2 + (dw != 1)) * (uint32_t)nplanes * (pitchB >> 4));
    }
    tap_shift[tap] = shift;
  }

  if (threadIdx.x == 0) {
    for (int i = 0; i < a.stages; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
    mbar_init(acc_full, 1);
    mbar_init(acc_empty, 4);
    mbar_fence_init();
  }
  pdl_trigger();
  if (warp == kMmaWarpW) {
    tmem_alloc(tmem_slot, 512);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();

  const int items = a.K * a.nmb * a.ngroups * a.nnblk * a.nsplit;
  // real planes of dY available for this M block (the rest of the 16 planes stays unloaded: rows >= Cout of D are
  // never read back)
  if (warp < kProdW) {
    // every producer warp walks all stages and loads the ones it owns; lane 0 waits for the free stage and posts the
    // byte count, all lanes issue the plane copies
    int st = 0;
    uint32_t ph = 0;
    for (int item = blockIdx.x; item < items; item += gridDim.x) {
      const int sp = item % a.nsplit;
      const int nbk = (item / a.nsplit) % a.nnblk;
      const int mb = (item / (a.nsplit * a.nnblk * a.ngroups)) % a.nmb;
      const int k = item / (a.nsplit * a.nnblk * a.ngroups * a.nmb);
      const int c0 = (int)((long long)a.body_chunks * sp / a.nsplit);
      const int c1 = (int)((long long)a.body_chunks * (sp + 1) / a.nsplit);
      const int mreal = min(16, (a.Cout >> 3) - mb * 16);
      const __nv_bfloat16* dyk = a.dY + (long long)k * a.dy_client + (long long)(mb * 16) * a.dy_plane;
      const __nv_bfloat16* xk = a.X + (long long)k * a.x_client + (long long)(nbk * nplanes) * a.x_plane;
      for (int ch = c0; ch < c1; ++ch) {
        const int mslots = a.vs ? a.vs * mreal : mreal;          // A planes filled per operand set
        if (st % kProdW != warp) {                               // another producer warp's stage
          if (++st == a.stages) { st = 0; ph ^= 1u; }
          continue;
        }
        if (lane == 0) {
          mbar_wait(&empty[st], ph ^ 1u);
          mbar_arrive_expect_tx(&full[st], NSET * (uint32_t)mslots * pitchA + bytesB);
        }
        __syncwarp();
        const long long row = (long long)a.lead + (long long)ch * kRowsK;
        const int per_set = mslots + npar * nplanes;
        for (int pp = lane; pp < NSET * per_set; pp += 32) {
          const int set = pp >= per_set ? 1 : 0, p = pp - set * per_set;           // 0 = hi (or the only) set, 1 = lo
          if (p < mslots) {
            const int dh = a.vs ? p / mreal : 1, pl = a.vs ? p - dh * mreal : p;   // copy dh: dY rows shifted by -(dh-1)*Wp
            const int slot = a.vs ? dh * 4 + pl : p;               // copy dh starts at row 32*dh of the A operand
            bulk_g2s(smA + (size_t)st * bytesA + (size_t)set * setA + (size_t)slot * pitchA,
                     dyk + (long long)set * a.dy_lo + (long long)pl * a.dy_plane + (row - (long long)(dh - 1) * a.Wp) * 8,
                     pitchA, &full[st]);
          } else {
            const int q = p - mslots, par = q / nplanes, j = q - par * nplanes;      // parity group, plane inside it
            bulk_g2s(smB + (size_t)st * bytesB + (size_t)set * setB + (size_t)q * pitchB,
                     xk + (long long)set * a.x_lo + ((long long)par * (a.Cin >> 3) + j) * a.x_plane + (row - a.halo) * 8,
                     pitchB, &full[st]);
          }
        }
        if (++st == a.stages) { st = 0; ph ^= 1u; }
      }
    }
  } else if (warp == kMmaWarpW) {
    if (elect_one()) {
      const uint32_t idesc = umma_idesc_bf16(128, a.N, 1, 1);       // both operands MN-major
      int st = 0, acc_it = 0;
      uint32_t ph = 0;
      for (int item = blockIdx.x; item < items; item += gridDim.x, ++acc_it) {
        const int sp = item % a.nsplit;
        const int grp = (item / (a.nsplit * a.nnblk)) % a.ngroups;
        const int c0 = (int)((long long)a.body_chunks * sp / a.nsplit);
        const int c1 = (int)((long long)a.body_chunks * (sp + 1) / a.nsplit);
        const int t0 = grp * a.tpg, t1 = min(a.taps, t0 + a.tpg);
        mbar_wait(acc_empty, (acc_it & 1) ^ 1u);
        tc_fence_after();
        for (int ch = c0; ch < c1; ++ch) {
          const int s = st;
          mbar_wait(&full[s], ph);
          if (++st == a.stages) { st = 0; ph ^= 1u; }
          tc_fence_after();
          // base descriptors once per stage; per MMA only the 14-bit start-address field moves (16-byte units)
          const uint64_t dA0 = umma_desc_noswz(smem_u32(smA + (size_t)s * bytesA), 128, pitchA);
          const uint64_t dB0 = umma_desc_noswz(smem_u32(smB + (size_t)s * bytesB), 128, pitchB) + (uint64_t)a.halo;
          const uint32_t acc0 = ch != c0 ? 1u : 0u;
          for (int tap = t0; tap < t1; ++tap) {
            const int shift = a.vs ? tap - 1 : tap_shift[tap];       // vs: "tap" = horizontal tap dw, all dh at once
            const uint32_t d_tmem = tmem_base + (uint32_t)((tap - t0) * a.N);
            const uint64_t dB1 = dB0 + (uint64_t)(int64_t)shift;
            umma_f16(d_tmem, dA0, dB1, idesc, acc0);
            if (SPLIT) {                            // + hi*lo + lo*hi
              umma_f16(d_tmem, dA0, dB1 + (uint64_t)(setB >> 4), idesc, 1u);
              umma_f16(d_tmem, dA0 + (uint64_t)(setA >> 4), dB1, idesc, 1u);
            }
#pragma unroll
            for (int ks = 1; ks < kRowsK / 16; ++ks) {
              const uint64_t da = dA0 + (uint64_t)(ks * 16), db = dB1 + (uint64_t)(ks * 16);
              umma_f16(d_tmem, da, db, idesc, 1u);
              if (SPLIT) {
                umma_f16(d_tmem, da, db + (uint64_t)(setB >> 4), idesc, 1u);
                umma_f16(d_tmem, da + (uint64_t)(setA >> 4), db, idesc, 1u);
              }
            }
          }
          umma_commit(&empty[s]);
        }
        umma_commit(acc_full);
      }
    }
  } else {
    const int q = warp & 3;
    int acc_it = 0;
    for (int item = blockIdx.x; item < items; item += gridDim.x, ++acc_it) {
      const int sp = item % a.nsplit;
      const int nbk = (item / a.nsplit) % a.nnblk;
      const int grp = (item / (a.nsplit * a.nnblk)) % a.ngroups;
      const int mb = (item / (a.nsplit * a.nnblk * a.ngroups)) % a.nmb;
      const int k = item / (a.nsplit * a.nnblk * a.ngroups * a.nmb);
      const int c0 = (int)((long long)a.body_chunks * sp / a.nsplit);
      const int c1 = (int)((long long)a.body_chunks * (sp + 1) / a.nsplit);
      const int t0 = grp * a.tpg, t1 = min(a.taps, t0 + a.tpg);
      mbar_wait(acc_full, acc_it & 1);
      tc_fence_after();
      const int co = mb * 128 + q * 32 + lane;
      float* gk = a.G + (long long)k * a.ldg + a.w_off;
      const bool vec = ((a.cin_real | a.w_off) & 3) == 0 && (a.ldg & 3) == 0 && a.npos == 0;
      if (c1 > c0) {
        for (int tap = t0; tap < t1; ++tap) {
          // vs: lane quadrant q holds vertical tap dh = q (rows 32*dh + co), accumulator `tap` is horizontal tap dw
          const int cow = a.vs ? lane : co;
          const int tapw = a.vs ? q * 3 + tap : tap;
          if (a.vs && q >= a.vs) break;
          float* dst = gk + ((long long)cow * a.taps + tapw) * a.cin_real + nbk * a.N;
          const int cin_left = a.cin_real - nbk * a.N;
          for (int n0 = 0; n0 < a.N; n0 += 16) {
            uint32_t v[16];
            tmem_ld_32x32b_x16(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)((tap - t0) * a.N + n0), v);
            tmem_ld_wait();
            if (cow < a.cout_real) {
              if (vec) {
#pragma unroll
                for (int j = 0; j < 16; j += 4)
                  if (n0 + j < cin_left)
                    red_add_v4(dst + n0 + j, __uint_as_float(v[j]), __uint_as_float(v[j + 1]),
                               __uint_as_float(v[j + 2]), __uint_as_float(v[j + 3]));
              } else if (a.npos == 0) {
#pragma unroll
                for (int j = 0; j < 16; ++j)
                  if (n0 + j < cin_left) atomicAdd(dst + n0 + j, __uint_as_float(v[j]));
              } else {          // flattened-map linear layer: channel n = pos*cgrp_pad + c -> [co][c*npos + pos]
#pragma unroll
                for (int j = 0; j < 16; ++j) {
                  const int n = nbk * a.N + n0 + j;
                  const int pos = n / a.cgrp_pad, c = n - pos * a.cgrp_pad;
                  if (c < a.cgrp_real && pos < a.npos)
                    atomicAdd(gk + (long long)co * (a.cgrp_real * a.npos) + c * a.npos + pos, __uint_as_float(v[j]));
                }
              }
            }
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(acc_empty);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == kMmaWarpW) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

}  // namespace

extern "C" {

// dY: [K][Cout/8][R][8] (zero outside the stored outputs), X: [K][Cin/8][R][8], both on the B x H x W grid
// (R = dba_conv_rows(B,H,W)).  Adds the weight gradient of a ksize x ksize convolution (ksize 1 or 3, pad ksize/2) into
// G[k*ldg + w_off + (co*taps + tap)*cin_real + ci] for co < cout_real, ci < cin_real (the gradient slab is zeroed by
// the caller once per step).  Cin / Cout are the padded channel counts of the planar buffers.
static int g_wgrad_vstack = 1;
void dba_set_wgrad_vstack(int on) { g_wgrad_vstack = on ? 1 : 0; }     // A/B switch (tests compare both forms)

static int wgrad_launch(const void* dY, const void* X, float* G, long long ldg, int w_off, int K, int B, int H, int W,
                        int Cin, int Cout, int cin_real, int cout_real, int ksize, int cgrp_pad, int cgrp_real,
                        int npos, void* stream, long long x_client_stride = 0, int s2d = 0) {
  DBA_REQUIRE(K > 0 && B > 0 && (ksize == 1 || ksize == 3 || ksize == 5), "dba_wgrad_tc: K=%d B=%d ksize=%d", K, B, ksize);
  // tf32 mode: dY / X are the bf16 hi/lo split copies ([K][2][C/8][R][8]) of the fp32 tensors
  const bool split = g_dba_act_mode == 1;
  const int kRowsK = split ? 64 : 128, nmax = split ? 128 : 256;
  DBA_REQUIRE(Cin % 16 == 0 && Cout % 8 == 0, "dba_wgrad_tc: Cin=%d Cout=%d", Cin, Cout);
  WgArgs a{};
  a.dY = (const __nv_bfloat16*)dY; a.X = (const __nv_bfloat16*)X; a.G = G; a.ldg = ldg; a.w_off = w_off;
  a.K = K; a.Cout = Cout; a.Cin = Cin; a.cin_real = cin_real; a.cout_real = cout_real; a.taps = ksize * ksize;
  a.Wp = W + 1; a.lead = a.Wp + 1; a.halo = (ksize / 2) * (a.Wp + 1);
  a.ksz = ksize; a.cgrp_pad = cgrp_pad; a.cgrp_real = cgrp_real; a.npos = npos; a.s2d = s2d;
  DBA_REQUIRE(!s2d || ksize == 3, "dba_wgrad_tc: the space-to-depth form is for 3x3 stride-2 convolutions");
  const long long S = (long long)(H + 1) * a.Wp;
  long long body = (long long)B * S;
  body = (body + 511) / 512 * 512;
  const long long R = a.lead + body + a.Wp + 2;
  const int nset = split ? 2 : 1;
  a.dy_plane = R * 8; a.dy_client = (long long)nset * (Cout / 8) * R * 8;
  a.x_plane = R * 8;
  a.x_client = x_client_stride ? x_client_stride : (long long)nset * (s2d ? 4 : 1) * (Cin / 8) * R * 8;
  a.dy_lo = a.dy_client / 2; a.x_lo = a.x_client / 2;       // the lo set follows the hi set inside a client's block
  a.body_chunks = (int)(((long long)B * S + kRowsK - 1) / kRowsK);
  a.nmb = (Cout + 127) / 128;
  // accumulator width: the largest multiple of 16 dividing Cin that fits the stage (four parity groups share it in the
  // space-to-depth form)
  const int ncap = s2d ? nmax / 4 : nmax;
  a.N = Cin <= ncap ? Cin : ncap;
  while (a.N > 16 && Cin % a.N != 0) a.N -= 16;
  a.nnblk = Cin / a.N;
  a.tpg = 512 / a.N;
  if (a.tpg > a.taps) a.tpg = a.taps;
  a.ngroups = (a.taps + a.tpg - 1) / a.tpg;
  a.tpg = (a.taps + a.ngroups - 1) / a.ngroups;            // balance the groups (e.g. 9 taps, 4 max -> 3+3+3)
  // vertical stack (see WgArgs::vs): three shifted copies of <= 32 output channels fill the 128 rows of A
  a.vs = (g_wgrad_vstack && ksize == 3 && !s2d && npos == 0 && Cout <= 32 && 3 * a.N <= 512) ? 3 : 0;
  if (a.vs) { a.tpg = 3; a.ngroups = 1; }
  const int base_items = K * a.nmb * a.ngroups * a.nnblk;
  int nsplit = (2 * kNumSMs + base_items - 1) / base_items;
  if (nsplit > a.body_chunks) nsplit = a.body_chunks;
  if (nsplit < 1) nsplit = 1;
  a.nsplit = nsplit;
  const size_t bytesA = (size_t)nset * 16 * kRowsK * 16;
  const size_t bytesB = (size_t)nset * (s2d ? 4 : 1) * (a.N / 8) * (kRowsK + 2 * a.halo) * 16;
  a.stages = 3;
  size_t need = a.stages * (bytesA + bytesB) + 1024;
  if (need > 220 * 1024) { a.stages = 2; need = a.stages * (bytesA + bytesB) + 1024; }
  DBA_REQUIRE(need <= 220 * 1024, "dba_wgrad_tc: shared memory need %zu too large", need);
  static bool attr_set[64] = {false};
  int dev = 0;
  DBA_CHECK_CUDA(cudaGetDevice(&dev));
  if (dev < 64 && !attr_set[dev]) {
    DBA_CHECK_CUDA(cudaFuncSetAttribute(wgrad_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024));
    DBA_CHECK_CUDA(cudaFuncSetAttribute(wgrad_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024));
    attr_set[dev] = true;
  }
  if (need < 120 * 1024) need = 120 * 1024;
  const int items = base_items * nsplit;
  const int grid = items < kNumSMs ? items : kNumSMs;
  if (split) launch_pdl(wgrad_tc_kernel<true>, grid, kThreadsW, need, (cudaStream_t)stream, a);
  else launch_pdl(wgrad_tc_kernel<false>, grid, kThreadsW, need, (cudaStream_t)stream, a);
  DBA_COUNT_LAUNCH(1);
  DBA_CHECK_LAUNCH();
  return DBA_OK;
}

int dba_wgrad_tc(const void* dY, const void* X, float* G, long long ldg, int w_off, int K, int B, int H, int W,
                 int Cin, int Cout, int cin_real, int cout_real, int ksize, void* stream) {
  return wgrad_launch(dY, X, G, ldg, w_off, K, B, H, W, Cin, Cout, cin_real, cout_real, ksize, 0, 0, 0, stream);
}

// x_client_stride (elements): distance between the clients' X tensors when X is a channel slice of a wider tensor (the
// parity-0 planes of a space-to-depth activation: weight gradient of a 1x1 stride-2 convolution on its output grid).
int dba_wgrad_tc_ex(const void* dY, const void* X, long long x_client_stride, float* G, long long ldg, int w_off, int K,
                    int B, int H, int W, int Cin, int Cout, int cin_real, int cout_real, int ksize, void* stream) {
  return wgrad_launch(dY, X, G, ldg, w_off, K, B, H, W, Cin, Cout, cin_real, cout_real, ksize, 0, 0, 0, stream,
                      x_client_stride);
}

// Weight gradient of a 3x3 stride-2 convolution on its OUTPUT grid (H, W): dY compact, X the space-to-depth form of the
// layer's input ([K][4*Cin/8][R][8], R = dba_conv_rows(B, H, W)) - a quarter of the full-grid work.
int dba_wgrad_tc_s2d(const void* dY, const void* X, float* G, long long ldg, int w_off, int K, int B, int H, int W,
                     int Cin, int Cout, int cin_real, int cout_real, void* stream) {
  return wgrad_launch(dY, X, G, ldg, w_off, K, B, H, W, Cin, Cout, cin_real, cout_real, 3, 0, 0, 0, stream, 0, 1);
}

// Weight gradient of a linear layer applied to a flattened [c_real][npos] feature map (torch's x.view(-1, C*h*w)) that
// the native path keeps as channels n = pos*c_pad + c of a 1x1 grid: G[w_off + co*(c_real*npos) + c*npos + pos].
int dba_wgrad_linear_chw(const void* dY, const void* X, float* G, long long ldg, int w_off, int K, int B, int Cout,
                         int cout_real, int c_pad, int c_real, int npos, void* stream) {
  return wgrad_launch(dY, X, G, ldg, w_off, K, B, 1, 1, c_pad * npos, Cout, c_pad * npos, cout_real, 1, c_pad, c_real,
                      npos, stream);
}

}  // extern "C"
