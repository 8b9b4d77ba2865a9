// Shared device helpers for the sm_100a kernels: error handling, warp reductions, and thin wrappers around the
// Blackwell PTX used here (mbarrier, cp.async.bulk (UBLKCP), tcgen05.{alloc,mma,commit,ld}, UMMA descriptors).
// No CUTLASS dependency: everything is inline PTX so the library builds with plain nvcc.
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>

#define DBA_OK 0
#define DBA_ERR_ARG 1
#define DBA_ERR_CUDA 2
#define DBA_ERR_LIMIT 3

// Last error text, readable through dba_last_error().
extern "C" const char* dba_last_error();
void dba_set_error(const char* fmt, ...);

#define DBA_CHECK_LAUNCH()                                                                     \
  do {                                                                                         \
    cudaError_t e__ = cudaGetLastError();                                                      \
    if (e__ != cudaSuccess) {                                                                  \
      dba_set_error("%s:%d CUDA launch error: %s", __FILE__, __LINE__, cudaGetErrorString(e__)); \
      return DBA_ERR_CUDA;                                                                     \
    }                                                                                          \
  } while (0)

#define DBA_CHECK_CUDA(x)                                                                      \
  do {                                                                                         \
    cudaError_t e__ = (x);                                                                     \
    if (e__ != cudaSuccess) {                                                                  \
      dba_set_error("%s:%d %s: %s", __FILE__, __LINE__, #x, cudaGetErrorString(e__));           \
      return DBA_ERR_CUDA;                                                                     \
    }                                                                                          \
  } while (0)

#define DBA_REQUIRE(cond, ...)                                                                 \
  do {                                                                                         \
    if (!(cond)) {                                                                             \
      dba_set_error(__VA_ARGS__);                                                              \
      return DBA_ERR_ARG;                                                                      \
    }                                                                                          \
  } while (0)

// Counts kernel launches made by this library (bench.py reports it as gpu_launches).
extern unsigned long long g_dba_launches;
// (atomic: entry points may be called from several host threads, e.g. dba_rows_sqdist_noalloc from a monitor thread)
#define DBA_COUNT_LAUNCH(n) ((void)__atomic_fetch_add(&g_dba_launches, (unsigned long long)(n), __ATOMIC_RELAXED))

// Activation element mode of the tensor-core network kernels (conv / wgrad / BatchNorm / pooling / input stage / head /
// weight packing).  0: bf16 activations, 8 channels per 16-byte chunk ("planar-8"), tcgen05.mma.kind::f16.
// 1: fp32 activations, 4 channels per 16-byte chunk ("planar-4"), tcgen05.mma.kind::tf32 - the precision class of the
// reference (fp32 tensors, TF32 convolutions: cuDNN's allow_tf32 default).  Both modes share ONE byte geometry: a
// "row" is a 16-byte chunk, planes are arrays of rows, UMMA core matrices are 8 rows x 16 bytes, and one MMA consumes
// 32 bytes of K per row, so every descriptor, pitch and shift below is mode-independent; only the number of channels
// per chunk (V) and the MMA kind change.  Set once by the executor (dba_set_act_dtype); a library-wide mode like a
// BLAS math mode, not a per-call argument.
extern int g_dba_act_mode;

// Programmatic dependent launch (1 = on): consecutive kernels of a stream / captured graph are launched with
// cudaLaunchAttributeProgrammaticStreamSerialization, so kernel N+1's blocks may be scheduled - and run their prologue
// (shared-memory carve-up, mbarrier init, TMEM allocation, tap tables) - while kernel N drains.
extern int g_dba_pdl;

namespace dba {

constexpr int kNumSMs = 148;

// Every kernel launched through launch_pdl() calls pdl_trigger() first (lets the NEXT kernel's blocks be scheduled as
// soon as all of this kernel's blocks have started) and pdl_wait() before it touches global memory (returns when the
// PREVIOUS kernel has completed and its writes are visible).  Both are no-ops for a kernel launched the plain way.
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st,
                              Args... args) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = g_dba_pdl ? 1 : 0;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

// 16-byte chunk <-> V floats (V = 8: bf16, V = 4: fp32)
template <int V> __device__ __forceinline__ void chunk_unpack(const uint4 u, float (&f)[V]);
template <int V> __device__ __forceinline__ uint4 chunk_pack(const float (&f)[V]);

// ReLU that propagates NaN like torch.relu (fmaxf(NaN, 0) would return 0): a model broken by a x100 scaled update - a
// negative running variance after extrapolation - must evaluate to NaN logits (arg-max = class 0) here as it does in the
// reference, not to a silently repaired network.
__device__ __forceinline__ float relu_nan(float v) { return v < 0.f ? 0.f : v; }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Block-wide sum; result valid in thread 0 (and broadcast through smem to all when bcast).  blockDim.x <= 1024.
template <typename T>
__device__ __forceinline__ T block_sum(T v, T* smem32, bool bcast = false) {
  v = warp_sum(v);
  int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  __syncthreads();
  if (lane == 0) smem32[w] = v;
  __syncthreads();
  int nw = (blockDim.x + 31) >> 5;
  v = (threadIdx.x < nw) ? smem32[threadIdx.x] : T(0);
  if (w == 0) v = warp_sum(v);
  if (bcast) {
    __syncthreads();
    if (threadIdx.x == 0) smem32[0] = v;
    __syncthreads();
    v = smem32[0];
  }
  return v;
}

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

__device__ __forceinline__ bool elect_one() {
  uint32_t pred = 0;
  asm volatile(
      "{\n\t.reg .pred P1;\n\t"
      "elect.sync _|P1, 0xffffffff;\n\t"
      "selp.b32 %0, 1, 0, P1;\n\t}\n"
      : "=r"(pred));
  return pred != 0;
}

// ------------------------------------------------------------------------------------------------ mbarrier
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n\t.reg .pred P1;\n\t"
      "WAIT_%=:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n\t"
      "@P1 bra DONE_%=;\n\t"
      "bra WAIT_%=;\n\t"
      "DONE_%=:\n\t}\n" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}

// ------------------------------------------------------------------------------- bulk async copy (TMA unit, 1-D)
// global -> shared, completion signalled on an mbarrier (SASS: UBLKCP).  bytes % 16 == 0, both addresses 16B aligned.
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(dst_smem)),
               "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// ------------------------------------------------------------------------------------------------- tcgen05
__device__ __forceinline__ void tmem_alloc(uint32_t* slot_in_smem, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(slot_in_smem)),
               "r"(ncols)
               : "memory");
}
__device__ __forceinline__ void tmem_relinquish() {
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// D[tmem] (+)= A[smem] * B[smem], kind::f16 (bf16/fp16 operands, fp32 accumulate); issued by ONE thread.
__device__ __forceinline__ void umma_f16(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                         uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}\n" ::"r"(tmem_d),
      "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
// Makes the mbarrier track completion of all prior tcgen05.mma of this thread (implies fence::before_thread_sync).
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
               : "memory");
}

// 32 lanes x 8 consecutive 32-bit columns -> 8 registers per thread (thread i of the warp = TMEM lane base+i).
__device__ __forceinline__ void tmem_ld_32x32b_x8(uint32_t taddr, uint32_t (&v)[8]) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7])
               : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld_32x32b_x16(uint32_t taddr, uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
      : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// UMMA shared-memory matrix descriptor, SWIZZLE_NONE ("interleaved" 8x16B core matrices).
//   K-major operand : 8 rows at 16 B pitch form a core matrix; SBO = byte distance between 8-row groups (M/N dir),
//                     LBO = byte distance between the two 8-element K chunks of one K=16 step.
//   MN-major operand: 8 K-rows at 16 B pitch form a core matrix (each row = 8 contiguous M/N elements);
//                     LBO = byte distance between 8-K-row groups, SBO = byte distance between groups of 8 M/N elements.
__device__ __forceinline__ uint64_t umma_desc_noswz(uint32_t smem_addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((smem_addr >> 4) & 0x3FFFu);
  d |= static_cast<uint64_t>((lbo_bytes >> 4) & 0x3FFFu) << 16;
  d |= static_cast<uint64_t>((sbo_bytes >> 4) & 0x3FFFu) << 32;
  d |= static_cast<uint64_t>(1) << 46;  // descriptor version for sm_100
  return d;
}

// Instruction descriptor for kind::f16: bf16 x bf16 -> fp32.  a_mn / b_mn: 1 = MN-major operand, 0 = K-major.
__host__ __device__ constexpr uint32_t umma_idesc_bf16(int M, int N, int a_mn, int b_mn) {
  return (1u << 4)                              // accumulator fp32
         | (1u << 7) | (1u << 10)               // A, B = bf16
         | (static_cast<uint32_t>(a_mn) << 15) | (static_cast<uint32_t>(b_mn) << 16) |
         (static_cast<uint32_t>(N >> 3) << 17) | (static_cast<uint32_t>(M >> 4) << 24);
}

// kind::tf32: fp32 operands in shared memory, read as tf32 (10-bit mantissa) by the tensor core; K = 8 per instruction.
__device__ __forceinline__ void umma_tf32(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                          uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}\n" ::"r"(tmem_d),
      "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
template <bool TF32>
__device__ __forceinline__ void umma(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                     uint32_t accumulate) {
  if (TF32) umma_tf32(tmem_d, desc_a, desc_b, idesc, accumulate);
  else umma_f16(tmem_d, desc_a, desc_b, idesc, accumulate);
}
// Instruction descriptor for kind::tf32: tf32 x tf32 -> fp32 (operand format code 2).
__host__ __device__ constexpr uint32_t umma_idesc_tf32(int M, int N, int a_mn, int b_mn) {
  return (1u << 4) | (2u << 7) | (2u << 10) | (static_cast<uint32_t>(a_mn) << 15) |
         (static_cast<uint32_t>(b_mn) << 16) | (static_cast<uint32_t>(N >> 3) << 17) |
         (static_cast<uint32_t>(M >> 4) << 24);
}

__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  __nv_bfloat162 t = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&t);
}
__device__ __forceinline__ float bf16lo(uint32_t u) { return __uint_as_float(u << 16); }
__device__ __forceinline__ float bf16hi(uint32_t u) { return __uint_as_float(u & 0xffff0000u); }

template <> __device__ __forceinline__ void chunk_unpack<8>(const uint4 u, float (&f)[8]) {
  f[0] = bf16lo(u.x); f[1] = bf16hi(u.x); f[2] = bf16lo(u.y); f[3] = bf16hi(u.y);
  f[4] = bf16lo(u.z); f[5] = bf16hi(u.z); f[6] = bf16lo(u.w); f[7] = bf16hi(u.w);
}
template <> __device__ __forceinline__ void chunk_unpack<4>(const uint4 u, float (&f)[4]) {
  f[0] = __uint_as_float(u.x); f[1] = __uint_as_float(u.y); f[2] = __uint_as_float(u.z); f[3] = __uint_as_float(u.w);
}
template <> __device__ __forceinline__ uint4 chunk_pack<8>(const float (&f)[8]) {
  uint4 p;
  p.x = pack_bf16x2(f[0], f[1]); p.y = pack_bf16x2(f[2], f[3]);
  p.z = pack_bf16x2(f[4], f[5]); p.w = pack_bf16x2(f[6], f[7]);
  return p;
}
template <> __device__ __forceinline__ uint4 chunk_pack<4>(const float (&f)[4]) {
  uint4 p;
  p.x = __float_as_uint(f[0]); p.y = __float_as_uint(f[1]); p.z = __float_as_uint(f[2]); p.w = __float_as_uint(f[3]);
  return p;
}
// fp32 -> nearest tf32 value (10-bit mantissa), kept in an fp32 register.  tcgen05 kind::tf32 reads fp32 operands by
// dropping the low 13 mantissa bits (truncation); cuDNN's TF32 convolutions round to nearest.  Every kernel that
// STORES a tensor which is later a tensor-core operand (activations, gradients w.r.t. conv outputs, packed weights)
// rounds it here, so that the truncation is exact and the error matches the reference's TF32 path.
__device__ __forceinline__ float tf32_rna(float x) {
  uint32_t u;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(u) : "f"(x));
  return __uint_as_float(u);
}
// chunk_pack for a tensor-core operand tensor: bf16 mode rounds by conversion, tf32 mode rounds to the tf32 grid.
template <int V> __device__ __forceinline__ uint4 chunk_pack_operand(const float (&f)[V]);
template <> __device__ __forceinline__ uint4 chunk_pack_operand<8>(const float (&f)[8]) { return chunk_pack<8>(f); }
template <> __device__ __forceinline__ uint4 chunk_pack_operand<4>(const float (&f)[4]) {
  uint4 p;
  p.x = __float_as_uint(tf32_rna(f[0])); p.y = __float_as_uint(tf32_rna(f[1]));
  p.z = __float_as_uint(tf32_rna(f[2])); p.w = __float_as_uint(tf32_rna(f[3]));
  return p;
}
// One element of a planar buffer addressed in 16-byte rows: channel c of row r lives in chunk (c / V) * R + r.
template <int V> __device__ __forceinline__ float elem_load(const uint4* base, long long row, int j);
template <> __device__ __forceinline__ float elem_load<8>(const uint4* base, long long row, int j) {
  return __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(base + row)[j]);
}
template <> __device__ __forceinline__ float elem_load<4>(const uint4* base, long long row, int j) {
  return reinterpret_cast<const float*>(base + row)[j];
}
// "Split" copy of one fp32 planar-4 chunk (tf32 mode): channels [4*chunk, 4*chunk + 4) of row `row` of client k as
// bf16 hi = bf16(x) and lo = bf16(x - hi) halves of a [K][2][C/8][R][8] bf16 tensor (hi set, then lo set) - the operand
// form of the weight-gradient kernel (csrc/wgrad_tc.cu).  `base` is addressed in 8-byte half chunks.
__device__ __forceinline__ void store_split(uint2* base, long long k, long long planes8, long long R, int chunk,
                                            long long row, const float (&f)[4]) {
  const long long hi = (((k * 2) * planes8 + (chunk >> 1)) * R + row) * 2 + (chunk & 1);
  uint2 ph, pl;
  ph.x = pack_bf16x2(f[0], f[1]);
  ph.y = pack_bf16x2(f[2], f[3]);
  pl.x = pack_bf16x2(f[0] - bf16lo(ph.x), f[1] - bf16hi(ph.x));
  pl.y = pack_bf16x2(f[2] - bf16lo(ph.y), f[3] - bf16hi(ph.y));
  base[hi] = ph;
  base[hi + planes8 * R * 2] = pl;
}
template <int V> __device__ __forceinline__ void elem_store(uint4* base, long long row, int j, float v);
template <> __device__ __forceinline__ void elem_store<8>(uint4* base, long long row, int j, float v) {
  reinterpret_cast<__nv_bfloat16*>(base + row)[j] = __float2bfloat16(v);
}
template <> __device__ __forceinline__ void elem_store<4>(uint4* base, long long row, int j, float v) {
  reinterpret_cast<float*>(base + row)[j] = v;
}

}  // namespace dba
