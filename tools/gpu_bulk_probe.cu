// Micro-benchmark: what does one cp.async.bulk (UBLKCP) request cost as a function of its size?
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o build/gpu_bulk_probe tools/gpu_bulk_probe.cu
//   build/gpu_bulk_probe            (on a B200)
//
// Every CTA streams `total` bytes from an L2-resident buffer into a ring of shared-memory stages.  A stage of
// `stage_bytes` is fetched as stage_bytes / piece copies of `piece` bytes, issued by the lanes of one warp in parallel
// (the way conv_tc's producer warp does); the consumer side only waits for the stage and hands it back.  Reported:
// cycles per request and bytes per clock per SM for grids of 1, 52 and 148 CTAs.  The question behind it: the
// streamed-weight layers of conv_tc run at ~11 B/clk per SM when the weights of an n-block arrive as 1 KB rows.
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e_), __LINE__); exit(1); } } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* b, uint32_t n) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(n));
}
__device__ __forceinline__ void mbar_expect(uint64_t* b, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(b)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* b) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(b)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* b, uint32_t parity) {
  asm volatile(
      "{\n.reg .pred p;\nW: mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra D;\nbra W;\nD:\n}" ::"r"(smem_u32(b)),
      "r"(parity)
      : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}

constexpr int kMaxStages = 16;

// mode 0: one producer warp, ring of `stages` (full / empty mbarriers per stage)           [conv_tc's scheme]
// mode 1: one barrier, expect_tx(total) once, every copy issued back to back, one wait      [what the copy engine can do]
// mode 2: like 0, but stage s is produced by warp s % nprod (nprod producer warps)
// mode 3: like 0, but the free-stage wait / expect_tx of stage it+1 is issued BEFORE the copies of stage it
__global__ void __launch_bounds__(192, 1) probe(const char* src, long long src_bytes, int piece, int stage_bytes, int stages,
                                                long long total, long long* cycles, int mode, int nprod) {
  extern __shared__ __align__(128) char smem[];
  __shared__ uint64_t full[kMaxStages], empty[kMaxStages];
  if (threadIdx.x == 0) {
    for (int i = 0; i < stages; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n_stage = (int)(total / stage_bytes), per = stage_bytes / piece;
  const long long t0 = clock64();
  const long long span = (long long)stage_bytes * 3;
  const long long wrap = (src_bytes - 2 * span) / span;       // stages before the source offset wraps
  if (mode == 1) {
    if (warp == 0) {
      if (lane == 0) mbar_expect(&full[0], (uint32_t)total);
      __syncwarp();
      for (int it = 0; it < n_stage; ++it) {
        const long long off = (long long)(it % wrap) * span;
        const int s = it % stages;
        for (int i = lane; i < per; i += 32)
          bulk_g2s(smem + (size_t)s * stage_bytes + (size_t)i * piece, src + off + (long long)i * piece * 3, (uint32_t)piece,
                   &full[0]);
      }
      if (lane == 0) mbar_wait(&full[0], 0);
    }
  } else if (warp < nprod) {
    const int first = mode == 2 ? warp : 0, step = mode == 2 ? nprod : 1;
    if (mode == 3 && lane == 0) { mbar_wait(&empty[0], 1u); mbar_expect(&full[0], (uint32_t)stage_bytes); }
    for (int it = first; it < n_stage; it += step) {
      const int s = it % stages;
      const uint32_t ph = (uint32_t)(it / stages) & 1u;
      if (mode != 3) {
        if (lane == 0) {
          mbar_wait(&empty[s], ph ^ 1u);
          mbar_expect(&full[s], (uint32_t)stage_bytes);
        }
        __syncwarp();
      }
      const long long off = (long long)(it % wrap) * span;
      for (int i = lane; i < per; i += 32)
        bulk_g2s(smem + (size_t)s * stage_bytes + (size_t)i * piece, src + off + (long long)i * piece * 3, (uint32_t)piece,
                 &full[s]);
      if (mode == 3) {
        const int it2 = it + 1;
        if (it2 < n_stage && lane == 0) {
          const int s2 = it2 % stages;
          mbar_wait(&empty[s2], ((uint32_t)(it2 / stages) & 1u) ^ 1u);
          mbar_expect(&full[s2], (uint32_t)stage_bytes);
        }
        __syncwarp();
      }
    }
  } else if (warp == 5 && lane == 0) {
    for (int it = 0; it < n_stage; ++it) {
      const int s = it % stages;
      const uint32_t ph = (uint32_t)(it / stages) & 1u;
      mbar_wait(&full[s], ph);
      mbar_arrive(&empty[s]);
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) cycles[blockIdx.x] = clock64() - t0;
}

int main() {
  const long long src_bytes = 8LL << 20;        // L2-resident
  char* src;
  long long* cyc;
  CK(cudaMalloc(&src, src_bytes));
  CK(cudaMemset(src, 1, src_bytes));
  CK(cudaMalloc(&cyc, 148 * sizeof(long long)));
  CK(cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  printf("%5s %6s %5s %7s %7s %6s | %9s %10s %9s %8s\n", "mode", "nprod", "grid", "piece", "stageB", "stages", "cyc/req",
         "cyc/stage", "B/clk/SM", "us");
  struct Cfg { int mode, nprod; };
  const Cfg cfgs[] = {{0, 1}, {1, 1}, {2, 2}, {2, 4}, {3, 1}};
  const int grids[] = {1, 148};
  const int pieces[] = {1024, 8192};
  for (const Cfg& c : cfgs) {
    for (int g : grids) {
      for (int stage_bytes : {8192, 32768}) {
        for (int piece : pieces) {
          if (stage_bytes < piece) continue;
          for (int stages : {4, 6}) {
            const long long total = c.mode == 1 ? (512LL << 10) : (2LL << 20);   // tx-count limit of one barrier: 2^20 - 1
            long long h[148];
            cudaEvent_t e0, e1;
            CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
            for (int rep = 0; rep < 2; ++rep) {
              CK(cudaEventRecord(e0));
              probe<<<g, 192, stages * stage_bytes>>>(src, src_bytes, piece, stage_bytes, stages, total, cyc, c.mode, c.nprod);
              CK(cudaEventRecord(e1));
              CK(cudaDeviceSynchronize());
            }
            float ms = 0;
            CK(cudaEventElapsedTime(&ms, e0, e1));
            CK(cudaMemcpy(h, cyc, g * sizeof(long long), cudaMemcpyDeviceToHost));
            long long mx = 0;
            for (int i = 0; i < g; ++i) mx = h[i] > mx ? h[i] : mx;
            const double reqs = (double)total / piece, nst = (double)total / stage_bytes;
            printf("%5d %6d %5d %7d %7d %6d | %9.1f %10.1f %9.1f %8.1f\n", c.mode, c.nprod, g, piece, stage_bytes, stages,
                   mx / reqs, mx / nst, (double)total / mx, ms * 1000.0);
            CK(cudaEventDestroy(e0)); CK(cudaEventDestroy(e1));
          }
        }
      }
    }
  }
  return 0;
}
