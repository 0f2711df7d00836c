// Microbenchmark (diagnostics): what HBM bandwidth does the access pattern of the container scan reach on its own?
// Every warp reads chunks of `UNROLL` x 512 bytes (128-bit loads, one chunk in flight ahead of the one being consumed),
// either scattered pseudo-randomly over the buffer (like work items) or in sequence.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o membench tools/membench.cu && ./membench
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

template <int UNROLL, int DEPTH>
__global__ void __launch_bounds__(256) scan(const uint4* __restrict__ buf, uint64_t n_chunks, uint32_t iters, int scattered, uint32_t* out) {
   extern __shared__ uint8_t pad[];
   const uint32_t lane = threadIdx.x & 31;
   const uint64_t gw = (uint64_t(blockIdx.x) * blockDim.x + threadIdx.x) >> 5, n_warps = (uint64_t(gridDim.x) * blockDim.x) >> 5;
   auto chunkOf = [&](uint32_t it) -> uint64_t {
      const uint64_t k = scattered ? ((gw * iters + it) * 2654435761ull + 12345ull) : (uint64_t(it) * n_warps + gw);
      return k % n_chunks;
   };
   uint4 v[DEPTH][UNROLL];
   uint32_t acc = 0;
#pragma unroll
   for (int d = 0; d < DEPTH; ++d) {
      const uint4* p = buf + chunkOf(d) * (UNROLL * 32) + lane;
#pragma unroll
      for (int k = 0; k < UNROLL; ++k) v[d][k] = __ldg(p + k * 32);
   }
   for (uint32_t it = 0; it < iters; it += DEPTH) {
#pragma unroll
      for (int d = 0; d < DEPTH; ++d) {
#pragma unroll
         for (int k = 0; k < UNROLL; ++k) acc += __popc(v[d][k].x) + __popc(v[d][k].y) + __popc(v[d][k].z) + __popc(v[d][k].w);
         if (it + d + DEPTH < iters) {
            const uint4* p = buf + chunkOf(it + d + DEPTH) * (UNROLL * 32) + lane;
#pragma unroll
            for (int k = 0; k < UNROLL; ++k) v[d][k] = __ldg(p + k * 32);
         }
      }
   }
   if (acc == 0x12345678u) out[0] = acc + pad[0];
}

template <int UNROLL, int DEPTH>
void run(const uint4* buf, size_t bytes, int ctas_per_sm, int scattered, uint32_t* out, size_t total_bytes) {
   const int sms = 148;
   const int smem = ctas_per_sm >= 8 ? 24 * 1024 : ctas_per_sm == 4 ? 49 * 1024 : 100 * 1024;
   cudaFuncSetAttribute(scan<UNROLL, DEPTH>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
   const uint64_t chunk_bytes = UNROLL * 512, n_chunks = bytes / chunk_bytes;
   const uint32_t grid = sms * ctas_per_sm, n_warps = grid * 8;
   uint32_t iters = uint32_t(total_bytes / chunk_bytes / n_warps);
   iters -= iters % DEPTH;
   cudaEvent_t a, b;
   cudaEventCreate(&a); cudaEventCreate(&b);
   float best = 1e9;
   for (int rep = 0; rep < 5; ++rep) {
      cudaEventRecord(a);
      scan<UNROLL, DEPTH><<<grid, 256, smem>>>(buf, n_chunks, iters, scattered, out);
      cudaEventRecord(b);
      cudaEventSynchronize(b);
      float ms; cudaEventElapsedTime(&ms, a, b);
      if (rep && ms < best) best = ms;
   }
   const double moved = double(iters) * n_warps * chunk_bytes;
   printf("chunk %5llu B depth %d ctas/sm %d %-9s moved %6.1f MB in %7.1f us  -> %6.0f GB/s\n", (unsigned long long)chunk_bytes, DEPTH, ctas_per_sm,
          scattered ? "scattered" : "sequence", moved / 1e6, best * 1e3, moved / best / 1e6);
}

int main() {
   const size_t bytes = 1ull << 30;
   uint4* buf; uint32_t* out;
   cudaMalloc(&buf, bytes); cudaMalloc(&out, 64);
   cudaMemset(buf, 1, bytes);
   for (size_t total : {size_t(130) << 20, size_t(1) << 30}) {
      printf("--- total %zu MB per launch\n", total >> 20);
      for (int scattered = 1; scattered >= 0; --scattered) {
         for (int cps : {4, 8}) {
            run<4, 1>(buf, bytes, cps, scattered, out, total);
            run<4, 2>(buf, bytes, cps, scattered, out, total);
            run<8, 1>(buf, bytes, cps, scattered, out, total);
            run<2, 2>(buf, bytes, cps, scattered, out, total);
         }
         run<4, 1>(buf, bytes, 2, scattered, out, total);
         run<8, 2>(buf, bytes, 2, scattered, out, total);
      }
   }
   if (cudaDeviceSynchronize() != cudaSuccess) { printf("CUDA error\n"); return 1; }
   return 0;
}
