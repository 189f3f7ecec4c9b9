// exp_store.cu -- which store flavour reaches the HBM write roofline for k_raster's output
// pattern?  N CTAs of 256 threads, each writing two 64 KB images (ogm[ai], dagm[ai]).
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -lineinfo -o profiles/_bin/exp_store profiles/exp_store.cu
//   profiles/_bin/exp_store [N=16384] [reps=20]
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cuda_runtime.h>

#define IMG 65536
#define THREADS 256

__device__ __forceinline__ uint32_t val(int i) { return 0x00ff00ffu * (i & 1); }

__global__ void __launch_bounds__(THREADS, 4) k_st16(uint8_t* a, uint8_t* b) {
  uint4* pa = reinterpret_cast<uint4*>(a + (size_t)blockIdx.x * IMG);
  uint4* pb = reinterpret_cast<uint4*>(b + (size_t)blockIdx.x * IMG);
  for (int i = threadIdx.x; i < IMG / 16; i += THREADS) pa[i] = make_uint4(val(i), val(i), val(i), val(i));
  for (int i = threadIdx.x; i < IMG / 16; i += THREADS) pb[i] = make_uint4(val(i), val(i), val(i), val(i));
}
__global__ void __launch_bounds__(THREADS, 4) k_st16cs(uint8_t* a, uint8_t* b) {
  uint4* pa = reinterpret_cast<uint4*>(a + (size_t)blockIdx.x * IMG);
  uint4* pb = reinterpret_cast<uint4*>(b + (size_t)blockIdx.x * IMG);
  for (int i = threadIdx.x; i < IMG / 16; i += THREADS) __stcs(pa + i, make_uint4(val(i), val(i), val(i), val(i)));
  for (int i = threadIdx.x; i < IMG / 16; i += THREADS) __stcs(pb + i, make_uint4(val(i), val(i), val(i), val(i)));
}
__device__ __forceinline__ void st32(uint8_t* p, uint32_t v) {
  asm volatile("st.global.v8.b32 [%0], {%1,%1,%1,%1,%1,%1,%1,%1};" ::"l"(p), "r"(v) : "memory");
}
__global__ void __launch_bounds__(THREADS, 4) k_st32(uint8_t* a, uint8_t* b) {
  uint8_t* pa = a + (size_t)blockIdx.x * IMG;
  uint8_t* pb = b + (size_t)blockIdx.x * IMG;
  for (int i = threadIdx.x; i < IMG / 32; i += THREADS) st32(pa + (size_t)i * 32, val(i));
  for (int i = threadIdx.x; i < IMG / 32; i += THREADS) st32(pb + (size_t)i * 32, val(i));
}
__device__ __forceinline__ void st32na(uint8_t* p, uint32_t v) {
  asm volatile("st.global.L1::no_allocate.v8.b32 [%0], {%1,%1,%1,%1,%1,%1,%1,%1};" ::"l"(p), "r"(v) : "memory");
}
__global__ void __launch_bounds__(THREADS, 4) k_st32na(uint8_t* a, uint8_t* b) {
  uint8_t* pa = a + (size_t)blockIdx.x * IMG;
  uint8_t* pb = b + (size_t)blockIdx.x * IMG;
  for (int i = threadIdx.x; i < IMG / 32; i += THREADS) st32na(pa + (size_t)i * 32, val(i));
  for (int i = threadIdx.x; i < IMG / 32; i += THREADS) st32na(pb + (size_t)i * 32, val(i));
}
// 4-byte stores, fully coalesced (128 B per warp instruction)
__global__ void __launch_bounds__(THREADS, 4) k_st4(uint8_t* a, uint8_t* b) {
  uint32_t* pa = reinterpret_cast<uint32_t*>(a + (size_t)blockIdx.x * IMG);
  uint32_t* pb = reinterpret_cast<uint32_t*>(b + (size_t)blockIdx.x * IMG);
  for (int i = threadIdx.x; i < IMG / 4; i += THREADS) pa[i] = val(i);
  for (int i = threadIdx.x; i < IMG / 4; i += THREADS) pb[i] = val(i);
}

// bulk (TMA) stores: the CTA fills a CHUNK-byte tile in shared memory, one thread issues
// cp.async.bulk shared -> global; two tiles in flight (double buffer)
template <int CHUNK>
__global__ void __launch_bounds__(THREADS, 4) k_bulk(uint8_t* a, uint8_t* b) {
  extern __shared__ __align__(128) uint8_t sm[];
  uint8_t* dst[2] = {a + (size_t)blockIdx.x * IMG, b + (size_t)blockIdx.x * IMG};
  int buf = 0;
  for (int img = 0; img < 2; ++img) {
    for (int off = 0; off < IMG; off += CHUNK, buf ^= 1) {
      uint8_t* tile = sm + buf * CHUNK;
      // the tile last used two iterations ago must have been read by the bulk engine
      if (threadIdx.x == 0) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
      __syncthreads();
      uint4* t4 = reinterpret_cast<uint4*>(tile);
      for (int i = threadIdx.x; i < CHUNK / 16; i += THREADS) t4[i] = make_uint4(val(i), val(i), val(i), val(i));
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      __syncthreads();
      if (threadIdx.x == 0) {
        uint32_t s = (uint32_t)__cvta_generic_to_shared(tile);
        asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst[img] + off), "r"(s),
                     "r"(CHUNK)
                     : "memory");
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      }
    }
  }
  if (threadIdx.x == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

template <class F>
void run(const char* name, F launch, int reps, size_t bytes) {
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  for (int i = 0; i < 3; ++i) launch();
  cudaDeviceSynchronize();
  float best = 1e30f, sum = 0.f;
  for (int r = 0; r < reps; ++r) {
    cudaEventRecord(e0);
    launch();
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    best = ms < best ? ms : best;
    sum += ms;
  }
  cudaError_t err = cudaGetLastError();
  printf("%-14s best %.4f ms  mean %.4f ms  %.1f GB/s (best)  %.1f GB/s (mean)  %s\n", name, best, sum / reps,
         bytes / (best * 1e6), bytes / (sum / reps * 1e6), err == cudaSuccess ? "" : cudaGetErrorString(err));
}

int main(int argc, char** argv) {
  int N = argc > 1 ? atoi(argv[1]) : 16384, reps = argc > 2 ? atoi(argv[2]) : 20;
  uint8_t *a, *b;
  size_t bytes = (size_t)N * IMG;
  cudaMalloc(&a, bytes);
  cudaMalloc(&b, bytes);
  cudaMemset(a, 0, bytes);
  cudaMemset(b, 0, bytes);
  run("st4", [&] { k_st4<<<N, THREADS>>>(a, b); }, reps, 2 * bytes);
  run("st16", [&] { k_st16<<<N, THREADS>>>(a, b); }, reps, 2 * bytes);
  run("st16cs", [&] { k_st16cs<<<N, THREADS>>>(a, b); }, reps, 2 * bytes);
  run("st32", [&] { k_st32<<<N, THREADS>>>(a, b); }, reps, 2 * bytes);
  run("st32na", [&] { k_st32na<<<N, THREADS>>>(a, b); }, reps, 2 * bytes);
  cudaFuncSetAttribute(k_bulk<8192>, cudaFuncAttributeMaxDynamicSharedMemorySize, 2 * 8192);
  cudaFuncSetAttribute(k_bulk<16384>, cudaFuncAttributeMaxDynamicSharedMemorySize, 2 * 16384);
  run("bulk8k", [&] { k_bulk<8192><<<N, THREADS, 2 * 8192>>>(a, b); }, reps, 2 * bytes);
  run("bulk16k", [&] { k_bulk<16384><<<N, THREADS, 2 * 16384>>>(a, b); }, reps, 2 * bytes);
  run("memset", [&] { cudaMemsetAsync(a, 1, bytes); cudaMemsetAsync(b, 1, bytes); }, reps, 2 * bytes);
  return 0;
}
