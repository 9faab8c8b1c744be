// Tuning aid: what does a pure 32 B/point copy reach with the access patterns annotate_kernel could use?
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/copy_probe tools/copy_probe.cu && /tmp/copy_probe
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
struct P8 { float x, y, z, w; uint32_t a, b, c, d; };
__device__ __forceinline__ P8 ld(const P8* p) { P8 r; asm volatile("ld.global.nc.L1::no_allocate.L2::evict_first.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];" : "=f"(r.x),"=f"(r.y),"=f"(r.z),"=f"(r.w),"=r"(r.a),"=r"(r.b),"=r"(r.c),"=r"(r.d) : "l"(p)); return r; }
__device__ __forceinline__ void st(P8* p, const P8& v) { asm volatile("st.global.L1::no_allocate.v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" :: "l"(p),"f"(v.x),"f"(v.y),"f"(v.z),"f"(v.w),"r"(v.a),"r"(v.b),"r"(v.c),"r"(v.d) : "memory"); }
// (a) one uint4 per thread, one element per thread, huge grid (torch-like)
__global__ void copy16(const uint4* in, uint4* out, size_t n) { size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; if (i < n) out[i] = in[i]; }
// (b) one 32-byte point per thread, huge grid
__global__ void copy32(const P8* in, P8* out, size_t n) { size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; if (i < n) st(out + i, ld(in + i)); }
// (c) persistent CTAs, 2048-point tiles strided by the grid, 2 points in flight + next-round prefetch (annotate_kernel's shape)
template <bool MASK>
__global__ void __launch_bounds__(256, 4) copy_tiles(const P8* in, P8* out, uint8_t* mask, unsigned n_tiles) {
  for (unsigned tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    size_t base = (size_t)tile * 2048 + threadIdx.x;
    P8 c0 = ld(in + base), c1 = ld(in + base + 256), n0, n1;
#pragma unroll 1
    for (int r = 0; r < 4; r++) {
      if (r < 3) { n0 = ld(in + base + 512); n1 = ld(in + base + 768); }
      c0.w = 1.f; c1.w = 1.f; c0.a = c0.a & 0x80 ? 0xffffff : 0; c1.a = c1.a & 0x80 ? 0xffffff : 0;
      st(out + base, c0); st(out + base + 256, c1);
      if (MASK) { mask[base] = c0.a ? 255 : 0; mask[base + 256] = c1.a ? 255 : 0; }
      c0 = n0; c1 = n1; base += 512;
    }
  }
}
// (d) persistent CTAs, each owning one contiguous chunk of the array
__global__ void __launch_bounds__(256, 4) copy_chunks(const P8* in, P8* out, size_t n) {
  size_t per = (n / gridDim.x / 512) * 512; size_t b0 = (size_t)blockIdx.x * per, b1 = blockIdx.x == gridDim.x - 1 ? n : b0 + per;
  for (size_t i = b0 + threadIdx.x; i + 256 < b1; i += 512) { P8 a = ld(in + i), b = ld(in + i + 256); st(out + i, a); st(out + i + 256, b); }
}
int main() {
  const size_t n = 276480000; P8 *in, *out; uint8_t* mask;
  cudaMalloc(&in, n * 32); cudaMalloc(&out, n * 32); cudaMalloc(&mask, n); cudaMemset(in, 1, n * 32);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  auto run = [&](const char* name, auto launch, double bytes) {
    for (int i = 0; i < 3; i++) launch();
    cudaEventRecord(e0); for (int i = 0; i < 10; i++) launch(); cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1); ms /= 10; printf("%-34s %.3f ms  %.0f GB/s  (%s)\n", name, ms, bytes / ms / 1e6, cudaGetErrorString(cudaGetLastError()));
  };
  run("memcpy D2D", [&] { cudaMemcpyAsync(out, in, n * 32, cudaMemcpyDeviceToDevice); }, 64.0 * n);
  run("copy16 grid=n", [&] { copy16<<<(unsigned)((2 * n + 255) / 256), 256>>>((uint4*)in, (uint4*)out, 2 * n); }, 64.0 * n);
  run("copy32 grid=n", [&] { copy32<<<(unsigned)((n + 255) / 256), 256>>>(in, out, n); }, 64.0 * n);
  run("copy32 grid=n block 1024", [&] { copy32<<<(unsigned)((n + 1023) / 1024), 1024>>>(in, out, n); }, 64.0 * n);
  run("tiles persistent 592", [&] { copy_tiles<false><<<592, 256>>>(in, out, mask, (unsigned)(n / 2048)); }, 64.0 * n);
  run("tiles persistent 592 +mask", [&] { copy_tiles<true><<<592, 256>>>(in, out, mask, (unsigned)(n / 2048)); }, 65.0 * n);
  run("tiles grid=n_tiles", [&] { copy_tiles<false><<<(unsigned)(n / 2048), 256>>>(in, out, mask, (unsigned)(n / 2048)); }, 64.0 * n);
  run("tiles grid=n_tiles +mask", [&] { copy_tiles<true><<<(unsigned)(n / 2048), 256>>>(in, out, mask, (unsigned)(n / 2048)); }, 65.0 * n);
  run("chunks persistent 592", [&] { copy_chunks<<<592, 256>>>(in, out, n); }, 64.0 * n);
  run("chunks persistent 1184", [&] { copy_chunks<<<1184, 256>>>(in, out, n); }, 64.0 * n);
  return 0;
}
