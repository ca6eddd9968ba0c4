// Micro-benchmark behind DESIGN.md "adjoint roofline": throughput of the 8-corner trilinear
// gradient scatter into a 256^3 fp32 volume (Z fastest) for different RED shapes.
//   mode 0: 8 x red.global.add.f32                    (one per corner)
//   mode 1: 4 x red.global.v2.f32.add on z pairs      (only legal when z0 is even; benchmark of the op)
//   mode 2: 4 x red.global.v4.f32.add, aligned quad containing z0,z0+1, zero padded
//   mode 3: mixed: v2 when z0 even else 2 scalar (what a real kernel can do with v2 only)
//   mode 4: mixed: v4 zero padded when (z0&3)!=3 else 2 scalar
//   mode 5: mixed: v2 when z0 even, v4 {0,a,b,0} when z0%4==1, scalar when z0%4==3
// Each lane walks a ray-like sequence of cells: warp lanes sit in an 8x4 patch (0.6 voxel pitch)
// and advance along a direction; every iteration flushes one cell (8 corners).
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o red_throughput red_throughput.cu
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

__device__ __forceinline__ void red1(float* p, float a) {
  asm volatile("red.global.add.f32 [%0], %1;" ::"l"(p), "f"(a) : "memory");
}
__device__ __forceinline__ void red2(float* p, float a, float b) {
  asm volatile("red.global.v2.f32.add [%0], {%1,%2};" ::"l"(p), "f"(a), "f"(b) : "memory");
}
__device__ __forceinline__ void red4(float* p, float a, float b, float c, float d) {
  asm volatile("red.global.v4.f32.add [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

template <int MODE>
__global__ void __launch_bounds__(128) k(float* vol, int n, int iters, float pitch, int dirsel) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int tx = blockIdx.x * 16 + (warp & 1) * 8 + (lane & 7);
  const int ty = blockIdx.y * 8 + (warp >> 1) * 4 + (lane >> 3);
  // ray through the volume, mostly along z for even blockIdx.z, along x for odd
  float px = tx * pitch, py = ty * pitch, pz = 0.3f;
  float dx = 0.11f, dy = 0.07f, dz = 0.9f;
  const int dsel = dirsel < 0 ? (blockIdx.z & 1) : dirsel;
  if (dsel == 1) { dx = 0.9f; dz = 0.11f; px = 0.3f; pz = tx * pitch; }
  if (dsel == 2) { dy = 0.9f; dz = 0.07f; py = 0.3f; pz = ty * pitch; dx = 0.11f; }
  const float a = 1.0f / 1024.f;
  for (int i = 0; i < iters; ++i) {
    int x0 = int(px), y0 = int(py), z0 = int(pz);
    if (x0 >= n - 1 || y0 >= n - 1 || z0 >= n - 1) { px = (tx * pitch) * 0.5f; py = ty * pitch * 0.5f; pz = 0.3f; if (dsel == 1) { px = 0.3f; pz = tx * pitch * 0.5f; } if (dsel == 2) { py = 0.3f; pz = ty * pitch * 0.5f; } continue; }
    float* c = vol + ((size_t)x0 * n + y0) * n + z0;
    const size_t sx = (size_t)n * n, sy = n;
    if (MODE == 6) {  // 4x4x2 bricks (128-byte lines), separable address f(x)+g(y)+h(z)
      const unsigned nb = n >> 2, nbz = n >> 1;
      auto fx = [&](int x) { return (unsigned(x) >> 2) * (nb * nbz * 32u) + (unsigned(x) & 3u) * 8u; };
      auto fy = [&](int y) { return (unsigned(y) >> 2) * (nbz * 32u) + (unsigned(y) & 3u) * 2u; };
      auto fz = [&](int z) { return (unsigned(z) >> 1) * 32u + (unsigned(z) & 1u); };
      const unsigned o = fx(x0) + fy(y0) + fz(z0);
      const unsigned ex = fx(x0 + 1) - fx(x0), ey = fy(y0 + 1) - fy(y0), ez = fz(z0 + 1) - fz(z0);
      float* b = vol + o;
      red1(b, a); red1(b + ez, a); red1(b + ey, a); red1(b + ey + ez, a);
      red1(b + ex, a); red1(b + ex + ez, a); red1(b + ex + ey, a); red1(b + ex + ey + ez, a);
    } else if (MODE == 0) {
      red1(c, a); red1(c + 1, a); red1(c + sy, a); red1(c + sy + 1, a);
      red1(c + sx, a); red1(c + sx + 1, a); red1(c + sx + sy, a); red1(c + sx + sy + 1, a);
    } else if (MODE == 1) {
      float* e = vol + ((size_t)x0 * n + y0) * n + (z0 & ~1);
      red2(e, a, a); red2(e + sy, a, a); red2(e + sx, a, a); red2(e + sx + sy, a, a);
    } else if (MODE == 2) {
      float* e = vol + ((size_t)x0 * n + y0) * n + (z0 & ~3);
      red4(e, a, a, 0.f, 0.f); red4(e + sy, a, a, 0.f, 0.f); red4(e + sx, a, a, 0.f, 0.f); red4(e + sx + sy, a, a, 0.f, 0.f);
    } else if (MODE == 3) {
      if ((z0 & 1) == 0) {
        red2(c, a, a); red2(c + sy, a, a); red2(c + sx, a, a); red2(c + sx + sy, a, a);
      } else {
        red1(c, a); red1(c + 1, a); red1(c + sy, a); red1(c + sy + 1, a);
        red1(c + sx, a); red1(c + sx + 1, a); red1(c + sx + sy, a); red1(c + sx + sy + 1, a);
      }
    } else if (MODE == 5) {
      const int q = z0 & 3;
      if ((q & 1) == 0) {
        red2(c, a, a); red2(c + sy, a, a); red2(c + sx, a, a); red2(c + sx + sy, a, a);
      } else if (q == 1) {
        float* e = c - 1;
        red4(e, 0.f, a, a, 0.f); red4(e + sy, 0.f, a, a, 0.f); red4(e + sx, 0.f, a, a, 0.f); red4(e + sx + sy, 0.f, a, a, 0.f);
      } else {
        red1(c, a); red1(c + 1, a); red1(c + sy, a); red1(c + sy + 1, a);
        red1(c + sx, a); red1(c + sx + 1, a); red1(c + sx + sy, a); red1(c + sx + sy + 1, a);
      }
    } else {
      const int q = z0 & 3;
      if (q != 3) {
        float* e = c - q;
        const float v0 = q == 0 ? a : 0.f, v1 = q <= 1 ? a : 0.f, v2 = (q == 1 || q == 2) ? a : 0.f, v3 = q == 2 ? a : 0.f;
        red4(e, v0, v1, v2, v3); red4(e + sy, v0, v1, v2, v3); red4(e + sx, v0, v1, v2, v3); red4(e + sx + sy, v0, v1, v2, v3);
      } else {
        red1(c, a); red1(c + 1, a); red1(c + sy, a); red1(c + sy + 1, a);
        red1(c + sx, a); red1(c + sx + 1, a); red1(c + sx + sy, a); red1(c + sx + sy + 1, a);
      }
    }
    px += dx; py += dy; pz += dz;
  }
}

template <int MODE>
float run(float* vol, int n, int iters, int dirsel = -1) {
  dim3 grid(32, 64, 8);  // 512x512 "pixels", 8 "views"
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  k<MODE><<<grid, 128>>>(vol, n, iters, 0.45f, dirsel);
  cudaDeviceSynchronize();
  cudaEventRecord(e0);
  k<MODE><<<grid, 128>>>(vol, n, iters, 0.45f, dirsel);
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms;
  cudaEventElapsedTime(&ms, e0, e1);
  const double cells = 32.0 * 64 * 8 * 128 * iters;
  printf("mode %d dir %d: %.3f ms, %.1f Gcells/s, %.1f Gcorner-adds/s\n", MODE, dirsel, ms, cells / ms * 1e-6, 8 * cells / ms * 1e-6);
  return ms;
}

int main() {
  const int n = 256, iters = 250;
  float* vol;
  cudaMalloc(&vol, sizeof(float) * n * n * n);
  cudaMemset(vol, 0, sizeof(float) * n * n * n);
  run<0>(vol, n, iters); run<1>(vol, n, iters); run<2>(vol, n, iters); run<3>(vol, n, iters); run<4>(vol, n, iters); run<5>(vol, n, iters);
  for (int d = 0; d < 3; ++d) { run<0>(vol, n, iters, d); run<6>(vol, n, iters, d); }
  run<6>(vol, n, iters);
  cudaError_t e = cudaDeviceSynchronize();
  printf("status: %s\n", cudaGetErrorString(e));
  return e != cudaSuccess;
}
