// Micro-benchmark behind DESIGN.md "what bounds the forward march": throughput of the 8-corner
// trilinear GATHER from a 256^3 fp32 volume for different load shapes / layouts, in the access pattern
// of the marcher (a warp = 8x4 pixel patch of rays advancing through the volume, one cell per step).
//   mode 0: 8 x LDG.32, linear layout (Z fastest)                      -- what the kernels do today
//   mode 1: 4 x LDG.64 on a PAIR layout  P[x][y][z] = (v[z], v[z+1])   -- 2x memory, always aligned
//   mode 2: 2 x LDG.128 on a QUAD layout Q[x][y][z] = (v[y][z], v[y][z+1], v[y+1][z], v[y+1][z+1]) -- 4x memory
//   mode 3: 8 x tex3D<float> point fetches from a 3D CUDA array
//   mode 4: 2 x tex2Dgather on a 2D CUDA array (width Z, height X*Y): the (y, z) 2x2 footprint per fetch
//   mode 5: linear layout, LDG.64 on the z pair when z0 is even, else 2 x LDG.32
//   mode 6: linear layout, LDG.128 on the aligned quad holding z0, z0+1 when (z0 & 3) != 3, else 2 x LDG.32
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o gather_throughput gather_throughput.cu
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>

struct Args {
  const float* lin;
  const float2* pair;
  const float4* quad;
  cudaTextureObject_t tex3, tex2;
  int n, iters;
  float pitch;
  int dirsel;
  float* out;
};

template <int MODE>
__global__ void __launch_bounds__(128) k(const Args a) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int tx = blockIdx.x * 16 + (warp & 1) * 8 + (lane & 7);
  const int ty = blockIdx.y * 8 + (warp >> 1) * 4 + (lane >> 3);
  const int n = a.n;
  float px = tx * a.pitch, py = ty * a.pitch, pz = 0.3f;
  float dx = 0.11f, dy = 0.07f, dz = 0.9f;
  if (a.dirsel == 1) { dx = 0.9f; dz = 0.11f; px = 0.3f; pz = tx * a.pitch; }
  if (a.dirsel == 2) { dy = 0.9f; dz = 0.07f; py = 0.3f; pz = ty * a.pitch; dx = 0.11f; }
  float acc = 0.f;
  const size_t sx = (size_t)n * n, sy = n;
  for (int i = 0; i < a.iters; ++i) {
    int x0 = int(px), y0 = int(py), z0 = int(pz);
    if (x0 >= n - 1 || y0 >= n - 1 || z0 >= n - 1) {
      px = tx * a.pitch * 0.5f; py = ty * a.pitch * 0.5f; pz = 0.3f;
      if (a.dirsel == 1) { px = 0.3f; pz = tx * a.pitch * 0.5f; }
      if (a.dirsel == 2) { py = 0.3f; pz = ty * a.pitch * 0.5f; }
      continue;
    }
    const size_t o = ((size_t)x0 * n + y0) * n + z0;
    if (MODE == 0) {
      const float* c = a.lin + o;
      acc += __ldg(c) + __ldg(c + 1) + __ldg(c + sy) + __ldg(c + sy + 1) + __ldg(c + sx) + __ldg(c + sx + 1) +
             __ldg(c + sx + sy) + __ldg(c + sx + sy + 1);
    } else if (MODE == 1) {
      const float2* c = a.pair + o;
      const float2 p0 = __ldg(c), p1 = __ldg(c + sy), p2 = __ldg(c + sx), p3 = __ldg(c + sx + sy);
      acc += p0.x + p0.y + p1.x + p1.y + p2.x + p2.y + p3.x + p3.y;
    } else if (MODE == 2) {
      const float4* c = a.quad + o;
      const float4 q0 = __ldg(c), q1 = __ldg(c + sx);
      acc += q0.x + q0.y + q0.z + q0.w + q1.x + q1.y + q1.z + q1.w;
    } else if (MODE == 3) {
      const float fx = x0 + 0.5f, fy = y0 + 0.5f, fz = z0 + 0.5f;  // texture coords (z fastest = tex x)
      acc += tex3D<float>(a.tex3, fz, fy, fx) + tex3D<float>(a.tex3, fz + 1, fy, fx) +
             tex3D<float>(a.tex3, fz, fy + 1, fx) + tex3D<float>(a.tex3, fz + 1, fy + 1, fx) +
             tex3D<float>(a.tex3, fz, fy, fx + 1) + tex3D<float>(a.tex3, fz + 1, fy, fx + 1) +
             tex3D<float>(a.tex3, fz, fy + 1, fx + 1) + tex3D<float>(a.tex3, fz + 1, fy + 1, fx + 1);
    } else if (MODE == 4) {
      const float u = z0 + 1.0f, v0 = float(x0 * n + y0) + 1.0f, v1 = float((x0 + 1) * n + y0) + 1.0f;
      const float4 g0 = tex2Dgather<float4>(a.tex2, u, v0, 0), g1 = tex2Dgather<float4>(a.tex2, u, v1, 0);
      acc += g0.x + g0.y + g0.z + g0.w + g1.x + g1.y + g1.z + g1.w;
    } else if (MODE == 5) {
      const float* c = a.lin + o;
      if ((z0 & 1) == 0) {
        const float2 p0 = __ldg((const float2*)c), p1 = __ldg((const float2*)(c + sy)), p2 = __ldg((const float2*)(c + sx)),
                     p3 = __ldg((const float2*)(c + sx + sy));
        acc += p0.x + p0.y + p1.x + p1.y + p2.x + p2.y + p3.x + p3.y;
      } else {
        acc += __ldg(c) + __ldg(c + 1) + __ldg(c + sy) + __ldg(c + sy + 1) + __ldg(c + sx) + __ldg(c + sx + 1) +
               __ldg(c + sx + sy) + __ldg(c + sx + sy + 1);
      }
    } else {
      const float* c = a.lin + o;
      const int q = z0 & 3;
      if (q != 3) {
        const float* e = c - q;
        const float4 p0 = __ldg((const float4*)e), p1 = __ldg((const float4*)(e + sy)), p2 = __ldg((const float4*)(e + sx)),
                     p3 = __ldg((const float4*)(e + sx + sy));
        auto pick = [&](const float4& p) { return q == 0 ? p.x + p.y : q == 1 ? p.y + p.z : p.z + p.w; };
        acc += pick(p0) + pick(p1) + pick(p2) + pick(p3);
      } else {
        acc += __ldg(c) + __ldg(c + 1) + __ldg(c + sy) + __ldg(c + sy + 1) + __ldg(c + sx) + __ldg(c + sx + 1) +
               __ldg(c + sx + sy) + __ldg(c + sx + sy + 1);
      }
    }
    px += dx; py += dy; pz += dz;
  }
  if (acc == 12345.678f) a.out[0] = acc;  // keep the loads alive
}

template <int MODE>
float run(Args a, int dirsel, float pitch) {
  a.dirsel = dirsel;
  a.pitch = pitch;
  dim3 grid(32, 64, 8);  // 512x512 "pixels", 8 "views"
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  k<MODE><<<grid, 128>>>(a);
  cudaDeviceSynchronize();
  cudaEventRecord(e0);
  k<MODE><<<grid, 128>>>(a);
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms;
  cudaEventElapsedTime(&ms, e0, e1);
  const double cells = 32.0 * 64 * 8 * 128 * a.iters;
  printf("mode %d dir %d pitch %.2f: %.3f ms, %.1f Gcells/s, %.1f Gcorner-loads/s  [%s]\n", MODE, dirsel, pitch, ms,
         cells / ms * 1e-6, 8 * cells / ms * 1e-6, cudaGetErrorString(cudaGetLastError()));
  return ms;
}

int main() {
  const int n = 256;
  const size_t N = (size_t)n * n * n;
  std::vector<float> h(N);
  for (size_t i = 0; i < N; ++i) h[i] = float((i * 2654435761u) >> 8 & 0xffff) / 65536.f;
  Args a{};
  a.n = n; a.iters = 250;
  float* lin; float2* pair; float4* quad;
  cudaMalloc(&lin, N * 4); cudaMalloc(&pair, N * 8); cudaMalloc(&quad, N * 16); cudaMalloc(&a.out, 4);
  cudaMemcpy(lin, h.data(), N * 4, cudaMemcpyHostToDevice);
  cudaMemset(pair, 0, N * 8); cudaMemset(quad, 0, N * 16);
  a.lin = lin; a.pair = pair; a.quad = quad;
  // 3D texture: extent (width = Z, height = Y, depth = X)
  cudaChannelFormatDesc cd = cudaCreateChannelDesc<float>();
  cudaArray_t arr3;
  cudaMalloc3DArray(&arr3, &cd, make_cudaExtent(n, n, n));
  cudaMemcpy3DParms cp = {};
  cp.srcPtr = make_cudaPitchedPtr(h.data(), n * 4, n, n);
  cp.dstArray = arr3; cp.extent = make_cudaExtent(n, n, n); cp.kind = cudaMemcpyHostToDevice;
  cudaMemcpy3D(&cp);
  cudaResourceDesc rd = {}; rd.resType = cudaResourceTypeArray; rd.res.array.array = arr3;
  cudaTextureDesc td = {};
  td.addressMode[0] = td.addressMode[1] = td.addressMode[2] = cudaAddressModeClamp;
  td.filterMode = cudaFilterModePoint; td.readMode = cudaReadModeElementType; td.normalizedCoords = 0;
  cudaCreateTextureObject(&a.tex3, &rd, &td, nullptr);
  // 2D gather texture: width Z, height X*Y
  cudaArray_t arr2;
  cudaMallocArray(&arr2, &cd, n, (size_t)n * n, cudaArrayTextureGather);
  cudaMemcpy2DToArray(arr2, 0, 0, h.data(), n * 4, n * 4, (size_t)n * n, cudaMemcpyHostToDevice);
  rd.res.array.array = arr2;
  cudaCreateTextureObject(&a.tex2, &rd, &td, nullptr);
  printf("setup: %s\n", cudaGetErrorString(cudaGetLastError()));
  for (float pitch : {0.85f, 0.45f}) {
    double tot[7] = {0, 0, 0, 0, 0, 0, 0};
    for (int d = 0; d < 3; ++d) {
      tot[0] += run<0>(a, d, pitch); tot[1] += run<1>(a, d, pitch); tot[2] += run<2>(a, d, pitch);
      tot[3] += run<3>(a, d, pitch); tot[4] += run<4>(a, d, pitch); tot[5] += run<5>(a, d, pitch);
      tot[6] += run<6>(a, d, pitch);
    }
    printf("pitch %.2f, sum over 3 directions (ms):", pitch);
    for (int m = 0; m < 7; ++m) printf(" m%d=%.2f", m, tot[m]);
    printf("\n");
  }
  cudaError_t e = cudaDeviceSynchronize();
  printf("status: %s\n", cudaGetErrorString(e));
  return e != cudaSuccess;
}
