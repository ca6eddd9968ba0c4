// Measured ceilings of the marchers' memory side on the layout they use today: per CELL VISIT
//   gather : 2 x LDG.128 from the QUAD copy of the volume          (forward, forward-mode, adjoint)
//   red    : 2 x red.global.v4.f32.add into the QUAD gradient scratch (adjoint with hasVolume)
//   both   : the two in one loop (what the adjoint's memory pipes see)
// in the access pattern of the kernels: a warp = 8x4 pixel patch of rays (pitch in voxels between neighbouring
// rays), every iteration advances ~0.9 voxel along the ray, i.e. one new cell per iteration -- nothing else in the
// loop, so the numbers are CEILINGS for cells/s, not models of the kernels. Three ray directions (mostly along z = the
// contiguous axis, along x, along y) because the lines touched per warp instruction depend on it; the headline figure
// is the harmonic mean over the three. Footprints: n = 256 (QUAD buffer 256 MiB) and n = 512 (2 GiB), both > L2.
// SURVEY.md 8(d) "gather_gbs / red_gops stored next to MEASURED_PEAKS.json"; tools/microbench/run_ceilings.py writes
// profiles/MEASURED_GATHER_RED.json from the JSON lines printed here.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o quad_ceilings quad_ceilings.cu
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

__device__ __forceinline__ void red4(float4* p, float a, float b, float c, float d) {
  asm volatile("red.global.v4.f32.add [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

// MODE 0 gather, 1 red, 2 both; XINNER: 0 = Q[x][y][z] (the two quads of a cell n*n entries apart), 1 = Q[y][z][x] (the
// two quads of a cell are neighbours), DVR_QUAD_XINNER of diffdvr_b200/csrc/dvr_params.h
template <int MODE, int XINNER>
__global__ void __launch_bounds__(128) k(const float4* __restrict__ quad, float4* grad, int n, int iters, float pitch,
                                         int dirsel, float* out) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int tx = blockIdx.x * 16 + (warp & 1) * 8 + (lane & 7);
  const int ty = blockIdx.y * 8 + (warp >> 1) * 4 + (lane >> 3);
  const float ox = blockIdx.z * 3.1f;  // "views": shifted copies so that CTAs of different views do not share lines
  float px = tx * pitch + ox, py = ty * pitch, pz = 0.3f;
  float dx = 0.11f, dy = 0.07f, dz = 0.9f;
  if (dirsel == 1) { dx = 0.9f; dz = 0.11f; px = 0.3f; pz = tx * pitch + ox; }
  if (dirsel == 2) { dy = 0.9f; dz = 0.07f; py = 0.3f; pz = ty * pitch; dx = 0.11f; px = tx * pitch + ox; }
  const float sx0 = px, sy0 = py, sz0 = pz;
  float acc = 0.f;
  const size_t sxx = (size_t)n * n;
  const float a = 1.0f / 1024.f;
  for (int i = 0; i < iters; ++i) {
    int x0 = int(px), y0 = int(py), z0 = int(pz);
    if (x0 >= n - 1 || y0 >= n - 1 || z0 >= n - 1) {  // left the volume: restart a little further in
      px = sx0 * 0.5f + 0.4f; py = sy0 * 0.5f + 0.4f; pz = sz0 * 0.5f + 0.4f;
      continue;
    }
    const size_t o = XINNER ? ((size_t)y0 * n + z0) * n + x0 : ((size_t)x0 * n + y0) * n + z0;
    const size_t sx = XINNER ? 1 : sxx;
    if (MODE == 0 || MODE == 2) {
      const float4 q0 = __ldg(quad + o), q1 = __ldg(quad + o + sx);
      acc += q0.x + q0.y + q0.z + q0.w + q1.x + q1.y + q1.z + q1.w;
    }
    if (MODE == 1 || MODE == 2) {
      red4(grad + o, a, a, a, a);
      red4(grad + o + sx, a, a, a, a);
    }
    px += dx; py += dy; pz += dz;
  }
  if (acc == 12345.678f) out[0] = acc;  // keep the loads alive
}

template <int MODE, int XINNER>
double run(const float4* quad, float4* grad, int n, int res, int views, int iters, float pitch, int dir, float* out) {
  dim3 grid(res / 16, res / 8, views);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  k<MODE, XINNER><<<grid, 128>>>(quad, grad, n, iters, pitch, dir, out);  // warm-up
  cudaDeviceSynchronize();
  float best = 1e30f;
  for (int rep = 0; rep < 3; ++rep) {
    cudaEventRecord(e0);
    k<MODE, XINNER><<<grid, 128>>>(quad, grad, n, iters, pitch, dir, out);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    best = ms < best ? ms : best;
  }
  cudaEventDestroy(e0); cudaEventDestroy(e1);
  const double cells = double(res) * res * views * iters;
  return cells / best * 1e-6;  // Gcells/s
}

template <int XINNER>
int measure(float* out) {
  for (int n : {256, 512}) {
    const size_t N = (size_t)n * n * n;
    float4 *quad, *grad;
    if (cudaMalloc(&quad, N * 16) != cudaSuccess || cudaMalloc(&grad, N * 16) != cudaSuccess) {
      printf("{\"n\": %d, \"error\": \"out of memory\"}\n", n);
      return 1;
    }
    cudaMemset(quad, 0, N * 16);
    cudaMemset(grad, 0, N * 16);
    const int res = n == 256 ? 512 : 1024, views = n == 256 ? 8 : 4, iters = 250;
    const float pitch = 0.6f;  // voxels between neighbouring rays at the volume centre for both C4 and C5
    double g[3], r[3], b[3], hg = 0, hr = 0, hb = 0;
    for (int d = 0; d < 3; ++d) {
      g[d] = run<0, XINNER>(quad, grad, n, res, views, iters, pitch, d, out);
      r[d] = run<1, XINNER>(quad, grad, n, res, views, iters, pitch, d, out);
      b[d] = run<2, XINNER>(quad, grad, n, res, views, iters, pitch, d, out);
      hg += 1.0 / g[d]; hr += 1.0 / r[d]; hb += 1.0 / b[d];
    }
    cudaError_t e = cudaDeviceSynchronize();
    printf("{\"layout\": \"%s\", \"n\": %d, \"footprint_mib\": %zu, \"rays\": %d, \"pitch_voxels\": %.2f, "
           "\"gather_gcells_s\": %.2f, \"red_gcells_s\": %.2f, \"gather_red_gcells_s\": %.2f, "
           "\"gather_by_dir_zxy\": [%.2f, %.2f, %.2f], \"red_by_dir_zxy\": [%.2f, %.2f, %.2f], "
           "\"gather_red_by_dir_zxy\": [%.2f, %.2f, %.2f], \"status\": \"%s\"}\n",
           XINNER ? "yzx" : "xyz", n, N * 16 >> 20, res * res * views, pitch, 3.0 / hg, 3.0 / hr, 3.0 / hb, g[0], g[1], g[2],
           r[0], r[1], r[2], b[0], b[1], b[2], cudaGetErrorString(e));
    cudaFree(quad); cudaFree(grad);
    if (e != cudaSuccess) return 1;
  }
  return 0;
}

// usage: quad_ceilings [xyz | yzx | both]   (default: both layouts, one JSON line per layout and volume size)
int main(int argc, char** argv) {
  float* out;
  cudaMalloc(&out, 4);
  const char* which = argc > 1 ? argv[1] : "both";
  int rc = 0;
  if (which[0] == 'x' || which[0] == 'b') rc |= measure<0>(out);
  if (which[0] == 'y' || which[0] == 'b') rc |= measure<1>(out);
  return rc;
}
