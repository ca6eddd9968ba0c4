// Forward-variables -> gradients kernel instance and launcher.
#include "dvr_fwdgrad.cuh"
#include "dvr_launch.h"

namespace dvr {

// ===============================================================================================
// forward variables -> gradients (renderer_kernels.cuh:504-561): per pixel <adjColor, J[:,d,:]>;
// stepsize / camera adjoints are STORED, TF adjoints are summed over all pixels. The TF sum is
// pre-reduced per warp (shuffle) and per CTA (shared memory) before R*C global atomics per CTA
// (the reference issues R*C global atomics per PIXEL).
// ===============================================================================================
__global__ void __launch_bounds__(256) k_fwdvars_to_gradients(const FwdVarsArgs a) {
  extern __shared__ float s_acc[];  // R*C
  const int entries = a.hasTf ? a.R * a.C : 0;
  for (int i = threadIdx.x; i < entries; i += blockDim.x) s_acc[i] = 0.f;
  __syncthreads();
  const int b = blockIdx.z;
  const long long pix = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const bool active = pix < (long long)a.W * a.H;
  const int y = active ? int(pix / a.W) : 0, x = active ? int(pix - (long long)y * a.W) : 0;
  f4 g = {0.f, 0.f, 0.f, 0.f};
  const float* fv = a.fv + b * a.fB + y * a.fY + x * a.fX;
  if (active) {
    const float* gp = a.gradColor.p + b * a.gradColor.b + y * a.gradColor.y + x * a.gradColor.x;
    g = {__ldg(gp), __ldg(gp + a.gradColor.c), __ldg(gp + 2 * a.gradColor.c), __ldg(gp + 3 * a.gradColor.c)};
  }
  auto contract = [&](int d) {
    const float* p = fv + d * a.fD;
    return g.x * __ldg(p) + g.y * __ldg(p + a.fC) + g.z * __ldg(p + 2 * a.fC) + g.w * __ldg(p + 3 * a.fC);
  };
  if (active) {
    if (a.dStep >= 0) a.adjStep[b * a.asB + y * a.asY + x * a.asX] = contract(a.dStep);
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      if (a.dRayStart[c] >= 0) a.adjRayStart[b * a.rsB + y * a.rsY + x * a.rsX + c * a.rsC] = contract(a.dRayStart[c]);
      if (a.dRayDir[c] >= 0) a.adjRayDir[b * a.rdB + y * a.rdY + x * a.rdX + c * a.rdC] = contract(a.dRayDir[c]);
    }
  }
  if (a.hasTf) {
    const int lane = threadIdx.x & 31;
    for (int e = 0; e < entries; ++e) {  // uniform loop, all lanes take part in the shuffles
      const int r = e / a.C, c = e - r * a.C;
      const int d = __ldg(a.dTf + b * a.dtB + r * a.dtR + c * a.dtC);
      if (d < 0) continue;  // uniform
      const float v = warp_sum(active ? contract(d) : 0.f);
      if (lane == 0 && v != 0.f) atomicAdd(s_acc + e, v);
    }
    __syncthreads();
    for (int e = threadIdx.x; e < entries; e += blockDim.x) {
      const int r = e / a.C, c = e - r * a.C;
      if (s_acc[e] != 0.f) atomicAdd(a.adjTf + b * a.atB + r * a.atR + c * a.atC, s_acc[e]);
    }
  }
}

// counts samples and cell visits of every ray (measurement utility, see dvr_count_cell_visits)
__global__ void __launch_bounds__(kThreads) k_count_cells(const RayInputs in, const int* term, long long tB,
                                                          long long tY, long long tX, unsigned long long* counts) {
  int x, y;
  thread_pixel(x, y);
  const int b = blockIdx.z;
  unsigned ns = 0, nc = 0;
  if (x < in.W && y < in.H) {
    const Ray ray = make_ray(in, x, y, b);
    const int n = __ldg(term + b * tB + y * tY + x * tX);
    int kx = -1, ky = -1, kz = -1;
    float fi = 0.f;
    for (int i = 0; i < n; ++i, fi += 1.0f) {
      const float t = sample_t(ray, fi);
      const int cx = __float_as_int(biased_trunc(fmaf(t, ray.pd.x, ray.ps.x)));
      const int cy = __float_as_int(biased_trunc(fmaf(t, ray.pd.y, ray.ps.y)));
      const int cz = __float_as_int(biased_trunc(fmaf(t, ray.pd.z, ray.ps.z)));
      if (cx != kx || cy != ky || cz != kz) { ++nc; kx = cx; ky = cy; kz = cz; }
    }
    ns = unsigned(n);
  }
  const unsigned ws = __reduce_add_sync(0xffffffffu, ns), wc = __reduce_add_sync(0xffffffffu, nc);
  if ((threadIdx.x & 31) == 0) {
    atomicAdd(counts, (unsigned long long)ws);
    atomicAdd(counts + 1, (unsigned long long)wc);
  }
}

// QUAD copy, see RayInputs::quads. One thread per quad: 4 strided reads (L1/L2 hits after the first touch of
// a line), one 16-byte write; a streaming pass over 5 x the volume size. Slice X duplicates slice X-1.
__global__ void __launch_bounds__(256) k_build_quads(const RayInputs in, float4* quads) {
  const unsigned YZ = unsigned(in.Y) * unsigned(in.Z);
  const unsigned n = (unsigned(in.X) + 1u) * YZ;
  const unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int b = blockIdx.z;
  const unsigned z = i % unsigned(in.Z), xy = i / unsigned(in.Z);
  const unsigned y = xy % unsigned(in.Y), x = min(xy / unsigned(in.Y), unsigned(in.X) - 1u);
  const unsigned y1 = min(y + 1, unsigned(in.Y) - 1), z1 = min(z + 1, unsigned(in.Z) - 1);
  const float* v = in.volume + b * in.volB + (long long)x * in.sx;
  float4 q;
  q.x = __ldg(v + (long long)y * in.sy + (long long)z * in.sz);
  q.y = __ldg(v + (long long)y * in.sy + (long long)z1 * in.sz);
  q.z = __ldg(v + (long long)y1 * in.sy + (long long)z * in.sz);
  q.w = __ldg(v + (long long)y1 * in.sy + (long long)z1 * in.sz);
  quads[(size_t)b * n + i] = q;
}

// Folds the QUAD gradient scratch into the caller's adj_volume (accumulating, like the atomics it replaces).
// Component k of G[x][y][z] targets voxel (min(x, X-1), y + (k>>1 clamped), z + (k&1 clamped)); as a gather per voxel
// and slice: .x of (y, z), .y of (y, z-1), .z of (y-1, z), .w of (y-1, z-1), plus the components that the clamp at
// the far faces folds back onto the voxel itself (z = Z-1: .y/.w of the same z; y = Y-1: .z/.w of the same y);
// voxels of the last slice also collect slice X (the x+1 corners of cells on the far x face).
__global__ void __launch_bounds__(256) k_fold_quads(const RayInputs in, const AdjointArgs aa) {
  const unsigned Z = unsigned(in.Z), Y = unsigned(in.Y), X = unsigned(in.X);
  const unsigned n = X * Y * Z;
  const unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int b = blockIdx.z;
  const unsigned z = i % Z, xy = i / Z;
  const unsigned y = xy % Y, x = xy / Y;
  const float4* g = aa.gradQuads + (size_t)b * aa.gradQuadB;
  const bool zl = z == Z - 1, yl = y == Y - 1;
  auto gather = [&](unsigned j) {
    const float4 q = g[j];
    float s = q.x;
    if (zl) s += q.y;
    if (yl) s += q.z;
    if (zl && yl) s += q.w;
    if (z > 0) {
      const float4 r = g[j - 1];
      s += r.y;
      if (yl) s += r.w;
    }
    if (y > 0) {
      const float4 r = g[j - Z];
      s += r.z;
      if (zl) s += r.w;
    }
    if (y > 0 && z > 0) s += g[j - Z - 1].w;
    return s;
  };
  float s = gather(i);
  if (x == X - 1) s += gather(i + Y * Z);
  if (s != 0.f) {
    float* p = aa.adjVolume + b * aa.avB + (long long)x * aa.avx + (long long)y * aa.avy + (long long)z * aa.avz;
    *p += s;
  }
}

// ---- the same two passes for the x-innermost layout Q[b][y][z][x] (DVR_QUAD_XINNER) ----------------------------------
// Both are transposes: the volume and adj_volume are z-fastest (contiguous torch tensors; any strides are handled, those
// are the coalesced ones), the quads x-fastest. One CTA of 256 threads owns a 32 (x) x 32 (z) tile of one y row and
// goes through shared memory, so both sides move whole lines.
constexpr int kTile = 32;

__global__ void __launch_bounds__(256) k_build_quads_xinner(const RayInputs in, float4* quads, unsigned tilesX,
                                                            unsigned tilesZ) {
  __shared__ float t[2][kTile][kTile + 1];  // [row y / y'][x][z0 .. z0 + 32]; odd pitch: conflict-free both ways
  const unsigned X = unsigned(in.X), Y = unsigned(in.Y), Z = unsigned(in.Z);
  unsigned blk = blockIdx.x;
  const unsigned tx = blk % tilesX; blk /= tilesX;
  const unsigned tz = blk % tilesZ;
  const unsigned y = blk / tilesZ;
  const unsigned x0 = tx * kTile, z0 = tz * kTile;
  const unsigned y1 = min(y + 1, Y - 1);
  const int b = blockIdx.z;
  const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const float* v = in.volume + b * in.volB;
  // load: lanes along z (33 values per row: z0 .. z0 + 32, clamped to Z - 1), rows = (y | y') x 32 x-values (slice X
  // duplicates slice X - 1)
  for (unsigned r = warp; r < 2 * kTile; r += 8) {
    const unsigned xs = min(x0 + (r & (kTile - 1)), X - 1), yy = r < kTile ? y : y1;
    const float* row = v + (long long)xs * in.sx + (long long)yy * in.sy;
    t[r >> 5][r & (kTile - 1)][lane] = __ldg(row + (long long)min(z0 + lane, Z - 1) * in.sz);
    if (lane == 0) t[r >> 5][r & (kTile - 1)][kTile] = __ldg(row + (long long)min(z0 + kTile, Z - 1) * in.sz);
  }
  __syncthreads();
  // store: lanes along x
  const unsigned x = x0 + lane;
  if (x > X) return;
  float4* out = quads + (size_t)b * ((size_t)(X + 1) * Y * Z) + ((size_t)y * Z) * (X + 1) + x;
  for (unsigned zl = warp; zl < kTile; zl += 8) {
    const unsigned z = z0 + zl;
    if (z >= Z) break;
    out[(size_t)z * (X + 1)] = make_float4(t[0][lane][zl], t[0][lane][zl + 1], t[1][lane][zl], t[1][lane][zl + 1]);
  }
}

// same sums in the same order as k_fold_quads (bit-identical results for the same scratch contents)
__global__ void __launch_bounds__(256) k_fold_quads_xinner(const RayInputs in, const AdjointArgs aa, unsigned tilesX,
                                                           unsigned tilesZ) {
  __shared__ float t[kTile][kTile + 1];  // [z][x]
  const unsigned X = unsigned(in.X), Y = unsigned(in.Y), Z = unsigned(in.Z);
  unsigned blk = blockIdx.x;
  const unsigned tx = blk % tilesX; blk /= tilesX;
  const unsigned tz = blk % tilesZ;
  const unsigned y = blk / tilesZ;
  const unsigned x0 = tx * kTile, z0 = tz * kTile;
  const int b = blockIdx.z;
  const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const unsigned SZ = X + 1, SY = SZ * Z;
  const float4* g = aa.gradQuads + (size_t)b * aa.gradQuadB;
  const bool yl = y == Y - 1;
  // gather: lanes along x (coalesced 16-byte reads)
  for (unsigned zl = warp; zl < kTile; zl += 8) {
    const unsigned z = z0 + zl, x = x0 + lane;
    float s = 0.f;
    if (z < Z && x < X) {
      const bool zlast = z == Z - 1;
      auto gather = [&](size_t j) {
        const float4 q = g[j];
        float a = q.x;
        if (zlast) a += q.y;
        if (yl) a += q.z;
        if (zlast && yl) a += q.w;
        if (z > 0) {
          const float4 r = g[j - SZ];
          a += r.y;
          if (yl) a += r.w;
        }
        if (y > 0) {
          const float4 r = g[j - SY];
          a += r.z;
          if (zlast) a += r.w;
        }
        if (y > 0 && z > 0) a += g[j - SY - SZ].w;
        return a;
      };
      const size_t j = (size_t)y * SY + (size_t)z * SZ + x;
      s = gather(j);
      if (x == X - 1) s += gather(j + 1);
    }
    t[zl][lane] = s;
  }
  __syncthreads();
  // scatter: lanes along z
  const unsigned z = z0 + lane;
  if (z >= Z) return;
  for (unsigned xl = warp; xl < kTile; xl += 8) {
    const unsigned x = x0 + xl;
    if (x >= X) break;
    const float s = t[lane][xl];
    if (s != 0.f) {
      float* p = aa.adjVolume + b * aa.avB + (long long)x * aa.avx + (long long)y * aa.avy + (long long)z * aa.avz;
      *p += s;
    }
  }
}

cudaError_t launch_fold_quads(const RayInputs& in, const AdjointArgs& aa, int nb, cudaStream_t s) {
#if DVR_QUAD_XINNER
  const unsigned tilesX = (unsigned(in.X) + kTile - 1) / kTile, tilesZ = (unsigned(in.Z) + kTile - 1) / kTile;
  return launch_kernel(k_fold_quads_xinner, dim3(tilesX * tilesZ * unsigned(in.Y), 1, nb), 256, 0, s, in, aa, tilesX, tilesZ);
#else
  const unsigned n = unsigned(in.X) * unsigned(in.Y) * unsigned(in.Z);
  return launch_kernel(k_fold_quads, dim3((n + 255) / 256, 1, nb), 256, 0, s, in, aa);
#endif
}

cudaError_t launch_build_quads(const RayInputs& in, int nb, float4* quads, cudaStream_t s) {
#if DVR_QUAD_XINNER
  const unsigned tilesX = (unsigned(in.X) + 1u + kTile - 1) / kTile, tilesZ = (unsigned(in.Z) + kTile - 1) / kTile;
  return launch_kernel(k_build_quads_xinner, dim3(tilesX * tilesZ * unsigned(in.Y), 1, nb), 256, 0, s, in, quads, tilesX,
                       tilesZ);
#else
  const unsigned n = (unsigned(in.X) + 1u) * unsigned(in.Y) * unsigned(in.Z);
  return launch_kernel(k_build_quads, dim3((n + 255) / 256, 1, nb), 256, 0, s, in, quads);
#endif
}

cudaError_t launch_count_cells(const RayInputs& in, const int* term, long long tB, long long tY, long long tX,
                               unsigned long long* counts, dim3 grid, cudaStream_t s) {
  return launch_kernel(k_count_cells, grid, kThreads, 0, s, in, term, tB, tY, tX, counts);
}

// zero-fill of a strided B*H*W*D*4 Jacobian (dense ones take cudaMemsetAsync)
__global__ void k_zero_jacobian(const ForwardGradArgs ga, int W, int H, int B) {
  const long long total = (long long)B * H * W * ga.D * 4;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    long long r = i;
    const int c = int(r & 3); r >>= 2;
    const int d = int(r % ga.D); r /= ga.D;
    const int x = int(r % W); r /= W;
    const int y = int(r % H); r /= H;
    ga.grad[r * ga.gB + y * ga.gY + x * ga.gX + d * ga.gD + c * ga.gC] = 0.f;
  }
}

cudaError_t launch_zero_jacobian(const ForwardGradArgs& ga, int W, int H, int B, cudaStream_t s) {
  const long long total = (long long)B * H * W * ga.D * 4;
  const unsigned blocks = unsigned(total / 256 + 1 < 148 * 16 ? total / 256 + 1 : 148 * 16);
  return launch_kernel(k_zero_jacobian, dim3(blocks), 256, 0, s, ga, W, H, B);
}

cudaError_t launch_fwdvars_to_gradients(const FwdVarsArgs& a, cudaStream_t s) {
  const long long pixels = (long long)a.W * a.H;
  const dim3 grid((unsigned)((pixels + 255) / 256), 1, a.B);
  const size_t smem = size_t(a.hasTf ? a.R * a.C : 0) * sizeof(float);
  return launch_kernel(k_fwdvars_to_gradients, grid, 256, smem, s, a);
}

}  // namespace dvr
