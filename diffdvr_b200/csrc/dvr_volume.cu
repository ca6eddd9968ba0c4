// renderer::Volume on the device: the producers of the 1*X*Y*Z density tensor that the
// marcher consumes.
//   k_implicit_volume   Volume::createImplicitDataset (renderer/volume_implicit.cpp:13-256 + the sampling loop of
//                       Volume::createSyntheticDataset, renderer/volume.cpp:369-387), one thread per voxel
//   k_volume_from_raw   the "convert" step of Volume::Volume(filename) (renderer/volume.cpp:232-276): the .cvol body
//                       is X fastest and u8 / u16 / f32; the marcher wants fp32 with Z fastest -> 32x32 (x, z) tiles
//                       transposed through shared memory so that both sides are coalesced
//   k_volume_to_raw     the inverse, for Volume::save (renderer/volume.cpp:86-151)
//   k_downsample_volume Volume::createMipmapLevel = adaptive average pooling (renderer/volume.cpp:359-365)
// This file is compiled with -fmad=false (csrc/Makefile): the reference evaluates the surfaces in plain fp32
// without contraction (x86-64 baseline has no FMA), and with the same operation order the 12 polynomial / sqrt
// surfaces come out BIT-IDENTICAL to the reference's; the three that call sin / cos (MarschnerLobb in fp64, Blobby,
// Tube) agree to the accuracy of the device's libm (tests/test_gpu_volume.py states the tolerances).
// All four are streaming kernels: 4 B written per voxel (implicit), 1|2|4 B read + 4 B written (from_raw).
#include "dvr_launch.h"

namespace dvr {

namespace {

__device__ __forceinline__ float sq(float s) { return s * s; }
__device__ __forceinline__ float cube(float s) { return s * s * s; }
// implicit surface f = 0 -> density 0.5, clamped (volume_implicit.cpp:81-86)
__device__ __forceinline__ float iso(float i) { return fmaxf(0.0f, fminf(1.0f, -i + 0.5f)); }

// NOTE "z2 = z * 2" (not z * z) in Barth, Heart, Kleine, Steiner, CrossCap, Kummer, Blobby is the reference's
// (volume_implicit.cpp:124, 137, 148, 173, 186, 199, 212); parity beats algebra.
__device__ float implicit_value(int eq, float x, float y, float z, const ImplicitArgs& a) {
  switch (eq) {
    case kEqMarschnerLobb: {  // :28-43, evaluated in fp64 from the fp32 position
      const double fM = a.fM, alpha = a.alpha;
      const double r = double(sqrtf(x * x + y * y));
      const double kPi = 3.14159265358979323846;
      const double pr = cos(2 * kPi * fM * cos(kPi * r / 2));
      const double num = (1 - sin(kPi * z / 2)) + alpha * (1 + pr);
      const double den = 2 * (1 + alpha);
      return float(num / den);
    }
    case kEqCube: {  // :45-58
      const float s = a.scale;
      const float d = sqrtf(sq(fmaxf(0.0f, fabsf(x) - s)) + sq(fmaxf(0.0f, fabsf(y) - s)) + sq(fmaxf(0.0f, fabsf(z) - s)));
      return 1 - d;
    }
    case kEqSphere: return 1 - sqrtf(sq(x) + sq(y) + sq(z));        // :60-68
    case kEqInverseSphere: return sqrtf(sq(x) + sq(y) + sq(z));     // :70-78
    case kEqDingDong: return iso(x * x + y * y - z * (1 - z * z));  // :88-96
    case kEqEndrass: {                                              // :98-112
      const float r2 = sqrtf(2.0f);
      const float p = sq(x + y) - 2;
      const float q = sq(x - y) - 2;
      const float c = -4 * (1 - r2);
      const float d = 8 * (2 - r2) * z * z + 2 * (2 - 7 * r2) * (x * x + y * y);
      const float e = -16 * sq(sq(z)) + 8 * (1 + 2 * r2) * sq(z) - 1 + 12 * r2;
      return 0.5f + (64 * (x * x - 1) * (y * y - 1) * p * q - sq(c + d + e));
    }
    case kEqBarth: {  // :114-127
      z += 0.5f;
      const float phi = (1 + sqrtf(5.0f)) / 2;
      const float x2 = x * x, y2 = y * y, z2 = z * 2, phi2 = phi * phi;
      return 0.5f + (4 * (phi2 * x2 - y2) * (phi2 * y2 - z2) * (phi2 * z2 - x2) - (1 + 2 * phi) * sq(x2 + y2 + z2 - 1));
    }
    case kEqHeart: {  // :129-138
      const float x2 = x * x, y2 = y * y, z2 = z * 2;
      return iso(cube(2 * x2 + 2 * y2 + z2 - 1) - 0.1f * x2 * z2 * z - y2 * z2 * z);
    }
    case kEqKleine: {  // :140-150
      const float x2 = x * x, y2 = y * y, z2 = z * 2;
      const float m = x2 + y2 + z2 - 2 * y - 1;
      return 0.5f + -((x2 + y2 + z2 + 2 * y - 1) * sq(m) - 8 * z2 + 16 * x * y * m);
    }
    case kEqCassini: {  // :152-162
      const float c = a.a;
      return iso((sq(x + c) + y * y) * (sq(x - c) + y * y) - z * z);
    }
    case kEqSteiner: {  // :164-174
      const float x2 = x * x, y2 = y * y, z2 = z * 2;
      return iso(x2 * y2 + x2 * z2 + y2 * z2 - 2 * x * y * z);
    }
    case kEqCrossCap: {  // :176-186
      const float x2 = x * x, y2 = y * y, z2 = z * 2;
      return iso(4 * x2 * (x2 + y2 * z2 + z) + y2 * (y2 + z2 - 1));
    }
    case kEqKummer: {  // :188-198
      const float x2 = x * x, y2 = y * y, z2 = z * 2;
      return iso(x2 * x2 + y2 * y2 + z2 * z2 - x2 - y2 - z2 - x2 * y2 - y2 * z2 - z2 * x2 + 1);
    }
    case kEqBlobby: {  // :200-210
      const float x2 = x * x, y2 = y * y, z2 = z * 2;
      return iso(x2 + y2 * z2 + sinf(4 * x) - cosf(4 * y) + sinf(4 * z) - 1);
    }
    default: {  // kEqTube, :212-221
      const float r = sqrtf(y * y + z * z);
      return ((1 - (r * cube(0.9f - 0.5f * cosf(7 * x)))) - 0.9f) * 10;
    }
  }
}

// VEC = 4: Z % 4 == 0, unit Z stride and a 16-byte aligned base -> one thread evaluates four consecutive z and
// issues one 128-bit store (the index arithmetic and the x / y positions are shared by the four voxels).
template <int VEC>
__global__ void __launch_bounds__(256) k_implicit_volume(const ImplicitArgs a) {
  const unsigned zq = unsigned(a.Z) / VEC;
  const unsigned n = unsigned(a.X) * unsigned(a.Y) * zq;
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const unsigned z = (i % zq) * VEC, xy = i / zq;
    const unsigned y = xy % unsigned(a.Y), x = xy / unsigned(a.Y);
    // volume.cpp:380-382: boxMin + index * scale, two roundings
    const float px = a.boxMin + float(x) * a.scaleStep, py = a.boxMin + float(y) * a.scaleStep;
    float v[VEC];
#pragma unroll
    for (int k = 0; k < VEC; ++k) v[k] = implicit_value(a.equation, px, py, a.boxMin + float(z + k) * a.scaleStep, a);
    float* o = a.out + (long long)x * a.sx + (long long)y * a.sy + (long long)z * a.sz;
    if constexpr (VEC == 4) *reinterpret_cast<float4*>(o) = make_float4(v[0], v[1], v[2], v[3]);
    else o[0] = v[0];
  }
}

template <typename T>
__device__ __forceinline__ float raw_to_float(T v);
template <>
__device__ __forceinline__ float raw_to_float<unsigned char>(unsigned char v) { return float(v) / 255.0f; }   // volume.cpp:244
template <>
__device__ __forceinline__ float raw_to_float<unsigned short>(unsigned short v) { return float(v) / 65535.0f; }  // :251
template <>
__device__ __forceinline__ float raw_to_float<float>(float v) { return v; }

// grid (ceil(X/32), Y, ceil(Z/32)), 256 threads = 32 x 8. TO_RAW = false: raw[x + X*(y + Y*z)] -> vol(x, y, z).
template <typename T, bool TO_RAW>
__global__ void __launch_bounds__(256) k_volume_raw(const RawVolumeArgs a) {
  __shared__ float tile[32][33];
  const int x0 = blockIdx.x * 32, z0 = blockIdx.z * 32, y = blockIdx.y;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const long long XY = (long long)a.X * a.Y;
  T* raw = static_cast<T*>(a.raw);
  if constexpr (!TO_RAW) {
    for (int j = ty; j < 32; j += 8) {  // read: lane along x (raw fastest)
      const int x = x0 + tx, z = z0 + j;
      if (x < a.X && z < a.Z) tile[j][tx] = raw_to_float<T>(raw[x + (long long)a.X * y + XY * z]);
    }
    __syncthreads();
    for (int j = ty; j < 32; j += 8) {  // write: lane along z
      const int x = x0 + j, z = z0 + tx;
      if (x < a.X && z < a.Z) a.vol[(long long)x * a.sx + (long long)y * a.sy + (long long)z * a.sz] = tile[tx][j];
    }
  } else {
    for (int j = ty; j < 32; j += 8) {
      const int x = x0 + j, z = z0 + tx;
      if (x < a.X && z < a.Z) tile[tx][j] = a.vol[(long long)x * a.sx + (long long)y * a.sy + (long long)z * a.sz];
    }
    __syncthreads();
    for (int j = ty; j < 32; j += 8) {
      const int x = x0 + tx, z = z0 + j;
      if (x < a.X && z < a.Z) reinterpret_cast<float*>(a.raw)[x + (long long)a.X * y + XY * z] = tile[j][tx];
    }
  }
}

// at::adaptive_avg_pool3d: output cell o covers input [floor(o*I/O), ceil((o+1)*I/O)) per axis
__device__ __forceinline__ int pool_lo(int o, int I, int O) { return int((long long)o * I / O); }
__device__ __forceinline__ int pool_hi(int o, int I, int O) { return int(((long long)(o + 1) * I + O - 1) / O); }

__global__ void __launch_bounds__(256) k_downsample_volume(const DownsampleArgs a) {
  const unsigned n = unsigned(a.oX) * unsigned(a.oY) * unsigned(a.oZ);
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const unsigned z = i % unsigned(a.oZ), xy = i / unsigned(a.oZ);
    const unsigned y = xy % unsigned(a.oY), x = xy / unsigned(a.oY);
    const int xa = pool_lo(x, a.X, a.oX), xb = pool_hi(x, a.X, a.oX);
    const int ya = pool_lo(y, a.Y, a.oY), yb = pool_hi(y, a.Y, a.oY);
    const int za = pool_lo(z, a.Z, a.oZ), zb = pool_hi(z, a.Z, a.oZ);
    float sum = 0.0f;
    for (int ix = xa; ix < xb; ++ix)
      for (int iy = ya; iy < yb; ++iy)
        for (int iz = za; iz < zb; ++iz) sum += __ldg(a.in + (long long)ix * a.sx + (long long)iy * a.sy + (long long)iz * a.sz);
    a.out[(long long)x * a.ox + (long long)y * a.oy + (long long)z * a.oz] = sum / float((xb - xa) * (yb - ya) * (zb - za));
  }
}

unsigned stream_blocks(unsigned n) {
  unsigned blocks = (n + 255) / 256;
  return blocks > 148u * 32u ? 148u * 32u : (blocks ? blocks : 1u);
}

}  // namespace

cudaError_t launch_implicit_volume(const ImplicitArgs& a, cudaStream_t s) {
  const bool vec = a.sz == 1 && a.Z % 4 == 0 && a.sx % 4 == 0 && a.sy % 4 == 0 &&
                   reinterpret_cast<unsigned long long>(a.out) % 16 == 0;
  const unsigned n = unsigned(a.X) * a.Y * a.Z;
  if (vec) return launch_kernel(k_implicit_volume<4>, dim3(stream_blocks(n / 4)), 256, 0, s, a);
  return launch_kernel(k_implicit_volume<1>, dim3(stream_blocks(n)), 256, 0, s, a);
}

cudaError_t launch_volume_raw(int dtype, bool toRaw, const RawVolumeArgs& a, cudaStream_t s) {
  const dim3 grid((a.X + 31) / 32, a.Y, (a.Z + 31) / 32);
  const int block = 256;
  if (toRaw) return launch_kernel(k_volume_raw<float, true>, grid, block, 0, s, a);
  switch (dtype) {
    case 0: return launch_kernel(k_volume_raw<unsigned char, false>, grid, block, 0, s, a);
    case 1: return launch_kernel(k_volume_raw<unsigned short, false>, grid, block, 0, s, a);
    case 2: return launch_kernel(k_volume_raw<float, false>, grid, block, 0, s, a);
    default: return kErrNoInstance;
  }
}

cudaError_t launch_downsample_volume(const DownsampleArgs& a, cudaStream_t s) {
  return launch_kernel(k_downsample_volume, dim3(stream_blocks(unsigned(a.oX) * a.oY * a.oZ)), 256, 0, s, a);
}

}  // namespace dvr
