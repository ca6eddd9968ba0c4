// TEST INFRASTRUCTURE ONLY -- never linked into the product.
//
// CPU restatement of renderer::Volume's producers of the density tensor:
//   create_implicit_volume  Volume::createImplicitDataset (renderer/volume_implicit.cpp:13-256) over the sampling loop
//                           of Volume::createSyntheticDataset (renderer/volume.cpp:369-387)
//   volume_from_raw / to_raw  the X-fastest u8 | u16 | f32 body of a .cvol file <-> the tensor
//                           (renderer/volume.cpp:86-151, 232-276)
//   downsample_volume       Volume::createMipmapLevel = at::adaptive_avg_pool3d (renderer/volume.cpp:359-365)
// The surfaces are evaluated in fp32 (MarschnerLobb: fp64) whatever R is, like the reference, whose eval() functions
// return float even in the USE_DOUBLE_PRECISION build. Compiled with -ffp-contract=off.
// Included by dvr_oracle.cpp. Pinned by tests/test_oracle_volume.py against oracle/_ref/libdvr_refvol.so
// (ref_shim_volume.cpp: the reference's own volume.cpp + volume_implicit.cpp, compiled unmodified) and the fixtures
// tests/golden/ref_volume_*.npz / *.cvol.
#pragma once
#include <cmath>

namespace dvro {

struct ImplicitParams { float fM, alpha, scale, a; };

static inline float vsq(float s) { return s * s; }
static inline float vcb(float s) { return s * s * s; }
static inline float surface_density(float i) { return std::max(0.0f, std::min(1.0f, -i + 0.5f)); }  // :81-86

// half extent of the sampling box per equation (boxMin = -h, boxMax = +h)
static const float kImplicitBox[15] = {1, 1, 1, 1, 2, 2, 1.5f, 1, 5, 1, 0.5f, 1, 2, 2, 1};

static float implicit_value(int eq, float x, float y, float z, const ImplicitParams& p) {
  const float x2 = x * x, y2 = y * y;
  // the reference's "z2" of equations 6..13 is z * 2 (volume_implicit.cpp:124, 137, 148, 173, 186, 199, 212)
  switch (eq) {
    case 0: {  // MarschnerLobb :28-43
      const double fM = p.fM, alpha = p.alpha;
      const double r = std::sqrt(x * x + y * y);  // float sqrt, widened
      const double pr = std::cos(2 * M_PI * fM * std::cos(M_PI * r / 2));
      return float(((1 - std::sin(M_PI * z / 2)) + alpha * (1 + pr)) / (2 * (1 + alpha)));
    }
    case 1: {  // Cube :45-58
      const float ex = std::max(0.0f, std::abs(x) - p.scale), ey = std::max(0.0f, std::abs(y) - p.scale),
                  ez = std::max(0.0f, std::abs(z) - p.scale);
      return 1 - std::sqrt(vsq(ex) + vsq(ey) + vsq(ez));
    }
    case 2: return 1 - std::sqrt(x2 + y2 + vsq(z));
    case 3: return std::sqrt(x2 + y2 + vsq(z));
    case 4: return surface_density(x2 + y2 - z * (1 - z * z));
    case 5: {  // Endrass :98-112
      const float s2 = sqrtf(2);
      const float a = vsq(x + y) - 2, b = vsq(x - y) - 2, c = -4 * (1 - s2);
      const float d = 8 * (2 - s2) * z * z + 2 * (2 - 7 * s2) * (x2 + y2);
      const float e = -16 * vsq(vsq(z)) + 8 * (1 + 2 * s2) * vsq(z) - 1 + 12 * s2;
      return 0.5f + (64 * (x2 - 1) * (y2 - 1) * a * b - vsq(c + d + e));
    }
    case 6: {  // Barth :114-127
      const float zs = z + 0.5f, phi = (1 + sqrtf(5)) / 2, z2 = zs * 2, phi2 = phi * phi;
      return 0.5f + (4 * (phi2 * x2 - y2) * (phi2 * y2 - z2) * (phi2 * z2 - x2) - (1 + 2 * phi) * vsq(x2 + y2 + z2 - 1));
    }
    case 7: {  // Heart
      const float z2 = z * 2;
      return surface_density(vcb(2 * x2 + 2 * y2 + z2 - 1) - 0.1f * x2 * z2 * z - y2 * z2 * z);
    }
    case 8: {  // Kleine
      const float z2 = z * 2, m = x2 + y2 + z2 - 2 * y - 1;
      return 0.5f + -((x2 + y2 + z2 + 2 * y - 1) * vsq(m) - 8 * z2 + 16 * x * y * m);
    }
    case 9: return surface_density((vsq(x + p.a) + y2) * (vsq(x - p.a) + y2) - z * z);  // Cassini
    case 10: {  // Steiner
      const float z2 = z * 2;
      return surface_density(x2 * y2 + x2 * z2 + y2 * z2 - 2 * x * y * z);
    }
    case 11: {  // CrossCap
      const float z2 = z * 2;
      return surface_density(4 * x2 * (x2 + y2 * z2 + z) + y2 * (y2 + z2 - 1));
    }
    case 12: {  // Kummer
      const float z2 = z * 2;
      return surface_density(x2 * x2 + y2 * y2 + z2 * z2 - x2 - y2 - z2 - x2 * y2 - y2 * z2 - z2 * x2 + 1);
    }
    case 13: {  // Blobby
      const float z2 = z * 2;
      return surface_density(x2 + y2 * z2 + sinf(4 * x) - cosf(4 * y) + sinf(4 * z) - 1);
    }
    case 14: {  // Tube
      const float r = sqrtf(y2 + z * z);
      return ((1 - (r * vcb(0.9f - 0.5f * cosf(7 * x)))) - 0.9f) * 10;
    }
    default: throw std::runtime_error("unknown implicit equation");
  }
}

template <typename R>
static void create_implicit_volume(int eq, const float* params, const DvrTensor& out) {
  if (eq < 0 || eq > 14) throw std::runtime_error("unknown implicit equation");
  const ImplicitParams p = params ? ImplicitParams{params[0], params[1], params[2], params[3]}
                                  : ImplicitParams{6.0f, 0.25f, 0.5f, 0.25f};
  const View<R> v(out, 4, false);
  const int n = int(v.size[1]);
  if (v.size[2] != n || v.size[3] != n) throw std::runtime_error("implicit volumes are cubes");
  const float lo = -kImplicitBox[eq], step = (kImplicitBox[eq] - lo) / (n - 1);
  for (int x = 0; x < n; ++x)
    for (int y = 0; y < n; ++y)
      for (int z = 0; z < n; ++z) v.at(0, x, y, z) = R(implicit_value(eq, lo + x * step, lo + y * step, lo + z * step, p));
}

template <typename R>
static void volume_from_raw(const void* raw, int dtype, const DvrTensor& out) {
  const View<R> v(out, 4, false);
  const int64_t X = v.size[1], Y = v.size[2], Z = v.size[3];
  for (int64_t z = 0; z < Z; ++z)
    for (int64_t y = 0; y < Y; ++y)
      for (int64_t x = 0; x < X; ++x) {
        const int64_t i = x + X * (y + Y * z);
        float f;
        switch (dtype) {
          case 0: f = static_cast<const unsigned char*>(raw)[i] / 255.0f; break;
          case 1: f = static_cast<const unsigned short*>(raw)[i] / 65535.0f; break;
          case 2: f = static_cast<const float*>(raw)[i]; break;
          default: throw std::runtime_error("unknown .cvol data type");
        }
        v.at(0, x, y, z) = R(f);
      }
}

template <typename R>
static void volume_to_raw(const DvrTensor& vol, float* raw) {
  const View<const R> v(vol, 4, false);
  const int64_t X = v.size[1], Y = v.size[2], Z = v.size[3];
  for (int64_t z = 0; z < Z; ++z)
    for (int64_t y = 0; y < Y; ++y)
      for (int64_t x = 0; x < X; ++x) raw[x + X * (y + Y * z)] = float(v.at(0, x, y, z));
}

template <typename R>
static void downsample_volume(const DvrTensor& vol, const DvrTensor& out) {
  const View<const R> v(vol, 4, false);
  const View<R> o(out, 4, false);
  auto lo = [](int64_t i, int64_t I, int64_t O) { return i * I / O; };
  auto hi = [](int64_t i, int64_t I, int64_t O) { return ((i + 1) * I + O - 1) / O; };
  for (int64_t x = 0; x < o.size[1]; ++x)
    for (int64_t y = 0; y < o.size[2]; ++y)
      for (int64_t z = 0; z < o.size[3]; ++z) {
        const int64_t xa = lo(x, v.size[1], o.size[1]), xb = hi(x, v.size[1], o.size[1]);
        const int64_t ya = lo(y, v.size[2], o.size[2]), yb = hi(y, v.size[2], o.size[2]);
        const int64_t za = lo(z, v.size[3], o.size[3]), zb = hi(z, v.size[3], o.size[3]);
        R sum = 0;
        for (int64_t i = xa; i < xb; ++i)
          for (int64_t j = ya; j < yb; ++j)
            for (int64_t k = za; k < zb; ++k) sum += v.at(0, i, j, k);
        o.at(0, x, y, z) = sum / R((xb - xa) * (yb - ya) * (zb - za));
      }
}

}  // namespace dvro
