// TEST INFRASTRUCTURE ONLY -- never linked into the product.
//
// extern "C" access to the REFERENCE's own renderer::Volume (renderer/volume.cpp, renderer/volume_implicit.cpp)
// compiled unmodified from /root/reference where it lies, together with the reference's vendored LZ4
// (third-party/lz4/lib/{lz4.c,lz4hc.c,lz4cpp.cpp}), against the installed libtorch (oracle/Makefile target
// `refvol` -> oracle/_ref/libdvr_refvol.so). Used by tests/ to pin the product's implicit-volume kernel and its
// .cvol reader / writer (incl. the LZ4 chunk stream) against the real thing.
// Arrays cross this boundary as dense float [X][Y][Z] (Z fastest); the reference's own in-memory and on-disk
// order is X fastest (renderer/volume.cpp:21-27).
#include <cstring>
#include <memory>
#include <string>
#include <unordered_map>

#include "volume.h"  // from /root/reference/renderer

namespace {
thread_local std::string g_error;
template <typename F>
int guarded(F&& f) {
  try {
    f();
    return 0;
  } catch (const std::exception& e) {
    g_error = e.what();
    return -1;
  }
}
void copy_out(const renderer::Volume& v, float* out) {
  const auto level = v.getLevel(0);
  const auto acc = level->dataCpu().accessor<real_t, 4>();
  const int X = level->sizeX(), Y = level->sizeY(), Z = level->sizeZ();
  for (int x = 0; x < X; ++x)
    for (int y = 0; y < Y; ++y)
      for (int z = 0; z < Z; ++z) out[(size_t(x) * Y + y) * Z + z] = float(acc[0][x][y][z]);
}
}  // namespace

extern "C" const char* refvol_last_error() { return g_error.c_str(); }

// Volume::createImplicitDataset (renderer/volume_implicit.cpp:214-256); out: resolution^3 floats
extern "C" int refvol_create_implicit(int equation, int resolution, int nparams, const char** keys,
                                      const float* values, float* out) {
  return guarded([&] {
    std::unordered_map<std::string, float> params;
    for (int i = 0; i < nparams; ++i) params[keys[i]] = values[i];
    auto v = renderer::Volume::createImplicitDataset(
        resolution, static_cast<renderer::Volume::ImplicitEquation>(equation), params);
    if (!v) throw std::runtime_error("unknown implicit equation");
    copy_out(*v, out);
  });
}

// Volume::createFromBuffer + setWorldSize + save (renderer/volume.cpp:84-151, 389-402)
extern "C" int refvol_save(const char* filename, const float* data, int X, int Y, int Z, double wx, double wy,
                           double wz, int compression) {
  return guarded([&] {
    long long sizes[3] = {X, Y, Z};
    long long strides[3] = {(long long)Y * Z, Z, 1};
    auto v = renderer::Volume::createFromBuffer(data, sizes, strides);
    v->setWorldSizeX(wx);
    v->setWorldSizeY(wy);
    v->setWorldSizeZ(wz);
    v->save(filename, compression);
  });
}

// Volume(filename) (renderer/volume.cpp:162-279). out == nullptr: only the header values are returned.
extern "C" int refvol_load(const char* filename, int* X, int* Y, int* Z, double* world, float* out) {
  return guarded([&] {
    renderer::Volume v{std::string(filename)};
    const int3 r = v.baseResolution();
    *X = r.x; *Y = r.y; *Z = r.z;
    world[0] = v.worldSizeX(); world[1] = v.worldSizeY(); world[2] = v.worldSizeZ();
    if (out) copy_out(v, out);
  });
}

// Volume::createMipmapLevel (renderer/volume.cpp:359-365) is NOT exported: it hands at::adaptive_avg_pool3d_out the
// 4-element sizes() of the 1*X*Y*Z level as output_size, which the installed libtorch (2.11) rejects
// ("output_size must be 3"). The tests pin mipmap levels against the operator it calls,
// torch.nn.functional.adaptive_avg_pool3d, with the three spatial sizes (tests/test_oracle_volume.py).
