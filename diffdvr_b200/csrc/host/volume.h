// renderer::Volume of the reference (renderer/volume.h:20-236): the container of the
// 1*X*Y*Z density tensor the marcher consumes, its .cvol file format and its synthetic data sets.
// B200-first differences, none of them visible in values or shapes:
//  * the DEVICE copy is the primary one. Implicit data sets are evaluated by a CUDA kernel
//    (dvr_create_implicit_volume), a .cvol body is uploaded as stored (u8 / u16 / f32, X fastest) and converted +
//    transposed on the device (dvr_volume_from_raw), mipmap levels are pooled on the device
//    (dvr_downsample_volume). dataCpu() is a lazily made host copy.
//  * tensors are contiguous with Z fastest (what the marcher's loads want); the reference's are X fastest
//    (renderer/volume.cpp:21-27). Strides are not part of the interface, the marcher takes either.
//  * getDataGpu works without a previous copy_to_gpu(); copy_to_gpu() is kept and is a no-op.
// Not provided: the histogram for the GUI's TF editor (extractHistogram) and the progress / logging callbacks.
#pragma once
#include <istream>
#include <memory>
#include <string>
#include <unordered_map>
#include <vector>

#include "renderer.h"

namespace renderer {

// the compressed body of a .cvol file: a stream of [int32 n][LZ4 block of n bytes] chunks (host code, no device)
namespace cvol {
std::vector<uint8_t> compressBody(const uint8_t* src, size_t bytes, int level);
void decompressBody(std::istream& s, uint8_t* dst, size_t bytes, const std::string& what);
}  // namespace cvol

class Volume {
 public:
  // renderer/volume.h:163-181
  enum class ImplicitEquation {
    MARSCHNER_LOBB, CUBE, SPHERE, INVERSE_SPHERE, DING_DONG, ENDRASS, BARTH, HEART, KLEINE, CASSINI, STEINER,
    CROSS_CAP, KUMMER, BLOBBY, TUBE, _NUM_IMPLICIT_EQUATION_
  };
  static constexpr int NO_COMPRESSION = 0;
  static constexpr int MAX_COMPRESSION = 9;

  Volume(int sizeX, int sizeY, int sizeZ);       // zero-filled
  explicit Volume(const std::string& filename);  // loads a .cvol file
  // compression 0: raw body; 1..9: LZ4 chunk stream (1, 2: fast match finder, 3..9: hash chains)
  void save(const std::string& filename, int compression = NO_COMPRESSION) const;

  static std::shared_ptr<Volume> createImplicitDataset(int resolution, ImplicitEquation equation,
                                                       const std::unordered_map<std::string, float>& params = {});
  // buffer[x * strides[0] + y * strides[1] + z * strides[2]], HOST memory (Volume.from_numpy)
  static std::shared_ptr<Volume> createFromBuffer(const float* buffer, const long long sizes[3],
                                                  const long long strides[3]);
  // wraps an existing 1*X*Y*Z (or X*Y*Z) float tensor; a CUDA tensor is used as it is, a CPU tensor is uploaded
  static std::shared_ptr<Volume> createFromTensor(const torch::Tensor& data);

  // level L has resolution max(1, base / (L + 1)) per axis, adaptive average pooling of level 0
  void createMipmapLevel(int level);
  void deleteAllMipmapLevels();
  bool hasLevel(int level) const { return level >= 0 && size_t(level) < levels_.size() && levels_[level].gpu.defined(); }

  const torch::Tensor& dataGpu(int level) const;
  const torch::Tensor& dataCpu(int level) const;
  void copyCpuToGpu() const {}

  double worldSizeX() const { return worldSize_[0]; }
  double worldSizeY() const { return worldSize_[1]; }
  double worldSizeZ() const { return worldSize_[2]; }
  void setWorldSize(double x, double y, double z) { worldSize_[0] = x; worldSize_[1] = y; worldSize_[2] = z; }
  // X, Y, Z of level 0
  int3 baseResolution() const;

 private:
  Volume() = default;
  struct Level {
    torch::Tensor gpu;
    mutable torch::Tensor cpu;
  };
  std::vector<Level> levels_;
  double worldSize_[3] = {1, 1, 1};
};

}  // namespace renderer
