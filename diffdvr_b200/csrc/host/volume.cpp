// renderer::Volume over the C-ABI; see volume.h. File format and behaviour follow renderer/volume.cpp:64-279.
#include "volume.h"

#include <c10/cuda/CUDAGuard.h>
#include <c10/cuda/CUDAStream.h>
#include <torch/torch.h>

#include <cstring>
#include <fstream>
#include <stdexcept>

#include "../../../include/dvr_b200.h"
#include "lz4_block.h"

namespace renderer {
namespace {

DvrTensor view_of(const torch::Tensor& t) {
  DvrTensor v{};
  v.ptr = t.data_ptr();
  v.ndim = int32_t(t.dim());
  for (int64_t i = 0; i < t.dim(); ++i) {
    v.size[i] = t.size(i);
    v.stride[i] = t.stride(i);
  }
  return v;
}
void raise_on_error(int rc) {
  if (rc == DVR_OK) return;
  const std::string msg = dvr_last_error();
  if (rc == DVR_ERR_INVALID_ARGUMENT) TORCH_CHECK(false, msg);
  throw std::runtime_error(msg);
}
void* current_stream() { return c10::cuda::getCurrentCUDAStream().stream(); }
at::TensorOptions on_gpu(c10::ScalarType t = c10::ScalarType::Float) {
  TORCH_CHECK(torch::cuda::is_available(),
              "renderer::Volume needs a CUDA device: this build has no CPU back-end (the hot path is CUDA for sm_100a only)");
  return at::TensorOptions().dtype(t).device(c10::kCUDA);
}

// ---- .cvol: 64-byte header + body (renderer/volume.cpp:64-84) ------------------------------------
const char kMagic[4] = {'c', 'v', 'o', 'l'};
enum CvolType : int32_t { kU8 = 0, kU16 = 1, kF32 = 2 };
const int kBytesPerType[3] = {1, 2, 4};
#pragma pack(push, 1)
struct CvolHeader {
  char magic[4];
  uint64_t size[3];
  double voxelSize[3];
  int32_t type;
  uint8_t compressed;
  uint8_t padding[7];
};
#pragma pack(pop)
static_assert(sizeof(CvolHeader) == 64, ".cvol header is 64 bytes");

}  // namespace

namespace cvol {

// [int32 n][LZ4 block of n bytes] per chunk of <= 64 KiB of input (lz4cpp.cpp:49-73; volume.cpp:128-139)
std::vector<uint8_t> compressBody(const uint8_t* src, size_t bytes, int level) {
  lz4::Encoder enc(src, bytes, level);
  std::vector<uint8_t> out, block;
  for (size_t off = 0; off < bytes; off += lz4::kMaxChunk) {
    block.clear();
    const int32_t n = int32_t(enc.encode_chunk(off, std::min(bytes, off + size_t(lz4::kMaxChunk)), block));
    const uint8_t* np = reinterpret_cast<const uint8_t*>(&n);
    out.insert(out.end(), np, np + 4);
    out.insert(out.end(), block.begin(), block.end());
  }
  return out;
}

// lz4cpp.cpp:100-110 driven by volume.cpp:205-218: chunks until `bytes` bytes are out
void decompressBody(std::istream& s, uint8_t* dst, size_t bytes, const std::string& what) {
  std::vector<uint8_t> block;
  size_t pos = 0;
  while (pos < bytes) {
    int32_t n = 0;
    s.read(reinterpret_cast<char*>(&n), 4);
    // a chunk holds at most 64 KiB of input, so at most LZ4's worst-case bound of that (the reference reads into a
    // buffer of exactly this size without checking, lz4cpp.cpp:92-104)
    constexpr int32_t kMaxBlock = lz4::kMaxChunk + lz4::kMaxChunk / 255 + 16;
    if (!s || n <= 0 || n > kMaxBlock) throw std::runtime_error("truncated or malformed LZ4 chunk in " + what);
    block.resize(size_t(n));
    s.read(reinterpret_cast<char*>(block.data()), n);
    if (!s) throw std::runtime_error("truncated LZ4 chunk in " + what);
    const long long got = lz4::decode_block(block.data(), size_t(n), dst, pos, bytes);
    if (got <= 0) throw std::runtime_error("malformed LZ4 chunk in " + what);
    pos += size_t(got);
  }
}

}  // namespace cvol

Volume::Volume(int sizeX, int sizeY, int sizeZ) {
  TORCH_CHECK(sizeX > 0 && sizeY > 0 && sizeZ > 0, "volume resolution must be positive");
  levels_.resize(1);
  levels_[0].gpu = torch::zeros({1, sizeX, sizeY, sizeZ}, on_gpu());
}

Volume::Volume(const std::string& filename) {
  std::ifstream s(filename, std::ifstream::binary);
  if (!s) throw std::runtime_error("cannot open " + filename);
  CvolHeader h;
  s.read(reinterpret_cast<char*>(&h), sizeof(h));
  if (!s || std::memcmp(h.magic, kMagic, 4) != 0) throw std::runtime_error("Illegal magic number");
  if (h.type < 0 || h.type > 2) throw std::runtime_error("unknown .cvol data type " + std::to_string(h.type));
  if (!h.size[0] || !h.size[1] || !h.size[2]) throw std::runtime_error("empty .cvol volume");
  if (h.size[0] > 0x7fffffffull || h.size[1] > 0x7fffffffull || h.size[2] > 0x7fffffffull ||
      h.size[0] * h.size[1] * h.size[2] > 0x7fffffffull)  // volume.cpp:192-193
    throw std::runtime_error("Dataset is too big, must fit into an 32-bit signed int");
  const size_t voxels = size_t(h.size[0] * h.size[1] * h.size[2]);
  const size_t bytes = voxels * kBytesPerType[h.type];
  for (int i = 0; i < 3; ++i) worldSize_[i] = h.voxelSize[i] * double(h.size[i]);

  // body -> host staging buffer (a torch tensor so that the upload is one .to())
  torch::Tensor raw = torch::empty({int64_t(bytes)}, at::TensorOptions().dtype(c10::ScalarType::Byte));
  uint8_t* dst = raw.data_ptr<uint8_t>();
  if (h.compressed) {
    cvol::decompressBody(s, dst, bytes, filename);
  } else {
    s.read(reinterpret_cast<char*>(dst), std::streamsize(bytes));
    if (size_t(s.gcount()) != bytes) throw std::runtime_error("truncated .cvol body in " + filename);
  }

  levels_.resize(1);
  levels_[0].gpu = torch::empty({1, int64_t(h.size[0]), int64_t(h.size[1]), int64_t(h.size[2])}, on_gpu());
  const torch::Tensor rawGpu = raw.to(c10::kCUDA);
  const DvrTensor v = view_of(levels_[0].gpu);
  raise_on_error(dvr_volume_from_raw(rawGpu.data_ptr(), h.type, &v, current_stream()));
}

void Volume::save(const std::string& filename, int compression) const {
  if (compression < 0 || compression > MAX_COMPRESSION) throw std::runtime_error("Illegal compression factor");
  const torch::Tensor& vol = dataGpu(0);
  const c10::cuda::CUDAGuard guard(vol.device());
  CvolHeader h{};
  std::memcpy(h.magic, kMagic, 4);
  for (int i = 0; i < 3; ++i) {
    h.size[i] = uint64_t(vol.size(i + 1));
    h.voxelSize[i] = worldSize_[i] / double(vol.size(i + 1));  // volume.cpp:108-110: (int) size -> double
  }
  h.type = kF32;
  h.compressed = compression > 0 ? 1 : 0;

  // X-fastest body on the device, then one download
  torch::Tensor rawGpu = torch::empty({vol.numel()}, vol.options());
  const DvrTensor v = view_of(vol);
  raise_on_error(dvr_volume_to_raw(&v, rawGpu.data_ptr<float>(), current_stream()));
  const torch::Tensor raw = rawGpu.cpu();
  const uint8_t* src = reinterpret_cast<const uint8_t*>(raw.data_ptr<float>());
  const size_t bytes = size_t(raw.numel()) * sizeof(float);

  std::ofstream s(filename, std::ofstream::binary);
  if (!s) throw std::runtime_error("cannot open " + filename + " for writing");
  s.write(reinterpret_cast<const char*>(&h), sizeof(h));
  if (h.compressed) {
    const std::vector<uint8_t> body = cvol::compressBody(src, bytes, compression);
    s.write(reinterpret_cast<const char*>(body.data()), std::streamsize(body.size()));
  } else {
    s.write(reinterpret_cast<const char*>(src), std::streamsize(bytes));
  }
  if (!s) throw std::runtime_error("error while writing " + filename);
}

std::shared_ptr<Volume> Volume::createImplicitDataset(int resolution, ImplicitEquation equation,
                                                      const std::unordered_map<std::string, float>& params) {
  if (int(equation) < 0 || equation >= ImplicitEquation::_NUM_IMPLICIT_EQUATION_) return {};  // volume_implicit.cpp:252
  TORCH_CHECK(resolution >= 2, "resolution must be at least 2");
  auto get = [&](const char* key, float def) {
    const auto it = params.find(key);
    return it == params.end() ? def : it->second;
  };
  const float p[4] = {get("fM", 6.0f), get("alpha", 0.25f), get("scale", 0.5f), get("a", 0.25f)};
  std::shared_ptr<Volume> vol(new Volume());
  vol->levels_.resize(1);
  vol->levels_[0].gpu = torch::empty({1, resolution, resolution, resolution}, on_gpu());
  const DvrTensor v = view_of(vol->levels_[0].gpu);
  raise_on_error(dvr_create_implicit_volume(int32_t(equation), p, &v, current_stream()));
  return vol;
}

std::shared_ptr<Volume> Volume::createFromBuffer(const float* buffer, const long long sizes[3],
                                                 const long long strides[3]) {
  TORCH_CHECK(buffer && sizes[0] > 0 && sizes[1] > 0 && sizes[2] > 0, "empty buffer");
  // strided host view (negative strides are not representable in a tensor: gather those on the host)
  torch::Tensor host;
  if (strides[0] >= 0 && strides[1] >= 0 && strides[2] >= 0) {
    host = torch::from_blob(const_cast<float*>(buffer), {1, sizes[0], sizes[1], sizes[2]},
                            {0, strides[0], strides[1], strides[2]}, at::TensorOptions().dtype(c10::ScalarType::Float));
  } else {
    host = torch::empty({1, sizes[0], sizes[1], sizes[2]}, at::TensorOptions().dtype(c10::ScalarType::Float));
    auto acc = host.accessor<float, 4>();
    for (long long x = 0; x < sizes[0]; ++x)
      for (long long y = 0; y < sizes[1]; ++y)
        for (long long z = 0; z < sizes[2]; ++z) acc[0][x][y][z] = buffer[x * strides[0] + y * strides[1] + z * strides[2]];
  }
  std::shared_ptr<Volume> vol(new Volume());
  vol->levels_.resize(1);
  vol->levels_[0].gpu = host.contiguous().to(on_gpu().device());
  return vol;
}

std::shared_ptr<Volume> Volume::createFromTensor(const torch::Tensor& data) {
  TORCH_CHECK(data.defined(), "data is not defined!");
  TORCH_CHECK(data.dtype() == c10::ScalarType::Float, "data must be of type c10::ScalarType::Float");
  torch::Tensor t = data.dim() == 3 ? data.unsqueeze(0) : data;
  TORCH_CHECK(t.dim() == 4 && t.size(0) == 1, "data must be of shape X*Y*Z or 1*X*Y*Z");
  std::shared_ptr<Volume> vol(new Volume());
  vol->levels_.resize(1);
  vol->levels_[0].gpu = t.device().is_cuda() ? t.detach() : t.detach().to(on_gpu().device());
  return vol;
}

void Volume::createMipmapLevel(int level) {
  TORCH_CHECK(level >= 0, "level has to be non-zero, but is ", level);
  if (hasLevel(level)) return;  // already available
  if (size_t(level) >= levels_.size()) levels_.resize(size_t(level) + 1);
  const torch::Tensor& src = levels_[0].gpu;
  const c10::cuda::CUDAGuard guard(src.device());
  auto dim = [&](int d) { return std::max<int64_t>(1, src.size(d) / (level + 1)); };  // volume.cpp:449-451
  torch::Tensor dst = torch::empty({1, dim(1), dim(2), dim(3)}, src.options());
  const DvrTensor a = view_of(src), b = view_of(dst);
  raise_on_error(dvr_downsample_volume(&a, &b, current_stream()));
  levels_[level].gpu = dst;
}

void Volume::deleteAllMipmapLevels() { levels_.resize(1); }

const torch::Tensor& Volume::dataGpu(int level) const {
  TORCH_CHECK(hasLevel(level), "Mipmap Level ", level, " was not created");
  return levels_[level].gpu;
}

const torch::Tensor& Volume::dataCpu(int level) const {
  TORCH_CHECK(hasLevel(level), "Mipmap Level ", level, " was not created");
  const Level& l = levels_[level];
  if (!l.cpu.defined()) l.cpu = l.gpu.cpu();
  return l.cpu;
}

int3 Volume::baseResolution() const {
  const torch::Tensor& t = levels_.at(0).gpu;
  return int3{int(t.size(1)), int(t.size(2)), int(t.size(3))};
}

}  // namespace renderer
