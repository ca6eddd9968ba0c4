// The LZ4 *block* format, written from its public specification (token = literal-length nibble | match-length
// nibble, 255-extension bytes, literals, 16-bit little-endian match offset, minimum match 4, last five bytes of a
// block literal, last match starting at least twelve bytes before the end of the block) -- the only part of LZ4
// that a compressed .cvol body needs. The reference links the LZ4 library (third-party/lz4, absent from this repo)
// through its LZ4Compressor / LZ4Decompressor wrappers (third-party/lz4/lib/lz4cpp.cpp:49-111), which frame the body
// as  [int32 compressedBytes][LZ4 block]  per chunk of at most 64 KiB of input and let matches reach back into the
// previous chunks (streaming mode with a contiguous buffer). Both directions here keep that contract:
//  * decode accepts match offsets that point before the current block, down to the start of the output buffer;
//  * encode searches the previous 64 KiB of the whole buffer, so later chunks compress against earlier ones.
// The encoder's match finder is its own (hash table + hash chains), so its bytes differ from liblz4's; every block
// it emits is a valid LZ4 block that liblz4 decodes (tests/test_oracle_volume.py: files written here are read back by
// the reference-built Volume and vice versa).
#pragma once
#include <cstddef>
#include <cstdint>
#include <vector>

namespace renderer {
namespace lz4 {

constexpr int kMaxChunk = 64 * 1024;  // LZ4Compressor::MAX_CHUNK_SIZE (lz4cpp.hpp:12)

// Decodes one block src[0, srcLen) to out[pos, ...), never writing at or beyond out[cap]. Matches may start anywhere
// in out[0, current). Returns the number of bytes produced, or -1 for a malformed block.
long long decode_block(const uint8_t* src, size_t srcLen, uint8_t* out, size_t pos, size_t cap);

// Streaming encoder over ONE contiguous input buffer, chunk after chunk in increasing order.
class Encoder {
 public:
  // level 0..2: one probe per position with acceleration (fast); 3..9: hash chains, 2^(level-1) candidates
  Encoder(const uint8_t* data, size_t size, int level);
  // Appends the block for data[begin, end) (end - begin <= kMaxChunk) to `dst`; returns its size in bytes.
  size_t encode_chunk(size_t begin, size_t end, std::vector<uint8_t>& dst);

 private:
  void insert(size_t p);
  const uint8_t* data_;
  size_t size_;
  int attempts_;
  bool fast_;
  std::vector<long long> head_;    // hash -> last position + 1 (0 = empty)
  std::vector<uint16_t> chain_;    // position & 0xffff -> distance to the previous position with the same hash
};

}  // namespace lz4
}  // namespace renderer
