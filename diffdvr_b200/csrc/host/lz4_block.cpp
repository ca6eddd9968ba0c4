// See lz4_block.h.
#include "lz4_block.h"

#include <cstring>

namespace renderer {
namespace lz4 {

namespace {
constexpr int kMinMatch = 4;
constexpr int kLastLiterals = 5;   // the last five bytes of a block are literals
constexpr int kMatchGuard = 12;    // no match starts within the last twelve bytes
constexpr size_t kWindow = 65535;  // largest offset a 16-bit field can hold
constexpr int kHashBits = 16;

inline uint32_t load32(const uint8_t* p) {
  uint32_t v;
  std::memcpy(&v, p, 4);
  return v;
}
inline uint32_t hash_at(const uint8_t* p) { return (load32(p) * 2654435761u) >> (32 - kHashBits); }

// a length >= 15 continues in bytes of 255 closed by one byte < 255
void put_length(std::vector<uint8_t>& dst, size_t rest) {
  while (rest >= 255) {
    dst.push_back(255);
    rest -= 255;
  }
  dst.push_back(uint8_t(rest));
}

void put_sequence(std::vector<uint8_t>& dst, const uint8_t* lit, size_t nlit, size_t offset, size_t matchLen) {
  const size_t ml = matchLen ? matchLen - kMinMatch : 0;
  dst.push_back(uint8_t((nlit >= 15 ? 15 : nlit) << 4 | (ml >= 15 ? 15 : ml)));
  if (nlit >= 15) put_length(dst, nlit - 15);
  dst.insert(dst.end(), lit, lit + nlit);
  if (!matchLen) return;  // closing sequence of the block: literals only
  dst.push_back(uint8_t(offset & 0xff));
  dst.push_back(uint8_t(offset >> 8));
  if (ml >= 15) put_length(dst, ml - 15);
}
}  // namespace

long long decode_block(const uint8_t* src, size_t srcLen, uint8_t* out, size_t pos, size_t cap) {
  size_t ip = 0, op = pos;
  auto more_length = [&](size_t& len) {
    uint8_t b;
    do {
      if (ip >= srcLen) return false;
      b = src[ip++];
      len += b;
    } while (b == 255);
    return true;
  };
  while (ip < srcLen) {
    const uint8_t token = src[ip++];
    size_t nlit = token >> 4;
    if (nlit == 15 && !more_length(nlit)) return -1;
    if (nlit > srcLen - ip || nlit > cap - op) return -1;
    std::memcpy(out + op, src + ip, nlit);
    ip += nlit;
    op += nlit;
    if (ip == srcLen) break;  // the closing sequence has no match part
    if (srcLen - ip < 2) return -1;
    const size_t offset = size_t(src[ip]) | size_t(src[ip + 1]) << 8;
    ip += 2;
    size_t len = token & 15;
    if (len == 15 && !more_length(len)) return -1;
    len += kMinMatch;
    if (offset == 0 || offset > op || len > cap - op) return -1;
    const uint8_t* from = out + op - offset;
    if (offset >= len) {
      std::memcpy(out + op, from, len);
    } else {
      for (size_t i = 0; i < len; ++i) out[op + i] = from[i];  // overlapping: the match repeats itself
    }
    op += len;
  }
  return (long long)(op - pos);
}

Encoder::Encoder(const uint8_t* data, size_t size, int level)
    : data_(data), size_(size), attempts_(level >= 3 ? 1 << (level - 1) : 1), fast_(level < 3),
      head_(size_t(1) << kHashBits, 0), chain_(fast_ ? 0 : 65536, 0) {}

void Encoder::insert(size_t p) {
  long long& h = head_[hash_at(data_ + p)];
  if (!fast_) {
    const size_t dist = h ? p - size_t(h - 1) : 0;
    chain_[p & 0xffff] = dist && dist <= kWindow ? uint16_t(dist) : 0;
  }
  h = (long long)p + 1;
}

size_t Encoder::encode_chunk(size_t begin, size_t end, std::vector<uint8_t>& dst) {
  const size_t start = dst.size();
  size_t anchor = begin, ip = begin;
  if (end - begin > kMatchGuard) {
    const size_t lastStart = end - kMatchGuard;  // a match may start at ip <= lastStart
    const size_t matchEnd = end - kLastLiterals;
    size_t misses = 0;
    while (ip <= lastStart) {
      // best candidate among the previous positions with the same hash
      size_t bestLen = 0, bestPos = 0;
      long long cand = head_[hash_at(data_ + ip)] - 1;
      for (int tries = 0; cand >= 0 && tries < attempts_; ++tries) {
        const size_t c = size_t(cand);
        if (ip - c > kWindow) break;
        if (load32(data_ + c) == load32(data_ + ip)) {
          size_t len = kMinMatch;
          while (ip + len < matchEnd && data_[c + len] == data_[ip + len]) ++len;
          if (len > bestLen) {
            bestLen = len;
            bestPos = c;
          }
          if (ip + len == matchEnd) break;  // cannot get longer (runs of equal bytes would walk the whole chain)
        }
        if (fast_) break;
        const uint16_t d = chain_[c & 0xffff];
        if (!d) break;
        cand -= d;
      }
      insert(ip);
      if (bestLen < size_t(kMinMatch)) {
        ip += fast_ ? 1 + (misses++ >> 6) : 1;  // acceleration: stride grows over incompressible data
        continue;
      }
      misses = 0;
      const size_t next = ip + bestLen;
      // the match may also extend backwards over pending literals (positions skipped by the acceleration)
      size_t from = ip;
      while (from > anchor && bestPos > 0 && data_[from - 1] == data_[bestPos - 1]) {
        --from;
        --bestPos;
      }
      put_sequence(dst, data_ + anchor, from - anchor, from - bestPos, next - from);
      // every position of the match stays findable: later chunks often repeat the middle of an earlier one
      for (size_t p = ip + 1; p < next && p + 4 <= size_; ++p) insert(p);
      ip = anchor = next;
    }
  }
  put_sequence(dst, data_ + anchor, end - anchor, 0, 0);
  return dst.size() - start;
}

}  // namespace lz4
}  // namespace renderer
