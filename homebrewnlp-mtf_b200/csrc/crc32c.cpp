// CRC-32C (Castagnoli) over host memory: the checksum of TFRecord frames (src/inputs.py reads them through
// tf.data.TFRecordDataset) and of TensorBundle checkpoints (tf.train.Saver, src/run/run.py:158-175).  Checkpoints
// checksum every tensor -- gigabytes per save -- so this is native: the SSE4.2 crc32 instruction where the host has
// it (three independent streams per block to cover the instruction's latency, recombined with a carry-less shift
// table), slicing-by-8 tables elsewhere.
#include <cstddef>
#include <cstdint>
#include <cstring>

#include "../../include/obst_b200.h"

namespace {

constexpr uint32_t POLY = 0x82F63B78u;  // reflected 0x1EDC6F41

struct Tables {
  uint32_t t[8][256];
  Tables() {
    for (uint32_t i = 0; i < 256; ++i) {
      uint32_t c = i;
      for (int k = 0; k < 8; ++k) c = (c & 1) ? (c >> 1) ^ POLY : c >> 1;
      t[0][i] = c;
    }
    for (uint32_t i = 0; i < 256; ++i)
      for (int s = 1; s < 8; ++s) t[s][i] = (t[s - 1][i] >> 8) ^ t[0][t[s - 1][i] & 0xFF];
  }
};
const Tables& tables() {
  static const Tables T;
  return T;
}

uint32_t crc_sw(uint32_t c, const uint8_t* p, size_t n) {
  const Tables& T = tables();
  while (n && (reinterpret_cast<uintptr_t>(p) & 7)) {
    c = T.t[0][(c ^ *p++) & 0xFF] ^ (c >> 8);
    --n;
  }
  while (n >= 8) {
    uint64_t w;
    std::memcpy(&w, p, 8);
    w ^= c;
    c = T.t[7][w & 0xFF] ^ T.t[6][(w >> 8) & 0xFF] ^ T.t[5][(w >> 16) & 0xFF] ^ T.t[4][(w >> 24) & 0xFF] ^
        T.t[3][(w >> 32) & 0xFF] ^ T.t[2][(w >> 40) & 0xFF] ^ T.t[1][(w >> 48) & 0xFF] ^ T.t[0][w >> 56];
    p += 8;
    n -= 8;
  }
  while (n--) c = T.t[0][(c ^ *p++) & 0xFF] ^ (c >> 8);
  return c;
}

#if defined(__x86_64__)
// x^(8 * len) mod P applied to a crc: advances `c` over `len` zero bytes.  Square-and-multiply over GF(2) 32x32
// matrices would be the general tool; with one fixed block length a 4 x 256 table of the operator does it in four
// look-ups.
struct Shift {
  uint32_t t[4][256];
  explicit Shift(size_t len) {
    // operator as its action on the 32 unit vectors, built by running the software crc over `len` zero bytes
    uint32_t col[32];
    for (int b = 0; b < 32; ++b) {
      uint32_t c = 1u << b;
      for (size_t i = 0; i < len; ++i) c = tables().t[0][c & 0xFF] ^ (c >> 8);
      col[b] = c;
    }
    for (int k = 0; k < 4; ++k)
      for (uint32_t v = 0; v < 256; ++v) {
        uint32_t r = 0;
        for (int b = 0; b < 8; ++b)
          if (v & (1u << b)) r ^= col[8 * k + b];
        t[k][v] = r;
      }
  }
  uint32_t operator()(uint32_t c) const {
    return t[0][c & 0xFF] ^ t[1][(c >> 8) & 0xFF] ^ t[2][(c >> 16) & 0xFF] ^ t[3][c >> 24];
  }
};

constexpr size_t LANE = 4096;  // bytes per stream of a 3-stream block

__attribute__((target("sse4.2"))) uint32_t crc_hw(uint32_t c, const uint8_t* p, size_t n) {
  static const Shift shift1(LANE), shift2(2 * LANE);
  while (n && (reinterpret_cast<uintptr_t>(p) & 7)) {
    c = __builtin_ia32_crc32qi(c, *p++);
    --n;
  }
  uint64_t c0 = c;
  while (n >= 3 * LANE) {
    uint64_t c1 = 0, c2 = 0;
    const uint64_t* a = reinterpret_cast<const uint64_t*>(p);
    const uint64_t* b = reinterpret_cast<const uint64_t*>(p + LANE);
    const uint64_t* d = reinterpret_cast<const uint64_t*>(p + 2 * LANE);
    for (size_t i = 0; i < LANE / 8; ++i) {
      c0 = __builtin_ia32_crc32di(c0, a[i]);
      c1 = __builtin_ia32_crc32di(c1, b[i]);
      c2 = __builtin_ia32_crc32di(c2, d[i]);
    }
    // crc(A|B|C) = shift_{2L}(crc A) ^ shift_L(crc B) ^ crc C  (the register form without the final inversion)
    c0 = shift2((uint32_t)c0) ^ shift1((uint32_t)c1) ^ (uint32_t)c2;
    p += 3 * LANE;
    n -= 3 * LANE;
  }
  while (n >= 8) {
    uint64_t w;
    std::memcpy(&w, p, 8);
    c0 = __builtin_ia32_crc32di(c0, w);
    p += 8;
    n -= 8;
  }
  c = (uint32_t)c0;
  while (n--) c = __builtin_ia32_crc32qi(c, *p++);
  return c;
}
#endif

}  // namespace

extern "C" uint32_t obst_crc32c(const void* host_data, int64_t nbytes, uint32_t crc) {
  if (!host_data || nbytes <= 0) return crc;
  uint32_t c = ~crc;
  const uint8_t* p = static_cast<const uint8_t*>(host_data);
#if defined(__x86_64__)
  static const bool hw = __builtin_cpu_supports("sse4.2");
  if (hw) return ~crc_hw(c, p, (size_t)nbytes);
#endif
  return ~crc_sw(c, p, (size_t)nbytes);
}

// test hook: the table path regardless of what the host supports
extern "C" uint32_t obst_crc32c_portable(const void* host_data, int64_t nbytes, uint32_t crc) {
  if (!host_data || nbytes <= 0) return crc;
  return ~crc_sw(~crc, static_cast<const uint8_t*>(host_data), (size_t)nbytes);
}
