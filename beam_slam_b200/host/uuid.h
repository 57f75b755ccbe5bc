// UUIDs and their hash, as the fuse graph containers see them.
//
// fuse_core::UUID = boost::uuids::uuid (dependencies/fuse/fuse_core/include/fuse_core/uuid.h:51); variables get
// deterministic SHA-1 name UUIDs (uuid.h:85-88,120-123; src/uuid.cpp:68-102), constraints random v4 ones
// (src/constraint.cpp:46-51). HashGraph keys its unordered_maps with boost::hash<UUID> (uuid.h:56,
// fuse_graphs/include/fuse_graphs/hash_graph.h:367-370), so the hash below fixes createProblem's iteration order.
#pragma once
#include <array>
#include <cstdint>
#include <cstring>
#include <random>
#include <string>

namespace bsfuse {

struct UUID {
  std::array<uint8_t, 16> data{};
  bool operator==(const UUID& o) const { return data == o.data; }
  bool operator!=(const UUID& o) const { return data != o.data; }
  bool operator<(const UUID& o) const { return data < o.data; }
};

// boost::uuids::hash_value (boost/uuid/uuid.hpp, Boost <= 1.85): byte-wise hash_combine
struct UuidHash {
  std::size_t operator()(const UUID& u) const noexcept {
    std::size_t seed = 0;
    for (uint8_t b : u.data) seed ^= static_cast<std::size_t>(b) + 0x9e3779b9 + (seed << 6) + (seed >> 2);
    return seed;
  }
};

namespace uuid {

class Sha1 {
 public:
  Sha1() { h_[0] = 0x67452301u; h_[1] = 0xEFCDAB89u; h_[2] = 0x98BADCFEu; h_[3] = 0x10325476u; h_[4] = 0xC3D2E1F0u; }
  void update(const void* p, size_t n) {
    const uint8_t* d = static_cast<const uint8_t*>(p);
    for (size_t i = 0; i < n; ++i) {
      buf_[len_++] = d[i]; ++total_;
      if (len_ == 64) { block(); len_ = 0; }
    }
  }
  void digest(uint32_t out[5]) {
    const uint64_t bits = total_ * 8;
    uint8_t pad = 0x80; update(&pad, 1);
    pad = 0; while (len_ != 56) update(&pad, 1);
    uint8_t lenb[8]; for (int i = 0; i < 8; ++i) lenb[i] = static_cast<uint8_t>(bits >> (56 - 8 * i));
    update(lenb, 8);
    for (int i = 0; i < 5; ++i) out[i] = h_[i];
  }
 private:
  static uint32_t rol(uint32_t v, int s) { return (v << s) | (v >> (32 - s)); }
  void block() {
    uint32_t w[80];
    for (int i = 0; i < 16; ++i)
      w[i] = (uint32_t(buf_[4 * i]) << 24) | (uint32_t(buf_[4 * i + 1]) << 16) | (uint32_t(buf_[4 * i + 2]) << 8) | buf_[4 * i + 3];
    for (int i = 16; i < 80; ++i) w[i] = rol(w[i - 3] ^ w[i - 8] ^ w[i - 14] ^ w[i - 16], 1);
    uint32_t a = h_[0], b = h_[1], c = h_[2], d = h_[3], e = h_[4];
    for (int i = 0; i < 80; ++i) {
      uint32_t f, k;
      if (i < 20) { f = (b & c) | (~b & d); k = 0x5A827999u; }
      else if (i < 40) { f = b ^ c ^ d; k = 0x6ED9EBA1u; }
      else if (i < 60) { f = (b & c) | (b & d) | (c & d); k = 0x8F1BBCDCu; }
      else { f = b ^ c ^ d; k = 0xCA62C1D6u; }
      const uint32_t t = rol(a, 5) + f + e + k + w[i];
      e = d; d = c; c = rol(b, 30); b = a; a = t;
    }
    h_[0] += a; h_[1] += b; h_[2] += c; h_[3] += d; h_[4] += e;
  }
  uint32_t h_[5]; uint8_t buf_[64]; size_t len_ = 0; uint64_t total_ = 0;
};

// boost::uuids::name_generator(ns)(data, n): RFC 4122 version 5
inline UUID name_uuid(const UUID& ns, const void* data, size_t n) {
  Sha1 sha; sha.update(ns.data.data(), 16); sha.update(data, n);
  uint32_t dg[5]; sha.digest(dg);
  UUID u;
  for (int i = 0; i < 4; ++i) {
    u.data[4 * i + 0] = static_cast<uint8_t>(dg[i] >> 24); u.data[4 * i + 1] = static_cast<uint8_t>(dg[i] >> 16);
    u.data[4 * i + 2] = static_cast<uint8_t>(dg[i] >> 8); u.data[4 * i + 3] = static_cast<uint8_t>(dg[i]);
  }
  u.data[8] = static_cast<uint8_t>((u.data[8] & 0xBF) | 0x80);
  u.data[6] = static_cast<uint8_t>((u.data[6] & 0x5F) | 0x50);
  return u;
}
inline UUID nil() { return UUID{}; }
// fuse_core::uuid::generate(const std::string&) (uuid.h:108-111)
inline UUID generate(const std::string& s) { return name_uuid(nil(), s.data(), s.size()); }
// generate(namespace_string, data, n) (uuid.h:120-123)
inline UUID generate(const std::string& ns, const void* data, size_t n) { return name_uuid(generate(ns), data, n); }
// generate(namespace_string, stamp, device_id) (src/uuid.cpp:82-97): {uint32 sec, uint32 nsec, 16-byte id}
inline UUID generate(const std::string& ns, uint32_t sec, uint32_t nsec, const UUID& device) {
  uint8_t buf[24];
  std::memcpy(buf, &sec, 4); std::memcpy(buf + 4, &nsec, 4); std::memcpy(buf + 8, device.data.data(), 16);
  return generate(ns, buf, sizeof(buf));
}
// generate(namespace_string, uint64 id) (src/uuid.cpp:99-102)
inline UUID generate(const std::string& ns, uint64_t id) { return generate(ns, &id, sizeof(id)); }
// generate(namespace_string, std::string) (uuid.h:133-148)
inline UUID generate(const std::string& ns, const std::string& s) { return generate(ns, s.data(), s.size()); }

// fuse_core::uuid::generate() is a random v4 from boost::uuids::random_generator (src/uuid.cpp:54-66) and is not
// reproducible in the reference; here the generator is seedable so a synthetic graph has one fixed order.
class RandomGenerator {
 public:
  explicit RandomGenerator(uint64_t seed) : rng_(seed) {}
  UUID operator()() {
    UUID u; const uint64_t a = rng_(), b = rng_();
    std::memcpy(u.data.data(), &a, 8); std::memcpy(u.data.data() + 8, &b, 8);
    u.data[8] = static_cast<uint8_t>((u.data[8] & 0xBF) | 0x80);
    u.data[6] = static_cast<uint8_t>((u.data[6] & 0x4F) | 0x40);
    return u;
  }
 private:
  std::mt19937_64 rng_;
};

inline std::string to_string(const UUID& u) {
  static const char* hex = "0123456789abcdef";
  std::string s;
  for (int i = 0; i < 16; ++i) {
    s.push_back(hex[u.data[i] >> 4]); s.push_back(hex[u.data[i] & 15]);
    if (i == 3 || i == 5 || i == 7 || i == 9) s.push_back('-');
  }
  return s;
}

}  // namespace uuid
}  // namespace bsfuse
