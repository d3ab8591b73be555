// Jubatus model file (".jubatus") container around the device table -- SURVEY.md §8(f) n2.
//
// Mirrors jubatus/server/framework/save_load.cpp:
//   header, 48 bytes (save_load.cpp:133-148):
//     [ 0.. 8) magic "jubatus\0"            [ 8..16) format_version = 1, big endian u64
//     [16..20) jubatus major  [20..24) minor  [24..28) maintenance   (big endian u32)
//     [28..32) crc32 (big endian u32)
//     [32..40) system data size  [40..48) user data size            (big endian u64)
//   crc32 covers header[0..28) + header[32..48) + system data + user data (save_load.cpp:84-92),
//   polynomial 0xEDB88320, initial/final xor 0xFFFFFFFF, chainable (common/crc32.cpp:25-52;
//   known answers in common/crc32_test.cpp:26-27: "" -> 0, "jubatus" -> 0x41918955).
//   system data = msgpack [version=1, timestamp, type, id, config]   (save_load.cpp:64-82)
//   user data   = msgpack [user_data_version, <driver pack>]          (save_load.cpp:118-131)
// load checks, in the reference's order and with its messages (save_load.cpp:170-283): magic,
// format version, jubatus version, crc32, system data version, server type, config (compared as
// JSON, not as text), user container arity, user data version.
//
// What is NOT the reference's: <driver pack>.  In the reference it is jubatus_core's msgpack
// serialisation of column_table, which is not available here; this build stores one raw msgpack
// string holding driver::nearest_neighbor::pack() (the "JUBANN1" blob of device_table::pack()
// followed by the converter's global-weight counts).  A file written by the
// reference therefore passes every container check here and then fails in unpack with a clear
// message (and the other way round).
#ifndef JUBANN_MODEL_FILE_HPP_
#define JUBANN_MODEL_FILE_HPP_

#include <stdint.h>
#include <string.h>

#include <stdio.h>

#include <ctime>
#include <istream>
#include <ostream>
#include <sstream>
#include <stdexcept>
#include <algorithm>
#include <string>
#include <vector>

#include "json_lite.hpp"

#ifndef JUBANN_JUBATUS_VERSION
#define JUBANN_JUBATUS_VERSION "0.9.2"   // wscript:7 of the reference this build stands in for
#endif

namespace jubatus_b200 {
namespace model_file {

class format_error : public std::runtime_error {
 public:
  explicit format_error(const std::string& m) : std::runtime_error(m) {}
};

// ---- crc32 (common/crc32.cpp:25-52) ------------------------------------------------------
struct crc32_table {
  uint32_t t[256];
  crc32_table() {
    for (uint32_t i = 0; i < 256; ++i) {
      uint32_t c = i;
      for (int j = 0; j < 8; ++j) c = (c >> 1) ^ ((c & 1u) ? 0xEDB88320u : 0u);
      t[i] = c;
    }
  }
};
inline uint32_t calc_crc32(const char* data, size_t size, uint32_t crc = 0) {
  static const crc32_table table;
  crc ^= 0xFFFFFFFFu;
  for (size_t i = 0; i < size; ++i) crc = (crc >> 8) ^ table.t[(static_cast<unsigned char>(data[i]) ^ crc) & 0xFFu];
  return crc ^ 0xFFFFFFFFu;
}

// ---- big endian fields (common/big_endian.hpp as used at save_load.cpp:135-148) ------------
inline void write_be32(uint32_t v, char* p) { for (int i = 0; i < 4; ++i) p[i] = static_cast<char>(v >> (24 - 8 * i)); }
inline void write_be64(uint64_t v, char* p) { for (int i = 0; i < 8; ++i) p[i] = static_cast<char>(v >> (56 - 8 * i)); }
inline uint32_t read_be32(const char* p) { uint32_t v = 0; for (int i = 0; i < 4; ++i) v = (v << 8) | static_cast<unsigned char>(p[i]); return v; }
inline uint64_t read_be64(const char* p) { uint64_t v = 0; for (int i = 0; i < 8; ++i) v = (v << 8) | static_cast<unsigned char>(p[i]); return v; }

// ---- the few msgpack forms the container uses ---------------------------------------------
// Strings are written as fixraw / raw16 / raw32 (the msgpack 0.5 family the reference links
// has no str8); the reader also accepts str8.  Unsigned integers take the shortest form.
struct mp_writer {
  std::string out;
  void array(uint32_t n) {
    if (n < 16) out.push_back(static_cast<char>(0x90 | n));
    else { out.push_back(static_cast<char>(0xdc)); out.push_back(static_cast<char>(n >> 8)); out.push_back(static_cast<char>(n)); }
  }
  void uint(uint64_t v) {
    char b[8];
    if (v < 128) out.push_back(static_cast<char>(v));
    else if (v < 256) { out.push_back(static_cast<char>(0xcc)); out.push_back(static_cast<char>(v)); }
    else if (v < 65536) { out.push_back(static_cast<char>(0xcd)); out.push_back(static_cast<char>(v >> 8)); out.push_back(static_cast<char>(v)); }
    else if (v < (1ull << 32)) { out.push_back(static_cast<char>(0xce)); write_be32(static_cast<uint32_t>(v), b); out.append(b, 4); }
    else { out.push_back(static_cast<char>(0xcf)); write_be64(v, b); out.append(b, 8); }
  }
  void raw(const std::string& s) {
    char b[4];
    if (s.size() < 32) out.push_back(static_cast<char>(0xa0 | s.size()));
    else if (s.size() < 65536) { out.push_back(static_cast<char>(0xda)); out.push_back(static_cast<char>(s.size() >> 8)); out.push_back(static_cast<char>(s.size())); }
    else { out.push_back(static_cast<char>(0xdb)); write_be32(static_cast<uint32_t>(s.size()), b); out.append(b, 4); }
    out.append(s);
  }
};

struct mp_reader {
  const char* p;
  const char* end;
  mp_reader(const char* b, size_t n) : p(b), end(b + n) {}
  unsigned char byte() { need(1); return static_cast<unsigned char>(*p++); }
  void need(size_t n) const { if (static_cast<size_t>(end - p) < n) throw format_error("truncated"); }
  uint32_t array() {
    const unsigned char t = byte();
    if ((t & 0xf0) == 0x90) return t & 0x0f;
    if (t == 0xdc) { need(2); uint32_t n = (static_cast<unsigned char>(p[0]) << 8) | static_cast<unsigned char>(p[1]); p += 2; return n; }
    if (t == 0xdd) { need(4); uint32_t n = read_be32(p); p += 4; return n; }
    throw format_error("not an array");
  }
  uint64_t uint() {
    const unsigned char t = byte();
    if (t < 0x80) return t;
    uint64_t v = 0;
    switch (t) {
      case 0xcc: case 0xd0: need(1); v = static_cast<unsigned char>(*p); p += 1; return v;
      case 0xcd: case 0xd1: need(2); v = (static_cast<unsigned char>(p[0]) << 8) | static_cast<unsigned char>(p[1]); p += 2; return v;
      case 0xce: case 0xd2: need(4); v = read_be32(p); p += 4; return v;
      case 0xcf: case 0xd3: need(8); v = read_be64(p); p += 8; return v;
      default: throw format_error("not an integer");
    }
  }
  std::string raw() {
    const unsigned char t = byte();
    size_t n;
    if ((t & 0xe0) == 0xa0) n = t & 0x1f;
    else if (t == 0xd9 || t == 0xc4) { need(1); n = static_cast<unsigned char>(*p); p += 1; }
    else if (t == 0xda || t == 0xc5) { need(2); n = (static_cast<unsigned char>(p[0]) << 8) | static_cast<unsigned char>(p[1]); p += 2; }
    else if (t == 0xdb || t == 0xc6) { need(4); n = read_be32(p); p += 4; }
    else throw format_error("not a string");
    need(n);
    std::string s(p, n);
    p += n;
    return s;
  }
};

// ---- container ------------------------------------------------------------------------------
struct system_data {          // save_load.cpp:64-82
  uint64_t version;
  uint64_t timestamp;
  std::string type, id, config;
  system_data() : version(0), timestamp(0) {}
};
static const uint64_t FORMAT_VERSION = 1;                  // save_load.cpp:45
static const uint64_t SYSTEM_DATA_CONTAINER_VERSION = 1;   // save_load.cpp:62

inline void jubatus_version(uint32_t v[3]) {
  int a = 0, b = 0, c = 0;
  sscanf(JUBANN_JUBATUS_VERSION, "%d.%d.%d", &a, &b, &c);
  v[0] = static_cast<uint32_t>(a); v[1] = static_cast<uint32_t>(b); v[2] = static_cast<uint32_t>(c);
}

inline uint32_t container_crc32(const char* header, const std::string& sys, const std::string& user) {   // save_load.cpp:84-92
  uint32_t crc = calc_crc32(header, 28);
  crc = calc_crc32(header + 32, 16, crc);
  crc = calc_crc32(sys.data(), sys.size(), crc);
  crc = calc_crc32(user.data(), user.size(), crc);
  return crc;
}

// save_server (save_load.cpp:113-166).  `driver_pack` is what the driver serialises.
inline void save(std::ostream& os, const std::string& type, const std::string& id, const std::string& config,
                 uint64_t user_data_version, const std::string& driver_pack, uint64_t timestamp = 0) {
  mp_writer sys;
  sys.array(5);
  sys.uint(SYSTEM_DATA_CONTAINER_VERSION);
  sys.uint(timestamp ? timestamp : static_cast<uint64_t>(std::time(NULL)));
  sys.raw(type); sys.raw(id); sys.raw(config);
  mp_writer user;
  user.array(2);
  user.uint(user_data_version);
  user.raw(driver_pack);
  char header[48];
  memset(header, 0, sizeof(header));
  memcpy(header, "jubatus", 8);
  write_be64(FORMAT_VERSION, header + 8);
  uint32_t v[3];
  jubatus_version(v);
  write_be32(v[0], header + 16); write_be32(v[1], header + 20); write_be32(v[2], header + 24);
  write_be64(sys.out.size(), header + 32);
  write_be64(user.out.size(), header + 40);
  write_be32(container_crc32(header, sys.out, user.out), header + 28);
  os.write(header, sizeof(header));
  os.write(sys.out.data(), static_cast<std::streamsize>(sys.out.size()));
  os.write(user.out.data(), static_cast<std::streamsize>(user.out.size()));
  if (!os) throw std::ios_base::failure("Failed to write model file.");
}

inline bool same_config(const std::string& a, const std::string& b) {   // compare_config, save_load.cpp:102-109
  return json::dump(json::parse(a)) == json::dump(json::parse(b));
}

inline void read_exact(std::istream& is, uint64_t n, std::string& out) {
  const uint64_t STEP = 16ull << 20;
  out.clear();
  while (out.size() < n) {
    const uint64_t want = std::min<uint64_t>(STEP, n - out.size());
    const size_t at = out.size();
    out.resize(at + want);
    is.read(&out[at], static_cast<std::streamsize>(want));
    if (static_cast<uint64_t>(is.gcount()) != want) throw std::ios_base::failure("cannot read input file");
  }
}

// load_server (save_load.cpp:168-283).  Returns the driver pack; `sys` receives the system data.
// expected_config is ignored when overwrite_config (load_file, server_base.cpp:207-209).
inline std::string load(std::istream& is, const std::string& expected_type, const std::string& expected_config,
                        bool overwrite_config, uint64_t expected_user_data_version, system_data* sys_out) {
  char header[48];
  is.read(header, 48);
  if (is.gcount() != 48 || memcmp(header, "jubatus", 8) != 0) throw format_error("invalid file format");
  const uint64_t fv = read_be64(header + 8);
  if (fv != FORMAT_VERSION) {
    std::ostringstream m; m << "invalid format version: " << fv << ", expected " << FORMAT_VERSION;
    throw format_error(m.str());
  }
  uint32_t v[3];
  jubatus_version(v);
  const uint32_t r0 = read_be32(header + 16), r1 = read_be32(header + 20), r2 = read_be32(header + 24);
  if (r0 != v[0] || r1 != v[1] || r2 != v[2]) {
    std::ostringstream m;
    m << "jubatus version mismatched: " << r0 << "." << r1 << "." << r2 << ", expected (current) version: " JUBANN_JUBATUS_VERSION;
    throw format_error(m.str());
  }
  const uint32_t crc_expected = read_be32(header + 28);
  const uint64_t sys_size = read_be64(header + 32), user_size = read_be64(header + 40);
  if (sys_size > (1ull << 32) || user_size > (1ull << 40)) throw format_error("invalid file format");
  // the sizes come from the file: memory is taken as the bytes actually arrive (16 MiB at a time), so a
  // header that claims a terabyte costs nothing before the stream runs dry
  std::string sys, user;
  read_exact(is, sys_size, sys);
  read_exact(is, user_size, user);
  const uint32_t crc_actual = container_crc32(header, sys, user);
  if (crc_actual != crc_expected) {
    std::ostringstream m; m << "invalid crc32 checksum: " << std::hex << crc_actual << ", expected " << crc_expected;
    throw format_error(m.str());
  }
  system_data sd;
  try {
    mp_reader r(sys.data(), sys.size());
    if (r.array() != 5) throw format_error("arity");
    sd.version = r.uint(); sd.timestamp = r.uint(); sd.type = r.raw(); sd.id = r.raw(); sd.config = r.raw();
  } catch (const format_error&) {
    throw format_error("system data is broken");
  }
  if (sys_out) *sys_out = sd;
  if (sd.version != SYSTEM_DATA_CONTAINER_VERSION) {
    std::ostringstream m; m << "invalid system data version: saved version: " << sd.version << ", expected " << SYSTEM_DATA_CONTAINER_VERSION;
    throw format_error(m.str());
  }
  if (sd.type != expected_type) throw format_error("server type mismatched: " + sd.type + ", expected " + expected_type);
  if (!overwrite_config) {
    bool same = false;
    try { same = same_config(sd.config, expected_config); } catch (const std::exception&) { same = false; }
    if (!same) throw format_error("server config mismatched: " + sd.config + ", expected " + expected_config);
  }
  try {
    mp_reader r(user.data(), user.size());
    if (r.array() != 2) throw format_error("invalid user container");
    const uint64_t uv = r.uint();
    if (uv != expected_user_data_version) {
      std::ostringstream m; m << "user data version mismatched: " << uv << ", expected " << expected_user_data_version;
      throw format_error(m.str());
    }
    return r.raw();
  } catch (const format_error& e) {
    const std::string w = e.what();
    if (w == "invalid user container" || w.compare(0, 28, "user data version mismatched") == 0) throw;
    throw format_error("user data is broken");
  }
}

}  // namespace model_file
}  // namespace jubatus_b200

#endif  // JUBANN_MODEL_FILE_HPP_
