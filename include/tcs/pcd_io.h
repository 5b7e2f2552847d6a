// include/tcs/pcd_io.h - PCD v0.7 reader / writer (ascii, binary, binary_compressed) for the headless tools.
//
// Replaces what the reference gets from pcl::io: load<PointT>() for the input cloud (src/io.cpp:32-85),
// pcl::io::loadPCDFile<PointXYZL> for the seeds (src/random_walker_segmentation_app.cpp:189) and
// pcl::io::savePCDFile for segmentation.pcd / cluster%i.pcd (:274-291; PCL's default there is ASCII).  All of the
// reference's data/*.pcd are `DATA binary_compressed`: two uint32 (compressed, uncompressed size) and an LZF stream
// whose decompressed bytes hold the fields as struct-of-arrays.  Header-only, no PCL needed.
#pragma once

#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <limits>
#include <sstream>
#include <stdexcept>
#include <string>
#include <vector>

#include "pcl_compat.h"

namespace tcs {
namespace pcd {

struct Field {
  std::string name;
  int size = 4;
  char type = 'F';
  int count = 1;
};

// A PCD file as columns of doubles would lose uint32 / float bit patterns; keep raw little-endian bytes per field.
struct Table {
  std::vector<Field> fields;
  std::vector<std::vector<unsigned char>> data;  // per field: n * count * size bytes
  size_t n = 0;
  uint32_t width = 0, height = 1;
  int find(const std::string& name) const {
    for (size_t i = 0; i < fields.size(); ++i)
      if (fields[i].name == name) return (int)i;
    return -1;
  }
};

// one LZF block (Marc Lehmann's liblzf format, what PCL's lzfDecompress reads)
inline bool lzf_decompress(const unsigned char* src, size_t n, unsigned char* out, size_t out_len) {
  size_t ip = 0, op = 0;
  while (ip < n) {
    unsigned ctrl = src[ip++];
    if (ctrl < 32) {  // literal run of ctrl + 1 bytes
      ++ctrl;
      if (ip + ctrl > n || op + ctrl > out_len) return false;
      std::memcpy(out + op, src + ip, ctrl);
      ip += ctrl;
      op += ctrl;
    } else {  // back reference
      size_t len = ctrl >> 5;
      if (len == 7) {
        if (ip >= n) return false;
        len += src[ip++];
      }
      if (ip >= n) return false;
      const size_t dist = ((size_t)(ctrl & 0x1F) << 8) + src[ip++] + 1;
      len += 2;
      if (dist > op || op + len > out_len) return false;
      for (size_t k = 0; k < len; ++k, ++op) out[op] = out[op - dist];  // may overlap: byte by byte
    }
  }
  return op == out_len;
}

// a valid LZF stream made of literal runs only - enough for a writer
inline std::vector<unsigned char> lzf_store(const std::vector<unsigned char>& raw) {
  std::vector<unsigned char> out;
  out.reserve(raw.size() + raw.size() / 32 + 1);
  for (size_t i = 0; i < raw.size(); i += 32) {
    const size_t c = std::min<size_t>(32, raw.size() - i);
    out.push_back((unsigned char)(c - 1));
    out.insert(out.end(), raw.begin() + i, raw.begin() + i + c);
  }
  return out;
}

namespace detail {
inline std::vector<std::string> split(const std::string& s) {
  std::istringstream is(s);
  std::vector<std::string> v;
  std::string t;
  while (is >> t) v.push_back(t);
  return v;
}
// parse one ascii token into the field's binary representation
inline void put_ascii(const Field& f, const std::string& tok, unsigned char* dst) {
  if (f.type == 'F') {
    double d = (tok == "nan" || tok == "NaN" || tok == "-nan") ? std::numeric_limits<double>::quiet_NaN() : std::strtod(tok.c_str(), nullptr);
    if (f.size == 4) { float v = (float)d; std::memcpy(dst, &v, 4); }
    else std::memcpy(dst, &d, 8);
  } else if (f.type == 'U') {
    unsigned long long u = std::strtoull(tok.c_str(), nullptr, 10);
    std::memcpy(dst, &u, f.size);  // little endian
  } else {
    long long i = std::strtoll(tok.c_str(), nullptr, 10);
    std::memcpy(dst, &i, f.size);
  }
}
}  // namespace detail

// Throws std::runtime_error on malformed files.
inline Table read(const std::string& path) {
  std::ifstream in(path, std::ios::binary);
  if (!in) throw std::runtime_error("cannot open " + path);
  std::string blob((std::istreambuf_iterator<char>(in)), std::istreambuf_iterator<char>());
  Table t;
  std::vector<std::string> names, sizes, types, counts;
  std::string mode;
  size_t pos = 0, points = 0;
  bool have_points = false;
  while (pos < blob.size()) {
    size_t end = blob.find('\n', pos);
    if (end == std::string::npos) end = blob.size();
    std::string line = blob.substr(pos, end - pos);
    pos = end + 1;
    std::vector<std::string> tok = detail::split(line);
    if (tok.empty() || tok[0][0] == '#') continue;
    const std::string key = tok[0];
    tok.erase(tok.begin());
    if (key == "FIELDS") names = tok;
    else if (key == "SIZE") sizes = tok;
    else if (key == "TYPE") types = tok;
    else if (key == "COUNT") counts = tok;
    else if (key == "WIDTH" && !tok.empty()) t.width = (uint32_t)std::strtoul(tok[0].c_str(), nullptr, 10);
    else if (key == "HEIGHT" && !tok.empty()) t.height = (uint32_t)std::strtoul(tok[0].c_str(), nullptr, 10);
    else if (key == "POINTS" && !tok.empty()) { points = std::strtoull(tok[0].c_str(), nullptr, 10); have_points = true; }
    else if (key == "DATA") { mode = tok.empty() ? "" : tok[0]; break; }
  }
  if (names.empty() || sizes.size() != names.size() || types.size() != names.size())
    throw std::runtime_error(path + ": malformed PCD header");
  t.n = have_points ? points : (size_t)t.width * t.height;
  size_t rec = 0;
  for (size_t i = 0; i < names.size(); ++i) {
    Field f;
    f.name = names[i];
    f.size = std::atoi(sizes[i].c_str());
    f.type = types[i][0];
    f.count = i < counts.size() ? std::atoi(counts[i].c_str()) : 1;
    if (f.size != 1 && f.size != 2 && f.size != 4 && f.size != 8) throw std::runtime_error(path + ": bad field size");
    if (f.count <= 0 || f.count > (1 << 20)) throw std::runtime_error(path + ": bad field count");
    if (f.type != 'F' && f.type != 'U' && f.type != 'I') throw std::runtime_error(path + ": bad field type");
    rec += (size_t)f.size * (size_t)f.count;
    t.fields.push_back(f);
  }
  // The header is untrusted: POINTS / WIDTH x HEIGHT decide how much is allocated and copied, so they are bounded by
  // what the file can hold before anything is sized from them (every point takes at least one byte of data in ascii
  // and binary; an LZF stream expands by at most ~255x, checked again against the stream's own size field below).
  const size_t data_bytes = blob.size() >= pos ? blob.size() - pos : 0;
  if (rec == 0 || t.n > std::numeric_limits<size_t>::max() / rec) throw std::runtime_error(path + ": POINTS x record size overflows");
  if (mode == "binary" && rec * t.n > data_bytes) throw std::runtime_error(path + ": binary data ends early");
  if (mode == "ascii" && t.n > data_bytes) throw std::runtime_error(path + ": ascii data ends early");
  if (mode == "binary_compressed" && (data_bytes < 8 || rec * t.n / 256 > data_bytes)) throw std::runtime_error(path + ": compressed data ends early");
  if (mode != "ascii" && mode != "binary" && mode != "binary_compressed") throw std::runtime_error(path + ": unknown DATA mode '" + mode + "'");
  t.data.resize(t.fields.size());
  for (size_t i = 0; i < t.fields.size(); ++i) t.data[i].assign(t.n * t.fields[i].size * t.fields[i].count, 0);
  const unsigned char* p = (const unsigned char*)blob.data() + pos;
  const size_t left = blob.size() >= pos ? blob.size() - pos : 0;
  if (mode == "ascii") {
    std::istringstream is(blob.substr(pos));
    std::string tok;
    for (size_t r = 0; r < t.n; ++r)
      for (size_t i = 0; i < t.fields.size(); ++i) {
        const Field& f = t.fields[i];
        for (int c = 0; c < f.count; ++c) {
          if (!(is >> tok)) throw std::runtime_error(path + ": ascii data ends early");
          detail::put_ascii(f, tok, &t.data[i][(r * f.count + c) * f.size]);
        }
      }
  } else if (mode == "binary") {
    if (left < rec * t.n) throw std::runtime_error(path + ": binary data ends early");
    for (size_t r = 0; r < t.n; ++r) {
      const unsigned char* q = p + r * rec;
      for (size_t i = 0; i < t.fields.size(); ++i) {
        const size_t b = (size_t)t.fields[i].size * t.fields[i].count;
        std::memcpy(&t.data[i][r * b], q, b);
        q += b;
      }
    }
  } else if (mode == "binary_compressed") {
    if (left < 8) throw std::runtime_error(path + ": compressed data ends early");
    uint32_t csize, usize;
    std::memcpy(&csize, p, 4);
    std::memcpy(&usize, p + 4, 4);
    if (left < 8 + (size_t)csize || usize != rec * t.n) throw std::runtime_error(path + ": compressed sizes do not match the header");
    std::vector<unsigned char> raw(usize);
    if (!lzf_decompress(p + 8, csize, raw.data(), usize)) throw std::runtime_error(path + ": corrupt LZF stream");
    size_t off = 0;
    for (size_t i = 0; i < t.fields.size(); ++i) {  // struct of arrays
      std::memcpy(t.data[i].data(), raw.data() + off, t.data[i].size());
      off += t.data[i].size();
    }
  } else {
    throw std::runtime_error(path + ": unknown DATA mode '" + mode + "'");
  }
  return t;
}

// mode: "ascii" (what pcl::io::savePCDFile writes by default), "binary" or "binary_compressed"
inline void write(const std::string& path, const Table& t, const std::string& mode = "ascii") {
  std::ofstream out(path, std::ios::binary);
  if (!out) throw std::runtime_error("cannot write " + path);
  std::ostringstream h;
  h << "# .PCD v0.7 - Point Cloud Data file format\nVERSION 0.7\nFIELDS";
  for (const Field& f : t.fields) h << ' ' << f.name;
  h << "\nSIZE";
  for (const Field& f : t.fields) h << ' ' << f.size;
  h << "\nTYPE";
  for (const Field& f : t.fields) h << ' ' << f.type;
  h << "\nCOUNT";
  for (const Field& f : t.fields) h << ' ' << f.count;
  const uint32_t width = t.width && (size_t)t.width * t.height == t.n ? t.width : (uint32_t)t.n;
  const uint32_t height = t.width && (size_t)t.width * t.height == t.n ? t.height : 1u;
  h << "\nWIDTH " << width << "\nHEIGHT " << height << "\nVIEWPOINT 0 0 0 1 0 0 0\nPOINTS " << t.n << "\nDATA " << mode << "\n";
  out << h.str();
  if (mode == "ascii") {
    std::string line;
    char buf[64];
    for (size_t r = 0; r < t.n; ++r) {
      line.clear();
      for (size_t i = 0; i < t.fields.size(); ++i) {
        const Field& f = t.fields[i];
        for (int c = 0; c < f.count; ++c) {
          const unsigned char* src = &t.data[i][(r * f.count + c) * f.size];
          if (f.type == 'F') {
            double d;
            if (f.size == 4) { float v; std::memcpy(&v, src, 4); d = v; } else std::memcpy(&d, src, 8);
            if (std::isnan(d)) std::snprintf(buf, sizeof buf, "nan");
            else std::snprintf(buf, sizeof buf, f.size == 4 ? "%.9g" : "%.17g", d);
          } else if (f.type == 'U') {
            unsigned long long u = 0;
            std::memcpy(&u, src, f.size);
            std::snprintf(buf, sizeof buf, "%llu", u);
          } else {
            long long v = 0;
            std::memcpy(&v, src, f.size);
            if (f.size < 8) {  // sign-extend
              const int sh = 64 - 8 * f.size;
              v = (long long)((unsigned long long)v << sh) >> sh;
            }
            std::snprintf(buf, sizeof buf, "%lld", v);
          }
          if (!line.empty()) line += ' ';
          line += buf;
        }
      }
      line += '\n';
      out << line;
    }
  } else if (mode == "binary") {
    std::vector<unsigned char> rec;
    for (size_t r = 0; r < t.n; ++r) {
      rec.clear();
      for (size_t i = 0; i < t.fields.size(); ++i) {
        const size_t b = (size_t)t.fields[i].size * t.fields[i].count;
        rec.insert(rec.end(), t.data[i].begin() + r * b, t.data[i].begin() + (r + 1) * b);
      }
      out.write((const char*)rec.data(), (std::streamsize)rec.size());
    }
  } else if (mode == "binary_compressed") {
    std::vector<unsigned char> raw;
    for (size_t i = 0; i < t.fields.size(); ++i) raw.insert(raw.end(), t.data[i].begin(), t.data[i].end());
    const std::vector<unsigned char> comp = lzf_store(raw);
    const uint32_t cs = (uint32_t)comp.size(), us = (uint32_t)raw.size();
    out.write((const char*)&cs, 4);
    out.write((const char*)&us, 4);
    out.write((const char*)comp.data(), (std::streamsize)comp.size());
  } else {
    throw std::runtime_error("unknown PCD DATA mode '" + mode + "'");
  }
}

namespace detail {
inline float get_f32(const Table& t, int f, size_t r) {
  const Field& fd = t.fields[f];
  const unsigned char* src = &t.data[f][r * fd.count * fd.size];
  if (fd.type == 'F') {
    if (fd.size == 4) { float v; std::memcpy(&v, src, 4); return v; }
    double d; std::memcpy(&d, src, 8); return (float)d;
  }
  unsigned long long u = 0;
  std::memcpy(&u, src, fd.size);
  return (float)u;
}
inline uint32_t get_bits32(const Table& t, int f, size_t r) {  // the field's low 32 bits as stored (rgb floats keep their pattern)
  const Field& fd = t.fields[f];
  uint32_t u = 0;
  std::memcpy(&u, &t.data[f][r * fd.count * fd.size], std::min(4, fd.size));
  return u;
}
inline void add_field(Table& t, const char* name, char type, const void* data, size_t stride, size_t n) {
  Field f;
  f.name = name; f.size = 4; f.type = type; f.count = 1;
  t.fields.push_back(f);
  std::vector<unsigned char> col(4 * n);
  for (size_t i = 0; i < n; ++i) std::memcpy(&col[4 * i], (const unsigned char*)data + i * stride, 4);
  t.data.push_back(col);
}
}  // namespace detail

// Loads x, y, z and, when present, rgb / rgba into a coloured cloud (points without colour stay opaque black, like
// pcl::copyPointCloud from PointXYZ).  Returns false when the file has no x/y/z.  *has_color mirrors hasColor()
// (src/io.cpp:13-30).  Ground-truth labels (data/*.pcd carry a `label` field) go to *labels when asked for.
template <typename PointT>
bool load_cloud(const std::string& path, pcl::PointCloud<PointT>& cloud, bool* has_color = nullptr,
                std::vector<uint32_t>* labels = nullptr) {
  const Table t = read(path);
  const int fx = t.find("x"), fy = t.find("y"), fz = t.find("z");
  if (fx < 0 || fy < 0 || fz < 0) return false;
  int fc = t.find("rgba");
  if (fc < 0) fc = t.find("rgb");
  const int fl = t.find("label");
  cloud.points.resize(t.n);
  cloud.width = t.width ? t.width : (uint32_t)t.n;
  cloud.height = t.width ? t.height : 1;
  cloud.is_dense = true;
  if (labels) labels->assign(t.n, 0u);
  for (size_t i = 0; i < t.n; ++i) {
    PointT& p = cloud.points[i];
    p.x = detail::get_f32(t, fx, i);
    p.y = detail::get_f32(t, fy, i);
    p.z = detail::get_f32(t, fz, i);
    if (!(std::isfinite(p.x) && std::isfinite(p.y) && std::isfinite(p.z))) cloud.is_dense = false;
    if (fc >= 0) p.rgba = detail::get_bits32(t, fc, i);
    if (labels && fl >= 0) (*labels)[i] = detail::get_bits32(t, fl, i);
  }
  if (has_color) *has_color = fc >= 0;
  return true;
}

// pcl::io::loadPCDFile<pcl::PointXYZL>: x, y, z, label (missing label -> 0)
inline bool load_labeled(const std::string& path, pcl::PointCloud<pcl::PointXYZL>& cloud) {
  const Table t = read(path);
  const int fx = t.find("x"), fy = t.find("y"), fz = t.find("z"), fl = t.find("label");
  if (fx < 0 || fy < 0 || fz < 0) return false;
  cloud.points.resize(t.n);
  cloud.width = t.width ? t.width : (uint32_t)t.n;
  cloud.height = t.width ? t.height : 1;
  for (size_t i = 0; i < t.n; ++i) {
    pcl::PointXYZL& p = cloud.points[i];
    p.x = detail::get_f32(t, fx, i);
    p.y = detail::get_f32(t, fy, i);
    p.z = detail::get_f32(t, fz, i);
    p.label = fl >= 0 ? detail::get_bits32(t, fl, i) : 0u;
  }
  return true;
}

// pcl::io::savePCDFile(name, PointCloud<PointXYZL>): FIELDS x y z label
inline void save_labeled(const std::string& path, const pcl::PointCloud<pcl::PointXYZL>& cloud, const std::string& mode = "ascii") {
  Table t;
  t.n = cloud.size();
  t.width = cloud.width; t.height = cloud.height;
  const pcl::PointXYZL* p = cloud.points.data();
  detail::add_field(t, "x", 'F', &p->x, sizeof *p, t.n);
  detail::add_field(t, "y", 'F', &p->y, sizeof *p, t.n);
  detail::add_field(t, "z", 'F', &p->z, sizeof *p, t.n);
  detail::add_field(t, "label", 'U', &p->label, sizeof *p, t.n);
  write(path, t, mode);
}

// pcl::io::savePCDFile(name, PointCloud<PointXYZRGB[A]>): FIELDS x y z rgba
template <typename PointT>
void save_colored(const std::string& path, const pcl::PointCloud<PointT>& cloud, const std::string& mode = "ascii") {
  Table t;
  t.n = cloud.size();
  t.width = cloud.width; t.height = cloud.height;
  const PointT* p = cloud.points.data();
  detail::add_field(t, "x", 'F', &p->x, sizeof *p, t.n);
  detail::add_field(t, "y", 'F', &p->y, sizeof *p, t.n);
  detail::add_field(t, "z", 'F', &p->z, sizeof *p, t.n);
  detail::add_field(t, "rgba", 'U', &p->rgba, sizeof *p, t.n);
  write(path, t, mode);
}

}  // namespace pcd
}  // namespace tcs
