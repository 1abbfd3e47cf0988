// PCD v0.7 reader / writer for LidarPoint clouds (see pcd_io.h).  Restates, from the published
// format, what the reference obtains from PCL 1.10 at PointCloudStorage.h:85-115 (save) and
// Slam.cxx:531 (pcl::io::loadPCDFile):
//  * header: FIELDS/SIZE/TYPE/COUNT of the registered fields (LidarPoint.h:68-77) in declaration
//    order, padding (`data[3]`) not written; WIDTH n, HEIGHT 1, identity VIEWPOINT;
//  * ascii: one point per line, fields separated by one blank, floating point with 8 significant
//    digits ("%.8g", also for the double `time`), integers in decimal, NaN as "nan";
//  * binary: the fields of each point packed back to back (28 B per point);
//  * binary_compressed: u32 compressed size, u32 raw size, LZF stream of the field-major (SoA)
//    rearrangement of the packed fields.
// The reader accepts any field order, `_` padding fields and unknown fields (skipped); a known
// field is taken only when its SIZE/TYPE match the registered ones (PCL: "Failed to find match
// for field"), missing fields keep the values of LidarPoint's default constructor.
#include "pcd_io.h"
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <sstream>

namespace lsb {
namespace pcd {

namespace {

struct Field { const char* name; int offset, size; char type; };
// POINT_CLOUD_REGISTER_POINT_STRUCT order (LidarPoint.h:68-77)
const Field kFields[] = {
  {"x", 0, 4, 'F'}, {"y", 4, 4, 'F'}, {"z", 8, 4, 'F'}, {"time", 16, 8, 'F'}, {"intensity", 24, 4, 'F'},
  {"laser_id", 28, 2, 'U'}, {"device_id", 30, 1, 'U'}, {"label", 31, 1, 'U'}};
const int kNbFields = 8;
const int kPacked = 28;   // bytes of the registered fields of one point

lsb_point DefaultPoint()
{
  lsb_point p;
  std::memset(&p, 0, sizeof(p));
  p.w = 1.f;
  return p;
}

void AppendReal(std::string& s, double v)
{
  if (std::isnan(v)) { s += "nan"; return; }
  char b[40];
  std::snprintf(b, sizeof(b), "%.8g", v);
  s += b;
}

}  // namespace

std::string Header(size_t n)
{
  std::ostringstream o;
  o << "# .PCD v0.7 - Point Cloud Data file format\nVERSION 0.7\nFIELDS";
  for (int f = 0; f < kNbFields; ++f) o << ' ' << kFields[f].name;
  o << "\nSIZE";
  for (int f = 0; f < kNbFields; ++f) o << ' ' << kFields[f].size;
  o << "\nTYPE";
  for (int f = 0; f < kNbFields; ++f) o << ' ' << kFields[f].type;
  o << "\nCOUNT";
  for (int f = 0; f < kNbFields; ++f) o << " 1";
  o << "\nWIDTH " << n << "\nHEIGHT 1\nVIEWPOINT 0 0 0 1 0 0 0\nPOINTS " << n << "\n";
  return o.str();
}

// ---- LZF ------------------------------------------------------------------------
// Stream grammar (liblzf): control byte c < 32: c + 1 literal bytes follow; otherwise a back
// reference of length (c >> 5) + 2 (if c >> 5 == 7 the next byte is added to the length) at
// distance ((c & 31) << 8 | next byte) + 1.
size_t LzfCompress(const unsigned char* in, size_t n, unsigned char* out, size_t cap)
{
  const int HLOG = 14;
  const size_t MAX_OFF = 1u << 13, MAX_REF = (1u << 8) + (1u << 3), MAX_LIT = 1u << 5;
  std::vector<uint32_t> table(1u << HLOG, 0xffffffffu);
  size_t ip = 0, op = 0, lit_start = 0;
  auto flush_literals = [&](size_t end) -> bool {
    while (lit_start < end) {
      const size_t run = std::min(MAX_LIT, end - lit_start);
      if (op + 1 + run > cap) return false;
      out[op++] = (unsigned char)(run - 1);
      std::memcpy(out + op, in + lit_start, run);
      op += run;
      lit_start += run;
    }
    return true;
  };
  while (ip + 2 < n) {
    const uint32_t v = (uint32_t)in[ip] << 16 | (uint32_t)in[ip + 1] << 8 | in[ip + 2];
    const uint32_t hs = ((v * 2654435761u) >> (32 - HLOG)) & ((1u << HLOG) - 1);
    const uint32_t ref = table[hs];
    table[hs] = (uint32_t)ip;
    if (ref != 0xffffffffu && ip - ref <= MAX_OFF && ip < 0xffffffffu && in[ref] == in[ip] && in[ref + 1] == in[ip + 1] &&
        in[ref + 2] == in[ip + 2]) {
      size_t len = 3;
      const size_t maxlen = std::min(MAX_REF, n - ip);
      while (len < maxlen && in[ref + len] == in[ip + len]) ++len;
      if (!flush_literals(ip)) return 0;
      const size_t off = ip - ref - 1, l = len - 2;
      if (op + 3 > cap) return 0;
      if (l < 7) out[op++] = (unsigned char)((l << 5) | (off >> 8));
      else { out[op++] = (unsigned char)((7u << 5) | (off >> 8)); out[op++] = (unsigned char)(l - 7); }
      out[op++] = (unsigned char)(off & 0xff);
      ip += len;
      lit_start = ip;
    } else {
      ++ip;
    }
  }
  if (!flush_literals(n)) return 0;
  return op;
}

size_t LzfDecompress(const unsigned char* in, size_t n, unsigned char* out, size_t cap)
{
  size_t ip = 0, op = 0;
  while (ip < n) {
    const unsigned c = in[ip++];
    if (c < 32) {
      const size_t run = c + 1;
      if (ip + run > n || op + run > cap) return 0;
      std::memcpy(out + op, in + ip, run);
      ip += run; op += run;
    } else {
      size_t len = c >> 5;
      if (len == 7) { if (ip >= n) return 0; len += in[ip++]; }
      if (ip >= n) return 0;
      const size_t dist = ((size_t)(c & 31) << 8 | in[ip++]) + 1;
      len += 2;
      if (dist > op || op + len > cap) return 0;
      for (size_t i = 0; i < len; ++i, ++op) out[op] = out[op - dist];   // overlapping copies are part of the format
    }
  }
  return op;
}

// ---- writer -----------------------------------------------------------------------
int Write(const std::string& path, const lsb_point* pts, size_t n, int format)
{
  if (n == 0) return -3;                                   // PointCloudStorage.h:91-92
  if (format < 0 || format > 2) return -4;                 // PointCloudStorage.h:111-113
  std::string body = Header(n);
  if (format == 0) {
    body += "DATA ascii\n";
    body.reserve(body.size() + n * 80);
    for (size_t i = 0; i < n; ++i) {
      const lsb_point& p = pts[i];
      AppendReal(body, p.x); body += ' ';
      AppendReal(body, p.y); body += ' ';
      AppendReal(body, p.z); body += ' ';
      AppendReal(body, p.time); body += ' ';
      AppendReal(body, p.intensity); body += ' ';
      body += std::to_string((unsigned)p.laser_id); body += ' ';
      body += std::to_string((unsigned)p.device_id); body += ' ';
      body += std::to_string((unsigned)p.label); body += '\n';
    }
  } else if (format == 1) {
    body += "DATA binary\n";
    const size_t at = body.size();
    body.resize(at + n * kPacked);
    char* o = &body[at];
    for (size_t i = 0; i < n; ++i)
      for (int f = 0; f < kNbFields; ++f) {
        std::memcpy(o, reinterpret_cast<const char*>(&pts[i]) + kFields[f].offset, kFields[f].size);
        o += kFields[f].size;
      }
  } else {
    body += "DATA binary_compressed\n";
    const size_t raw = n * kPacked;
    if (raw > 0xffffffffu) return -1;                      // the two sizes of the wire format are 32-bit
    std::vector<unsigned char> soa(raw);
    unsigned char* o = soa.data();
    for (int f = 0; f < kNbFields; ++f)
      for (size_t i = 0; i < n; ++i, o += kFields[f].size)
        std::memcpy(o, reinterpret_cast<const char*>(&pts[i]) + kFields[f].offset, kFields[f].size);
    std::vector<unsigned char> comp(raw + raw / 16 + 64);
    const size_t c = LzfCompress(soa.data(), raw, comp.data(), comp.size());
    if (c == 0) return -1;
    const uint32_t sizes[2] = {(uint32_t)c, (uint32_t)raw};
    body.append(reinterpret_cast<const char*>(sizes), 8);
    body.append(reinterpret_cast<const char*>(comp.data()), c);
  }
  FILE* fp = std::fopen(path.c_str(), "wb");
  if (!fp) return -1;
  const size_t w = std::fwrite(body.data(), 1, body.size(), fp);
  const int rc = std::fclose(fp);
  return (w == body.size() && rc == 0) ? 0 : -1;
}

// ---- reader -----------------------------------------------------------------------
namespace {

struct FileField { std::string name; int size = 0; char type = 0; int count = 1; int target = -1; };

std::vector<std::string> Tokens(const std::string& line)
{
  std::vector<std::string> t;
  std::istringstream is(line);
  for (std::string w; is >> w;) t.push_back(w);
  return t;
}

int Fail(std::string* err, const std::string& why) { if (err) *err = why; return -1; }

void StoreAscii(const Field& f, const std::string& tok, lsb_point& p)
{
  char* dst = reinterpret_cast<char*>(&p) + f.offset;
  if (f.type == 'F') {
    const bool nan = tok == "nan" || tok == "-nan" || tok == "NaN";
    if (f.size == 4) { const float v = nan ? std::nanf("") : std::strtof(tok.c_str(), nullptr); std::memcpy(dst, &v, 4); }
    else { const double v = nan ? std::nan("") : std::strtod(tok.c_str(), nullptr); std::memcpy(dst, &v, 8); }
  } else {
    const unsigned long v = std::strtoul(tok.c_str(), nullptr, 10);
    if (f.size == 1) { const uint8_t b = (uint8_t)v; std::memcpy(dst, &b, 1); }
    else { const uint16_t b = (uint16_t)v; std::memcpy(dst, &b, 2); }
  }
}

}  // namespace

int Read(const std::string& path, std::vector<lsb_point>& out, std::string* err)
{
  out.clear();
  FILE* fp = std::fopen(path.c_str(), "rb");
  if (!fp) return Fail(err, "cannot open " + path);
  std::string buf;
  {
    char chunk[1 << 16];
    for (size_t r; (r = std::fread(chunk, 1, sizeof(chunk), fp)) > 0;) buf.append(chunk, r);
    std::fclose(fp);
  }
  // header: text lines up to and including DATA
  std::vector<FileField> ff;
  size_t pos = 0, width = 0, height = 1, npoints = 0;
  bool have_points = false, have_width = false;
  std::string data;
  while (pos < buf.size() && data.empty()) {
    size_t eol = buf.find('\n', pos);
    if (eol == std::string::npos) eol = buf.size();
    std::string line = buf.substr(pos, eol - pos);
    pos = std::min(eol + 1, buf.size());
    if (!line.empty() && line.back() == '\r') line.pop_back();
    const std::vector<std::string> t = Tokens(line);
    if (t.empty() || t[0][0] == '#') continue;
    const std::string& key = t[0];
    auto need_fields = [&]() { return ff.size() == t.size() - 1; };
    if (key == "VERSION" || key == "VIEWPOINT") continue;
    else if (key == "FIELDS" || key == "COLUMNS") { ff.assign(t.size() - 1, FileField()); for (size_t i = 1; i < t.size(); ++i) ff[i - 1].name = t[i]; }
    else if (key == "SIZE") { if (!need_fields()) return Fail(err, "SIZE does not match FIELDS"); for (size_t i = 1; i < t.size(); ++i) ff[i - 1].size = std::atoi(t[i].c_str()); }
    else if (key == "TYPE") { if (!need_fields()) return Fail(err, "TYPE does not match FIELDS"); for (size_t i = 1; i < t.size(); ++i) ff[i - 1].type = t[i][0]; }
    else if (key == "COUNT") { if (!need_fields()) return Fail(err, "COUNT does not match FIELDS"); for (size_t i = 1; i < t.size(); ++i) ff[i - 1].count = std::atoi(t[i].c_str()); }
    else if (key == "WIDTH" && t.size() > 1) { width = std::strtoull(t[1].c_str(), nullptr, 10); have_width = true; }
    else if (key == "HEIGHT" && t.size() > 1) height = std::strtoull(t[1].c_str(), nullptr, 10);
    else if (key == "POINTS" && t.size() > 1) { npoints = std::strtoull(t[1].c_str(), nullptr, 10); have_points = true; }
    else if (key == "DATA" && t.size() > 1) data = t[1];
    else return Fail(err, "unknown header entry '" + key + "'");
  }
  if (data.empty()) return Fail(err, "no DATA entry");
  if (ff.empty()) return Fail(err, "no FIELDS entry");
  if (!have_points) { if (!have_width) return Fail(err, "neither POINTS nor WIDTH"); npoints = width * height; }
  size_t stride = 0;
  for (FileField& f : ff) {
    if (f.size <= 0 || f.count < 0) return Fail(err, "bad SIZE/COUNT of field '" + f.name + "'");
    if (f.count == 0) f.count = 1;
    for (int k = 0; k < kNbFields; ++k)
      if (f.name == kFields[k].name && f.count == 1 && f.size == kFields[k].size && f.type == kFields[k].type) f.target = k;
    stride += (size_t)f.size * f.count;
  }
  if (npoints == 0) return 0;
  // a header cannot promise more points than the payload can hold (checked before anything is allocated)
  const size_t payload = buf.size() - pos;
  if (stride == 0 || npoints > payload) return Fail(err, "POINTS exceeds what the file can hold");
  if (data == "binary" && npoints > payload / stride) return Fail(err, "binary payload shorter than POINTS x point size");
  if (data == "binary_compressed" && npoints > (size_t)0xffffffffu / stride) return Fail(err, "raw size differs from POINTS x point size");
  out.assign(npoints, DefaultPoint());

  if (data == "ascii") {
    size_t i = 0;
    while (i < npoints && pos < buf.size()) {
      size_t eol = buf.find('\n', pos);
      if (eol == std::string::npos) eol = buf.size();
      const std::vector<std::string> t = Tokens(buf.substr(pos, eol - pos));
      pos = eol + 1;
      if (t.empty()) continue;
      size_t col = 0;
      for (const FileField& f : ff) {
        if (col + f.count > t.size()) return Fail(err, "short ascii line");
        if (f.target >= 0) StoreAscii(kFields[f.target], t[col], out[i]);
        col += f.count;
      }
      ++i;
    }
    if (i != npoints) return Fail(err, "fewer ascii lines than POINTS");
    return 0;
  }
  if (data == "binary") {
    if (buf.size() - pos < npoints * stride) return Fail(err, "binary payload shorter than POINTS x point size");
    const char* src = buf.data() + pos;
    for (size_t i = 0; i < npoints; ++i)
      for (const FileField& f : ff) {
        if (f.target >= 0) std::memcpy(reinterpret_cast<char*>(&out[i]) + kFields[f.target].offset, src, f.size);
        src += (size_t)f.size * f.count;
      }
    return 0;
  }
  if (data == "binary_compressed") {
    if (buf.size() - pos < 8) return Fail(err, "missing compressed sizes");
    uint32_t sizes[2];
    std::memcpy(sizes, buf.data() + pos, 8);
    pos += 8;
    if (buf.size() - pos < sizes[0]) return Fail(err, "compressed payload truncated");
    if ((size_t)sizes[1] != npoints * stride) return Fail(err, "raw size differs from POINTS x point size");
    std::vector<unsigned char> raw(sizes[1]);
    if (LzfDecompress(reinterpret_cast<const unsigned char*>(buf.data()) + pos, sizes[0], raw.data(), raw.size()) != raw.size())
      return Fail(err, "LZF stream malformed");
    const unsigned char* src = raw.data();
    for (const FileField& f : ff) {                          // field-major planes
      const size_t fs = (size_t)f.size * f.count;
      if (f.target >= 0)
        for (size_t i = 0; i < npoints; ++i)
          std::memcpy(reinterpret_cast<char*>(&out[i]) + kFields[f.target].offset, src + i * fs, f.size);
      src += fs * npoints;
    }
    return 0;
  }
  return Fail(err, "unknown DATA type '" + data + "'");
}

}  // namespace pcd
}  // namespace lsb
