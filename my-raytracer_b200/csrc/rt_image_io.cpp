// rt_image_io.cpp — PPM / PNG codecs (see rt_image_io.h).  RFC 1950/1951 inflate, RFC 2083 PNG.
#include "rt_image_io.h"
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>

namespace rtb {
namespace {

// ---------------------------------------------------------------- inflate ----------------------------------
struct Bits {
  const uint8_t *p;
  size_t n, pos = 0;
  uint32_t buf = 0;
  int cnt = 0;
  bool bad = false;
  uint32_t get(int k) {
    while (cnt < k) {
      if (pos >= n) { bad = true; return 0; }
      buf |= (uint32_t)p[pos++] << cnt;
      cnt += 8;
    }
    const uint32_t v = buf & ((k == 32) ? 0xffffffffu : ((1u << k) - 1u));
    buf >>= k;
    cnt -= k;
    return v;
  }
  void align() { buf = 0; cnt = 0; }
};

struct Huff {
  uint16_t count[16];
  uint16_t symbol[288];
  bool build(const uint8_t *len, int n) {
    std::memset(count, 0, sizeof(count));
    for (int i = 0; i < n; ++i) ++count[len[i]];
    count[0] = 0;
    int left = 1;
    for (int l = 1; l < 16; ++l) {
      left <<= 1;
      left -= count[l];
      if (left < 0) return false; // over-subscribed
    }
    uint16_t offs[16];
    offs[1] = 0;
    for (int l = 1; l < 15; ++l) offs[l + 1] = offs[l] + count[l];
    for (int i = 0; i < n; ++i)
      if (len[i]) symbol[offs[len[i]]++] = (uint16_t)i;
    return true;
  }
  int decode(Bits &b) const {
    int code = 0, first = 0, index = 0;
    for (int l = 1; l < 16; ++l) {
      code |= (int)b.get(1);
      if (b.bad) return -1;
      const int c = count[l];
      if (code - c < first) return symbol[index + (code - first)];
      index += c;
      first += c;
      first <<= 1;
      code <<= 1;
    }
    return -1;
  }
};

const uint16_t kLenBase[29] = {3, 4, 5, 6, 7, 8, 9, 10, 11, 13, 15, 17, 19, 23, 27, 31, 35, 43, 51, 59, 67, 83, 99, 115, 131, 163, 195, 227, 258};
const uint16_t kLenExtra[29] = {0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 2, 2, 2, 2, 3, 3, 3, 3, 4, 4, 4, 4, 5, 5, 5, 5, 0};
const uint16_t kDistBase[30] = {1, 2, 3, 4, 5, 7, 9, 13, 17, 25, 33, 49, 65, 97, 129, 193, 257, 385, 513, 769, 1025, 1537, 2049, 3073, 4097, 6145, 8193, 12289, 16385, 24577};
const uint16_t kDistExtra[30] = {0, 0, 0, 0, 1, 1, 2, 2, 3, 3, 4, 4, 5, 5, 6, 6, 7, 7, 8, 8, 9, 9, 10, 10, 11, 11, 12, 12, 13, 13};

bool inflate_block(Bits &b, const Huff &lit, const Huff &dist, std::vector<uint8_t> &out) {
  for (;;) {
    const int s = lit.decode(b);
    if (s < 0) return false;
    if (s < 256) { out.push_back((uint8_t)s); continue; }
    if (s == 256) return true;
    const int li = s - 257;
    if (li >= 29) return false;
    const int len = kLenBase[li] + (int)b.get(kLenExtra[li]);
    const int ds = dist.decode(b);
    if (ds < 0 || ds >= 30) return false;
    const size_t d = kDistBase[ds] + (size_t)b.get(kDistExtra[ds]);
    if (b.bad || d > out.size()) return false;
    const size_t from = out.size() - d;
    for (int k = 0; k < len; ++k) out.push_back(out[from + k]);
  }
}

uint32_t crc_table[256];
bool crc_ready = false;
uint32_t crc32(const uint8_t *p, size_t n, uint32_t c = 0) {
  if (!crc_ready) {
    for (uint32_t i = 0; i < 256; ++i) {
      uint32_t x = i;
      for (int k = 0; k < 8; ++k) x = (x & 1) ? 0xedb88320u ^ (x >> 1) : x >> 1;
      crc_table[i] = x;
    }
    crc_ready = true;
  }
  c = ~c;
  for (size_t i = 0; i < n; ++i) c = crc_table[(c ^ p[i]) & 0xff] ^ (c >> 8);
  return ~c;
}
uint32_t adler32(const uint8_t *p, size_t n) {
  uint32_t a = 1, b = 0;
  for (size_t i = 0; i < n; ++i) {
    a = (a + p[i]) % 65521u;
    b = (b + a) % 65521u;
  }
  return (b << 16) | a;
}
uint32_t be32(const uint8_t *p) { return ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | p[3]; }
void put32(std::vector<uint8_t> &v, uint32_t x) {
  v.push_back((uint8_t)(x >> 24)); v.push_back((uint8_t)(x >> 16)); v.push_back((uint8_t)(x >> 8)); v.push_back((uint8_t)x);
}
void chunk(std::vector<uint8_t> &out, const char *type, const std::vector<uint8_t> &body) {
  put32(out, (uint32_t)body.size());
  std::vector<uint8_t> t(type, type + 4);
  t.insert(t.end(), body.begin(), body.end());
  out.insert(out.end(), t.begin(), t.end());
  put32(out, crc32(t.data(), t.size()));
}
int paeth(int a, int b, int c) {
  const int p = a + b - c, pa = std::abs(p - a), pb = std::abs(p - b), pc = std::abs(p - c);
  return (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c);
}

} // namespace

bool inflate_zlib(const uint8_t *src, size_t n, std::vector<uint8_t> &out, std::string &err) {
  if (n < 6 || (src[0] & 0x0f) != 8 || ((src[0] << 8) | src[1]) % 31 != 0 || (src[1] & 0x20)) {
    err = "zlib: bad header";
    return false;
  }
  Bits b{src + 2, n - 2};
  int last = 0;
  while (!last) {
    last = (int)b.get(1);
    const int type = (int)b.get(2);
    if (b.bad) { err = "zlib: truncated"; return false; }
    if (type == 0) {
      b.align();
      if (b.pos + 4 > b.n) { err = "zlib: truncated stored block"; return false; }
      const unsigned len = b.p[b.pos] | (b.p[b.pos + 1] << 8), nlen = b.p[b.pos + 2] | (b.p[b.pos + 3] << 8);
      b.pos += 4;
      if ((len ^ 0xffffu) != nlen || b.pos + len > b.n) { err = "zlib: bad stored block"; return false; }
      out.insert(out.end(), b.p + b.pos, b.p + b.pos + len);
      b.pos += len;
    } else if (type == 1 || type == 2) {
      Huff lit, dist;
      uint8_t len[320];
      if (type == 1) {
        for (int i = 0; i < 144; ++i) len[i] = 8;
        for (int i = 144; i < 256; ++i) len[i] = 9;
        for (int i = 256; i < 280; ++i) len[i] = 7;
        for (int i = 280; i < 288; ++i) len[i] = 8;
        lit.build(len, 288);
        for (int i = 0; i < 30; ++i) len[i] = 5;
        dist.build(len, 30);
      } else {
        const int nlen = (int)b.get(5) + 257, ndist = (int)b.get(5) + 1, ncode = (int)b.get(4) + 4;
        static const uint8_t order[19] = {16, 17, 18, 0, 8, 7, 9, 6, 10, 5, 11, 4, 12, 3, 13, 2, 14, 1, 15};
        uint8_t cl[19] = {0};
        for (int i = 0; i < ncode; ++i) cl[order[i]] = (uint8_t)b.get(3);
        Huff code;
        if (b.bad || nlen > 286 || ndist > 30 || !code.build(cl, 19)) { err = "zlib: bad dynamic header"; return false; }
        int i = 0;
        while (i < nlen + ndist) {
          const int s = code.decode(b);
          if (s < 0) { err = "zlib: bad code lengths"; return false; }
          if (s < 16) { len[i++] = (uint8_t)s; continue; }
          int rep, val = 0;
          if (s == 16) {
            if (i == 0) { err = "zlib: repeat without a previous length"; return false; }
            val = len[i - 1];
            rep = 3 + (int)b.get(2);
          } else if (s == 17) rep = 3 + (int)b.get(3);
          else rep = 11 + (int)b.get(7);
          if (i + rep > nlen + ndist) { err = "zlib: too many code lengths"; return false; }
          while (rep--) len[i++] = (uint8_t)val;
        }
        if (!lit.build(len, nlen) || !dist.build(len + nlen, ndist)) { err = "zlib: bad code"; return false; }
      }
      if (!inflate_block(b, lit, dist, out)) { err = "zlib: corrupt block"; return false; }
    } else {
      err = "zlib: reserved block type";
      return false;
    }
  }
  return true;
}

bool decode_png(const uint8_t *d, size_t n, Image8 &img, std::string &err) {
  static const uint8_t sig[8] = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a};
  if (n < 8 || std::memcmp(d, sig, 8) != 0) { err = "png: bad signature"; return false; }
  size_t pos = 8;
  int w = 0, h = 0, depth = 0, ctype = 0, interlace = 0;
  std::vector<uint8_t> idat, plte, trns;
  bool seen_ihdr = false;
  while (pos + 12 <= n) {
    const uint32_t len = be32(d + pos);
    if (pos + 12 + (size_t)len > n) { err = "png: truncated chunk"; return false; }
    const uint8_t *type = d + pos + 4, *body = d + pos + 8;
    if (crc32(type, 4 + (size_t)len) != be32(body + len)) { err = "png: chunk CRC mismatch"; return false; }
    if (!std::memcmp(type, "IHDR", 4) && len == 13) {
      w = (int)be32(body); h = (int)be32(body + 4); depth = body[8]; ctype = body[9]; interlace = body[12];
      seen_ihdr = true;
    } else if (!std::memcmp(type, "PLTE", 4)) plte.assign(body, body + len);
    else if (!std::memcmp(type, "tRNS", 4)) trns.assign(body, body + len);
    else if (!std::memcmp(type, "IDAT", 4)) idat.insert(idat.end(), body, body + len);
    else if (!std::memcmp(type, "IEND", 4)) break;
    pos += 12 + (size_t)len;
  }
  if (!seen_ihdr || w <= 0 || h <= 0) { err = "png: missing IHDR"; return false; }
  if (depth != 8 || interlace != 0 || !(ctype == 0 || ctype == 2 || ctype == 3 || ctype == 4 || ctype == 6)) {
    err = "png: only 8-bit non-interlaced grey/RGB/palette/RGBA images are supported";
    return false;
  }
  const int ch = ctype == 0 ? 1 : ctype == 2 ? 3 : ctype == 3 ? 1 : ctype == 4 ? 2 : 4;
  std::vector<uint8_t> raw;
  raw.reserve(((size_t)w * ch + 1) * h);
  if (!inflate_zlib(idat.data(), idat.size(), raw, err)) return false;
  const size_t stride = (size_t)w * ch;
  if (raw.size() < (stride + 1) * (size_t)h) { err = "png: image data too short"; return false; }
  std::vector<uint8_t> cur(stride), prev(stride, 0);
  img.w = w; img.h = h;
  img.rgba.assign((size_t)w * h * 4, 255);
  for (int y = 0; y < h; ++y) {
    const uint8_t *row = raw.data() + (stride + 1) * (size_t)y;
    const int f = row[0];
    for (size_t i = 0; i < stride; ++i) {
      const int a = i >= (size_t)ch ? cur[i - ch] : 0, b = prev[i], c = i >= (size_t)ch ? prev[i - ch] : 0;
      int v = row[1 + i];
      switch (f) {
        case 0: break;
        case 1: v += a; break;
        case 2: v += b; break;
        case 3: v += (a + b) / 2; break;
        case 4: v += paeth(a, b, c); break;
        default: err = "png: bad filter type"; return false;
      }
      cur[i] = (uint8_t)v;
    }
    uint8_t *o = img.rgba.data() + (size_t)y * w * 4;
    for (int x = 0; x < w; ++x) {
      const uint8_t *s = cur.data() + (size_t)x * ch;
      if (ctype == 0) { o[0] = o[1] = o[2] = s[0]; }
      else if (ctype == 2) { o[0] = s[0]; o[1] = s[1]; o[2] = s[2]; }
      else if (ctype == 3) {
        if ((size_t)s[0] * 3 + 2 >= plte.size()) { err = "png: palette index out of range"; return false; }
        o[0] = plte[s[0] * 3]; o[1] = plte[s[0] * 3 + 1]; o[2] = plte[s[0] * 3 + 2];
        if (s[0] < trns.size()) o[3] = trns[s[0]];
      } else if (ctype == 4) { o[0] = o[1] = o[2] = s[0]; o[3] = s[1]; }
      else { o[0] = s[0]; o[1] = s[1]; o[2] = s[2]; o[3] = s[3]; }
      o += 4;
    }
    prev.swap(cur);
  }
  return true;
}

bool decode_ppm(const uint8_t *d, size_t n, Image8 &img, std::string &err) {
  size_t pos = 0;
  auto token = [&](long &v) -> bool {
    for (;;) {
      while (pos < n && (d[pos] == ' ' || d[pos] == '\n' || d[pos] == '\r' || d[pos] == '\t')) ++pos;
      if (pos < n && d[pos] == '#') { while (pos < n && d[pos] != '\n') ++pos; continue; }
      break;
    }
    if (pos >= n || d[pos] < '0' || d[pos] > '9') return false;
    v = 0;
    while (pos < n && d[pos] >= '0' && d[pos] <= '9') v = v * 10 + (d[pos++] - '0');
    return true;
  };
  if (n < 2 || d[0] != 'P' || d[1] != '6') { err = "ppm: only binary P6 is supported"; return false; }
  pos = 2;
  long w, h, mx;
  if (!token(w) || !token(h) || !token(mx) || w <= 0 || h <= 0 || mx != 255) { err = "ppm: bad header (8-bit P6 expected)"; return false; }
  ++pos; // the single whitespace after maxval
  if (pos + (size_t)w * h * 3 > n) { err = "ppm: truncated"; return false; }
  img.w = (int)w; img.h = (int)h;
  img.rgba.resize((size_t)w * h * 4);
  for (size_t i = 0; i < (size_t)w * h; ++i) {
    img.rgba[4 * i] = d[pos + 3 * i]; img.rgba[4 * i + 1] = d[pos + 3 * i + 1]; img.rgba[4 * i + 2] = d[pos + 3 * i + 2];
    img.rgba[4 * i + 3] = 255;
  }
  return true;
}

bool read_image(const std::string &path, Image8 &out, std::string &err) {
  FILE *f = std::fopen(path.c_str(), "rb");
  if (!f) { err = "cannot open " + path; return false; }
  std::vector<uint8_t> buf;
  uint8_t tmp[65536];
  size_t k;
  while ((k = std::fread(tmp, 1, sizeof(tmp), f)) > 0) buf.insert(buf.end(), tmp, tmp + k);
  std::fclose(f);
  if (buf.size() >= 2 && buf[0] == 'P' && buf[1] == '6') return decode_ppm(buf.data(), buf.size(), out, err);
  return decode_png(buf.data(), buf.size(), out, err);
}

void encode_ppm(const uint8_t *rgba, int w, int h, std::vector<uint8_t> &out) {
  char hdr[64];
  const int n = std::snprintf(hdr, sizeof(hdr), "P6\n%d %d\n255\n", w, h);
  out.assign(hdr, hdr + n);
  out.reserve(out.size() + (size_t)w * h * 3);
  for (size_t i = 0; i < (size_t)w * h; ++i) {
    out.push_back(rgba[4 * i]); out.push_back(rgba[4 * i + 1]); out.push_back(rgba[4 * i + 2]);
  }
}

void encode_png(const uint8_t *rgba, int w, int h, std::vector<uint8_t> &out) {
  static const uint8_t sig[8] = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a};
  out.assign(sig, sig + 8);
  std::vector<uint8_t> ihdr;
  put32(ihdr, (uint32_t)w); put32(ihdr, (uint32_t)h);
  ihdr.push_back(8); ihdr.push_back(6); ihdr.push_back(0); ihdr.push_back(0); ihdr.push_back(0);
  chunk(out, "IHDR", ihdr);
  // filter type 0 on every row, zlib stream of stored blocks (no compression: a frame is written once)
  std::vector<uint8_t> raw;
  raw.reserve(((size_t)w * 4 + 1) * h);
  for (int y = 0; y < h; ++y) {
    raw.push_back(0);
    raw.insert(raw.end(), rgba + (size_t)y * w * 4, rgba + (size_t)(y + 1) * w * 4);
  }
  std::vector<uint8_t> z;
  z.push_back(0x78); z.push_back(0x01);
  size_t pos = 0;
  do {
    const size_t len = std::min<size_t>(65535, raw.size() - pos);
    z.push_back(pos + len == raw.size() ? 1 : 0);
    z.push_back((uint8_t)(len & 0xff)); z.push_back((uint8_t)(len >> 8));
    z.push_back((uint8_t)(~len & 0xff)); z.push_back((uint8_t)((~len >> 8) & 0xff));
    z.insert(z.end(), raw.begin() + pos, raw.begin() + pos + len);
    pos += len;
  } while (pos < raw.size());
  put32(z, adler32(raw.data(), raw.size()));
  chunk(out, "IDAT", z);
  chunk(out, "IEND", {});
}

} // namespace rtb
