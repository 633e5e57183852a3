// rt_image.cpp — image writers of librtcore.so: PPM, PFM and a baseline JPEG encoder.
// Replaces stbi_write_jpg in WriteImage (utils.cu:87-96).  Host-only, not on the hot path.
//
// The JPEG writer is a plain baseline sequential encoder (ITU-T T.81): YCbCr 4:4:4, the Annex K
// quantisation tables scaled by `quality`, the Annex K typical Huffman tables.
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/rtcore.h"

namespace rtc {
void set_error(const std::string &msg);
}

namespace {

const uint8_t kZigzag[64] = {0,  1,  8,  16, 9,  2,  3,  10, 17, 24, 32, 25, 18, 11, 4,  5,  12, 19, 26, 33, 40, 48,
                             41, 34, 27, 20, 13, 6,  7,  14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23,
                             30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};
const uint8_t kQLuma[64] = {16, 11, 10, 16, 24,  40,  51,  61,  12, 12, 14, 19, 26,  58,  60,  55,
                            14, 13, 16, 24, 40,  57,  69,  56,  14, 17, 22, 29, 51,  87,  80,  62,
                            18, 22, 37, 56, 68,  109, 103, 77,  24, 35, 55, 64, 81,  104, 113, 92,
                            49, 64, 78, 87, 103, 121, 120, 101, 72, 92, 95, 98, 112, 100, 103, 99};
const uint8_t kQChroma[64] = {17, 18, 24, 47, 99, 99, 99, 99, 18, 21, 26, 66, 99, 99, 99, 99, 24, 26, 56, 99, 99, 99,
                              99, 99, 47, 66, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99,
                              99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99};
// Annex K.3 Huffman specifications: BITS (codes per length 1..16) and HUFFVAL
const uint8_t kDcLumaBits[16] = {0, 1, 5, 1, 1, 1, 1, 1, 1, 0, 0, 0, 0, 0, 0, 0};
const uint8_t kDcChromaBits[16] = {0, 3, 1, 1, 1, 1, 1, 1, 1, 1, 1, 0, 0, 0, 0, 0};
const uint8_t kDcVals[12] = {0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11};
const uint8_t kAcLumaBits[16] = {0, 2, 1, 3, 3, 2, 4, 3, 5, 5, 4, 4, 0, 0, 1, 0x7d};
const uint8_t kAcLumaVals[162] = {
    0x01, 0x02, 0x03, 0x00, 0x04, 0x11, 0x05, 0x12, 0x21, 0x31, 0x41, 0x06, 0x13, 0x51, 0x61, 0x07, 0x22, 0x71,
    0x14, 0x32, 0x81, 0x91, 0xa1, 0x08, 0x23, 0x42, 0xb1, 0xc1, 0x15, 0x52, 0xd1, 0xf0, 0x24, 0x33, 0x62, 0x72,
    0x82, 0x09, 0x0a, 0x16, 0x17, 0x18, 0x19, 0x1a, 0x25, 0x26, 0x27, 0x28, 0x29, 0x2a, 0x34, 0x35, 0x36, 0x37,
    0x38, 0x39, 0x3a, 0x43, 0x44, 0x45, 0x46, 0x47, 0x48, 0x49, 0x4a, 0x53, 0x54, 0x55, 0x56, 0x57, 0x58, 0x59,
    0x5a, 0x63, 0x64, 0x65, 0x66, 0x67, 0x68, 0x69, 0x6a, 0x73, 0x74, 0x75, 0x76, 0x77, 0x78, 0x79, 0x7a, 0x83,
    0x84, 0x85, 0x86, 0x87, 0x88, 0x89, 0x8a, 0x92, 0x93, 0x94, 0x95, 0x96, 0x97, 0x98, 0x99, 0x9a, 0xa2, 0xa3,
    0xa4, 0xa5, 0xa6, 0xa7, 0xa8, 0xa9, 0xaa, 0xb2, 0xb3, 0xb4, 0xb5, 0xb6, 0xb7, 0xb8, 0xb9, 0xba, 0xc2, 0xc3,
    0xc4, 0xc5, 0xc6, 0xc7, 0xc8, 0xc9, 0xca, 0xd2, 0xd3, 0xd4, 0xd5, 0xd6, 0xd7, 0xd8, 0xd9, 0xda, 0xe1, 0xe2,
    0xe3, 0xe4, 0xe5, 0xe6, 0xe7, 0xe8, 0xe9, 0xea, 0xf1, 0xf2, 0xf3, 0xf4, 0xf5, 0xf6, 0xf7, 0xf8, 0xf9, 0xfa};
const uint8_t kAcChromaBits[16] = {0, 2, 1, 2, 4, 4, 3, 4, 7, 5, 4, 4, 0, 1, 2, 0x77};
const uint8_t kAcChromaVals[162] = {
    0x00, 0x01, 0x02, 0x03, 0x11, 0x04, 0x05, 0x21, 0x31, 0x06, 0x12, 0x41, 0x51, 0x07, 0x61, 0x71, 0x13, 0x22,
    0x32, 0x81, 0x08, 0x14, 0x42, 0x91, 0xa1, 0xb1, 0xc1, 0x09, 0x23, 0x33, 0x52, 0xf0, 0x15, 0x62, 0x72, 0xd1,
    0x0a, 0x16, 0x24, 0x34, 0xe1, 0x25, 0xf1, 0x17, 0x18, 0x19, 0x1a, 0x26, 0x27, 0x28, 0x29, 0x2a, 0x35, 0x36,
    0x37, 0x38, 0x39, 0x3a, 0x43, 0x44, 0x45, 0x46, 0x47, 0x48, 0x49, 0x4a, 0x53, 0x54, 0x55, 0x56, 0x57, 0x58,
    0x59, 0x5a, 0x63, 0x64, 0x65, 0x66, 0x67, 0x68, 0x69, 0x6a, 0x73, 0x74, 0x75, 0x76, 0x77, 0x78, 0x79, 0x7a,
    0x82, 0x83, 0x84, 0x85, 0x86, 0x87, 0x88, 0x89, 0x8a, 0x92, 0x93, 0x94, 0x95, 0x96, 0x97, 0x98, 0x99, 0x9a,
    0xa2, 0xa3, 0xa4, 0xa5, 0xa6, 0xa7, 0xa8, 0xa9, 0xaa, 0xb2, 0xb3, 0xb4, 0xb5, 0xb6, 0xb7, 0xb8, 0xb9, 0xba,
    0xc2, 0xc3, 0xc4, 0xc5, 0xc6, 0xc7, 0xc8, 0xc9, 0xca, 0xd2, 0xd3, 0xd4, 0xd5, 0xd6, 0xd7, 0xd8, 0xd9, 0xda,
    0xe2, 0xe3, 0xe4, 0xe5, 0xe6, 0xe7, 0xe8, 0xe9, 0xea, 0xf2, 0xf3, 0xf4, 0xf5, 0xf6, 0xf7, 0xf8, 0xf9, 0xfa};

struct HuffTable {
  uint16_t code[256];
  uint8_t len[256];
  void build(const uint8_t bits[16], const uint8_t *vals) {
    memset(len, 0, sizeof(len));
    uint16_t c = 0;
    int k = 0;
    for (int l = 1; l <= 16; l++) {
      for (int i = 0; i < bits[l - 1]; i++) {
        code[vals[k]] = c++;
        len[vals[k]] = (uint8_t)l;
        k++;
      }
      c <<= 1;
    }
  }
};

struct BitWriter {
  std::vector<uint8_t> &out;
  uint32_t acc = 0;
  int n = 0;
  explicit BitWriter(std::vector<uint8_t> &o) : out(o) {}
  void put(uint32_t bits, int len) {
    acc = (acc << len) | (bits & ((1u << len) - 1));
    n += len;
    while (n >= 8) {
      uint8_t b = (uint8_t)(acc >> (n - 8));
      out.push_back(b);
      if (b == 0xff) out.push_back(0);
      n -= 8;
    }
  }
  void flush() {
    if (n > 0) put(0x7f, 8 - n);
  }
};

void fdct8x8(float *b) {  // separable direct DCT-II, double accumulation
  static double c[8][8];
  static bool init = false;
  if (!init) {
    for (int u = 0; u < 8; u++)
      for (int x = 0; x < 8; x++) c[u][x] = (u == 0 ? std::sqrt(0.125) : 0.5) * std::cos((2 * x + 1) * u * M_PI / 16.0);
    init = true;
  }
  double t[64];
  for (int y = 0; y < 8; y++)
    for (int u = 0; u < 8; u++) {
      double s = 0;
      for (int x = 0; x < 8; x++) s += c[u][x] * b[y * 8 + x];
      t[y * 8 + u] = s;
    }
  for (int u = 0; u < 8; u++)
    for (int v = 0; v < 8; v++) {
      double s = 0;
      for (int y = 0; y < 8; y++) s += c[v][y] * t[y * 8 + u];
      b[v * 8 + u] = (float)s;
    }
}

void put_marker(std::vector<uint8_t> &o, uint8_t m) {
  o.push_back(0xff);
  o.push_back(m);
}
void put16(std::vector<uint8_t> &o, int v) {
  o.push_back((uint8_t)(v >> 8));
  o.push_back((uint8_t)v);
}

int encode_block(BitWriter &bw, float *blk, const uint8_t *q, int prev_dc, const HuffTable &dc, const HuffTable &ac) {
  fdct8x8(blk);
  int z[64];
  for (int i = 0; i < 64; i++) z[i] = (int)std::lrint(blk[kZigzag[i]] / q[kZigzag[i]]);
  auto magnitude = [](int v, int *bits) {
    int a = v < 0 ? -v : v, n = 0;
    while (a >> n) n++;
    *bits = v < 0 ? v - 1 : v;
    return n;
  };
  int bits, diff = z[0] - prev_dc;
  int n = magnitude(diff, &bits);
  bw.put(dc.code[n], dc.len[n]);
  if (n) bw.put((uint32_t)bits, n);
  int last = 63;
  while (last > 0 && z[last] == 0) last--;
  int run = 0;
  for (int i = 1; i <= last; i++) {
    if (z[i] == 0) {
      run++;
      continue;
    }
    while (run >= 16) {
      bw.put(ac.code[0xf0], ac.len[0xf0]);
      run -= 16;
    }
    n = magnitude(z[i], &bits);
    bw.put(ac.code[(run << 4) | n], ac.len[(run << 4) | n]);
    bw.put((uint32_t)bits, n);
    run = 0;
  }
  if (last != 63) bw.put(ac.code[0], ac.len[0]);
  return z[0];
}

bool write_file(const char *path, const void *data, size_t n) {
  FILE *f = fopen(path, "wb");
  if (!f) return false;
  bool ok = fwrite(data, 1, n, f) == n;
  return fclose(f) == 0 && ok;
}

}  // namespace

extern "C" int rt_write_jpeg(const char *path, int width, int height, const uint8_t *rgb, int quality) {
  if (!path || !rgb || width <= 0 || height <= 0 || width > 65535 || height > 65535) {
    rtc::set_error("rt_write_jpeg: bad argument");
    return RT_ERR_INVALID;
  }
  quality = quality < 1 ? 1 : (quality > 100 ? 100 : quality);
  int scale = quality < 50 ? 5000 / quality : 200 - 2 * quality;
  uint8_t ql[64], qc[64];
  for (int i = 0; i < 64; i++) {
    int a = (kQLuma[i] * scale + 50) / 100, b = (kQChroma[i] * scale + 50) / 100;
    ql[i] = (uint8_t)(a < 1 ? 1 : (a > 255 ? 255 : a));
    qc[i] = (uint8_t)(b < 1 ? 1 : (b > 255 ? 255 : b));
  }
  HuffTable dcl, dcc, acl, acc;
  dcl.build(kDcLumaBits, kDcVals);
  dcc.build(kDcChromaBits, kDcVals);
  acl.build(kAcLumaBits, kAcLumaVals);
  acc.build(kAcChromaBits, kAcChromaVals);

  std::vector<uint8_t> o;
  put_marker(o, 0xd8);  // SOI
  put_marker(o, 0xe0);  // APP0 JFIF
  put16(o, 16);
  const uint8_t jfif[] = {'J', 'F', 'I', 'F', 0, 1, 1, 0, 0, 1, 0, 1, 0, 0};
  o.insert(o.end(), jfif, jfif + sizeof(jfif));
  for (int t = 0; t < 2; t++) {  // DQT
    put_marker(o, 0xdb);
    put16(o, 67);
    o.push_back((uint8_t)t);
    for (int i = 0; i < 64; i++) o.push_back((t ? qc : ql)[kZigzag[i]]);
  }
  put_marker(o, 0xc0);  // SOF0
  put16(o, 17);
  o.push_back(8);
  put16(o, height);
  put16(o, width);
  o.push_back(3);
  for (int c = 0; c < 3; c++) {
    o.push_back((uint8_t)(c + 1));
    o.push_back(0x11);
    o.push_back(c ? 1 : 0);
  }
  struct {
    int cls_id;
    const uint8_t *bits, *vals;
    int nvals;
  } hts[4] = {{0x00, kDcLumaBits, kDcVals, 12},
              {0x10, kAcLumaBits, kAcLumaVals, 162},
              {0x01, kDcChromaBits, kDcVals, 12},
              {0x11, kAcChromaBits, kAcChromaVals, 162}};
  for (auto &h : hts) {  // DHT
    put_marker(o, 0xc4);
    put16(o, 19 + h.nvals);
    o.push_back((uint8_t)h.cls_id);
    o.insert(o.end(), h.bits, h.bits + 16);
    o.insert(o.end(), h.vals, h.vals + h.nvals);
  }
  put_marker(o, 0xda);  // SOS
  put16(o, 12);
  o.push_back(3);
  for (int c = 0; c < 3; c++) {
    o.push_back((uint8_t)(c + 1));
    o.push_back(c ? 0x11 : 0x00);
  }
  o.push_back(0);
  o.push_back(63);
  o.push_back(0);

  BitWriter bw(o);
  int dc[3] = {0, 0, 0};
  float blk[3][64];
  for (int by = 0; by < height; by += 8)
    for (int bx = 0; bx < width; bx += 8) {
      for (int y = 0; y < 8; y++)
        for (int x = 0; x < 8; x++) {
          int sy = by + y < height ? by + y : height - 1, sx = bx + x < width ? bx + x : width - 1;
          const uint8_t *p = rgb + 3 * ((size_t)sy * width + sx);
          float r = p[0], g = p[1], b = p[2];
          blk[0][y * 8 + x] = 0.299f * r + 0.587f * g + 0.114f * b - 128.f;
          blk[1][y * 8 + x] = -0.168736f * r - 0.331264f * g + 0.5f * b;
          blk[2][y * 8 + x] = 0.5f * r - 0.418688f * g - 0.081312f * b;
        }
      dc[0] = encode_block(bw, blk[0], ql, dc[0], dcl, acl);
      dc[1] = encode_block(bw, blk[1], qc, dc[1], dcc, acc);
      dc[2] = encode_block(bw, blk[2], qc, dc[2], dcc, acc);
    }
  bw.flush();
  put_marker(o, 0xd9);  // EOI
  if (!write_file(path, o.data(), o.size())) {
    rtc::set_error(std::string("rt_write_jpeg: cannot write ") + path);
    return RT_ERR_INVALID;
  }
  return RT_OK;
}

extern "C" int rt_write_ppm(const char *path, int width, int height, const uint8_t *rgb) {
  if (!path || !rgb || width <= 0 || height <= 0) {
    rtc::set_error("rt_write_ppm: bad argument");
    return RT_ERR_INVALID;
  }
  std::string hdr = "P6\n" + std::to_string(width) + " " + std::to_string(height) + "\n255\n";
  std::vector<uint8_t> o(hdr.begin(), hdr.end());
  o.insert(o.end(), rgb, rgb + (size_t)width * height * 3);
  if (!write_file(path, o.data(), o.size())) {
    rtc::set_error(std::string("rt_write_ppm: cannot write ") + path);
    return RT_ERR_INVALID;
  }
  return RT_OK;
}

// PFM stores rows bottom-to-top; negative scale = little endian.
extern "C" int rt_write_pfm(const char *path, int width, int height, const float *rgb) {
  if (!path || !rgb || width <= 0 || height <= 0) {
    rtc::set_error("rt_write_pfm: bad argument");
    return RT_ERR_INVALID;
  }
  std::string hdr = "PF\n" + std::to_string(width) + " " + std::to_string(height) + "\n-1.0\n";
  std::vector<uint8_t> o(hdr.begin(), hdr.end());
  for (int y = height - 1; y >= 0; y--) {
    const uint8_t *row = reinterpret_cast<const uint8_t *>(rgb + (size_t)y * width * 3);
    o.insert(o.end(), row, row + (size_t)width * 12);
  }
  if (!write_file(path, o.data(), o.size())) {
    rtc::set_error(std::string("rt_write_pfm: cannot write ") + path);
    return RT_ERR_INVALID;
  }
  return RT_OK;
}
