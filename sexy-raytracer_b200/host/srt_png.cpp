// srt_png.cpp — PNG read/write for the host side (libsrt_host.so); zlib does inflate/deflate/crc32, the rest is here.
//
//   srt_png_read  : what `stbi_load(filename, &w, &h, &n, reqComp)` gives the reference's imagePNG (texture.h:112-123):
//                   8-bit samples, row 0 = top, converted to reqComp channels with stb_image's rules
//                   (16-bit -> high byte; 1/2/4-bit grey scaled to 0..255; palette expanded; grey <-> RGB by replication /
//                   luma (77 r + 150 g + 29 b) >> 8; alpha dropped).  Non-interlaced and Adam7 images.
//   srt_png_write : the output stage after the path - stbi_write_png("test.png", w, h, 4, target, 4*w) (main.cpp:235-237):
//                   8-bit, 1-4 channels, filter 0 rows, one IDAT.  The FILE bytes differ from stb's (other deflate), the
//                   decoded pixels do not.
#include <zlib.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/srt_host.h"

namespace srt_host_detail {
srt_status fail(srt_status s, const std::string& m);
}
using srt_host_detail::fail;

namespace {

uint32_t be32(const uint8_t* p) { return ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | p[3]; }

int paeth(int a, int b, int c) {
  const int p = a + b - c, pa = abs(p - a), pb = abs(p - b), pc = abs(p - c);
  return (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c);
}

// undo the per-row filters of one (sub-)image: in = (1 + stride) * h bytes, out = stride * h bytes
bool unfilter(const uint8_t* in, uint8_t* out, size_t stride, size_t h, size_t bpp) {
  for (size_t y = 0; y < h; y++) {
    const uint8_t ft = in[y * (stride + 1)];
    const uint8_t* src = in + y * (stride + 1) + 1;
    uint8_t* cur = out + y * stride;
    const uint8_t* up = y ? out + (y - 1) * stride : nullptr;
    for (size_t x = 0; x < stride; x++) {
      const int a = x >= bpp ? cur[x - bpp] : 0, b = up ? up[x] : 0, c = (up && x >= bpp) ? up[x - bpp] : 0;
      int v;
      switch (ft) {
        case 0: v = src[x]; break;
        case 1: v = src[x] + a; break;
        case 2: v = src[x] + b; break;
        case 3: v = src[x] + ((a + b) >> 1); break;
        case 4: v = src[x] + paeth(a, b, c); break;
        default: return false;
      }
      cur[x] = (uint8_t)v;
    }
  }
  return true;
}

struct Png {
  uint32_t w = 0, h = 0;
  int depth = 0, ctype = 0, interlace = 0;
  std::vector<uint8_t> idat, plte;
};

int channels_of(int ctype) { return ctype == 0 ? 1 : ctype == 2 ? 3 : ctype == 3 ? 1 : ctype == 4 ? 2 : 4; }

// one decoded sample row set (w x h, `ch` samples per pixel at `depth` bits) -> 8-bit samples at (x0 + x*dx, y0 + y*dy)
void scatter_pass(const Png& P, const uint8_t* px, uint32_t pw, uint32_t ph, uint32_t x0, uint32_t y0, uint32_t dx, uint32_t dy,
                  std::vector<uint8_t>& img, int ch) {
  const size_t stride = ((size_t)pw * ch * P.depth + 7) / 8;
  static const int scale[9] = {0, 0xff, 0x55, 0, 0x11, 0, 0, 0, 0x01};
  for (uint32_t y = 0; y < ph; y++) {
    const uint8_t* row = px + y * stride;
    for (uint32_t x = 0; x < pw; x++)
      for (int c = 0; c < ch; c++) {
        const size_t s = (size_t)x * ch + c;
        uint8_t v;
        if (P.depth == 8) v = row[s];
        else if (P.depth == 16) v = row[2 * s];                      // stb: high byte
        else {
          const int per = 8 / P.depth;
          const int sh = (per - 1 - (int)(s % per)) * P.depth;
          v = (row[s / per] >> sh) & ((1 << P.depth) - 1);
          if (P.ctype == 0) v = (uint8_t)(v * scale[P.depth]);       // stb scales grey, leaves palette indices
        }
        img[((size_t)(y0 + y * dy) * P.w + (x0 + x * dx)) * ch + c] = v;
      }
  }
}

}  // namespace

extern "C" {

void srt_host_free(void* p) { free(p); }

srt_status srt_png_read_memory(const uint8_t* buf, int64_t len, int32_t req_comp, int32_t* out_w, int32_t* out_h, uint8_t** out) try {
  static const uint8_t sig[8] = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a};
  if (!buf || !out_w || !out_h || !out) return fail(SRT_ERR_INVALID, "NULL argument");
  if (req_comp < 1 || req_comp > 4) return fail(SRT_ERR_INVALID, "req_comp must be 1..4");
  *out = nullptr; *out_w = *out_h = 0;
  if (len < 8 || memcmp(buf, sig, 8) != 0) return fail(SRT_ERR_INVALID, "not a PNG file");
  Png P;
  bool have_ihdr = false, end = false;
  for (int64_t pos = 8; pos + 12 <= len && !end;) {
    const uint32_t n = be32(buf + pos);
    const uint8_t* type = buf + pos + 4;
    const uint8_t* data = buf + pos + 8;
    if (pos + 12 + (int64_t)n > len) return fail(SRT_ERR_INVALID, "truncated PNG chunk");
    if (!memcmp(type, "IHDR", 4)) {
      if (n != 13) return fail(SRT_ERR_INVALID, "bad IHDR");
      P.w = be32(data); P.h = be32(data + 4); P.depth = data[8]; P.ctype = data[9]; P.interlace = data[12];
      if (data[10] || data[11] || P.interlace > 1) return fail(SRT_ERR_UNSUPPORTED, "PNG compression/filter/interlace method");
      have_ihdr = true;
    } else if (!memcmp(type, "PLTE", 4)) {
      P.plte.assign(data, data + n);
    } else if (!memcmp(type, "IDAT", 4)) {
      P.idat.insert(P.idat.end(), data, data + n);
    } else if (!memcmp(type, "IEND", 4)) {
      end = true;
    }
    pos += 12 + (int64_t)n;
  }
  if (!have_ihdr || P.idat.empty() || !P.w || !P.h) return fail(SRT_ERR_INVALID, "PNG without IHDR/IDAT");
  if (P.w > (1u << 24) || P.h > (1u << 24)) return fail(SRT_ERR_UNSUPPORTED, "PNG too large");
  const bool depth_ok = (P.ctype == 0 && (P.depth == 1 || P.depth == 2 || P.depth == 4 || P.depth == 8 || P.depth == 16)) ||
                        (P.ctype == 3 && (P.depth == 1 || P.depth == 2 || P.depth == 4 || P.depth == 8)) ||
                        ((P.ctype == 2 || P.ctype == 4 || P.ctype == 6) && (P.depth == 8 || P.depth == 16));
  if (!depth_ok) return fail(SRT_ERR_INVALID, "bad PNG colour type / bit depth");
  if (P.ctype == 3 && P.plte.size() < 3) return fail(SRT_ERR_INVALID, "palette PNG without PLTE");
  const int ch = channels_of(P.ctype);
  const size_t bpp = (size_t)(ch * P.depth + 7) / 8;

  // passes: (x0, y0, dx, dy)
  static const uint32_t adam7[7][4] = {{0, 0, 8, 8}, {4, 0, 8, 8}, {0, 4, 4, 8}, {2, 0, 4, 4}, {0, 2, 2, 4}, {1, 0, 2, 2}, {0, 1, 1, 2}};
  static const uint32_t whole[1][4] = {{0, 0, 1, 1}};
  const uint32_t(*pass)[4] = P.interlace ? adam7 : whole;
  const int n_pass = P.interlace ? 7 : 1;
  size_t raw_len = 0;
  for (int k = 0; k < n_pass; k++) {
    const uint32_t pw = (P.w - pass[k][0] + pass[k][2] - 1) / pass[k][2], ph = (P.h - pass[k][1] + pass[k][3] - 1) / pass[k][3];
    if (pw && ph) raw_len += (((size_t)pw * ch * P.depth + 7) / 8 + 1) * ph;
  }
  // deflate expands at most ~1032:1: a header that promises more than its IDAT bytes can hold is rejected before anything of
  // that size is allocated (width and height are only bounded by 2^24 each)
  if (raw_len > P.idat.size() * (size_t)1040 + 65536) return fail(SRT_ERR_INVALID, "PNG: image data too short for its dimensions");
  std::vector<uint8_t> raw(raw_len);
  uLongf got = (uLongf)raw_len;
  const int zr = uncompress(raw.data(), &got, P.idat.data(), (uLong)P.idat.size());
  if (zr != Z_OK || got != raw_len) return fail(SRT_ERR_INVALID, "PNG: corrupt image data");

  std::vector<uint8_t> img((size_t)P.w * P.h * ch);
  size_t off = 0;
  for (int k = 0; k < n_pass; k++) {
    const uint32_t pw = (P.w - pass[k][0] + pass[k][2] - 1) / pass[k][2], ph = (P.h - pass[k][1] + pass[k][3] - 1) / pass[k][3];
    if (!pw || !ph) continue;
    const size_t stride = ((size_t)pw * ch * P.depth + 7) / 8;
    std::vector<uint8_t> px(stride * ph);
    if (!unfilter(raw.data() + off, px.data(), stride, ph, bpp)) return fail(SRT_ERR_INVALID, "PNG: bad filter type");
    off += (stride + 1) * ph;
    scatter_pass(P, px.data(), pw, ph, pass[k][0], pass[k][1], pass[k][2], pass[k][3], img, ch);
  }

  // -> `src_n` interleaved 8-bit channels
  int src_n = ch;
  if (P.ctype == 3) {
    std::vector<uint8_t> rgb((size_t)P.w * P.h * 3);
    const size_t np = P.plte.size() / 3;
    for (size_t i = 0; i < (size_t)P.w * P.h; i++) {
      const size_t q = img[i] < np ? img[i] : 0;
      rgb[3 * i] = P.plte[3 * q]; rgb[3 * i + 1] = P.plte[3 * q + 1]; rgb[3 * i + 2] = P.plte[3 * q + 2];
    }
    img.swap(rgb);
    src_n = 3;
  }
  // stb_image's stbi__convert_format
  const size_t npx = (size_t)P.w * P.h;
  uint8_t* o = (uint8_t*)malloc(npx * req_comp + 4);     // + 4: a 1-byte map is read 2 bytes past its last texel (material.h:197)
  if (!o) return fail(SRT_ERR_NOMEM, "PNG pixels");
  memset(o + npx * req_comp, 0, 4);
  for (size_t i = 0; i < npx; i++) {
    const uint8_t* s = img.data() + i * src_n;
    uint8_t r, g, b, a = 255;
    if (src_n <= 2) { r = g = b = s[0]; if (src_n == 2) a = s[1]; }
    else { r = s[0]; g = s[1]; b = s[2]; if (src_n == 4) a = s[3]; }
    const uint8_t y = src_n <= 2 ? s[0] : (uint8_t)(((r * 77) + (g * 150) + (29 * b)) >> 8);
    uint8_t* d = o + i * req_comp;
    switch (req_comp) {
      case 1: d[0] = y; break;
      case 2: d[0] = y; d[1] = a; break;
      case 3: d[0] = r; d[1] = g; d[2] = b; break;
      default: d[0] = r; d[1] = g; d[2] = b; d[3] = a; break;
    }
  }
  *out = o; *out_w = (int32_t)P.w; *out_h = (int32_t)P.h;
  return SRT_OK;
} catch (const std::bad_alloc&) { return fail(SRT_ERR_NOMEM, "out of memory"); } catch (const std::exception& e_) { return fail(SRT_ERR_INVALID, e_.what()); } catch (...) { return fail(SRT_ERR_INVALID, "unknown exception"); }

srt_status srt_png_read(const char* path, int32_t req_comp, int32_t* out_w, int32_t* out_h, uint8_t** out) try {
  if (!path) return fail(SRT_ERR_INVALID, "NULL path");
  FILE* f = fopen(path, "rb");
  if (!f) return fail(SRT_ERR_INVALID, std::string("cannot open ") + path);
  std::vector<uint8_t> buf;
  uint8_t tmp[65536];
  size_t n;
  while ((n = fread(tmp, 1, sizeof tmp, f)) > 0) buf.insert(buf.end(), tmp, tmp + n);
  fclose(f);
  return srt_png_read_memory(buf.data(), (int64_t)buf.size(), req_comp, out_w, out_h, out);
} catch (const std::bad_alloc&) { return fail(SRT_ERR_NOMEM, "out of memory"); } catch (const std::exception& e_) { return fail(SRT_ERR_INVALID, e_.what()); } catch (...) { return fail(SRT_ERR_INVALID, "unknown exception"); }

srt_status srt_png_write(const char* path, int32_t w, int32_t h, int32_t comp, const uint8_t* pixels, int32_t stride_bytes) try {
  if (!path || !pixels) return fail(SRT_ERR_INVALID, "NULL argument");
  if (w <= 0 || h <= 0 || comp < 1 || comp > 4) return fail(SRT_ERR_INVALID, "bad image size / channel count");
  if (stride_bytes == 0) stride_bytes = w * comp;
  const size_t row = (size_t)w * comp;
  std::vector<uint8_t> raw((row + 1) * h);
  for (int32_t y = 0; y < h; y++) {
    raw[(row + 1) * y] = 0;
    memcpy(raw.data() + (row + 1) * y + 1, pixels + (size_t)stride_bytes * y, row);
  }
  uLongf clen = compressBound((uLong)raw.size());
  std::vector<uint8_t> z(clen);
  if (compress2(z.data(), &clen, raw.data(), (uLong)raw.size(), 6) != Z_OK) return fail(SRT_ERR_NOMEM, "deflate failed");
  FILE* f = fopen(path, "wb");
  if (!f) return fail(SRT_ERR_INVALID, std::string("cannot create ") + path);
  auto chunk = [&](const char* type, const uint8_t* data, uint32_t n) {
    uint8_t hd[8] = {(uint8_t)(n >> 24), (uint8_t)(n >> 16), (uint8_t)(n >> 8), (uint8_t)n, (uint8_t)type[0], (uint8_t)type[1],
                     (uint8_t)type[2], (uint8_t)type[3]};
    uint32_t c = (uint32_t)crc32(0, hd + 4, 4);
    if (n) c = (uint32_t)crc32(c, data, n);
    const uint8_t tl[4] = {(uint8_t)(c >> 24), (uint8_t)(c >> 16), (uint8_t)(c >> 8), (uint8_t)c};
    return fwrite(hd, 1, 8, f) == 8 && (n == 0 || fwrite(data, 1, n, f) == n) && fwrite(tl, 1, 4, f) == 4;
  };
  static const uint8_t sig[8] = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a};
  static const uint8_t ctype[5] = {0, 0, 4, 2, 6};
  const uint8_t ihdr[13] = {(uint8_t)(w >> 24), (uint8_t)(w >> 16), (uint8_t)(w >> 8), (uint8_t)w, (uint8_t)(h >> 24), (uint8_t)(h >> 16),
                            (uint8_t)(h >> 8), (uint8_t)h, 8, ctype[comp], 0, 0, 0};
  bool ok = fwrite(sig, 1, 8, f) == 8 && chunk("IHDR", ihdr, 13) && chunk("IDAT", z.data(), (uint32_t)clen) && chunk("IEND", nullptr, 0);
  ok = (fclose(f) == 0) && ok;
  return ok ? SRT_OK : fail(SRT_ERR_INVALID, std::string("write error on ") + path);
} catch (const std::bad_alloc&) { return fail(SRT_ERR_NOMEM, "out of memory"); } catch (const std::exception& e_) { return fail(SRT_ERR_INVALID, e_.what()); } catch (...) { return fail(SRT_ERR_INVALID, "unknown exception"); }


// Checkpoint file "SRTF": version 1 = {w, h, samples_done, seed, sums}; version 2 adds {sampler, max_bounce} after the seed so
// that a resumed render can be refused when it would mix two sample streams (b200Device::rtFrameProgressive).
srt_status srt_pfm_write(const char* path, int32_t w, int32_t h, const float* rgb, float scale) try {
  if (!path || !rgb) return fail(SRT_ERR_INVALID, "NULL argument");
  if (w <= 0 || h <= 0) return fail(SRT_ERR_INVALID, "bad image size");
  FILE* f = fopen(path, "wb");
  if (!f) return fail(SRT_ERR_INVALID, std::string("cannot open ") + path);
  fprintf(f, "PF\n%d %d\n-1.0\n", w, h);                      // negative scale = little endian
  std::vector<float> row((size_t)w * 3);
  bool ok = true;
  for (int y = h - 1; y >= 0 && ok; y--) {                      // PFM rows run bottom to top
    const float* src = rgb + (size_t)y * w * 3;
    for (size_t k = 0; k < row.size(); k++) row[k] = src[k] * scale;
    ok = fwrite(row.data(), sizeof(float), row.size(), f) == row.size();
  }
  ok = (fclose(f) == 0) && ok;
  return ok ? SRT_OK : fail(SRT_ERR_INVALID, std::string("short write to ") + path);
} catch (const std::bad_alloc&) { return fail(SRT_ERR_NOMEM, "out of memory"); } catch (const std::exception& e_) { return fail(SRT_ERR_INVALID, e_.what()); } catch (...) { return fail(SRT_ERR_INVALID, "unknown exception"); }

srt_status srt_pfm_read(const char* path, int32_t* out_w, int32_t* out_h, float** out_rgb) try {
  if (!path || !out_w || !out_h || !out_rgb) return fail(SRT_ERR_INVALID, "NULL argument");
  *out_rgb = nullptr;
  FILE* f = fopen(path, "rb");
  if (!f) return fail(SRT_ERR_INVALID, std::string("cannot open ") + path);
  char magic[3] = {0, 0, 0};
  int w = 0, h = 0;
  float scale = 0;
  if (fscanf(f, "%2s %d %d %f", magic, &w, &h, &scale) != 4 || strcmp(magic, "PF") != 0 || fgetc(f) == EOF) { fclose(f); return fail(SRT_ERR_INVALID, "not a colour PFM file"); }
  if (w <= 0 || h <= 0 || (int64_t)w * h > ((int64_t)1 << 28) || !(scale < 0)) { fclose(f); return fail(SRT_ERR_UNSUPPORTED, "PFM: bad size or big-endian data"); }
  float* px = static_cast<float*>(malloc((size_t)w * h * 3 * sizeof(float)));
  if (!px) { fclose(f); return fail(SRT_ERR_NOMEM, "out of memory"); }
  bool ok = true;
  for (int y = h - 1; y >= 0 && ok; y--) ok = fread(px + (size_t)y * w * 3, sizeof(float), (size_t)w * 3, f) == (size_t)w * 3;
  fclose(f);
  if (!ok) { free(px); return fail(SRT_ERR_INVALID, "PFM: truncated file"); }
  const float m = -scale;
  if (m != 1.0f) for (size_t k = 0; k < (size_t)w * h * 3; k++) px[k] *= m;
  *out_w = w; *out_h = h; *out_rgb = px;
  return SRT_OK;
} catch (const std::bad_alloc&) { return fail(SRT_ERR_NOMEM, "out of memory"); } catch (const std::exception& e_) { return fail(SRT_ERR_INVALID, e_.what()); } catch (...) { return fail(SRT_ERR_INVALID, "unknown exception"); }

srt_status srt_checkpoint_write2(const char* path, int32_t w, int32_t h, int32_t samples_done, uint64_t seed, int32_t sampler,
                                 int32_t max_bounce, const float* sum) try {
  if (!path || !sum || w <= 0 || h <= 0 || samples_done < 0) return fail(SRT_ERR_INVALID, "bad argument");
  FILE* f = fopen(path, "wb");
  if (!f) return fail(SRT_ERR_INVALID, std::string("cannot create ") + path);
  const uint32_t ver = 2;
  const size_t n = (size_t)w * h * 3;
  bool ok = fwrite("SRTF", 1, 4, f) == 4 && fwrite(&ver, 4, 1, f) == 1 && fwrite(&w, 4, 1, f) == 1 && fwrite(&h, 4, 1, f) == 1 &&
            fwrite(&samples_done, 4, 1, f) == 1 && fwrite(&seed, 8, 1, f) == 1 && fwrite(&sampler, 4, 1, f) == 1 &&
            fwrite(&max_bounce, 4, 1, f) == 1 && fwrite(sum, sizeof(float), n, f) == n;
  ok = (fclose(f) == 0) && ok;
  return ok ? SRT_OK : fail(SRT_ERR_INVALID, std::string("write error on ") + path);
} catch (const std::bad_alloc&) { return fail(SRT_ERR_NOMEM, "out of memory"); } catch (const std::exception& e_) { return fail(SRT_ERR_INVALID, e_.what()); } catch (...) { return fail(SRT_ERR_INVALID, "unknown exception"); }

srt_status srt_checkpoint_read2(const char* path, int32_t* w, int32_t* h, int32_t* samples_done, uint64_t* seed, int32_t* sampler,
                                int32_t* max_bounce, float** out) try {
  if (!path || !w || !h || !samples_done || !seed || !out) return fail(SRT_ERR_INVALID, "NULL argument");
  *out = nullptr;
  FILE* f = fopen(path, "rb");
  if (!f) return fail(SRT_ERR_INVALID, std::string("cannot open ") + path);
  char magic[4]; uint32_t ver = 0;
  int32_t smp = -1, mb = -1;                       // version 1 files do not say
  bool ok = fread(magic, 1, 4, f) == 4 && !memcmp(magic, "SRTF", 4) && fread(&ver, 4, 1, f) == 1 && (ver == 1 || ver == 2) &&
            fread(w, 4, 1, f) == 1 && fread(h, 4, 1, f) == 1 && fread(samples_done, 4, 1, f) == 1 && fread(seed, 8, 1, f) == 1;
  if (ok && ver == 2) ok = fread(&smp, 4, 1, f) == 1 && fread(&mb, 4, 1, f) == 1;
  if (!ok || *w <= 0 || *h <= 0 || *samples_done < 0 || (int64_t)*w * *h > (1ll << 32)) { fclose(f); return fail(SRT_ERR_INVALID, "not a checkpoint file"); }
  if (sampler) *sampler = smp;
  if (max_bounce) *max_bounce = mb;
  const size_t n = (size_t)*w * *h * 3;
  float* p = (float*)malloc(n * sizeof(float));
  if (!p) { fclose(f); return fail(SRT_ERR_NOMEM, "checkpoint buffer"); }
  const bool full = fread(p, sizeof(float), n, f) == n;
  fclose(f);
  if (!full) { free(p); return fail(SRT_ERR_INVALID, "truncated checkpoint file"); }
  *out = p;
  return SRT_OK;
} catch (const std::bad_alloc&) { return fail(SRT_ERR_NOMEM, "out of memory"); } catch (const std::exception& e_) { return fail(SRT_ERR_INVALID, e_.what()); } catch (...) { return fail(SRT_ERR_INVALID, "unknown exception"); }

srt_status srt_checkpoint_write(const char* path, int32_t w, int32_t h, int32_t samples_done, uint64_t seed, const float* sum) {
  return srt_checkpoint_write2(path, w, h, samples_done, seed, -1, -1, sum);
}
srt_status srt_checkpoint_read(const char* path, int32_t* w, int32_t* h, int32_t* samples_done, uint64_t* seed, float** out) {
  return srt_checkpoint_read2(path, w, h, samples_done, seed, nullptr, nullptr, out);
}

}  // extern "C"
