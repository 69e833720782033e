// rt_image_io.h — minimal image codecs for the caller side of the path: the reference's
// frames end up in PNG files (outputs/*.png) and its meshes carry PNG textures, both through the course
// framework's absent Image class.  Own code, no third-party library: binary PPM (P6) and PNG (8-bit grey /
// RGB / palette / RGBA, non-interlaced) readers, PPM writer, PNG writer (stored deflate blocks).
#ifndef RTB_IMAGE_IO_H
#define RTB_IMAGE_IO_H
#include <cstdint>
#include <string>
#include <vector>

namespace rtb {

struct Image8 {
  int w = 0, h = 0;
  std::vector<uint8_t> rgba; // 4 bytes per pixel, row 0 = TOP row of the file
};

bool read_image(const std::string &path, Image8 &out, std::string &err);       // by content: "P6" or PNG signature
bool decode_png(const uint8_t *data, size_t n, Image8 &out, std::string &err);
bool decode_ppm(const uint8_t *data, size_t n, Image8 &out, std::string &err);
void encode_png(const uint8_t *rgba_top_first, int w, int h, std::vector<uint8_t> &out);   // 8-bit RGBA
void encode_ppm(const uint8_t *rgba_top_first, int w, int h, std::vector<uint8_t> &out);   // P6, alpha dropped
bool inflate_zlib(const uint8_t *src, size_t n, std::vector<uint8_t> &out, std::string &err);

} // namespace rtb
#endif
