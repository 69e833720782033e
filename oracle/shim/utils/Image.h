// TEST INFRASTRUCTURE (oracle shim). Pixel store addressed (x, y), row-major y*W+x.
#ifndef SHIM_IMAGE_H
#define SHIM_IMAGE_H
#include "utils/vec4.h"
#include <vector>
class Image {
public:
  Image() {}
  Image(unsigned w, unsigned h) { resize(w, h); }
  void resize(unsigned w, unsigned h) { w_ = w; h_ = h; px_.assign(size_t(w) * h, vec4()); }
  unsigned width() const { return w_; }
  unsigned height() const { return h_; }
  vec4 &operator()(unsigned x, unsigned y) { return px_[size_t(y) * w_ + x]; }
  const vec4 &operator()(unsigned x, unsigned y) const { return px_[size_t(y) * w_ + x]; }
  vec4 *data() { return px_.data(); }
private:
  unsigned w_ = 0, h_ = 0;
  std::vector<vec4> px_;
};
#endif
