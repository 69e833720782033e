// TEST INFRASTRUCTURE (oracle shim). Pinhole camera (build-defined; the reference
// only shows the call sites mytracer_gpu.cu:141,208-209 and the public width_/height_).
// primary_ray(x, y): origin eye_, direction lower_left_ + x*x_dir_ + y*y_dir_ - eye_.
#ifndef SHIM_CAMERA_H
#define SHIM_CAMERA_H
#include "utils/Ray.h"
#include "utils/vec4.h"
struct Camera {
  vec4 eye_, x_dir_, y_dir_, lower_left_;
  int width_ = 0, height_ = 0;
  Camera() {}
  // fovy in degrees; image plane through `center`, y up along `up`
  Camera(const vec4 &eye, const vec4 &center, const vec4 &up, double fovy, int width, int height)
      : eye_(eye), width_(width), height_(height) {
    const vec4 view = normalize(center - eye);
    const double dist = norm(center - eye);
    const double img_h = 2.0 * dist * std::tan(0.5 * fovy / 180.0 * M_PI);
    const double img_w = (double(width) / double(height)) * img_h;
    const vec4 right = normalize(cross(view, up));
    const vec4 new_up = normalize(cross(right, view));
    x_dir_ = (img_w / double(width)) * right;
    y_dir_ = (img_h / double(height)) * new_up;
    lower_left_ = center - (0.5 * double(width)) * x_dir_ - (0.5 * double(height)) * y_dir_;
  }
  Ray primary_ray(double x, double y) const {
    return Ray(eye_, lower_left_ + x * x_dir_ + y * y_dir_ - eye_);
  }
};
#endif
