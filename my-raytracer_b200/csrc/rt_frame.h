// rt_frame.h — C-ABI PODs -> kernel parameter records (shared by rt_api.cu and the test build).
#ifndef RTB_FRAME_H
#define RTB_FRAME_H
#include "../../include/rt_b200.h"
#include "rt_types.h"
#include <cstring>
#include <string>

namespace rtb {

inline int make_frame_params(const rt_camera *cam, int num_lights, const rt_render_params *p,
                             const rt_frame_out *out, FrameParams &fp, std::string &err) {
  if (!cam || !p || !out) { err = "render: null camera/params/out"; return RT_ERR_INVALID_ARG; }
  if (cam->width <= 0 || cam->height <= 0) { err = "render: camera width/height must be positive"; return RT_ERR_INVALID_ARG; }
  if ((long long)cam->width * cam->height > 0x7fffffffLL / 64) { err = "render: image too large"; return RT_ERR_LIMIT; }
  if (num_lights < 0) { err = "render: negative light count"; return RT_ERR_INVALID_ARG; }
  if (p->max_depth < 0) { err = "render: negative max_depth"; return RT_ERR_INVALID_ARG; }
  const int s = p->spp_grid < 1 ? 1 : p->spp_grid;
  if (s > 16) { err = "render: spp_grid > 16"; return RT_ERR_LIMIT; }
  if ((out->hit_id || out->hit_t) && s != 1) { err = "render: hit_id/hit_t need spp_grid == 1"; return RT_ERR_INVALID_ARG; }
  const int world = p->tile_world < 1 ? 1 : p->tile_world;
  const int chunk = p->tile_chunk < 1 ? 1 : p->tile_chunk;
  if (p->tile_rank < 0 || p->tile_rank >= world) { err = "render: tile_rank outside [0, tile_world)"; return RT_ERR_INVALID_ARG; }
  std::memset(&fp, 0, sizeof(fp));
  const rt_vec4 *src[4] = {&cam->eye, &cam->x_dir, &cam->y_dir, &cam->lower_left};
  double *dst[4] = {fp.cam.eye, fp.cam.x_dir, fp.cam.y_dir, fp.cam.lower_left};
  for (int k = 0; k < 4; ++k) { dst[k][0] = src[k]->x; dst[k][1] = src[k]->y; dst[k][2] = src[k]->z; }
  fp.cam.width = cam->width; fp.cam.height = cam->height;
  fp.background[0] = p->background.x; fp.background[1] = p->background.y; fp.background[2] = p->background.z;
  fp.ambience[0] = p->ambience.x; fp.ambience[1] = p->ambience.y; fp.ambience[2] = p->ambience.z;
  fp.max_depth = p->max_depth;
  fp.spp_grid = s;
  fp.num_lights = num_lights;
  fp.tile_rank = p->tile_rank; fp.tile_world = world; fp.tile_chunk = chunk;
  fp.tiles_x = (cam->width + kTileW - 1) / kTileW;
  fp.tiles_y = (cam->height + kTileH - 1) / kTileH;
  fp.flags = p->flags;
  const long long ntiles = (long long)fp.tiles_x * fp.tiles_y;
  if (p->dyn_tiles < 0 || p->dyn_chunk < 0) { err = "render: negative dyn_tiles/dyn_chunk"; return RT_ERR_INVALID_ARG; }
  if (p->dyn_tiles > 0 && (!p->dyn_counter || p->dyn_chunk < 1)) { err = "render: dyn_tiles needs dyn_counter and dyn_chunk >= 1"; return RT_ERR_INVALID_ARG; }
  const long long dyn = p->dyn_tiles < ntiles ? p->dyn_tiles : ntiles;
  fp.static_tiles = (int)(ntiles - dyn);
  fp.dyn_chunk = p->dyn_chunk > 0 ? p->dyn_chunk : 1;
  return RT_OK;
}

} // namespace rtb
#endif
