// rt_scene_io.cpp — the caller side of the boundary: scene / OBJ / texture ingestion into
// the reference's `Data` SoA, and the image writer.
//
// What the reference shows of this layer: Raytracer::pre_read_scene (mytracer.cpp:302-350: tokens, '#'
// comments, `mesh <file> <mode>` with the mesh path relative to the scene file), pre_read_obj
// (mytracer.cpp:424-500: v / vt / vn / mtllib / usemtl / f records, three indices per face) and build_Data
// (mytracer.cpp:166-296: per-mesh concatenation with vbase / tbase / ibase offsets).  The parser proper
// (read_scene, read_obj, Image) belongs to the absent course framework; the token set below is that
// framework's published scene format and is a stated convention of this build (DESIGN.md "Conventions"):
//
//   depth N | camera eye(3) center(3) up(3) fovy width height | background r g b | ambience r g b
//   light pos(3) color(3) | sphere center(3) radius MATERIAL | plane center(3) normal(3) MATERIAL
//   mesh file.obj FLAT|PHONG MATERIAL          MATERIAL = ambient(3) diffuse(3) specular(3) shininess mirror [shadowable 0|1]
//
// OBJ: v, vt, f (v, v/vt, v//vn, v/vt/vn; negative = relative; polygons are fanned), mtllib -> map_Kd texture
// (PPM or PNG, rt_image_io.h).  vn records are ignored: the reference recomputes normals (mymesh.cpp:103-163).
#include "../../include/rt_b200.h"
#include "rt_image_io.h"
#include "rt_math.cuh"
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <sstream>
#include <string>
#include <vector>

namespace rtb {
void set_error(const std::string &msg);
}

struct rt_scene_file {
  std::vector<int> vertexMeshId, vertexIdx, textureIdx;
  std::vector<rt_vec4> vertexPos, vertexNormals, normals, texels, ka, kd, ks;
  std::vector<double> U, V, shin, mirror;
  std::vector<int> firstVertex, vertexCount, firstVertexIdx, vertexIdxCount, firstTextCoord, textCoordCount,
      firstTextIdx, textIdxCount, texW, texH, drawMode;
  std::vector<size_t> firstMeshTex;
  std::vector<unsigned char> shadowable;
  std::vector<rt_vec4> bbmin, bbmax;
  std::vector<int> left, first, count;
  std::vector<rt_object> objects;
  std::vector<rt_vec4> lpos, lcol;
  rt_scene_file_view view;
};

namespace {
using namespace rtb;

struct MeshIn {
  std::vector<rt_vec4> pos;
  std::vector<double> u, v;
  std::vector<int> idx, uvidx; // 3 per triangle, local ids
  Image8 tex;
  bool has_tex = false;
  int mode = 0;
  rt_material mat;
};

std::string dir_of(const std::string &p) {
  const size_t k = p.find_last_of('/');
  return k == std::string::npos ? std::string() : p.substr(0, k + 1);
}

bool read_material(std::istream &is, rt_material &m) {
  double a[11];
  for (double &x : a)
    if (!(is >> x)) return false;
  m.ambient = rt_vec4{a[0], a[1], a[2], 0}; m.diffuse = rt_vec4{a[3], a[4], a[5], 0}; m.specular = rt_vec4{a[6], a[7], a[8], 0};
  m.shininess = a[9]; m.mirror = a[10]; m.shadowable = 1; m.pad_ = 0;
  // optional 12th number (extension of this build): shadowable 0/1 — Material::shadowable exists in the
  // reference (mytracer.cpp:589) but how the framework's parser sets it is not visible; default true
  is >> std::ws;
  const int c = is.peek();
  if (c == '0' || c == '1') {
    int flag = 1;
    is >> flag;
    m.shadowable = flag != 0;
  }
  return true;
}

bool load_obj(const std::string &path, MeshIn &m, std::string &err) {
  std::ifstream ifs(path);
  if (!ifs) { err = "cannot open " + path; return false; }
  std::string line;
  while (std::getline(ifs, line)) {
    if (!line.empty() && line.back() == '\r') line.pop_back();
    std::istringstream ls(line);
    std::string head;
    if (!(ls >> head) || head[0] == '#') continue;
    if (head == "v") {
      double x, y, z;
      if (!(ls >> x >> y >> z)) { err = path + ": bad v record"; return false; }
      m.pos.push_back(rt_vec4{x, y, z, 0});
    } else if (head == "vt") {
      double u, v;
      if (!(ls >> u >> v)) { err = path + ": bad vt record"; return false; }
      m.u.push_back(u); m.v.push_back(v);
    } else if (head == "f") {
      std::vector<int> vi, ti;
      std::string tok;
      while (ls >> tok) {
        int a = 0, b = 0;
        const size_t s1 = tok.find('/');
        a = std::atoi(tok.substr(0, s1).c_str());
        if (s1 != std::string::npos) {
          const size_t s2 = tok.find('/', s1 + 1);
          const std::string mid = tok.substr(s1 + 1, s2 == std::string::npos ? std::string::npos : s2 - s1 - 1);
          if (!mid.empty()) b = std::atoi(mid.c_str());
        }
        if (a == 0) { err = path + ": bad face index"; return false; }
        a = a > 0 ? a - 1 : (int)m.pos.size() + a;
        b = b > 0 ? b - 1 : (b < 0 ? (int)m.u.size() + b : 0);
        if (a < 0 || a >= (int)m.pos.size() || b < 0 || (b > 0 && b >= (int)m.u.size())) {
          err = path + ": face index out of range";
          return false;
        }
        vi.push_back(a); ti.push_back(b);
      }
      if (vi.size() < 3) { err = path + ": face with fewer than 3 vertices"; return false; }
      for (size_t k = 1; k + 1 < vi.size(); ++k) {
        m.idx.push_back(vi[0]); m.idx.push_back(vi[k]); m.idx.push_back(vi[k + 1]);
        m.uvidx.push_back(ti[0]); m.uvidx.push_back(ti[k]); m.uvidx.push_back(ti[k + 1]);
      }
    } else if (head == "mtllib") {
      std::string name;
      ls >> name;
      std::ifstream mf(dir_of(path) + name);
      std::string ml;
      while (mf && std::getline(mf, ml)) {
        if (!ml.empty() && ml.back() == '\r') ml.pop_back();
        std::istringstream ms(ml);
        std::string h2, file;
        if ((ms >> h2 >> file) && h2 == "map_Kd" && !m.has_tex) {
          if (!read_image(dir_of(path) + file, m.tex, err)) return false;
          m.has_tex = true;
        }
      }
    }
  }
  if (m.has_tex && m.u.empty()) m.has_tex = false; // a texture without coordinates cannot be sampled
  return true;
}

} // namespace

extern "C" int rt_scene_file_free(rt_scene_file *f) {
  delete f;
  return RT_OK;
}

extern "C" int rt_scene_file_load(const char *path, unsigned flags, rt_scene_file **out) {
  if (!path || !out) { set_error("rt_scene_file_load: null argument"); return RT_ERR_INVALID_ARG; }
  std::ifstream ifs(path);
  if (!ifs) { set_error(std::string("rt_scene_file_load: cannot open ") + path); return RT_ERR_INVALID_ARG; }
  const std::string base = dir_of(path);
  std::vector<MeshIn> meshes;
  rt_scene_file *f = new rt_scene_file();
  std::memset(&f->view, 0, sizeof(f->view));
  f->view.max_depth = 5;
  f->view.ambience = rt_vec4{0, 0, 0, 0};
  f->view.background = rt_vec4{0, 0, 0, 0};
  double eye[3] = {0, 0, 5}, center[3] = {0, 0, 0}, up[3] = {0, 1, 0}, fovy = 45;
  int width = 0, height = 0;
  std::string tok, err;
  auto fail = [&](const std::string &m, int code) { set_error("rt_scene_file_load: " + m); delete f; return code; };
  while (ifs >> tok) {
    if (tok[0] == '#') { std::getline(ifs, tok); continue; }
    if (tok == "depth") {
      if (!(ifs >> f->view.max_depth)) return fail("bad depth", RT_ERR_BAD_SCENE);
    } else if (tok == "camera") {
      if (!(ifs >> eye[0] >> eye[1] >> eye[2] >> center[0] >> center[1] >> center[2] >> up[0] >> up[1] >> up[2] >> fovy >>
            width >> height))
        return fail("bad camera", RT_ERR_BAD_SCENE);
    } else if (tok == "background" || tok == "ambience") {
      rt_vec4 c{0, 0, 0, 0};
      if (!(ifs >> c.x >> c.y >> c.z)) return fail("bad " + tok, RT_ERR_BAD_SCENE);
      (tok == "background" ? f->view.background : f->view.ambience) = c;
    } else if (tok == "light") {
      rt_vec4 p{0, 0, 0, 0}, c{0, 0, 0, 0};
      if (!(ifs >> p.x >> p.y >> p.z >> c.x >> c.y >> c.z)) return fail("bad light", RT_ERR_BAD_SCENE);
      f->lpos.push_back(p); f->lcol.push_back(c);
    } else if (tok == "sphere" || tok == "plane") {
      rt_object o;
      std::memset(&o, 0, sizeof(o));
      o.kind = tok == "sphere" ? RT_SPHERE : RT_PLANE;
      bool ok = (bool)(ifs >> o.center.x >> o.center.y >> o.center.z);
      if (o.kind == RT_SPHERE) ok = ok && (ifs >> o.radius);
      else {
        D3 n;
        ok = ok && (ifs >> n.x >> n.y >> n.z);
        n = normalize(n);
        o.normal = rt_vec4{n.x, n.y, n.z, 0};
      }
      if (!ok || !read_material(ifs, o.material)) return fail("bad " + tok, RT_ERR_BAD_SCENE);
      f->objects.push_back(o);
    } else if (tok == "mesh") {
      std::string file, mode;
      MeshIn m;
      if (!(ifs >> file >> mode) || !read_material(ifs, m.mat)) return fail("bad mesh record", RT_ERR_BAD_SCENE);
      if (mode != "FLAT" && mode != "PHONG") return fail("mesh draw mode must be FLAT or PHONG", RT_ERR_BAD_SCENE);
      m.mode = mode == "PHONG" ? 1 : 0;
      if (!load_obj(base + file, m, err)) return fail(err, RT_ERR_BAD_SCENE);
      meshes.push_back(std::move(m));
    } else {
      return fail("unknown token '" + tok + "'", RT_ERR_BAD_SCENE);
    }
  }
  if (width <= 0 || height <= 0) return fail("no camera record", RT_ERR_BAD_SCENE);
  int st = rt_camera_look_at(eye, center, up, fovy, width, height, &f->view.camera);
  if (st != RT_OK) { delete f; return st; }

  // ---- build_Data (mytracer.cpp:166-296) ---------------------------------------------------------------
  const int M = (int)meshes.size();
  int vbase = 0, tbase = 0, ibase = 0;
  size_t texoff = 0;
  for (int mi = 0; mi < M; ++mi) {
    MeshIn &m = meshes[mi];
    const int nv = (int)m.pos.size(), nt = (int)m.idx.size() / 3, nuv = m.has_tex ? (int)m.u.size() : 0;
    std::vector<rt_vec4> fn(nt), vn(nv);
    st = (flags & RT_LOAD_DEVICE_NORMALS)
             ? rt_compute_normals_device(m.pos.data(), nv, m.idx.data(), nt, fn.data(), vn.data(), nullptr)
             : rt_compute_normals(m.pos.data(), nv, m.idx.data(), nt, fn.data(), vn.data());
    if (st != RT_OK) { delete f; return st; }
    f->vertexPos.insert(f->vertexPos.end(), m.pos.begin(), m.pos.end());
    f->vertexNormals.insert(f->vertexNormals.end(), vn.begin(), vn.end());
    f->vertexMeshId.insert(f->vertexMeshId.end(), nv, mi);
    f->firstVertex.push_back(vbase); f->vertexCount.push_back(nv);
    if (nuv) { f->U.insert(f->U.end(), m.u.begin(), m.u.end()); f->V.insert(f->V.end(), m.v.begin(), m.v.end()); }
    f->firstTextCoord.push_back(tbase); f->textCoordCount.push_back(nuv);
    for (int k = 0; k < 3 * nt; ++k) {
      f->vertexIdx.push_back(vbase + m.idx[k]);
      f->textureIdx.push_back(tbase + (nuv ? m.uvidx[k] : 0));
    }
    f->normals.insert(f->normals.end(), fn.begin(), fn.end());
    f->firstVertexIdx.push_back(ibase); f->vertexIdxCount.push_back(3 * nt);
    f->firstTextIdx.push_back(ibase); f->textIdxCount.push_back(3 * nt);
    if (m.has_tex) {
      f->texW.push_back(m.tex.w); f->texH.push_back(m.tex.h); f->firstMeshTex.push_back(texoff);
      for (size_t p = 0; p < (size_t)m.tex.w * m.tex.h; ++p)
        f->texels.push_back(rt_vec4{m.tex.rgba[4 * p] / 255.0, m.tex.rgba[4 * p + 1] / 255.0, m.tex.rgba[4 * p + 2] / 255.0, 0});
      texoff += (size_t)m.tex.w * m.tex.h;
    } else {
      f->texW.push_back(-1); f->texH.push_back(-1); f->firstMeshTex.push_back((size_t)-1);
    }
    f->drawMode.push_back(m.mode);
    f->ka.push_back(m.mat.ambient); f->kd.push_back(m.mat.diffuse); f->ks.push_back(m.mat.specular);
    f->shin.push_back(m.mat.shininess); f->mirror.push_back(m.mat.mirror); f->shadowable.push_back(m.mat.shadowable ? 1 : 0);
    vbase += nv; ibase += 3 * nt; tbase += nuv;
  }
  rt_data_soa &d = f->view.data;
  d.tMeshCount_ = M; d.tVertexCount_ = vbase; d.tVertexIdxCount_ = ibase; d.tTextCoordCount_ = tbase; d.tTextIdxCount_ = ibase;
  d.tTexelCount_ = texoff;
  d.vertexMeshId_ = f->vertexMeshId.data(); d.vertexPos_ = f->vertexPos.data(); d.vertexNormals_ = f->vertexNormals.data();
  d.normals_ = f->normals.data(); d.vertexIdx_ = f->vertexIdx.data(); d.textureIdx_ = f->textureIdx.data();
  d.textureCoordinatesU_ = f->U.data(); d.textureCoordinatesV_ = f->V.data();
  d.firstVertex_ = f->firstVertex.data(); d.vertexCount_ = f->vertexCount.data();
  d.firstVertexIdx_ = f->firstVertexIdx.data(); d.vertexIdxCount_ = f->vertexIdxCount.data();
  d.firstTextCoord_ = f->firstTextCoord.data(); d.textCoordCount_ = f->textCoordCount.data();
  d.firstTextIdx_ = f->firstTextIdx.data(); d.textIdxCount_ = f->textIdxCount.data();
  d.meshTexels_ = f->texels.data(); d.meshTexWidth_ = f->texW.data(); d.meshTexHeight_ = f->texH.data();
  d.firstMeshTex_ = f->firstMeshTex.data(); d.meshDrawMode_ = f->drawMode.data();
  d.materialAmbient_ = f->ka.data(); d.materialDiffuse_ = f->kd.data(); d.materialSpecular_ = f->ks.data();
  d.materialShininess_ = f->shin.data(); d.materialMirror_ = f->mirror.data(); d.materialShadowable_ = f->shadowable.data();

  // ---- BVH::initSoA (mybvh.cpp:375-406) ------------------------------------------------------------------
  const int N = ibase / 3;
  const int cap = N > 0 ? 2 * N - 1 : 1;
  f->bbmin.assign(cap, rt_vec4{0, 0, 0, 0}); f->bbmax.assign(cap, rt_vec4{0, 0, 0, 0});
  f->left.assign(cap, 0); f->first.assign(cap, 0); f->count.assign(cap, 0);
  rt_bvh_nodes_soa &b = f->view.bvh;
  b.bb_min_ = f->bbmin.data(); b.bb_max_ = f->bbmax.data(); b.leftChildIdx_ = f->left.data();
  b.firstTriIdx_ = f->first.data(); b.triCount_ = f->count.data();
  f->view.nodes_used = 0;
  if (N > 0) {
    st = (flags & RT_LOAD_DEVICE_BVH) ? rt_bvh_build_median_device(&d, &b, &f->view.nodes_used, nullptr)
                                      : rt_bvh_build_median(&d, &b, &f->view.nodes_used);
    if (st != RT_OK) { delete f; return st; }
  }
  f->view.objects = f->objects.data(); f->view.num_objects = (int)f->objects.size();
  f->view.lights_pos = f->lpos.data(); f->view.lights_color = f->lcol.data(); f->view.num_lights = (int)f->lpos.size();
  *out = f;
  return RT_OK;
}

extern "C" int rt_scene_file_get(const rt_scene_file *f, rt_scene_file_view *out) {
  if (!f || !out) { set_error("rt_scene_file_get: null argument"); return RT_ERR_INVALID_ARG; }
  *out = f->view;
  return RT_OK;
}

// Frame (index y*W + x, y = 0 at the BOTTOM: lower_left + y*y_dir) -> upright image file, by extension.
extern "C" int rt_image_write(const char *path, const uint8_t *rgba8, int width, int height) {
  if (!path || !rgba8 || width <= 0 || height <= 0) { set_error("rt_image_write: invalid argument"); return RT_ERR_INVALID_ARG; }
  std::vector<uint8_t> top((size_t)width * height * 4), bytes;
  for (int y = 0; y < height; ++y)
    std::memcpy(top.data() + (size_t)y * width * 4, rgba8 + (size_t)(height - 1 - y) * width * 4, (size_t)width * 4);
  const std::string p(path);
  const bool png = p.size() >= 4 && (p.compare(p.size() - 4, 4, ".png") == 0 || p.compare(p.size() - 4, 4, ".PNG") == 0);
  if (png) encode_png(top.data(), width, height, bytes);
  else encode_ppm(top.data(), width, height, bytes);
  FILE *fp = std::fopen(path, "wb");
  if (!fp) { set_error(std::string("rt_image_write: cannot open ") + path); return RT_ERR_INVALID_ARG; }
  const size_t n = std::fwrite(bytes.data(), 1, bytes.size(), fp);
  std::fclose(fp);
  if (n != bytes.size()) { set_error("rt_image_write: short write"); return RT_ERR_INVALID_ARG; }
  return RT_OK;
}

// PPM (P6) or PNG file -> RGBA8, row 0 = top row of the file.  Release with rt_image_free.
extern "C" int rt_image_read(const char *path, int *width, int *height, uint8_t **rgba8) {
  if (!path || !width || !height || !rgba8) { set_error("rt_image_read: null argument"); return RT_ERR_INVALID_ARG; }
  Image8 img;
  std::string err;
  if (!read_image(path, img, err)) { set_error("rt_image_read: " + err); return RT_ERR_BAD_SCENE; }
  *width = img.w; *height = img.h;
  *rgba8 = (uint8_t *)std::malloc(img.rgba.size());
  if (!*rgba8) return RT_ERR_OUT_OF_MEMORY;
  std::memcpy(*rgba8, img.rgba.data(), img.rgba.size());
  return RT_OK;
}

extern "C" int rt_image_free(uint8_t *rgba8) {
  std::free(rgba8);
  return RT_OK;
}
