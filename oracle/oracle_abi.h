/* TEST INFRASTRUCTURE — C ABI shared by the two oracle libraries:
 *   oracle/_ref/libref_oracle.so  = the reference's own CPU sources (compiled by path from
 *                                   /root/reference) + shim + ref_glue.cpp
 *   oracle/librt_oracle.so        = the independent plain-C++ restatement (rt_oracle.cpp)
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
 * may load these.  The product (my-raytracer_b200/) never does.
 *
 * All vectors are 3 packed doubles unless stated.  Images are row-major, index y*W + x.
 * material = 12 doubles: ambient[3] diffuse[3] specular[3] shininess mirror shadowable(0/1).
 */
#ifndef ORACLE_ABI_H
#define ORACLE_ABI_H
#ifdef __cplusplus
extern "C" {
#endif

void *orc_new(void);
void orc_free(void *h);
const char *orc_kind(void); /* "reference" or "port" */

/* Scene assembly (what the absent read_scene would do).  uvidx/u/v/texels may be NULL.
 * texels: texW*texH*3 doubles, row-major y*texW+x.  draw_mode 0 FLAT, 1 PHONG. */
int orc_add_mesh(void *h, int nv, const double *pos, int nt, const int *idx, const int *uvidx, int nuv,
                 const double *u, const double *v, int texW, int texH, const double *texels,
                 const double *material, int draw_mode);
int orc_add_sphere(void *h, const double *center, double radius, const double *material);
int orc_add_plane(void *h, const double *center, const double *normal, const double *material);
int orc_add_light(void *h, const double *pos, const double *color);
void orc_set_camera(void *h, const double *eye, const double *center, const double *up, double fovy,
                    int width, int height);
/* out: eye, x_dir, y_dir, lower_left (12 doubles) */
void orc_get_camera(void *h, double *out12);
void orc_set_globals(void *h, const double *background, const double *ambience, int max_depth);
/* compute_normals() on every mesh, then the median-split BVH build */
void orc_build(void *h);

/* Introspection of the built structures */
int orc_num_triangles(void *h);
int orc_num_nodes(void *h);
void orc_get_bvh(void *h, double *bbmin3, double *bbmax3, int *left, int *first, int *count);
/* perm[i] = original (mesh-order) global triangle index of the triangle at BVH position i */
void orc_get_tri_order(void *h, int *perm);
void orc_get_normals(void *h, int mesh, double *face_normals3, double *vertex_normals3);

/* Rendering.  rgb: W*H*3 doubles after min(c,1).  ids/t may be NULL.  ids: BVH-order triangle
 * index, N + object index for analytic primitives, -1 for a miss.  Rows y0 <= y < y1 with
 * stride ystep are rendered (others untouched).  spp_grid s > 1 averages an s x s regular
 * sub-pixel grid (offset formula of mytracer_gpu.cu:203-207) before the clamp; ids/t then
 * describe the centre ray.  Returns seconds spent in the pixel loop. */
double orc_render(void *h, double *rgb, int *ids, double *t, int y0, int y1, int ystep, int spp_grid,
                  int nthreads);
/* rays cast by the last orc_render: primary, shadow, reflection */
void orc_ray_counts(void *h, long long *out3);
/* arbitrary rays: origins/dirs n*3 (dirs are normalised by Ray()); rgb n*3 (unclamped trace()),
 * ids n, t n */
void orc_trace_rays(void *h, int n, const double *origins, const double *dirs, double *rgb, int *ids,
                    double *t);

/* Known-answer kernels (one primitive, one ray). Each returns hit (0/1). */
/* out: t, alpha, beta, gamma(=1-alpha-beta) are not exposed by the reference; we return t, px,py,pz */
int orc_tri_test(const double *p0, const double *p1, const double *p2, const double *o, const double *d,
                 double *out4);
int orc_aabb_test(const double *bbmin, const double *bbmax, const double *o, const double *d,
                  double *tmin_out);
int orc_plane_test(const double *center, const double *normal, const double *o, const double *d,
                   double *out4);
int orc_sphere_test(const double *center, double radius, const double *o, const double *d, double *out7);
/* lighting() at one surface point against the scene's lights; material 12 doubles; out rgb */
void orc_lighting(void *h, const double *point, const double *normal, const double *view,
                  const double *material, double *rgb);
/* nearest-texel lookup of mesh `mesh` for barycentric (alpha,beta,gamma) of BVH-order triangle tri */
void orc_texture(void *h, int tri, double alpha, double beta, double gamma, double *rgb);

#ifdef __cplusplus
}
#endif
#endif
