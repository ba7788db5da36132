/*
 * oracle.c - TEST INFRASTRUCTURE ONLY.  See oracle.h.
 *
 * Plain-C restatement of the reference hot path.  Build: oracle/Makefile
 * (gcc -O2 -ffp-contract=off; no -march=native so no FMA is ever formed).
 * Citations are relative to /root/reference.
 */
#define _GNU_SOURCE
#include "oracle.h"

#include <math.h>
#include <pthread.h>
#include <stdlib.h>
#include <string.h>

/* ---- Eigen 3.3.7 fixed-size float semantics -------------------------------------------- */
typedef struct { float x, y, z; } v3;

static inline v3 V(float x, float y, float z) { v3 r = {x, y, z}; return r; }
static inline v3 vadd(v3 a, v3 b) { return V(a.x + b.x, a.y + b.y, a.z + b.z); }
static inline v3 vsub(v3 a, v3 b) { return V(a.x - b.x, a.y - b.y, a.z - b.z); }
static inline v3 vmuls(v3 a, float s) { return V(a.x * s, a.y * s, a.z * s); }
static inline v3 vdivs(v3 a, float s) { return V(a.x / s, a.y / s, a.z / s); }
static inline v3 vneg(v3 a) { return V(-a.x, -a.y, -a.z); }
/* 3-term redux = a0 + (a1 + a2): Eigen/src/Core/Redux.h:91-105 (HalfLength split) */
static inline float vdot(v3 a, v3 b) { return a.x * b.x + (a.y * b.y + a.z * b.z); }
/* Eigen/src/Geometry/OrthoMethods.h:43-47 */
static inline v3 vcross(v3 a, v3 b) {
  return V(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
static inline float vnorm(v3 a) { return sqrtf(vdot(a, a)); }
/* Eigen/src/Core/Dot.h:124-134: identity when squaredNorm is not > 0 */
static inline v3 vnormalized(v3 a) {
  float z = vdot(a, a);
  if (z > 0.f) return vdivs(a, sqrtf(z));
  return a;
}
/* Affine3f * Vector3f (Eigen/src/Geometry/Transform.h:1372-1392): the vector is extended to
 * (v,1) and multiplied by the 4x4 matrix; with SSE2 that product is evaluated column by
 * column, res = ((col0*x + col1*y) + col2*z) + col3*1 (Eigen/src/Core/ProductEvaluators.h,
 * etor_product_packet_impl<ColMajor>), i.e. sequential with the translation added LAST.
 * m: 3x4 row-major. */
static inline v3 affine_mul(const float *m, v3 p) {
  return V(((m[0] * p.x + m[1] * p.y) + m[2] * p.z) + m[3] * 1.0f,
           ((m[4] * p.x + m[5] * p.y) + m[6] * p.z) + m[7] * 1.0f,
           ((m[8] * p.x + m[9] * p.y) + m[10] * p.z) + m[11] * 1.0f);
}
static inline v3 linear_mul(const float *m, v3 p) { /* 3x3 block of a 3x4 */
  return V(m[0] * p.x + (m[1] * p.y + m[2] * p.z), m[4] * p.x + (m[5] * p.y + m[6] * p.z),
           m[8] * p.x + (m[9] * p.y + m[10] * p.z));
}
/* Matrix3f inverse by cofactors: Eigen/src/LU/InverseImpl.h:120-170.  a, inv: 3x3 row-major
 * with row stride `lda` (4 for the linear part of a 3x4). */
static inline float cof(const float *a, int lda, int i, int j) {
  int i1 = (i + 1) % 3, i2 = (i + 2) % 3, j1 = (j + 1) % 3, j2 = (j + 2) % 3;
  return a[i1 * lda + j1] * a[i2 * lda + j2] - a[i1 * lda + j2] * a[i2 * lda + j1];
}
static void inverse3(const float *a, int lda, float *inv, int ldi) {
  float c0 = cof(a, lda, 0, 0), c1 = cof(a, lda, 1, 0), c2 = cof(a, lda, 2, 0);
  float det = c0 * a[0] + (c1 * a[lda] + c2 * a[2 * lda]);
  float invdet = 1.0f / det;
  inv[0] = c0 * invdet; inv[1] = c1 * invdet; inv[2] = c2 * invdet;
  inv[ldi + 0] = cof(a, lda, 0, 1) * invdet;
  inv[ldi + 1] = cof(a, lda, 1, 1) * invdet;
  inv[ldi + 2] = cof(a, lda, 2, 1) * invdet;
  inv[2 * ldi + 0] = cof(a, lda, 0, 2) * invdet;
  inv[2 * ldi + 1] = cof(a, lda, 1, 2) * invdet;
  inv[2 * ldi + 2] = cof(a, lda, 2, 2) * invdet;
}
/* Affine3f::inverse(): Eigen/src/Geometry/Transform.h:1215-1226; translation = (-Linv) * t */
static void affine_inverse(const float *m, float *inv) {
  inverse3(m, 4, inv, 4);
  float n[9];
  for (int r = 0; r < 3; ++r)
    for (int c = 0; c < 3; ++c) n[r * 3 + c] = -inv[r * 4 + c];
  for (int r = 0; r < 3; ++r)
    inv[r * 4 + 3] = n[r * 3] * m[3] + (n[r * 3 + 1] * m[7] + n[r * 3 + 2] * m[11]);
}
static inline float fminf_std(float a, float b) { return (b < a) ? b : a; } /* std::min */
static inline float fmaxf_std(float a, float b) { return (a < b) ? b : a; } /* std::max */

static inline v3 ld3(const float *p) { return V(p[0], p[1], p[2]); }
/* normalized() of a *dynamic-size* expression (`vertices[i].head(3) - ...`, mesh.hpp:475-476,
 * objimporter.hpp:94-95): the redux is the sequential loop (x*x + y*y) + z*z
 * (Eigen/src/Core/Redux.h:183-200), not the unrolled a0 + (a1 + a2). */
static inline v3 vnormalized_dyn(v3 a) {
  float z = (a.x * a.x + a.y * a.y) + a.z * a.z;
  if (z > 0.f) return vdivs(a, sqrtf(z));
  return a;
}

/* ---- mesh preparation --------------------------------------------------------------- */
void oracle_face_normals(const float *verts, const int *faces, int n_faces, float *out) {
  for (int f = 0; f < n_faces; ++f) { /* mesh.hpp:475-477 */
    v3 a = ld3(verts + 4 * faces[3 * f]), b = ld3(verts + 4 * faces[3 * f + 1]),
       c = ld3(verts + 4 * faces[3 * f + 2]);
    v3 v1 = vnormalized_dyn(vsub(c, a));
    v3 v0 = vnormalized_dyn(vsub(b, a));
    v3 n = vnormalized(vcross(v0, v1));
    out[3 * f] = n.x; out[3 * f + 1] = n.y; out[3 * f + 2] = n.z;
  }
}

void oracle_compute_normals(const float *verts, int n_verts, const int *faces, int n_faces,
                            float *out) {
  memset(out, 0, sizeof(float) * 3 * (size_t)n_verts); /* objimporter.hpp:84-85 */
  for (int f = 0; f < n_faces; ++f) {                  /* objimporter.hpp:91-100 */
    v3 a = ld3(verts + 4 * faces[3 * f]), b = ld3(verts + 4 * faces[3 * f + 1]),
       c = ld3(verts + 4 * faces[3 * f + 2]);
    v3 v0 = vnormalized_dyn(vsub(b, a));
    v3 v1 = vnormalized_dyn(vsub(c, a));
    v3 n = vnormalized(vcross(v0, v1));
    for (int k = 0; k < 3; ++k) {
      float *o = out + 3 * faces[3 * f + k];
      o[0] += n.x; o[1] += n.y; o[2] += n.z;
    }
  }
  for (int i = 0; i < n_verts; ++i) { /* objimporter.hpp:104-105 */
    v3 n = vnormalized(ld3(out + 3 * i));
    out[3 * i] = n.x; out[3 * i + 1] = n.y; out[3 * i + 2] = n.z;
  }
}

void oracle_shape_matrix(const float *verts, int n_verts, float *m) {
  /* mesh.hpp:611-628 */
  float xMax = verts[0], xMin = verts[0], yMax = verts[1], yMin = verts[1], zMax = verts[2],
        zMin = verts[2];
  for (int i = 0; i < n_verts; ++i) {
    const float *v = verts + 4 * i;
    xMax = fmaxf_std(v[0], xMax); yMax = fmaxf_std(v[1], yMax); zMax = fmaxf_std(v[2], zMax);
    xMin = fminf_std(v[0], xMin); yMin = fminf_std(v[1], yMin); zMin = fminf_std(v[2], zMin);
  }
  /* 1.0 / float -> double -> float */
  float ext = fmaxf_std(fmaxf_std(fabsf(xMax - xMin), fabsf(yMax - yMin)), fabsf(zMax - zMin));
  float scale = (float)(1.0 / (double)ext);
  float cx = (float)((double)(xMax + xMin) / 2.0), cy = (float)((double)(yMax + yMin) / 2.0),
        cz = (float)((double)(zMax + zMin) / 2.0);
  /* model.hpp:169-173: shape.scale(s) then shape.translate(-c) on an identity Affine3f:
   * linear = I * s; translation += linear * (-c) */
  float lin[9] = {scale, 0, 0, 0, scale, 0, 0, 0, scale};
  /* Identity * s : each coefficient 1*s or 0*s */
  for (int k = 0; k < 9; ++k) lin[k] = (k % 4 == 0 ? 1.0f : 0.0f) * scale;
  v3 t = V(-cx, -cy, -cz);
  for (int r = 0; r < 3; ++r) {
    m[r * 4 + 0] = lin[r * 3]; m[r * 4 + 1] = lin[r * 3 + 1]; m[r * 4 + 2] = lin[r * 3 + 2];
    m[r * 4 + 3] = 0.0f + (lin[r * 3] * t.x + (lin[r * 3 + 1] * t.y + lin[r * 3 + 2] * t.z));
  }
}

/* ---- BoundingBox build ---------------------------------------------------------------- */
typedef struct {
  float low[3], high[3], shape[3];
  int failed[3];
  int *faces;
  int n;
} obox;

static void ob_reshape(obox *b) { /* BoundingBox.cpp:18-24 */
  for (int k = 0; k < 3; ++k) b->shape[k] = b->high[k] - b->low[k];
}
static int ob_has_face(const obox *b, const float *verts, const int *f3) { /* :26-39 */
  for (int k = 0; k < 3; ++k) {
    const float *v = verts + 4 * f3[k];
    if (!(v[0] >= b->low[0] && v[0] <= b->high[0] && v[1] >= b->low[1] && v[1] <= b->high[1] &&
          v[2] >= b->low[2] && v[2] <= b->high[2]))
      return 0;
  }
  return 1;
}
static void ob_fit_faces(obox *b, const float *verts, const int *faces) { /* :48-90 */
  if (b->n <= 0) return;
  const float *t = verts + 4 * faces[3 * b->faces[0]];
  float mn[3] = {t[0], t[1], t[2]}, mx[3] = {t[0], t[1], t[2]};
  for (int i = 0; i < b->n; ++i)
    for (int j = 0; j < 3; ++j) {
      const float *v = verts + 4 * faces[3 * b->faces[i] + j];
      for (int k = 0; k < 3; ++k) {
        if (v[k] < mn[k]) mn[k] = v[k];
        else if (v[k] > mx[k]) mx[k] = v[k];
      }
    }
  for (int k = 0; k < 3; ++k) { b->low[k] = mn[k]; b->high[k] = mx[k]; }
  ob_reshape(b);
}
static float ob_average(const obox *b, const float *verts, const int *faces, int axis) {
  float average = 0; /* :151-161, FP32 accumulation, divide by (float)(size*3) */
  for (int i = 0; i < b->n; ++i) {
    const int *f = faces + 3 * b->faces[i];
    average += verts[4 * f[0] + axis];
    average += verts[4 * f[1] + axis];
    average += verts[4 * f[2] + axis];
  }
  average /= (float)((size_t)b->n * 3);
  return average;
}

typedef struct { obox **v; int n, cap; } oboxlist;
static obox *ob_new(oboxlist *l) {
  obox *b = (obox *)calloc(1, sizeof(obox));
  if (l->n == l->cap) {
    l->cap = l->cap ? l->cap * 2 : 64;
    l->v = (obox **)realloc(l->v, sizeof(obox *) * (size_t)l->cap);
  }
  l->v[l->n++] = b;
  return b;
}

/* BoundingBox::splitBox, :109-149. Returns b itself (axis failed), NULL (all failed) or the
 * new box. */
static obox *ob_split(obox *b, oboxlist *l, const float *verts, const int *faces) {
  float oldLow[3], oldHigh[3];
  memcpy(oldLow, b->low, sizeof oldLow);
  memcpy(oldHigh, b->high, sizeof oldHigh);
  float width = b->shape[0], height = b->shape[1], depth = b->shape[2];
  int choice;
  if ((width >= height || b->failed[1]) && (width >= depth || b->failed[2]) && !b->failed[0]) choice = 0;
  else if ((height >= width || b->failed[0]) && (height >= depth || b->failed[2]) && !b->failed[1]) choice = 1;
  else if (!(b->failed[0] && b->failed[1] && b->failed[2])) choice = 2;
  else return NULL;

  b->high[choice] = ob_average(b, verts, faces, choice);
  ob_reshape(b);
  /* outsideFaces(), :92-107 */
  int *inside = (int *)malloc(sizeof(int) * (size_t)(b->n ? b->n : 1));
  int *outside = (int *)malloc(sizeof(int) * (size_t)(b->n ? b->n : 1));
  int ni = 0, no = 0;
  for (int i = 0; i < b->n; ++i) {
    if (!ob_has_face(b, verts, faces + 3 * b->faces[i])) outside[no++] = b->faces[i];
    else inside[ni++] = b->faces[i];
  }
  free(b->faces);
  b->faces = inside;
  b->n = ni;
  int impossible = (ni == 0 || no == 0);
  if (impossible) {
    if (ni == 0) { free(b->faces); b->faces = outside; b->n = no; }
    else free(outside);
    b->failed[choice] = 1;
    memcpy(b->low, oldLow, sizeof oldLow);
    memcpy(b->high, oldHigh, sizeof oldHigh);
    ob_reshape(b);
    return b;
  }
  b->failed[0] = b->failed[1] = b->failed[2] = 0;
  obox *nb = ob_new(l);
  nb->faces = outside;
  nb->n = no;
  ob_fit_faces(nb, verts, faces);
  ob_fit_faces(b, verts, faces);
  return nb;
}

int oracle_build_boxes(const float *verts, const int *faces, int n_faces, int min_faces,
                       int max_boxes, float **boxes, int **box_off, int **box_faces) {
  oboxlist l = {0, 0, 0};
  obox *root = ob_new(&l); /* flyscene.cpp:401-402, BoundingBox.cpp:41-46 */
  root->faces = (int *)malloc(sizeof(int) * (size_t)(n_faces ? n_faces : 1));
  root->n = n_faces;
  for (int i = 0; i < n_faces; ++i) root->faces[i] = i;
  ob_fit_faces(root, verts, faces);
  int notDone = 1;
  while (notDone && l.n < max_boxes) { /* flyscene.cpp:405-418 */
    notDone = 0;
    int cur = l.n;
    for (int i = 0; i < cur; ++i) {
      obox *b = l.v[i];
      if (b->n > min_faces && (!b->failed[0] || !b->failed[1] || !b->failed[2])) {
        obox *nb = ob_split(b, &l, verts, faces);
        while (nb == b) nb = ob_split(b, &l, verts, faces);
        notDone = 1;
      }
    }
  }
  *boxes = (float *)malloc(sizeof(float) * 6 * (size_t)l.n);
  *box_off = (int *)malloc(sizeof(int) * (size_t)(l.n + 1));
  *box_faces = (int *)malloc(sizeof(int) * (size_t)(n_faces ? n_faces : 1));
  int off = 0;
  for (int i = 0; i < l.n; ++i) {
    memcpy(*boxes + 6 * i, l.v[i]->low, 12);
    memcpy(*boxes + 6 * i + 3, l.v[i]->high, 12);
    (*box_off)[i] = off;
    memcpy(*box_faces + off, l.v[i]->faces, sizeof(int) * (size_t)l.v[i]->n);
    off += l.v[i]->n;
    free(l.v[i]->faces);
    free(l.v[i]);
  }
  (*box_off)[l.n] = off;
  int n = l.n;
  free(l.v);
  return n;
}

void oracle_free(void *p) { free(p); }

/* ---- camera ------------------------------------------------------------------------- */
void oracle_camera_derive(const oracle_camera *cam, float eye[3], float inv_view[12]) {
  /* camera.hpp:115-118: view.linear().inverse() * (-view.translation()) */
  float li[9];
  inverse3(cam->view, 4, li, 3);
  v3 nt = V(-cam->view[3], -cam->view[7], -cam->view[11]);
  eye[0] = li[0] * nt.x + (li[1] * nt.y + li[2] * nt.z);
  eye[1] = li[3] * nt.x + (li[4] * nt.y + li[5] * nt.z);
  eye[2] = li[6] * nt.x + (li[7] * nt.y + li[8] * nt.z);
  affine_inverse(cam->view, inv_view);
}

static v3 screen_to_world(const oracle_camera *cam, const float *inv_view, float ri, float rj) {
  /* camera.hpp:155-173; NDC in double, then float */
  const float *vp = cam->viewport;
  v3 nc;
  nc.x = (float)(2.0 * (double)(ri - vp[0]) / (double)vp[2] - 1.0);
  nc.y = (float)(1.0 - 2.0 * (double)(rj - vp[1]) / (double)vp[3]);
  nc.z = (float)-1.0;
  /* camera.hpp:263-266: (float)1.0f/tan((fovy/2.0f)*(M_PI/180.0f)) -> float */
  float pscale = (float)((double)1.0f / tan((double)(cam->fovy / 2.0f) * (M_PI / (double)180.0f)));
  float scale = (float)(1.0 / (double)pscale);
  nc.x *= cam->aspect * scale;
  nc.y *= scale;
  return affine_mul(inv_view, nc);
}

void oracle_primary_dir(const oracle_camera *cam, int i, int j, float dir[3]) {
  float eye[3], inv[12];
  oracle_camera_derive(cam, eye, inv);
  v3 w = screen_to_world(cam, inv, (float)i, (float)j);
  v3 d = vnormalized(vsub(w, V(eye[0], eye[1], eye[2])));
  dir[0] = d.x; dir[1] = d.y; dir[2] = d.z;
}

/* ---- the tracer ----------------------------------------------------------------------- */
/* Mutable "member" state of Flyscene that leaks between rays (flyscene.hpp:179-184, Q4/Q5) */
typedef struct {
  const oracle_scene *s;
  const oracle_lights *l;
  int max_depth;
  v3 background;
  v3 ka, kd, ks;
  float shininess;
  float minv[12]; /* getShapeModelMatrix().inverse(), constant per scene */
  oracle_counts c;
} tracer;

static void tracer_init(tracer *t, const oracle_scene *s, const oracle_lights *l, int max_depth,
                        const float *bg) {
  memset(t, 0, sizeof *t);
  t->s = s; t->l = l; t->max_depth = max_depth;
  t->background = bg ? ld3(bg) : V((float)0.9, (float)0.9, (float)0.9);
  t->ka = V((float)0.2, (float)0.2, (float)0.2); /* flyscene.hpp:179-182 */
  t->kd = V((float)0.9, (float)0.9, 0.f);
  t->ks = V(0.f, 0.f, 0.f);
  t->shininess = 0.f;
  affine_inverse(s->shape, t->minv);
}

/* interpolateNormal, flyscene.cpp:572-600 */
static v3 interpolate_normal(const tracer *t, int f, v3 P) {
  const oracle_scene *s = t->s;
  const int *id = s->faces + 3 * f;
  v3 vert0 = affine_mul(s->shape, ld3(s->verts + 4 * id[0]));
  v3 vert1 = affine_mul(s->shape, ld3(s->verts + 4 * id[1]));
  v3 vert2 = affine_mul(s->shape, ld3(s->verts + 4 * id[2]));
  v3 facenormal = vnormalized(ld3(s->face_normal + 3 * f));
  v3 edge0 = vsub(vert1, vert0), edge1 = vsub(vert2, vert1), edge2 = vsub(vert0, vert2);
  v3 in0 = vsub(P, vert0), in1 = vsub(P, vert1), in2 = vsub(P, vert2);
  v3 a0v = vcross(edge0, in0), a1v = vcross(edge1, in1), a2v = vcross(edge2, in2);
  if (vdot(facenormal, a0v) < 0 || vdot(facenormal, a1v) < 0 || vdot(facenormal, a2v) < 0)
    return V(0, 0, 0);
  float area0 = vnorm(a0v) / 2, area1 = vnorm(a1v) / 2, area2 = vnorm(a2v) / 2;
  float area = vnorm(vcross(edge0, vneg(edge2))) / 2;
  v3 n0 = vnormalized(ld3(s->normals + 3 * id[0]));
  v3 n1 = vnormalized(ld3(s->normals + 3 * id[1]));
  v3 n2 = vnormalized(ld3(s->normals + 3 * id[2]));
  v3 sum = vadd(vadd(vdivs(vmuls(n0, area1), area), vdivs(vmuls(n1, area2), area)),
                vdivs(vmuls(n2, area0), area));
  return vnormalized(sum);
}

/* calculateDistance, flyscene.cpp:444-478. Returns t (-1 = miss) and P. */
static float calculate_distance(const tracer *t, v3 o, v3 d, int f, v3 *Pout) {
  const oracle_scene *s = t->s;
  v3 vert0 = affine_mul(s->shape, ld3(s->verts + 4 * s->faces[3 * f]));
  v3 facenormal = vnormalized(ld3(s->face_normal + 3 * f));
  if (vdot(facenormal, d) == 0) return -1.f;
  float distancePlane = vdot(facenormal, vert0);
  float orth = distancePlane - vdot(o, facenormal);
  float tt = orth / vdot(d, facenormal);
  v3 P = vadd(o, vmuls(d, tt));
  if (vnorm(interpolate_normal(t, f, P)) == 0) return -1.f;
  *Pout = P;
  return tt;
}

/* intersectBox, flyscene.cpp:484-507 */
static int intersect_box(const tracer *t, v3 o, v3 d, const float *box) {
  v3 o2 = affine_mul(t->minv, o);
  v3 d2 = vnormalized(linear_mul(t->minv, d));
  v3 tmin = vsub(ld3(box), o2), tmax = vsub(ld3(box + 3), o2);
  tmin.x /= d2.x; tmax.x /= d2.x;
  tmin.y /= d2.y; tmax.y /= d2.y;
  tmin.z /= d2.z; tmax.z /= d2.z;
  v3 tinv = V(fminf_std(tmin.x, tmax.x), fminf_std(tmin.y, tmax.y), fminf_std(tmin.z, tmax.z));
  v3 toutv = V(fmaxf_std(tmin.x, tmax.x), fmaxf_std(tmin.y, tmax.y), fmaxf_std(tmin.z, tmax.z));
  float tin = fmaxf_std(tinv.x, fmaxf_std(tinv.y, tinv.z));
  float tout = fminf_std(toutv.x, fminf_std(toutv.y, toutv.z));
  return !(tin > tout || tout < 0);
}

/* calculateMinimumFace, flyscene.cpp:373-396 */
static int minimum_face(const tracer *t, v3 o, v3 d, v3 *Pout, float *tout) {
  const oracle_scene *s = t->s;
  float best = INFINITY;
  int bestf = -1;
  v3 bestP = V(0, 0, 0);
  for (int b = 0; b < s->n_boxes; ++b) {
    if (!intersect_box(t, o, d, s->boxes + 6 * b)) continue;
    for (int i = s->box_off[b]; i < s->box_off[b + 1]; ++i) {
      v3 P;
      float tt = calculate_distance(t, o, d, s->box_faces[i], &P);
      if (0 <= tt && tt < best) { best = tt; bestP = P; bestf = s->box_faces[i]; }
    }
  }
  *Pout = bestP;
  *tout = best;
  return bestf;
}

/* shadow, flyscene.cpp:510-526 */
static int shadow(const tracer *t, v3 P, v3 L) {
  const oracle_scene *s = t->s;
  v3 inter = vadd(P, vmuls(L, (float)0.003));
  for (int b = 0; b < s->n_boxes; ++b) {
    if (!intersect_box(t, P, L, s->boxes + 6 * b)) continue;
    for (int i = s->box_off[b]; i < s->box_off[b + 1]; ++i) {
      v3 Q;
      if (calculate_distance(t, inter, L, s->box_faces[i], &Q) >= 0) return 1;
    }
  }
  return 0;
}

/* calcSingleColor, flyscene.cpp:542-566 */
static v3 calc_single_color(tracer *t, int f, v3 origin, v3 L, v3 I, v3 P) {
  t->c.shadow++;
  if (shadow(t, P, L)) return V(0, 0, 0);
  int m = t->s->face_mat[f];
  if (m != -1) {
    const float *mt = t->s->materials + 12 * m;
    t->ka = ld3(mt); t->kd = ld3(mt + 3); t->ks = ld3(mt + 6);
    t->shininess = mt[9];
  }
  v3 normal = interpolate_normal(t, f, P);
  v3 lightReflection = vsub(L, vmuls(normal, 2 * vdot(normal, L)));
  v3 eye = vnormalized(vsub(origin, P));
  v3 ambient = V(I.x * t->ka.x, I.y * t->ka.y, I.z * t->ka.z);
  v3 diffuse = vmuls(V(I.x * t->kd.x, I.y * t->kd.y, I.z * t->kd.z), fmaxf_std(vdot(L, normal), 0.f));
  v3 specular = vmuls(V(I.x * t->ks.x, I.y * t->ks.y, I.z * t->ks.z),
                      fmaxf_std(powf(vdot(lightReflection, eye), t->shininess), 0.f));
  return vadd(vadd(ambient, diffuse), specular);
}

/* calculateColor, flyscene.cpp:603-614 */
static v3 calculate_color(tracer *t, int f, v3 origin, v3 P) {
  v3 sum = V(0, 0, 0);
  const oracle_lights *l = t->l;
  for (int i = 0; i < l->n_point; ++i) {
    v3 L = vneg(vnormalized(vsub(P, ld3(l->point + 6 * i))));
    sum = vadd(sum, calc_single_color(t, f, origin, L, ld3(l->point + 6 * i + 3), P));
  }
  for (int i = 0; i < l->n_dir; ++i)
    sum = vadd(sum, calc_single_color(t, f, origin, ld3(l->dir + 6 * i), ld3(l->dir + 6 * i + 3), P));
  return V(fmaxf_std(fminf_std(sum.x, 1.f), 0.f), fmaxf_std(fminf_std(sum.y, 1.f), 0.f),
           fmaxf_std(fminf_std(sum.z, 1.f), 0.f));
}

/* traceRay, flyscene.cpp:317-371 */
static v3 trace_ray(tracer *t, v3 origin, v3 dir, int depth, int *hit0, float *t0) {
  if (depth == t->max_depth) return V(0, 0, 0);
  v3 P;
  float tt;
  int f = minimum_face(t, origin, dir, &P, &tt);
  if (depth == 0) { if (hit0) *hit0 = f; if (t0) *t0 = tt; }
  if (tt == INFINITY) return depth == 0 ? t->background : V(0, 0, 0);
  t->c.hits++;
  v3 direct = calculate_color(t, f, origin, P);
  if (t->s->face_mat[f] != -1) t->ks = ld3(t->s->materials + 12 * t->s->face_mat[f] + 6);
  v3 n = interpolate_normal(t, f, P);
  v3 dn = vnormalized(dir);
  v3 r = vnormalized(vsub(dn, vmuls(vmuls(n, vdot(dn, n)), 2)));
  v3 o2 = vadd(P, vmuls(r, (float)0.001));
  if (depth + 1 < t->max_depth) t->c.reflection++;
  v3 refl = trace_ray(t, o2, r, depth + 1, 0, 0);
  v3 c = vadd(direct, V(refl.x * t->ks.x, refl.y * t->ks.y, refl.z * t->ks.z));
  c.x = fmaxf_std(fminf_std(c.x, 1.f), 0.f);
  c.y = fmaxf_std(fminf_std(c.y, 1.f), 0.f);
  c.z = fmaxf_std(fminf_std(c.z, 1.f), 0.f);
  return c;
}

/* ---- drivers --------------------------------------------------------------------------- */
typedef struct {
  const oracle_scene *s; const oracle_camera *cam; const oracle_lights *l;
  const int *pixels; long long n; int max_depth; const float *bg;
  float *rgb; int *hit; float *tprim; float *dirs;
  const float *rays; int *iout; float *fout; int mode; /* 0 render 1 shadow 2 closest */
  int tid, nt; oracle_counts c;
} job;

static void *worker(void *arg) {
  job *j = (job *)arg;
  tracer t;
  oracle_lights nol = {0, 0, 0, 0};
  tracer_init(&t, j->s, j->l ? j->l : &nol, j->max_depth, j->bg);
  if (j->mode == 0) {
    float eye[3], inv[12];
    oracle_camera_derive(j->cam, eye, inv);
    v3 origin = ld3(eye);
    for (long long k = j->tid; k < j->n; k += j->nt) { /* flyscene.cpp:305-311 */
      v3 w = screen_to_world(j->cam, inv, (float)j->pixels[2 * k], (float)j->pixels[2 * k + 1]);
      v3 d = vnormalized(vsub(w, origin));
      int h = -1; float tp = INFINITY;
      t.c.primary++;
      v3 c = trace_ray(&t, origin, d, 0, &h, &tp);
      if (j->rgb) { j->rgb[3 * k] = c.x; j->rgb[3 * k + 1] = c.y; j->rgb[3 * k + 2] = c.z; }
      if (j->hit) j->hit[k] = h;
      if (j->tprim) j->tprim[k] = tp;
      if (j->dirs) { j->dirs[3 * k] = d.x; j->dirs[3 * k + 1] = d.y; j->dirs[3 * k + 2] = d.z; }
    }
  } else {
    for (long long k = j->tid; k < j->n; k += j->nt) {
      v3 o = ld3(j->rays + 6 * k), d = ld3(j->rays + 6 * k + 3);
      if (j->mode == 1) j->iout[k] = shadow(&t, o, d);
      else {
        v3 P; float tt;
        j->iout[k] = minimum_face(&t, o, d, &P, &tt);
        if (j->fout) j->fout[k] = tt;
      }
    }
  }
  j->c = t.c;
  return 0;
}

static void run_jobs(job *proto, int n_threads, oracle_counts *counts) {
  if (n_threads < 1) n_threads = 1;
  job *js = (job *)malloc(sizeof(job) * (size_t)n_threads);
  pthread_t *th = (pthread_t *)malloc(sizeof(pthread_t) * (size_t)n_threads);
  for (int i = 0; i < n_threads; ++i) {
    js[i] = *proto; js[i].tid = i; js[i].nt = n_threads;
    if (n_threads == 1) worker(&js[i]);
    else pthread_create(&th[i], 0, worker, &js[i]);
  }
  oracle_counts c = {0, 0, 0, 0};
  for (int i = 0; i < n_threads; ++i) {
    if (n_threads > 1) pthread_join(th[i], 0);
    c.primary += js[i].c.primary; c.shadow += js[i].c.shadow;
    c.reflection += js[i].c.reflection; c.hits += js[i].c.hits;
  }
  if (counts) *counts = c;
  free(js); free(th);
}

void oracle_render(const oracle_scene *s, const oracle_camera *cam, const oracle_lights *l,
                   const int *pixels, long long n, int max_depth, const float background[3],
                   int n_threads, float *rgb, int *hit, float *tprim, float *dirs,
                   oracle_counts *counts) {
  job j; memset(&j, 0, sizeof j);
  j.s = s; j.cam = cam; j.l = l; j.pixels = pixels; j.n = n; j.max_depth = max_depth;
  j.bg = background; j.rgb = rgb; j.hit = hit; j.tprim = tprim; j.dirs = dirs; j.mode = 0;
  run_jobs(&j, n_threads, counts);
}

void oracle_trace_rays(const oracle_scene *s, const oracle_lights *l, const float *rays,
                       long long n, int max_depth, const float background[3], float *rgb) {
  tracer t;
  tracer_init(&t, s, l, max_depth, background);
  for (long long k = 0; k < n; ++k) {
    v3 c = trace_ray(&t, ld3(rays + 6 * k), ld3(rays + 6 * k + 3), 0, 0, 0);
    rgb[3 * k] = c.x; rgb[3 * k + 1] = c.y; rgb[3 * k + 2] = c.z;
  }
}

void oracle_shadow_rays(const oracle_scene *s, const float *rays, long long n, int n_threads,
                        int *out) {
  job j; memset(&j, 0, sizeof j);
  j.s = s; j.rays = rays; j.n = n; j.iout = out; j.mode = 1; j.max_depth = 1;
  run_jobs(&j, n_threads, 0);
}

void oracle_closest_rays(const oracle_scene *s, const float *rays, long long n, int n_threads,
                         int *face, float *t) {
  job j; memset(&j, 0, sizeof j);
  j.s = s; j.rays = rays; j.n = n; j.iout = face; j.fout = t; j.mode = 2; j.max_depth = 1;
  run_jobs(&j, n_threads, 0);
}

void oracle_quantise(const float *rgb, long long n, unsigned char *out) {
  for (long long k = 0; k < n; ++k) { /* ppmIO.hpp:150 */
    float c = 255 * rgb[k];
    int q = (c != c) ? 0 : (int)c;
    out[k] = (unsigned char)(q < 255 ? (q < 0 ? 0 : q) : 255);
  }
}
