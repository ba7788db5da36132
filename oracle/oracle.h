/*
 * oracle.h - TEST INFRASTRUCTURE ONLY.
 *
 * Plain-C CPU restatement of the reference's ray-tracing hot path
 * (plindhorst/Ray-Tracing-Project: src/flyscene.cpp:299-614, src/BoundingBox.cpp:7-161,
 * tucano/camera.hpp:115-118,155-173,263-266, tucano/mesh.hpp:448-482,592-644,
 * tucano/model.hpp:102-105,169-173, tucano/utils/objimporter.hpp:81-106,
 * tucano/utils/ppmIO.hpp:135-156).  Every float operation follows the evaluation order
 * Eigen 3.3.7 produces for the reference build flags (-O2 -ffp-contract=off, SSE2 only),
 * so results are bit-identical to the reference binary; oracle/_ref/ref_harness (the
 * reference's own code) pins that in tests/test_oracle_vs_reference.py and through the
 * committed fixtures in tests/golden/.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
 * may load this library.  The product (libflyscene_b200.so) never does.
 */
#ifndef RTP_ORACLE_H
#define RTP_ORACLE_H

#ifdef __cplusplus
extern "C" {
#endif

typedef struct oracle_scene {
  int n_verts, n_faces, n_mats, n_boxes;
  const float *verts;       /* V x 4, object space (w = 1)                mesh.hpp:387 */
  const float *normals;     /* V x 3, per-vertex normals as loaded        mesh.hpp:395 */
  const int *faces;         /* F x 3 vertex ids                           mesh.hpp:238-248 */
  const int *face_mat;      /* F material ids (-1 = none)                                */
  const float *face_normal; /* F x 3 Face::normal                         mesh.hpp:475-477 */
  const float *materials;   /* M x 12: ka[3] kd[3] ks[3] Ns Ni d          mtl.hpp:17-37 */
  float shape[12];          /* getShapeModelMatrix(), 3x4 row-major       model.hpp:102 */
  const float *boxes;       /* B x 6: low[3] high[3] (object space)       BoundingBox.h:16-17 */
  const int *box_off;       /* B + 1 offsets into box_faces                              */
  const int *box_faces;     /* face ids in box order                      BoundingBox.h:26 */
} oracle_scene;

typedef struct oracle_camera {
  float view[12];    /* view matrix, 3x4 row-major (Camera::view_matrix)  */
  float viewport[4]; /* x, y, w, h                                        */
  float fovy;        /* degrees                                           */
  float aspect;      /* aspect_ratio as passed to setPerspectiveMatrix    */
} oracle_camera;

typedef struct oracle_lights {
  int n_point, n_dir;
  const float *point; /* n_point x 6: pos[3] rgb[3]                 flyscene.hpp:132 */
  const float *dir;   /* n_dir x 6: vector[3] (used verbatim) rgb[3] flyscene.hpp:135 */
} oracle_lights;

typedef struct oracle_counts {
  long long primary, shadow, reflection, hits;
} oracle_counts;

/* Face::normal as Mesh::createFaces builds it (mesh.hpp:475-477). out: F x 3 */
void oracle_face_normals(const float *verts, const int *faces, int n_faces, float *out);
/* Smooth normals as MeshImporter::computeNormals (objimporter.hpp:81-106); one index group. */
void oracle_compute_normals(const float *verts, int n_verts, const int *faces, int n_faces, float *out);
/* normalization_scale / objectCenter (mesh.hpp:611-628) then shape = scale * translate
 * (model.hpp:169-173). out_shape: 3x4 row-major. */
void oracle_shape_matrix(const float *verts, int n_verts, float *out_shape);

/* generateBoundingBoxes (flyscene.cpp:399-418) + BoundingBox::{fitMesh,fitFaces,splitBox,
 * outsideFaces,hasFace,hasVertex,averageVertexCoord} (BoundingBox.cpp:26-161).
 * Outputs are malloc'ed; free with oracle_free. Returns the number of boxes. */
int oracle_build_boxes(const float *verts, const int *faces, int n_faces, int min_faces,
                       int max_boxes, float **boxes, int **box_off, int **box_faces);
void oracle_free(void *p);

/* Camera::getCenter (camera.hpp:115-118) and getViewMatrix().inverse() (3x4 row-major). */
void oracle_camera_derive(const oracle_camera *cam, float eye[3], float inv_view[12]);
/* (screenToWorld(i,j) - getCenter()).normalized()  (flyscene.cpp:308, camera.hpp:155-173) */
void oracle_primary_dir(const oracle_camera *cam, int i, int j, float dir[3]);

/* The per-pixel loop of traceRayThread over an explicit pixel list (flyscene.cpp:299-314),
 * full W x H instead of the reference's H x H square.
 * pixels: n x 2 (i, j).  Outputs (any may be NULL): rgb n x 3, hit n (primary face id or -1),
 * tprim n, dirs n x 3.  n_threads > 1 interleaves pixels over pthreads, each with its own
 * copy of the reference's material member state. */
void oracle_render(const oracle_scene *s, const oracle_camera *cam, const oracle_lights *l,
                   const int *pixels, long long n, int max_depth, const float background[3],
                   int n_threads, float *rgb, int *hit, float *tprim, float *dirs,
                   oracle_counts *counts);

/* Flyscene::traceRay for arbitrary rays (flyscene.cpp:317-371). rays: n x 6 (o, d). */
void oracle_trace_rays(const oracle_scene *s, const oracle_lights *l, const float *rays,
                       long long n, int max_depth, const float background[3], float *rgb);
/* Flyscene::shadow (flyscene.cpp:510-526): out[r] = 1 if occluded. rays: n x 6 (P, L). */
void oracle_shadow_rays(const oracle_scene *s, const float *rays, long long n, int n_threads,
                        int *out);
/* calculateMinimumFace (flyscene.cpp:373-396): face id (-1 = miss) and t. */
void oracle_closest_rays(const oracle_scene *s, const float *rays, long long n, int n_threads,
                         int *face, float *t);

/* writePPMImage quantiser (ppmIO.hpp:150): min(255, (int)(255*c)); NaN -> 0 here (the
 * reference's cast is undefined for NaN; excluded from comparisons). */
void oracle_quantise(const float *rgb, long long n_values, unsigned char *out);

#ifdef __cplusplus
}
#endif
#endif
