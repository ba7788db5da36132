// tucano_camera.hpp - Tucano::Camera / Tucano::Flycamera without OpenGL (included by tucano_lite.hpp).
// Only state and math: view matrix, viewport, fovy / aspect and the two functions the ray tracer
// calls per pixel, getCenter (camera.hpp:115-118) and screenToWorld (camera.hpp:155-173).  The CUDA
// library derives the same quantities once per frame from (view, viewport, fovy, aspect) - see
// rtp_set_camera in include/rtp.h - so these host versions serve user code (e.g. directional
// lights, flyscene.cpp:242) and the parity tests.
#pragma once
#include "eigen_lite.hpp"

#ifndef M_PI
#define M_PI 3.14159265358979323846
#endif

namespace Tucano {

class Camera {
 protected:
  Eigen::Affine3f view_matrix = Eigen::Affine3f::Identity();
  Eigen::Affine3f default_view = Eigen::Affine3f::Identity();
  Eigen::Vector4f viewport = Eigen::Vector4f(0.f, 0.f, 1.f, 1.f);
  float near_plane = 0.1f, far_plane = 100.f;
  float fovy = 60.f;
  float aspect_ratio = 1.f;
  bool use_perspective = true;

 public:
  virtual ~Camera() {}
  void resetViewMatrix() { view_matrix = default_view; }  // camera.hpp:88-91
  virtual void reset() { resetViewMatrix(); }
  // camera.hpp:115-118
  Eigen::Vector3f getCenter() const { return view_matrix.linear().inverse() * (-view_matrix.translation()); }
  // camera.hpp:263-266: float / float tan argument promoted to double by M_PI, result cast to float
  float getPerspectiveScale() const { return (float)(1.0f / std::tan((double)(fovy / 2.0f) * (M_PI / 180.0f))); }
  // camera.hpp:155-173: pixel CORNER (no +0.5), NDC in double, image plane at z = -1
  Eigen::Vector3f screenToWorld(const Eigen::Vector2f &raster_coords) const {
    Eigen::Vector3f norm_coords((float)(2.0 * (raster_coords[0] - viewport[0]) / viewport[2] - 1.0),
                                (float)(1.0 - 2.0 * (raster_coords[1] - viewport[1]) / viewport[3]), -1.0f);
    const float scale = (float)(1.0 / getPerspectiveScale());
    norm_coords[0] *= aspect_ratio * scale;
    norm_coords[1] *= scale;
    return getViewMatrix().inverse() * norm_coords;
  }
  Eigen::Affine3f getViewMatrix() const { return view_matrix; }
  Eigen::Affine3f *viewMatrix() { return &view_matrix; }
  void setViewMatrix(const Eigen::Affine3f &mat) { view_matrix = mat; }
  Eigen::Matrix3f getRotationMatrix() const { return view_matrix.rotation(); }
  Eigen::Vector3f getTranslationMatrix() const { return view_matrix.translation(); }
  Eigen::Vector4f getViewport() const { return viewport; }
  Eigen::Vector2i getViewportSize() const { return Eigen::Vector2i((int)viewport[2], (int)viewport[3]); }
  float getViewportAspectRatio() const { return viewport[2] / viewport[3]; }
  void setViewport(const Eigen::Vector4f &vp) { viewport = vp; }
  void setViewport(const Eigen::Vector2i &vp) { viewport = Eigen::Vector4f(0.f, 0.f, (float)vp[0], (float)vp[1]); }
  void setViewport(const Eigen::Vector2f &vp) { viewport = Eigen::Vector4f(0.f, 0.f, vp[0], vp[1]); }
  // camera.hpp:433-444 (the projection matrix itself only feeds the GL preview)
  void setPerspectiveMatrix(float fy, float in_aspect_ratio, float in_near_plane, float in_far_plane) {
    fovy = fy; aspect_ratio = in_aspect_ratio; near_plane = in_near_plane; far_plane = in_far_plane;
    use_perspective = true;
  }
  float getFovy() const { return fovy; }
  float getAspectRatio() const { return aspect_ratio; }
  float getNearPlane() const { return near_plane; }
  float getFarPlane() const { return far_plane; }
};

// utils/flycamera.hpp:41-318: 6-DOF fly-through camera
class Flycamera : public Camera {
 protected:
  float speed = 0.05f;
  Eigen::Vector2f start_mouse_pos;
  Eigen::Matrix3f rotation_matrix, default_rotation;
  Eigen::Vector3f translation_vector, default_translation;
  float rotation_Y_axis = 0.f, rotation_X_axis = 0.f;

 public:
  Flycamera() { reset(); }
  // flycamera.hpp:76-86: eye at (0,0,2) looking down -z
  void reset() override {
    start_mouse_pos = Eigen::Vector2f::Zero();
    translation_vector = Eigen::Vector3f::Zero();
    rotation_matrix = Eigen::Matrix3f::Identity();
    default_translation = Eigen::Vector3f(0.0, 0.0, -2.0);
    default_rotation = Eigen::Matrix3f::Identity();
    rotation_X_axis = 0.0;
    rotation_Y_axis = 0.0;
    updateViewMatrix();
  }
  void setDefaultView(const Eigen::Affine3f &view) {
    default_rotation = view.rotation();
    default_translation = view.translation();
  }
  Eigen::Vector3f getDefaultTranslation() const { return default_translation; }
  // flycamera.hpp:166-191
  virtual void updateViewMatrix() {
    resetViewMatrix();
    Eigen::Vector3f rotX = Eigen::AngleAxisf(rotation_Y_axis, Eigen::Vector3f::UnitY()) * Eigen::Vector3f::UnitX();
    rotX.normalize();
    Eigen::Vector3f rotZ = Eigen::AngleAxisf(rotation_Y_axis, Eigen::Vector3f::UnitY()) * Eigen::Vector3f::UnitZ();
    rotZ = Eigen::AngleAxisf(rotation_X_axis, rotX) * rotZ;
    rotZ.normalize();
    Eigen::Vector3f rotY = Eigen::AngleAxisf(rotation_X_axis, rotX) * Eigen::Vector3f::UnitY();
    rotY.normalize();
    rotation_matrix.setRow(0, rotX);
    rotation_matrix.setRow(1, rotY);
    rotation_matrix.setRow(2, rotZ);
    view_matrix.rotate(default_rotation);
    view_matrix.rotate(rotation_matrix);
    view_matrix.translate(default_translation);
    view_matrix.translate(translation_vector);
  }
  // flycamera.hpp:196-203 (x and y negated, as in the original)
  virtual void translate(float dx, float dy, float dz) {
    Eigen::AngleAxisf yaw(rotation_Y_axis, Eigen::Vector3f::UnitY());
    translation_vector += (yaw * Eigen::Vector3f(-dx, -dy, dz)) * speed;
  }
  virtual void strideLeft() { translation_vector += (Eigen::AngleAxisf(rotation_Y_axis, Eigen::Vector3f::UnitY()) * Eigen::Vector3f(1.0, 0.0, 0.0)) * speed; }
  virtual void strideRight() { translation_vector += (Eigen::AngleAxisf(rotation_Y_axis, Eigen::Vector3f::UnitY()) * Eigen::Vector3f(-1.0, 0.0, 0.0)) * speed; }
  virtual void moveBack() { translation_vector += (Eigen::AngleAxisf(rotation_Y_axis, Eigen::Vector3f::UnitY()) * Eigen::Vector3f(0.0, 0.0, -1.0)) * speed; }
  virtual void moveForward() { translation_vector += (Eigen::AngleAxisf(rotation_Y_axis, Eigen::Vector3f::UnitY()) * Eigen::Vector3f(0.0, 0.0, 1.0)) * speed; }
  virtual void moveDown() { translation_vector += Eigen::Vector3f::UnitY() * speed; }
  virtual void moveUp() { translation_vector -= Eigen::Vector3f::UnitY() * speed; }
  // flycamera.hpp:258-262
  Eigen::Vector2f normalizePosition(const Eigen::Vector2f &pos) const {
    return Eigen::Vector2f((float)((pos[0] / ((viewport[2] - viewport[0]) / 2.0)) - 1.0),
                           (float)(1.0 - (pos[1] / ((viewport[3] - viewport[1]) / 2.0))));
  }
  void startRotation(Eigen::Vector2f pos) { start_mouse_pos = normalizePosition(pos); }
  // flycamera.hpp:277-302
  virtual void rotate(Eigen::Vector2f new_mouse_pos) {
    Eigen::Vector2f new_position = normalizePosition(new_mouse_pos);
    Eigen::Vector2f dir2d = new_position - start_mouse_pos;
    start_mouse_pos = new_position;
    float anglex = (float)(dir2d[1] * M_PI);
    float angley = (float)(-dir2d[0] * M_PI);
    rotation_X_axis += anglex;
    rotation_Y_axis += angley;
    if (rotation_X_axis > 2 * M_PI) rotation_X_axis -= 2 * M_PI;
    if (rotation_X_axis < 0) rotation_X_axis += 2 * M_PI;
    if (rotation_Y_axis > 2 * M_PI) rotation_Y_axis -= 2 * M_PI;
    if (rotation_Y_axis < 0) rotation_Y_axis += 2 * M_PI;
  }
  virtual void rotateZ(Eigen::Vector2f) {}
  void setSpeed(const float &s) { speed = s; }
};

}  // namespace Tucano
