// Camera (engine/renderer/camera.h:9-45). set_camera fills the reference's fields through atray_camera_set so the
// basis bits are the library's (= the reference's arithmetic).
#pragma once
#include "settings.h"
#include "atray_b200.h"
struct Camera
{
    RenderSettings render_settings;
    f32 aspect_ratio;
    vec3f camera_z, camera_x, camera_y;
    vec3f eye = {0, 0, 0};
    vec3f frame_center;
    f32 h_fov = 1;
    f32 half_pixel_width = 0, half_pixel_height = 0;
};
inline void camera_to_abi(const Camera& cm, AtrayCamera& a)
{
    a.width = cm.render_settings.resolution.x; a.height = cm.render_settings.resolution.y;
    a.anti_aliasing = cm.render_settings.anti_aliasing; a.samples_per_pixel = cm.render_settings.samples_per_pixel;
    a.bounce_limit = cm.render_settings.bounce_limit; a.aspect_ratio = cm.aspect_ratio;
    const vec3f* src[5] = {&cm.camera_z, &cm.camera_x, &cm.camera_y, &cm.eye, &cm.frame_center};
    float* dst[5] = {a.camera_z, a.camera_x, a.camera_y, a.eye, a.frame_center};
    for (int i = 0; i < 5; ++i) { dst[i][0] = src[i]->x; dst[i][1] = src[i]->y; dst[i][2] = src[i]->z; }
    a.h_fov = cm.h_fov; a.half_pixel_width = cm.half_pixel_width; a.half_pixel_height = cm.half_pixel_height;
}
// camera.h:23-34: new eye and orientation for a camera set_camera has initialised (the basis is recomputed by the library with the
// reference's arithmetic; resolution, samples, h_fov and the half-pixel sizes do not depend on the position and stay)
inline void update_camera_pos(Camera& cm, vec3f eye, vec3f facing_towards)
{
    AtrayCamera a;
    atray_camera_set(&a, cm.render_settings.resolution.x, cm.render_settings.resolution.y, cm.render_settings.anti_aliasing,
                     cm.render_settings.samples_per_pixel, cm.render_settings.bounce_limit, &eye.x, &facing_towards.x, cm.h_fov);
    cm.camera_z = {a.camera_z[0], a.camera_z[1], a.camera_z[2]};
    cm.camera_x = {a.camera_x[0], a.camera_x[1], a.camera_x[2]};
    cm.camera_y = {a.camera_y[0], a.camera_y[1], a.camera_y[2]};
    cm.eye = {a.eye[0], a.eye[1], a.eye[2]};
    cm.frame_center = {a.frame_center[0], a.frame_center[1], a.frame_center[2]};
}
inline void set_camera(Camera& cm, vec3f eye, vec3f facing_towards, RenderSettings& render_settings, f32 h_fov_)
{
    AtrayCamera a;
    atray_camera_set(&a, render_settings.resolution.x, render_settings.resolution.y, render_settings.anti_aliasing,
                     render_settings.samples_per_pixel, render_settings.bounce_limit, &eye.x, &facing_towards.x, h_fov_);
    cm.render_settings = render_settings;
    cm.aspect_ratio = a.aspect_ratio;
    cm.camera_z = {a.camera_z[0], a.camera_z[1], a.camera_z[2]};
    cm.camera_x = {a.camera_x[0], a.camera_x[1], a.camera_x[2]};
    cm.camera_y = {a.camera_y[0], a.camera_y[1], a.camera_y[2]};
    cm.eye = {a.eye[0], a.eye[1], a.eye[2]};
    cm.frame_center = {a.frame_center[0], a.frame_center[1], a.frame_center[2]};
    cm.h_fov = a.h_fov; cm.half_pixel_width = a.half_pixel_width; cm.half_pixel_height = a.half_pixel_height;
}
