// sensor.h -- PerspectiveCamera::sampleRayDifferential (reference: src/sensors/perspective.cpp:286-317,
// projection set-up :127-150, sensor.cpp:225-309).  The offset rays of the differential (perspective.cpp:160-163,
// 310-314) are recomputed from the stored sample position where a bitmap texture needs them (sensor_ray_differentials).
#pragma once
#include "../../include/mtsb200.h"
#include "traverse.h"

namespace mtsb {

MTSB_HD V3 xform_point(const float *m, V3 p) {
    return V3(m[0] * p.x + m[1] * p.y + m[2] * p.z + m[3], m[4] * p.x + m[5] * p.y + m[6] * p.z + m[7],
              m[8] * p.x + m[9] * p.y + m[10] * p.z + m[11]);
}
MTSB_HD V3 xform_vec(const float *m, V3 v) {
    return V3(m[0] * v.x + m[1] * v.y + m[2] * v.z, m[4] * v.x + m[5] * v.y + m[6] * v.z,
              m[8] * v.x + m[9] * v.y + m[10] * v.z);
}

// frac_x, frac_y: position on the FULL film in [0,1) (sample position + crop offset, divided by the film size)
MTSB_HD void sensor_ray_frac(const MtsbSensor &se, float near_x, float near_y, float frac_x, float frac_y, RayIn &ray) {
    // near-plane point of the sample: closed form of m_sampleToCamera (perspective.cpp:127-136, 296-298)
    V3 nearP(fmul(fsub(1.0f, fmul(2.0f, frac_x)), near_x), fmul(fsub(1.0f, fmul(2.0f, frac_y)), near_y), se.near_clip);
    float len = fsqrt(fadd(fadd(fmul(nearP.x, nearP.x), fmul(nearP.y, nearP.y)), fmul(nearP.z, nearP.z)));
    float rl = fdiv(1.0f, len);
    V3 d(fmul(nearP.x, rl), fmul(nearP.y, rl), fmul(nearP.z, rl));
    float invZ = fdiv(1.0f, d.z);
    ray.mint = fmul(se.near_clip, invZ);
    ray.maxt = fmul(se.far_clip, invZ);
    const float *m = se.to_world;
    ray.ox = m[3]; ray.oy = m[7]; ray.oz = m[11];
    ray.dx = fadd(fadd(fmul(m[0], d.x), fmul(m[1], d.y)), fmul(m[2], d.z));
    ray.dy = fadd(fadd(fmul(m[4], d.x), fmul(m[5], d.y)), fmul(m[6], d.z));
    ray.dz = fadd(fadd(fmul(m[8], d.x), fmul(m[9], d.y)), fmul(m[10], d.z));
}

MTSB_HD void sensor_ray(const MtsbSensor &se, float near_x, float near_y, float px, float py, RayIn &ray) {
    float a = fdiv(fadd(px, (float) se.crop_x), (float) se.film_width);
    float b = fdiv(fadd(py, (float) se.crop_y), (float) se.film_height);
    sensor_ray_frac(se, near_x, near_y, a, b, ray);
}

// Directions of the two offset rays of a camera ray (both start at the camera position): perspective.cpp:312-313 with
// m_dx / m_dy = one pixel on the near plane (:160-163), then RayDifferential::scaleDifferential(1 / sqrt(sampleCount))
// (ray.h:174-179, integrator.cpp:200-201, 237).
MTSB_HD void sensor_ray_differentials(const MtsbSensor &se, float near_x, float near_y, float px, float py, int sample_count,
                                      V3 &rxd, V3 &ryd) {
    const float a = fdiv(fadd(px, (float) se.crop_x), (float) se.film_width);
    const float b = fdiv(fadd(py, (float) se.crop_y), (float) se.film_height);
    const V3 nearP(fmul(fsub(1.0f, fmul(2.0f, a)), near_x), fmul(fsub(1.0f, fmul(2.0f, b)), near_y), se.near_clip);
    const float dxx = fdiv(fmul(-2.0f, near_x), (float) se.film_width), dyy = fdiv(fmul(-2.0f, near_y), (float) se.film_height);
    const V3 d = xform_vec(se.to_world, normalize(nearP));
    const V3 rx = xform_vec(se.to_world, normalize(V3(fadd(nearP.x, dxx), nearP.y, nearP.z)));
    const V3 ry = xform_vec(se.to_world, normalize(V3(nearP.x, fadd(nearP.y, dyy), nearP.z)));
    const float amount = fdiv(1.0f, fsqrt((float) (sample_count > 1 ? sample_count : 1)));
    rxd = d + (rx - d) * amount;
    ryd = d + (ry - d) * amount;
}

}  // namespace mtsb
