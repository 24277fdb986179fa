// camera.cuh — Camera ctor + Camera::GenerateRay (Camera.hpp:6-35), inUnitDisk (Random.hpp:60-64).
#pragma once
#include <cmath>
#include "shade_math.cuh"

namespace ptb {

struct CamParams {      // Camera ctor, Camera.hpp:7-20
    V3 from, u, v, w;
    float half_w, half_h, defocus_radius, focus_distance, focus_angle, shutter0, shutter1;
    int medium;
};


inline void make_camera(const pt_camera_desc* c, int W, int H, CamParams& C){
    const V3 from = v3(c->look_from[0], c->look_from[1], c->look_from[2]), at = v3(c->look_at[0], c->look_at[1], c->look_at[2]);
    C.from = from;
    C.w = normalize(from - at);
    C.u = normalize(cross(v3(0, 1, 0), C.w));
    C.v = cross(C.w, C.u);
    C.defocus_radius = (float)(c->focus_distance * std::tan(c->focus_angle / 2.0));
    C.half_w = (float)std::tan(c->fov * 0.5);
    C.half_h = C.half_w * (float)H / (float)W;
    C.focus_distance = c->focus_distance; C.focus_angle = c->focus_angle;
    C.shutter0 = c->shutter_start; C.shutter1 = c->shutter_end;
    C.medium = c->medium;
}


// p = pixel position (already jittered), time in [0,1), (lx, ly) the lens sample
PT_HDI void camera_ray(const CamParams& C, int W, int H, float px, float py, float time, float lx, float ly, V3& org, V3& dir, float& t){
    const float uc = px / (float)W, vc = py / (float)H;
    dir = normalize(-C.w + ((2.0f * uc - 1.0f) * C.half_w) * C.u + ((2.0f * vc - 1.0f) * C.half_h) * C.v);
    org = C.from;
    t = C.shutter0 + (C.shutter1 - C.shutter0) * time;
    if(!(C.focus_distance == 0.0f || C.focus_angle == 0.0f)){
        const float r = sqrtf(lx), th = 2.0f * kPi * ly;          // inUnitDisk, Random.hpp:60-64
        const V3 off = (r * cosf(th)) * (C.u * C.defocus_radius) + (r * sinf(th)) * (C.v * C.defocus_radius);
        dir = dir * C.focus_distance;
        org = C.from + off;
        dir = normalize(dir - off);
    }
}

} // namespace ptb
