// bvh4q_encode.hpp -- PT_LAYOUT_FAST: one Bvh4Node -> Bvh4QNode, shared by the host encoder (bvh_build.cpp encode_bvh4q)
// and the device one (scene_pack_gpu.cu).
// Per cluster and axis: step = 2^e with (extent + margins) / step <= 255 and step >= 2^-20 of the largest coordinate
// (so that a quarter step stays far above the float rounding of the traversal's plane distances), origin a quarter
// step below the cluster's lower corner.  Child planes: lower ones rounded down, upper ones up, each by at least a
// quarter step; the result is checked in float arithmetic (origin + q * step) against the exact box and widened if
// a rounding ever went the wrong way -- so the encoding is conservative whatever the accuracy of log2 / ceil on the
// side that runs it.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>
#include "pt_types.h"
#include "bvh_flatten4.hpp"     // PTB_HD

namespace ptb {

PTB_HD inline void encode_node4q(const Bvh4Node& a, Bvh4QNode& q){
    memset(&q, 0, sizeof q);
    const float* mins[3] = { a.xmin, a.ymin, a.zmin };
    const float* maxs[3] = { a.xmax, a.ymax, a.zmax };
    uint32_t words[6] = { 0, 0, 0, 0, 0, 0 };
    for(int ax = 0; ax < 3; ax++){
        float lo = INFINITY, hi = -INFINITY;
        for(int c = 0; c < 4; c++)if(a.ref[c] != REF_EMPTY){ lo = fminf(lo, mins[ax][c]); hi = fmaxf(hi, maxs[ax][c]); }
        if(!(lo <= hi)){ lo = 0.0f; hi = 0.0f; }              // no child at all: never visited
        const double mag = fmax(fmax(fabs((double)lo), fabs((double)hi)), 1e-30);
        int e = (int)ceil(log2(fmax((double)hi - (double)lo, 1e-300) / 254.0));
        const int e_mag = (int)floor(log2(mag)) - 20;
        if(e < e_mag)e = e_mag;
        if(e < -120)e = -120;
        for(;; e++){
            const float step = ldexpf(1.0f, e);
            const float org = lo - 0.25f * step;
            bool ok = e <= 120;
            uint32_t wl = 0, wh = 0;
            for(int c = 0; c < 4 && ok; c++){
                if(a.ref[c] == REF_EMPTY){ wl |= 255u << (8 * c); continue; }      // lo planes at 255, hi planes at 0 steps: whichever of the two is the near plane for a ray, the slab interval comes out empty
                long ql = (long)floor(((double)mins[ax][c] - (double)org) / (double)step - 0.25);
                long qh = (long)ceil(((double)maxs[ax][c] - (double)org) / (double)step + 0.25);
                if(ql < 0)ql = 0;
                // float check, the way a decoder would form the planes
                while(ql > 0 && !(org + (float)ql * step <= mins[ax][c]))ql--;
                while(qh < 256 && !(org + (float)qh * step >= maxs[ax][c]))qh++;
                if(!(org + (float)ql * step <= mins[ax][c]) || qh > 255){ ok = false; break; }
                wl |= (uint32_t)ql << (8 * c); wh |= (uint32_t)qh << (8 * c);
            }
            if(ok){
                q.org[ax] = org; q.ex[ax] = (uint8_t)(e + 127);
                words[ax] = wl; words[3 + ax] = wh;
                break;
            }
            if(e > 120){ q.org[ax] = lo; q.ex[ax] = 254; words[ax] = 0; words[3 + ax] = 0xFFFFFFFFu; break; }   // absurd extents: boxes that always pass
        }
    }
    for(int k = 0; k < 6; k++)q.q[k] = words[k];
    for(int c = 0; c < 4; c++)q.ref[c] = a.ref[c];
    for(int k = 0; k < 8; k++)q.order[k] = a.order[k];
}

} // namespace ptb
