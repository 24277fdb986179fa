// render.cu — wavefront path integrator (placeholder until the kernels land).
#include <cuda_runtime.h>
#include "scene_host.hpp"

namespace ptb {

int render_passes(pt_scene*, const pt_camera_desc*, const pt_render_params*, int, pt_render_stats*){
    set_error("pt_render: not implemented yet");
    return PT_ERR_INVALID;
}

} // namespace ptb
