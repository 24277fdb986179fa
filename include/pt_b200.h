/*
 * pt_b200.h — C-ABI of the B200-native wavefront path tracer.
 *
 * Drop-in boundary for ONE hot path of marko176/PathTracing (citations are
 * file:line in the reference tree): BVH2/BVH4 TLAS+BLAS closest-hit and
 * shadow-ray traversal plus the path-integrator bounce loop.  The reference has
 * no FFI layer of its own: its API is a set of C++ classes
 * (Scene / Mesh / BLASx / TLASx / Camera / Film / *Integrator).  This header
 * is what those classes bind to once they are re-implemented as thin host
 * objects (see pathtracing_b200/host/pt_host.hpp and INTEGRATION.md).
 *
 * Conventions: every function returns PT_OK (0) or a negative pt_status, except
 * the *_add_* calls which return the new object's id (>= 0) or a negative
 * status.  pt_last_error() returns a thread-local message for the last failure.
 * Host arrays are borrowed for the duration of the call and copied; no torch
 * types, no C++ types, plain pointers and sizes only.
 */
#ifndef PT_B200_H
#define PT_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PT_API __attribute__((visibility("default")))

typedef enum pt_status {
    PT_OK = 0,
    PT_ERR_INVALID = -1,  /* bad argument / bad id / wrong state            */
    PT_ERR_CUDA = -2,     /* CUDA runtime error (message in pt_last_error)  */
    PT_ERR_NO_DEVICE = -3,/* no CUDA device: there is NO CPU fallback       */
    PT_ERR_STATE = -4,    /* scene not committed / already committed        */
    PT_ERR_NOMEM = -5,
    PT_ERR_OVERFLOW = -6  /* some ray needed more traversal-stack entries than the kernel has (96): results unreliable */
} pt_status;

/* ---- ray casting ------------------------------------------------------------
 * replaces Scene::Intersect / Scene::IntersectPred (Scene.hpp:10-14,
 * Scene.cpp:4-6,31-33) called per ray; here called per batch.               */

/* 32 B ray.  replaces struct Ray (Ray.hpp:14-41): inv_dir is recomputed on the
 * device with the reference's clamp (|d|<1e-32 -> 1e32).                    */
typedef struct pt_ray {
    float ox, oy, oz, tmax;
    float dx, dy, dz, time;
} pt_ray;

#define PT_MISS 0xFFFFFFFFu

/* 16 B hit record.  Carries what SurfaceInteraction (Interaction.hpp:36-51)
 * is rebuilt from in the shade stage: t, the two barycentrics returned by
 * glm::intersectRayTriangle (glm/gtx/intersect.inl:29-94) and the triangle
 * index inside its BLAS (meshes concatenated in pt_scene_add_blas order).
 * prim == PT_MISS and t == +inf on a miss.                                  */
typedef struct pt_hit {
    float t, u, v;
    uint32_t prim;
} pt_hit;

/* ---- scene description -------------------------------------------------------- */

typedef enum pt_tex_kind {
    PT_TEX_SOLID = 0,     /* SolidColor        Texture.hpp:122-133  */
    PT_TEX_IMAGE_U8 = 1,  /* ImageTexture      Texture.hpp:135-165  */
    PT_TEX_IMAGE_F32 = 2, /* FloatImageTexture Texture.hpp:167-195  */
    PT_TEX_CHECKER = 3,   /* CheckerTexture    Texture.hpp:197-214  */
    PT_TEX_UV = 4,        /* UVTexture         Texture.hpp:216-223  */
    PT_TEX_NORMAL = 5     /* NormalTexture     Texture.hpp:225-232  */
} pt_tex_kind;

typedef struct pt_texture_desc {
    int32_t kind;           /* pt_tex_kind */
    float color[3];         /* SOLID: albedo */
    float color_scale[3];   /* Texture::colorScale (Texture.hpp:110) */
    int32_t width, height, channels; /* IMAGE_*: texel grid, 1..4 channels  */
    const void* pixels;     /* IMAGE_U8: uint8 (already sRGB->linear LUT'ed if
                               the reference would load with gammaCorrection,
                               Texture.cpp:10-18); IMAGE_F32: float          */
    int32_t tex_a, tex_b;   /* CHECKER: the two child textures */
    float uv_scale[2];      /* CHECKER: uvscale (invScale = 1/uvscale)       */
} pt_texture_desc;

typedef enum pt_mat_kind {
    PT_MAT_DIFFUSE = 0,        /* MicrofacetDiffuse    Material.hpp:200-361 */
    PT_MAT_DIELECTRIC = 1,     /* MicrofacetDielectric Material.hpp:363-598 */
    PT_MAT_THIN_DIELECTRIC = 2,/* ThinDielectric       Material.hpp:601-657 */
    PT_MAT_CONDUCTOR = 3       /* SpecularConductor    Material.hpp:660-673 */
} pt_mat_kind;

typedef enum pt_alpha_mode { /* AlphaMode, Material.hpp:176-198 */
    PT_ALPHA_OPAQUE = 0,
    PT_ALPHA_BLEND = 1,       /* the reference's default for Diffuse/Dielectric */
    PT_ALPHA_MASK = 2
} pt_alpha_mode;

typedef struct pt_material_desc {
    int32_t kind;        /* pt_mat_kind */
    int32_t tex;         /* albedo texture id                                  */
    int32_t norm;        /* normal-map texture id or -1                        */
    int32_t rough;       /* roughness texture (.g used) or -1 -> default       */
    int32_t metal;       /* metallic texture (.b used) or -1 -> default (0)    */
    int32_t alpha;       /* alpha-mask texture (.r used) or -1                 */
    float ri;            /* DIELECTRIC / THIN: refractive index                */
    float albedo[3];     /* CONDUCTOR: Schlick F0 tint                         */
    int32_t alpha_mode;  /* pt_alpha_mode (after setAlphaTester's demotion rule,
                            Material.hpp:350-353, applied by the caller)       */
    float alpha_cutoff;
} pt_material_desc;

typedef struct pt_medium_desc { /* HomogeneusMedium ctor, Medium.hpp:18-20 */
    float sigma_a[3], sigma_s[3];
    float g;             /* HenyeyGreenstein(g), clamped to +-0.99 */
    float density;
    float Le[3];
    float le_density;
} pt_medium_desc;

/* One Mesh (Mesh.hpp:16-22).  normals/texcoords are per vertex and required;
 * tangents may be NULL (Shape.cpp:228-238 fallback).  material < 0 makes the
 * mesh a material-less medium boundary (Integrators.cpp:228-232).
 * emissive_tex >= 0 turns every triangle into a two-sided AreaLight as
 * Model::BuildBlas does (Model.hpp:47-59).                                   */
typedef struct pt_mesh_desc {
    const float* positions;   /* 3 * n_vertices */
    const float* normals;     /* 3 * n_vertices */
    const float* texcoords;   /* 2 * n_vertices */
    const float* tangents;    /* 3 * n_vertices or NULL */
    const uint32_t* indices;  /* 3 * n_triangles */
    uint32_t n_vertices, n_triangles;
    int32_t material, emissive_tex, medium;
} pt_mesh_desc;

/* An analytic shape placed directly in the TLAS: GeometricPrimitive(shape, material, areaLight, medium)
 * handed to Scene::Add (Primitive.hpp:17-33, examples/example_1.cpp:37-66) with a SphereShape
 * (Shape.hpp:21-45, Shape.cpp:3-81) or a QuadShape (Shape.hpp:118-175, Shape.cpp:316-359).          */
typedef enum pt_shape_kind { PT_SHAPE_SPHERE = 0, PT_SHAPE_QUAD = 1 } pt_shape_kind;

typedef struct pt_shape_desc {
    int32_t kind;           /* pt_shape_kind */
    float p[3];             /* SPHERE: center; QUAD: origin Q */
    float u[3], v[3];       /* QUAD: edge vectors */
    float radius;           /* SPHERE */
    int32_t material;       /* -1: medium boundary (Integrators.cpp:228-232) */
    int32_t medium;         /* -1: none */
    int32_t emissive_tex;   /* >= 0: the primitive carries AreaLight(shape, texture, one_sided) (Light.hpp:179-206) */
    int32_t one_sided;
} pt_shape_desc;

typedef enum pt_light_kind {
    PT_LIGHT_UNIFORM_INFINITE = 0, /* UniformInfiniteLight  Light.cpp:27-50   */
    PT_LIGHT_GRADIENT_INFINITE = 1,/* FunctionInfiniteLight Light.cpp:53-108 with
                                      the gradient sky every reference scene uses
                                      (main.cpp:108-111):
                                      scale*((1-a)*color + a*color_b), a=.5*(d.y+1);
                                      tex >= 0 instead: any other sky function, baked
                                      by the host into a lat-long float texture
                                      (GetSphereUV parametrisation) that Le reads;
                                      sampling and pdf stay uniform on the sphere   */
    PT_LIGHT_TEXTURE_INFINITE = 2, /* TextureInfiniteLight  Light.cpp:110-200 */
    PT_LIGHT_DISTANT = 3,          /* DistantLight          Light.cpp:202-227 */
    PT_LIGHT_POINT = 4             /* PointLight            Light.cpp:229-251 */
} pt_light_kind;

typedef struct pt_light_desc {
    int32_t kind;
    float color[3];
    float color_b[3];
    float scale;        /* GRADIENT: multiplier; TEXTURE: LeScale */
    float dir_or_pos[3];/* DISTANT: dir; POINT: p */
    int32_t tex;        /* TEXTURE_INFINITE */
    int32_t power_func; /* 0: sqrt(r) (default), 1: 4*r (PointLight default), 2: 4*sqrt(r) */
} pt_light_desc;

typedef struct pt_scene pt_scene;

/* lifetime */
PT_API int pt_device_count(void);
PT_API pt_scene* pt_scene_create(int device);            /* NULL on failure   */
PT_API void pt_scene_destroy(pt_scene*);
PT_API const char* pt_last_error(void);

/* build (before commit).  replaces the user-side object graph of
 * examples/example_1.cpp:17-101.                                            */
PT_API int pt_scene_add_texture(pt_scene*, const pt_texture_desc*);
PT_API int pt_scene_add_material(pt_scene*, const pt_material_desc*);
PT_API int pt_scene_add_medium(pt_scene*, const pt_medium_desc*);
PT_API int pt_scene_add_mesh(pt_scene*, const pt_mesh_desc*);
/* a BLAS = Model::BuildBlas<BLASx> over these meshes (Model.hpp:41-60)       */
PT_API int pt_scene_add_blas(pt_scene*, const int32_t* mesh_ids, int n_meshes);
/* transform: column-major 4x4 (glm::mat4) -> TransformedPrimitive
 * (Primitive.hpp:37-50); NULL -> the BLAS itself is the TLAS primitive.      */
PT_API int pt_scene_add_instance(pt_scene*, int blas_id, const float* transform16);
/* returns the primitive's id in the same sequence as pt_scene_add_instance (Scene::Add order).
 * A hit on a shape reports prim = 0 and inst = this id; u, v = the quad's (alpha, beta), 0 for a sphere. */
PT_API int pt_scene_add_shape(pt_scene*, const pt_shape_desc*);
/* TransformedPrimitive(GeometricPrimitive(shape, ...), transform) handed to Scene::Add
 * (Primitive.hpp:34-49, Primitive.cpp:32-73): the shape is intersected in its own space, the hit comes back
 * through `transform` (column-major glm::mat4), its AreaLight becomes a TransformedLight (Light.cpp:296-336). */
PT_API int pt_scene_add_shape_xf(pt_scene*, const pt_shape_desc*, const float* transform16);
/* AnimatedPrimitive(primitive, direction, timeBounds) (Primitive.hpp:52-66, Primitive.cpp:77-96; main.cpp:189)
 * around the top-level primitive `prim_id` (any id returned by pt_scene_add_instance / _shape / _shape_xf):
 * per ray, the primitive is seen under translate(direction * t), t = clamp(ray.time - tb.x, tb.x, tb.y) /
 * (tb.y - tb.x); its lights become AnimatedLights (Light.cpp:338-365).  One wrapper per primitive.            */
PT_API int pt_scene_animate(pt_scene*, int prim_id, const float direction[3], const float time_bounds[2]);
PT_API int pt_scene_add_light(pt_scene*, const pt_light_desc*);
PT_API int pt_scene_set_medium(pt_scene*, int medium_id); /* Scene::SetMedium  */

/* Scene::BuildTlas<TLAS2|TLAS4|TLAS8> + BLAS2|BLAS4|BLAS8 build + upload.
 * bvh_width in {2,4,8} (BLAS8/TLAS8, BVH.hpp:1228-1552).  light_sampler: 0 Uniform, 1 Power (LightSampler.cpp). */
PT_API int pt_scene_commit(pt_scene*, int bvh_width, int light_sampler);

/* Where the binary binned-SAH build (BVHBase::BuildBaseThreaded, BVH.hpp:290-390) of every BLAS and of the
 * TLAS runs.  PT_BUILDER_GPU takes the same decisions with the same float operations on the device
 * (csrc/bvh_build_gpu.cu) and yields the same tree and leaf order as PT_BUILDER_HOST, so the packed
 * BVH2 / BVH4 / BVH8 arrays are byte-identical; hierarchies under 4096 items stay on the host (launch
 * overhead).  With PT_BUILDER_GPU a BLAS of 4096 triangles or more (any width) is packed on the device altogether
 * (csrc/scene_pack_gpu.cu: triangle boxes from the mesh arrays, the binary build, the BVH2 / BVH4 / BVH8 collapse in the host
 * flatten's DFS numbering, the triangle records, the quantised clusters): the same bytes as the host path
 * (tests/test_bvh_build.py::test_blas_packed_on_the_device_gives_the_host_arrays), no tree download, no
 * per-triangle host array.  Call before pt_scene_commit.  Default: PT_BUILDER_GPU (a 10 M-triangle commit
 * 4.05 s on the host -> 0.34 s, profiles/r02_build_check_v4.log); PTB_BUILDER=host in
 * the environment selects the host builder at pt_scene_create, PTB_GPU_PACK=0 keeps the GPU builder but
 * flattens on the host.                                                                                       */
typedef enum pt_builder { PT_BUILDER_HOST = 0, PT_BUILDER_GPU = 1 } pt_builder;
PT_API int pt_scene_set_builder(pt_scene*, int builder);

/* Node layout the traversal kernels walk (before pt_scene_commit; bvh_width 4 only).
 * PT_LAYOUT_EXACT (default): the reference's BVH4_CLUSTER content (BVH.hpp:45-60) in
 * 128 B, box tests in the reference's own arithmetic: the same nodes are visited in
 * the same order, exact ties resolve to the same triangle.
 * PT_LAYOUT_FAST: the SAME tree with the child boxes quantised to 8 bits per plane
 * around a per-cluster origin (64 B per cluster, 64 B per triangle record), rounded
 * outwards, so every box contains the exact one; the closest-hit kernel visits a
 * cluster's children by entry distance instead of the reference's per-octant order.
 * Triangle tests are unchanged, so the closest hit (primitive / instance id, bits of
 * t) is the reference's except where two candidates tie exactly or the reference's
 * own box test rounds a hit away that is an ulp or two closer (12 of 33.5 M rays,
 * profiles/r02_tie_check.log); t is never larger.  Half the 32 B sectors per cluster
 * visit and 15 % fewer instructions: 1.62 against 1.07 Grays/s.   */
typedef enum pt_layout { PT_LAYOUT_EXACT = 0, PT_LAYOUT_FAST = 1 } pt_layout;
PT_API int pt_scene_set_layout(pt_scene*, int layout);

typedef struct pt_build_stats {   /* wall-clock split of the last pt_scene_commit */
    double total_s;
    double pack_s;          /* triangle records + BVH builds + flattening (includes bvh2_s, flatten_s) */
    double bvh2_s;          /* binary SAH builds (host: CPU time; GPU: upload + kernels + download)    */
    double flatten_s;       /* BuildBVH2 / buildBVH4 / buildBVH8 collapse (host)                      */
    double upload_s;        /* traversal arrays H2D                                                   */
    double shade_tables_s;  /* shading / light tables + their upload                                  */
    float gpu_build_ms;     /* PT_BUILDER_GPU: build kernels (CUDA events)                            */
    float gpu_upload_ms, gpu_download_ms;
    int32_t builder, gpu_levels;
    uint32_t gpu_subtrees;
    uint64_t gpu_items;     /* items that went through the GPU builder                                */
} pt_build_stats;
PT_API int pt_scene_build_stats(const pt_scene*, pt_build_stats*);

/* Stand-alone binary build = BVHBase::BuildBVHBase (BVH.hpp:95-105) over n boxes {lo[3], hi[3]}.
 * nodes[0] is the root, nodes are numbered in DFS preorder (left subtree first); inner: count = 0,
 * left / right = children, axis = split axis; leaf: right = first item (position in `order`), count > 0.
 * order[n]: item index per leaf-order position.  node_capacity >= 2n - 1.  builder as above; PT_BUILDER_HOST
 * needs no device.                                                                                       */
typedef struct pt_bvh2_node { float lo[3], hi[3]; uint32_t left, right, count, axis; } pt_bvh2_node;
PT_API int pt_build_bvh2(int device, int builder, const float* boxes6, uint32_t n,
                         pt_bvh2_node* nodes, uint32_t node_capacity, uint32_t* n_nodes, uint32_t* order,
                         float* gpu_build_ms);

/* scene bounds after commit: Scene::BoundingBox (Scene.cpp:45-47)           */
PT_API int pt_scene_bounds(const pt_scene*, float bmin[3], float bmax[3]);

/* ---- ray casting entry points ------------------------------------------------- */

/* host buffers: H2D, kernel, D2H inside the call.  inst may be NULL.        */
PT_API int pt_trace_closest(pt_scene*, const pt_ray* rays, size_t n,
                            pt_hit* hits, uint32_t* inst);
PT_API int pt_trace_any(pt_scene*, const pt_ray* rays, size_t n, uint8_t* occluded);

/* Compact transfer form of pt_trace_closest for callers whose rays all have tmax = inf and time = 0 and who need
 * only {t, primitive id} back (plus the optional instance ids): 24 B in and 8 B out per ray instead of 32 + 16.
 * Same traversal, same kernel family, same t bits and ids as pt_trace_closest on the expanded rays.  It exists
 * because host-buffer ray casts on 4 or more GPUs of one box are bound by the HOST memory system (every rank
 * moves its rays and hits through the same NUMA node): a third fewer bytes is a third more
 * rays.  No counterpart in the reference (its Scene::Intersect takes one Ray at a time, Scene.cpp:4-6).         */
typedef struct pt_ray24 { float o[3], d[3]; } pt_ray24;
typedef struct pt_hit8 { float t; uint32_t prim; } pt_hit8;     /* miss: t = +inf, prim = 0xFFFFFFFF */
PT_API int pt_trace_closest_compact(pt_scene*, const pt_ray24* rays, size_t n, pt_hit8* hits, uint32_t* inst);

/* device buffers (already resident in HBM), launched on `stream`
 * (a cudaStream_t passed as void*; NULL = default stream); asynchronous.    */
PT_API int pt_trace_closest_dev(pt_scene*, const void* d_rays, size_t n,
                                void* d_hits, void* d_inst, void* stream);
PT_API int pt_trace_any_dev(pt_scene*, const void* d_rays, size_t n,
                            void* d_occluded, void* stream);

/* The reference's traversal stacks are 32 entries and unchecked (BVH.hpp:478,505,
 * 1035,1128: overflow is undefined behaviour there).  Here the unified TLAS+BLAS
 * stack holds 96 entries and a push that finds it full is COUNTED: the
 * synchronous entry points (pt_trace_closest / pt_trace_any / pt_render) return
 * PT_ERR_OVERFLOW when one of their rays overflowed; after asynchronous *_dev
 * launches read the scene's running total here (synchronises the device).     */
PT_API int pt_scene_stack_overflows(pt_scene*, uint64_t* count);

/* NO COUNTERPART IN THE REFERENCE: added
 * for callers that must not allocate device memory (tight serving loops,
 * harnesses that forbid cudaMalloc).  Same result as pt_trace_closest, bit for
 * bit (the same kernel), but the call makes no cudaMalloc / cudaFree /
 * cudaMallocHost or other allocating call: the staged rays, hits, instance ids
 * and the ray-dispenser cursor live in a fixed `static __device__` scratch
 * array inside the library (3.4 MB, held by every GPU that loads the library
 * until the process exits).  n > PT_STATIC_MAX_RAYS is refused with
 * PT_ERR_INVALID and a message.  Host buffers in, host buffers out, `inst` may
 * be NULL; synchronous.  All calls on one GPU share the scratch (it keeps the
 * last call's contents; what must start at zero is cleared per call): calls
 * that overlap on one GPU (two host threads, two scenes) race on it.          */
#define PT_STATIC_MAX_RAYS 65536
PT_API int pt_trace_closest_static(pt_scene*, const pt_ray* rays, size_t n,
                                   pt_hit* hits, uint32_t* inst);

/* traversal statistics of the last counted launch (for the roofline's
 * algorithmic bytes): runs a counting build of the SAME
 * traversal over device rays and returns totals.                            */
typedef struct pt_trace_stats {
    uint64_t rays;
    uint64_t tlas_nodes;   /* TLAS inner-node (child-pair / cluster) visits  */
    uint64_t blas_nodes;   /* BLAS inner-node visits                         */
    uint64_t instances;    /* instance records entered                       */
    uint64_t triangles;    /* triangle tests                                 */
    uint64_t hits;
} pt_trace_stats;
PT_API int pt_trace_closest_stats(pt_scene*, const void* d_rays, size_t n,
                                  pt_trace_stats* out);

/* Counted renders (per-sample algorithmic bytes): while counting is on, pt_render
 * launches the counting build of the same traversal kernels (slower) and the scene sums the
 * cluster / instance / triangle visits of its closest-hit and shadow rays.
 * pt_scene_render_counts returns the totals since the last call and clears them; `rays`
 * is left 0 (pt_render_stats has the ray counts).  bvh_width 4 (either layout) only.      */
PT_API int pt_scene_set_counting(pt_scene*, int on);
PT_API int pt_scene_render_counts(pt_scene*, pt_trace_stats* closest, pt_trace_stats* shadow);

/* ---- rendering ---------------------------------------------------------------- */

typedef struct pt_camera_desc { /* Camera ctor, Camera.hpp:6-20 */
    float look_from[3], look_at[3];
    float fov;             /* radians, horizontal (Camera.hpp:12) */
    float focus_angle, focus_distance;
    int32_t medium;        /* camera medium id or -1 */
    float shutter_start, shutter_end;
} pt_camera_desc;

typedef enum pt_integrator_kind {
    PT_INTEGRATOR_SIMPLE = 0, /* SimplePathIntegrator::Li Integrators.cpp:131-180 */
    PT_INTEGRATOR_PATH = 1,   /* PathIntegrator::Li       Integrators.cpp:182-294 */
    PT_INTEGRATOR_VOLPATH = 2 /* VolPathIntegrator::Li    Integrators.cpp:296-479 */
} pt_integrator_kind;

typedef enum pt_filter_kind {
    PT_FILTER_BOX = 0, PT_FILTER_GAUSSIAN = 1, PT_FILTER_MITCHELL = 2, PT_FILTER_LANCZOS = 3
} pt_filter_kind;

typedef struct pt_render_params {
    int32_t width, height;
    int32_t spp;            /* exactly spp samples per pixel (adaptive loop of
                               Integrators.cpp:59-86 disabled unless adaptive=1) */
    int32_t max_depth;
    int32_t integrator;     /* pt_integrator_kind */
    int32_t sampler;        /* 0 UniformSampler, 1 StratifiedSampler */
    int32_t strata_x, strata_y;
    int32_t filter;         /* pt_filter_kind */
    float filter_radius[2];
    float filter_param[2];  /* gaussian sigma | mitchell b,c | lanczos tau */
    uint64_t seed;
    int32_t pass_begin;     /* sample passes [pass_begin, pass_end) with stride */
    int32_t pass_stride;    /*   pass_stride: rank r of G renders r, r+G, ...   */
    int32_t adaptive;       /* 1: per-pixel adaptive loop, up to 128 batches    */
} pt_render_params;

typedef struct pt_render_stats {
    uint64_t samples;       /* camera samples traced                          */
    uint64_t closest_rays;  /* extend rays                                    */
    uint64_t shadow_rays;   /* NEE rays (any-hit or transmittance march legs) */
    uint64_t kernel_launches;
    double gpu_ms;          /* CUDA-event time of the whole render            */
    uint64_t nonfinite_samples; /* samples whose radiance came out NaN/inf (e.g. the inf * 0 of a
                               grazing GGX D, Material.hpp:73-86) and were left out of the film:
                               the reference would add them and lose the pixel               */
} pt_render_stats;

/* Integrator::Render (Integrators.cpp:23-129) for the params' sample passes.
 * Accumulates into the scene's device film (double RGBW, W*H*4), creating or
 * clearing it when `clear` != 0.                                             */
/* Multi-GPU adaptive sampling: when the sample passes of an adaptive render are sharded over
 * ranks (pass_begin / pass_stride), every rank must apply the stopping rule of Integrators.cpp:82-85 to the
 * SAME per-pixel statistics.  After each round pt_render calls `sum(user, d_buf, n, stream)`, which must add
 * the n doubles at device pointer d_buf over all ranks in place (ncclAllReduce / torch.distributed.all_reduce)
 * and return 0 once the result is visible to work queued on `stream` (a cudaStream_t).  Not needed for
 * fixed-spp renders, whose only collective is the final film reduce.                                       */
typedef int (*pt_allreduce_fn)(void* user, void* d_buf, size_t n_doubles, void* stream);
PT_API int pt_scene_set_allreduce(pt_scene*, pt_allreduce_fn sum, void* user);

PT_API int pt_render(pt_scene*, const pt_camera_desc*, const pt_render_params*,
                     int clear, pt_render_stats* stats);
/* device film pointer (double[4*W*H], RGB sums * w and weight sums) for the
 * caller's own collective (one ncclReduce).                   */
PT_API int pt_film_device_ptr(pt_scene*, void** d_film, size_t* n_doubles);
/* Film readback: rgbw[4*W*H] doubles, row 0 first (Film.hpp:125-132 layout). */
PT_API int pt_film_download(pt_scene*, double* rgbw);
/* per-pixel sum of squared sample radiance, sumsq[3*W*H] (sample assigned to pixel floor(p)): what
 * VarianceEstimator (Util.hpp:8-43) accumulates per pixel in Integrators.cpp:78-81.             */
PT_API int pt_film_download_sumsq(pt_scene*, double* sumsq);
PT_API int pt_film_upload(pt_scene*, const double* rgbw, int width, int height);

/* ---- multi-GPU rendering -----------------------------------------
 * replaces the worker threads of Integrator::Render(threadCount) (Integrators.hpp:13,
 * Integrators.cpp:23-41,117-123) and their merge into one film (Film::Merge,
 * Film.hpp:125-132): a worker is a GPU.  The committed scene is replicated device
 * to device, rank r of G renders the sample passes r, r+G, ... of the frame into its
 * own film, and ONE reduce(sum) of the fp64 accumulators (NCCL over NVLink) brings
 * the frame to rank 0; adaptive renders add one all-reduce of the per-pixel
 * estimator sums per round.  Sample streams are keyed by sample id, so the frame
 * equals the single-GPU frame up to the order of the fp64 sums.  NCCL is bound at
 * run time (libnccl.so.2, the copy already in the process if there is one).        */

/* a second committed copy of `src` on `device` (tables copied GPU to GPU, no build)  */
PT_API int pt_scene_replicate(pt_scene* src, int device, pt_scene** out);

/* one process, G GPUs, one host thread per GPU.  devices[0] must be the primary's
 * device; devices == NULL: every visible GPU.  All entries different (NCCL), or all
 * equal ("virtual ranks" sharing one GPU: the sharding logic without NCCL, reduced
 * by an add kernel).  The primary stays owned by the caller and must outlive the
 * pt_multi; replicas are owned by it.                                              */
typedef struct pt_multi pt_multi;
PT_API pt_multi* pt_multi_create(pt_scene* primary, const int* devices, int n_devices);
PT_API void pt_multi_destroy(pt_multi*);
PT_API int pt_multi_gpus(const pt_multi*);
PT_API pt_scene* pt_multi_scene(pt_multi*, int rank);     /* rank 0 = the primary   */
/* Integrator::Render over all ranks; the frame is the primary scene's film
 * (pt_film_download(primary)).  stats: sums over the ranks, gpu_ms = the slowest.  */
PT_API int pt_multi_render(pt_multi*, const pt_camera_desc*, const pt_render_params*,
                           int clear, pt_render_stats* stats);

/* one process per GPU (torchrun, mpirun): rank 0 calls pt_dist_unique_id and the
 * launcher carries the 128 bytes to the other ranks; every rank then calls
 * pt_dist_init on its committed scene (collective) and pt_dist_render per frame
 * (collective; the frame is rank 0's film).  world == 1 needs no id and no NCCL.   */
PT_API int pt_dist_unique_id(void* id128);
PT_API int pt_dist_init(pt_scene*, const void* id128, int rank, int world);
PT_API int pt_dist_render(pt_scene*, const pt_camera_desc*, const pt_render_params*,
                          int clear, pt_render_stats* stats);
PT_API int pt_dist_finalize(pt_scene*);

/* ---- diagnostic probes of the device build of the shading code (csrc/unit.cu) ------
 * One record per thread through the SAME device functions k_shade calls, compiled with
 * the same flags (fast math): the interaction Shape::Intersect + Primitive::Intersect
 * build (Shape.cpp:205-243, Primitive.cpp:20-24,54-63), Material::{scatter,
 * calc_attenuation, PDF} (Material.hpp:200-673), the AreaLight seen by a hit
 * (Light.cpp:254-291), Light::{sample, PDF, L} (Light.cpp:9-365) and
 * HenyeyGreenstein::{Sample, PDF} (PhaseFunction.cpp:4-26).  Host buffers, synchronous.
 * Record layouts: csrc/unit.cu (those of oracle/ref_driver.cpp ref_unit_*).             */
PT_API int pt_unit_shade(pt_scene*, const pt_ray* rays, size_t n, const float* in6,
                         float* out, int stride /* >= 43 */);
PT_API int pt_unit_light_count(pt_scene*);
PT_API int pt_unit_light_sample(pt_scene*, int light, size_t n, const float* in5,
                                float* out, int stride /* >= 22 */, float time);
PT_API int pt_unit_hg(int device, float g, size_t n, const float* in5, float* out5);
/* A committed scene's traversal tables read back (tests compare the device-packed arrays with the host-packed ones byte for
 * byte): table 0 the nodes of the commit's width (BVH2 64 B, BVH4 128 B, BVH8 256 B), 1 triangles (48 B), 2 quantised clusters (64 B, PT_LAYOUT_FAST), 3 padded triangles
 * (64 B), 4 instance records (96 B); the shading geometry: 5 positions, 6 normals, 7 texcoords, 8 tangents, 9 indices, 10 shading
 * triangles (16 B), 11 emissive ordinals.  *bytes = the table's size; copied only if it fits cap_bytes.                     */
PT_API int pt_unit_table(pt_scene*, int table, void* dst, size_t cap_bytes, size_t* bytes);

#ifdef __cplusplus
}
#endif
#endif /* PT_B200_H */
