// shade_math.cuh — small float3 helpers, the PCG32 stream and the texture
// evaluators shared by the shade kernels (device) and by the commit-time light
// pre-processing (host).  Everything here is __host__ __device__.
#pragma once
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include "../../include/pt_b200.h"
#include "shade_types.h"

#define PT_HDI __host__ __device__ __forceinline__
// The big evaluators (texture, material, light, interaction).  k_shade stalls on instruction fetch (ncu:
// stalled_no_instruction dominates), so code size matters; measured on B200 (Cornell 1024^2 x 16 spp, shade
// stage): IEEE inline 16.3 ms (460 KB of SASS), IEEE calls 22.1, fast-math calls 17.5 (92 KB), fast-math
// inline 11.2 (256 KB).  Calls lose to inlining despite the smaller code (ShadeScene is copied to a
// stack frame, live registers spill around each call), so inline + fast math is the default.
#if !defined(PTB_SHADE_NOINLINE)
#define PT_HDN __host__ __device__ __forceinline__
#elif defined(__CUDACC__)
#define PT_HDN static __host__ __device__ __noinline__
#else
#define PT_HDN static __attribute__((noinline))
#endif

namespace ptb {

constexpr float kPi = 3.14159265358979323846f;
constexpr float kInvPi = 0.31830988618379067154f;

struct V3 { float x, y, z; };
PT_HDI V3 v3(float x, float y, float z){ V3 r; r.x = x; r.y = y; r.z = z; return r; }
PT_HDI V3 v3(float s){ return v3(s, s, s); }
PT_HDI V3 operator+(V3 a, V3 b){ return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
PT_HDI V3 operator-(V3 a, V3 b){ return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
PT_HDI V3 operator-(V3 a){ return v3(-a.x, -a.y, -a.z); }
PT_HDI V3 operator*(V3 a, V3 b){ return v3(a.x * b.x, a.y * b.y, a.z * b.z); }
PT_HDI V3 operator*(V3 a, float s){ return v3(a.x * s, a.y * s, a.z * s); }
PT_HDI V3 operator*(float s, V3 a){ return v3(a.x * s, a.y * s, a.z * s); }
PT_HDI V3 operator/(V3 a, float s){ return v3(a.x / s, a.y / s, a.z / s); }
PT_HDI V3& operator+=(V3& a, V3 b){ a.x += b.x; a.y += b.y; a.z += b.z; return a; }
PT_HDI V3& operator*=(V3& a, V3 b){ a.x *= b.x; a.y *= b.y; a.z *= b.z; return a; }
PT_HDI float dot(V3 a, V3 b){ return a.x * b.x + a.y * b.y + a.z * b.z; }
PT_HDI V3 cross(V3 a, V3 b){ return v3(a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y); }
PT_HDI float length(V3 a){ return sqrtf(dot(a, a)); }
PT_HDI V3 normalize(V3 a){ return a * (1.0f / sqrtf(dot(a, a))); }   // glm::normalize = v * inversesqrt(dot)
PT_HDI bool is_zero(V3 a){ return a.x == 0.0f && a.y == 0.0f && a.z == 0.0f; }
PT_HDI V3 ld3(const float* p){ return v3(p[0], p[1], p[2]); }
PT_HDI float clampf(float x, float lo, float hi){ return fminf(fmaxf(x, lo), hi); }
PT_HDI float lum(V3 c){ return 0.2126f * c.x + 0.7152f * c.y + 0.0722f * c.z; }   // Util.hpp:4-6
PT_HDI V3 reflect3(V3 I, V3 N){ return I - N * (dot(N, I) * 2.0f); }              // glm::reflect
PT_HDI V3 refract3(V3 I, V3 N, float eta){                                         // glm::refract
    const float d = dot(N, I);
    const float k = 1.0f - eta * eta * (1.0f - d * d);
    return k >= 0.0f ? (eta * I - (eta * d + sqrtf(k)) * N) : v3(0.0f);
}

// onb (Onb.hpp:5-24)
struct Onb { V3 a0, a1, a2; };
PT_HDI Onb onb_from_normal(V3 n){
    Onb o; o.a2 = n;
    const V3 up = fabsf(n.x) > 0.9999f ? v3(0, 1, 0) : v3(1, 0, 0);
    o.a1 = normalize(cross(o.a2, up));
    o.a0 = cross(o.a1, o.a2);
    return o;
}
PT_HDI Onb onb_from_frame(V3 ns, V3 tangent){
    Onb o; o.a2 = ns; o.a0 = tangent; o.a1 = cross(o.a2, o.a0);
    return o;
}
PT_HDI V3 to_world(const Onb& o, V3 v){ return v.x * o.a0 + v.y * o.a1 + v.z * o.a2; }
PT_HDI V3 to_local(const Onb& o, V3 v){ return v3(dot(v, o.a0), dot(v, o.a1), dot(v, o.a2)); }

// ---- PCG32 (pcg32 = setseq_xsh_rr_64_32, pcg-cpp/pcg_random.hpp:1866): one stream per camera sample ----
struct Pcg { uint64_t state, inc; };
PT_HDI uint32_t pcg_next(Pcg& g){
    const uint64_t old = g.state;
    g.state = old * 6364136223846793005ULL + g.inc;
    const uint32_t xorshifted = (uint32_t)(((old >> 18u) ^ old) >> 27u);
    const uint32_t rot = (uint32_t)(old >> 59u);
    return (xorshifted >> rot) | (xorshifted << ((32u - rot) & 31u));
}
// One generator per camera sample.  PCG streams that share a seed and differ only in a small increment
// are strongly correlated (the states stay an affine function of each other), which showed up as a
// per-pixel variance ~3x what the sample variance predicts; so both the start state and the increment
// come out of a 64-bit finaliser (splitmix64) of (seed, sample stream).
PT_HDI uint64_t mix64(uint64_t x){
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}
PT_HDI uint64_t pcg_inc(uint64_t seed, uint64_t stream){ return (mix64(mix64(stream) ^ (seed * 0xD1B54A32D192ED03ull)) << 1u) | 1u; }
PT_HDI Pcg pcg_seed(uint64_t seed, uint64_t stream){
    Pcg g; g.state = 0; g.inc = pcg_inc(seed, stream);
    pcg_next(g); g.state += mix64(mix64(seed) + stream); pcg_next(g);
    return g;
}
// RandomFloatx8 lane (Random.hpp:22-44): float(u32) * (1/0xFFFFFFFF) — may return exactly 1
PT_HDI float rnd_x8(Pcg& g){ return (float)pcg_next(g) * (1.0f / 4294967295.0f); }
// random_float (Random.hpp:16-20): uniform_real_distribution<float> over pcg32, in [0,1)
PT_HDI float rnd_float(Pcg& g){
    const float r = (float)pcg_next(g) * 2.3283064365386963e-10f;
    return r >= 1.0f ? 0.99999994f : r;
}
// random_double (Random.hpp:10-14): 53 random bits in [0,1)
PT_HDI double rnd_double(Pcg& g){
    const uint64_t hi = pcg_next(g), lo = pcg_next(g);
    return (double)(((hi << 32) | lo) >> 11) * (1.0 / 9007199254740992.0);
}

// ---- StratifiedSampler helpers (Util.hpp:45-73, 75-168) ----------------------------------------
// Hash(px, py, dimension): MurmurHash64A over {uint32 px, uint32 py, uint64 dimension}, seed 0
PT_HDI uint64_t hash_ppd(uint32_t px, uint32_t py, uint64_t dim){
    const uint64_t m = 0xc6a4a7935bd1e995ull;
    uint64_t h = 0ull ^ (16ull * m);
    uint64_t k = (uint64_t)px | ((uint64_t)py << 32);
    k *= m; k ^= k >> 47; k *= m; h ^= k; h *= m;
    k = dim;
    k *= m; k ^= k >> 47; k *= m; h ^= k; h *= m;
    h ^= h >> 47; h *= m; h ^= h >> 47;
    return h;
}
PT_HDI uint32_t permutation_element(uint32_t i, uint32_t l, uint32_t p){
    uint32_t w = l - 1;
    w |= w >> 1; w |= w >> 2; w |= w >> 4; w |= w >> 8; w |= w >> 16;
    do{
        i ^= p; i *= 0xe170893du; i ^= p >> 16; i ^= (i & w) >> 4; i ^= p >> 8; i *= 0x0929eb3fu; i ^= p >> 23;
        i ^= (i & w) >> 1; i *= 1u | p >> 27; i *= 0x6935fa69u; i ^= (i & w) >> 11; i *= 0x74dcb303u;
        i ^= (i & w) >> 2; i *= 0x9e501cc3u; i ^= (i & w) >> 2; i *= 0xc860a3dfu; i &= w; i ^= i >> 5;
    } while(i >= l);
    return (i + p) % l;
}

// ---- textures (Texture.hpp:108-231, Texture.cpp:38-94) ------------------------------------
PT_HDI int wrap_index(int i, int n){ int m = i % n; if(m < 0)m += n; return m; }

PT_HDI float texel_ch(const ShadeScene& S, const DevTexture& t, int x, int y, int ch){
    x = wrap_index(x, t.w); y = wrap_index(y, t.h);
    const size_t o = ((size_t)y * t.w + x) * t.ch + ch;
    if(t.kind == PT_TEX_IMAGE_U8)return (float)S.texels[t.texels + o] / 255.0f;
    return reinterpret_cast<const float*>(S.texels + t.texels)[o];
}

// ImageTexture::Evaluate / FloatImageTexture::Evaluate: bilinear over the four texels around
// uv * res - 0.5 with repeat wrap.  GetChannelAt(p, 1..3): channels beyond the image's own count read
// neighbouring bytes in the reference; synthetic textures always carry >= 3 channels.
PT_HDI V3 tex_image(const uint8_t* texels, const DevTexture& t, float u, float v){
    const float x = u * t.w - 0.5f, y = v * t.h - 0.5f;
    const int xi = (int)floorf(x), yi = (int)floorf(y);
    const float dx = x - xi, dy = y - yi;
    const int x0 = wrap_index(xi, t.w), x1 = x0 + 1 == t.w ? 0 : x0 + 1;
    const int y0 = wrap_index(yi, t.h), y1 = y0 + 1 == t.h ? 0 : y0 + 1;
    const int c1 = t.ch > 1 ? 1 : 0, c2 = t.ch > 2 ? 2 : 0;
    const size_t o00 = ((size_t)y0 * t.w + x0) * t.ch, o10 = ((size_t)y0 * t.w + x1) * t.ch;
    const size_t o01 = ((size_t)y1 * t.w + x0) * t.ch, o11 = ((size_t)y1 * t.w + x1) * t.ch;
    V3 a, b, c, d;
    if(t.kind == PT_TEX_IMAGE_U8){
        const uint8_t* px = texels + t.texels;
        a = v3(px[o00] / 255.0f, px[o00 + c1] / 255.0f, px[o00 + c2] / 255.0f);
        b = v3(px[o10] / 255.0f, px[o10 + c1] / 255.0f, px[o10 + c2] / 255.0f);
        c = v3(px[o01] / 255.0f, px[o01 + c1] / 255.0f, px[o01 + c2] / 255.0f);
        d = v3(px[o11] / 255.0f, px[o11 + c1] / 255.0f, px[o11 + c2] / 255.0f);
    } else{
        const float* px = reinterpret_cast<const float*>(texels + t.texels);
        a = v3(px[o00], px[o00 + c1], px[o00 + c2]); b = v3(px[o10], px[o10 + c1], px[o10 + c2]);
        c = v3(px[o01], px[o01 + c1], px[o01 + c2]); d = v3(px[o11], px[o11 + c1], px[o11 + c2]);
    }
    return ((1 - dx) * (1 - dy)) * a + (dx * (1 - dy)) * b + ((1 - dx) * dy) * c + (dx * dy) * d;
}

// Texture::Evaluate(interaction): uv and (for NormalTexture) the shading normal
// One real function on the device (PT_TEX_CALL): inlined at its ~12 call sites it was most of k_shade's
// code.  It takes the two table pointers by value so that no ShadeScene copy is needed for the call.
#if defined(__CUDACC__) && !defined(PTB_TEX_INLINE)
#define PT_TEX_CALL static __host__ __device__ __noinline__
#else
#define PT_TEX_CALL PT_HDI
#endif
PT_TEX_CALL V3 tex_eval_tables(const DevTexture* textures, const uint8_t* texels, int id, float u, float v, V3 ns){
    V3 scale = v3(1.0f);
#pragma unroll 1
    for(int guard = 0; guard < 8; guard++){
        const DevTexture& t = textures[id];
        const V3 sc = scale * ld3(t.scale);
        switch(t.kind){
        case PT_TEX_SOLID: return sc * ld3(t.color);
        case PT_TEX_IMAGE_U8: case PT_TEX_IMAGE_F32: return sc * tex_image(texels, t, u, v);
        case PT_TEX_UV: return sc * v3(u, v, 0.0f);
        case PT_TEX_NORMAL: return sc * ns;
        case PT_TEX_CHECKER: {
            const int iu = (int)floorf(u * t.inv_scale[0]), iv = (int)floorf(v * t.inv_scale[1]);
            id = ((iu + iv) % 2 == 0) ? t.tex_a : t.tex_b;
            scale = sc;
            break;
        }
        default: return v3(0.0f);
        }
    }
    return v3(0.0f);
}
PT_HDI V3 tex_eval(const ShadeScene& S, int id, float u, float v, V3 ns){ return tex_eval_tables(S.textures, S.texels, id, u, v, ns); }

// Texture::alpha(uv) (Texture.cpp:48-94): 1 unless a 4-channel image
PT_HDN float tex_alpha(const ShadeScene& S, int id, float u, float v){
    for(int guard = 0; guard < 8; guard++){
        const DevTexture& t = S.textures[id];
        if(t.kind == PT_TEX_CHECKER){
            const int iu = (int)floorf(u * t.inv_scale[0]), iv = (int)floorf(v * t.inv_scale[1]);
            id = ((iu + iv) % 2 == 0) ? t.tex_a : t.tex_b;
            continue;
        }
        if((t.kind != PT_TEX_IMAGE_U8 && t.kind != PT_TEX_IMAGE_F32) || t.ch != 4)return 1.0f;
        const float x = u * t.w - 0.5f, y = v * t.h - 0.5f;
        const int xi = (int)floorf(x), yi = (int)floorf(y);
        const float dx = x - xi, dy = y - yi;
        return (1 - dx) * (1 - dy) * texel_ch(S, t, xi, yi, 3) + dx * (1 - dy) * texel_ch(S, t, xi + 1, yi, 3) +
               (1 - dx) * dy * texel_ch(S, t, xi, yi + 1, 3) + dx * dy * texel_ch(S, t, xi + 1, yi + 1, 3);
    }
    return 1.0f;
}

// SphereShape::GetSphereUV (Shape.hpp:35-43)
PT_HDI void sphere_uv(V3 p, float& u, float& v){
    p = normalize(p);
    const float theta = acosf(clampf(p.y, -1.0f, 1.0f));
    float phi = atan2f(p.z, p.x);
    if(phi < 0.0f)phi += 2.0f * kPi;
    u = kInvPi * phi * 0.5f; v = kInvPi * theta;
}

// uniform direction on the sphere used by every infinite light's sample() (Light.cpp:35-42)
PT_HDI V3 sphere_dir(float ux, float uy){
    const float z = 2.0f * ux - 1.0f;
    const float theta = 2.0f * kPi * uy;
    const float r = sqrtf(1.0f - z * z);
    return v3(r * cosf(theta), r * sinf(theta), z);
}

// InfiniteLight::Le(ray) for the three infinite kinds
PT_HDI V3 inf_Le(const ShadeScene& S, const DevLight& L, V3 dir){
    if(L.kind == DL_UNIFORM_INF)return ld3(L.color);
    if(L.kind == DL_GRADIENT_INF){
        if(L.tex >= 0){     // a FunctionInfiniteLight whose sky function the host baked into a lat-long table (host/PathTracing.hpp)
            float u, v;
            sphere_uv(dir, u, v);
            return tex_eval(S, L.tex, u, v, v3(0.0f));
        }
        const float a = 0.5f * (dir.y + 1.0f);
        return L.scale * ((1.0f - a) * ld3(L.color) + a * ld3(L.color_b));
    }
    float u, v;
    sphere_uv(dir, u, v);
    return L.scale * tex_eval(S, L.tex, u, v, v3(0.0f));
}

} // namespace ptb
