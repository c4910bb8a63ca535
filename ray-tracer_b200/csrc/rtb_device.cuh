// rtb_device.cuh -- device-side data layout and per-ray arithmetic of the B200 path tracer.
//
// The arithmetic here follows the reference's expressions term by term (same operand order, same
// float/double types, same libm entry points) because parity is defined on its results:
// primary-hit ids bit-exact, radiance at matched RNG streams.  What differs is
// everything around the arithmetic: one LBVH over all primitives instead of a linear object scan,
// SoA float4 records instead of 5 KB objects, a material table instead of per-primitive copies.
// Reference file:line citations are relative to /root/reference/src.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace rtb {

// per-lane traversal stack entries (local memory); rtb_scene_commit refuses trees that would need more
#ifndef RTB_STACK
#define RTB_STACK 96
#endif

// -DRTB_CHECKS: device-side asserts on every queue / slot / stack index (the stand-in for compute-sanitizer,
// which is closed on this GPU pool); a failed assert traps the kernel and the call returns an error
#ifdef RTB_CHECKS
#include <cassert>
#define RTB_ASSERT(x) assert(x)
#else
#define RTB_ASSERT(x)
#endif

#define RTB_INF_F 1073741824.0f /* INF = 1 << 31 - 1 == 1 << 30, objects.cu:6 */
#define RTB_EPS 0.000001f      /* FLOAT_PRECISION_ERROR, objects.cu:7 */

// leaf reference kinds (low 3 bits of an encoded leaf id).  A quad is two consecutive triangles of
// the same array; the one-way kinds carry OneWayQuad's direction test (objects.cu:273-280).
enum { KIND_TRI = 0, KIND_QUAD = 1, KIND_SPHERE = 2, KIND_QUAD_ONEWAY = 3, KIND_QUAD_ONEWAY_INV = 4 };
#define RTB_KIND_BITS 3

// ---- material table entry (64 B): Material + Texture of material.cu:4-186, flattened -----------
struct DMaterial {
    int type;          // Material::STANDARD / EMISSIVE / REFRACTIVE
    int tex_type;      // Texture::COLOUR / GRADIENT / CHECKERBOARD / IMAGE
    float smoothness;
    float refractive_index;
    float a[3];        // COLOUR: colour; CHECKERBOARD: light
    int num_squares;
    float b[3];        // CHECKERBOARD: dark; EMISSIVE: emitted_light
    int need_uv;
    int img_w, img_h;
    int img_offset;    // first float of the image in DScene.image_rgb
    int pad;
};

// ---- device scene (all pointers device memory, 16-byte aligned) --------------------------------
#define RTB_MAX_LARGE 32  // the candidates of a ray are a 32-bit mask

struct DScene {
    // LBVH: internal node i = 4 float4: {lo0.xyz, ref0} {hi0.xyz, -} {lo1.xyz, ref1} {hi1.xyz, -}
    // ref >= 0: internal node index; ref < 0: leaf, ~ref = (index << 3) | kind
    const float4 *nodes;
    int num_prims;       // all primitives (leaves of the LBVH + large primitives)
    int num_bvh_prims;   // leaves of the LBVH
    int root_ref;        // encoded like a child ref (a 1-primitive tree has a leaf root)
    float4 root_node[4]; // copy of the root node: lives in the kernel parameters (uniform loads)
    // Large primitives (walls, floors: boxes that span a big share of the scene) are kept out of
    // the tree and tested first, in a warp-uniform loop; their hits bound the tree traversal.
    const int *large_refs;
    const float4 *large_boxes;  // 2 float4 per large primitive: padded lo, hi
    float4 large_box_c[2 * RTB_MAX_LARGE];  // the same boxes in the kernel parameters (constant bank: operands, not loads)
    int num_large;
    unsigned int large_sphere_mask;  // bit i: large primitive i is a sphere
    // triangles, 3 float4 each: {p0.xyz, s1.x} {s1.yz, s2.xy} {s2.z, n.xyz}.  KIND_TRI leaves
    // name one triangle, KIND_QUAD* leaves the first of two consecutive ones (t1, t2 of a Quad).
    const float4 *tris;
    const int2 *tri_ids;     // {object index, sub id}
    const float2 *tri_uv;    // 3 per triangle (Triangle::texture_points)
    const int *tri_refleaf;  // reference-BVH leaf node of a mesh triangle, -1 otherwise
    // spheres (KIND_SPHERE): {center.xyz, radius}
    const float4 *spheres;
    const int *sphere_ids;   // object index
    // per object
    const int *obj_material;
    const DMaterial *materials;
    const float *image_rgb;
    // reference-BVH boxes for the degenerate-case path check (see ref_path_hit): per node
    // 2 float4 {lo.xyz, parent} {hi.xyz, -}
    const float4 *ref_nodes;
    // brute-force list (option "traversal" = 1): every leaf ref in creation order
    const int *all_refs;
    int num_tris, num_spheres, num_objects;
};

// Screen-tile sharding with the exchange fused into the render kernels: a finished pixel is stored into the
// accumulation image of EVERY GPU of the job (peer memory over NVLink / NVSwitch: plain global stores through
// cudaDeviceEnablePeerAccess or CUDA IPC mappings), so the frame needs no gather afterwards -- the transfer rides
// along with the rendering, 12 bytes per peer per finished pixel.  n == 0: single GPU, store to `out` only.
#define RTB_MAX_PEERS 8
struct PeerOut {
    float *image[RTB_MAX_PEERS];  // every rank's accumulation image, this rank's own included
    int n;
};

struct DCamera {  // == DeviceCamData, camera.cu:12-18
    float cam_pos[3], tl[3], du[3], dv[3];
};

struct DRender {  // == RenderData, raytracer.cu:4-12
    int rays_per_pixel, reflection_limit, antialias;
    float sky[3];
};

// ---- Vec3 (utils.cu:13-163): same expression order as the reference's operators ----------------
struct V3 {
    float x, y, z;
};
__device__ __forceinline__ V3 v3(float x, float y, float z) { V3 r; r.x = x; r.y = y; r.z = z; return r; }
__device__ __forceinline__ V3 operator+(V3 a, V3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
__device__ __forceinline__ V3 operator-(V3 a, V3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
__device__ __forceinline__ V3 operator*(V3 a, float s) { return v3(a.x * s, a.y * s, a.z * s); }
__device__ __forceinline__ V3 operator*(V3 a, V3 b) { return v3(a.x * b.x, a.y * b.y, a.z * b.z); }
__device__ __forceinline__ V3 operator/(V3 a, float s) { return v3(a.x / s, a.y / s, a.z / s); }
// utils.cu:132-138: three products, then left-to-right sum
__device__ __forceinline__ float dot(V3 a, V3 b) {
    float nx = a.x * b.x, ny = a.y * b.y, nz = a.z * b.z;
    return nx + ny + nz;
}
// utils.cu:148-155
__device__ __forceinline__ V3 cross(V3 a, V3 b) {
    return v3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
// utils.cu:120-130: v * (1 / sqrt(x*x + y*y + z*z)), IEEE sqrt and divide
__device__ __forceinline__ V3 normalised(V3 a) {
    float mag_sq = a.x * a.x + a.y * a.y + a.z * a.z;
    float inv_mag = 1 / sqrtf(mag_sq);
    return v3(a.x * inv_mag, a.y * inv_mag, a.z * inv_mag);
}

// ---- RNG (utils.cu:220-239) ----------------------------------------------------------------------
__device__ __forceinline__ uint32_t rng_seed(int array_index, int time_ms) {
    return (uint32_t)array_index * 3145739u + (uint32_t)time_ms * 6291469u;  // raytracer.cu:127
}

// result / 4294967295.0 rounded to float (utils.cu:228).  RTB_RNG_DIV keeps the literal FP64
// divide; the default evaluates the same correctly-rounded double quotient with one FMA:
//   r / (2^32 - 1) = r * 2^-32 * (1 + 2^-32 + 2^-64 + ...),  x = r * 2^-32 exact,
//   fma(x, 2^-32, x) = RN(x + x * 2^-32), and the neglected tail (< 2^-64 relative) never moves
//   the double rounding because x + x*2^-32 has at most 64 significant bits whose low part is far
//   from a double rounding boundary -- verified exhaustively for all 2^32 inputs by
//   tests/rng_exhaustive.c (run by tests/test_cpu_oracle.py) before being enabled.
__device__ __forceinline__ float rng_to_float(uint32_t r) {
#ifdef RTB_RNG_FMA
    double x = (double)r * 2.3283064365386963e-10;  // 2^-32, exact
    return (float)fma(x, 2.3283064365386963e-10, x);
#else
    return (float)((double)r / 4294967295.0);
#endif
}

__device__ __forceinline__ float rng_next(uint32_t &state) {
    uint32_t ns = state * 747796405u + 2891336453u;
    state = ns;
    uint32_t r = ((ns >> ((ns >> 28) + 4)) ^ ns) * 277803737u;
    r = (r >> 22) ^ r;
    return rng_to_float(r);
}

// Box-Muller, utils.cu:234-239: theta in double (2 * 3.14159 * u), log/sqrt/cos in float
__device__ __forceinline__ float rng_normal(uint32_t &state) {
    float theta = (float)(2 * 3.14159 * (double)rng_next(state));
    float rho = sqrtf(-2 * logf(rng_next(state)));
    return rho * cosf(theta);
}

}  // namespace rtb
