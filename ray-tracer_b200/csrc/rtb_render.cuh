// rtb_render.cuh -- the render kernels.  Included by rtb_render_fast.cu (default -fmad=true, like
// the reference's own nvcc build) and rtb_render_strict.cu (-fmad=false: + - * / sqrt round exactly
// as written, which makes primary-hit ids and distances bit-identical to the host oracle).
// RTB_NS names the variant's namespace.
#pragma once
#include "rtb_trace.cuh"
#include "rtb_launch.h"

namespace rtb {
namespace RTB_NS {

// float(3.141592653589793), camera.cu:9
#define RTB_PI_F 3.14159274101257324f

struct Surface {
    V3 point, normal;
    float u, v;
    int material;
};

// RayHitData of the winning primitive (objects.cu:55-68,154-160) + the object's material index
__device__ __forceinline__ void make_surface(const DScene &s, const Ray &r, const Best &b, Surface &sf) {
    int idx = b.code >> 1;
    sf.point = r.d * b.t + r.o;  // Ray::get_pos, ray.cu:63-65
    sf.u = 0; sf.v = 0;
    int obj;
    if (b.code & 1) {
        float4 sp = __ldg(s.spheres + idx);
        V3 c = v3(sp.x, sp.y, sp.z);
        sf.normal = normalised(sf.point - c);
        obj = __ldg(s.sphere_ids + idx);
        sf.material = __ldg(s.obj_material + obj);
        if (__ldg(&s.materials[sf.material].need_uv)) {  // assign_texture_coords, objects.cu:82-97
            float theta = asinf((sf.point.y - c.y) / sp.w);
            float phi = acosf((sf.point.x - c.x) / sp.w);
            float u = (theta + RTB_PI_F / 2) / RTB_PI_F;
            float v_ratio = (1 - phi / RTB_PI_F) / 2;
            int behind = sf.point.z > c.z;
            int mult = 1 - 2 * behind;
            sf.u = u;
            sf.v = 1 * behind + mult * v_ratio;
        }
        return;
    }
    const float4 *T = s.tris + 3 * idx;
    obj = __ldg(s.tri_ids + idx).x;
    float4 c = __ldg(T + 2);
    V3 n = v3(c.y, c.z, c.w);
    sf.normal = n * (float)(1 - 2 * (dot(n, r.d) > 0));  // objects.cu:158
    sf.material = __ldg(s.obj_material + obj);
    if (__ldg(&s.materials[sf.material].need_uv)) {
        // assign_texture_coords(hit, w, u, v): tp0*w + tp1*u + tp2*v, objects.cu:160,196-199
        float w = 1 - b.u - b.v;
        float2 t0 = __ldg(s.tri_uv + 3 * idx), t1 = __ldg(s.tri_uv + 3 * idx + 1), t2 = __ldg(s.tri_uv + 3 * idx + 2);
        sf.u = (t0.x * w + t1.x * b.u) + t2.x * b.v;
        sf.v = (t0.y * w + t1.y * b.u) + t2.y * b.v;
    }
}

__device__ __forceinline__ void change_direction(Ray &r, V3 d) {  // ray.cu:198-202
    r.d = d;
    r.inv = v3(1 / d.x, 1 / d.y, 1 / d.z);
}

// Ray::apply_antialias, ray.cu:130-142: (u - 0.5) * 2 * RANGE in double, RANGE = float(0.001)
__device__ __forceinline__ void apply_antialias(Ray &r, uint32_t &rng) {
    const double range = (double)0.001f;
    V3 off;
    off.x = (float)(((double)rng_next(rng) - 0.5) * 2 * range);
    off.y = (float)(((double)rng_next(rng) - 0.5) * 2 * range);
    off.z = (float)(((double)rng_next(rng) - 0.5) * 2 * range);
    change_direction(r, normalised(r.d + off));
}

// The reference recomputes direction_inv at every change of direction (ray.cu:198-202), but only
// the box tests read it, and only for the ray as it enters get_ray_collision: scatter and
// sample start set the direction alone and the inverse is taken once, here, after the antialias
// jitter (which is itself a change_direction when enabled).  Same IEEE divides, same values.
__device__ __forceinline__ void ray_finalise(Ray &r, bool aa, uint32_t &rng) {
    if (aa) apply_antialias(r, rng);
    else change_direction(r, r.d);
}

// Ray::reflect, ray.cu:67-75 with diffuse_reflect :157-170, true_lambertian_reflect :172-178,
// perfect_reflect :180-186, lerp :32-34.  The three Box-Muller draws (logf + cosf each) run as a
// rolled loop and the function has ONE call site (path_shade): the shading code is the larger half
// of the instruction footprint, and the SM's instruction cache (32 KB) is what the kernels that
// hold both shading and traversal run out of first.
__device__ __forceinline__ void reflect(Ray &r, const Surface &sf, float smoothness, uint32_t &rng) {
    float g0 = 0, g1 = 0, g2 = 0;  // draw order x, y, z (ray.cu:159-161): g2 ends up holding the first draw
#pragma unroll 1
    for (int k = 0; k < 3; k++) {
        g2 = g1; g1 = g0;
        g0 = rng_normal(rng);
    }
    V3 rv = v3(g2, g1, g0);
    if (dot(rv, sf.normal) < 0) rv = rv * -1.f;
    rv = normalised(rv);
    V3 diffuse = normalised(sf.normal + rv);
    float dn = dot(r.d, sf.normal);
    V3 spec = normalised(r.d - sf.normal * 2.f * dn);
    r.o = sf.point;
    r.d = normalised(diffuse + (spec - diffuse) * smoothness);  // direction_inv: see ray_finalise
}

// Ray::refract, ray.cu:77-128 + get_reflection_coeff :188-196.  Returns true when the ray reflects instead
// (past the critical angle, or the Schlick coefficient wins the draw): the caller then runs reflect(), exactly
// where the reference calls it (ray.cu:108-111), so the draw order is unchanged.
static __device__ __noinline__ bool refract(Ray &r, const Surface &sf, float mat_n, float &cur_n, uint32_t &rng) {
    float n1, n2;
    V3 refn;
    if (dot(sf.normal, r.d) > 0) {
        n1 = mat_n; n2 = cur_n; refn = sf.normal;
    } else {
        n1 = cur_n; n2 = mat_n; refn = sf.normal * -1.f;
    }
    cur_n = n2;
    float theta1 = acos(min(dot(r.d, refn), 1.0));           // double overloads, as in the reference
    float theta2 = asin(min(n1 * sin(theta1) / n2, 1.0));
    float critical = asin(n2 / n1);
    float sqrt_r0 = (n1 - n2) / (n1 + n2);
    float r0 = sqrt_r0 * sqrt_r0;
    float cos_theta = cos(theta1);
    float coeff = r0 + (1 - r0) * pow((1 - cos_theta), 5);
    if (theta1 > critical || coeff > rng_next(rng)) return true;
    V3 perp;
    if (theta1 != 0) perp = (r.d - refn * cos(theta1)) / sin(theta1);
    else perp = v3(0, 0, 0);
    V3 res = refn * cos(theta2) + perp * sin(theta2);
    r.o = sf.point;
    r.d = normalised(res);  // direction_inv: see ray_finalise
    return false;
}

// Texture::get_texture_colour, material.cu:53-124
__device__ __forceinline__ V3 texture_colour(const DScene &s, const DMaterial &m, float u, float v) {
    switch (m.tex_type) {
        case 0: return v3(m.a[0], m.a[1], m.a[2]);
        case 1: return v3(u, v, 0);
        case 2: {
            int uc = u * m.num_squares;
            int vc = v * m.num_squares;
            if ((uc + vc) % 2 == 0) return v3(m.a[0], m.a[1], m.a[2]);
            return v3(m.b[0], m.b[1], m.b[2]);
        }
        case 3: {
            int uc = (m.img_w - 1) * u;
            int vc = (m.img_h - 1) * v;
            int idx = vc * m.img_w + uc;
            idx = max(0, min(idx, m.img_w * m.img_h - 1));  // the reference reads out of bounds here
            const float *px = s.image_rgb + m.img_offset + 3 * idx;
            return v3(__ldg(px), __ldg(px + 1), __ldg(px + 2));
        }
    }
    return v3(0, 0, 0);
}

// cam_pixel_to_world (camera.cu:24-29) + Ray::set_direction_origin (ray.cu:147-155)
__device__ __forceinline__ V3 primary_direction(const DCamera &c, int x, int y) {
    V3 du = v3(c.du[0], c.du[1], c.du[2]), dv = v3(c.dv[0], c.dv[1], c.dv[2]);
    V3 plane_point = du * (float)x + dv * (float)y;
    V3 view_pos = v3(c.tl[0], c.tl[1], c.tl[2]) + plane_point;
    V3 o = v3(c.cam_pos[0], c.cam_pos[1], c.cam_pos[2]);
    return normalised(view_pos - o);
}

template <bool BRUTE> __device__ __forceinline__ Best trace(const DScene &s, const Ray &r, unsigned mask, Counters &cnt) {
    if (BRUTE) return trace_brute(s, r, cnt);
    return trace_bvh(s, r, mask, cnt);
}

// local row index -> global row for the row-cyclic band sharding (SURVEY.md 8(e) partitioning A)
__device__ __forceinline__ int shard_row(int rank, int world, int band_rows, int local_row) {
    if (world <= 1) return local_row;
    int band = local_row / band_rows, within = local_row - band * band_rows;
    return (band * world + rank) * band_rows + within;
}

// ---- per-pixel path state (what the reference keeps on the thread's stack across the spp and
// bounce loops, raytracer.cu:67-113) -- shared by the megakernel and the wavefront kernels so that
// both evaluate the same expressions in the same order ----------------------------------------
struct Path {
    uint32_t rng;      // per pixel, carried from sample to sample (raytracer.cu:127-129)
    int sample, bounce;
    V3 thr, col, acc;  // ray_colour, final_colour of trace_ray (:68-69); sum over samples (:102-105)
    float cur_n;       // Ray::current_refractive_index
};

struct FrameConsts {
    int spp, limit;
    bool aa;
    V3 sky, cam_o;
};

__device__ __forceinline__ FrameConsts frame_consts(const DRender &rd, const DCamera &cam) {
    FrameConsts f;
    f.spp = rd.rays_per_pixel; f.limit = rd.reflection_limit; f.aa = rd.antialias != 0;
    f.sky = v3(rd.sky[0], rd.sky[1], rd.sky[2]);
    f.cam_o = v3(cam.cam_pos[0], cam.cam_pos[1], cam.cam_pos[2]);
    return f;
}

// raytracer.cu:74-91 for a ray whose closest hit is `best`: sky on a miss, else scatter + light /
// throughput; closes the sample when the path ends.  Returns true when the segment hit something.
__device__ __forceinline__ bool path_shade(const DScene &s, const FrameConsts &f, Ray &ray, const Best &best, Path &P) {
    bool end_sample, hit = best.code >= 0;
    if (!hit) {
        P.col = P.col + f.sky * P.thr;
        end_sample = true;
    } else {
        Surface sf;
        make_surface(s, ray, best, sf);
        const DMaterial &m = s.materials[sf.material];
        int mtype = __ldg(&m.type);
        bool reflects = mtype == 0 || mtype == 1;  // update_ray, raytracer.cu:49-64
        if (mtype == 2) reflects = refract(ray, sf, __ldg(&m.refractive_index), P.cur_n, P.rng);
        if (reflects) reflect(ray, sf, __ldg(&m.smoothness), P.rng);
        if (mtype == 1) P.col = P.col + v3(__ldg(&m.b[0]), __ldg(&m.b[1]), __ldg(&m.b[2])) * P.thr;
        else P.thr = P.thr * texture_colour(s, m, sf.u, sf.v);
        P.bounce++;
        end_sample = P.bounce >= f.limit;
    }
    if (end_sample) {
        P.acc = P.acc + P.col;
        P.sample++;
        P.bounce = 0;
    }
    return hit;
}

__device__ __forceinline__ void path_begin_pixel(Path &P, int array_index, int time_ms) {
    P.rng = rng_seed(array_index, time_ms);
    P.acc = v3(0, 0, 0);
    P.sample = 0;
    P.bounce = 0;
}

// get_ray_colour's tail, raytracer.cu:107-112: mean over the samples, then the progressive mean
__device__ __forceinline__ void path_write_pixel(const Path &P, int spp, int frame_num, const float *prev, float *out, const PeerOut &peers, int array_index) {
    V3 colour = P.acc / (float)spp;
    V3 previous = v3(prev[array_index], prev[array_index + 1], prev[array_index + 2]);
    V3 previous_sum = previous * (float)frame_num;
    V3 res = (colour + previous_sum) / (float)(frame_num + 1);
    if (peers.n == 0) {
        out[array_index] = res.x;
        out[array_index + 1] = res.y;
        out[array_index + 2] = res.z;
        return;
    }
    // fused exchange of the tile sharding: the pixel goes to every GPU's image (this GPU's own is one of them)
    for (int r = 0; r < peers.n; r++) {
        float *img = peers.image[r];
        img[array_index] = res.x;
        img[array_index + 1] = res.y;
        img[array_index + 2] = res.z;
    }
}

// trace_ray(ray) starts from a copy of the primary ray, raytracer.cu:67-69,103
__device__ __forceinline__ void path_begin_sample(const FrameConsts &f, V3 pdir, Ray &ray, Path &P) {
    ray.o = f.cam_o;
    ray.d = pdir;  // direction_inv: see ray_finalise
    P.cur_n = 1;
    P.thr = v3(1, 1, 1);
    P.col = v3(0, 0, 0);
}

// ---------------------------------------------------------------------------------------------
// Persistent megakernel with dynamic fetch.
//
// One lane owns one pixel at a time and runs the reference's per-pixel function (get_pixel_colour
// -> get_ray_colour -> trace_ray, raytracer.cu:67-136).  The samples of a pixel stay serial in
// their lane because the RNG state is per pixel and carried from sample to sample (SURVEY.md
// F10/H1).  Pixels are handed out in 8x4 tiles through one atomic counter.
//
// The warp alternates between two phases so that neither runs with mostly idle lanes:
//   SHADE    lanes whose ray has finished shade the hit, advance their sample / bounce / pixel
//            state machine, generate the next ray (antialias jitter) and test it against the large
//            primitives;
//   TRAVERSE lanes with a live ray walk the LBVH (resumable state in `Trav`).
// A lane that finishes its ray early waits only until `p.refill` lanes of the warp are waiting
// (or nobody traverses any more); then the SHADE phase runs for all of them at once while the long
// rays keep their traversal state, and everybody re-enters TRAVERSE together.
// ---------------------------------------------------------------------------------------------
// node steps between two ballots of the megakernel's NODE phase.  Measured with 7 resident blocks per SM (72 registers):
// 2 steps -3.5 % on the 1 M-triangle tile share, -2 % on the Cornell scenes, +-0 on the cube scene; 3 steps +13 % on
// the cube scene, 4 steps spill.  (With 6 blocks and 80 registers, round 1, two steps were 4 % slower.)
#ifndef RTB_MEGA_NODE_STEPS
#define RTB_MEGA_NODE_STEPS 2
#endif
template <bool BRUTE, bool COUNT, bool EARLY_POP>
__global__ void __launch_bounds__(RTB_BLOCK_THREADS, RTB_MIN_BLOCKS) render_megakernel(const RenderParams p) {
    const DScene &s = p.scene;
    const FrameConsts f = frame_consts(p.rd, p.cam);
    const int tiles_x = (p.width + 7) >> 3, tiles_y = (p.local_rows + 3) >> 2;
    const unsigned int total = (unsigned int)tiles_x * tiles_y * 32u;
    const unsigned int FULL = 0xffffffffu;
    const int refill = p.refill, walk_min = p.walk_min;

    Counters cnt;
    cnt = Counters{};
    unsigned long long n_seg = 0, n_hit = 0, n_samples = 0;

    // lane state
    enum { IDLE = 0, WAITING = 1, TRAVERSING = 2 };
    int mode = WAITING;    // WAITING: needs the SHADE phase; IDLE: out of work
    bool pending = false;  // a finished ray whose hit has not been shaded yet
    bool fresh = false;    // a new ray that has not entered the tree yet
    int array_index = -1;  // (y * W + x) * 3 of the pixel this lane owns, -1 = none
    Path P;
    P.rng = 0; P.sample = 0; P.bounce = 0; P.cur_n = 1;
    P.thr = v3(1, 1, 1); P.col = v3(0, 0, 0); P.acc = v3(0, 0, 0);
    V3 pdir = v3(0, 0, 1);
    Ray ray;
    ray.o = f.cam_o; ray.d = pdir; ray.inv = pdir;
    Trav T;
    T.cur = RTB_DONE; T.leaf = 0; T.sp = 0; T.top = RTB_DONE;
    T.best.t = RTB_INF_F; T.best.u = 0; T.best.v = 0; T.best.code = -1;

    for (;;) {
        unsigned int waiting = __ballot_sync(FULL, mode == WAITING);
        unsigned int traversing = __ballot_sync(FULL, mode == TRAVERSING);
        if (!waiting && !traversing) break;  // every lane is out of work
        if (waiting && (!traversing || __popc(waiting) >= refill)) {
            // ------------------------------- SHADE phase -------------------------------------
            if (mode == WAITING) {
                if (pending) {  // raytracer.cu:74-91 for the ray that just finished
                    pending = false;
                    bool hit = path_shade(s, f, ray, T.best, P);
                    if (COUNT) { n_seg++; n_hit += hit; }
                }
                // advance the state machine until this lane holds a ray to trace (or runs dry)
                for (;;) {
                    if (array_index < 0) {
                        unsigned int idx = atomicAdd(p.work_counter, 1u);
                        if (idx >= total) { mode = IDLE; break; }
                        unsigned int tile = idx >> 5, within = idx & 31u;
                        int x = (int)(tile % tiles_x) * 8 + (int)(within & 7u);
                        int ly = (int)(tile / tiles_x) * 4 + (int)(within >> 3);
                        if (x >= p.width || ly >= p.local_rows) continue;
                        int y = shard_row(p.shard_rank, p.shard_world, p.band_rows, ly);
                        if (y >= p.height) continue;
                        array_index = (y * p.width + x) * 3;  // raytracer.cu:123
                        RTB_ASSERT(array_index >= 0 && array_index + 2 < p.width * p.height * 3);
                        path_begin_pixel(P, array_index, p.time_ms);
                        pdir = primary_direction(p.cam, x, y);
                    }
                    if (P.bounce == 0) {
                        if (P.sample >= f.spp) {
                            path_write_pixel(P, f.spp, p.frame_num, p.prev, p.out, p.peers, array_index);
                            array_index = -1;
                            continue;
                        }
                        path_begin_sample(f, pdir, ray, P);
                        if (COUNT) n_samples++;
                        if (f.limit <= 0) { P.sample++; continue; }
                    }
                    break;  // a ray is ready
                }
                if (mode != IDLE) {
                    ray_finalise(ray, f.aa, P.rng);
                    fresh = true;
                    mode = TRAVERSING;
                }
            }
            // new rays start together: large primitives, then the cursor goes to the root
            unsigned starting = __ballot_sync(FULL, fresh);
            if (fresh) {
                fresh = false;
                if (BRUTE) {
                    T.best = trace_brute(s, ray, cnt);
                    T.cur = RTB_DONE; T.leaf = 0;
                } else {
                    trav_begin(s, ray, T, starting, cnt);
                }
            }
        }
        // --------------------------------- TRAVERSE phase --------------------------------------
        unsigned tmask = __ballot_sync(FULL, mode == TRAVERSING);
        // leave the tree once `refill` lanes of the warp wait for the SHADE phase
        int quota = refill - __popc(__ballot_sync(FULL, mode == WAITING));
#ifdef RTB_DYNAMIC_MASK
        if (mode == TRAVERSING) {
            if (!BRUTE) trav_run<RTB_MEGA_NODE_STEPS, EARLY_POP>(s, ray, T, tmask, quota < 1 ? 1 : quota, walk_min, cnt);
            if (trav_done(T)) {
                mode = WAITING;
                pending = true;
            }
        }
#else
        // all 32 lanes enter together: WAITING and IDLE lanes hold a finished traversal state
        // quota over the whole warp: the lanes that are not traversing now, plus max(1, refill - waiting) more
        if (!BRUTE) trav_run<RTB_MEGA_NODE_STEPS, EARLY_POP>(s, ray, T, FULL, (quota < 1 ? 1 : quota) + 32 - __popc(tmask), walk_min, cnt);
        if (mode == TRAVERSING && trav_done(T)) {
            mode = WAITING;
            pending = true;
        }
#endif
    }
    if (COUNT && p.counters) {
        atomicAdd(p.counters + 0, n_seg);
        atomicAdd(p.counters + 1, n_hit);
        atomicAdd(p.counters + 2, cnt.box);
        atomicAdd(p.counters + 3, cnt.tri);
        atomicAdd(p.counters + 4, cnt.sph);
        atomicAdd(p.counters + 5, n_samples);
    }
}

// primary-hit ids, antialias off; one thread per pixel
template <bool BRUTE> __global__ void primary_ids_kernel(const IdsParams p) {
    int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y * blockDim.y + threadIdx.y;
    // out-of-frame lanes trace too (the traversal is warp-synchronous) and write nothing
    const bool valid = x < p.width && y < p.height;
    const unsigned mask = 0xffffffffu;
    Ray ray;
    ray.o = v3(p.cam.cam_pos[0], p.cam.cam_pos[1], p.cam.cam_pos[2]);
    change_direction(ray, primary_direction(p.cam, x, y));
    Counters cnt;
#ifdef RTB_COUNTERS
    cnt = Counters{};
#endif
    Best best = trace<BRUTE>(p.scene, ray, mask, cnt);
    size_t i = (size_t)y * p.width + x;
    if (!valid) return;
    if (best.code < 0) {
        p.obj_id[i] = -1; p.sub_id[i] = -1; p.t[i] = RTB_INF_F;
    } else {
        int2 ids = code_ids(p.scene, best.code);
        p.obj_id[i] = ids.x; p.sub_id[i] = ids.y; p.t[i] = best.t;
    }
}

// early_pop: the variant of the traversal loop that fetches the stack entries with the node (trav_run, EARLY_POP)
cudaError_t launch_render(const RenderParams &p, bool brute, bool count, bool early_pop, LaunchConfig cfg, cudaStream_t st) {
    if (brute) {
        if (count) render_megakernel<true, true, false><<<cfg.blocks, cfg.threads, 0, st>>>(p);
        else render_megakernel<true, false, false><<<cfg.blocks, cfg.threads, 0, st>>>(p);
    } else if (early_pop) {
        if (count) render_megakernel<false, true, true><<<cfg.blocks, cfg.threads, 0, st>>>(p);
        else render_megakernel<false, false, true><<<cfg.blocks, cfg.threads, 0, st>>>(p);
    } else {
        if (count) render_megakernel<false, true, false><<<cfg.blocks, cfg.threads, 0, st>>>(p);
        else render_megakernel<false, false, false><<<cfg.blocks, cfg.threads, 0, st>>>(p);
    }
    return cudaGetLastError();
}

cudaError_t launch_primary_ids(const IdsParams &p, bool brute, cudaStream_t st) {
    dim3 block(8, 8), grid((p.width + 7) / 8, (p.height + 7) / 8);
    if (brute) primary_ids_kernel<true><<<grid, block, 0, st>>>(p);
    else primary_ids_kernel<false><<<grid, block, 0, st>>>(p);
    return cudaGetLastError();
}

cudaError_t render_occupancy(int threads, bool early_pop, int *blocks_per_sm) {
    if (early_pop) return cudaOccupancyMaxActiveBlocksPerMultiprocessor(blocks_per_sm, render_megakernel<false, false, true>, threads, 0);
    return cudaOccupancyMaxActiveBlocksPerMultiprocessor(blocks_per_sm, render_megakernel<false, false, false>, threads, 0);
}

#include "rtb_wavefront.cuh"
#include "rtb_pool.cuh"
#include "rtb_stream.cuh"

}  // namespace RTB_NS
}  // namespace rtb
