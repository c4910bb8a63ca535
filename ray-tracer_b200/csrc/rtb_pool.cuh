// rtb_pool.cuh -- the pooled persistent kernel (option "kernel" = 3): ONE launch per frame.
// Included by rtb_render.cuh inside namespace rtb::RTB_NS, after the shared path helpers.
//
// The wavefront kernels (rtb_wavefront.cuh) keep shading and traversal converged, but they pay for
// it with a frame-wide barrier per path segment: two launches per segment (8 956 per 256-spp frame),
// a 96-byte record per pixel that does not fit the L2 (1.1 TB of DRAM traffic per C3 step), a queue
// rebuild, and the tail of the longest ray of every iteration.  The megakernel has none of that, but
// a lane owns a pixel, so its traversal loop runs with the lanes that are not waiting to be shaded.
//
// Here every WARP owns a pool of 64 pixel slots whose records live in global memory, 6 KB per
// warp: 29 MB for the whole grid, resident in the 126 MB L2, never evicted by another warp's
// traffic (accessed with .cg, so they stay out of L1 too, which the tree keeps).  Lanes are decoupled
// from pixels:
//   SHADE     32 slots whose ray has finished are shaded together (the same straight-line code as
//             wf_shade: consume the hit, scatter, advance the sample / bounce / pixel state machine,
//             new ray, large primitives, root step); a finished pixel's slot takes the next pixel of
//             the frame from one global counter;
//   REFILL    lanes without a ray take the next rays of the warp's own ready list;
//   TRAVERSE  the warp-synchronous walk of rtb_trace.cuh, left as soon as `refill` lanes are free.
// With 64 slots and at most 32 rays in flight, at least 32 slots are always ready for one of the two
// phases, so both run with full warps until the frame runs out of pixels.  There is no barrier
// wider than a warp, no queue shared between warps, no DRAM traffic but the pixel's final write.
// The samples of a pixel stay strictly serial (one live path per slot), so the image is bit-identical
// to the megakernel's and the wavefront kernels' (SURVEY.md F10 / H1).
//
// Per-slot record, 6 float4 planes of 64 entries per warp (same fields as the wavefront record):
//   A {thr.xyz, cur_n}   B {col.xyz, rng bits}   C {acc.xyz, sample << 8 | bounce}
//   O {origin.xyz, first node below the root}   D {direction.xyz, second node}   H {t, u, v, code}
// plus the slot's pixel (x | y << 16) in shared memory.

#ifndef RTB_POOL_THREADS
#define RTB_POOL_THREADS 128
#endif
#ifndef RTB_POOL_MIN_BLOCKS
#define RTB_POOL_MIN_BLOCKS 6
#endif
#define RTB_POOL_SLOTS 64
#define RTB_POOL_WARPS (RTB_POOL_THREADS / 32)

__device__ __forceinline__ float4 pool_ld(const float4 *p) { return __ldcg(p); }
__device__ __forceinline__ void pool_st(float4 *p, float4 v) { __stcg(p, v); }

__device__ __forceinline__ unsigned long long pool_or64(unsigned long long v) {
    unsigned int lo = __reduce_or_sync(0xffffffffu, (unsigned int)v);
    unsigned int hi = __reduce_or_sync(0xffffffffu, (unsigned int)(v >> 32));
    return ((unsigned long long)hi << 32) | lo;
}

// The lowest n set bits of `m` (warp-uniform), and take[r] = position of the r-th of them for r < min(n, 32).
// Lane l looks at bits l and l + 32.
__device__ __forceinline__ unsigned long long pool_select(unsigned long long m, int n, unsigned char *take, unsigned int lane) {
    unsigned int lo = (unsigned int)m, hi = (unsigned int)(m >> 32);
    unsigned int below = (1u << lane) - 1u;
    int r0 = __popc(lo & below), r1 = __popc(lo) + __popc(hi & below);
    bool t0 = ((lo >> lane) & 1u) && r0 < n, t1 = ((hi >> lane) & 1u) && r1 < n;
    __syncwarp();
    if (t0 && r0 < 32) take[r0] = (unsigned char)lane;
    if (t1 && r1 < 32) take[r1] = (unsigned char)(lane + 32u);
    __syncwarp();
    return ((unsigned long long)__ballot_sync(0xffffffffu, t1) << 32) | __ballot_sync(0xffffffffu, t0);
}

template <bool BRUTE, bool COUNT>
__global__ void __launch_bounds__(RTB_POOL_THREADS, RTB_POOL_MIN_BLOCKS) pool_render_kernel(const PoolParams p) {
    const DScene &s = p.scene;
    const FrameConsts f = frame_consts(p.rd, p.cam);
    const unsigned int FULL = 0xffffffffu;
    const unsigned int lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const int tiles_x = (p.width + 7) >> 3, tiles_y = (p.local_rows + 3) >> 2;
    const unsigned int total = (unsigned int)tiles_x * tiles_y * 32u;
    const int refill = p.refill, walk_min = p.walk_min;

    __shared__ unsigned int s_pix[RTB_POOL_WARPS][RTB_POOL_SLOTS];
    __shared__ unsigned char s_take[RTB_POOL_WARPS][32];
    unsigned int *pix = s_pix[warp];
    unsigned char *take = s_take[warp];

    const size_t warp_gid = (size_t)blockIdx.x * RTB_POOL_WARPS + warp;
    float4 *st = p.pool + warp_gid * (6 * RTB_POOL_SLOTS);
    float4 *stA = st, *stB = st + RTB_POOL_SLOTS, *stC = st + 2 * RTB_POOL_SLOTS, *stO = st + 3 * RTB_POOL_SLOTS,
           *stD = st + 4 * RTB_POOL_SLOTS, *stH = st + 5 * RTB_POOL_SLOTS;

    Counters cnt;
    cnt = Counters{};
    unsigned long long n_seg = 0, n_hit = 0, n_samples = 0;

    // warp-uniform slot sets
    unsigned long long shade_ready = 0, trav_ready = 0, empty = ~0ull;
    bool exhausted = false;  // the frame has no pixels left

    // the ray this lane walks through the tree
    int tslot = -1;
    Ray ray;
    ray.o = f.cam_o; ray.d = v3(0, 0, 1); ray.inv = ray.d;
    Trav T;
    T.cur = RTB_DONE; T.leaf = 0; T.sp = 0; T.top = RTB_DONE;
    T.best.t = RTB_INF_F; T.best.u = 0; T.best.v = 0; T.best.code = -1;

    for (;;) {
        bool busy = tslot >= 0 && !trav_done(T);
        unsigned int busy_mask = __ballot_sync(FULL, busy);
        int n_free = 32 - __popc(busy_mask);
        if (n_free >= refill || !busy_mask) {
            // ---- retire: finished lanes store their hit; the slot waits for the SHADE phase ----
            bool retire = tslot >= 0 && !busy;
            if (retire) pool_st(stH + tslot, make_float4(T.best.t, T.best.u, T.best.v, __int_as_float(T.best.code)));
            shade_ready |= pool_or64(retire ? 1ull << tslot : 0ull);
            if (retire) tslot = -1;
            // ---- SHADE while the ready list cannot fill the free lanes --------------------------
            while (__popcll(trav_ready) < n_free && (shade_ready || (empty && !exhausted))) {
                // up to 32 slots: finished rays first, then (while the frame has pixels) empty slots
                int n_ready = min(32, __popcll(shade_ready));
                int n_batch = exhausted ? n_ready : min(32, n_ready + __popcll(empty));
                unsigned long long from_ready = pool_select(shade_ready, n_ready, take, lane);
                bool live = (int)lane < n_batch;
                int slot = (int)lane < n_ready ? take[lane] : 0;
                unsigned long long from_empty = pool_select(empty, n_batch - n_ready, take, lane);
                bool fresh = live && (int)lane >= n_ready;
                if (fresh) slot = take[lane - n_ready];
                unsigned long long batch = from_ready | from_empty;
                RTB_ASSERT(!live || (slot >= 0 && slot < RTB_POOL_SLOTS && ((batch >> slot) & 1ull)));
                RTB_ASSERT(__popcll(batch) == n_batch && !(from_ready & from_empty) && !(batch & trav_ready));
                int array_index = -1;
                V3 pdir = v3(0, 0, 1);
                Path P;
                P.rng = 0; P.sample = 0; P.bounce = 0; P.cur_n = 1;
                P.thr = v3(1, 1, 1); P.col = v3(0, 0, 0); P.acc = v3(0, 0, 0);
                Ray sr;
                sr.o = f.cam_o; sr.d = pdir; sr.inv = pdir;
                if (live && !fresh) {
                    float4 a = pool_ld(stA + slot), b = pool_ld(stB + slot), c = pool_ld(stC + slot), o = pool_ld(stO + slot),
                           d = pool_ld(stD + slot), h = pool_ld(stH + slot);
                    unsigned int xy = pix[slot];
                    int x = (int)(xy & 0xffffu), y = (int)(xy >> 16);
                    array_index = (y * p.width + x) * 3;
                    pdir = primary_direction(p.cam, x, y);
                    P.thr = v3(a.x, a.y, a.z); P.cur_n = a.w;
                    P.col = v3(b.x, b.y, b.z); P.rng = __float_as_uint(b.w);
                    P.acc = v3(c.x, c.y, c.z);
                    unsigned int sb = __float_as_uint(c.w);
                    P.sample = (int)(sb >> 8); P.bounce = (int)(sb & 255u);
                    sr.o = v3(o.x, o.y, o.z);
                    sr.d = v3(d.x, d.y, d.z);
                    sr.inv = sr.d;  // never read before the next change of direction
                    Best best;
                    best.t = h.x; best.u = h.y; best.v = h.z; best.code = __float_as_int(h.w);
                    bool hit = path_shade(s, f, sr, best, P);
                    if (COUNT) { n_seg++; n_hit += hit; }
                }
                // advance until the slot holds a ray to trace, or the frame has no pixel for it
                bool ran_dry = false;
                if (live) {
                    for (;;) {
                        if (array_index < 0) {
                            unsigned int idx = atomicAdd(p.work_counter, 1u);
                            if (idx >= total) { ran_dry = true; live = false; break; }
                            unsigned int tile = idx >> 5, within = idx & 31u;
                            int x = (int)(tile % tiles_x) * 8 + (int)(within & 7u);
                            int ly = (int)(tile / tiles_x) * 4 + (int)(within >> 3);
                            if (x >= p.width || ly >= p.local_rows) continue;
                            int y = shard_row(p.shard_rank, p.shard_world, p.band_rows, ly);
                            if (y >= p.height) continue;
                            array_index = (y * p.width + x) * 3;  // raytracer.cu:123
                            RTB_ASSERT(array_index >= 0 && array_index + 2 < p.width * p.height * 3);
                            pix[slot] = (unsigned int)x | ((unsigned int)y << 16);
                            path_begin_pixel(P, array_index, p.time_ms);
                            pdir = primary_direction(p.cam, x, y);
                        }
                        if (P.bounce == 0) {
                            if (P.sample >= f.spp) {
                                path_write_pixel(P, f.spp, p.frame_num, p.prev, p.out, p.peers, array_index);
                                array_index = -1;
                                continue;
                            }
                            path_begin_sample(f, pdir, sr, P);
                            if (COUNT) n_samples++;
                            if (f.limit <= 0) { P.sample++; continue; }
                        }
                        break;
                    }
                }
                if (__any_sync(FULL, ran_dry)) exhausted = true;
                if (live) {
                    ray_finalise(sr, f.aa, P.rng);  // antialias jitter, direction_inv
                    pool_st(stA + slot, make_float4(P.thr.x, P.thr.y, P.thr.z, P.cur_n));
                    pool_st(stB + slot, make_float4(P.col.x, P.col.y, P.col.z, __uint_as_float(P.rng)));
                    pool_st(stC + slot, make_float4(P.acc.x, P.acc.y, P.acc.z, __uint_as_float(((unsigned int)P.sample << 8) | (unsigned int)P.bounce)));
                }
                // the new rays meet the large primitives and the root with the whole warp converged
                unsigned int alive = __ballot_sync(FULL, live);
                bool enter = false;
                if (live) {
                    Best first;
                    int start = RTB_DONE, then = 0;
                    if (BRUTE) {
                        first = trace_brute(s, sr, cnt);
                    } else {
                        large_prims_best(s, sr, first, alive, cnt);
                        root_step(s, sr, first, start, then, cnt);
                        enter = start != RTB_DONE;
                    }
                    pool_st(stO + slot, make_float4(sr.o.x, sr.o.y, sr.o.z, __int_as_float(start)));
                    pool_st(stD + slot, make_float4(sr.d.x, sr.d.y, sr.d.z, __int_as_float(then)));
                    pool_st(stH + slot, make_float4(first.t, first.u, first.v, __int_as_float(first.code)));
                }
                unsigned long long bit = 1ull << slot;
                unsigned long long entering = pool_or64(enter ? bit : 0ull);
                unsigned long long staying = pool_or64(live && !enter ? bit : 0ull);  // the hit is final already: shade again
                trav_ready |= entering;
                shade_ready = (shade_ready & ~batch) | staying;
                empty = (empty & ~batch) | (batch & ~(entering | staying));
            }
            // ---- REFILL: free lanes take the next rays of the ready list -------------------------
            if (trav_ready) {
                unsigned int want = __ballot_sync(FULL, tslot < 0);
                int n_take = min(__popc(want), __popcll(trav_ready));
                unsigned long long taken = pool_select(trav_ready, n_take, take, lane);
                int rank = __popc(want & ((1u << lane) - 1u));
                if (tslot < 0 && rank < n_take) {
                    tslot = take[rank];
                    RTB_ASSERT(tslot >= 0 && tslot < RTB_POOL_SLOTS && ((taken >> tslot) & 1ull));
                    float4 o = pool_ld(stO + tslot), d = pool_ld(stD + tslot), h = pool_ld(stH + tslot);
                    ray.o = v3(o.x, o.y, o.z);
                    ray.d = v3(d.x, d.y, d.z);
                    ray.inv = v3(1 / d.x, 1 / d.y, 1 / d.z);  // ray.cu:198-202
                    T.best.t = h.x; T.best.u = h.y; T.best.v = h.z; T.best.code = __float_as_int(h.w);
                    trav_start_below_root(T, __float_as_int(o.w), __float_as_int(d.w));
                }
                trav_ready &= ~taken;
            }
            busy = tslot >= 0 && !trav_done(T);
            busy_mask = __ballot_sync(FULL, busy);
            if (!busy_mask) {
                // nothing in flight: either rays that ended at once (retired next round) or the end of the frame
                if (__any_sync(FULL, tslot >= 0)) continue;
                if (!trav_ready && !shade_ready && (exhausted || !empty)) break;
                continue;
            }
        }
        // ---- TRAVERSE: all 32 lanes enter together (lanes without a ray hold a finished state) ------
        // leave when `refill` lanes are free, as long as something can be done with free lanes
        bool more = trav_ready || shade_ready || (empty && !exhausted);
        if (!BRUTE) trav_run<2>(s, ray, T, FULL, more ? max(refill, 33 - __popc(busy_mask)) : 0, walk_min, cnt);
    }
    if (COUNT && p.counters) {
        atomicAdd(p.counters + 0, n_seg);
        atomicAdd(p.counters + 1, n_hit);
        atomicAdd(p.counters + 2, cnt.box);
        atomicAdd(p.counters + 3, cnt.tri);
        atomicAdd(p.counters + 4, cnt.sph);
        atomicAdd(p.counters + 5, n_samples);
#ifdef RTB_COUNTERS
        if (lane == 0) {  // the lane statistics are warp-uniform
            atomicAdd(p.counters + 8, cnt.node_iters); atomicAdd(p.counters + 9, cnt.node_lanes);
            atomicAdd(p.counters + 10, cnt.node_blocked); atomicAdd(p.counters + 11, cnt.node_done);
            atomicAdd(p.counters + 12, cnt.leaf_iters); atomicAdd(p.counters + 13, cnt.leaf_lanes);
            atomicAdd(p.counters + 14, cnt.mask_lanes);
        }
#endif
    }
}

cudaError_t launch_pool(const PoolParams &p, bool brute, bool count, int blocks, cudaStream_t st) {
    if (brute) {
        if (count) pool_render_kernel<true, true><<<blocks, RTB_POOL_THREADS, 0, st>>>(p);
        else pool_render_kernel<true, false><<<blocks, RTB_POOL_THREADS, 0, st>>>(p);
    } else {
        if (count) pool_render_kernel<false, true><<<blocks, RTB_POOL_THREADS, 0, st>>>(p);
        else pool_render_kernel<false, false><<<blocks, RTB_POOL_THREADS, 0, st>>>(p);
    }
    return cudaGetLastError();
}

cudaError_t pool_occupancy(int *blocks_per_sm) {
    return cudaOccupancyMaxActiveBlocksPerMultiprocessor(blocks_per_sm, pool_render_kernel<false, false>, RTB_POOL_THREADS, 0);
}

int pool_block_threads() { return RTB_POOL_THREADS; }
