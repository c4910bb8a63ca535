// rtb_wavefront.cuh -- the wavefront variant of the render path (option "kernel" = 1).
// Included by rtb_render.cuh inside namespace rtb::RTB_NS, after the shared path helpers.
//
// The megakernel keeps a pixel's whole state in one lane and has to run its two very different
// halves -- shading (long, straight-line, RNG / libm heavy) and tree traversal (short, data
// dependent loops) -- with whichever lanes of the warp happen to need them.  Here the halves are
// separate kernels over a queue of live pixels:
//
//   wf_shade     one thread per live pixel: consume the closest hit of the pixel's current ray
//                (raytracer.cu:74-91), advance the sample / bounce state machine, emit the next
//                ray and test it against the large primitives (walls, floors: a converged loop),
//                or finish the pixel (raytracer.cu:107-112) and drop it from the queue
//                (warp-ballot compaction, order preserving inside a warp);
//   wf_traverse  persistent warps with per-lane dynamic fetch: a lane whose ray is finished
//                stores the hit and pulls the next ray from the queue (cheap: two float4 loads),
//                so the traversal loop runs with full warps.
//
// One iteration = one path segment of every live pixel.  The samples of a pixel stay strictly
// serial (one live path per pixel), because the reference's RNG state is per pixel and carried
// from sample to sample: the images are bit-identical to the megakernel's.
//
// Per-pixel record (6 float4 = 96 B, SoA by field, indexed by the pixel's local slot):
//   A {thr.xyz, cur_n}   B {col.xyz, rng bits}   C {acc.xyz, sample << 8 | bounce}
//   O {origin.xyz, first node below the root}   D {direction.xyz, second node}   H {t, u, v, code}

#ifndef RTB_WF_NODE_STEPS
#define RTB_WF_NODE_STEPS 2   // node steps between two ballots in wf_traverse (measured: 1 is 1-2 % slower, 3 see DESIGN.md)
#endif
#ifndef RTB_SHADE_MIN_BLOCKS
#define RTB_SHADE_MIN_BLOCKS 4
#endif
#ifndef RTB_TRAV_MIN_BLOCKS
#define RTB_TRAV_MIN_BLOCKS 8
#endif

__device__ __forceinline__ void wf_slot_to_pixel(const WfParams &p, int slot, int &x, int &y) {
    int ly = slot / p.width;
    x = slot - ly * p.width;
    y = shard_row(p.shard_rank, p.shard_world, p.band_rows, ly);
}

template <bool BRUTE, bool COUNT>
__global__ void __launch_bounds__(RTB_SHADE_THREADS, RTB_SHADE_MIN_BLOCKS) wf_shade_kernel(const WfParams p) {
    const DScene &s = p.scene;
    const FrameConsts f = frame_consts(p.rd, p.cam);
    const unsigned int n = p.first ? (unsigned int)p.slot_count : *p.n_in;
    const unsigned int lane = threadIdx.x & 31u;
    unsigned long long n_seg = 0, n_hit = 0, n_samples = 0;
    Counters cnt;
    cnt = Counters{};
    if (blockIdx.x == 0 && threadIdx.x == 0) *p.fetch = 0;  // for the traversal launch that follows
    // whole warps walk the queue so that the ballot compaction below is converged
    for (unsigned int base = (blockIdx.x * blockDim.x + threadIdx.x) & ~31u; base < n; base += gridDim.x * blockDim.x) {
        unsigned int i = base + lane;
        bool live = i < n;
        int slot = 0;
        Ray ray;
        ray.o = f.cam_o; ray.d = v3(0, 0, 1); ray.inv = ray.d;
        if (live) {
            slot = p.first ? p.slot_base + (int)i : p.queue_in[i];
            RTB_ASSERT(slot >= p.slot_base && slot < p.slot_base + p.slot_count);
            int x, y;
            wf_slot_to_pixel(p, slot, x, y);
            int array_index = (y * p.width + x) * 3;  // raytracer.cu:123
            V3 pdir = primary_direction(p.cam, x, y);
            Path P;
            if (p.first) {
                p.alive[slot] = 1;
                path_begin_pixel(P, array_index, p.time_ms);
                P.cur_n = 1; P.thr = v3(1, 1, 1); P.col = v3(0, 0, 0);
                ray.o = f.cam_o; ray.d = pdir; ray.inv = pdir;
            } else {
                float4 a = p.st.A[slot], b = p.st.B[slot], c = p.st.C[slot], o = p.st.O[slot], d = p.st.D[slot], h = p.st.H[slot];
                P.thr = v3(a.x, a.y, a.z); P.cur_n = a.w;
                P.col = v3(b.x, b.y, b.z); P.rng = __float_as_uint(b.w);
                P.acc = v3(c.x, c.y, c.z);
                unsigned int sb = __float_as_uint(c.w);
                P.sample = (int)(sb >> 8); P.bounce = (int)(sb & 255u);
                ray.o = v3(o.x, o.y, o.z);
                ray.d = v3(d.x, d.y, d.z);
                ray.inv = ray.d;  // never read here: every direction change recomputes it (ray.cu:198-202)
                Best best;
                best.t = h.x; best.u = h.y; best.v = h.z; best.code = __float_as_int(h.w);
                bool hit = path_shade(s, f, ray, best, P);
                if (COUNT) { n_seg++; n_hit += hit; }
            }
            // advance until the pixel holds a ray to trace, or is finished
            for (;;) {
                if (P.bounce == 0) {
                    if (P.sample >= f.spp) {
                        path_write_pixel(P, f.spp, p.frame_num, p.prev, p.out, p.peers, array_index);
                        live = false;
                        p.alive[slot] = 0;
                        break;
                    }
                    path_begin_sample(f, pdir, ray, P);
                    if (COUNT) n_samples++;
                    if (f.limit <= 0) { P.sample++; continue; }
                }
                break;
            }
            if (live) {
                ray_finalise(ray, f.aa, P.rng);  // antialias jitter, direction_inv
                p.st.A[slot] = make_float4(P.thr.x, P.thr.y, P.thr.z, P.cur_n);
                p.st.B[slot] = make_float4(P.col.x, P.col.y, P.col.z, __uint_as_float(P.rng));
                p.st.C[slot] = make_float4(P.acc.x, P.acc.y, P.acc.z, __uint_as_float(((unsigned int)P.sample << 8) | (unsigned int)P.bounce));
            }
        }
        // the new rays meet the large primitives here, with the whole warp converged, instead of
        // at the start of the traversal, where only the few freshly fetched lanes would run it
        // ... and take the root's node step (the root node sits in the kernel parameters): a ray
        // that cannot touch the tree closer than its large-primitive hit never enters wf_traverse
        unsigned int alive = __ballot_sync(0xffffffffu, live);
        bool enter = false;
        if (live) {
            Best first;
            int start = RTB_DONE, then = 0;
            if (BRUTE) {
                first.t = RTB_INF_F; first.u = 0; first.v = 0; first.code = -1;
                enter = true;
            } else {
                large_prims_best(s, ray, first, alive, cnt);
                root_step(s, ray, first, start, then, cnt);
                enter = start != RTB_DONE;
            }
            p.st.O[slot] = make_float4(ray.o.x, ray.o.y, ray.o.z, __int_as_float(start));
            p.st.D[slot] = make_float4(ray.d.x, ray.d.y, ray.d.z, __int_as_float(then));
            p.st.H[slot] = make_float4(first.t, first.u, first.v, __int_as_float(first.code));
        }
        unsigned int entering = __ballot_sync(0xffffffffu, enter);
        unsigned int tpos = 0;
        if (lane == 0 && entering) tpos = atomicAdd(p.tn, (unsigned int)__popc(entering));
        tpos = __shfl_sync(0xffffffffu, tpos, 0);
        RTB_ASSERT(!enter || tpos + __popc(entering) <= (unsigned int)p.slot_count);
        if (enter) p.tqueue[tpos + __popc(entering & ((1u << lane) - 1u))] = slot;
        // compaction: the survivors of this warp go to consecutive queue positions, in order
        unsigned int pos = 0;
        if (lane == 0 && alive) pos = atomicAdd(p.n_out, (unsigned int)__popc(alive));
        pos = __shfl_sync(0xffffffffu, pos, 0);
        RTB_ASSERT(!live || pos + __popc(alive) <= (unsigned int)p.slot_count);
        if (live) p.queue_out[pos + __popc(alive & ((1u << lane) - 1u))] = slot;
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

// Persistent traversal: every lane owns one ray at a time; finished lanes store their hit and
// fetch the next ray as soon as `refill` of them are waiting (or nobody traverses any more).
template <bool BRUTE, bool COUNT>
__global__ void __launch_bounds__(RTB_BLOCK_THREADS, RTB_TRAV_MIN_BLOCKS) wf_traverse_kernel(const WfParams p) {
    const DScene &s = p.scene;
    const unsigned int FULL = 0xffffffffu;
    const unsigned int n = *p.tn;
    const unsigned int lane = threadIdx.x & 31u;
    if (blockIdx.x == 0 && threadIdx.x == 0) {  // the counters the next wf_shade appends to
        *p.n_out = 0;
        *p.tn_next = 0;
    }
    Counters cnt;
    cnt = Counters{};
    Ray ray;
    ray.o = v3(0, 0, 0); ray.d = v3(0, 0, 1); ray.inv = ray.d;
    Trav T;
    T.cur = RTB_DONE; T.leaf = 0; T.sp = 0; T.top = RTB_DONE;
    T.best.t = RTB_INF_F; T.best.u = 0; T.best.v = 0; T.best.code = -1;
    int slot = -1;           // the ray this lane owns, -1 = none
    bool exhausted = false;  // the queue has nothing left (warp-uniform)
    for (;;) {
        bool busy = slot >= 0 && !trav_done(T);
        unsigned int busy_mask = __ballot_sync(FULL, busy);
        int free_lanes = 32 - __popc(busy_mask);
        if (!exhausted && (free_lanes >= p.refill || !busy_mask)) {
            // finished lanes store their result ...
            if (slot >= 0 && !busy) {
                p.st.H[slot] = make_float4(T.best.t, T.best.u, T.best.v, __int_as_float(T.best.code));
                slot = -1;
            }
            // ... and take the next rays of the queue
            unsigned int want = __ballot_sync(FULL, slot < 0);
            unsigned int pos = 0;
            if (lane == 0) pos = atomicAdd(p.fetch, (unsigned int)__popc(want));
            pos = __shfl_sync(FULL, pos, 0);
            unsigned int mine = pos + __popc(want & ((1u << lane) - 1u));
            bool fresh = slot < 0 && mine < n;
            if (pos + __popc(want) >= n) exhausted = true;
            if (fresh) {
                RTB_ASSERT(mine < (unsigned int)p.slot_count);
                slot = p.tqueue[mine];
                RTB_ASSERT(slot >= p.slot_base && slot < p.slot_base + p.slot_count);
                float4 o = p.st.O[slot], d = p.st.D[slot], h = p.st.H[slot];
                ray.o = v3(o.x, o.y, o.z);
                ray.d = v3(d.x, d.y, d.z);
                ray.inv = v3(1 / d.x, 1 / d.y, 1 / d.z);  // ray.cu:198-202
                if (BRUTE) {
                    T.best = trace_brute(s, ray, cnt);
                    T.cur = RTB_DONE; T.leaf = 0;
                } else {
                    // wf_shade left the best large-primitive hit in H: only the tree remains
                    T.best.t = h.x; T.best.u = h.y; T.best.v = h.z; T.best.code = __float_as_int(h.w);
                    trav_start_below_root(T, __float_as_int(o.w), __float_as_int(d.w));
                }
            }
            busy = slot >= 0 && !trav_done(T);
            busy_mask = __ballot_sync(FULL, busy);
        }
        if (!busy_mask) {
            if (slot >= 0) {  // rays that ended inside trav_begin (or brute force)
                p.st.H[slot] = make_float4(T.best.t, T.best.u, T.best.v, __int_as_float(T.best.code));
                slot = -1;
            }
            if (exhausted) break;
            continue;
        }
#ifdef RTB_DYNAMIC_MASK
        if (busy) {
            // while the queue has rays, leave the tree when `refill` lanes are free; afterwards run to the end
            int quota = exhausted ? 0 : p.refill - (32 - __popc(busy_mask));
            if (!BRUTE) trav_run(s, ray, T, busy_mask, exhausted ? 0 : (quota < 1 ? 1 : quota), p.walk_min, cnt);
        }
#else
        // all 32 lanes enter together (lanes without a ray hold a finished state): while the queue has
        // rays, leave the tree when `refill` lanes are free; afterwards run to the end
        // (the quota counts finished lanes of the whole warp: the lanes that are free now, plus at least one)
        if (!BRUTE) trav_run<RTB_WF_NODE_STEPS>(s, ray, T, FULL, exhausted ? 0 : max(p.refill, 33 - __popc(busy_mask)), p.walk_min, cnt);
#endif
    }
    if (COUNT && p.counters) {
        atomicAdd(p.counters + 2, cnt.box);
        atomicAdd(p.counters + 3, cnt.tri);
        atomicAdd(p.counters + 4, cnt.sph);
        if (lane == 0) {  // the lane statistics are warp-uniform
            atomicAdd(p.counters + 8, cnt.node_iters); atomicAdd(p.counters + 9, cnt.node_lanes);
            atomicAdd(p.counters + 10, cnt.node_blocked); atomicAdd(p.counters + 11, cnt.node_done);
            atomicAdd(p.counters + 12, cnt.leaf_iters); atomicAdd(p.counters + 13, cnt.leaf_lanes);
            atomicAdd(p.counters + 14, cnt.mask_lanes);
        }
    }
}

// ---- queue rebuild -----------------------------------------------------------------------------
// The ballot compaction in wf_shade keeps the order inside a warp but appends the warps in arrival
// order, so while pixels keep finishing the queue slowly turns into short runs of neighbouring
// slots in random order, and the per-slot records are no longer read in whole sectors.  Every few
// iterations the queue is rebuilt in slot order from the per-slot alive flags: block counts, a
// one-block scan, an order-preserving scatter.
constexpr int WF_QT = 256;

__global__ void wf_queue_count_kernel(int base, int n, const unsigned char *alive, int *block_count) {
    int i = blockIdx.x * WF_QT + threadIdx.x;
    int c = __syncthreads_count(i < n && alive[base + i]);
    if (threadIdx.x == 0) block_count[blockIdx.x] = c;
}

__global__ void wf_queue_scan_kernel(int num_blocks, int *block_count, unsigned int *total) {
    __shared__ int part[1024];
    __shared__ int carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (int base = 0; base < num_blocks; base += 1024) {
        int i = base + threadIdx.x;
        int v = i < num_blocks ? block_count[i] : 0;
        part[threadIdx.x] = v;
        __syncthreads();
        for (int off = 1; off < 1024; off <<= 1) {
            int t = threadIdx.x >= off ? part[threadIdx.x - off] : 0;
            __syncthreads();
            part[threadIdx.x] += t;
            __syncthreads();
        }
        if (i < num_blocks) block_count[i] = carry + part[threadIdx.x] - v;  // exclusive
        __syncthreads();
        if (threadIdx.x == 1023) carry += part[1023];
        __syncthreads();
    }
    if (threadIdx.x == 0) *total = (unsigned int)carry;
}

__global__ void wf_queue_scatter_kernel(int base, int n, const unsigned char *alive, const int *block_offset, int *queue) {
    __shared__ int warp_sum[WF_QT / 32];
    int i = blockIdx.x * WF_QT + threadIdx.x;
    bool k = i < n && alive[base + i];
    unsigned int b = __ballot_sync(0xffffffffu, k);
    int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) warp_sum[warp] = __popc(b);
    __syncthreads();
    int off = block_offset[blockIdx.x];
    for (int w = 0; w < warp; w++) off += warp_sum[w];
    off += __popc(b & ((1u << lane) - 1u));
    if (k) queue[off] = base + i;
}

cudaError_t launch_wf_rebuild(int slot_base, int n_slots, const unsigned char *alive, int *block_count, int *queue, unsigned int *total, cudaStream_t st) {
    int blocks = (n_slots + WF_QT - 1) / WF_QT;
    wf_queue_count_kernel<<<blocks, WF_QT, 0, st>>>(slot_base, n_slots, alive, block_count);
    wf_queue_scan_kernel<<<1, 1024, 0, st>>>(blocks, block_count, total);
    wf_queue_scatter_kernel<<<blocks, WF_QT, 0, st>>>(slot_base, n_slots, alive, block_count, queue);
    return cudaGetLastError();
}

cudaError_t launch_wf_shade(const WfParams &p, bool brute, bool count, int blocks, cudaStream_t st) {
    if (brute) {
        if (count) wf_shade_kernel<true, true><<<blocks, RTB_SHADE_THREADS, 0, st>>>(p);
        else wf_shade_kernel<true, false><<<blocks, RTB_SHADE_THREADS, 0, st>>>(p);
    } else {
        if (count) wf_shade_kernel<false, true><<<blocks, RTB_SHADE_THREADS, 0, st>>>(p);
        else wf_shade_kernel<false, false><<<blocks, RTB_SHADE_THREADS, 0, st>>>(p);
    }
    return cudaGetLastError();
}

cudaError_t launch_wf_traverse(const WfParams &p, bool brute, bool count, int blocks, cudaStream_t st) {
    if (brute) {
        if (count) wf_traverse_kernel<true, true><<<blocks, RTB_BLOCK_THREADS, 0, st>>>(p);
        else wf_traverse_kernel<true, false><<<blocks, RTB_BLOCK_THREADS, 0, st>>>(p);
    } else {
        if (count) wf_traverse_kernel<false, true><<<blocks, RTB_BLOCK_THREADS, 0, st>>>(p);
        else wf_traverse_kernel<false, false><<<blocks, RTB_BLOCK_THREADS, 0, st>>>(p);
    }
    return cudaGetLastError();
}

cudaError_t wf_occupancy(int *shade_blocks_per_sm, int *traverse_blocks_per_sm) {
    cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(shade_blocks_per_sm, wf_shade_kernel<false, false>, RTB_SHADE_THREADS, 0);
    if (e != cudaSuccess) return e;
    return cudaOccupancyMaxActiveBlocksPerMultiprocessor(traverse_blocks_per_sm, wf_traverse_kernel<false, false>, RTB_BLOCK_THREADS, 0);
}
