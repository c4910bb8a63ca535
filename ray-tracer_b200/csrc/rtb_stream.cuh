// rtb_stream.cuh -- the streaming kernel (option "kernel" = 4): ONE launch per frame, SMs specialised by role.
// Included by rtb_render.cuh inside namespace rtb::RTB_NS, after the shared path helpers.
//
// What the other variants pay for (measured):
//   wavefront   a frame-wide barrier per path segment: two launches per segment, a traversal launch that drains
//               with ever emptier warps (9.7 of 32 lanes idle on average), 96-byte records of ALL pixels streaming
//               through DRAM (1.1 TB per C3 step), a periodic queue rebuild;
//   pooled      shading and traversal interleaved by the warps of one SM: the two code paths together (40 KB
//               executed) overflow the SM's 32 KB instruction cache -- 6.5 stall cycles per issued instruction
//               waiting for instructions, and the kernel loses to the wavefront pair although its warps are fuller.
// Here every SM runs ONE of the two roles for the whole frame, so its instruction cache holds one code path:
//   SHADE SMs     take slots whose ray has finished from the hit queue, 32 at a time: consume the hit, scatter,
//                 advance the sample / bounce / pixel state machine, new ray, large primitives, root step (the
//                 code of wf_shade); rays that must enter the tree go to the ray queue, the others straight back
//                 to the hit queue; a finished pixel's slot takes the next pixel of the frame;
//   TRAVERSE SMs  persistent warps with per-lane dynamic fetch from the ray queue (the code of wf_traverse); a
//                 finished ray's slot goes to the hit queue.  The queues never run dry before the frame does, so
//                 the warps stay full: no per-segment drain.
// The queues are rings of slot indices in global memory, multi-producer / multi-consumer, sharded 8 ways to spread
// the atomics; a slot sits in at most one queue, so a ring as large as the slot count cannot overflow.  The per-slot
// records (the wavefront's six float4 planes) belong to a bounded number of slots in flight (a few hundred
// thousand: tens of MB), not to every pixel, so they stay in the 126 MB L2: no DRAM traffic but the pixel's final
// write.  One live path per slot keeps a pixel's samples serial: the image is bit-identical to the other kernels'.
// The SHADE : TRAVERSE split of the SMs is a launch parameter; the timed kernel choice tries a few.

#ifndef RTB_STREAM_THREADS
#define RTB_STREAM_THREADS 128
#endif
#ifndef RTB_STREAM_MIN_BLOCKS
#define RTB_STREAM_MIN_BLOCKS 8
#endif
// Sub-queues per queue, to spread the atomics on the head / tail words (a single FIFO per queue measured 1.7x
// slower than eight).  Every consumer warp belongs to ONE sub-queue for the whole frame and draws its tickets only
// there; producers rotate over the sub-queues.  (Consumers that rotate as well can all end up waiting on sub-queues
// that get no items while the items pile up where nobody holds a ticket: measured as a hang on small frames.)
// Within a sub-queue unsatisfied tickets ([tail, head)) and unclaimed items ([head, tail)) cannot both exist.
// (RTB_STREAM_SHARDS is defined in rtb_launch.h: the host lays the rings out.)
#define RTB_STREAM_MAX_SMID 1024

// control words (unsigned int each)
enum { SC_WORK = 0, SC_NEXT_SLOT = 1, SC_PIXELS_DONE = 2, SC_DONE = 3, SC_SM_COUNT = 4, SC_ERROR = 5, SC_WORDS = 8 };

// A warp that has waited this many idle polls (seconds) declares the frame stalled: every warp leaves, the host
// reports an error instead of a hung device.  Cannot happen by design; it is the seat belt of a kernel whose warps
// wait for each other.
#define RTB_STREAM_WATCHDOG (1 << 21)
__device__ __forceinline__ void stream_stalled(const StreamParams &p) {
    atomicExch(p.ctl + SC_ERROR, 1u);
    __threadfence();
    atomicExch(p.ctl + SC_DONE, 1u);
}
// an idle warp's watchdog: `spins` idle polls in a row; the count starts again whenever the hit queue has moved (a
// role can be idle for a whole long frame -- a scene that never enters the tree -- while the frame makes progress)
__device__ __forceinline__ bool stream_watchdog(const StreamParams &p, int &spins, unsigned int &seen) {
    if ((++spins & 1023) == 0) {
        unsigned int now = *(const volatile unsigned int *)p.hq.tail;
        if (now != seen) { seen = now; spins = 0; }
    }
    return spins > RTB_STREAM_WATCHDOG;
}

__device__ __forceinline__ unsigned int ld_vol(const unsigned int *p) { return *(const volatile unsigned int *)p; }
__device__ __forceinline__ int ld_vol(const int *p) { return *(const volatile int *)p; }
__device__ __forceinline__ void st_vol(unsigned int *p, unsigned int v) { *(volatile unsigned int *)p = v; }

// The queues.  A first version let consumers CAS the head up to the tail: with ~4 700 warps retrying on eight words
// the kernel spent 80 % of its stall samples inside that loop (ncu) and ran 30x slower than the wavefront pair.
// Now a consumer never retries: it draws TICKETS -- positions of the ring, possibly ahead of the tail -- with one
// atomicAdd and polls its own entries; the k-th item ever pushed goes to whoever holds ticket k.  A lane keeps a
// ticket until its item arrives (or the frame ends), so no item is ever left without a taker.
// Entry = 0 (empty) or (slot + 1) | lap << 31, lap = bit `shift` of the position: a ticket that runs a whole ring ahead
// of an item still waiting in its entry cannot take that item by mistake; a producer waits for its entry to be empty.
// sub-queue q of a StreamQueue: ring[q * cap .. (q + 1) * cap), counters head/tail[q * 32] (own 128-byte lines)
__device__ __forceinline__ unsigned int *sq_entry(const StreamQueue &Q, int q, unsigned int pos) {
    return (unsigned int *)Q.ring + (size_t)q * (Q.mask + 1u) + (pos & Q.mask);
}
__device__ __forceinline__ unsigned int sq_lap(const StreamQueue &Q, unsigned int pos) { return ((pos & (Q.mask + 1u)) != 0u) ? 0x80000000u : 0u; }

// Lanes with `has` append `slot` (warp-collective, all 32 lanes call).  The slot's record must have been written
// with pool_st (.cg): the fence below orders those stores before the ring entry that publishes the slot.
__device__ __forceinline__ void sq_push(const StreamQueue &Q, int q, bool has, int slot, unsigned int lane) {
    unsigned int m = __ballot_sync(0xffffffffu, has);
    if (!m) return;
    unsigned int base = 0;
    if (lane == 0) base = atomicAdd(Q.tail + 32 * q, (unsigned int)__popc(m));
    base = __shfl_sync(0xffffffffu, base, 0);
    if (has) {
        unsigned int pos = base + __popc(m & ((1u << lane) - 1u));
        unsigned int *e = sq_entry(Q, q, pos);
        // the item of the previous lap is still there: only when the queue holds a whole ring (it cannot: a slot sits in
        // at most one queue and the ring has room for all of them); bounded so that a logic error cannot hang the GPU
        for (int w = 0; ld_vol(e) != 0u && w < (1 << 24); w++) {}
        __threadfence();
        st_vol(e, (unsigned int)(slot + 1) | sq_lap(Q, pos));
    }
}

// Lanes with `need` draw a ticket of sub-queue q (warp-collective): tq / tpos name the entry the lane waits for.
__device__ __forceinline__ void sq_claim(const StreamQueue &Q, int q, bool need, int &tq, unsigned int &tpos, unsigned int lane) {
    unsigned int m = __ballot_sync(0xffffffffu, need);
    if (!m) return;
    unsigned int base = 0;
    if (lane == 0) base = atomicAdd(Q.head + 32 * q, (unsigned int)__popc(m));
    base = __shfl_sync(0xffffffffu, base, 0);
    if (need) { tq = q; tpos = base + __popc(m & ((1u << lane) - 1u)); }
}

// Has this lane's item arrived?  Returns its slot and gives the ticket up, or -1.
__device__ __forceinline__ int sq_poll(const StreamQueue &Q, int &tq, unsigned int tpos) {
    if (tq < 0) return -1;
    unsigned int *e = sq_entry(Q, tq, tpos);
    unsigned int v = ld_vol(e);
    if (v == 0u || (v & 0x80000000u) != sq_lap(Q, tpos)) return -1;
    st_vol(e, 0u);
    // no fence: the record loads that follow take their address from `v` (a dependent load cannot pass the load it
    // depends on) and bypass the L1 (.cg), and the producer fenced its record stores before it wrote the entry
    tq = -1;
    return (int)(v & 0x7fffffffu) - 1;
}

// ---- SHADE role: one warp --------------------------------------------------------------------------------------
template <bool BRUTE, bool COUNT>
__device__ __forceinline__ void stream_shade_warp(const StreamParams &p, unsigned int lane, unsigned int warp_gid) {
    const DScene &s = p.scene;
    const FrameConsts f = frame_consts(p.rd, p.cam);
    const unsigned int FULL = 0xffffffffu;
    const int tiles_x = (p.width + 7) >> 3, tiles_y = (p.local_rows + 3) >> 2;
    const unsigned int total_idx = (unsigned int)tiles_x * tiles_y * 32u;
    unsigned long long n_seg = 0, n_hit = 0, n_samples = 0;
    Counters cnt;
    cnt = Counters{};
    bool fresh_left = true;   // slots never used yet may remain (warp-uniform)
    bool frame_dry = false;   // the frame has handed out its last pixel (warp-uniform)
    unsigned int turn = warp_gid;
    int idle = 0;
    int tq = -1;              // this lane's ticket of the hit queue (sub-queue, position), -1 = none
    unsigned int tpos = 0;
    bool is_fresh = false;
    int spins = 0;
    unsigned int seen = 0;
    // utilisation of this role (the host re-balances the SM split with it): lane-cycles busy / lane-cycles total
    unsigned long long busy_cycles = 0;
    const long long t_start = clock64();
    long long t_batch = t_start;
    int n_prev = 0;
    for (;;) {
        {   // the batch that just ended kept n_prev lanes busy
            long long now = clock64();
            busy_cycles += (unsigned long long)(now - t_batch) * (unsigned int)n_prev;
            t_batch = now;
            n_prev = 0;
        }
        // ---- up to 32 slots: finished rays first, then (while the frame has pixels) slots never used ----
        // every lane holds a ticket of the hit queue; the lanes whose item has arrived form the batch
        sq_claim(p.hq, (int)(warp_gid & (RTB_STREAM_SHARDS - 1)), tq < 0, tq, tpos, lane);
        int slot = -1;
        unsigned int have = 0;
        for (int tries = 0;; tries++) {
            if (slot < 0) slot = sq_poll(p.hq, tq, tpos);
            have = __ballot_sync(FULL, slot >= 0);
            if (have == FULL) break;
            if (have && tries >= 4) break;            // a partial batch after a short wait: better than idling
            if (!have && tries >= 2) break;
        }
        int n_hq = __popc(have);
        int n_fresh = 0;
        if (n_hq < 32 && fresh_left && !frame_dry) {
            unsigned int base = 0;
            if (lane == 0) base = atomicAdd(p.ctl + SC_NEXT_SLOT, (unsigned int)(32 - n_hq));
            base = __shfl_sync(FULL, base, 0);
            if (base >= (unsigned int)p.num_slots) fresh_left = false;
            else n_fresh = min(32 - n_hq, p.num_slots - (int)base);
            int rank = __popc(~have & ((1u << lane) - 1u));   // my place among the lanes without a slot
            if (slot < 0 && rank < n_fresh) { slot = (int)base + rank; is_fresh = true; }
        }
        const int n_batch = n_hq + n_fresh;
        if (n_batch == 0) {
            if (ld_vol(p.ctl + SC_DONE)) break;
            if (stream_watchdog(p, spins, seen)) { stream_stalled(p); break; }
            idle = min(idle + 1, 6);
            __nanosleep(32u << idle);
            continue;
        }
        idle = 0;
        spins = 0;
        turn++;
        n_prev = n_batch;
        t_batch = clock64();
        bool live = slot >= 0;
        const bool fresh = is_fresh;
        is_fresh = false;
        RTB_ASSERT(!live || slot < p.num_slots);
        int array_index = -1;
        V3 pdir = v3(0, 0, 1);
        Path P;
        P.rng = 0; P.sample = 0; P.bounce = 0; P.cur_n = 1;
        P.thr = v3(1, 1, 1); P.col = v3(0, 0, 0); P.acc = v3(0, 0, 0);
        Ray sr;
        sr.o = f.cam_o; sr.d = pdir; sr.inv = pdir;
        unsigned int xy = 0;
        if (live && !fresh) {
            float4 a = pool_ld(p.st.A + slot), b = pool_ld(p.st.B + slot), c = pool_ld(p.st.C + slot), o = pool_ld(p.st.O + slot),
                   d = pool_ld(p.st.D + slot), h = pool_ld(p.st.H + slot);
            xy = __ldcg(p.pix + slot);
            int x = (int)(xy & 0xffffu), y = (int)(xy >> 16);
            array_index = (y * p.width + x) * 3;
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
        unsigned int wrote = 0;   // pixels this lane finished in this batch (several when a sample needs no ray at all)
        if (live) {
            for (;;) {
                if (array_index < 0) {
                    unsigned int idx = atomicAdd(p.ctl + SC_WORK, 1u);
                    if (idx >= total_idx) { ran_dry = true; live = false; break; }
                    unsigned int tile = idx >> 5, within = idx & 31u;
                    int x = (int)(tile % tiles_x) * 8 + (int)(within & 7u);
                    int ly = (int)(tile / tiles_x) * 4 + (int)(within >> 3);
                    if (x >= p.width || ly >= p.local_rows) continue;
                    int y = shard_row(p.shard_rank, p.shard_world, p.band_rows, ly);
                    if (y >= p.height) continue;
                    array_index = (y * p.width + x) * 3;  // raytracer.cu:123
                    RTB_ASSERT(array_index >= 0 && array_index + 2 < p.width * p.height * 3);
                    xy = (unsigned int)x | ((unsigned int)y << 16);
                    path_begin_pixel(P, array_index, p.time_ms);
                }
                if (P.bounce == 0) {
                    if (P.sample >= f.spp) {
                        path_write_pixel(P, f.spp, p.frame_num, p.prev, p.out, p.peers, array_index);
                        wrote++;
                        array_index = -1;
                        continue;
                    }
                    // the primary direction is needed at the start of a sample only (raytracer.cu:103)
                    pdir = primary_direction(p.cam, (int)(xy & 0xffffu), (int)(xy >> 16));
                    path_begin_sample(f, pdir, sr, P);
                    if (COUNT) n_samples++;
                    if (f.limit <= 0) { P.sample++; continue; }
                }
                break;
            }
        }
        if (__any_sync(FULL, ran_dry)) frame_dry = true;
        // finished pixels: the last one ends the frame for everybody
        {
            unsigned int w = __reduce_add_sync(FULL, wrote);
            if (w && lane == 0) {
                __threadfence();
                unsigned int before = atomicAdd(p.ctl + SC_PIXELS_DONE, w);
                if (before + w >= p.total_pixels) { __threadfence(); atomicExch(p.ctl + SC_DONE, 1u); }
            }
        }
        if (live) {
            ray_finalise(sr, f.aa, P.rng);  // antialias jitter, direction_inv
            pool_st(p.st.A + slot, make_float4(P.thr.x, P.thr.y, P.thr.z, P.cur_n));
            pool_st(p.st.B + slot, make_float4(P.col.x, P.col.y, P.col.z, __uint_as_float(P.rng)));
            pool_st(p.st.C + slot, make_float4(P.acc.x, P.acc.y, P.acc.z, __uint_as_float(((unsigned int)P.sample << 8) | (unsigned int)P.bounce)));
            __stcg(p.pix + slot, xy);
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
            pool_st(p.st.O + slot, make_float4(sr.o.x, sr.o.y, sr.o.z, __int_as_float(start)));
            pool_st(p.st.D + slot, make_float4(sr.d.x, sr.d.y, sr.d.z, __int_as_float(then)));
            pool_st(p.st.H + slot, make_float4(first.t, first.u, first.v, __int_as_float(first.code)));
        }
        sq_push(p.rq, (int)(turn & (RTB_STREAM_SHARDS - 1)), enter, slot, lane);
        sq_push(p.hq, (int)((turn + 3) & (RTB_STREAM_SHARDS - 1)), live && !enter, slot, lane);  // the hit is final already: shade again
    }
    if (lane == 0) {
        atomicAdd(p.stats + 0, busy_cycles);
        atomicAdd(p.stats + 1, (unsigned long long)(clock64() - t_start) * 32ull);
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

// ---- TRAVERSE role: one warp -----------------------------------------------------------------------------------
template <bool BRUTE, bool COUNT>
__device__ __forceinline__ void stream_traverse_warp(const StreamParams &p, unsigned int lane, unsigned int warp_gid) {
    const DScene &s = p.scene;
    const unsigned int FULL = 0xffffffffu;
    Counters cnt;
    cnt = Counters{};
    Ray ray;
    ray.o = v3(0, 0, 0); ray.d = v3(0, 0, 1); ray.inv = ray.d;
    Trav T;
    T.cur = RTB_DONE; T.leaf = 0; T.sp = 0; T.top = RTB_DONE;
    T.best.t = RTB_INF_F; T.best.u = 0; T.best.v = 0; T.best.code = -1;
    int slot = -1;   // the ray this lane owns, -1 = none
    unsigned int turn = warp_gid;
    int idle = 0;
    int tq = -1;     // this lane's ticket of the ray queue (sub-queue, position), -1 = none
    unsigned int tpos = 0;
    int spins = 0;
    unsigned int seen = 0;
    unsigned long long busy_cycles = 0;
    const long long t_start = clock64();
    long long t_last = t_start;
    int n_prev = 0;
    for (;;) {
        bool busy = slot >= 0 && !trav_done(T);
        unsigned int busy_mask = __ballot_sync(FULL, busy);
        {   // since the last look, n_prev lanes were walking
            long long now = clock64();
            busy_cycles += (unsigned long long)(now - t_last) * (unsigned int)n_prev;
            t_last = now;
            n_prev = __popc(busy_mask);
        }
        int free_lanes = 32 - __popc(busy_mask);
        if (free_lanes >= p.refill || !busy_mask) {
            // finished lanes hand their slot to the shading SMs ...
            bool retire = slot >= 0 && !busy;
            if (retire) pool_st(p.st.H + slot, make_float4(T.best.t, T.best.u, T.best.v, __int_as_float(T.best.code)));
            sq_push(p.hq, (int)(turn & (RTB_STREAM_SHARDS - 1)), retire, slot, lane);
            if (retire) slot = -1;
            // ... and the lanes without a ray draw tickets of the ray queue (those that hold one keep it)
            sq_claim(p.rq, (int)(warp_gid & (RTB_STREAM_SHARDS - 1)), slot < 0 && tq < 0, tq, tpos, lane);
            turn++;
        }
        // lanes with a ticket look for their ray (one coalesced load for the warp)
        if (__any_sync(FULL, tq >= 0)) {
            int got = sq_poll(p.rq, tq, tpos);
            if (got >= 0) {
                slot = got;
                RTB_ASSERT(slot < p.num_slots);
                float4 o = pool_ld(p.st.O + slot), d = pool_ld(p.st.D + slot), h = pool_ld(p.st.H + slot);
                ray.o = v3(o.x, o.y, o.z);
                ray.d = v3(d.x, d.y, d.z);
                ray.inv = v3(1 / d.x, 1 / d.y, 1 / d.z);  // ray.cu:198-202
                if (BRUTE) {
                    T.best = trace_brute(s, ray, cnt);
                    T.cur = RTB_DONE; T.leaf = 0;
                } else {
                    // the shading SM left the best large-primitive hit in H: only the tree remains
                    T.best.t = h.x; T.best.u = h.y; T.best.v = h.z; T.best.code = __float_as_int(h.w);
                    trav_start_below_root(T, __float_as_int(o.w), __float_as_int(d.w));
                }
            }
        }
        {
            busy = slot >= 0 && !trav_done(T);
            busy_mask = __ballot_sync(FULL, busy);
            n_prev = __popc(busy_mask);
            if (!busy_mask) {
                if (__any_sync(FULL, slot >= 0)) continue;   // rays that ended at once: retire them next round
                if (ld_vol(p.ctl + SC_DONE)) break;
                if (stream_watchdog(p, spins, seen)) { stream_stalled(p); break; }
                idle = min(idle + 1, 5);
                __nanosleep(32u << idle);
                continue;
            }
            idle = 0;
            spins = 0;
        }
        // all 32 lanes enter together (lanes without a ray hold a finished state); leave when `refill` lanes are free.
        // A warp the queue could not fill comes back after a while even if nobody finished, to look again.
        const bool waiting = __any_sync(FULL, tq >= 0);
        if (!BRUTE) trav_run<2>(s, ray, T, FULL, max(p.refill, 33 - __popc(busy_mask)), p.walk_min, cnt, waiting ? 24 : 0);
    }
    if (lane == 0) {
        atomicAdd(p.stats + 2, busy_cycles);
        atomicAdd(p.stats + 3, (unsigned long long)(clock64() - t_start) * 32ull);
    }
    if (COUNT && p.counters) {
        atomicAdd(p.counters + 2, cnt.box);
        atomicAdd(p.counters + 3, cnt.tri);
        atomicAdd(p.counters + 4, cnt.sph);
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

template <bool BRUTE, bool COUNT>
__global__ void __launch_bounds__(RTB_STREAM_THREADS, RTB_STREAM_MIN_BLOCKS) stream_render_kernel(const StreamParams p) {
    __shared__ int s_role;
    if (threadIdx.x == 0) {
        // the first block to arrive on an SM draws the SM's role; the SHADE SMs are spread evenly over the draw order
        unsigned int smid;
        asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
        int *role = p.sm_role + (smid & (RTB_STREAM_MAX_SMID - 1));
        int r = atomicCAS(role, 0, -1);
        if (r == 0) {
            unsigned int k = atomicAdd(p.ctl + SC_SM_COUNT, 1u);
            unsigned long long a = (unsigned long long)(k + 1) * p.shade_sms / p.total_sms, b = (unsigned long long)k * p.shade_sms / p.total_sms;
            r = (a != b) ? 1 : 2;   // 1 = SHADE, 2 = TRAVERSE
            __threadfence();
            atomicExch(role, r);
        } else {
            while (r == -1) r = ld_vol(role);
        }
        s_role = r;
    }
    __syncthreads();
    const unsigned int lane = threadIdx.x & 31u;
    const unsigned int warp_gid = blockIdx.x * (RTB_STREAM_THREADS / 32) + (threadIdx.x >> 5);
    if (s_role == 1) stream_shade_warp<BRUTE, COUNT>(p, lane, warp_gid);
    else stream_traverse_warp<BRUTE, COUNT>(p, lane, warp_gid);
}

cudaError_t launch_stream(const StreamParams &p, bool brute, bool count, int blocks, cudaStream_t st) {
    if (brute) {
        if (count) stream_render_kernel<true, true><<<blocks, RTB_STREAM_THREADS, 0, st>>>(p);
        else stream_render_kernel<true, false><<<blocks, RTB_STREAM_THREADS, 0, st>>>(p);
    } else {
        if (count) stream_render_kernel<false, true><<<blocks, RTB_STREAM_THREADS, 0, st>>>(p);
        else stream_render_kernel<false, false><<<blocks, RTB_STREAM_THREADS, 0, st>>>(p);
    }
    return cudaGetLastError();
}

cudaError_t stream_occupancy(int *blocks_per_sm) {
    return cudaOccupancyMaxActiveBlocksPerMultiprocessor(blocks_per_sm, stream_render_kernel<false, false>, RTB_STREAM_THREADS, 0);
}

int stream_block_threads() { return RTB_STREAM_THREADS; }
