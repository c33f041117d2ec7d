// K4w: wavefront form of the glTF path tracer (same arithmetic as gltf_kernel.cuh, different
// scheduling).  ncu on the megakernel showed 7.5 of 32 lanes active on average (profiles/
// r01_megakernel_*): incoherent rays make per-warp lockstep traversal idle most lanes.  Here every
// bounce is split into stages that exchange path indices through device queues:
//
//   generate  : camera rays of one wave (all pixels x S_w sample indices)           -> queue A
//   extend    : closest hit, PERSISTENT warps that refill idle lanes from the queue -> hit[]
//   shade     : material/textures/BRDF/sampling; emission into rad[]               -> queue M
//   light     : FiguresMix::getTotalPdf query, persistent like extend; then the pdf mean, weight
//               and firefly clamp                                                  -> queue B (next bounce)
//   accumulate: total[pixel] += sum over the wave's samples of rad[] (fixed order, no float atomics)
//
// Inside the persistent kernels a warp alternates between an "inner node" phase and a "leaf"
// phase chosen by ballot, so the two code paths are never both executed for one step, and lanes
// whose query finished are refilled every B200RT_WF_BATCH steps.
// Reference code replaced: same as gltf_kernel.cuh (hw8/src/scene.cpp:84-187, bvh.h:111-142,
// distributions.h:122-165).
#pragma once
#include "gltf_kernel.cuh"
#include "hw6_kernel.cuh"      // hw6_surface / hw6_finish: the hw6 integrator as a shade stage
#include "wave_state.h"

namespace b200rt {

#ifndef B200RT_EXT_MINB
#define B200RT_EXT_MINB 10       /* min resident blocks per SM asked of ptxas for extend (48 registers; profiles/r02_sweep_hit_on_accept.log) */
#endif
#ifndef B200RT_LIGHT_MINB
#define B200RT_LIGHT_MINB 9      /* ... and for light (56 registers; sweep: profiles/r01_sweep_quantised_nodes.log) */
#endif
#ifndef B200RT_SHADE_MINB
#define B200RT_SHADE_MINB 4
#endif
#ifndef B200RT_EXT_UNROLL
#define B200RT_EXT_UNROLL 2      /* box steps per phase decision in wf_extend's inner phase */
#endif
#define B200RT_WF_MISS 0x7FFFFFFFu
#define B200RT_FULL 0xffffffffu

// Wave state accessors: plain caching.  Evict-first (.cs) accesses were measured: extend and light gain 1 %, shade
// loses 15 % (it re-reads what the previous stage just wrote).
__device__ __forceinline__ float4 ld_state(const float4 *p) {
    return *p;
}
__device__ __forceinline__ void st_state(float4 *p, const float4 v) {
    *p = v;
}

__device__ __forceinline__ bool slot_to_pixel(const WaveParams &wp, int slot, int &px, int &py, int &s_local) {
    const int per = wp.n_tiles * 32;
    s_local = slot / per;
    const int r = slot - s_local * per;
    const int tile = r >> 5, lane = r & 31;
    px = (tile % wp.tiles_x) * 8 + (lane & 7);
    py = (tile / wp.tiles_x) * 4 + (lane >> 3);
    return px < wp.cam.width && py < wp.cam.height;
}

// Warp-aggregated append; every lane of the warp must call it (converged).
__device__ __forceinline__ void queue_push(int *queue, int *counter, bool want, int value) {
    const unsigned m = __ballot_sync(B200RT_FULL, want);
    if (m == 0) return;
    const int lane = threadIdx.x & 31;
    const int leader = __ffs(m) - 1;
    int base = 0;
    if (lane == leader) base = atomicAdd(counter, __popc(m));
    base = __shfl_sync(B200RT_FULL, base, leader);
    if (want) __stcs(queue + base + __popc(m & ((1u << lane) - 1u)), value);
}

// Append to two queues whose counters are the low / high half of one aligned 64-bit word: one atomic round trip
// per warp instead of two (ncu: the returning atomics were 10 % of wf_shade's stall samples).
__device__ __forceinline__ void queue_push2(int *q_lo, int *q_hi, int *counter_lo, bool want_lo, bool want_hi, int value_lo, int value_hi) {
    const unsigned m_lo = __ballot_sync(B200RT_FULL, want_lo), m_hi = __ballot_sync(B200RT_FULL, want_hi);
    if ((m_lo | m_hi) == 0) return;
    const int lane = threadIdx.x & 31;
    unsigned long long base = 0;
    if (lane == 0) base = atomicAdd(reinterpret_cast<unsigned long long *>(counter_lo), (unsigned long long)__popc(m_lo) | ((unsigned long long)__popc(m_hi) << 32));
    base = __shfl_sync(B200RT_FULL, base, 0);
    const unsigned below = (1u << lane) - 1u;
    if (want_lo) __stcs(q_lo + (int)(unsigned)base + __popc(m_lo & below), value_lo);
    if (want_hi) __stcs(q_hi + (int)(unsigned)(base >> 32) + __popc(m_hi & below), value_hi);
}

// Prefetch one 32-byte sector into L2 (the wave state streams through DRAM; whoever knows a path index early
// turns the later demand load into an L2 hit).
__device__ __forceinline__ void prefetch_l2(const void *p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }
#ifndef B200RT_REFILL_PF
#define B200RT_REFILL_PF 2048   /* wf_extend / wf_light: a lane that takes queue entry i prefetches the ray of entry i + this (0 = off) */
#endif

// Traversal stack of the persistent kernels: per-thread local memory.  ncu on wf_extend: it makes 46 % of the L1
// requests and its write-through stores 20 % of the L2 sectors -- yet keeping the first 8-32 entries in shared memory
// was 7 % SLOWER (profiles/r01_sweep_smem_stack.log), as was storing the entry distance for pruning on pop.
struct TravStack {
    uint32_t deep[B200RT_STACK];
    __device__ __forceinline__ void push(int &sp, uint32_t v) { deep[sp++] = v; }
    __device__ __forceinline__ uint32_t pop(int &sp) { return deep[--sp]; }
};

// full = 0: only the ray direction is stored -- the first extend / shade of the wave (`first`) know the rest of a
// camera path's state without reading it (origin = camera, throughput 1, radiance 0, Philox key / counter from the slot).
__global__ void wf_generate(const WaveParams wp, const WaveBuffers wb, DCounters *counters, const bool dense, const bool full, const bool unit_dir = true) {
    const int n_slots = wp.n_tiles * 32 * wp.s_count;
    const int warps = (gridDim.x * blockDim.x) >> 5;
    const int lane = threadIdx.x & 31;
    const V3 cpos = ld3(wp.cam.pos), cright = ld3(wp.cam.right), cup = ld3(wp.cam.up), cfwd = ld3(wp.cam.forward);
    unsigned made = 0;
    for (int base = ((blockIdx.x * blockDim.x + threadIdx.x) >> 5) * 32; base < n_slots; base += warps * 32) {
        const int slot = base + lane;
        int px, py, sl;
        const bool ok = slot_to_pixel(wp, slot, px, py, sl);
        if (ok) {
            const uint32_t pixel = (uint32_t)(py * wp.cam.width + px);
            const uint4 j = philox4x32_10((uint32_t)(wp.s_begin + sl), 0u, pixel, wp.seed);
            const float fx = (float)px + u01(j.x), fy = (float)py + u01(j.y);
            const float nx = wp.cam.tan_x * (2.f * fx / (float)wp.cam.width - 1.f);
            const float ny = wp.cam.tan_y * (2.f * fy / (float)wp.cam.height - 1.f);
            const V3 dv = nx * cright - ny * cup + cfwd;
            const V3 d = unit_dir ? normalize(dv) : dv;      // hw6's camera does not normalise (hw6/src/scene.cpp:107-126)
            st_state(&wb.ray_d[slot], make_float4(d.x, d.y, d.z, 0.f));
            if (full) {
            st_state(&wb.ray_o[slot], make_float4(cpos.x, cpos.y, cpos.z, __uint_as_float(pixel)));
            // whole 32-byte sectors ({hit, thr} and {rad, bsdf}): a half-written sector costs L2 a fill from DRAM first
            st_state(&wb.hit[slot], make_float4(0.f, 0.f, 0.f, 0.f));
            st_state(&wb.thr[slot], make_float4(1.f, 1.f, 1.f, 0.f));
            st_state(&wb.rad[slot], make_float4(0.f, 0.f, 0.f, __uint_as_float((uint32_t)(wp.s_begin + sl))));
            st_state(&wb.bsdf[slot], make_float4(0.f, 0.f, 0.f, 0.f));
            }
            made++;
        }
        // Frames whose width / height are multiples of the 8x4 tile have no dead slots: the first queue is the
        // identity and its length is known, no atomics needed (260 k same-address atomics per 4K wave otherwise).
        if (dense) { if (ok) __stcs(wb.queue[0] + slot, slot); }
        else queue_push(wb.queue[0], wb.ctr - 1, ok, slot);
    }
    if (dense && blockIdx.x == 0 && threadIdx.x == 0) wb.ctr[-1] = n_slots;
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) made += __shfl_down_sync(B200RT_FULL, made, off);
    if (lane == 0 && made) atomicAdd(&counters->paths, (unsigned long long)made);
}

#ifdef B200RT_DBG_PHASE
// development instrumentation (off by default): warp-level phase statistics of wf_extend
//   0 inner warp-steps, 1 lanes at inner in those, 2 leaf warp-steps, 3 lanes at leaf in those,
//   4 triangle iterations (warp), 5 refills, 6 lanes waiting at a leaf during inner steps, 7 idle lanes during inner steps
__device__ unsigned long long g_phase[8];
#define DBG_PHASE(i, v) do { if (lane == 0) dbg[i] += (v); } while (0)
#else
#define DBG_PHASE(i, v) do { } while (0)
#endif

// ---- persistent closest-hit traversal ------------------------------------------------------------
template <bool COUNT>
__global__ void __launch_bounds__(128, B200RT_EXT_MINB) wf_extend(const DGltfScene sc, const WaveBuffers wb, const int *__restrict__ q_in,
                                                 const int *__restrict__ n_in, int *cursor, DCounters *counters,
                                                 const int batch, const int leaf_t, const int first = 0,
                                                 const float cam_x = 0.f, const float cam_y = 0.f, const float cam_z = 0.f) {
    const int n = *n_in;
    const int lane = threadIdx.x & 31;
    TravStack stack;
    int sp = 0;
    bool active = false, exhausted = false;
    int p = 0;
    V3 o = mk(0, 0, 0), d = mk(0, 0, 1);
    QRay qr = {};
    float tbest = 0.f;
    uint32_t cur = 0;
    unsigned node_tests = 0, tri_tests = 0;
#ifdef B200RT_DBG_PHASE
    unsigned long long dbg[8] = {};
#endif

    while (true) {
        const unsigned idle = __ballot_sync(B200RT_FULL, !active);
        if (!exhausted && idle) {
            const int n_idle = __popc(idle), leader = __ffs(idle) - 1;
            int base = 0;
            if (lane == leader) base = atomicAdd(cursor, n_idle);
            base = __shfl_sync(B200RT_FULL, base, leader);
            if (!active) {
                const int idx = base + __popc(idle & ((1u << lane) - 1u));
                if (idx < n) {
                    p = __ldcs(q_in + idx);
#if B200RT_REFILL_PF
                    const int ahead = idx + B200RT_REFILL_PF < n ? __ldcs(q_in + idx + B200RT_REFILL_PF) : -1;
#endif
                    const float4 rd = ld_state(&wb.ray_d[p]);
                    if (first) o = mk(cam_x, cam_y, cam_z);      // camera rays: the origin is not stored (wf_generate)
                    else { const float4 ro = ld_state(&wb.ray_o[p]); o = mk(ro.x, ro.y, ro.z); }
#if B200RT_REFILL_PF
                    if (ahead >= 0) prefetch_l2(&wb.ray_o[ahead]);     // arrived together with p: no extra wait
#endif
                    d = mk(rd.x, rd.y, rd.z);
                    qr = make_qray(sc.q, o, d);
                    tbest = B200RT_TMAX;
                    sp = 0;
                    cur = sc.root;
                    active = true;
                }
            }
            if (base + n_idle >= n) exhausted = true;
            DBG_PHASE(5, 1);
        }
        if (__ballot_sync(B200RT_FULL, active) == 0) break;

        for (int it = 0; it < batch; it++) {
            const bool at_leaf = active && (cur & B200RT_LEAF);
            const bool at_inner = active && !(cur & B200RT_LEAF);
            const unsigned leaf_mask = __ballot_sync(B200RT_FULL, at_leaf);
            const unsigned inner_mask = __ballot_sync(B200RT_FULL, at_inner);
            if ((leaf_mask | inner_mask) == 0) break;
#ifdef B200RT_DBG_PHASE
            if (inner_mask == 0 || __popc(leaf_mask) >= leaf_t) { DBG_PHASE(2, 1); DBG_PHASE(3, __popc(leaf_mask)); }
            else { DBG_PHASE(0, 1); DBG_PHASE(1, __popc(inner_mask)); DBG_PHASE(6, __popc(leaf_mask)); DBG_PHASE(7, 32 - __popc(leaf_mask | inner_mask)); }
#endif
            // end of a subtree: next stacked node, or the ray is finished.  The hit record is written whenever a closer
            // hit is accepted, not carried to the end of the ray: four registers less through the box steps, which is
            // what lets the kernel run 10 blocks per SM at 48 registers without spills (a miss is written here)
            auto do_pop = [&]() {
                bool got = false;
                if (sp > 0) { cur = stack.pop(sp); got = true; }
                if (!got) {
                    if (tbest == B200RT_TMAX) st_state(&wb.hit[p], make_float4(B200RT_TMAX, 0.f, 0.f, __uint_as_float(B200RT_WF_MISS)));
                    active = false;
                }
            };
            if (inner_mask == 0 || __popc(leaf_mask) >= leaf_t) {
                if (at_leaf) {
                    const uint32_t first = cur & 0x07FFFFFFu, count = (cur >> 27) & 15u;
                    for (uint32_t i = first; i < first + count; i++) {
                        const float4 *q = reinterpret_cast<const float4 *>(sc.tris + i);
                        const float4 v0 = __ldg(q), e1 = __ldg(q + 1), e2 = __ldg(q + 2);
                        if (COUNT) tri_tests++;
                        float t, u, v, det;
                        if (tri_hit(v0, e1, e2, o, d, tbest, t, u, v, det)) {
                            tbest = t;
                            st_state(&wb.hit[p], make_float4(t, u, v, __uint_as_float(i | (det > 0.f ? 0x80000000u : 0u))));
                        }
                    }
                    do_pop();
                }
            } else {
                // inner phase: B200RT_EXT_UNROLL box steps per phase decision (the ballots and the phase choice are
                // a quarter of a step's instructions); lanes that reach a leaf or finish sit out the remaining ones
#pragma unroll
                for (int k = 0; k < B200RT_EXT_UNROLL; k++) {
                    if (active && !(cur & B200RT_LEAF)) {
                        float tl, tr;
                        bool hl, hr;
                        uint32_t ml, mr;
                        pair_test(sc.nodes, cur, qr, tbest, hl, hr, tl, tr, ml, mr);
                        if (COUNT) node_tests += 2;
                        if (hl && hr) {
                            const bool left_first = tl <= tr;
                            // no bound check: b200rt_upload_scene rejects trees deeper than B200RT_STACK
                            stack.push(sp, left_first ? mr : ml);
                            cur = left_first ? ml : mr;
                        } else if (hl) cur = ml;
                        else if (hr) cur = mr;
                        else do_pop();
                    }
                }
            }
        }
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(&counters->rays, (unsigned long long)n);
#ifdef B200RT_DBG_PHASE
    if (lane == 0) for (int k = 0; k < 8; k++) if (dbg[k]) atomicAdd(&g_phase[k], dbg[k]);
#endif
    if (COUNT) {
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) {
            node_tests += __shfl_down_sync(B200RT_FULL, node_tests, off);
            tri_tests += __shfl_down_sync(B200RT_FULL, tri_tests, off);
        }
        if (lane == 0) {
            atomicAdd(&counters->node_tests, (unsigned long long)node_tests);
            atomicAdd(&counters->tri_tests, (unsigned long long)tri_tests);
        }
    }
}

// ---- shade --------------------------------------------------------------------------------------
// Two latency-hiding variants of this kernel were built, measured slower and removed from the source (numbers in
// DESIGN.md section 4): a cp.async software pipeline of the queue entry -> path state -> triangle / shading record
// hops through shared memory (+20 % shade time: the material -> texel hops that remain dominate, the staging shrinks
// L1 and pushes the 128-register build into spills), and prefetch.global.L2 of the next items' state records and hit
// records (+8 % / +24 %).
// LAST = the path's final bounce: only the emission of the hit is added (surface_emission), nothing continues.
// HW6 = hw6's integrator (hw6_kernel.cuh) instead of the metallic-roughness surface: a diffuse hit goes on to the light
// stage like every hw7 / hw8 hit, a mirror / dielectric hit has its weight already and joins the next bounce directly.
template <bool HW8, bool COUNT, bool LAST = false, bool HW6 = false>
__global__ void __launch_bounds__(128, COUNT ? 3 : B200RT_SHADE_MINB) wf_shade(const DGltfScene sc, const WaveBuffers wb, const WaveParams wp, const int *__restrict__ q_in,
                                                const int *__restrict__ n_in, int *q_out, int *n_out, int *q_direct, int *n_direct, int bounce,
                                                const float *__restrict__ lut_srgb_g, DCounters *counters, const int drop_dark = 0, const int first = 0,
                                                const int pack_mask = 1) {
    __shared__ float lut_srgb[256];
    __shared__ float lut_lin[256];
    for (int i = threadIdx.x; i < 256; i += blockDim.x) {
        lut_srgb[i] = lut_srgb_g[i];
        lut_lin[i] = 0.00392156862745098f * (float)i;
    }
    __syncthreads();
    const int n = *n_in;
    const int warps = (gridDim.x * blockDim.x) >> 5;
    const int lane = threadIdx.x & 31;
    const bool has_lights = sc.n_light_tris > 0;
    ThreadCounters tc = {};
    unsigned answered = 0;      // light queries answered here by the pre-filter (they stay light queries in the stats)
    for (int base = ((blockIdx.x * blockDim.x + threadIdx.x) >> 5) * 32; base < n; base += warps * 32) {
        const bool live = base + lane < n;
        bool cont = false, direct = false;
        int p = 0;
        uint32_t light_mask = 0u;     // pierced boxes of the light cut: rides in the top bits of the light-queue entry
        if (live) {
            p = __ldcs(q_in + base + lane);
            const float4 h = ld_state(&wb.hit[p]);
            const float4 rd = ld_state(&wb.ray_d[p]);
            float4 ro, th, ra;
            if (first) {
                // camera paths: origin = camera, throughput 1, radiance 0; Philox key = pixel, counter = sample index
                int px, py, sl;
                slot_to_pixel(wp, p, px, py, sl);
                ro = make_float4(wp.cam.pos[0], wp.cam.pos[1], wp.cam.pos[2], __uint_as_float((uint32_t)(py * wp.cam.width + px)));
                th = make_float4(1.f, 1.f, 1.f, 0.f);
                ra = make_float4(0.f, 0.f, 0.f, __uint_as_float((uint32_t)(wp.s_begin + sl)));
            } else {
                ro = ld_state(&wb.ray_o[p]);
                th = ld_state(&wb.thr[p]);
                ra = ld_state(&wb.rad[p]);
            }
            const V3 o = mk(ro.x, ro.y, ro.z), d = mk(rd.x, rd.y, rd.z);
            V3 T = mk(th.x, th.y, th.z), L = mk(ra.x, ra.y, ra.z);
            const uint32_t bits = __float_as_uint(h.w);
            if (bits == B200RT_WF_MISS) {
                V3 bg = ld3(sc.bg);
                if (HW8 && sc.env) {
                    const float tx = 0.5f + 0.5f * atan2f(d.z, d.x) * 0.318309886183791f;
                    const float ty = 0.5f - asinf(d.y) * 0.318309886183791f;
                    bg = sample_texture(sc.env, sc.env_w, sc.env_h, tx, ty, lut_srgb);
                }
                L = L + T * bg;
            } else if (LAST) {
                L = L + T * surface_emission<HW8, COUNT>(sc, lut_srgb, (int)(bits & 0x7FFFFFFFu), h.y, h.z, tc);
            } else {
                // key = pixel, counter = (global sample index, block of this bounce); both ride in spare state lanes
                const uint4 rnd = philox4x32_10(__float_as_uint(ra.w), 16u * (uint32_t)(bounce + 1), __float_as_uint(ro.w), wp.seed);
                BounceOut bo;
                bool specular = false;
                if constexpr (HW6) {
                    cont = hw6_surface<COUNT>(sc, o, d, h.x, (int)(bits & 0x7FFFFFFFu), (bits >> 31) != 0, rnd, L, T, tc, bo, specular);
                    if (cont && specular && q_direct) direct = true;      // (without emitters q_out IS the next bounce's queue)
                    if (cont && !specular && !has_lights) hw6_finish(bo.brdf, bo.pdf_partial, bo.absdot, 0.f, 0, T);
                } else {
                    SurfRec rec;
                    load_surf_rec<HW8>(sc, (int)(bits & 0x7FFFFFFFu), rec);
                    cont = shade_surface<HW8, COUNT>(sc, lut_srgb, lut_lin, o, d, h.x, h.y, h.z, rec, (bits >> 31) != 0, L, T, rnd, tc, bo);
                    if (cont && !has_lights) cont = finish_bounce(bo.brdf, bo.pdf_partial, bo.absdot, 0.f, 0, T, sc.weight_clamp);
                }
                if (cont && !specular && has_lights && q_direct) {
                    // A ray that misses the boxes around the emitters has a light pdf of exactly 0
                    // (light_boxes_hit): the bounce is finished here and the path skips the light stage
                    // (its queue entry, bsdf round trip and refill slot).
                    light_mask = light_boxes_mask(sc, bo.xo, bo.nd);
                    if (!light_mask) {
                        if (COUNT) tc.light_node_tests++;
                        answered++;
                        direct = true;
                        if constexpr (HW6) hw6_finish(bo.brdf, bo.pdf_partial, bo.absdot, 0.f, sc.n_light_tris, T);
                        else cont = finish_bounce(bo.brdf, bo.pdf_partial, bo.absdot, 0.f, sc.n_light_tris, T, sc.weight_clamp);
                        bo.pdf_partial = 0.f;
                        // drop_dark (set when the NEXT bounce is the last one and the background is black): that
                        // bounce only adds the emission of what it hits, and a ray that misses the emitters' boxes
                        // cannot hit an emitter -- the path would add exactly zero, so it ends here
                        if (drop_dark) cont = false;
                    }
                }
                if (cont) {
                    st_state(&wb.ray_o[p], make_float4(bo.xo.x, bo.xo.y, bo.xo.z, ro.w));
                    st_state(&wb.ray_d[p], make_float4(bo.nd.x, bo.nd.y, bo.nd.z, bo.absdot));
                    st_state(&wb.thr[p], make_float4(T.x, T.y, T.z, bo.pdf_partial));
                    if (has_lights && !direct) st_state(&wb.bsdf[p], make_float4(bo.brdf.x, bo.brdf.y, bo.brdf.z, 0.f));
                }
            }
            st_state(&wb.rad[p], make_float4(L.x, L.y, L.z, ra.w));
        }
        if (q_direct) queue_push2(q_out, q_direct, n_out, cont && !direct, cont && direct, (int)((uint32_t)p | (pack_mask ? light_mask << B200RT_WF_MASK_SHIFT : 0u)), p);   // n_direct == n_out + 1
        else queue_push(q_out, n_out, cont && !direct, p);
    }
    if (q_direct) {
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) answered += __shfl_down_sync(B200RT_FULL, answered, off);
        if (lane == 0 && answered) atomicAdd(&counters->light_queries, (unsigned long long)answered);
    }
    if (COUNT) flush_counters(counters, tc);
}

// ---- persistent light-pdf traversal + end of the bounce ---------------------------------------------
template <bool COUNT>
__global__ void __launch_bounds__(128, B200RT_LIGHT_MINB) wf_light(const DGltfScene sc, const WaveBuffers wb, const int *__restrict__ q_in,
                                                const int *__restrict__ n_in, int *cursor, int *q_out, int *n_out, DCounters *counters,
                                                const int batch, const int leaf_t, const bool filtered, const int drop_dark = 0) {
    const int n = *n_in;
    const int lane = threadIdx.x & 31;
    TravStack stack;
    int sp = 0;
    bool active = false, exhausted = false;
    int p = 0;
    V3 o = mk(0, 0, 0), d = mk(0, 0, 1);
    QRay qr = {};
    float total = 0.f;
    uint32_t cur = 0;
    unsigned node_tests = 0, tri_tests = 0;
    bool push_pending = false;
    int push_p = 0;

    while (true) {
        // converged point: publish the paths that finished during the last batch
        queue_push(q_out, n_out, push_pending, push_p);
        push_pending = false;
        const unsigned idle = __ballot_sync(B200RT_FULL, !active);
        if (!exhausted && idle) {
            const int n_idle = __popc(idle), leader = __ffs(idle) - 1;
            int base = 0;
            if (lane == leader) base = atomicAdd(cursor, n_idle);
            base = __shfl_sync(B200RT_FULL, base, leader);
            if (!active) {
                const int idx = base + __popc(idle & ((1u << lane) - 1u));
                if (idx < n) {
                    // `filtered`: the entry carries the pierced-cut mask above bit 26 (the host packs it only for waves
                    // below 2^26 slots); otherwise the entry IS the slot and must not be masked (a frame above 67 M
                    // pixels has slots >= 2^26: masking aliased them onto low slots)
                    const uint32_t slot_mask = filtered ? (1u << B200RT_WF_MASK_SHIFT) - 1u : 0xFFFFFFFFu;
                    const uint32_t entry = (uint32_t)__ldcs(q_in + idx);
                    p = (int)(entry & slot_mask);
#if B200RT_REFILL_PF
                    const int ahead = idx + B200RT_REFILL_PF < n ? (int)((uint32_t)__ldcs(q_in + idx + B200RT_REFILL_PF) & slot_mask) : -1;
#endif
                    const float4 ro = ld_state(&wb.ray_o[p]), rd = ld_state(&wb.ray_d[p]);
#if B200RT_REFILL_PF
                    if (ahead >= 0) prefetch_l2(&wb.ray_o[ahead]);
                    prefetch_l2(&wb.thr[p]); prefetch_l2(&wb.bsdf[p]);   // read when this query ends
#endif
                    o = mk(ro.x, ro.y, ro.z); d = mk(rd.x, rd.y, rd.z);
                    qr = make_qray(sc.lq, o, d);
                    total = 0.f; sp = 0;
                    active = true;
                    // the light root's box is a real filter (unlike the scene root); `filtered`: shade already
                    // applied light_boxes_hit to every query in this queue
                    if (COUNT) node_tests++;
                    // (the mask of pierced cut boxes came with the queue entry); the traversal starts at those boxes
                    const uint32_t mask = filtered ? entry >> B200RT_WF_MASK_SHIFT : light_boxes_mask(sc, o, d);
                    if (mask) {      // light_start() through this kernel's own stack type
                        if (!sc.light_top_start || sc.n_light_top == 0) cur = sc.light_root;
                        else {
#pragma unroll
                            for (int k = B200RT_LIGHT_TOP - 1; k >= 0; k--)
                                if (mask & (1u << k)) stack.push(sp, sc.light_top_word[k]);
                            cur = stack.pop(sp);
                        }
                    }
                    else cur = B200RT_LEAF;      // empty leaf: finishes on the next leaf phase
                }
            }
            if (base + n_idle >= n) exhausted = true;
        }
        if (__ballot_sync(B200RT_FULL, active) == 0) break;

        for (int it = 0; it < batch; it++) {
            const bool at_leaf = active && (cur & B200RT_LEAF);
            const bool at_inner = active && !(cur & B200RT_LEAF);
            const unsigned leaf_mask = __ballot_sync(B200RT_FULL, at_leaf);
            const unsigned inner_mask = __ballot_sync(B200RT_FULL, at_inner);
            if ((leaf_mask | inner_mask) == 0) break;
            bool pop = false;
            if (inner_mask == 0 || __popc(leaf_mask) >= leaf_t) {
                if (at_leaf) {
                    const uint32_t first = cur & 0x07FFFFFFu, count = (cur >> 27) & 15u;
                    for (uint32_t i = first; i < first + count; i++) {
                        const float4 *q = reinterpret_cast<const float4 *>(sc.light_tris + i);
                        const float4 v0 = __ldg(q), e1 = __ldg(q + 1), e2 = __ldg(q + 2);
                        if (COUNT) tri_tests++;
                        float t, u, v, det;
                        if (tri_hit(v0, e1, e2, o, d, B200RT_TMAX, t, u, v, det)) total = __fadd_rn(total, light_term(sc, i, v0, e1, e2, d, t, u, v));
                    }
                    pop = true;
                }
            } else if (at_inner) {
                float tl, tr;
                bool hl, hr;
                uint32_t ml, mr;
                pair_test(sc.light_nodes, cur, qr, INFINITY, hl, hr, tl, tr, ml, mr);
                if (COUNT) node_tests += 2;
                if (hl && hr) {
                    stack.push(sp, mr);
                    cur = ml;
                } else if (hl) cur = ml;
                else if (hr) cur = mr;
                else pop = true;
            }
            if (pop) {
                if (sp > 0) cur = stack.pop(sp);
                else {
                    // end of the bounce: mean pdf, weight, firefly clamp (scene.cpp:158-163)
                    const float4 th = ld_state(&wb.thr[p]), bs = ld_state(&wb.bsdf[p]);
                    V3 T = mk(th.x, th.y, th.z);
                    // drop_dark: see wf_shade.  total == 0 means the ray pierces no emissive triangle within T_MAX (every
                    // hit adds a positive term), so the last bounce would add nothing: the path ends here
                    bool go;
                    if (sc.hw6) { hw6_finish(mk(bs.x, bs.y, bs.z), th.w, ld_state(&wb.ray_d[p]).w, total, sc.n_light_tris, T); go = true; }
                    else go = finish_bounce(mk(bs.x, bs.y, bs.z), th.w, ld_state(&wb.ray_d[p]).w, total, sc.n_light_tris, T, sc.weight_clamp);
                    if (go && !(drop_dark && total == 0.f)) {
                        st_state(&wb.thr[p], make_float4(T.x, T.y, T.z, 0.f));
                        push_pending = true;
                        push_p = p;
                    }
                    active = false;
                }
            }
        }
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(&counters->light_queries, (unsigned long long)n);
    if (COUNT) {
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) {
            node_tests += __shfl_down_sync(B200RT_FULL, node_tests, off);
            tri_tests += __shfl_down_sync(B200RT_FULL, tri_tests, off);
        }
        if (lane == 0) {
            atomicAdd(&counters->light_node_tests, (unsigned long long)node_tests);
            atomicAdd(&counters->light_tri_tests, (unsigned long long)tri_tests);
        }
    }
}

// total[pixel] += sum of the wave's samples, in sample order (deterministic)
__global__ void wf_accumulate(const WaveParams wp, const WaveBuffers wb, float *__restrict__ total) {
    const int per = wp.n_tiles * 32;
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= per) return;
    int px, py, sl;
    if (!slot_to_pixel(wp, r, px, py, sl)) return;
    float3 acc = make_float3(0.f, 0.f, 0.f);
    for (int s = 0; s < wp.s_count; s++) {
        const float4 v = ld_state(&wb.rad[(size_t)s * per + r]);
        acc.x += v.x; acc.y += v.y; acc.z += v.z;
    }
    float *dst = total + ((size_t)py * wp.cam.width + px) * 3;
    dst[0] += acc.x; dst[1] += acc.y; dst[2] += acc.z;
}

}  // namespace b200rt
