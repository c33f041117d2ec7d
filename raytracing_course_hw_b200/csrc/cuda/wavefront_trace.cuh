// K4w `wf_trace`: ONE persistent traversal kernel per bounce that answers both queries a ray carries:
//   phase 0  closest hit in the scene BVH                      (Scene::intersect, hw8/src/include/bvh.h:111-142)
//   phase 1  FiguresMix::getTotalPdf over the light BVH         (hw8/src/include/distributions.h:122-165)
//            for the SAME ray -- the direction sampled at the previous bounce is both the ray whose
//            light pdf the estimator needs and the next ray to trace -- followed by the end of that
//            bounce: mean pdf, weight, firefly clamp             (hw8/src/scene.cpp:158-163).
// Fusing the two removes a kernel, a queue and a ray reload per bounce, and lets light-BVH steps
// share warp steps with scene-BVH steps (the inner-node step is identical code on a different array).
// A path whose weight is clamped gets hit = B200RT_WF_DEAD; its (already traced) ray is discarded,
// which is what the reference does by returning before the recursive call.
#pragma once
#include "wavefront.cuh"

namespace b200rt {

template <bool COUNT>
__global__ void __launch_bounds__(128, B200RT_EXT_MINB) wf_trace(const DGltfScene sc, const WaveBuffers wb, const int *__restrict__ q_in,
                                                                  const int *__restrict__ n_in, int *cursor, DCounters *counters,
                                                                  const int batch, const int leaf_t, const int do_light) {
    const int n = *n_in;
    const int lane = threadIdx.x & 31;
    uint32_t stack[B200RT_STACK];
    int sp = 0;
    bool active = false, exhausted = false, light = false;
    int p = 0;
    V3 o = mk(0, 0, 0), d = mk(0, 0, 1);
    QRay qr = {};
    float tbest = 0.f, hu = 0.f, hv = 0.f, hdet = 0.f, total = 0.f;
    uint32_t htri = B200RT_WF_MISS, cur = 0;
    unsigned node_tests = 0, tri_tests = 0, lnode_tests = 0, ltri_tests = 0;
    unsigned dead = 0;      // paths the clamp ended: their ray was traced speculatively and is not counted

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
                    const float4 ro = ld_state(&wb.ray_o[p]), rd = ld_state(&wb.ray_d[p]);
                    o = mk(ro.x, ro.y, ro.z); d = mk(rd.x, rd.y, rd.z);
                    qr = make_qray(sc.q, o, d);
                    tbest = B200RT_TMAX; htri = B200RT_WF_MISS; hu = hv = hdet = 0.f;
                    sp = 0;
                    cur = sc.root;
                    light = false;
                    active = true;
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
                    const DTri *tris = light ? sc.light_tris : sc.tris;
                    for (uint32_t i = first; i < first + count; i++) {
                        const float4 *q = reinterpret_cast<const float4 *>(tris + i);
                        const float4 v0 = __ldg(q), e1 = __ldg(q + 1), e2 = __ldg(q + 2);
                        if (COUNT) { if (light) ltri_tests++; else tri_tests++; }
                        float t, u, v, det;
                        if (tri_hit(v0, e1, e2, o, d, light ? B200RT_TMAX : tbest, t, u, v, det)) {
                            const V3 E1 = mk(e1.x, e1.y, e1.z), E2 = mk(e2.x, e2.y, e2.z);

                            if (!light) {
                                tbest = t; hu = u; hv = v; htri = i; hdet = det;
                            } else {
                                V3 yn;
                                if (sc.hw8) {   // interpolated shading normal of the emitter (hw8 distributions.h:140-145)
                                    const float4 *nq = reinterpret_cast<const float4 *>(sc.light_normals + i);
                                    const float4 n0 = __ldg(nq), n1 = __ldg(nq + 1), n2 = __ldg(nq + 2);
                                    yn = normalize(mk(n2.x + u * (n0.x - n2.x) + v * (n1.x - n2.x),
                                                      n2.y + u * (n0.y - n2.y) + v * (n1.y - n2.y),
                                                      n2.z + u * (n0.z - n2.z) + v * (n1.z - n2.z)));
                                } else {        // geometric normal (hw7 distributions.h:145)
                                    yn = normalize(cross(E2, E1));
                                }
                                const V3 td = t * d;
                                total += v0.w * len2(td) / fabsf(dot(d, yn));
                            }
                        }
                    }
                    pop = true;
                }
            } else if (at_inner) {
                float tl, tr;
                bool hl, hr;
                uint32_t ml, mr;
                pair_test(light ? sc.light_nodes : sc.nodes, cur, qr, light ? INFINITY : tbest, hl, hr, tl, tr, ml, mr);
                if (COUNT) { if (light) lnode_tests += 2; else node_tests += 2; }
                if (hl && hr) {
                    const bool left_first = light || tl <= tr;      // light sum: fixed left-to-right order (as wf_light)
                    stack[sp++] = left_first ? mr : ml;
                    cur = left_first ? ml : mr;
                } else if (hl) cur = ml;
                else if (hr) cur = mr;
                else pop = true;
            }
            if (pop) {
                if (sp > 0) cur = stack[--sp];
                else if (!light) {
                    // scene query done
                    const uint32_t bits = htri == B200RT_WF_MISS ? B200RT_WF_MISS : (htri | (hdet > 0.f ? 0x80000000u : 0u));
                    st_state(&wb.hit[p], make_float4(tbest, hu, hv, __uint_as_float(bits)));
                    if (do_light) {
                        // same ray against the light BVH; its root box is a real filter
                        light = true; total = 0.f;
                        if (COUNT) lnode_tests++;
                        if (light_boxes_hit(sc, o, d)) { qr = make_qray(sc.lq, o, d); cur = sc.light_root; }
                        else cur = B200RT_LEAF;      // empty leaf: ends on the next leaf phase
                    } else active = false;
                } else {
                    // end of the previous bounce: mean pdf, weight, firefly clamp (scene.cpp:158-163)
                    const float4 th = ld_state(&wb.thr[p]), bs = ld_state(&wb.bsdf[p]);
                    V3 T = mk(th.x, th.y, th.z);
                    if (finish_bounce(mk(bs.x, bs.y, bs.z), th.w, ld_state(&wb.ray_d[p]).w, total, sc.n_light_tris, T))
                        st_state(&wb.thr[p], make_float4(T.x, T.y, T.z, 0.f));
                    else {
                        wb.hit[p].w = __uint_as_float(B200RT_WF_DEAD);
                        dead++;
                    }
                    active = false;
                }
            }
        }
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        atomicAdd(&counters->rays, (unsigned long long)n);
        if (do_light) atomicAdd(&counters->light_queries, (unsigned long long)n);
    }
    if (do_light) {
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) dead += __shfl_down_sync(B200RT_FULL, dead, off);
        if (lane == 0 && dead) atomicAdd(&counters->rays, 0ull - (unsigned long long)dead);
    }
    if (COUNT) {
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) {
            node_tests += __shfl_down_sync(B200RT_FULL, node_tests, off);
            tri_tests += __shfl_down_sync(B200RT_FULL, tri_tests, off);
            lnode_tests += __shfl_down_sync(B200RT_FULL, lnode_tests, off);
            ltri_tests += __shfl_down_sync(B200RT_FULL, ltri_tests, off);
        }
        if (lane == 0) {
            atomicAdd(&counters->node_tests, (unsigned long long)node_tests);
            atomicAdd(&counters->tri_tests, (unsigned long long)tri_tests);
            atomicAdd(&counters->light_node_tests, (unsigned long long)lnode_tests);
            atomicAdd(&counters->light_tri_tests, (unsigned long long)ltri_tests);
        }
    }
}

}  // namespace b200rt
