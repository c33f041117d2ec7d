/* TEST INFRASTRUCTURE ONLY (oracle/): CPU walk of the FLATTENED scene BVH exactly as it sits in device memory
 * (b200rt_scene_bvh_download: 32-byte sibling-pair records + 48-byte triangle records), for SURVEY 8(d)'s cross-check
 * of the device counters ("N_* counted by the kernel and cross-checked against a CPU traversal of the same flattened
 * BVH on a fixed ray set").  It restates the closest-hit traversal that replaces BVH::intersect_
 * (/root/reference/hw8/src/include/bvh.h:111-142) in the form the device runs it: near child first, distance pruning,
 * explicit stack (raytracing_course_hw_b200/csrc/cuda/gltf_kernel.cuh: make_qray, slab_q, pair_test, tri_hit, and
 * wavefront.cuh: wf_extend).  Box and triangle arithmetic are bit-identical to the device's: every operation there is
 * explicitly rounded or explicitly fused (__fmul_rn / fmaf ...), and restated here with the same operations.
 * Build: gcc -O2 -ffp-contract=off -shared -fPIC (raytracing_course_hw_b200/build.py: build_oracle). */
#include <math.h>
#include <stdint.h>
#include <string.h>

#define LEAF 0x80000000u
#define TMAX 1e4f
#define MISS 0x7FFFFFFFu

typedef struct { uint32_t w[8]; } Pair;           /* a.x a.y a.z a.w b.x b.y b.z b.w */
typedef struct { float v0[4], e1[4], e2[4]; } Tri;
typedef struct {
    uint64_t n_records, n_tris;
    uint32_t root;
    int32_t quant_bits;
    float guard;
    float base[3], scale[3], inv_scale[3];
} Layout;                                        /* = b200rt_bvh_layout */

static float as_float(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }
static uint32_t as_uint(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }

/* u16 half of a record word spliced under the exponent of 0.5 (the device does it with one PRMT): 0.5 + q * 2^-16 */
static float q15(uint32_t w, int hi) { return as_float(0x3F000000u | (((hi ? w >> 16 : w) & 0xFFFFu) << 8)); }

typedef struct { float inv[3], bn[3], bf[3]; int hi[3]; } QRay;

static float safe_rcp(float d) {
    const float r = 1.f / d;
    return fabsf(r) < 1e30f ? r : copysignf(1e30f, d);
}
static void make_qray(const Layout *L, const float *o, const float *d, QRay *r) {
    for (int a = 0; a < 3; a++) {
        const float oq = (o[a] - L->base[a]) * L->inv_scale[a];
        const float ig = L->scale[a] * safe_rcp(d[a]);
        const float c = -(oq * ig);
        const float g = fmaf(fabsf(oq), 4.76837158203125e-7f, L->guard) * fabsf(ig);      /* see make_qray on the device */
        const float off = 32768.f * ig;
        r->bn[a] = (c - g) - off;
        r->bf[a] = (c + g) - off;
        r->inv[a] = 65536.f * ig;
        r->hi[a] = d[a] < 0.f;
    }
}
static int slab(uint32_t wx, uint32_t wy, uint32_t wz, const QRay *r, float tbest, float *tnear) {
    const float tx0 = fmaf(q15(wx, r->hi[0]), r->inv[0], r->bn[0]), tx1 = fmaf(q15(wx, !r->hi[0]), r->inv[0], r->bf[0]);
    const float ty0 = fmaf(q15(wy, r->hi[1]), r->inv[1], r->bn[1]), ty1 = fmaf(q15(wy, !r->hi[1]), r->inv[1], r->bf[1]);
    const float tz0 = fmaf(q15(wz, r->hi[2]), r->inv[2], r->bn[2]), tz1 = fmaf(q15(wz, !r->hi[2]), r->inv[2], r->bf[2]);
    const float tmin = fmaxf(fmaxf(tx0, ty0), tz0);
    const float tmax = fminf(fminf(tx1, ty1), tz1);
    *tnear = tmin;
    return tmin <= tmax && tmax >= 0.f && tmin <= tbest;
}
/* the device's cross_p / dot_p: products rounded, then one fused multiply-add per remaining term */
static void cross3(const float *a, const float *b, float *c) {
    c[0] = fmaf(a[1], b[2], -(a[2] * b[1])); c[1] = fmaf(a[2], b[0], -(a[0] * b[2])); c[2] = fmaf(a[0], b[1], -(a[1] * b[0]));
}
static float dot3(const float *a, const float *b) { return fmaf(a[2], b[2], fmaf(a[1], b[1], a[0] * b[0])); }
static float flip(float x, uint32_t sb) { return as_float(as_uint(x) ^ sb); }

static int tri_hit(const Tri *T, const float *o, const float *d, float tmax, float *t, float *u, float *v, float *det) {
    float pv[3], tv[3], qv[3];
    cross3(d, T->e2, pv);
    *det = dot3(T->e1, pv);
    for (int a = 0; a < 3; a++) tv[a] = o[a] - T->v0[a];
    cross3(tv, T->e1, qv);
    const uint32_t sb = as_uint(*det) & 0x80000000u;
    const float ad = fabsf(*det);
    const float un = flip(dot3(tv, pv), sb), vn = flip(dot3(d, qv), sb), tn = flip(dot3(T->e2, qv), sb);
    if (!(un >= 0.f && vn >= 0.f && un + vn <= ad && tn > 0.f && tn < tmax * ad)) return 0;
    const float idet = 1.f / ad;
    *t = tn * idet; *u = un * idet; *v = vn * idet;
    return 1;
}

/* rays: n x 6 floats; hits: n x 4 floats {t, u, v, bits}; per_ray (optional): n x 2 uint32 {boxes, triangles tested}.
 * counts[0] += child boxes tested (two per record fetched, like the device counter), counts[1] += triangles tested. */
int pairs_traverse(const Layout *L, const Pair *rec, const Tri *tris, const float *rays, int n, float *hits,
                   uint32_t *per_ray, uint64_t *counts) {
    if (!L || !rec || !tris || !rays || !hits || !counts || L->quant_bits != 15) return -1;
    uint64_t boxes = 0, tests = 0;
    for (int i = 0; i < n; i++) {
        const float *o = rays + 6 * (size_t)i, *d = o + 3;
        QRay qr;
        make_qray(L, o, d, &qr);
        uint32_t stack[64];
        int sp = 0;
        float tbest = TMAX, hu = 0.f, hv = 0.f, hdet = 0.f;
        uint32_t htri = MISS, cur = L->root, nb = 0, nt = 0;
        for (;;) {
            if (!(cur & LEAF)) {
                if (cur >= L->n_records) return -2;
                const Pair *p = rec + cur;
                float tl, tr;
                const int hl = slab(p->w[0], p->w[1], p->w[2], &qr, tbest, &tl);
                const int hr = slab(p->w[3], p->w[4], p->w[5], &qr, tbest, &tr);
                const uint32_t ml = p->w[6], mr = p->w[7];
                nb += 2;
                if (hl && hr) {
                    const int left_first = tl <= tr;
                    if (sp >= 64) return -3;
                    stack[sp++] = left_first ? mr : ml;
                    cur = left_first ? ml : mr;
                    continue;
                }
                if (hl) { cur = ml; continue; }
                if (hr) { cur = mr; continue; }
            } else {
                const uint32_t first = cur & 0x07FFFFFFu, count = (cur >> 27) & 15u;
                if ((uint64_t)first + count > L->n_tris) return -4;
                for (uint32_t k = first; k < first + count; k++) {
                    float t, u, v, det;
                    nt++;
                    if (tri_hit(tris + k, o, d, tbest, &t, &u, &v, &det)) { tbest = t; hu = u; hv = v; htri = k; hdet = det; }
                }
            }
            if (sp == 0) break;
            cur = stack[--sp];
        }
        float *h = hits + 4 * (size_t)i;
        h[0] = tbest; h[1] = hu; h[2] = hv;
        h[3] = as_float(htri == MISS ? MISS : (htri | (hdet > 0.f ? 0x80000000u : 0u)));
        if (per_ray) { per_ray[2 * (size_t)i] = nb; per_ray[2 * (size_t)i + 1] = nt; }
        boxes += nb; tests += nt;
    }
    counts[0] += boxes; counts[1] += tests;
    return 0;
}
