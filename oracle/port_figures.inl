// TEST INFRASTRUCTURE ONLY.  Analytic primitive intersection of hw2..hw5, restated.
// Included inside a namespace by each port TU so that sqrt/fabs bind the way they do in
// the matching reference TU (<math.h> for hw2, <cmath> for hw4+).
//   Figure::intersect        hw2/src/primitives.cpp:25-34,  hw4/src/primitives.cpp:9-27
//   quadratic                hw2/src/primitives.cpp:36-57,  hw4/src/primitives.cpp:29-49
//   ellipsoid / plane / box  hw4/src/primitives.cpp:51-118 (plane gains T_MAX = 1e4 from hw4 on)

struct Hit {
    float t = 0;
    V norma;
    bool inside = false;
    bool valid = false;
};

static const float kTMax = 1e4;

static inline bool smallest_positive_root(float a, float b, float c, float &t, bool &inside) {
    float d = b * b - 4 * a * c;
    if (d <= 0) return false;
    float x1 = (-b - sqrt(d)) / (2 * a);
    float x2 = (-b + sqrt(d)) / (2 * a);
    if (x1 > x2) std::swap(x1, x2);
    if (x2 < 0) return false;
    if (x1 < 0) { t = x2; inside = true; }
    else { t = x1; inside = false; }
    return true;
}

static inline Hit hit_ellipsoid(const V &r, const Ray &ray) {
    float c = (ray.o / r).len2() - 1;
    float b = 2. * (ray.o / r).dot(ray.d / r);
    float a = (ray.d / r).len2();
    Hit h;
    if (!smallest_positive_root(a, b, c, h.t, h.inside)) return h;
    V point = ray.o + h.t * ray.d;
    V norma = point / (r * r);
    if (h.inside) norma = -1. * norma;
    h.norma = norma.normalize();
    h.valid = true;
    return h;
}

static inline Hit hit_plane(const V &n, const Ray &ray, bool use_tmax) {
    Hit h;
    float t = -ray.o.dot(n) / ray.d.dot(n);
    if (t > 0 && (!use_tmax || t < kTMax)) {
        h.valid = true; h.t = t;
        if (ray.d.dot(n) > 0) { h.norma = -1. * n; h.inside = true; }
        else { h.norma = n; h.inside = false; }
    }
    return h;
}

static inline Hit hit_box(const V &s, const Ray &ray) {
    Hit h;
    V ts1 = (-1. * s - ray.o) / ray.d;
    V ts2 = (s - ray.o) / ray.d;
    float t1x = std::min(ts1.x, ts2.x), t2x = std::max(ts1.x, ts2.x);
    float t1y = std::min(ts1.y, ts2.y), t2y = std::max(ts1.y, ts2.y);
    float t1z = std::min(ts1.z, ts2.z), t2z = std::max(ts1.z, ts2.z);
    float t1 = std::max(std::max(t1x, t1y), t1z);
    float t2 = std::min(std::min(t2x, t2y), t2z);
    if (t1 > t2 || t2 < 0) return h;
    if (t1 < 0) { h.inside = true; h.t = t2; }
    else { h.inside = false; h.t = t1; }
    V p = ray.o + h.t * ray.d;
    V norma = p / s;
    float mx = std::max(std::max(fabs(norma.x), fabs(norma.y)), fabs(norma.z));
    if (fabs(norma.x) != mx) norma.x = 0;
    if (fabs(norma.y) != mx) norma.y = 0;
    if (fabs(norma.z) != mx) norma.z = 0;
    if (h.inside) norma = -1. * norma;
    h.norma = norma;
    h.valid = true;
    return h;
}

// World-space query: world -> local by (ray - position).rotate(rotation); normal back by the
// conjugate, then normalised.
static inline Hit hit_figure(const b200rt_primitive &f, const Ray &ray, bool plane_tmax) {
    Q rot(f.rotation);
    Ray local = ray.shifted(V(f.position)).rotated(rot);
    Hit h;
    if (f.type == B200RT_PRIM_ELLIPSOID) h = hit_ellipsoid(V(f.data), local);
    else if (f.type == B200RT_PRIM_PLANE) h = hit_plane(V(f.data), local, plane_tmax);
    else h = hit_box(V(f.data), local);
    if (!h.valid) return h;
    h.norma = rot.conjugate().transform(h.norma).normalize();
    return h;
}
