// prior.cpp - host half of the planar-prior construction: 2-D Delaunay
// triangulation of the support points.
//
// The reference delegates this to cv::Subdiv2D (ACMMP::DelaunayTriangulation,
// ACMMP.cpp:896-918: incremental insertion + getTriangleList, vertices truncated
// to int).  OpenCV C++ is not available here and a quad-edge incremental
// structure is a poor fit anyway; this is a divide-and-conquer triangulation
// (Guibas & Stolfi 1985) over the integer pixel coordinates with exact 64-bit
// orientation / in-circle predicates: O(n log n), no floating point, no
// degeneracy failures (support points sit on a 5-pixel quasi-grid, so co-circular
// quadruples are the common case, not the exception).  The Delaunay triangulation
// is unique up to the diagonal chosen inside co-circular quads; triangle ORDER
// (which only decides who wins a raster pixel shared by two triangles,
// acmmp_definitions.cpp:346-352) is by smallest edge index here.
//
// The top levels of the recursion run on separate threads, each half in its own
// edge arena; appending the right arena to the left one reproduces exactly the edge
// numbering of the sequential recursion (left subtree, right subtree, merge), so the
// triangles and their order do not depend on the number of threads.
#include "prior.h"

#include <algorithm>
#include <atomic>
#include <cstdint>
#include <future>
#include <numeric>
#include <thread>
#include <vector>

#include <stdlib.h>

namespace acmmp {
namespace {

struct Pt {
    int x, y, id;
};

// Quad-edge structure: edge e = 4 * q + r, r in 0..3.
struct QuadEdges {
    std::vector<int> onext;   // per directed/dual edge
    std::vector<int> org;     // per edge (only r = 0, 2 meaningful)
    std::vector<char> dead;   // per quad
    static int rot(int e) { return (e & ~3) | ((e + 1) & 3); }
    static int sym(int e) { return (e & ~3) | ((e + 2) & 3); }
    static int irot(int e) { return (e & ~3) | ((e + 3) & 3); }
    int on(int e) const { return onext[e]; }
    int oprev(int e) const { return rot(onext[rot(e)]); }
    int lnext(int e) const { return rot(onext[irot(e)]); }
    int rprev(int e) const { return onext[sym(e)]; }
    int dest(int e) const { return org[sym(e)]; }
    int make_edge(int a, int b) {
        const int e = (int)onext.size();
        onext.resize(e + 4);
        org.resize(e + 4, -1);
        dead.push_back(0);
        onext[e] = e;
        onext[e + 1] = e + 3;
        onext[e + 2] = e + 2;
        onext[e + 3] = e + 1;
        org[e] = a;
        org[e + 2] = b;
        return e;
    }
    void splice(int a, int b) {
        const int alpha = rot(onext[a]), beta = rot(onext[b]);
        std::swap(onext[a], onext[b]);
        std::swap(onext[alpha], onext[beta]);
    }
    int connect(int a, int b) {
        const int e = make_edge(dest(a), org[b]);
        splice(e, lnext(a));
        splice(sym(e), b);
        return e;
    }
    void remove(int e) {
        splice(e, oprev(e));
        splice(sym(e), oprev(sym(e)));
        dead[e >> 2] = 1;
    }
};

struct Triangulator {
    const std::vector<Pt> &p;
    QuadEdges q;
    // npts: points this arena will hold edges for (~3 edges per point + removed ones, 4 slots per edge)
    Triangulator(const std::vector<Pt> &pts, size_t npts) : p(pts) {
        q.onext.reserve(npts * 16);
        q.org.reserve(npts * 16);
        q.dead.reserve(npts * 4);
    }
    // > 0 when a, b, c make a left turn (in a y-up frame)
    int64_t orient(int a, int b, int c) const {
        return (int64_t)(p[b].x - p[a].x) * (p[c].y - p[a].y) - (int64_t)(p[b].y - p[a].y) * (p[c].x - p[a].x);
    }
    bool ccw(int a, int b, int c) const { return orient(a, b, c) > 0; }
    // d strictly inside the circle through a, b, c (a, b, c counter-clockwise)
    bool in_circle(int a, int b, int c, int d) const {
        const int64_t ax = p[a].x - p[d].x, ay = p[a].y - p[d].y;
        const int64_t bx = p[b].x - p[d].x, by = p[b].y - p[d].y;
        const int64_t cx = p[c].x - p[d].x, cy = p[c].y - p[d].y;
        const int64_t a2 = ax * ax + ay * ay, b2 = bx * bx + by * by, c2 = cx * cx + cy * cy;
        const int64_t det = ax * (by * c2 - b2 * cy) - ay * (bx * c2 - b2 * cx) + a2 * (bx * cy - by * cx);
        return det > 0;
    }
    bool right_of(int x, int e) const { return ccw(x, q.dest(e), q.org[e]); }
    bool left_of(int x, int e) const { return ccw(x, q.org[e], q.dest(e)); }
    bool valid(int e, int basel) const { return right_of(q.dest(e), basel); }

    // other's edges appended behind this arena's; returns the offset added to other's edge ids
    int append(const QuadEdges &o) {
        const int off = (int)q.onext.size();
        q.onext.resize(off + o.onext.size());
        for (size_t i = 0; i < o.onext.size(); ++i) q.onext[off + i] = o.onext[i] + off;
        q.org.insert(q.org.end(), o.org.begin(), o.org.end());
        q.dead.insert(q.dead.end(), o.dead.begin(), o.dead.end());
        return off;
    }

    // build(lo, hi) with the two halves of the top `depth` levels triangulated concurrently
    void build_parallel(int lo, int hi, int depth, int &le, int &re) {
        const int n = hi - lo;
        if (depth <= 0 || n < 4096 || !q.onext.empty()) {
            build(lo, hi, le, re);
            return;
        }
        const int mid = lo + n / 2;
        Triangulator R(p, (size_t)(hi - mid));
        int ldo, ldi, rdi, rdo;
        auto right = std::async(std::launch::async, [&] { R.build_parallel(mid, hi, depth - 1, rdi, rdo); });
        build_parallel(lo, mid, depth - 1, ldo, ldi);
        right.get();
        const int off = append(R.q);
        rdi += off;
        rdo += off;
        merge(ldo, ldi, rdi, rdo, le, re);
    }

    void build(int lo, int hi, int &le, int &re) {
        const int n = hi - lo;
        if (n == 2) {
            const int a = q.make_edge(lo, lo + 1);
            le = a;
            re = QuadEdges::sym(a);
            return;
        }
        if (n == 3) {
            const int a = q.make_edge(lo, lo + 1), b = q.make_edge(lo + 1, lo + 2);
            q.splice(QuadEdges::sym(a), b);
            if (ccw(lo, lo + 1, lo + 2)) {
                q.connect(b, a);
                le = a;
                re = QuadEdges::sym(b);
            } else if (ccw(lo, lo + 2, lo + 1)) {
                const int c = q.connect(b, a);
                le = QuadEdges::sym(c);
                re = c;
            } else {
                le = a;
                re = QuadEdges::sym(b);
            }
            return;
        }
        const int mid = lo + n / 2;
        int ldo, ldi, rdi, rdo;
        build(lo, mid, ldo, ldi);
        build(mid, hi, rdi, rdo);
        merge(ldo, ldi, rdi, rdo, le, re);
    }

    // joins the triangulations of two x-separated halves (Guibas & Stolfi, fig. 23)
    void merge(int ldo, int ldi, int rdi, int rdo, int &le, int &re) {
        for (;;) {   // lower common tangent
            if (left_of(q.org[rdi], ldi)) ldi = q.lnext(ldi);
            else if (right_of(q.org[ldi], rdi)) rdi = q.rprev(rdi);
            else break;
        }
        int basel = q.connect(QuadEdges::sym(rdi), ldi);
        if (q.org[ldi] == q.org[ldo]) ldo = QuadEdges::sym(basel);
        if (q.org[rdi] == q.org[rdo]) rdo = basel;
        for (;;) {   // merge upwards
            int lcand = q.on(QuadEdges::sym(basel));
            if (valid(lcand, basel)) {
                while (in_circle(q.dest(basel), q.org[basel], q.dest(lcand), q.dest(q.on(lcand)))) {
                    const int t = q.on(lcand);
                    q.remove(lcand);
                    lcand = t;
                }
            }
            int rcand = q.oprev(basel);
            if (valid(rcand, basel)) {
                while (in_circle(q.dest(basel), q.org[basel], q.dest(rcand), q.dest(q.oprev(rcand)))) {
                    const int t = q.oprev(rcand);
                    q.remove(rcand);
                    rcand = t;
                }
            }
            const bool lv = valid(lcand, basel), rv = valid(rcand, basel);
            if (!lv && !rv) break;
            if (!lv || (rv && in_circle(q.dest(lcand), q.org[lcand], q.org[rcand], q.dest(rcand))))
                basel = q.connect(rcand, QuadEdges::sym(basel));
            else
                basel = q.connect(QuadEdges::sym(basel), QuadEdges::sym(lcand));
        }
        le = ldo;
        re = rdo;
    }
};

std::atomic<int> g_threads{0};   // 0 = from the host's core count

// 2^depth threads for the top of the recursion
int delaunay_parallel_depth() {
    int t = g_threads.load();
    if (t <= 0) {
        // the scan runs up to three triangulations at once (software pipeline, capi_scan.cu.inc)
        static const int hw = [] {
            if (const char *e = getenv("ACMMP_DELAUNAY_THREADS")) return atoi(e);
            return (int)(std::thread::hardware_concurrency() / 3);
        }();
        t = hw;
    }
    int d = 0;
    while (d < 3 && (2 << d) <= t) ++d;
    return d;
}

}  // namespace

void delaunay_set_threads(int threads) { g_threads.store(threads < 0 ? 0 : threads); }

int delaunay_triangulate(const int32_t *xy, int n, std::vector<int32_t> &tris) {
    tris.clear();
    if (n < 3) return 0;
    std::vector<Pt> pts(n);
    for (int i = 0; i < n; ++i) pts[i] = Pt{xy[2 * i], xy[2 * i + 1], i};
    std::sort(pts.begin(), pts.end(), [](const Pt &a, const Pt &b) { return a.x != b.x ? a.x < b.x : a.y < b.y; });
    pts.erase(std::unique(pts.begin(), pts.end(), [](const Pt &a, const Pt &b) { return a.x == b.x && a.y == b.y; }),
              pts.end());
    const int m = (int)pts.size();
    if (m < 3) return 0;
    Triangulator T(pts, pts.size());
    int le, re;
    T.build_parallel(0, m, delaunay_parallel_depth(), le, re);
    const QuadEdges &q = T.q;
    const int ne = (int)q.onext.size();
    tris.reserve((size_t)m * 6);
    for (int e = 0; e < ne; e += 2) {   // primal directed edges: r = 0 and r = 2
        if (q.dead[e >> 2]) continue;
        const int e1 = q.lnext(e), e2 = q.lnext(e1);
        if (q.lnext(e2) != e) continue;           // left face is not a triangle (outer face)
        if (e > e1 || e > e2) continue;           // emit each face once, from its smallest edge
        const int a = q.org[e], b = q.org[e1], c = q.org[e2];
        if (T.orient(a, b, c) <= 0) continue;     // the outer face of a 3-point hull
        tris.push_back(pts[a].x); tris.push_back(pts[a].y);
        tris.push_back(pts[b].x); tris.push_back(pts[b].y);
        tris.push_back(pts[c].x); tris.push_back(pts[c].y);
    }
    return (int)(tris.size() / 6);
}

}  // namespace acmmp
