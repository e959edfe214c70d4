// Device-side polygon map queries and two-ship env constants (RLEnv/MSRL_env_ex.py,
// simulators/ship_in_transit/obstacle.py).  The map lives in the kernel-argument constant bank: every
// lane of a warp walks the same edge at the same time, so edge data are uniform operands.
//
// shapely / GEOS is an un-vendored, un-pinned dependency of the reference; these follow the
// published GEOS algorithms (RayCrossingCounter, Distance::pointToSegment) as restated in
// oracle/geometry.py -- PARITY UNPINNED for this sub-path.
#pragma once
#include "ship_core.cuh"

namespace sit {

template <typename R>
struct MapP {
    R x1[SIT_MAX_EDGES], y1[SIT_MAX_EDGES], x2[SIT_MAX_EDGES], y2[SIT_MAX_EDGES];
    R dx[SIT_MAX_EDGES], dy[SIT_MAX_EDGES];
    R len2[SIT_MAX_EDGES], inv_len2[SIT_MAX_EDGES];
    uint8_t poly_end[SIT_MAX_EDGES];  // 1 on the last segment of each polygon
    int32_t n_edges, n_poly;
};

template <typename R>
struct EnvP {
    // PolygonObstacle.map_boundaries obstacle.py:111-124 and the horizons built from them
    R min_n, max_n, min_e, max_e;                 // is_route_outside_horizon MSRL_env_ex.py:517-542
    R lo_n[2], hi_n[2], lo_e[2], hi_e[2];         // is_pos_outside_horizon :460-488, per role (L/2 margin)
    R inv_max_north;                              // reward normalisation :650,697,725
    R inv_e_tol, e_tol;                           // e_tolerance = 1000 :119
    double arrive2;                               // largest double x with sqrt(x) <= 200 (:754,829): the squared distance is compared in double
    R rpm_max;                                    // 2000 :557
    double collide2;                              // 50^2 :592
    R near2;                                      // (L/2 * sqrt2 + slack)^2: beyond it all hull corners share the centre's side
    double theta, sampling_frequency;             // :125,130
    R half_len[2];
};

// env flag bits (state field "flags")
enum { F_STOP_TEST = 1, F_STOP_OBS = 2, F_NSTORED_SHIFT = 2, F_NSTORED_MASK = 3 << 2, F_IW_BAD = 16, F_FRESH = 32,
       F_IW_VALID = 64, F_PENDING = 128, F_HAVE_ACT = 256 };

// One ring segment as the kernels read it: staged once per CTA from the constant bank into shared memory so the
// edge loop issues two 128-bit broadcast LDS per edge (float) instead of eight indexed constant loads.
template <typename R> struct IntOf;
template <> struct IntOf<float> { typedef int32_t type; };
template <> struct IntOf<double> { typedef int64_t type; };

template <typename R>
struct __align__(16) EdgeRec {
    R x1, y1, dx, dy;
    R len2, inv_len2, y2;
    typename IntOf<R>::type pbit;   // 1 << (polygon id & 31) ... per-polygon crossing parity lives in one 32-bit mask
    R x2, _p0, _p1, _p2;            // third 128-bit word: only the containment test reads it
};

template <typename R>
__device__ __forceinline__ void stage_edges(const MapP<R>& m, EdgeRec<R>* sm) {
    int poly = 0;
    // polygon ids are a running count of poly_end flags: recomputed per thread from the (tiny) uniform array
    for (int i = threadIdx.x; i < m.n_edges; i += blockDim.x) {
        poly = 0;
        for (int j = 0; j < i; ++j) poly += m.poly_end[j];
        EdgeRec<R> e;
        e.x1 = m.x1[i]; e.y1 = m.y1[i]; e.dx = m.dx[i]; e.dy = m.dy[i];
        e.len2 = m.len2[i]; e.inv_len2 = m.inv_len2[i]; e.y2 = m.y2[i];
        e.x2 = m.x2[i]; e._p0 = e._p1 = e._p2 = (R)0;
        e.pbit = (typename IntOf<R>::type)(1u << (poly & 31));
        sm[i] = e;
    }
    __syncthreads();
}

// Point-in-any-polygon for NP points at once: obstacle.py:126-129 any(poly.contains(Point(e, n))), interior only
// (a point on a ring segment or vertex is boundary, hence not contained -- GEOS RayCrossingCounter, restated in
// oracle/geometry.py ring_locate).  Branch-free form of the per-edge if / else-if chain: the three boundary cases are
// mutually exclusive with a counted crossing, so they are OR-ed into a per-polygon boundary mask that vetoes the
// polygon's parity bit at the end.
template <typename R, int NP>
__device__ __forceinline__ void map_contains_n(const EdgeRec<R>* __restrict__ E, const int n_edges, const R (&py)[NP],
                                               const R (&px)[NP], bool (&inside)[NP]) {
    uint32_t parity[NP], bmask[NP];
#pragma unroll
    for (int q = 0; q < NP; ++q) { parity[q] = 0u; bmask[q] = 0u; }
#pragma unroll 2
    for (int i = 0; i < n_edges; ++i) {
        const EdgeRec<R> e = E[i];
        const uint32_t pbit = (uint32_t)e.pbit;
        const bool down = e.y2 < e.y1, horiz = e.y1 == e.y2;
        const R lo = e.x1 <= e.x2 ? e.x1 : e.x2, hi = e.x1 <= e.x2 ? e.x2 : e.x1;
#pragma unroll
        for (int q = 0; q < NP; ++q) {
            const bool right = !(e.x1 < px[q] && e.x2 < px[q]);     // segments strictly left of the point never count
            const bool straddle = (e.y1 > py[q] && e.y2 <= py[q]) || (e.y2 > py[q] && e.y1 <= py[q]);
            const R d = e.dx * (py[q] - e.y1) - e.dy * (px[q] - e.x1);
            const bool on = (px[q] == e.x2 && py[q] == e.y2) || (horiz && e.y1 == py[q] && lo <= px[q] && px[q] <= hi) ||
                            (straddle && d == (R)0);
            bmask[q] |= (right && on) ? pbit : 0u;
            parity[q] ^= (right && straddle && ((d > (R)0) != down)) ? pbit : 0u;
        }
    }
#pragma unroll
    for (int q = 0; q < NP; ++q) inside[q] = (parity[q] & ~bmask[q]) != 0u;
}

template <typename R>
__device__ __forceinline__ bool map_contains(const EdgeRec<R>* __restrict__ E, const int n_edges, R py /*north*/, R px /*east*/) {
    const R ay[1] = {py}, ax[1] = {px};
    bool in[1];
    map_contains_n<R, 1>(E, n_edges, ay, ax, in);
    return in[0];
}

// Same test over a subset of the ring segments (indices in `list`).  Exact whenever every skipped segment can contribute
// neither a crossing nor a boundary hit for the point: its y-range does not reach py, or it lies strictly left of px
// (see arc_edge_list in shipsim.cu, which builds such a subset once for all candidate waypoints of one decision).
template <typename R>
__device__ __forceinline__ bool map_contains_list(const EdgeRec<R>* __restrict__ E, const uint8_t* __restrict__ list,
                                                  const int n_list, const R py, const R px) {
    uint32_t parity = 0u, bmask = 0u;
#pragma unroll 2
    for (int j = 0; j < n_list; ++j) {
        const EdgeRec<R> e = E[list[j]];
        const uint32_t pbit = (uint32_t)e.pbit;
        const bool down = e.y2 < e.y1, horiz = e.y1 == e.y2;
        const R lo = e.x1 <= e.x2 ? e.x1 : e.x2, hi = e.x1 <= e.x2 ? e.x2 : e.x1;
        const bool right = !(e.x1 < px && e.x2 < px);
        const bool straddle = (e.y1 > py && e.y2 <= py) || (e.y2 > py && e.y1 <= py);
        const R d = e.dx * (py - e.y1) - e.dy * (px - e.x1);
        const bool on = (px == e.x2 && py == e.y2) || (horiz && e.y1 == py && lo <= px && px <= hi) || (straddle && d == (R)0);
        bmask |= (right && on) ? pbit : 0u;
        parity ^= (right && straddle && ((d > (R)0) != down)) ? pbit : 0u;
    }
    return (parity & ~bmask) != 0u;
}

// One pass over the ring segments for the hull centre: squared distance to the nearest ring
// (obstacle.py:138-141 min(poly.exterior.distance)) and containment of the centre, sharing the
// per-edge cross product.  Valid as the containment of all four hull corners whenever
// dist2 > near2 (no ring can pass between the centre and a corner).  Polygons sharing a parity bit
// (more than 32 polygons) would alias: sit_create rejects such maps.
template <typename R>
__device__ __forceinline__ void map_scan(const EdgeRec<R>* __restrict__ E, const int n_edges, R py, R px, R& dist2,
                                         bool& centre_inside) {
    R best0 = (R)3.0e38, best1 = (R)3.0e38;
    uint32_t parity = 0;
#ifndef SIT_SCAN_UNROLL
#define SIT_SCAN_UNROLL 8
#endif
    constexpr int kUnroll = SIT_SCAN_UNROLL;
#pragma unroll kUnroll
    for (int i = 0; i < n_edges; ++i) {
        const EdgeRec<R> e = E[i];
        const R ex = px - e.x1, ey = py - e.y1;
        const R v2 = ex * ex + ey * ey;               // distance^2 to the segment's first vertex
        const R dot = ex * e.dx + ey * e.dy;
        const R cr = e.dx * ey - e.dy * ex;           // orientation of (a, b, p)
        const R pr = cr * cr * e.inv_len2;            // distance^2 to the supporting line
        const R cand = (dot > (R)0 && dot < e.len2) ? pr : v2;
        if (i & 1) best1 = min_(best1, cand); else best0 = min_(best0, cand);
        const bool straddle = (e.y1 > py) != (e.y2 > py);
        parity ^= (straddle && ((cr > (R)0) != (e.dy < (R)0))) ? (uint32_t)e.pbit : 0u;
    }
    dist2 = min_(best0, best1);
    centre_inside = parity != 0u;
}

// ---- nearest-ring cache ------------------------------------------------------------------------------------------
// A ship moves a few metres per step while the rings are hundreds of metres away, so the edge that realises
// obstacles_distance (obstacle.py:138-141) stays among the same few edges for tens of steps.  Each lane keeps the 4
// edges nearest to the position p0 of its last scan and a displacement budget b = min((d5 - d1) / 2, d1) - 1 m
// (d1, d5 = nearest / fifth-nearest segment distance at p0).  While |p - p0| < b the nearest segment at p is one of the
// cached four (every other segment is at least d5 - |p - p0| away, segment distance being 1-Lipschitz) and the ship
// cannot have crossed a ring (|p - p0| < d1), so the distance is the minimum over four segments and the containment
// of the centre is the cached one -- bit-identical to the full scan.  A lane whose budget is spent is rescanned by the
// WHOLE warp: lanes take 32 edges at a time, top-5 by five REDUX.MIN over (distance bits | edge index) keys, crossing
// parity by one REDUX.XOR.
struct MapCache {
    uint32_t idx;      // four edge indices, 8 bits each
    float p0n, p0e;    // position of the last scan
    float b2;          // squared displacement budget; negative: the centre was inside a polygon at p0; 0: invalid
};

// squared distance to segment i (GEOS Distance::pointToSegment cases) and the orientation product
template <typename R>
__device__ __forceinline__ R segment_dist2(const EdgeRec<R>& e, R py, R px, R& cr) {
    const R ex = px - e.x1, ey = py - e.y1;
    const R v2 = ex * ex + ey * ey;
    const R dot = ex * e.dx + ey * e.dy;
    cr = e.dx * ey - e.dy * ex;
    const R pr = cr * cr * e.inv_len2;
    const R fx = px - e.x2, fy = py - e.y2;
    const R w2 = fx * fx + fy * fy;                   // the next segment's first vertex: same bits as its v2
    return dot <= (R)0 ? v2 : (dot >= e.len2 ? w2 : pr);
}

__device__ __forceinline__ float f32_down(double x) { return __double2float_rd(x); }
__device__ __forceinline__ float f32_down(float x) { return x; }

// all 32 lanes: scan the map for the point (py, px) of one lane; every lane gets the same result
template <typename R>
__device__ __forceinline__ MapCache coop_scan(const EdgeRec<R>* __restrict__ E, const int n_edges, R py, R px, int lane) {
    const unsigned FULL = 0xffffffffu;
    uint32_t key[SIT_MAX_EDGES / 32];
    uint32_t parity = 0u;
#pragma unroll
    for (int m = 0; m < SIT_MAX_EDGES / 32; ++m) {
        const int j = lane + 32 * m;
        key[m] = 0xffffffffu;
        if (32 * m < n_edges) {             // warp-uniform
            if (j < n_edges) {
                const EdgeRec<R> e = E[j];
                R cr;
                const R d2 = segment_dist2(e, py, px, cr);
                key[m] = (__float_as_uint(f32_down(d2)) & 0xffffff80u) | (uint32_t)j;
                const bool straddle = (e.y1 > py) != (e.y2 > py);
                parity ^= (straddle && ((cr > (R)0) != (e.dy < (R)0))) ? (uint32_t)e.pbit : 0u;
            }
        }
    }
    uint32_t g[5];
#pragma unroll
    for (int r = 0; r < 5; ++r) {
        uint32_t loc = key[0];
#pragma unroll
        for (int m = 1; m < SIT_MAX_EDGES / 32; ++m) loc = min(loc, key[m]);
        g[r] = __reduce_min_sync(FULL, loc);
#pragma unroll
        for (int m = 0; m < SIT_MAX_EDGES / 32; ++m) key[m] = key[m] == g[r] ? 0xffffffffu : key[m];
    }
    parity = __reduce_xor_sync(FULL, parity);
    MapCache c;
    const uint32_t i0 = g[0] & 0x7fu;
    c.idx = i0 | ((g[1] == 0xffffffffu ? i0 : (g[1] & 0x7fu)) << 8) | ((g[2] == 0xffffffffu ? i0 : (g[2] & 0x7fu)) << 16) |
            ((g[3] == 0xffffffffu ? i0 : (g[3] & 0x7fu)) << 24);
    c.p0n = (float)py; c.p0e = (float)px;
    const float d1 = sqrtf(__uint_as_float(g[0] | 0x7fu));                                   // rounded up
    const float d5 = g[4] == 0xffffffffu ? 1.0e6f : sqrtf(__uint_as_float(g[4] & 0xffffff80u));   // rounded down
    const float b = fminf(0.5f * (d5 - d1), d1) - 1.0f;     // 1 m covers the float copies of p0 and of the keys
    const float b2 = b > 0.0f ? b * b : 0.0f;
    c.b2 = copysignf(b2, parity != 0u ? -1.0f : 1.0f);   // the sign bit survives b2 == 0
    if (n_edges == 0 || g[0] == 0xffffffffu) { c.idx = 0u; c.b2 = 0.0f; }
    return c;
}

// per lane: obstacles_distance^2 from the four cached segments
template <typename R>
__device__ __forceinline__ R cached_dist2(const EdgeRec<R>* __restrict__ E, uint32_t idx, R py, R px) {
    R best = (R)3.0e38;
#pragma unroll
    for (int r = 0; r < 4; ++r) {
        R cr;
        best = min_(best, segment_dist2(E[(idx >> (8 * r)) & 0xffu], py, px, cr));
    }
    return best;
}

// all 32 lanes: is any of the four hull corners of one lane's ship inside a polygon (MSRL_env_ex.py:490-515)
template <typename R>
__device__ __forceinline__ bool coop_hull_inside(const EdgeRec<R>* __restrict__ E, const int n_edges, R n, R e_, R half, int lane) {
    const unsigned FULL = 0xffffffffu;
    const R py[4] = {n - half, n - half, n + half, n + half}, px[4] = {e_ - half, e_ + half, e_ - half, e_ + half};
    uint32_t parity[4] = {0u, 0u, 0u, 0u}, bmask[4] = {0u, 0u, 0u, 0u};
    for (int j = lane; j < n_edges; j += 32) {
        const EdgeRec<R> e = E[j];
        const uint32_t pbit = (uint32_t)e.pbit;
        const bool down = e.y2 < e.y1, horiz = e.y1 == e.y2;
        const R lo = e.x1 <= e.x2 ? e.x1 : e.x2, hi = e.x1 <= e.x2 ? e.x2 : e.x1;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const bool right = !(e.x1 < px[q] && e.x2 < px[q]);
            const bool straddle = (e.y1 > py[q] && e.y2 <= py[q]) || (e.y2 > py[q] && e.y1 <= py[q]);
            const R d = e.dx * (py[q] - e.y1) - e.dy * (px[q] - e.x1);
            const bool on = (px[q] == e.x2 && py[q] == e.y2) || (horiz && e.y1 == py[q] && lo <= px[q] && px[q] <= hi) ||
                            (straddle && d == (R)0);
            bmask[q] |= (right && on) ? pbit : 0u;
            parity[q] ^= (right && straddle && ((d > (R)0) != down)) ? pbit : 0u;
        }
    }
    bool in = false;
#pragma unroll
    for (int q = 0; q < 4; ++q)
        in |= (__reduce_xor_sync(FULL, parity[q]) & ~__reduce_or_sync(FULL, bmask[q])) != 0u;
    return in;
}

}  // namespace sit
