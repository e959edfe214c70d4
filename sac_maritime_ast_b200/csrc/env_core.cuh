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
    R arrive2[2];                                 // largest x with sqrt(x) <= 200 (:754,829), per precision
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

}  // namespace sit
