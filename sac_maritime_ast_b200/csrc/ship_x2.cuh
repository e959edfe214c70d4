// Packed-FP32 form of the ship step: ONE THREAD OWNS TWO SHIPS and every FP32 operation of the step is issued once for
// both as a Blackwell f32x2 instruction (SASS FFMA2 / FMUL2 / FADD2, one issue slot for two ships).  The fp32 K1 kernel
// was bound by instruction issue (200 warp instructions per ship-step, issue slots 90 % busy, FMA pipe 60 %); packing
// halves the FP32 instruction count per ship, uniform parameters ride as uniform-register / immediate broadcast operands
// of the packed instructions, and the transcendental pieces are written so that their polynomials pack too:
//   * sin / cos of the heading: Cody-Waite reduction by pi (not pi/2: no quadrant swap), sign applied by one XOR per ship,
//     degree-9 / degree-10 minimax polynomials on [-pi/2, pi/2] (|error| < 1.3e-7 evaluated in float);
//   * atan of the LOS term: one MUFU.RCP per ship for |x| > 1, odd degree-15 minimax polynomial on [-1, 1] (< 1.1e-7);
//   * the machinery-mode formulas of ship_engine.py:32-76 are folded into one branch-free form (ModeX) so that ships in
//     different modes share the packed instructions.
//   * positions: the state in HBM stays float64 (north, east); inside the loop each ship carries its position RELATIVE TO THE
//     START OF ITS ROUTE SEGMENT as a compensated float pair (X - CX, Kahan summation: ~48 significant bits, as good as the
//     float64 accumulation of float increments it replaces) so that the per-step position update, the LOS cross-track error
//     and the waypoint pre-test are packed FP32 too -- no FP64 and no F2F conversion in the steady-state step.  The
//     waypoint switch itself is still DECIDED IN DOUBLE on the reconstructed float64 position with the reference's
//     operation order: a packed float pre-test with a 1 % band around the acceptance circle selects the (rare) steps on
//     which that exact test runs, for both ships of the thread in one branch.
// Same physics, same reference citations as ship_core.cuh::ship_step; the arithmetic differs from the scalar float step
// only in rounding (different but equivalent operation order).
//
// Reference citations are relative to /root/reference/.
#pragma once
#include "ship_core.cuh"

namespace sit {
namespace x2 {

typedef float2 V;

__device__ __forceinline__ V bc(float a) { return make_float2(a, a); }
__device__ __forceinline__ V neg(V a) { return make_float2(-a.x, -a.y); }
__device__ __forceinline__ V vabs(V a) { return make_float2(fabsf(a.x), fabsf(a.y)); }
__device__ __forceinline__ V fma(V a, V b, V c) { return __ffma2_rn(a, b, c); }
__device__ __forceinline__ V fma(V a, float b, V c) { return __ffma2_rn(a, bc(b), c); }
__device__ __forceinline__ V fma(V a, float b, float c) { return __ffma2_rn(a, bc(b), bc(c)); }
__device__ __forceinline__ V fma(V a, V b, float c) { return __ffma2_rn(a, b, bc(c)); }
__device__ __forceinline__ V mul(V a, V b) { return __fmul2_rn(a, b); }
__device__ __forceinline__ V mul(V a, float b) { return __fmul2_rn(a, bc(b)); }
__device__ __forceinline__ V add(V a, V b) { return __fadd2_rn(a, b); }
__device__ __forceinline__ V add(V a, float b) { return __fadd2_rn(a, bc(b)); }
__device__ __forceinline__ V sub(V a, V b) { return __fadd2_rn(a, neg(b)); }

// ---- uniform parameters of the packed step (floats; ptxas keeps them as uniform-register operands) ----------------
// One machinery mode in branch-free form.  ship_engine.py:46-76 distribute_load:
//   MOTOR:  load_me = min(total, me_cap);            load_el = (total + hotel) - load_me
//   other:  load_me = total + load_me_add;           load_el = load_el_const
// both are  load_me = min(total + add, cap),  load_el = a_el * ((total + hotel) - load_me) + c_el.
struct ModeX {
    float avail_total, avail_me, avail_el;   // update_available_propulsion_power :32-44
    float cap_tq_me, cap_tq_el;              // torque caps :374-375,383-384
    float add_me, cap_me;                    // load_me = min(total + add_me, cap_me)
    float a_el, c_el;                        // load_el = a_el * ((total + hotel) - load_me) + c_el
    float inv_me_cap, inv_el_cap;            // 0 where the reference guards a zero capacity
    float me_cap_kw, _pad;                   // main_engine_capacity / 1000: 'available power me [kw]' (ship_model.py:662-664)
    // specific fuel consumption as a polynomial in the LOAD instead of the load fraction (ship_engine.py:256-292 with
    // pct = load / capacity folded in, dt / 3.6e9 too): rate dt = load ((a load + b) load + c)
    float fme_a, fme_b, fel_a, fel_b;
};

struct ShipPX {
    double ra2;                                            // radius_of_acceptance^2, decided in double
    float ra2_band;                                        // float pre-test: d2 <= 1.01 ra^2 + 16 -> run the exact test
    float c_inv_pi, c_s3, c_c4, c_a6;                      // second constants of two-constant FMAs (uniform operands)
    float dt, look2, sat_e, neg_ki_los, windup;            // LOS_guidance.py:105-121
    float kp_h, kd_idt, ki_h, rud_max, rud_bias;           // controllers.py:72-100 (kd / dt folded)
    float kp_kd;                                           // kp + kd / dt: the PID on (err, e_prev) instead of (err, err - e_prev)
    float kp_s, ki_s, kp_w, ki_w;                          // controllers.py:138-143
    float hotel, sfc_me[3], sfc_dg[3];                     // ship_engine.py:256-292 (dt / 3.6e9 folded)
    float sh_me, sh_hsg, sh_lin, sh_quad;                  // shaft_eq :355-361
    float au_vr, au_vrr, au_lin, au_nl, au_wind, au_thr;   // ship_model.py:576-606, accelerations
    float av_ur, av_urr, av_lin, av_nl, av_wind, av_rud;
    float ar_ca, ar_lin, ar_nl, ar_wind, ar_rud;
    ModeX mode[SIT_MAX_MODES];
    int32_t n_modes, _pad;
};

// ---- per-thread register state: two ships --------------------------------------------------------------------------
struct Ship2 {
    V X, Y, CX, CY;                         // position relative to the segment start; true value = X - CX (Kahan)
    V SX, SY;                               // segment vector wp[k] - wp[k-1]
    V thr2;                                 // pre-test threshold: ra2_band while a next waypoint exists, -1 at the end of the route
    V psi, u, v, r, om;                     // ship_model.py:161-164, ship_engine.py:327
    V ei_s, ei_w, ei_h, e_prev, e_ct, e_int;
    V f_me, f_el;                           // fuel_cons = its value at load + the increments of both (formed at store)
    V u_sp, vcn, vce, wcb, wsb;
    V alpha, sa, ca;
    int32_t k[2], n_wp[2];
};

// per-ship mode constants as lanes of packed registers (several modes in one fleet) or as uniform scalars (one mode)
struct ModeV { V avail_total, avail_me, avail_el, cap_tq_me, cap_tq_el, add_me, cap_me, a_el, c_el, inv_me_cap, inv_el_cap, me_cap_kw,
                   fme_a, fme_b, fel_a, fel_b; };

// the stored (pre-step) observables the env reads back (MSRL_env_ex.py:267-269,298-300): shaft speed [rad/s] before the
// step and the load on the main engine [W]
struct StepOut2 { V om_pre, load_me; };
enum { AVOID_NONE = 0, AVOID_BOTH = 1, AVOID_LANE0 = 2, AVOID_LANE0_RT = 3 };

__device__ __forceinline__ float& lane(V& a, int l) { return l == 0 ? a.x : a.y; }
__device__ __forceinline__ float lane(const V& a, int l) { return l == 0 ? a.x : a.y; }
__device__ __forceinline__ float lane(const float& a, int) { return a; }

// float64 position of ship l from its segment start and the compensated relative pair
__device__ __forceinline__ double abs_pos(double p, float x, float c) { return p + ((double)x - (double)c); }
// relative pair of a float64 offset: x = nearest float, c = -(remainder)
__device__ __forceinline__ void split_rel(double rel, float& x, float& c) { x = (float)rel; c = -(float)(rel - (double)x); }

// direction cosines and angle atan2(sy, sx) of a segment vector: one MUFU.RSQ + a Newton step, and a branch-free atan2 on
// the same odd polynomial as atan2v.  Runs once per waypoint switch, for one ship of the warp at a time.
struct SegGeo { float alpha, sa, ca; };
__device__ __forceinline__ SegGeo seg_geometry(const float sx, const float sy) {
    const float l2 = fmaf(sx, sx, sy * sy);
    float il = rsqrt_approx(l2);
    il = il * fmaf(-0.5f * l2 * il, il, 1.5f);
    const bool degenerate = !(l2 > 0.0f);                     // zero-length segment: atan2(0, 0) = 0
    SegGeo q;
    q.sa = degenerate ? 0.0f : sy * il; q.ca = degenerate ? 1.0f : sx * il;
    const float ax = fabsf(sx), ay = fabsf(sy);
    const float z = fminf(ax, ay) * rcp_(fmaxf(fmaxf(ax, ay), 1e-30f));
    const float z2 = z * z;
    float pa = fmaf(z2, -0.004355357028543949f, 0.023039964959025383f);
    pa = fmaf(pa, z2, -0.05777335166931152f); pa = fmaf(pa, z2, 0.09794218093156815f);
    pa = fmaf(pa, z2, -0.13976575434207916f); pa = fmaf(pa, z2, 0.19962702691555023f);
    pa = fmaf(pa, z2, -0.3333165943622589f);
    pa = fmaf(pa * z2, z, z);
    if (ay > ax) pa = 1.57079632679489662f - pa;
    if (sx < 0.0f) pa = 3.14159265358979324f - pa;
    q.alpha = copysignf(pa, sy);
    return q;
}
// the same as a real call: the env kernel reaches this from nine places, and a compact instruction footprint matters more there
// than the call overhead on a path taken once per waypoint switch
__device__ __noinline__ SegGeo seg_geometry_call(const float sx, const float sy) { return seg_geometry(sx, sy); }

// LOS_guidance.py:110-112: segment k-1 -> k of ship l: vector, angle (atan2(dE, dN)), threshold of the waypoint pre-test
template <bool COMPACT = false>
__device__ __forceinline__ void set_segment2(Ship2& s, const ShipPX& P, const double pN, const double pE, const double cN,
                                             const double cE, const int l) {
    const float sx = (float)(cN - pN), sy = (float)(cE - pE);
    const SegGeo q = COMPACT ? seg_geometry_call(sx, sy) : seg_geometry(sx, sy);
    lane(s.SX, l) = sx; lane(s.SY, l) = sy;
    lane(s.alpha, l) = q.alpha; lane(s.sa, l) = q.sa; lane(s.ca, l) = q.ca;
    lane(s.thr2, l) = (l == 0 ? s.n_wp[0] > s.k[0] + 1 : s.n_wp[1] > s.k[1] + 1) ? P.ra2_band : -1.0f;
}

// LOS_guidance.py:88-103 next_wpt for ship l, decided in double on the float64 position with the reference's operation order
// (squared distance, <=, end-of-route hold).  On a switch the relative pair is re-based on the new segment start.
template <bool COMPACT = false>
__device__ __forceinline__ void next_wpt2(Ship2& s, const ShipPX& P, const ShipPtrs<float>& g, const int64_t i, const int l) {
    const int64_t S = g.n_ships;
    const int k = l == 0 ? s.k[0] : s.k[1], nw = l == 0 ? s.n_wp[0] : s.n_wp[1];
    const double pN = g.wp_n[(int64_t)(k - 1) * S + i], pE = g.wp_e[(int64_t)(k - 1) * S + i];
    const double cN = g.wp_n[(int64_t)k * S + i], cE = g.wp_e[(int64_t)k * S + i];
    const double N = abs_pos(pN, lane(s.X, l), lane(s.CX, l)), E = abs_pos(pE, lane(s.Y, l), lane(s.CY, l));
    const double dn = cN - N, de = cE - E;
    const double d2 = __dadd_rn(__dmul_rn(dn, dn), __dmul_rn(de, de));
    if (d2 <= P.ra2 && nw > k + 1) {
        if (l == 0) s.k[0] = k + 1; else s.k[1] = k + 1;
        split_rel(N - cN, lane(s.X, l), lane(s.CX, l));
        split_rel(E - cE, lane(s.Y, l), lane(s.CY, l));
        set_segment2<COMPACT>(s, P, cN, cE, g.wp_n[(int64_t)(k + 1) * S + i], g.wp_e[(int64_t)(k + 1) * S + i], l);
    }
}

// sin and cos of two headings.  t = x / pi + 1.5 * 2^23 holds round(x / pi) in its low mantissa bits; r = x - j pi in two
// Cody-Waite steps (pi = 3.140625 + 9.67653589793e-4: the product j * 3.140625 is exact for |j| < 2^13, i.e. |x| < 2.5e4 rad);
// sin x = (-1)^j sin r, cos x = (-1)^j cos r: the parity bit of j moves to the sign position and is XOR-ed in.
__device__ __forceinline__ void sincos2(const ShipPX& P, const V x, V& sn, V& cs) {
    const V t = fma(x, P.c_inv_pi, 12582912.0f);
    const V j = add(t, -12582912.0f);
    V r = fma(j, -3.140625f, x);
    r = fma(j, -9.67653589793e-4f, r);
    const V r2 = mul(r, r);
    V ps = fma(r2, P.c_s3, -0.00019806601630989462f);
    ps = fma(ps, r2, 0.008333017118275166f);
    ps = fma(ps, r2, -0.16666656732559204f);
    ps = mul(ps, r2);
    ps = fma(ps, r, r);
    V pc = fma(r2, P.c_c4, 2.476187000866048e-05f);
    pc = fma(pc, r2, -0.00138884037733078f);
    pc = fma(pc, r2, 0.04166664183139801f);
    pc = fma(pc, r2, -0.5f);
    pc = fma(pc, r2, 1.0f);
    const uint32_t fx = (uint32_t)__float_as_int(t.x) << 31, fy = (uint32_t)__float_as_int(t.y) << 31;
    sn.x = __int_as_float(__float_as_int(ps.x) ^ fx); sn.y = __int_as_float(__float_as_int(ps.y) ^ fy);
    cs.x = __int_as_float(__float_as_int(pc.x) ^ fx); cs.y = __int_as_float(__float_as_int(pc.y) ^ fy);
}

// atan of two arguments: z = x for |x| <= 1, 1 / x beyond (atan x = sign(x) pi/2 - atan(1 / x)); atan z = z + z^3 P(z^2)
__device__ __forceinline__ V atan2v(const ShipPX& P, const V x) {
    const bool ix = fabsf(x.x) > 1.0f, iy = fabsf(x.y) > 1.0f;
    V z;
    z.x = ix ? rcp_approx(x.x) : x.x;
    z.y = iy ? rcp_approx(x.y) : x.y;
    const V z2 = mul(z, z);
    V p = fma(z2, P.c_a6, 0.023039964959025383f);
    p = fma(p, z2, -0.05777335166931152f);
    p = fma(p, z2, 0.09794218093156815f);
    p = fma(p, z2, -0.13976575434207916f);
    p = fma(p, z2, 0.19962702691555023f);
    p = fma(p, z2, -0.3333165943622589f);
    p = mul(p, z2);
    p = fma(p, z, z);
    V h;   // copysign(pi / 2, x)
    h.x = __int_as_float((__float_as_int(x.x) & 0x80000000) | 0x3fc90fdb);
    h.y = __int_as_float((__float_as_int(x.y) & 0x80000000) | 0x3fc90fdb);
    const V alt = sub(h, p);
    V res;
    res.x = ix ? alt.x : p.x;
    res.y = iy ? alt.y : p.y;
    return res;
}

// One step of two ships.  MP = ModeX: one machinery mode for the whole fleet (uniform operands); MP = ModeV: per-ship modes.
// idx(l): the ship index of lane l (only read on the rare waypoint-switch path).  AVOID: the collision-avoidance branch on
// no ship, both, or lane 0 only (the two-ship env: lane 0 = test ship, MSRL_env_ex.py:240-249).  STORE = false reproduces
// init_step (MSRL_env_ex.py:188-215): no store_simulation_data, i.e. no fuel accumulation.
// AVOID_LANE0_RT: lane-0 avoidance and the store are switched by the run-time flags avoid0_rt / store_rt, so that ONE instance
// of the step serves both the normal step and init_step (the env kernel's instruction footprint); COMPACT: the rare
// waypoint-switch path as a two-trip loop around real calls instead of two inlined copies.
template <int AVOID, bool STORE, bool COMPACT = false, typename MP, typename IdxFn>
__device__ __forceinline__ void ship_step2(const ShipPX& P, const MP& M, Ship2& s, const ShipPtrs<float>& g, const IdxFn& idx,
                                           StepOut2& o, const bool avoid0_rt = true, const bool store_rt = true) {
    // ---- LOS_guidance.py:88-103 next_wpt: packed float pre-test; the decision itself in double (next_wpt2) -----------------
    {
        const V dn = sub(s.SX, s.X), de = sub(s.SY, s.Y);
        const V d2 = fma(de, de, mul(dn, dn));
        const bool h0 = d2.x <= s.thr2.x, h1 = d2.y <= s.thr2.y;
        if (h0 | h1) {
            if (COMPACT) {
#pragma unroll 1
                for (int l = 0; l < 2; ++l)
                    if (l == 0 ? h0 : h1) next_wpt2<true>(s, P, g, idx(l), l);
            } else {
                if (h0) next_wpt2(s, P, g, idx(0), 0);
                if (h1) next_wpt2(s, P, g, idx(1), 1);
            }
        }
    }
    // ---- LOS_guidance.py:105-121 los_guidance ---------------------------------------------------------------------
    V e = fma(s.Y, s.ca, neg(mul(s.X, s.sa)));
    s.e_ct = e;                                             // |e| is taken where it is stored
    const V e2 = mul(e, e);
    if (e2.x >= P.look2) e.x = P.sat_e;                      // sign lost on purpose (0.99 * lookahead)
    if (e2.y >= P.look2) e.y = P.sat_e;
    const V w = fma(neg(e), e, P.look2);
    const V q = mul(e, make_float2(rsqrt_approx(w.x), rsqrt_approx(w.y)));
    const V ti = add(s.e_int, q);
    if (fabsf(ti.x) <= P.windup) s.e_int.x = ti.x;
    if (fabsf(ti.y) <= P.windup) s.e_int.y = ti.y;
    const V href = add(s.alpha, atan2v(P, fma(s.e_int, P.neg_ki_los, neg(q))));

    // ---- controllers.py:81-93,188-189 heading PID: no wrap, no anti-windup -------------------------------------------
    const V err = sub(href, s.psi);
    s.ei_h = fma(err, P.dt, s.ei_h);
    // kp err + kd (err - e_prev) / dt + ki ei  =  (kp + kd / dt) err - (kd / dt) e_prev + ki ei: one packed op fewer
    V rud = fma(s.e_prev, -P.kd_idt, mul(err, P.kp_kd));
    s.e_prev = err;
    rud = fma(s.ei_h, P.ki_h, rud);                          // negated in the clamp below
    rud.x = fmaxf(-P.rud_max, fminf(-rud.x, P.rud_max));
    rud.y = fmaxf(-P.rud_max, fminf(-rud.y, P.rud_max));

    // ---- controllers.py:138-143 throttle cascade; the shaft PI is fed u (MSRL_env_ex.py:233-237) ----------------------
    const V e1 = sub(s.u_sp, s.u);
    s.ei_s = fma(e1, P.dt, s.ei_s);
    const V e2s = fma(e1, P.kp_s, fma(s.ei_s, P.ki_s, neg(s.u)));
    s.ei_w = fma(e2s, P.dt, s.ei_w);
    V thr = fma(e2s, P.kp_w, mul(s.ei_w, P.ki_w));

    // ---- MSRL_env_ex.py:240-249 collision-avoidance branch (test ship: always) ---------------------------------------
    if (AVOID != AVOID_NONE && (AVOID != AVOID_LANE0_RT || avoid0_rt)) {
        thr.x = fminf(fmaxf(thr.x * 0.5f, 0.0f), 1.1f);
        rud.x = fminf(fmaxf(rud.x + P.rud_bias, -P.rud_max), P.rud_max);
    }
    if (AVOID == AVOID_BOTH) {
        thr.y = fminf(fmaxf(thr.y * 0.5f, 0.0f), 1.1f);
        rud.y = fminf(fmaxf(rud.y + P.rud_bias, -P.rud_max), P.rud_max);
    }

    // ---- ship_engine.py:363-384 torques (capped), thrust from the pre-update omega ------------------------------------
    const V om = s.om;
    const V omp = add(om, 0.1f);
    const V tio = mul(thr, make_float2(rcp_approx(omp.x), rcp_approx(omp.y)));
    V tq_me = mul(tio, M.avail_me), tq_hsg = mul(tio, M.avail_el);
    tq_me.x = fminf(tq_me.x, lane(M.cap_tq_me, 0)); tq_me.y = fminf(tq_me.y, lane(M.cap_tq_me, 1));
    tq_hsg.x = fminf(tq_hsg.x, lane(M.cap_tq_el, 0)); tq_hsg.y = fminf(tq_hsg.y, lane(M.cap_tq_el, 1));
    const V thrust_w = mul(om, vabs(om));

    // ---- ship_model.py:645-684 store_simulation_data: load split + fuel accumulators (ship_engine.py:239-292) -----------
    o.om_pre = om;
    o.load_me = bc(0.0f);
    if (STORE && (AVOID != AVOID_LANE0_RT || store_rt)) {
        // total = thr * available power is never needed on its own: it enters the two loads through fused multiply-adds
        V load_me = fma(thr, M.avail_total, M.add_me);
        load_me.x = fminf(load_me.x, lane(M.cap_me, 0)); load_me.y = fminf(load_me.y, lane(M.cap_me, 1));
        const V load_el = fma(sub(fma(thr, M.avail_total, P.hotel), load_me), M.a_el, M.c_el);
        // rate dt = load (a pct^2 + b pct + c) dt / 3.6e9 with pct = load / capacity folded into the coefficients (ModeX); a zero
        // load gives a zero rate without the reference's guards
        const V pme = fma(fma(load_me, M.fme_a, M.fme_b), load_me, P.sfc_me[2]);
        const V pel = fma(fma(load_el, M.fel_a, M.fel_b), load_el, P.sfc_dg[2]);
        o.load_me = load_me;
        s.f_me = fma(load_me, pme, s.f_me);
        s.f_el = fma(load_el, pel, s.f_el);   // fuel_cons (the sum of both, ship_engine.py:288-292) is formed where the state is stored
    }

    // ---- ship_model.py:233-250 kinematics ---------------------------------------------------------------------------
    V sn, c;
    sincos2(P, s.psi, sn, c);
    const V d_north = fma(c, s.u, neg(mul(sn, s.v)));      // [m/s]; the dt factor rides the Kahan update below
    const V d_east = fma(sn, s.u, mul(c, s.v));

    // ---- ship_engine.py:355-361,391-395 shaft equation ------------------------------------------------------------------
    V d_om = mul(tq_me, P.sh_me);
    d_om = fma(tq_hsg, P.sh_hsg, d_om);
    d_om = fma(neg(fma(om, P.sh_quad, P.sh_lin)), om, d_om);

    // ---- ship_model.py:576-622 kinetics: current in the body frame, wind, rudder -----------------------------------------
    const V ur = fma(neg(sn), s.vce, fma(neg(c), s.vcn, s.u));      // u - (c vcn + s vce)
    const V vr = fma(sn, s.vcn, fma(neg(c), s.vce, s.v));           // v - (c vce - s vcn)
    const V urw = fma(s.wsb, sn, fma(s.wcb, c, neg(s.u)));          // W cos(b - psi) - u
    const V vrw = fma(neg(s.wcb), sn, fma(s.wsb, c, neg(s.v)));     // W sin(b - psi) - v
    const V w2 = fma(vrw, vrw, mul(urw, urw));
    const V Vw = make_float2(sqrt_approx(w2.x), sqrt_approx(w2.y));
    const V ru = mul(rud, ur);
    V d_u = mul(fma(vr, -P.au_vrr, mul(s.v, P.au_vr)), s.r);
    d_u = fma(neg(fma(s.u, P.au_nl, P.au_lin)), ur, d_u);
    d_u = fma(mul(Vw, urw), -P.au_wind, d_u);
    d_u = fma(thrust_w, P.au_thr, d_u);
    V d_v = mul(fma(s.u, -P.av_ur, mul(ur, P.av_urr)), s.r);
    d_v = fma(neg(fma(s.v, P.av_nl, P.av_lin)), vr, d_v);
    d_v = fma(mul(Vw, vrw), -P.av_wind, d_v);
    d_v = fma(ru, -P.av_rud, d_v);
    V d_r = mul(neg(fma(s.r, P.ar_nl, P.ar_lin)), s.r);
    d_r = fma(mul(ur, vr), -P.ar_ca, d_r);
    d_r = fma(mul(urw, vrw), -P.ar_wind, d_r);
    d_r = fma(ru, -P.ar_rud, d_r);

    // ---- ship_model.py:632-643, utils.py:50-53 explicit Euler, all derivatives from the old state ---------------------------
    {   // position: Kahan-compensated float accumulation of the float increments
        const V yn = fma(d_north, P.dt, neg(s.CX)), ye = fma(d_east, P.dt, neg(s.CY));
        const V tn = add(s.X, yn), te = add(s.Y, ye);
        s.CX = sub(sub(tn, s.X), yn); s.CY = sub(sub(te, s.Y), ye);
        s.X = tn; s.Y = te;
    }
    s.psi = fma(s.r, P.dt, s.psi);
    s.u = fma(d_u, P.dt, s.u);
    s.v = fma(d_v, P.dt, s.v);
    s.r = fma(d_r, P.dt, s.r);
    s.om = fma(d_om, P.dt, om);
}

}  // namespace x2
}  // namespace sit
