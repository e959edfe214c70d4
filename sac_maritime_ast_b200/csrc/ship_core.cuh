// Device-side model of one ship-step, templated on the arithmetic type R (double | float).
//
// One thread owns one ship; its state lives in registers (struct Ship) for the K steps of a
// launch and is read / written as coalesced SoA rows at the launch boundaries.  The step is the
// algebraically reduced form of the reference's scalar Python step (the CPU oracle keeps the
// literal form, so a transcription error in either shows up as a parity failure):
//   * the route segment in use (previous / current waypoint, segment angle and its sin / cos) is
//     cached in registers and rebuilt only when the waypoint index changes;
//   * the wind force uses cos(gamma) = u_rw / V, sin(gamma) = -v_rw / V instead of
//     atan2 -> cos / sin / sin(2 gamma);  wind direction enters as (W cos beta, W sin beta);
//   * M is diagonal (x_g = 0, ship_model.py:79), so the three 3x3 inversions of the reference
//     collapse to pre-divided coefficients.
// Positions (north, east) are carried in double in BOTH precisions so that waypoint switching and
// the other distance predicates are decided on the same numbers the float64 reference sees.
//
// Reference citations are relative to /root/reference/.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/shipsim.h"

namespace sit {

// ---- math wrappers ---------------------------------------------------------------------------
// double: IEEE-rounded library calls (the 1e-9 parity path).  float: branch-free forms with <= 2 ulp error
// (MUFU approximations, Cody-Waite sin/cos) -- no slow-path calls, so ptxas keeps the uniform parameters in
// uniform registers across the K-step loop.
__device__ __forceinline__ float rcp_approx(float x) { float r; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
__device__ __forceinline__ float rsqrt_approx(float x) { float r; asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
__device__ __forceinline__ float sqrt_approx(float x) { float r; asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }

__device__ __forceinline__ void sincos_(double x, double& s, double& c) { sincos(x, &s, &c); }
__device__ __forceinline__ void sincos_(float x, float& s, float& c) {
    // |x| stays far below 1e5 (the heading is unwrapped but bounded by the turns of a route): three-constant
    // Cody-Waite reduction by pi/2, then the cephes minimax polynomials on [-pi/4, pi/4]
    const float t = fmaf(x, 0.636619772f, 12582912.0f);
    const int q = __float_as_int(t);
    const float j = t - 12582912.0f;
    float r = fmaf(j, -1.57079601e+00f, x);
    r = fmaf(j, -3.13916473e-07f, r);
    r = fmaf(j, -5.39030253e-15f, r);
    const float r2 = r * r;
    float ps = fmaf(r2, -1.9515295891e-4f, 8.3321608736e-3f);
    ps = fmaf(ps, r2, -1.6666654611e-1f);
    ps = fmaf(ps * r2, r, r);
    float pc = fmaf(r2, 2.443315711809948e-5f, -1.388731625493765e-3f);
    pc = fmaf(pc, r2, 4.166664568298827e-2f);
    pc = fmaf(pc, r2, -0.5f);
    pc = fmaf(pc, r2, 1.0f);
    const bool swap = q & 1;
    const float ss = swap ? pc : ps, cc = swap ? ps : pc;
    s = (q & 2) ? -ss : ss;
    c = ((q + 1) & 2) ? -cc : cc;
}
__device__ __forceinline__ double atan_(double x) { return atan(x); }
__device__ __forceinline__ float atan_(float x) { return atanf(x); }
__device__ __forceinline__ double atan2_(double y, double x) { return atan2(y, x); }
__device__ __forceinline__ float atan2_(float y, float x) { return atan2f(y, x); }
__device__ __forceinline__ double sqrt_(double x) { return sqrt(x); }
__device__ __forceinline__ float sqrt_(float x) { return sqrt_approx(x); }
__device__ __forceinline__ double abs_(double x) { return fabs(x); }
__device__ __forceinline__ float abs_(float x) { return fabsf(x); }
__device__ __forceinline__ double min_(double a, double b) { return fmin(a, b); }
__device__ __forceinline__ float min_(float a, float b) { return fminf(a, b); }
__device__ __forceinline__ double max_(double a, double b) { return fmax(a, b); }
__device__ __forceinline__ float max_(float a, float b) { return fmaxf(a, b); }
// e / sqrt(x): one MUFU.RSQ in float, exact sqrt + div in double
__device__ __forceinline__ double over_sqrt_(double e, double x) { return e / sqrt(x); }
__device__ __forceinline__ float over_sqrt_(float e, float x) { return e * rsqrt_approx(x); }
__device__ __forceinline__ double rcp_(double x) { return 1.0 / x; }
__device__ __forceinline__ float rcp_(float x) { const float r = rcp_approx(x); return fmaf(fmaf(-x, r, 1.0f), r, r); }
// load / 1000 (ship_model.py:661): exact division on the float64 parity path, reciprocal multiply in float
__device__ __forceinline__ double kilo_(double x) { return x / 1000.0; }
__device__ __forceinline__ float kilo_(float x) { return x * 0.001f; }

// ---- uniform parameters (kernel arguments -> constant bank, one copy per role) -----------------
template <typename R>
struct ModeP {  // MachineryMode, ship_engine.py:23-76, pre-digested
    R avail_total, avail_me, avail_el;  // update_available_propulsion_power :32-44
    R cap_tq_me, cap_tq_el;             // avail / 5 * pi / 30, :374-375,383-384
    R me_cap, me_cap_kw, el_cap_kw;     // capacity, and /1000 for the logged channels
    R inv_me_cap, inv_el_cap;           // 0 where the reference guards a zero capacity
    R load_me_add, load_el_const;       // distribute_load :46-76 folded per generator state
    int32_t sg, _pad;
};

template <typename R>
struct ShipP {
    R dt, inv_dt;
    double ra2;                                   // radius_of_acceptance^2 (decided in double)
    R look, look2, sat_e, ki_los, windup;         // LOS_guidance.py:105-121
    R kp_h, kd_h, ki_h, rud_max, rud_bias;        // controllers.py:72-100; 3 deg bias MSRL_env_ex.py:248
    R kp_s, ki_s, kp_w, ki_w;                     // controllers.py:138-143
    R hotel;
    R sfc_me[3], sfc_dg[3];                       // ship_engine.py:256-261 (already / 3.6e9)
    R sh_me, sh_hsg, sh_lin, sh_quad;             // shaft_eq ship_engine.py:355-361, / J_p folded
    R thrust_coef;                                // D^4 k_t, ship_engine.py:363-366
    R rad2deg, rpm_per_rads;
    // accelerations = forces / (m + added mass), ship_model.py:576-606 with :252-276 expanded
    R au_vr, au_vrr, au_lin, au_nl, au_wind, au_thr;
    R av_ur, av_urr, av_lin, av_nl, av_wind, av_rud;
    R ar_ca, ar_lin, ar_nl, ar_wind, ar_rud;
    R half_len;                                   // L / 2, MSRL_env_ex.py:476,498
    // SimplifiedMachineryModel (ship_engine.py:398-433) + ThrottleFromSpeedSetPointSimplifiedPropulsion (controllers.py:154-172)
    R k_thrust, inv_thrust_T, kp_sp, ki_sp, inv_m11;
    int32_t mach, _pad_m;
    ModeP<R> mode[SIT_MAX_MODES];
    int32_t n_modes, _pad;
};

// ---- SoA views of the caller-owned state buffer ----------------------------------------------
template <typename R>
struct ShipPtrs {
    double *north, *east, *init_north, *init_east;
    R *psi, *u, *v, *r, *omega, *ei_speed, *ei_shaft, *ei_heading, *e_prev, *e_ct, *e_ct_int;
    R *fuel_me, *fuel_el, *fuel, *u_sp, *vc_n, *vc_e, *wind_speed, *wind_dir;
    R *init_psi, *init_u, *init_v, *init_r, *init_vc_n, *init_vc_e, *init_wind_speed, *init_wind_dir,
        *init_u_sp, *init_omega, *init_ei_shaft;
    int32_t *k, *n_wp, *mode, *n_wp0, *steps;
    R *wp_n, *wp_e, *wp0_n, *wp0_e;  // [max_wp][n_ships]
    int32_t* mc_idx; float *mc_n, *mc_e, *mc_b2;   // nearest-ring cache (env kernel)
    float *rel_n, *rel_cn, *rel_e, *rel_ce;        // packed fp32 fleet kernel: compensated position relative to the segment start
    int64_t n_ships;
    int32_t max_wp, _pad;
};

// ---- per-ship register state --------------------------------------------------------------------
template <typename R>
struct Ship {
    double N, E;                                   // ship_model.py:159-160
    R psi, u, v, r, om;                            // :161-164, ship_engine.py:327
    R ei_s, ei_w, ei_h, e_prev, e_ct, e_int;       // controllers.py:49,77-78; LOS_guidance.py:59-60
    R f_me, f_el, f_tot;                           // ship_engine.py:336-338
    R u_sp, vcn, vce, wcb, wsb;                    // set point; current; wind as W cos b, W sin b
    double pN, pE, cN, cE;                         // cached route segment k-1 -> k (widened once per switch)
    R alpha, sa, ca;                               // its angle, sin, cos
    int32_t k, n_wp, mode, steps;
};

template <typename R>
__device__ __forceinline__ void load_segment(Ship<R>& s, const ShipPtrs<R>& g, int64_t i) {
    // LOS_guidance.py:110-112: segment k-1 -> k and its angle (atan2(dE, dN))
    const int64_t S = g.n_ships;
    s.pN = g.wp_n[(int64_t)(s.k - 1) * S + i];
    s.pE = g.wp_e[(int64_t)(s.k - 1) * S + i];
    s.cN = g.wp_n[(int64_t)s.k * S + i];
    s.cE = g.wp_e[(int64_t)s.k * S + i];
    s.alpha = atan2_((R)(s.cE - s.pE), (R)(s.cN - s.pN));
    sincos_(s.alpha, s.sa, s.ca);
}

template <typename R>
__device__ __forceinline__ void set_wind(Ship<R>& s, R wind_speed, R wind_dir) {
    R sb, cb;
    sincos_(wind_dir, sb, cb);
    s.wcb = wind_speed * cb;
    s.wsb = wind_speed * sb;
}

template <typename R>
__device__ __forceinline__ void load_ship(Ship<R>& s, const ShipPtrs<R>& g, int64_t i) {
    s.N = g.north[i]; s.E = g.east[i];
    s.psi = g.psi[i]; s.u = g.u[i]; s.v = g.v[i]; s.r = g.r[i]; s.om = g.omega[i];
    s.ei_s = g.ei_speed[i]; s.ei_w = g.ei_shaft[i]; s.ei_h = g.ei_heading[i]; s.e_prev = g.e_prev[i];
    s.e_ct = g.e_ct[i]; s.e_int = g.e_ct_int[i];
    s.f_me = g.fuel_me[i]; s.f_el = g.fuel_el[i]; s.f_tot = g.fuel[i];
    s.u_sp = g.u_sp[i]; s.vcn = g.vc_n[i]; s.vce = g.vc_e[i];
    set_wind(s, g.wind_speed[i], g.wind_dir[i]);
    s.k = g.k[i]; s.n_wp = g.n_wp[i]; s.mode = g.mode[i]; s.steps = g.steps[i];
    load_segment(s, g, i);
}

template <typename R>
__device__ __forceinline__ void store_ship(const Ship<R>& s, const ShipPtrs<R>& g, int64_t i) {
    g.north[i] = s.N; g.east[i] = s.E;
    g.psi[i] = s.psi; g.u[i] = s.u; g.v[i] = s.v; g.r[i] = s.r; g.omega[i] = s.om;
    g.ei_speed[i] = s.ei_s; g.ei_shaft[i] = s.ei_w; g.ei_heading[i] = s.ei_h; g.e_prev[i] = s.e_prev;
    g.e_ct[i] = s.e_ct; g.e_ct_int[i] = s.e_int;
    g.fuel_me[i] = s.f_me; g.fuel_el[i] = s.f_el; g.fuel[i] = s.f_tot;
    g.k[i] = s.k; g.n_wp[i] = s.n_wp; g.steps[i] = s.steps;
    // u_sp, current, wind and mode are inputs: written by the caller (or by the disturbance path)
}

template <typename R>
struct StepOut {
    R rud, thr;                // commanded this step
    R rpm, power_me_kw, avail_me_kw;  // stored (pre-step) observables the env reads back
};

// One ship-step.  avoid: MSRL_env_ex.py:240-249.  store: store_simulation_data + next_time
// (false reproduces init_step, MSRL_env_ex.py:188-215).  lg: 27 logged channels or nullptr.
// MACH = SIT_MACH_SIMPLIFIED: the machinery state s.om is the THRUST FORCE (first-order lag of the commanded power) and the
// throttle is the saturated speed PI; guidance, heading PID, load split, fuel and the hull are the same code.
template <typename R, bool LOG, bool SINGLE_MODE = false, int MACH = SIT_MACH_FULL>
__device__ __forceinline__ void ship_step(const ShipP<R>& P, Ship<R>& s, const ShipPtrs<R>& g, int64_t i,
                                          bool avoid, bool store, StepOut<R>& o, R* lg) {
    // ---- LOS_guidance.py:88-103 next_wpt: squared distance, <=, end-of-route hold --------------
    {
        const double dn = s.cN - s.N, de = s.cE - s.E;
        const double d2 = __dadd_rn(__dmul_rn(dn, dn), __dmul_rn(de, de));
        if (d2 <= P.ra2 && s.n_wp > s.k + 1) {
            s.k += 1;
            load_segment(s, g, i);
        }
    }
    // ---- LOS_guidance.py:105-121 los_guidance ----------------------------------------------------
    const R en = (R)(s.N - s.pN), ee = (R)(s.E - s.pE);
    R e = ee * s.ca - en * s.sa;
    s.e_ct = abs_(e);
    if (e * e >= P.look2) e = P.sat_e;  // sign lost on purpose (0.99 * lookahead)
    const R q = over_sqrt_(e, P.look2 - e * e);
    if (abs_(s.e_int + q) <= P.windup) s.e_int += q;
    const R href = s.alpha + atan_(-q - s.e_int * P.ki_los);

    // ---- controllers.py:81-93,188-189 heading PID: no wrap, no anti-windup ----------------------
    const R err = href - s.psi;
    const R d_err = (err - s.e_prev) * P.inv_dt;
    s.ei_h += err * P.dt;
    s.e_prev = err;
    R rud = -(err * P.kp_h + d_err * P.kd_h + s.ei_h * P.ki_h);
    rud = max_(-P.rud_max, min_(rud, P.rud_max));

    // ---- controllers.py:138-143 throttle cascade; the shaft PI is fed u (MSRL_env_ex.py:233-237) -
    const R e1 = s.u_sp - s.u;
    s.ei_s += e1 * P.dt;
    R thr;
    if (MACH == SIT_MACH_SIMPLIFIED) {
        // controllers.py:170-172: sat(pi_ctrl(set point, speed), 0, 1.1) with sat = max(low, min(val, hi)) (:63-69)
        thr = max_((R)0, min_(e1 * P.kp_sp + s.ei_s * P.ki_sp, (R)1.1));
    } else {
        const R e2 = (e1 * P.kp_s + s.ei_s * P.ki_s) - s.u;
        s.ei_w += e2 * P.dt;
        thr = e2 * P.kp_w + s.ei_w * P.ki_w;
    }

    // ---- MSRL_env_ex.py:240-249 collision-avoidance branch (test ship: always) -------------------
    if (avoid) {
        thr = min_(max_(thr * (R)0.5, (R)0), (R)1.1);
        rud = min_(max_(rud + P.rud_bias, -P.rud_max), P.rud_max);
    }
    o.rud = rud; o.thr = thr;

    // the reference's default MachineryModes holds ONE mode (main_ast.py:159-161): its constants are then
    // uniform operands; with several modes the per-ship index makes them per-lane constant loads
    const ModeP<R>& M = SINGLE_MODE ? P.mode[0] : P.mode[s.mode];
    const R om = s.om;
    R tq_me = (R)0, tq_hsg = (R)0, thrust_w = (R)0;
    if (MACH == SIT_MACH_FULL) {
        const R inv_om = rcp_(om + (R)0.1);
        tq_me = min_(thr * M.avail_me * inv_om, M.cap_tq_me);       // ship_engine.py:368-375
        tq_hsg = min_(thr * M.avail_el * inv_om, M.cap_tq_el);      // :377-384
        thrust_w = om * abs_(om);                                   // :363-366 (pre-update omega)
    }

    // ---- ship_model.py:645-684 store_simulation_data: observables + fuel accumulators ------------
    if (store) {
        const R total = thr * M.avail_total;                        // ship_engine.py:239-254
        R load_me, load_el;
        if (M.sg == SIT_SG_MOTOR) {
            load_me = min_(total, M.me_cap);
            load_el = total + P.hotel - load_me;
        } else {
            load_me = total + M.load_me_add;
            load_el = M.load_el_const;
        }
        const R pct_me = load_me * M.inv_me_cap, pct_el = load_el * M.inv_el_cap;
        // ship_engine.py:263-292 fuel_consumption
        const R rate_me = (load_me == (R)0) ? (R)0 : load_me * ((P.sfc_me[0] * pct_me + P.sfc_me[1]) * pct_me + P.sfc_me[2]);
        const R rate_el = (pct_el == (R)0) ? (R)0 : load_el * ((P.sfc_dg[0] * pct_el + P.sfc_dg[1]) * pct_el + P.sfc_dg[2]);
        s.f_me += rate_me * P.dt;
        s.f_el += rate_el * P.dt;
        s.f_tot += (rate_me + rate_el) * P.dt;
        o.rpm = MACH == SIT_MACH_FULL ? om * P.rpm_per_rads : (R)0;   // the simplified model has no shaft
        o.power_me_kw = kilo_(load_me);
        o.avail_me_kw = M.me_cap_kw;
        if (LOG && lg) {
            lg[0] = (R)s.steps * P.dt;          lg[1] = (R)s.N;                 lg[2] = (R)s.E;
            lg[3] = s.psi * P.rad2deg;          lg[4] = rud * P.rad2deg;        lg[5] = s.u;
            lg[6] = s.v;                        lg[7] = s.r * P.rad2deg;        lg[8] = o.rpm;
            lg[9] = pct_me;                     lg[10] = pct_el;                lg[11] = o.power_me_kw;
            lg[12] = M.me_cap_kw;               lg[13] = kilo_(load_el);     lg[14] = M.el_cap_kw;
            lg[15] = kilo_(load_el + load_me);                                  lg[16] = kilo_(total);
            lg[17] = rate_me;                   lg[18] = rate_el;               lg[19] = rate_me + rate_el;
            lg[20] = s.f_me;                    lg[21] = s.f_el;                lg[22] = s.f_tot;
            lg[23] = tq_me;                     lg[24] = MACH == SIT_MACH_FULL ? kilo_(P.thrust_coef * thrust_w) : kilo_(om);
            lg[25] = s.e_ct;                    lg[26] = abs_(s.psi - href);    // radians (controllers.py:327-328)
        }
        s.steps += 1;                                               // utils.py:42-48 next_time
    }

    // ---- ship_model.py:233-250 kinematics ---------------------------------------------------------
    R sn, c;
    sincos_(s.psi, sn, c);
    const R d_north = c * s.u - sn * s.v;
    const R d_east = sn * s.u + c * s.v;

    // ---- ship_engine.py:355-361,391-395 shaft equation --------------------------------------------
    // simplified: ship_engine.py:423-428 update_thrust_force, d_thrust = (-k_thrust thrust + load_perc (P_me + P_el)) / T
    const R d_om = MACH == SIT_MACH_FULL ? tq_me * P.sh_me + tq_hsg * P.sh_hsg - (P.sh_lin + P.sh_quad * om) * om
                                         : (thr * (M.avail_me + M.avail_el) - P.k_thrust * om) * P.inv_thrust_T;

    // ---- ship_model.py:576-622 kinetics ------------------------------------------------------------
    const R uc = c * s.vcn + sn * s.vce, vc = c * s.vce - sn * s.vcn;   // inv(R(psi)) * v_current
    const R ur = s.u - uc, vr = s.v - vc;
    // wind, ship_model.py:211-231: (uw, vw) = W (cos(b - psi), sin(b - psi))
    const R urw = (s.wcb * c + s.wsb * sn) - s.u;
    const R vrw = (s.wsb * c - s.wcb * sn) - s.v;
    const R Vw = sqrt_(urw * urw + vrw * vrw);
    const R d_u = P.au_vr * s.v * s.r - P.au_vrr * vr * s.r - (P.au_lin + P.au_nl * s.u) * ur -
                  P.au_wind * Vw * urw + (MACH == SIT_MACH_FULL ? P.au_thr * thrust_w : P.inv_m11 * om);
    const R d_v = P.av_urr * ur * s.r - P.av_ur * s.u * s.r - (P.av_lin + P.av_nl * s.v) * vr -
                  P.av_wind * Vw * vrw - P.av_rud * rud * ur;
    const R d_r = -P.ar_ca * ur * vr - (P.ar_lin + P.ar_nl * s.r) * s.r - P.ar_wind * urw * vrw -
                  P.ar_rud * rud * ur;

    // ---- ship_model.py:632-643, utils.py:50-53 explicit Euler, all derivatives from the old state --
    s.N += (double)(d_north * P.dt);
    s.E += (double)(d_east * P.dt);
    s.psi += s.r * P.dt;
    s.u += d_u * P.dt;
    s.v += d_v * P.dt;
    s.r += d_r * P.dt;
    s.om = om + d_om * P.dt;
}

// ShipModelAST.reset (ship_model.py:340-400,708-712), throttle reset no-op (controllers.py:145-151),
// auto-pilot + navigation reset (controllers.py:333-350, LOS_guidance.py:123-136), set point
// (MSRL_env_ex.py:165).  omega, fuel and the three controller integrators keep their values unless
// `hard`.  Writes straight to the state arrays (rare path).
template <typename R>
__device__ __forceinline__ void reset_ship_arrays(const ShipPtrs<R>& g, int64_t i, bool hard) {
    const int64_t S = g.n_ships;
    g.north[i] = g.init_north[i]; g.east[i] = g.init_east[i];
    g.psi[i] = g.init_psi[i]; g.u[i] = g.init_u[i]; g.v[i] = g.init_v[i]; g.r[i] = g.init_r[i];
    g.vc_n[i] = g.init_vc_n[i]; g.vc_e[i] = g.init_vc_e[i];
    g.wind_speed[i] = g.init_wind_speed[i]; g.wind_dir[i] = g.init_wind_dir[i];
    g.u_sp[i] = g.init_u_sp[i];
    g.steps[i] = 0; g.k[i] = 1;
    g.e_ct[i] = (R)0; g.e_ct_int[i] = (R)0;
    const int32_t n0 = g.n_wp0[i];
    g.n_wp[i] = n0;
    for (int j = 0; j < g.max_wp; ++j) {
        g.wp_n[(int64_t)j * S + i] = g.wp0_n[(int64_t)j * S + i];
        g.wp_e[(int64_t)j * S + i] = g.wp0_e[(int64_t)j * S + i];
    }
    if (hard) {
        g.omega[i] = g.init_omega[i];
        g.ei_speed[i] = (R)0; g.ei_shaft[i] = g.init_ei_shaft[i];
        g.ei_heading[i] = (R)0; g.e_prev[i] = (R)0;
        g.fuel_me[i] = (R)0; g.fuel_el[i] = (R)0; g.fuel[i] = (R)0;
    }
}

}  // namespace sit
