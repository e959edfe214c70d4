// libshipsim.so -- kernels and C ABI of the batched ship-in-transit simulator (see include/shipsim.h).
// sm_100a only; no CPU fallback.  Reference citations are relative to /root/reference/.
#include <cuda_runtime.h>

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <new>
#include <string>
#include <vector>

#include "env_core.cuh"
#include "ship_x2.cuh"

using namespace sit;

// =================================================================================================
// Kernel argument blocks
// =================================================================================================
template <typename R>
struct FleetIO {
    const R* dist; int32_t dist_steps; int32_t avoid; int32_t store; int32_t traj_steps;
    R* traj; R* log27; R* cmd;
    int8_t* k8;              // optional: waypoint index of the last step as int8 [n_ships] (host-buffer path)
    int64_t first, count;    // ship sub-range [first, first + count) of this launch (the host-buffer pipeline launches chunks)
};

template <typename R>
struct EnvPtrs {
    double *prev_n, *prev_e, *eps_dist, *samp_dist;
    R *ab_seg, *act_n, *act_e, *ep_return;
    R *ab_alpha, *cur_obs;                       // rollout: AB heading (MSRL_env_ex.py:125), observation held for the policy [10][B]
    int32_t *flags, *ep_steps;
    int32_t *n_dec, *dec_serial, *tr_slot;       // rollout: decisions this episode / ever, open replay row
    double* draw_t;                              // time_since_last_ship_drawing (MSRL_env_ex.py:60,416-421)
    int64_t n_envs;
};

struct RollArgs {   // device-resident rollout (include/shipsim.h SitRollout), by value
    float* replay; int64_t capacity; unsigned long long* counters;
    uint64_t seed; int64_t env_id_offset;
    int32_t max_episode_steps, max_route_resampling, sampling_frequency, truncate;
    int64_t env_first, env_count;   // env sub-range of this launch (sit_rollout_collect pipelines env groups)
};

template <typename R>
struct EnvIO {
    const R* action; const uint8_t* sac_update; int32_t act_steps; int32_t init; int32_t auto_reset; int32_t rec_steps;
    R* obs; R* reward; uint8_t* done; int32_t* status; R* reward_sum; uint8_t* done_any; unsigned long long* stats;
    const uint8_t* mask;   // init_step only
    R* rew_parts;          // [rec][SIT_N_REWARD_PARTS][B] per-step increments of reward_results (MSRL_env_ex.py:924-964)
    uint8_t* snap;         // [rec][B]: ship_snap_shot() taken in this step (MSRL_env_ex.py:416-421)
    double draw_dt;        // ship_model.int.dt as the reference adds it to the timer (double)
};

template <typename R>
struct EnvKArgs {
    ShipP<R> P[2];
    EnvP<R> EP;
    MapP<R> M;
};

// =================================================================================================
// K1: fleet of bare ships, K steps per launch, state in registers
// =================================================================================================
template <typename R, bool LOG, int MACH = SIT_MACH_FULL>
__global__ void __launch_bounds__(256) fleet_step_kernel(const __grid_constant__ ShipP<R> P, const ShipPtrs<R> g,
                                                         const int K, const FleetIO<R> io) {
    const int64_t li = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t S = g.n_ships, i = io.first + li;
    if (li >= io.count) return;
    Ship<R> s;
    load_ship(s, g, i);
    R w_speed = (R)0, w_dir = (R)0;
    for (int t = 0; t < K; ++t) {
        if (io.dist && (t == 0 || io.dist_steps > 1)) {
            // ship_model.py:154-156: vel_c / wind_speed / wind_dir are plain attributes, assigned per step
            const R* d = io.dist + (int64_t)(io.dist_steps > 1 ? t : 0) * 4 * S;
            s.vcn = d[i]; s.vce = d[S + i];
            w_speed = d[2 * S + i]; w_dir = d[3 * S + i];
            set_wind(s, w_speed, w_dir);
        }
        StepOut<R> o;
        R lg[SIT_N_LOG];
        ship_step<R, LOG, false, MACH>(P, s, g, i, io.avoid != 0, io.store != 0, o, (LOG && io.log27) ? lg : nullptr);
        if (io.traj && (io.traj_steps != 1 || t == K - 1)) {
            R* tr = io.traj + (int64_t)(io.traj_steps != 1 ? t : 0) * SIT_N_TRAJ * S + i;
            // streaming stores (st.global.cs): the trajectory is written once and never read back by the kernel
            __stcs(tr, (R)s.N); __stcs(tr + S, (R)s.E); __stcs(tr + 2 * S, s.psi); __stcs(tr + 3 * S, s.u);
            __stcs(tr + 4 * S, s.v); __stcs(tr + 5 * S, s.r); __stcs(tr + 6 * S, s.om); __stcs(tr + 7 * S, (R)s.k);
        }
        if (LOG && io.log27) {
            R* lo = io.log27 + (int64_t)t * SIT_N_LOG * S + i;
#pragma unroll
            for (int c = 0; c < SIT_N_LOG; ++c) __stcs(lo + (int64_t)c * S, lg[c]);
        }
        if (io.cmd) {
            R* co = io.cmd + (int64_t)t * 2 * S + i;
            co[0] = o.rud; co[S] = o.thr;
        }
    }
    store_ship(s, g, i);
    if (io.dist) {
        g.vc_n[i] = s.vcn; g.vc_e[i] = s.vce; g.wind_speed[i] = w_speed; g.wind_dir[i] = w_dir;
    }
}

// Fast variant: nothing is read or written per step (disturbances held for the launch, trajectory row of the last
// step only, no log / command streams), so the K-step loop is pure register arithmetic with compile-time flags.
template <typename R, bool AVOID, bool SINGLE_MODE, bool TRAJ = false>
__global__ void __launch_bounds__(256) fleet_step_fast_kernel(const __grid_constant__ ShipP<R> P, const ShipPtrs<R> g,
                                                              const int K, const FleetIO<R> io) {
    const int64_t li = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t S = g.n_ships, i = io.first + li;
    if (li >= io.count) return;
    Ship<R> s;
    load_ship(s, g, i);
    if (io.dist) {
        s.vcn = io.dist[i]; s.vce = io.dist[S + i];
        const R w_speed = io.dist[2 * S + i], w_dir = io.dist[3 * S + i];
        set_wind(s, w_speed, w_dir);
        g.vc_n[i] = s.vcn; g.vc_e[i] = s.vce; g.wind_speed[i] = w_speed; g.wind_dir[i] = w_dir;
    }
    StepOut<R> o;
    R* tr = io.traj + i;   // running row pointer of the every-step trajectory: one 64-bit add chain per step, no multiplies
#pragma unroll 1
    for (int t = 0; t < K; ++t) {
        ship_step<R, false, SINGLE_MODE>(P, s, g, i, AVOID, true, o, nullptr);
        if (TRAJ) {   // every-step trajectory [K][8][S]: coalesced streaming stores, the only memory traffic of the loop
            R* q = tr;
            __stcs(q, (R)s.N); q += S; __stcs(q, (R)s.E); q += S; __stcs(q, s.psi); q += S; __stcs(q, s.u); q += S;
            __stcs(q, s.v); q += S; __stcs(q, s.r); q += S; __stcs(q, s.om); q += S; __stcs(q, (R)s.k);
            tr = q + S;
        }
    }
    if (!TRAJ && io.traj) {
        tr[0] = (R)s.N; tr[S] = (R)s.E; tr[2 * S] = s.psi; tr[3 * S] = s.u;
        tr[4 * S] = s.v; tr[5 * S] = s.r; tr[6 * S] = s.om; tr[7 * S] = (R)s.k;
    }
    if (io.k8) io.k8[i] = (int8_t)s.k;
    store_ship(s, g, i);
}

// fp32 fast variant, packed: one thread owns the two ships (i0, i0 + 1) and runs the f32x2 step of ship_x2.cuh (SASS FFMA2 /
// FMUL2 / FADD2: one issue slot advances both ships).  State rows are read / written as 64-bit (float2 / double2) accesses,
// so a warp still touches whole 256-byte / 512-byte lines.  An odd tail ship is paired with a copy of itself that is
// never stored.
#ifndef SIT_X2_BLOCK
#define SIT_X2_BLOCK 128
#endif
#ifndef SIT_X2_MINB
#define SIT_X2_MINB 3
#endif
#ifndef SIT_X2_UNROLL
#define SIT_X2_UNROLL 2   /* two steps per loop trip: the loop-carried register moves disappear (2.19e11 -> 2.24e11 ship-steps/s) */
#endif
#define SIT_PRAGMA_(x) _Pragma(#x)
#define SIT_PRAGMA(x) SIT_PRAGMA_(x)
template <bool AVOID, bool SINGLE_MODE, bool TRAJ>
__global__ void __launch_bounds__(SIT_X2_BLOCK, SIT_X2_MINB)
fleet_step_x2_kernel(const __grid_constant__ x2::ShipPX P, const ShipPtrs<float> g, const int K, const FleetIO<float> io) {
    using namespace x2;
    const int64_t li = 2 * ((int64_t)blockIdx.x * blockDim.x + threadIdx.x);
    if (li >= io.count) return;
    const int64_t S = g.n_ships, i0 = io.first + li;
    const bool two = li + 1 < io.count;
    const int64_t i1 = two ? i0 + 1 : i0;
    // 64-bit accesses need an even element offset: state rows start 256-byte aligned (i0 is even), rows of [c][S] blocks
    // (disturbances, trajectory) only when S is even
    const bool vec = two && !(S & 1);
    auto ld = [&](const float* p) -> V { return two ? *reinterpret_cast<const float2*>(p + i0) : make_float2(p[i0], p[i0]); };
    auto st = [&](float* p, const V v) { if (two) *reinterpret_cast<float2*>(p + i0) = v; else p[i0] = v.x; };
    auto ldr = [&](const float* p) -> V { return vec ? *reinterpret_cast<const float2*>(p + i0) : make_float2(p[i0], p[i1]); };
    auto str = [&](float* p, const V v) {
        if (vec) *reinterpret_cast<float2*>(p + i0) = v;
        else { p[i0] = v.x; if (two) p[i1] = v.y; }
    };
    // ship index of lane l, recomputed where it is needed (the rare waypoint-switch path) instead of living in registers
    auto idx = [&](int l) -> int64_t {
        const int64_t j = io.first + 2 * ((int64_t)blockIdx.x * blockDim.x + threadIdx.x);
        return (l == 1 && 2 * ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) + 1 < io.count) ? j + 1 : j;
    };
    Ship2 s;
    s.psi = ld(g.psi); s.u = ld(g.u); s.v = ld(g.v); s.r = ld(g.r); s.om = ld(g.omega);
    s.ei_s = ld(g.ei_speed); s.ei_w = ld(g.ei_shaft); s.ei_h = ld(g.ei_heading); s.e_prev = ld(g.e_prev);
    s.e_ct = ld(g.e_ct); s.e_int = ld(g.e_ct_int);
    s.f_me = ld(g.fuel_me); s.f_el = ld(g.fuel_el);
    // fuel_cons (ship_engine.py:288-292) accumulates the sum of both rates: carried as fuel_me + fuel_el plus whatever offset
    // the caller's state had (zero unless a caller set the three accumulators inconsistently)
    const V f_off = sub(ld(g.fuel), add(s.f_me, s.f_el));
    s.u_sp = ld(g.u_sp);
    V w_speed, w_dir;
    if (io.dist) {   // disturbances held for the launch (ship_model.py:154-156: plain attributes)
        s.vcn = ldr(io.dist); s.vce = ldr(io.dist + S); w_speed = ldr(io.dist + 2 * S); w_dir = ldr(io.dist + 3 * S);
        st(g.vc_n, s.vcn); st(g.vc_e, s.vce); st(g.wind_speed, w_speed); st(g.wind_dir, w_dir);
    } else {
        s.vcn = ld(g.vc_n); s.vce = ld(g.vc_e); w_speed = ld(g.wind_speed); w_dir = ld(g.wind_dir);
    }
    {
        float sb, cb;
        sincos_(w_dir.x, sb, cb); s.wcb.x = w_speed.x * cb; s.wsb.x = w_speed.x * sb;
        sincos_(w_dir.y, sb, cb); s.wcb.y = w_speed.y * cb; s.wsb.y = w_speed.y * sb;
    }
    s.k[0] = g.k[i0]; s.k[1] = g.k[i1]; s.n_wp[0] = g.n_wp[i0]; s.n_wp[1] = g.n_wp[i1];
    double pN[2], pE[2];   // segment start of each ship (float64): live across the loop only in the TRAJ instantiation
    int kk[2];
#pragma unroll
    for (int l = 0; l < 2; ++l) {
        const int64_t i = l == 0 ? i0 : i1;
        const int k = s.k[l];
        kk[l] = k;
        pN[l] = g.wp_n[(int64_t)(k - 1) * S + i]; pE[l] = g.wp_e[(int64_t)(k - 1) * S + i];
        const double cN = g.wp_n[(int64_t)k * S + i], cE = g.wp_e[(int64_t)k * S + i];
        // the relative pair written by the previous launch is used as it is when it still describes the float64 position
        // (bit-identical continuation across launches); anything else (reset, a caller that moved the ship or its route)
        // re-derives it from the float64 position
        const double N = g.north[i], E = g.east[i];
        float x = g.rel_n[i], cx = g.rel_cn[i], y = g.rel_e[i], cy = g.rel_ce[i];
        if (!(abs_pos(pN[l], x, cx) == N)) split_rel(N - pN[l], x, cx);
        if (!(abs_pos(pE[l], y, cy) == E)) split_rel(E - pE[l], y, cy);
        lane(s.X, l) = x; lane(s.CX, l) = cx; lane(s.Y, l) = y; lane(s.CY, l) = cy;
        set_segment2(s, P, pN[l], pE[l], cN, cE, l);
    }
    ModeV MV;
    if (!SINGLE_MODE) {
        const ModeX &a = P.mode[g.mode[i0]], &b = P.mode[g.mode[i1]];
#define MV_(f) MV.f = make_float2(a.f, b.f)
        MV_(avail_total); MV_(avail_me); MV_(avail_el); MV_(cap_tq_me); MV_(cap_tq_el); MV_(add_me); MV_(cap_me);
        MV_(a_el); MV_(c_el); MV_(inv_me_cap); MV_(inv_el_cap); MV_(me_cap_kw);
        MV_(fme_a); MV_(fme_b); MV_(fel_a); MV_(fel_b);
#undef MV_
    }
    float* tr = io.traj + i0;   // running row pointer of the every-step trajectory
    SIT_PRAGMA(unroll SIT_X2_UNROLL)
    for (int t = 0; t < K; ++t) {
        StepOut2 so;
        if (SINGLE_MODE) ship_step2<AVOID ? AVOID_BOTH : AVOID_NONE, true>(P, P.mode[0], s, g, idx, so);
        else ship_step2<AVOID ? AVOID_BOTH : AVOID_NONE, true>(P, MV, s, g, idx, so);
        if (TRAJ) {   // every-step trajectory [K][8][S]: coalesced 64-bit streaming stores, the only memory traffic of the loop
#pragma unroll
            for (int l = 0; l < 2; ++l)
                if (s.k[l] != kk[l]) {   // the ship switched segments: new float64 base
                    kk[l] = s.k[l];
                    pN[l] = g.wp_n[(int64_t)(kk[l] - 1) * S + idx(l)]; pE[l] = g.wp_e[(int64_t)(kk[l] - 1) * S + idx(l)];
                }
            float* q = tr;
            const V vn = make_float2((float)abs_pos(pN[0], s.X.x, s.CX.x), (float)abs_pos(pN[1], s.X.y, s.CX.y));
            const V ve = make_float2((float)abs_pos(pE[0], s.Y.x, s.CY.x), (float)abs_pos(pE[1], s.Y.y, s.CY.y));
            const V vk = make_float2((float)s.k[0], (float)s.k[1]);
            if (vec) {
                __stcs(reinterpret_cast<float2*>(q), vn); q += S; __stcs(reinterpret_cast<float2*>(q), ve); q += S;
                __stcs(reinterpret_cast<float2*>(q), s.psi); q += S; __stcs(reinterpret_cast<float2*>(q), s.u); q += S;
                __stcs(reinterpret_cast<float2*>(q), s.v); q += S; __stcs(reinterpret_cast<float2*>(q), s.r); q += S;
                __stcs(reinterpret_cast<float2*>(q), s.om); q += S; __stcs(reinterpret_cast<float2*>(q), vk);
            } else {
                const int64_t o = two ? 1 : 0;   // the odd tail ship writes its own element twice
                __stcs(q, vn.x); __stcs(q + o, vn.y); q += S; __stcs(q, ve.x); __stcs(q + o, ve.y); q += S;
                __stcs(q, s.psi.x); __stcs(q + o, s.psi.y); q += S; __stcs(q, s.u.x); __stcs(q + o, s.u.y); q += S;
                __stcs(q, s.v.x); __stcs(q + o, s.v.y); q += S; __stcs(q, s.r.x); __stcs(q + o, s.r.y); q += S;
                __stcs(q, s.om.x); __stcs(q + o, s.om.y); q += S; __stcs(q, vk.x); __stcs(q + o, vk.y);
            }
            tr = q + S;
        }
    }
    // ---- write back: float64 position from the (possibly new) segment start, the relative pair beside it ------------------
    double N[2], E[2];
#pragma unroll
    for (int l = 0; l < 2; ++l) {
        const int64_t i = l == 0 ? i0 : i1;
        const double bN = g.wp_n[(int64_t)(s.k[l] - 1) * S + i], bE = g.wp_e[(int64_t)(s.k[l] - 1) * S + i];
        N[l] = abs_pos(bN, lane(s.X, l), lane(s.CX, l)); E[l] = abs_pos(bE, lane(s.Y, l), lane(s.CY, l));
    }
    if (!TRAJ && io.traj) {
        float* t0 = io.traj;
        str(t0, make_float2((float)N[0], (float)N[1])); str(t0 + S, make_float2((float)E[0], (float)E[1]));
        str(t0 + 2 * S, s.psi); str(t0 + 3 * S, s.u); str(t0 + 4 * S, s.v); str(t0 + 5 * S, s.r); str(t0 + 6 * S, s.om);
        str(t0 + 7 * S, make_float2((float)s.k[0], (float)s.k[1]));
    }
    if (io.k8) { io.k8[i0] = (int8_t)s.k[0]; if (two) io.k8[i1] = (int8_t)s.k[1]; }
    g.north[i0] = N[0]; g.east[i0] = E[0];
    if (two) { g.north[i1] = N[1]; g.east[i1] = E[1]; }
    st(g.rel_n, s.X); st(g.rel_cn, s.CX); st(g.rel_e, s.Y); st(g.rel_ce, s.CY);
    st(g.psi, s.psi); st(g.u, s.u); st(g.v, s.v); st(g.r, s.r); st(g.omega, s.om);
    st(g.ei_speed, s.ei_s); st(g.ei_shaft, s.ei_w); st(g.ei_heading, s.ei_h); st(g.e_prev, s.e_prev);
    st(g.e_ct, vabs(s.e_ct)); st(g.e_ct_int, s.e_int);
    st(g.fuel_me, s.f_me); st(g.fuel_el, s.f_el);
    st(g.fuel, add(add(s.f_me, s.f_el), f_off));
    g.k[i0] = s.k[0]; g.n_wp[i0] = s.n_wp[0]; g.steps[i0] += K;   // utils.py:42-48 next_time, once per stored step
    if (two) { g.k[i1] = s.k[1]; g.n_wp[i1] = s.n_wp[1]; g.steps[i1] += K; }
}

// =================================================================================================
// K3: reset / capture_init
// =================================================================================================
template <typename R>
__global__ void capture_init_kernel(const ShipPtrs<R> g, const EnvPtrs<R> ge, const double sampling_frequency,
                                    const int ships_per_env) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t S = g.n_ships;
    if (i >= S) return;
    g.init_north[i] = g.north[i]; g.init_east[i] = g.east[i];
    g.init_psi[i] = g.psi[i]; g.init_u[i] = g.u[i]; g.init_v[i] = g.v[i]; g.init_r[i] = g.r[i];
    g.init_vc_n[i] = g.vc_n[i]; g.init_vc_e[i] = g.vc_e[i];
    g.init_wind_speed[i] = g.wind_speed[i]; g.init_wind_dir[i] = g.wind_dir[i];
    g.init_u_sp[i] = g.u_sp[i]; g.init_omega[i] = g.omega[i]; g.init_ei_shaft[i] = g.ei_shaft[i];
    const int32_t n = g.n_wp[i];
    g.n_wp0[i] = n;
    for (int j = 0; j < g.max_wp; ++j) {
        g.wp0_n[(int64_t)j * S + i] = g.wp_n[(int64_t)j * S + i];
        g.wp0_e[(int64_t)j * S + i] = g.wp_e[(int64_t)j * S + i];
    }
    if (ships_per_env == 2 && (i & 1)) {
        // reward_function_params, MSRL_env_ex.py:122-127: AB = first -> last waypoint of the obstacle route
        const double dn = (double)g.wp_n[(int64_t)(n - 1) * S + i] - (double)g.wp_n[i];
        const double de = (double)g.wp_e[(int64_t)(n - 1) * S + i] - (double)g.wp_e[i];
        ge.ab_seg[i >> 1] = (R)(sqrt(dn * dn + de * de) / sampling_frequency);
        ge.ab_alpha[i >> 1] = (R)atan2(de, dn);
    }
}

template <typename R>
__global__ void reset_kernel(const ShipPtrs<R> g, const EnvPtrs<R> ge, const uint8_t* mask, const int hard,
                             const int ships_per_env) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= g.n_ships) return;
    const int64_t env = ships_per_env == 2 ? (i >> 1) : i;
    if (mask && !mask[env]) return;
    reset_ship_arrays(g, i, hard != 0);
    if (ships_per_env == 2 && (i & 1)) {
        // MSRL_env_ex.py:168-183: stop flags, distance counters; the stored-position history is
        // cleared by ShipModelAST.reset (ship_model.py:708-712)
        ge.flags[env] = (ge.flags[env] & (F_IW_BAD | F_IW_VALID)) | F_FRESH;
        ge.eps_dist[env] = 0.0; ge.samp_dist[env] = 0.0;
        ge.prev_n[env] = 0.0; ge.prev_e[env] = 0.0;
        ge.ep_return[env] = (R)0; ge.ep_steps[env] = 0;
        ge.n_dec[env] = 0;
    }
}

// =================================================================================================
// K2: two-ship AST env.  Lane pair (2j, 2j+1) = (test ship, obstacle ship) of one env.
// =================================================================================================
template <typename R>
__device__ __forceinline__ void reset_ship_regs(Ship<R>& s, const ShipPtrs<R>& g, int64_t i, bool hard) {
    const int64_t S = g.n_ships;
    s.N = g.init_north[i]; s.E = g.init_east[i];
    s.psi = g.init_psi[i]; s.u = g.init_u[i]; s.v = g.init_v[i]; s.r = g.init_r[i];
    s.vcn = g.init_vc_n[i]; s.vce = g.init_vc_e[i];
    const R ws = g.init_wind_speed[i], wd = g.init_wind_dir[i];
    set_wind(s, ws, wd);
    g.vc_n[i] = s.vcn; g.vc_e[i] = s.vce; g.wind_speed[i] = ws; g.wind_dir[i] = wd;
    s.u_sp = g.init_u_sp[i]; g.u_sp[i] = s.u_sp;
    s.steps = 0; s.k = 1; s.e_ct = (R)0; s.e_int = (R)0;
    s.n_wp = g.n_wp0[i];
    for (int j = 0; j < g.max_wp; ++j) {
        g.wp_n[(int64_t)j * S + i] = g.wp0_n[(int64_t)j * S + i];
        g.wp_e[(int64_t)j * S + i] = g.wp0_e[(int64_t)j * S + i];
    }
    load_segment(s, g, i);
    if (hard) {   // what the reference's reset() leaves alone (quirk Q3): a fresh construction
        s.om = g.init_omega[i];
        s.ei_s = (R)0; s.ei_w = g.init_ei_shaft[i]; s.ei_h = (R)0; s.e_prev = (R)0;
        s.f_me = (R)0; s.f_el = (R)0; s.f_tot = (R)0;
    }
}

#ifndef SIT_ENV_MINB_F32
#define SIT_ENV_MINB_F32 4
#endif
// ROLL: device-resident rollout (sit_rollout_step_k): envs waiting for a policy action do not step, the decision
// step's transition is completed in the replay ring, finished episodes restart in-kernel and new decision events are
// detected after every step.  Lanes of one env always take the same branches, so the pair exchanges use a two-lane mask.
template <typename R, bool SAME, bool ROLL, bool PARTS = false>
__global__ void __launch_bounds__(128, sizeof(R) == 4 ? SIT_ENV_MINB_F32 : 2)
env_step_kernel(const __grid_constant__ EnvKArgs<R> A, const ShipPtrs<R> g, const EnvPtrs<R> ge, const int K,
                const EnvIO<R> io, const RollArgs ro) {
    const unsigned FULL = 0xffffffffu;
    __shared__ EdgeRec<R> s_edges[SIT_MAX_EDGES];
    stage_edges(A.M, s_edges);
    const int n_edges = A.M.n_edges;
    const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x + 2 * ro.env_first;   // env sub-range of this launch
    const int64_t S = g.n_ships, B = ge.n_envs;
    const bool active = tid < 2 * (ro.env_first + ro.env_count);
    const int64_t i = active ? tid : (tid & 1);   // idle lanes of the last warp shadow env 0 and never store
    const int role = (int)(i & 1);
    const int64_t env = i >> 1;
    const bool is_obs = role == SIT_ROLE_OBS;
    const ShipP<R>& P = SAME ? A.P[0] : A.P[role];
    const EnvP<R>& EP = A.EP;
    const int lane = threadIdx.x & 31;
    // The exchanges between the two ships of an env sit where EVERY lane of the warp arrives (in the rollout too: an env that
    // waits for an action skips the phases around them, not the exchanges), so they use the full mask: no per-pair WARPSYNC
    // on a computed mask, and the warp reconverges every step (env leg 1.34e10 -> 1.43e10 env-steps/s, rollout 3.13e7 -> 3.26e7).

    Ship<R> s;
    load_ship(s, g, i);
    // route end point (constant: insertions go before it, controllers.py:298-303)
    R endN = g.wp_n[(int64_t)(s.n_wp - 1) * S + i], endE = g.wp_e[(int64_t)(s.n_wp - 1) * S + i];
    int32_t flags = ge.flags[env];
    bool stop = (flags & (is_obs ? F_STOP_OBS : F_STOP_TEST)) != 0;
    int nstored = (flags & F_NSTORED_MASK) >> F_NSTORED_SHIFT;
    bool fresh = (flags & F_FRESH) != 0;
    bool iw_bad = (flags & F_IW_BAD) != 0, iw_valid = (flags & F_IW_VALID) != 0;
    bool pending = ROLL && (flags & F_PENDING) != 0, have_act = ROLL && (flags & F_HAVE_ACT) != 0, tr_open = false;
    double prev_n = ge.prev_n[env], prev_e = ge.prev_e[env];
    double eps_dist = ge.eps_dist[env], samp_dist = ge.samp_dist[env];
    R ab_seg = ge.ab_seg[env], act_n = ge.act_n[env], act_e = ge.act_e[env];
    R ep_return = ge.ep_return[env];
    int32_t ep_steps = ge.ep_steps[env];
    int32_t n_dec = ROLL ? ge.n_dec[env] : 0;
    R rsum = (R)0;
    bool any_done = false;
    unsigned my_cnt = 0;   // lane b < 12 counts status bit b of this warp's envs; lane 12 counts done steps
    unsigned n_live = 0;   // env-steps actually executed by this lane (ROLL: pending envs do not step)
    unsigned n_eps = 0;    // ROLL: episodes this env ended in the launch (terminal state; num_steps_episode when truncating)
    unsigned n_closed = 0; // ROLL: replay rows this env completed in the launch
    double draw_t = (PARTS && io.snap) ? ge.draw_t[env] : 0.0;   // ship_draw timer: compiled into the PARTS instantiation only
    MapCache mc;           // nearest-ring cache, carried across launches in the state buffer
    mc.idx = (uint32_t)g.mc_idx[i]; mc.p0n = g.mc_n[i]; mc.p0e = g.mc_e[i]; mc.b2 = g.mc_b2[i];

    for (int t = 0; t < K; ++t) {
        // ---- inputs: converted_action and SAC_update (MSRL_env_ex.py:402-405) -----------------------
        bool su = false, live = true;
        if (ROLL) {
            if (pending) {
                if (have_act) { su = true; pending = false; have_act = false; tr_open = true; }   // action from sit_rollout_policy
                else live = false;
            }
            if (!__any_sync(FULL, live && active)) break;
        } else if (io.action && (t == 0 || io.act_steps > 1)) {   // an action held for the launch is read (and validated) once
            const int64_t ta = io.act_steps > 1 ? t : 0;
            const R a_n = io.action[(ta * 2 + 0) * B + env], a_e = io.action[(ta * 2 + 1) * B + env];
            if (!iw_valid || a_n != act_n || a_e != act_e) {
                // is_route_outside_horizon :517-542 or is_route_inside_obstacles :544-552 of the
                // action, cached until the action changes
                act_n = a_n; act_e = a_e;
                iw_bad = (a_n < EP.min_n || a_n > EP.max_n || a_e < EP.min_e || a_e > EP.max_e) ||
                         map_contains(s_edges, n_edges, a_n, a_e);
                iw_valid = true;
            }
            if (io.sac_update && (io.act_steps > 1 || t == 0)) su = io.sac_update[ta * B + env] != 0;
        }
        int status = 0;
        bool done = false;
        StepOut<R> o;
        o.rpm = (R)0; o.power_me_kw = (R)0; o.avail_me_kw = (R)0; o.rud = (R)0; o.thr = (R)0;
        double ship_d2 = 0.0;
        if (live) {   // ---- phase A: the two ship steps ------------------------------------------------------
        n_live += 1;
        const bool init = (ROLL || io.init < 0) ? fresh : (t == 0 && io.init != 0);
        const bool frozen = is_obs && stop;
        const double pre_n = s.N, pre_e = s.E;
        if (!frozen) {
            if (is_obs && su) {
                // obs_step :324-338 -> update_route controllers.py:298-303: list.insert(-1, wp)
                if (s.n_wp < g.max_wp) {
                    g.wp_n[(int64_t)s.n_wp * S + i] = endN; g.wp_e[(int64_t)s.n_wp * S + i] = endE;
                    g.wp_n[(int64_t)(s.n_wp - 1) * S + i] = act_n; g.wp_e[(int64_t)(s.n_wp - 1) * S + i] = act_e;
                    if (s.k == s.n_wp - 1) {   // the segment in use ended at the old last point
                        s.cN = act_n; s.cE = act_e;
                        s.alpha = atan2_((R)(s.cE - s.pE), (R)(s.cN - s.pN));
                        sincos_(s.alpha, s.sa, s.ca);
                    }
                    s.n_wp += 1;
                }
                samp_dist = 0.0;
            }
            // test_step :217-283 (avoidance branch always taken) | obs_step :346-373
            ship_step<R, false>(P, s, g, i, !is_obs, true, o, nullptr);
        } else {
            s.steps += 2;   // frozen ship: next_time twice, :294,307
        }
        if (is_obs) {
            if (frozen) {
                if (nstored >= 1) nstored = 2;   // store_last_simulation_data repeats the last row :292
            } else {
                if (!init && nstored >= 1) {      // :391-395, last two *stored* (pre-step) positions
                    const double dn = pre_n - prev_n, de = pre_e - prev_e;
                    const double d = sqrt(__dadd_rn(__dmul_rn(dn, dn), __dmul_rn(de, de)));
                    eps_dist += d; samp_dist += d;
                }
                prev_n = pre_n; prev_e = pre_e;
                nstored = nstored < 2 ? nstored + 1 : 2;
            }
        }

        }
        // ---- exchange post-step positions inside the lane pair (whole warp: an env that waits for an action exchanges its
        // held positions, nothing reads the result) ----------------------------------------------------------------
        {
            const double oth_n = __shfl_xor_sync(FULL, s.N, 1), oth_e = __shfl_xor_sync(FULL, s.E, 1);
            const double n_t = is_obs ? oth_n : s.N, e_t = is_obs ? oth_e : s.E;
            const double n_o = is_obs ? s.N : oth_n, e_o = is_obs ? s.E : oth_e;
            const double ddn = n_t - n_o, dde = e_t - e_o;
            ship_d2 = __dadd_rn(__dmul_rn(ddn, ddn), __dmul_rn(dde, dde));
        }

        // ---- phase B (whole warp): distance to the nearest ring + hull containment, nearest-ring cache -------
        const R pn = (R)s.N, pe = (R)s.E;
        R d2map = (R)3.0e38;
        bool inside = false;
        if (n_edges > 0) {
            const float ddn_ = (float)pn - mc.p0n, dde_ = (float)pe - mc.p0e;
            unsigned nb = __ballot_sync(FULL, live && !(ddn_ * ddn_ + dde_ * dde_ < fabsf(mc.b2)));
            while (nb) {   // budget spent: the whole warp rescans that lane's position
                const int L = __ffs(nb) - 1;
                nb &= nb - 1;
                const MapCache c = coop_scan(s_edges, n_edges, __shfl_sync(FULL, pn, L), __shfl_sync(FULL, pe, L), lane);
                if (lane == L) mc = c;
            }
            d2map = cached_dist2(s_edges, mc.idx, pn, pe);
            inside = (__float_as_uint(mc.b2) >> 31) != 0u;   // far from every ring all four corners share the centre's side
            unsigned hb = __ballot_sync(FULL, live && d2map <= EP.near2);
            while (hb) {   // is_pos_inside_obstacles, MSRL_env_ex.py:490-515: the four corners of the +-L/2 square
                const int L = __ffs(hb) - 1;
                hb &= hb - 1;
                const bool in = coop_hull_inside(s_edges, n_edges, __shfl_sync(FULL, pn, L), __shfl_sync(FULL, pe, L),
                                                 __shfl_sync(FULL, EP.half_len[role], L), lane);
                if (lane == L) inside = in;
            }
        }

        int st = 0; bool dn_ = false; R part = (R)0; bool collide = false; R ect = (R)0, dmap = (R)0; bool stop_in = false, dec = false;
        if (live) {   // ---- phase C: rewards, terminal checks ------------------------------------------------------------
        dmap = sqrt_(d2map);
        const bool outside = pn < EP.lo_n[role] || pn > EP.hi_n[role] || pe < EP.lo_e[role] || pe > EP.hi_e[role];
        const double dnE = s.N - (double)endN, deE = s.E - (double)endE;
        const bool arrived = __dadd_rn(__dmul_rn(dnE, dnE), __dmul_rn(deE, deE)) <= EP.arrive2;   // decided in double in both precisions
        ect = s.e_ct;

        // ---- per-ship reward terms and terminal checks (sequential stop_flag gating) -------------------
        stop_in = PARTS && stop;      // the obstacle ship's non-terminal terms see the flag from before this step
        if (!is_obs) {
            // test_ship_non_terminal_state_reward :628-664
            const R ntt = abs_(ect) * EP.inv_e_tol + ((R)1 - dmap * EP.inv_max_north) * (R)0.01;
            R rt = (R)0;
            // test_ship_terminal_state_reward :734-809
            if (arrived) { stop = true; dn_ = true; st |= SIT_ST_TEST_ENDPOINT; }
            if (outside) { stop = true; dn_ = true; st |= SIT_ST_TEST_HORIZON; }
            if (inside) { if (!stop) rt += (R)1000; stop = true; dn_ = true; st |= SIT_ST_TEST_TERRAIN; }
            if (abs_(o.rpm) > EP.rpm_max) { if (!stop) rt += (R)1000; stop = true; dn_ = true; st |= SIT_ST_TEST_MECH; }
            if (abs_(ect) > EP.e_tol) { if (!stop) rt += (R)1000; stop = true; dn_ = true; st |= SIT_ST_TEST_NAV; }
            if (o.power_me_kw > o.avail_me_kw) { if (!stop) rt += (R)1000; stop = true; dn_ = true; st |= SIT_ST_TEST_BLACKOUT; }
            part = ntt + rt;
        } else {
            // obs_ship_non_terminal_state_reward :666-710 (stop flag from before this step's checks)
            R nto = (R)0;
            if (!stop) nto = (R)0.1 + (-(abs_(ect) * EP.inv_e_tol) * (R)0.01) + (-((R)1 - dmap * EP.inv_max_north) * (R)0.01);
            R ro_ = (R)0;
            // obs_ship_terminal_state_reward :811-881
            if (arrived) { stop = true; st |= SIT_ST_OBS_ENDPOINT; }                       // no done
            if (outside) { stop = true; dn_ = true; st |= SIT_ST_OBS_HORIZON; }
            if (inside) { if (!stop) ro_ -= (R)1000; dn_ = true; st |= SIT_ST_OBS_TERRAIN; } // no stop flag
            if (iw_bad) { if (!stop) ro_ -= (R)1000; stop = true; dn_ = true; st |= SIT_ST_OBS_IW; }
            if (abs_(ect) > EP.e_tol || samp_dist > (double)ab_seg * EP.theta) {
                if (!stop) ro_ -= (R)1000; stop = true; dn_ = true; st |= SIT_ST_OBS_NAV;
            }
            // shared_non_terminal_state_reward :712-731 (sees the stop flag after the checks above)
            R snt = (R)0;
            // (the distance itself in the arithmetic type: only the collision DECISION below needs the float64 square)
            if (!stop) snt = ((R)1 - sqrt_((R)ship_d2) * EP.inv_max_north) * (R)0.001;
            part = nto;      // assembled below as ((((ntt+rt) + nto) + ro) + snt) + rs
            // stash ro, snt in o.* free slots
            o.rud = ro_; o.thr = snt;
        }
        // shared_terminal_state_reward :883-904, is_ship_collision :584-603
        collide = ship_d2 < EP.collide2;
        if (collide) { stop = true; dn_ = true; st |= SIT_ST_COLLISION; }
        }
        // partial reward, status bits and done flag of the other ship of the env (whole warp, see above)
        const R part_t = __shfl_xor_sync(FULL, part, 1);
        const int st_o = __shfl_xor_sync(FULL, st, 1);
        const bool dn_o = __shfl_xor_sync(FULL, (int)dn_, 1) != 0;
        if (live) {   // ---- phase C, continued: outputs, replay rows, episode restart ----------------------------------------
        done = dn_ || dn_o;
        status = st | st_o;
        R reward = (R)0;
        if (is_obs) reward = ((((part_t + part) + o.rud) + o.thr)) + (collide ? (R)2000 : (R)0);   // :970-972
        rsum += reward; ep_return += reward; ep_steps += 1;
        any_done |= done;
        fresh = false;
        // MSRL_env_ex.py:416-421: `if time_since_last_ship_drawing > 30: snap both ships; timer = 0`, then `timer += dt`
        // (after both ship steps, before the observation is assembled: the outline is drawn on this step's obs pose)
        bool snapped = false;
        if (PARTS && io.snap) {
            if (draw_t > 30.0) { snapped = true; draw_t = 0.0; }
            draw_t += io.draw_dt;
        }

        // ---- outputs: coalesced SoA rows [step][channel][env] -------------------------------------------
        if (active && (io.rec_steps > 1 || t == K - 1)) {
            const int64_t rec = io.rec_steps > 1 ? t : 0;
            if (io.obs) {
                R* ob = io.obs + rec * SIT_N_OBS * B + env;
                if (!is_obs) {   // :424-435: post-step pose, pre-step stored rpm / e_ct / power
                    ob[0] = pn; ob[B] = pe; ob[2 * B] = s.psi; ob[3 * B] = o.rpm; ob[4 * B] = ect; ob[5 * B] = o.power_me_kw;
                } else {
                    ob[6 * B] = pn; ob[7 * B] = pe; ob[8 * B] = s.psi; ob[9 * B] = ect;
                }
            }
            if (PARTS && io.rew_parts) {   // reward_results increments: compiled into the PARTS instantiation only
                R* rp = io.rew_parts + rec * SIT_N_REWARD_PARTS * B + env;
                const R r_ect = abs_(ect) * EP.inv_e_tol, r_col = ((R)1 - dmap * EP.inv_max_north) * (R)0.01;
                if (!is_obs) {
                    rp[0] = r_ect; rp[B] = r_col; rp[2 * B] = r_ect + r_col;
                } else {
                    const bool on = !stop_in;
                    rp[3 * B] = on ? (R)0.1 : (R)0; rp[4 * B] = on ? -r_ect * (R)0.01 : (R)0; rp[5 * B] = on ? -r_col : (R)0;
                    rp[6 * B] = on ? (R)0.1 + (-r_ect * (R)0.01) + (-r_col) : (R)0; rp[7 * B] = o.thr;
                }
            }
            if (is_obs) {
                if (io.reward) io.reward[rec * B + env] = reward;
                if (io.done) io.done[rec * B + env] = done ? 1 : 0;
                if (io.status) io.status[rec * B + env] = status;
                if (PARTS && io.snap) io.snap[rec * B + env] = snapped ? 1 : 0;
            }
        }
        bool episode_end = done && io.auto_reset != 0;
        if (ROLL) {
            // main_ast.py:387: mask = 1 if episode_steps == num_steps_episode else float(not done); the reference's
            // episode_steps counts init_step as 1, so it equals ep_steps + 1 here
            // (== : only the row that closes AT that step is masked; the reference's loop is `while not done`, :322, so the
            // episode goes on unless the caller opted into truncation -- then >= and == coincide)
            const bool timeup = ro.max_episode_steps > 0 && ep_steps + 1 == ro.max_episode_steps;
            if (tr_open) {   // close the replay row opened by sit_rollout_policy: (reward, next_state, mask), :395-396
                if (active) {
                    float* row = ro.replay + (int64_t)ge.tr_slot[env] * SIT_TR_FLOATS;
                    if (!is_obs) {
                        row[12] = (float)pn; row[13] = (float)pe; row[14] = (float)s.psi; row[15] = (float)o.rpm;
                        row[16] = (float)ect; row[17] = (float)o.power_me_kw;
                    } else {
                        row[18] = (float)pn; row[19] = (float)pe; row[20] = (float)s.psi; row[21] = (float)ect;
                        row[11] = (float)reward; row[22] = timeup ? 1.0f : (done ? 0.0f : 1.0f);
                    }
                }
                tr_open = false;
                if (active && is_obs) n_closed += 1;
            }
            episode_end = done || (timeup && ro.truncate);
        }
        // ---- in-kernel reset() + init_step() right after a terminal step ---------------------------------
        if (episode_end) {
            if (ROLL && active && is_obs) n_eps += 1;
            reset_ship_regs(s, g, i, io.auto_reset == 2);   // MSRL_env_ex.py:145-186 (omega, integrators, fuel kept unless hard)
            endN = g.wp_n[(int64_t)(s.n_wp - 1) * S + i]; endE = g.wp_e[(int64_t)(s.n_wp - 1) * S + i];
            stop = false; nstored = 0; fresh = true;
            eps_dist = 0.0; samp_dist = 0.0; prev_n = 0.0; prev_e = 0.0;
            ep_return = (R)0; ep_steps = 0;
            StepOut<R> o2;
            ship_step<R, false>(P, s, g, i, false, false, o2, nullptr);   // init_step :188-215
            if (ROLL) {
                // the new episode's first decision acts on reset()'s float32 initial state (MSRL_env_ex.py:86-90,186)
                pending = true; n_dec = 0;
                if (active) {
                    R* co = ge.cur_obs + env;
                    if (!is_obs) {
                        co[0] = (R)(float)g.init_north[i]; co[B] = (R)(float)g.init_east[i]; co[2 * B] = (R)(float)g.init_psi[i];
                        co[3 * B] = (R)0; co[4 * B] = (R)0; co[5 * B] = (R)0;
                    } else {
                        co[6 * B] = (R)(float)g.init_north[i]; co[7 * B] = (R)(float)g.init_east[i];
                        co[8 * B] = (R)(float)g.init_psi[i]; co[9 * B] = (R)0;
                    }
                }
            }
        } else if (ROLL) {
            // decision event: the obstacle ship is inside the radius of acceptance of its current intermediate
            // waypoint (controllers.py:317-325; the same squared test as the waypoint switch, LOS_guidance.py:96)
            if (is_obs && !stop && n_dec < ro.sampling_frequency && s.k < s.n_wp - 1) {
                const double dn = s.cN - s.N, de = s.cE - s.E;
                dec = __dadd_rn(__dmul_rn(dn, dn), __dmul_rn(de, de)) <= P.ra2;
            }
        }
        }   // live
        if (ROLL) {   // the obstacle lane's decision flag reaches the test lane (whole warp; false for envs that did not step)
            dec = __shfl_sync(FULL, (int)dec, lane | 1) != 0;
            if (dec) {
                pending = true;
                if (active) {
                    R* co = ge.cur_obs + env;
                    if (!is_obs) {
                        co[0] = pn; co[B] = pe; co[2 * B] = s.psi; co[3 * B] = o.rpm; co[4 * B] = ect; co[5 * B] = o.power_me_kw;
                    } else {
                        co[6 * B] = pn; co[7 * B] = pe; co[8 * B] = s.psi; co[9 * B] = ect;
                    }
                }
            }
        }
        // ---- warp-level reduction of termination / collision flags --------------------------------------
        if (io.stats) {
            const bool cnt_lane = active && is_obs && live;
            if (__any_sync(FULL, cnt_lane && status != 0)) {
#pragma unroll
                for (int b = 0; b < 12; ++b) {
                    const unsigned bal = __ballot_sync(FULL, cnt_lane && ((status >> b) & 1));
                    if (lane == b) my_cnt += __popc(bal);
                }
                const unsigned bal = __ballot_sync(FULL, cnt_lane && done);
                if (lane == 12) my_cnt += __popc(bal);
            }
        }
    }

    // ---- write back ------------------------------------------------------------------------------------------
    const bool stop_t = __shfl_xor_sync(FULL, (int)stop, 1) != 0;   // the test lane's flag, seen by the obs lane
    if (active) {
        store_ship(s, g, i);
        g.mc_idx[i] = (int32_t)mc.idx; g.mc_n[i] = mc.p0n; g.mc_e[i] = mc.p0e; g.mc_b2[i] = mc.b2;
        if (is_obs) {
            int32_t f = (stop_t ? F_STOP_TEST : 0) | (stop ? F_STOP_OBS : 0) | (nstored << F_NSTORED_SHIFT) |
                        (iw_bad ? F_IW_BAD : 0) | (iw_valid ? F_IW_VALID : 0) | (fresh ? F_FRESH : 0) |
                        (pending ? F_PENDING : 0) | (have_act ? F_HAVE_ACT : 0);
            ge.flags[env] = f;
            ge.prev_n[env] = prev_n; ge.prev_e[env] = prev_e;
            ge.eps_dist[env] = eps_dist; ge.samp_dist[env] = samp_dist;
            ge.act_n[env] = act_n; ge.act_e[env] = act_e;
            ge.ep_return[env] = ep_return; ge.ep_steps[env] = ep_steps;
            if (ROLL) ge.n_dec[env] = n_dec;
            if (PARTS && io.snap) ge.draw_t[env] = draw_t;
            if (io.reward_sum) io.reward_sum[env] = rsum;
            if (io.done_any) io.done_any[env] = any_done ? 1 : 0;
        }
    }
    if (io.stats) {
        if (lane <= 12 && my_cnt) atomicAdd(&io.stats[lane], (unsigned long long)my_cnt);
        unsigned steps = (active && is_obs) ? n_live : 0u;   // env-steps executed by this warp
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) steps += __shfl_xor_sync(FULL, steps, d);
        if (lane == 13 && steps) atomicAdd(&io.stats[13], (unsigned long long)steps);
    }
    if (ROLL) {
        unsigned eps = n_eps;
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) eps += __shfl_xor_sync(FULL, eps, d);
        if (lane == 14 && eps) atomicAdd(&ro.counters[SIT_RC_EPISODES], (unsigned long long)eps);
        unsigned cl = n_closed;
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) cl += __shfl_xor_sync(FULL, cl, d);
        if (lane == 15 && cl) atomicAdd(&ro.counters[SIT_RC_CLOSED], (unsigned long long)cl);
    }
}

// init_step, MSRL_env_ex.py:188-215: both ships advance one dt without storing, without next_time and
// without the avoidance branch
template <typename R, bool SAME>
__global__ void __launch_bounds__(256) env_init_step_kernel(const __grid_constant__ ShipP<R> P0,
                                                            const __grid_constant__ ShipP<R> P1, const ShipPtrs<R> g,
                                                            const uint8_t* mask) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= g.n_ships) return;
    if (mask && !mask[i >> 1]) return;
    const ShipP<R>& P = (SAME || !(i & 1)) ? P0 : P1;
    Ship<R> s;
    load_ship(s, g, i);
    StepOut<R> o;
    ship_step<R, false>(P, s, g, i, false, false, o, nullptr);
    store_ship(s, g, i);
}

#include "env_x2.cuh"

// =================================================================================================
// K4: rollout policy -- for every pending env: tanh-Gaussian / uniform policy on its held observation, scoping
// angle -> intermediate waypoint, validity resampling, replay row head.  One CTA scans 256 envs, compacts the
// pending ones (deterministic order) and runs the MLP in tiles of 16 envs: thread j owns hidden unit j, the
// 16 activations of each unit sit in shared memory as one 64-byte row read back with broadcast 128-bit loads.
// =================================================================================================
__device__ __forceinline__ void philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1,
                                              uint32_t out[4]) {
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
        const uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
        const uint32_t n0 = hi1 ^ c1 ^ k0, n1 = lo1, n2 = hi0 ^ c3 ^ k1, n3 = lo0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

struct PolicyArgs { int32_t mode, hidden; const float *w0, *b0, *w1, *b1, *w2, *b2, *obs_scale; };
// scratch of one sit_rollout_policy call (device, owned by the handle): compact list of the pending envs
struct PolScratch {   // env / retry [B], obs [B][10], out [B][2]
    unsigned int *n_pending, *n_retry, *ticket; unsigned long long* slot_base;
    int32_t *env, *retry, *cta_cnt; float *obs, *out; unsigned char* w1p;
};

// ---- K4a: compact the pending envs, open their replay rows.  DETERMINISTIC ORDER: pass 1 counts the pending envs of every
// 256-env CTA; the last CTA to finish (ticket) turns the counts into exclusive offsets in blockIdx order and reserves the
// replay slots [TRANSITIONS, TRANSITIONS + total); pass 2 places env e at offset[cta] + rank inside the CTA.  The compact
// list and the ring rows are therefore in ascending env order whatever the CTA scheduling: two runs with one seed fill
// the ring identically (the learner draws batches by row index).  With several env groups in flight (SIT_ROLLOUT_GROUPS > 1)
// each group's block of rows is ordered, the order of the blocks is not.
template <typename R>
__global__ void __launch_bounds__(256) rollout_count_kernel(const EnvPtrs<R> ge, const RollArgs ro, const PolScratch sc) {
    __shared__ int s_warp_cnt[8];
    __shared__ bool s_last;
    __shared__ unsigned int s_scan[256];
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int64_t env = ro.env_first + (int64_t)blockIdx.x * 256 + tid;
    const int32_t fl = env < ro.env_first + ro.env_count ? ge.flags[env] : 0;
    const bool pend = (fl & F_PENDING) && !(fl & F_HAVE_ACT);
    const unsigned bal = __ballot_sync(0xffffffffu, pend);
    if (lane == 0) s_warp_cnt[wid] = __popc(bal);
    __syncthreads();
    if (tid == 0) {
        int total = 0;
        for (int w = 0; w < 8; ++w) total += s_warp_cnt[w];
        sc.cta_cnt[blockIdx.x] = total;
        __threadfence();
        s_last = atomicAdd(sc.ticket, 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    // exclusive scan of cta_cnt[0 .. gridDim.x) in place: thread t owns a contiguous run of `per` entries
    const int nb = (int)gridDim.x, per = (nb + 255) / 256;
    const int lo = min(tid * per, nb), hi = min(lo + per, nb);
    unsigned int mine = 0;
    for (int j = lo; j < hi; ++j) mine += (unsigned int)__ldcg(&sc.cta_cnt[j]);
    s_scan[tid] = mine;
    __syncthreads();
    for (int d = 1; d < 256; d <<= 1) {
        const unsigned int v = tid >= d ? s_scan[tid - d] : 0u;
        __syncthreads();
        s_scan[tid] += v;
        __syncthreads();
    }
    unsigned int run = s_scan[tid] - mine;
    for (int j = lo; j < hi; ++j) { const unsigned int c = (unsigned int)__ldcg(&sc.cta_cnt[j]); sc.cta_cnt[j] = (int32_t)run; run += c; }
    if (tid == 255) {
        const unsigned int total = s_scan[255];
        *sc.n_pending = total;
        *sc.slot_base = atomicAdd(&ro.counters[SIT_RC_TRANSITIONS], (unsigned long long)total);
        atomicAdd(&ro.counters[SIT_RC_DECISIONS], (unsigned long long)total);
        *sc.ticket = 0u;
    }
}

template <typename R>
__global__ void __launch_bounds__(256) rollout_gather_kernel(const EnvPtrs<R> ge, const RollArgs ro, const PolScratch sc,
                                                             const float* __restrict__ obs_scale) {
    __shared__ int s_warp_cnt[8];
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int64_t B = ge.n_envs;
    const int64_t env = ro.env_first + (int64_t)blockIdx.x * 256 + tid;
    const int32_t fl = env < ro.env_first + ro.env_count ? ge.flags[env] : 0;
    const bool pend = (fl & F_PENDING) && !(fl & F_HAVE_ACT);
    const unsigned bal = __ballot_sync(0xffffffffu, pend);
    if (lane == 0) s_warp_cnt[wid] = __popc(bal);
    __syncthreads();
    if (!pend) return;
    int off = 0;
    for (int w = 0; w < wid; ++w) off += s_warp_cnt[w];
    const unsigned int r = (unsigned int)sc.cta_cnt[blockIdx.x] + off + __popc(bal & ((1u << lane) - 1u));
    sc.env[r] = (int32_t)env;
    const long long slot = (long long)((*sc.slot_base + r) % (unsigned long long)ro.capacity);
    ge.tr_slot[env] = (int32_t)slot;
    float* row = ro.replay + slot * SIT_TR_FLOATS;
    // held observation (MSRL_env_ex.py:424-435 layout) as float32, like the reference's policy input and replay state; the
    // policy network sees it scaled per channel (SitPolicy.obs_scale), the replay row keeps the raw values
#pragma unroll
    for (int c = 0; c < SIT_N_OBS; ++c) {
        const float v = (float)ge.cur_obs[(int64_t)c * B + env];
        sc.obs[(int64_t)r * SIT_N_OBS + c] = obs_scale ? v * obs_scale[c] : v;
        row[c] = v;
    }
    row[23] = __int_as_float((int32_t)(ro.env_id_offset + env));
}

// ---- K4b: policy MLP on the compact list, tiles of 16 rows: thread j owns hidden unit j, the 16 activations of each
// unit sit in shared memory as one 64-byte row read back with broadcast 128-bit loads ------------------------------
#define POL_TILE 16
__global__ void __launch_bounds__(256) rollout_mlp_kernel(const PolicyArgs pol, const PolScratch sc) {
    __shared__ float s_obs[POL_TILE][SIT_N_OBS];
    __shared__ __align__(16) float s_h[2][256][POL_TILE];   // [layer][unit][row]
    const int tid = threadIdx.x;
    const int n = (int)*sc.n_pending;
    const int H = pol.hidden;
    for (int t0 = blockIdx.x * POL_TILE; t0 < n; t0 += gridDim.x * POL_TILE) {
        const int nt = min(POL_TILE, n - t0);
        if (tid < POL_TILE * SIT_N_OBS) {
            const int e = tid / SIT_N_OBS, c = tid % SIT_N_OBS;
            s_obs[e][c] = e < nt ? sc.obs[(int64_t)(t0 + e) * SIT_N_OBS + c] : 0.0f;
        }
        __syncthreads();
        // layer 0: relu(b0 + obs . w0)   (nn_models/mlp.py:121-133)
        if (tid < H) {
            float acc[POL_TILE];
            const float b = pol.b0[tid];
#pragma unroll
            for (int e = 0; e < POL_TILE; ++e) acc[e] = b;
            for (int c = 0; c < SIT_N_OBS; ++c) {
                const float w = pol.w0[c * H + tid];
#pragma unroll
                for (int e = 0; e < POL_TILE; ++e) acc[e] = fmaf(s_obs[e][c], w, acc[e]);
            }
#pragma unroll
            for (int e = 0; e < POL_TILE; ++e) s_h[0][tid][e] = fmaxf(acc[e], 0.0f);
        }
        __syncthreads();
        // layer 1: relu(b1 + h0 . w1)   (:135-142)
        if (tid < H) {
            float acc[POL_TILE];
            const float b = pol.b1[tid];
#pragma unroll
            for (int e = 0; e < POL_TILE; ++e) acc[e] = b;
#pragma unroll 4
            for (int k = 0; k < H; ++k) {
                const float w = pol.w1[(int64_t)k * H + tid];
                const float4* hp = reinterpret_cast<const float4*>(&s_h[0][k][0]);
#pragma unroll
                for (int q = 0; q < POL_TILE / 4; ++q) {
                    const float4 hv = hp[q];
                    acc[4 * q + 0] = fmaf(hv.x, w, acc[4 * q + 0]); acc[4 * q + 1] = fmaf(hv.y, w, acc[4 * q + 1]);
                    acc[4 * q + 2] = fmaf(hv.z, w, acc[4 * q + 2]); acc[4 * q + 3] = fmaf(hv.w, w, acc[4 * q + 3]);
                }
            }
#pragma unroll
            for (int e = 0; e < POL_TILE; ++e) s_h[1][tid][e] = fmaxf(acc[e], 0.0f);
        }
        __syncthreads();
        // output layer: (mu, log_sigma) = b2 + h1 . w2, no output nonlinearity (distributions/normal.py:78-83)
        if (tid < POL_TILE * 2) {
            const int e = tid >> 1, c = tid & 1;
            float acc = pol.b2[c];
            for (int k = 0; k < H; ++k) acc = fmaf(s_h[1][k][e], pol.w2[k * 2 + c], acc);
            if (e < nt) sc.out[(int64_t)(t0 + e) * 2 + c] = acc;
        }
        __syncthreads();
    }
}

// ---- K4b (tensor cores): the same MLP with the H x H hidden layer on tcgen05.  One CTA = one 128-row tile of pending
// envs (persistent over tiles).  Layer 0 (K = 10) runs on the CUDA cores and writes relu(h0) as bf16 straight into shared
// memory in the UMMA canonical K-major layout (8 x 16-byte core matrices, no swizzle); W1 is staged once per CTA as the
// bf16 B operand (N x K, K-major, i.e. transposed on the way in).  One elected thread issues H/16 tcgen05.mma
// (M = 128, N = H, K = 16, fp32 accumulate in TMEM) and commits to an mbarrier; the epilogue reads the accumulator rows
// back with tcgen05.ld (thread = TMEM lane = env), applies bias + ReLU and folds the 2-wide output layer into the same
// pass.  Numerics: bf16 inputs to the hidden-layer product (h0 and W1 rounded to nearest even), fp32 everywhere else.
#define TC_M 128
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint64_t umma_desc_kmajor_noswizzle(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
    // cute::UMMA::SmemDescriptor: start >> 4 [0,14) | LBO >> 4 [16,30) | SBO >> 4 [32,46) | version 1 [46,48) | layout NONE [61,64)
    return (uint64_t)((saddr >> 4) & 0x3FFFu) | ((uint64_t)((lbo_bytes >> 4) & 0x3FFFu) << 16) |
           ((uint64_t)((sbo_bytes >> 4) & 0x3FFFu) << 32) | (1ull << 46);
}
__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
    uint32_t r;
    asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
    return r;
}

// W1 [k][n] fp32 (n contiguous) -> the bf16 B-operand image (N x K, K-major canonical layout) in global scratch, once per
// policy call (the learner may have changed the weights); every CTA then pulls the whole image with ONE bulk async copy.
__global__ void __launch_bounds__(256) rollout_pack_w1_kernel(const PolicyArgs pol, const PolScratch sc) {
    const int H = pol.hidden;
    const uint32_t sbo = 16u * (uint32_t)H, lbo = 128u;
    for (int item = blockIdx.x * 256 + threadIdx.x; item < H * (H >> 3); item += gridDim.x * 256) {
        const int nn = item % H, kc = item / H;
        const float* w = pol.w1 + (size_t)(kc * 8) * H + nn;
        uint4 v;
        v.x = pack_bf16x2(w[0], w[H]); v.y = pack_bf16x2(w[2 * H], w[3 * H]);
        v.z = pack_bf16x2(w[4 * H], w[5 * H]); v.w = pack_bf16x2(w[6 * H], w[7 * H]);
        *reinterpret_cast<uint4*>(sc.w1p + (size_t)(nn >> 3) * sbo + (size_t)kc * lbo + (nn & 7) * 16) = v;
    }
}

// <= 64 registers per thread (one 256-thread CTA beside three resident env-step CTAs of 128 x 128 registers): in a chained
// collect (rollout_chains) the policy chain of one env group runs while another group's step kernel fills the SMs
#ifndef SIT_POL_TC_MINB
#define SIT_POL_TC_MINB 4
#endif
__global__ void __launch_bounds__(256, SIT_POL_TC_MINB) rollout_mlp_tc_kernel(const PolicyArgs pol, const PolScratch sc, const int tmem_cols) {
    extern __shared__ __align__(128) unsigned char tc_smem[];
    const int H = pol.hidden;
    const int tid = threadIdx.x, warp = tid >> 5;
    const int n = (int)*sc.n_pending;
    if ((int)blockIdx.x * TC_M >= n) return;                         // nothing for this CTA: leave before staging W1
    unsigned char* sA = tc_smem;                                      // [128/8][H/8][8][8] bf16
    unsigned char* sB = sA + (size_t)TC_M * H * 2;                    // [H/8][H/8][8][8] bf16, rows = output unit n
    float* s_w2 = reinterpret_cast<float*>(sB + (size_t)H * H * 2);   // [H][2]
    float* s_b1 = s_w2 + 2 * H;                                       // [H]
    float* s_part = s_b1 + H;                                         // [128][2] partial output sums of the upper column half
    uint64_t* s_bar = reinterpret_cast<uint64_t*>(s_part + 2 * TC_M);   // [0]: MMAs of a tile done, [1]: W1 image landed
    uint32_t* s_tmem = reinterpret_cast<uint32_t*>(s_bar + 2);
    const uint32_t sbo = 16u * (uint32_t)H, lbo = 128u;               // next 8-row group | next 16-byte K chunk

    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(s_tmem)), "r"(tmem_cols));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
    }
    if (tid == 32) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(s_bar)));
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(s_bar + 1)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::);
        // B operand: the packed bf16 image of W1 (rollout_pack_w1_kernel) in one bulk async copy, landing on s_bar[1] while
        // the CTA computes layer 0 of its first tile
        const uint32_t bytes = (uint32_t)H * (uint32_t)H * 2u;
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(s_bar + 1)), "r"(bytes) : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     ::"r"(smem_u32(sB)), "l"(sc.w1p), "r"(bytes), "r"(smem_u32(s_bar + 1)) : "memory");
    }
    for (int i = tid; i < 2 * H; i += 256) s_w2[i] = pol.w2[i];
    for (int i = tid; i < H; i += 256) s_b1[i] = pol.b1[i];
    asm volatile("tcgen05.fence::before_thread_sync;" ::);
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::);
    const uint32_t tmem_base = *s_tmem;
    const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(H >> 3) << 17) | ((uint32_t)(TC_M >> 4) << 24);
    const int row = tid & (TC_M - 1), half = tid >> 7;                // TMEM lane / env of the tile | column half
    const int kh = H >> 1;                                            // columns per half (multiple of 16)
    uint32_t parity = 0;
    bool b_ready = false;
    for (int t0 = blockIdx.x * TC_M; t0 < n; t0 += gridDim.x * TC_M) {
        // layer 0 on the CUDA cores: relu(b0 + obs . w0) for this thread's row and column half (nn_models/mlp.py:121-133)
        float o[SIT_N_OBS];
#pragma unroll
        for (int c = 0; c < SIT_N_OBS; ++c) o[c] = (t0 + row) < n ? sc.obs[(int64_t)(t0 + row) * SIT_N_OBS + c] : 0.0f;
        for (int k0 = half * kh; k0 < (half + 1) * kh; k0 += 8) {
            float a[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) a[j] = pol.b0[k0 + j];
#pragma unroll
            for (int c = 0; c < SIT_N_OBS; ++c) {
                const float4 wa = *reinterpret_cast<const float4*>(pol.w0 + c * H + k0);
                const float4 wb = *reinterpret_cast<const float4*>(pol.w0 + c * H + k0 + 4);
                a[0] = fmaf(o[c], wa.x, a[0]); a[1] = fmaf(o[c], wa.y, a[1]); a[2] = fmaf(o[c], wa.z, a[2]); a[3] = fmaf(o[c], wa.w, a[3]);
                a[4] = fmaf(o[c], wb.x, a[4]); a[5] = fmaf(o[c], wb.y, a[5]); a[6] = fmaf(o[c], wb.z, a[6]); a[7] = fmaf(o[c], wb.w, a[7]);
            }
            uint4 v;
            v.x = pack_bf16x2(fmaxf(a[0], 0.f), fmaxf(a[1], 0.f)); v.y = pack_bf16x2(fmaxf(a[2], 0.f), fmaxf(a[3], 0.f));
            v.z = pack_bf16x2(fmaxf(a[4], 0.f), fmaxf(a[5], 0.f)); v.w = pack_bf16x2(fmaxf(a[6], 0.f), fmaxf(a[7], 0.f));
            *reinterpret_cast<uint4*>(sA + (size_t)(row >> 3) * sbo + (size_t)(k0 >> 3) * lbo + (row & 7) * 16) = v;
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy smem writes -> visible to the MMA's async proxy
        __syncthreads();
        // layer 1 on the tensor cores: D[128 x H] (TMEM, fp32) = A[128 x H] . B[H x H]
        if (tid == 0) {
            if (!b_ready) {   // first tile: wait for the W1 image
                asm volatile("{\n\t.reg .pred P1;\n\tLAB_WAITB:\n\tmbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n\t@P1 bra DONEB;\n\t"
                             "bra LAB_WAITB;\n\tDONEB:\n\t}" ::"r"(smem_u32(s_bar + 1)), "r"(0u) : "memory");
                b_ready = true;
            }
            asm volatile("tcgen05.fence::after_thread_sync;" ::);
            const uint64_t da = umma_desc_kmajor_noswizzle(smem_u32(sA), lbo, sbo), db = umma_desc_kmajor_noswizzle(smem_u32(sB), lbo, sbo);
            for (int ks = 0; ks < (H >> 4); ++ks) {
                const uint64_t adv = (uint64_t)((ks * 2 * lbo) >> 4);           // 16 bf16 = two core matrices along K
                const uint32_t acc = ks > 0 ? 1u : 0u;
                asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                             "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
                             ::"r"(tmem_base), "l"(da + adv), "l"(db + adv), "r"(idesc), "r"(acc) : "memory");
            }
            asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(s_bar)) : "memory");
        }
        asm volatile("{\n\t.reg .pred P1;\n\tLAB_WAIT:\n\tmbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n\t@P1 bra DONE;\n\t"
                     "bra LAB_WAIT;\n\tDONE:\n\t}" ::"r"(smem_u32(s_bar)), "r"(parity) : "memory");
        parity ^= 1u;
        asm volatile("tcgen05.fence::after_thread_sync;" ::);
        // epilogue: relu(b1 + D) folded into the output layer (mlp.py:135-142, normal.py:78-83)
        float mu = 0.0f, ls = 0.0f;
        const uint32_t taddr = tmem_base + ((uint32_t)((warp & 3) * 32) << 16);
        for (int c0 = half * kh; c0 < (half + 1) * kh; c0 += 16) {
            uint32_t r[16];
            asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
                         : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
                           "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
                         : "r"(taddr + (uint32_t)c0));
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
            for (int j = 0; j < 16; ++j) {
                const float h1 = fmaxf(__uint_as_float(r[j]) + s_b1[c0 + j], 0.0f);
                const float2 w = reinterpret_cast<const float2*>(s_w2)[c0 + j];
                mu = fmaf(h1, w.x, mu);
                ls = fmaf(h1, w.y, ls);
            }
        }
        if (half == 1) { s_part[2 * row] = mu; s_part[2 * row + 1] = ls; }
        asm volatile("tcgen05.fence::before_thread_sync;" ::);
        __syncthreads();                                                // accumulator fully read: the next tile may overwrite it
        if (half == 0 && (t0 + row) < n) {
            sc.out[(int64_t)(t0 + row) * 2] = mu + s_part[2 * row] + pol.b2[0];
            sc.out[(int64_t)(t0 + row) * 2 + 1] = ls + s_part[2 * row + 1] + pol.b2[1];
        }
    }
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(tmem_cols));
}

// ---- K4c / K4d: sample, convert, validate, publish.  The first valid try in draw order wins, exactly as a sequential
// redraw loop would pick it.  K4c: one warp per pending env, lanes = tries 0..31 (most envs finish here).  K4d: one CTA
// per env that found no valid waypoint yet, threads = the next 256 tries per round (main_ast.py:49-51 allows 1000). -------
template <typename R>
struct TryCtx { int64_t gid; int32_t serial; R base_n, base_e, L, alpha; float mu, sig; };

template <typename R>
__device__ __forceinline__ void load_try_ctx(TryCtx<R>& c, const ShipPtrs<R>& g, const EnvPtrs<R>& ge, const PolicyArgs& pol,
                                             const RollArgs& ro, const PolScratch& sc, int p, int64_t e) {
    c.gid = ro.env_id_offset + e;
    c.serial = ge.dec_serial[e];
    const int64_t io_ = 2 * e + 1;                        // obstacle ship slot
    // previous intermediate waypoint, or the route start for the first decision of the episode
    const bool first = ge.n_dec[e] == 0;
    c.base_n = first ? g.wp0_n[io_] : ge.act_n[e];
    c.base_e = first ? g.wp0_e[io_] : ge.act_e[e];
    c.L = ge.ab_seg[e]; c.alpha = ge.ab_alpha[e];
    c.mu = 0.0f; c.sig = 0.0f;
    if (pol.mode != SIT_POLICY_UNIFORM) {
        c.mu = sc.out[(int64_t)p * 2];
        c.sig = expf(fminf(fmaxf(sc.out[(int64_t)p * 2 + 1], -20.0f), 2.0f));   // LOG_SIG_CAP_MIN / MAX, normal.py:14-15,86-87
    }
}

// Every candidate waypoint of one decision lies on the arc of radius L about (base_n, base_e) between the headings
// alpha -/+ pi/6.  One warp lists the ring segments that can matter for a point of that arc: a segment whose y-range misses
// the arc's bounding box can neither be crossed by nor touch the +x ray of such a point, and a segment strictly left of
// the box is ignored by the crossing test itself.  The box is the hull of the arc's end points and mid point grown by the
// sagitta of a half arc, L (1 - cos(pi/12)) < 0.035 L, plus slack for rounding.  The re-draw loop then tests ~5-10 segments
// instead of all 39 with an identical verdict.  Returns the list length (all lanes); list is per-warp shared memory.
template <typename R>
__device__ __forceinline__ int arc_edge_list(const TryCtx<R>& c, const EdgeRec<R>* E, const int n_edges, uint8_t* list,
                                             const int lane) {
    const R sixth = (R)0.52359877559829887308;
    R s0, c0, s1, c1, sm, cm;
    sincos_(c.alpha - sixth, s0, c0); sincos_(c.alpha + sixth, s1, c1); sincos_(c.alpha, sm, cm);
    const R grow = (R)0.04 * abs_(c.L) + (R)2;
    const R ylo = c.base_n + min_(min_(c.L * c0, c.L * c1), c.L * cm) - grow, yhi = c.base_n + max_(max_(c.L * c0, c.L * c1), c.L * cm) + grow;
    const R xlo = c.base_e + min_(min_(c.L * s0, c.L * s1), c.L * sm) - grow;
    int n = 0;
    for (int i0 = 0; i0 < n_edges; i0 += 32) {
        const int i = i0 + lane;
        bool keep = false;
        if (i < n_edges) {
            const EdgeRec<R> e = E[i];
            const bool y_miss = max_(e.y1, e.y2) < ylo || min_(e.y1, e.y2) > yhi;
            const bool left = e.x1 < xlo && e.x2 < xlo;
            keep = !(y_miss || left);                       // NaN box: every comparison is false, every segment is kept
        }
        const unsigned b = __ballot_sync(0xffffffffu, keep);
        if (keep) list[n + __popc(b & ((1u << lane) - 1u))] = (uint8_t)i;
        n += __popc(b);
    }
    __syncwarp();
    return n;
}

// try number tr of one decision: action, waypoint, validity
template <typename R>
__device__ __forceinline__ bool try_waypoint(const TryCtx<R>& c, const EnvKArgs<R>& A, const EdgeRec<R>* E,
                                             const uint8_t* list, int n_list,
                                             int mode, uint64_t seed, int tr, float& la, R& ln, R& le) {
    uint32_t x[4];
    philox4x32_10((uint32_t)c.gid, (uint32_t)((uint64_t)c.gid >> 32), (uint32_t)c.serial, (uint32_t)tr, (uint32_t)seed,
                  (uint32_t)(seed >> 32), x);
    const double u1 = ((double)x[0] + 0.5) * (1.0 / 4294967296.0), u2 = ((double)x[1] + 0.5) * (1.0 / 4294967296.0);
    if (mode == SIT_POLICY_UNIFORM) {
        la = (float)(2.0 * u1 - 1.0);                     // policies/uniform_policy.py:21-22
    } else if (mode == SIT_POLICY_MEAN) {
        la = tanhf(c.mu);                                 // deterministic branch, gaussian_policy.py:129-141
    } else {
        const float eps = (float)(sqrt(-2.0 * log(u1)) * cospi(2.0 * u2));
        la = tanhf(fmaf(c.sig, eps, c.mu));               // normal.py:91-93, gaussian_policy.py:116-117
    }
    R sn, cs;
    sincos_(c.alpha + (R)la * (R)0.52359877559829887308, sn, cs);   // scoping angle a * pi/6 about AB_alpha
    ln = c.base_n + c.L * cs; le = c.base_e + c.L * sn;
    return (ln < A.EP.min_n || ln > A.EP.max_n || le < A.EP.min_e || le > A.EP.max_e) || map_contains_list(E, list, n_list, ln, le);
}

template <typename R>
__device__ __forceinline__ void publish_action(const EnvPtrs<R>& ge, const RollArgs& ro, int64_t e, float a, R wn, R we, bool bad,
                                               int tries) {
    if (tries > 1) atomicAdd(&ro.counters[SIT_RC_RESAMPLES], (unsigned long long)(tries - 1));
    ge.act_n[e] = wn; ge.act_e[e] = we;
    ge.flags[e] = (ge.flags[e] & ~F_IW_BAD) | F_HAVE_ACT | F_IW_VALID | (bad ? F_IW_BAD : 0);
    ge.n_dec[e] += 1; ge.dec_serial[e] += 1;
    ro.replay[(int64_t)ge.tr_slot[e] * SIT_TR_FLOATS + 10] = a;
}

template <typename R>
__global__ void __launch_bounds__(256) rollout_sample_kernel(const __grid_constant__ EnvKArgs<R> A, const ShipPtrs<R> g,
                                                             const EnvPtrs<R> ge, const PolicyArgs pol, const RollArgs ro,
                                                             const PolScratch sc) {
    __shared__ EdgeRec<R> s_edges[SIT_MAX_EDGES];
    stage_edges(A.M, s_edges);
    __shared__ uint8_t s_list[8][SIT_MAX_EDGES];
    const int n_edges = A.M.n_edges;
    const int lane = threadIdx.x & 31;
    uint8_t* list = s_list[threadIdx.x >> 5];
    const int n = (int)*sc.n_pending;
    const int warps = gridDim.x * (blockDim.x >> 5);
    const int max_tries = pol.mode == SIT_POLICY_MEAN ? 1 : 1 + max(ro.max_route_resampling, 0);
    for (int p = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); p < n; p += warps) {
        const int64_t e = sc.env[p];
        TryCtx<R> c;
        load_try_ctx(c, g, ge, pol, ro, sc, p, e);
        __syncwarp();                                      // the previous env's list is no longer read
        const int n_list = arc_edge_list(c, s_edges, n_edges, list, lane);
        float la; R ln, le;
        const bool lbad = try_waypoint(c, A, s_edges, list, n_list, pol.mode, ro.seed, lane, la, ln, le);
        const unsigned ok = __ballot_sync(0xffffffffu, !lbad && lane < max_tries);
        if (!ok && max_tries > 32) {              // no valid waypoint among the first 32 draws: CTA-wide rounds (K4d)
            if (lane == 0) sc.retry[atomicAdd(sc.n_retry, 1u)] = p;
            continue;
        }
        const int src = ok ? (__ffs(ok) - 1) : (max_tries - 1);
        if (lane == src) publish_action(ge, ro, e, la, ln, le, ok == 0u, src + 1);
    }
}

template <typename R>
__global__ void __launch_bounds__(256) rollout_resample_kernel(const __grid_constant__ EnvKArgs<R> A, const ShipPtrs<R> g,
                                                               const EnvPtrs<R> ge, const PolicyArgs pol, const RollArgs ro,
                                                               const PolScratch sc) {
    __shared__ EdgeRec<R> s_edges[SIT_MAX_EDGES];
    __shared__ int s_best, s_nlist;
    __shared__ uint8_t s_list[SIT_MAX_EDGES];
    stage_edges(A.M, s_edges);
    const int n_edges = A.M.n_edges;
    const int n = (int)*sc.n_retry;
    const int max_tries = 1 + max(ro.max_route_resampling, 0);
    for (int q = blockIdx.x; q < n; q += gridDim.x) {
        const int p = sc.retry[q];
        const int64_t e = sc.env[p];
        TryCtx<R> c;
        load_try_ctx(c, g, ge, pol, ro, sc, p, e);
        if (threadIdx.x < 32) {
            const int nl = arc_edge_list(c, s_edges, n_edges, s_list, threadIdx.x);
            if (threadIdx.x == 0) { s_best = 0x7fffffff; s_nlist = nl; }
        }
        __syncthreads();
        const int n_list = s_nlist;
        for (int r0 = 32; r0 < max_tries; r0 += 256) {
            const int tr = r0 + threadIdx.x;
            float la = 0.0f; R ln = (R)0, le = (R)0;
            bool lbad = true;
            if (tr < max_tries) {
                lbad = try_waypoint(c, A, s_edges, s_list, n_list, pol.mode, ro.seed, tr, la, ln, le);
                if (!lbad) atomicMin(&s_best, tr);
            }
            __syncthreads();
            const int best = s_best;
            __syncthreads();   // everyone has read this round's verdict before the next round's atomicMin
            if (best != 0x7fffffff) {
                if (tr == best) publish_action(ge, ro, e, la, ln, le, false, tr + 1);
                break;
            }
            if (tr == max_tries - 1) publish_action(ge, ro, e, la, ln, le, true, max_tries);   // every draw was invalid
        }
        __syncthreads();
    }
}

// arm a rollout: every env pending on reset()'s float32 initial state (MSRL_env_ex.py:86-90,186)
template <typename R>
__global__ void rollout_begin_kernel(const ShipPtrs<R> g, const EnvPtrs<R> ge) {
    const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t B = ge.n_envs;
    if (e >= B) return;
    ge.flags[e] = (ge.flags[e] & ~F_HAVE_ACT) | F_PENDING;
    ge.n_dec[e] = 0;
    R* co = ge.cur_obs + e;
    const int64_t it = 2 * e, io_ = 2 * e + 1;
    co[0] = (R)(float)g.init_north[it]; co[B] = (R)(float)g.init_east[it]; co[2 * B] = (R)(float)g.init_psi[it];
    co[3 * B] = (R)0; co[4 * B] = (R)0; co[5 * B] = (R)0;
    co[6 * B] = (R)(float)g.init_north[io_]; co[7 * B] = (R)(float)g.init_east[io_]; co[8 * B] = (R)(float)g.init_psi[io_];
    co[9 * B] = (R)0;
}

// Map queries for a batch of candidate waypoints: is_route_outside_horizon (MSRL_env_ex.py:517-542), is_route_inside_obstacles
// (:544-552 -> obstacle.py:131-136) and obstacles_distance (obstacle.py:138-141).  Used by the waypoint sampler to
// reject invalid intermediate waypoints before they reach step().
template <typename R>
__global__ void __launch_bounds__(256) map_query_kernel(const __grid_constant__ EnvKArgs<R> A, const int64_t n, const R* north,
                                                        const R* east, uint8_t* flags, R* dist) {
    __shared__ EdgeRec<R> s_edges[SIT_MAX_EDGES];
    stage_edges(A.M, s_edges);
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const R pn = north[i], pe = east[i];
    const EnvP<R>& EP = A.EP;
    uint8_t f = 0;
    if (map_contains(s_edges, A.M.n_edges, pn, pe)) f |= 1;
    if (pn < EP.min_n || pn > EP.max_n || pe < EP.min_e || pe > EP.max_e) f |= 2;
    if (flags) flags[i] = f;
    if (dist) {
        R d2; bool in;
        map_scan(s_edges, A.M.n_edges, pn, pe, d2, in);
        dist[i] = sqrt_(d2);
    }
}

// StaticObstacle (obstacle.py:8-90): circular obstacles.  obstacles_distance :55-67 = min_i (sqrt(|d_i^2|) - r_i);
// if_ship_inside_obstacles / if_route_inside_obstacles :69-85 = any_i (d_i^2 <= r_i^2) with r_i^2 formed by the same product.
// The reference's operation order (squares summed north first, no FMA contraction) is kept so the fp64 result is the
// reference's bit for bit.
#define SIT_MAX_CIRCLES 64
template <typename R> struct CircleSet { int32_t n, _pad; R cn[SIT_MAX_CIRCLES], ce[SIT_MAX_CIRCLES], cr[SIT_MAX_CIRCLES]; };
__device__ __forceinline__ double mul_rn_(double a, double b) { return __dmul_rn(a, b); }
__device__ __forceinline__ float mul_rn_(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ double add_rn_(double a, double b) { return __dadd_rn(a, b); }
__device__ __forceinline__ float add_rn_(float a, float b) { return __fadd_rn(a, b); }
template <typename R>
__global__ void __launch_bounds__(256) circle_query_kernel(const __grid_constant__ CircleSet<R> cs, const int64_t n,
                                                           const R* __restrict__ north, const R* __restrict__ east,
                                                           uint8_t* __restrict__ flags, R* __restrict__ dist) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const R pn = north[i], pe = east[i];
    bool inside = false;
    R dmin = (R)0;
    for (int c = 0; c < cs.n; ++c) {
        const R dn = pn - cs.cn[c], de = pe - cs.ce[c];
        const R d2 = add_rn_(mul_rn_(dn, dn), mul_rn_(de, de));
        inside |= d2 <= mul_rn_(cs.cr[c], cs.cr[c]);
        const R d = sizeof(R) == 8 ? (R)sqrt((double)abs_(d2)) - cs.cr[c] : (R)sqrtf((float)abs_(d2)) - cs.cr[c];
        dmin = c == 0 ? d : min_(dmin, d);
    }
    if (flags) flags[i] = inside ? 1 : 0;
    if (dist) dist[i] = dmin;
}

// ship_snap_shot (ship_model.py:326-338) for n poses: ShipDraw.local_coords -> rotate_coords(psi) -> translate_coords
// (utils.py:67-101); six outline points per pose, x_ned / y_ned [n][6].
template <typename R>
__global__ void __launch_bounds__(256) ship_outline_kernel(const int64_t n, const R* __restrict__ north, const R* __restrict__ east,
                                                           const R* __restrict__ psi, R* __restrict__ x_out, R* __restrict__ y_out) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const R l = (R)80, b = (R)20;
    const R lx[6] = {-l / 2, l / 4, l / 2, l / 4, -l / 2, -l / 2};
    const R ly[6] = {-b / 2, -b / 2, (R)0, b / 2, b / 2, -b / 2};
    R sn, cs;
    if (sizeof(R) == 8) { double s_, c_; sincos((double)psi[i], &s_, &c_); sn = (R)s_; cs = (R)c_; }
    else { float s_, c_; sincosf((float)psi[i], &s_, &c_); sn = (R)s_; cs = (R)c_; }
    const R pn = north[i], pe = east[i];
#pragma unroll
    for (int j = 0; j < 6; ++j) {
        x_out[i * 6 + j] = add_rn_(add_rn_(mul_rn_(cs, lx[j]), -mul_rn_(sn, ly[j])), pn);   // cos x - sin y + north
        y_out[i * 6 + j] = add_rn_(add_rn_(mul_rn_(sn, lx[j]), mul_rn_(cs, ly[j])), pe);    // sin x + cos y + east
    }
}

// =================================================================================================
// FMA-peak microbenchmark (roofline denominator for the compute-bound step kernels)
// =================================================================================================
template <typename R>
__global__ void __launch_bounds__(256) fma_peak_kernel(R* out, int iters, R a, R b) {
    R x0 = (R)threadIdx.x, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int k = 0; k < 16; ++k) {
            x0 = x0 * a + b; x1 = x1 * a + b; x2 = x2 * a + b; x3 = x3 * a + b;
            x4 = x4 * a + b; x5 = x5 * a + b; x6 = x6 * a + b; x7 = x7 * a + b;
        }
    }
    const R sum = ((x0 + x1) + (x2 + x3)) + ((x4 + x5) + (x6 + x7));
    if (sum == (R)123456789) out[0] = sum;   // never true: keeps the chain alive
}

// the same chain as packed f32x2 FMAs (FFMA2: two FMAs per lane per instruction)
__global__ void __launch_bounds__(256) fma_peak_x2_kernel(float* out, int iters, float a, float b) {
    const float t = (float)threadIdx.x;
    float2 x0 = make_float2(t, t + 8), x1 = make_float2(t + 1, t + 9), x2 = make_float2(t + 2, t + 10), x3 = make_float2(t + 3, t + 11),
           x4 = make_float2(t + 4, t + 12), x5 = make_float2(t + 5, t + 13), x6 = make_float2(t + 6, t + 14), x7 = make_float2(t + 7, t + 15);
    const float2 a2 = make_float2(a, a), b2 = make_float2(b, b);
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int k = 0; k < 16; ++k) {
            x0 = __ffma2_rn(x0, a2, b2); x1 = __ffma2_rn(x1, a2, b2); x2 = __ffma2_rn(x2, a2, b2); x3 = __ffma2_rn(x3, a2, b2);
            x4 = __ffma2_rn(x4, a2, b2); x5 = __ffma2_rn(x5, a2, b2); x6 = __ffma2_rn(x6, a2, b2); x7 = __ffma2_rn(x7, a2, b2);
        }
    }
    const float2 s = __fadd2_rn(__fadd2_rn(__fadd2_rn(x0, x1), __fadd2_rn(x2, x3)), __fadd2_rn(__fadd2_rn(x4, x5), __fadd2_rn(x6, x7)));
    if (s.x + s.y == 123456789.0f) out[0] = s.x;   // never true: keeps the chain alive
}

// =================================================================================================
// Host side
// =================================================================================================
static thread_local std::string g_err;
static int fail(int code, const std::string& msg) { g_err = msg; return code; }
#define CUDA_TRY(x)                                                                                   \
    do {                                                                                              \
        cudaError_t e_ = (x);                                                                         \
        if (e_ != cudaSuccess) return fail(SIT_E_CUDA, std::string(#x) + ": " + cudaGetErrorString(e_)); \
    } while (0)

struct FieldDef { const char* name; char kind; int rows; bool env; };   // kind: D f64, R real, I i32
static const FieldDef kShipFields[] = {
    {"north", 'D', 1, false}, {"east", 'D', 1, false}, {"init_north", 'D', 1, false}, {"init_east", 'D', 1, false},
    {"psi", 'R', 1, false}, {"u", 'R', 1, false}, {"v", 'R', 1, false}, {"r", 'R', 1, false}, {"omega", 'R', 1, false},
    {"ei_speed", 'R', 1, false}, {"ei_shaft", 'R', 1, false}, {"ei_heading", 'R', 1, false}, {"e_prev", 'R', 1, false},
    {"e_ct", 'R', 1, false}, {"e_ct_int", 'R', 1, false},
    {"fuel_me", 'R', 1, false}, {"fuel_el", 'R', 1, false}, {"fuel", 'R', 1, false},
    {"u_sp", 'R', 1, false}, {"vc_n", 'R', 1, false}, {"vc_e", 'R', 1, false}, {"wind_speed", 'R', 1, false},
    {"wind_dir", 'R', 1, false},
    {"init_psi", 'R', 1, false}, {"init_u", 'R', 1, false}, {"init_v", 'R', 1, false}, {"init_r", 'R', 1, false},
    {"init_vc_n", 'R', 1, false}, {"init_vc_e", 'R', 1, false}, {"init_wind_speed", 'R', 1, false},
    {"init_wind_dir", 'R', 1, false}, {"init_u_sp", 'R', 1, false}, {"init_omega", 'R', 1, false},
    {"init_ei_shaft", 'R', 1, false},
    {"k", 'I', 1, false}, {"n_wp", 'I', 1, false}, {"mode", 'I', 1, false}, {"n_wp0", 'I', 1, false},
    {"steps", 'I', 1, false},
    {"wp_n", 'R', -1, false}, {"wp_e", 'R', -1, false}, {"wp0_n", 'R', -1, false}, {"wp0_e", 'R', -1, false},
    // nearest-ring cache of the env kernel (env_core.cuh MapCache): always float32
    {"mc_idx", 'I', 1, false}, {"mc_n", 'F', 1, false}, {"mc_e", 'F', 1, false}, {"mc_b2", 'F', 1, false},
    // packed fp32 fleet kernel (ship_x2.cuh): position relative to the segment start as a compensated float pair, carried
    // across launches beside north / east (which stay authoritative: a pair that no longer matches them is re-derived)
    {"rel_n", 'F', 1, false}, {"rel_cn", 'F', 1, false}, {"rel_e", 'F', 1, false}, {"rel_ce", 'F', 1, false},
};
static const FieldDef kEnvFields[] = {
    {"prev_n", 'D', 1, true}, {"prev_e", 'D', 1, true}, {"eps_dist", 'D', 1, true}, {"samp_dist", 'D', 1, true},
    {"ab_seg", 'R', 1, true}, {"act_n", 'R', 1, true}, {"act_e", 'R', 1, true}, {"ep_return", 'R', 1, true},
    {"flags", 'I', 1, true}, {"ep_steps", 'I', 1, true},
    {"ab_alpha", 'R', 1, true}, {"cur_obs", 'R', SIT_N_OBS, true},
    {"n_dec", 'I', 1, true}, {"dec_serial", 'I', 1, true}, {"tr_slot", 'I', 1, true},
    {"draw_t", 'D', 1, true},
};

#define SIT_PIPE 4
struct SitHandle_ {
    SitConfig cfg;
    int64_t n_envs, n_ships;
    int device, ships_per_env, sm_count;
    bool same_params;
    std::vector<SitField> fields;
    size_t state_bytes;
    ShipP<double> pd[2]; ShipP<float> pf[2];
    x2::ShipPX px;                               // packed fp32 fleet step (fleet handles, ship[0])
    EnvKArgs<double>* kd; EnvKArgs<float>* kf;   // host copies passed by value at launch
    // scratch for the *_host entry point
    void* d_scratch; size_t scratch_bytes;
    void* d_pol; size_t pol_bytes;               // sit_rollout_policy scratch
    void* h_pinned; size_t pinned_bytes;
    // sit_rollout_collect: instantiated CUDA graph of one (policy + K steps) iteration and the arguments it was captured for
    cudaGraphExec_t roll_graph; std::string roll_key;
    std::string roll_key_seen;                   // chained collect: arguments of the last eager run (captured on their second use)
    cudaEvent_t cnt_ev[SIT_PIPE + 1];            // chained collect: row-slot reservation order between the env groups
    size_t tc_smem_set;                          // dynamic shared memory opted in for rollout_mlp_tc_kernel
    // sit_fleet_step_host: chunk pipeline over internal streams
    bool pipe_ready; cudaStream_t pipe[SIT_PIPE]; cudaEvent_t pipe_done[SIT_PIPE], pipe_in;
};

template <typename R> struct HostSel;
template <> struct HostSel<double> {
    static const ShipP<double>& ship(const SitHandle_* h) { return h->pd[0]; }
    static const EnvKArgs<double>& env(const SitHandle_* h) { return *h->kd; }
};
template <> struct HostSel<float> {
    static const ShipP<float>& ship(const SitHandle_* h) { return h->pf[0]; }
    static const EnvKArgs<float>& env(const SitHandle_* h) { return *h->kf; }
};

static bool params_equal(const SitShipParams& a, const SitShipParams& b) { return std::memcmp(&a, &b, sizeof(a)) == 0; }

template <typename R>
static R largest_with_sqrt_le(double bound) {
    // largest x (in R) such that sqrt_R(x) <= bound: lets `sqrt(d2) <= bound` (MSRL_env_ex.py:754,829) be decided
    // without the square root, bit-for-bit
    R x = (R)(bound * bound);
    while (std::sqrt(x) > (R)bound) x = std::nextafter(x, (R)0);
    for (;;) {
        R y = std::nextafter(x, (R)INFINITY);
        if (std::sqrt(y) <= (R)bound) x = y; else break;
    }
    return x;
}

template <typename R>
static void derive_ship(const SitShipParams& c, ShipP<R>& P) {
    const double PI = 3.14159265358979323846;
    // ship_model.py:71-80
    const double payload = 0.9 * (c.dead_weight_tonnage - c.bunkers);
    const double lsw = c.dead_weight_tonnage / c.coefficient_of_deadweight_to_displacement - c.dead_weight_tonnage;
    const double mass = lsw + payload + c.bunkers + c.ballast;
    const double L = c.length_of_ship, W = c.width_of_ship;
    const double i_z = mass * (L * L + W * W) / 12;
    // ship_model.py:193-209
    const double x_du = mass * c.added_mass_coefficient_in_surge, y_dv = mass * c.added_mass_coefficient_in_sway,
                 n_dr = i_z * c.added_mass_coefficient_in_yaw;
    // ship_model.py:123-130
    const double rho = 1.2, cx = 0.5, cy = 0.7, cn = 0.08, a_f = W * 8.0, a_l = L * 8.0;
    const double m11 = mass + x_du, m22 = mass + y_dv, m33 = i_z + n_dr;
    std::memset(&P, 0, sizeof(P));
    P.dt = (R)c.dt; P.inv_dt = (R)(1.0 / c.dt);
    P.ra2 = c.radius_of_acceptance * c.radius_of_acceptance;
    P.look = (R)c.lookahead_distance; P.look2 = (R)(c.lookahead_distance * c.lookahead_distance);
    P.sat_e = (R)(0.99 * c.lookahead_distance);
    P.ki_los = (R)c.los_integral_gain; P.windup = (R)c.los_integrator_windup_limit;
    P.kp_h = (R)c.kp_heading; P.kd_h = (R)c.kd_heading; P.ki_h = (R)c.ki_heading;
    P.rud_max = (R)(c.max_rudder_angle_degrees * PI / 180); P.rud_bias = (R)(3.0 * PI / 180.0);
    P.kp_s = (R)c.kp_ship_speed; P.ki_s = (R)c.ki_ship_speed; P.kp_w = (R)c.kp_shaft_speed; P.ki_w = (R)c.ki_shaft_speed;
    P.hotel = (R)c.hotel_load;
    for (int j = 0; j < 3; ++j) { P.sfc_me[j] = (R)(c.sfc_me[j] / 3.6e9); P.sfc_dg[j] = (R)(c.sfc_dg[j] / 3.6e9); }
    const double jp = c.propeller_inertia;
    P.sh_me = (R)(1.0 / (c.gear_ratio_between_main_engine_and_propeller * jp));
    P.sh_hsg = (R)(1.0 / (c.gear_ratio_between_hybrid_shaft_generator_and_propeller * jp));
    P.sh_lin = (R)((c.linear_friction_main_engine / c.gear_ratio_between_main_engine_and_propeller +
                    c.linear_friction_hybrid_shaft_generator / c.gear_ratio_between_hybrid_shaft_generator_and_propeller) / jp);
    P.sh_quad = (R)(c.propeller_speed_to_torque_coefficient / jp);
    const double thrust_coef = std::pow(c.propeller_diameter, 4) * c.propeller_speed_to_thrust_force_coefficient;
    P.thrust_coef = (R)thrust_coef;
    P.rad2deg = (R)(180.0 / PI); P.rpm_per_rads = (R)(30.0 / PI);
    P.au_vr = (R)(mass / m11); P.au_vrr = (R)(y_dv / m11);
    P.au_lin = (R)(mass / c.mass_over_linear_friction_coefficient_in_surge / m11);
    P.au_nl = (R)(c.nonlinear_friction_coefficient__in_surge / m11);
    P.au_wind = (R)(0.5 * rho * cx * a_f / m11); P.au_thr = (R)(thrust_coef / m11);
    P.av_ur = (R)(mass / m22); P.av_urr = (R)(x_du / m22);
    P.av_lin = (R)(mass / c.mass_over_linear_friction_coefficient_in_sway / m22);
    P.av_nl = (R)(c.nonlinear_friction_coefficient__in_sway / m22);
    P.av_wind = (R)(0.5 * rho * cy * a_l / m22); P.av_rud = (R)(c.rudder_angle_to_sway_force_coefficient / m22);
    P.ar_ca = (R)((x_du - y_dv) / m33);
    P.ar_lin = (R)(i_z / c.mass_over_linear_friction_coefficient_in_yaw / m33);
    P.ar_nl = (R)(c.nonlinear_friction_coefficient__in_yaw / m33);
    P.ar_wind = (R)(2.0 * 0.5 * rho * cn * a_l * L / m33); P.ar_rud = (R)(c.rudder_angle_to_yaw_force_coefficient / m33);
    P.half_len = (R)(L / 2);
    P.mach = c.machinery_kind;
    P.k_thrust = (R)(2160.0 / 790.0);                                  // ship_engine.py:420
    P.inv_thrust_T = (R)(c.machinery_kind == SIT_MACH_SIMPLIFIED ? 1.0 / c.thrust_force_dynamic_time_constant : 0.0);
    P.kp_sp = (R)c.kp_simplified; P.ki_sp = (R)c.ki_simplified;
    P.inv_m11 = (R)(1.0 / m11);
    P.n_modes = c.n_modes;
    for (int j = 0; j < c.n_modes && j < SIT_MAX_MODES; ++j) {
        ModeP<R>& M = P.mode[j];
        const double me = c.mode_main_engine_capacity[j], el = c.mode_electrical_capacity[j], hotel = c.hotel_load;
        const int sg = c.mode_shaft_generator_state[j];
        double at, ame, ael;   // ship_engine.py:32-44
        if (sg == SIT_SG_MOTOR) { at = me + el - hotel; ame = me; ael = el - hotel; }
        else if (sg == SIT_SG_GEN) { at = me - hotel; ame = me - hotel; ael = 0; }
        else { at = me; ame = me; ael = 0; }
        M.avail_total = (R)at; M.avail_me = (R)ame; M.avail_el = (R)ael;
        M.cap_tq_me = (R)(ame / 5 * PI / 30); M.cap_tq_el = (R)(ael / 5 * PI / 30);
        M.me_cap = (R)me; M.me_cap_kw = (R)(me / 1000); M.el_cap_kw = (R)(el / 1000);
        M.inv_me_cap = (R)((sg == SIT_SG_MOTOR && me == 0) ? 0.0 : 1.0 / me);
        M.inv_el_cap = (R)((sg == SIT_SG_GEN && el == 0) ? 0.0 : 1.0 / el);
        if (sg == SIT_SG_GEN) { const double le = hotel < el ? hotel : el; M.load_me_add = (R)(hotel - le); M.load_el_const = (R)le; }
        else { M.load_me_add = (R)0; M.load_el_const = (R)hotel; }
        M.sg = sg;
    }
}

// uniform parameters of the packed fp32 step (ship_x2.cuh) from the scalar ones: folded gains, branch-free machinery modes
static void derive_shipx(const ShipP<float>& P, x2::ShipPX& X) {
    std::memset(&X, 0, sizeof(X));
    X.ra2 = P.ra2;
    X.ra2_band = (float)(1.01 * P.ra2 + 16.0);
    X.c_inv_pi = 0.318309886183790672f; X.c_s3 = 2.600024345156271e-06f; X.c_c4 = -2.607681608424173e-07f;
    X.c_a6 = -0.004355357028543949f;
    X.dt = P.dt; X.look2 = P.look2; X.sat_e = P.sat_e; X.neg_ki_los = -P.ki_los; X.windup = P.windup;
    X.kp_h = P.kp_h; X.kd_idt = P.kd_h * P.inv_dt; X.ki_h = P.ki_h; X.rud_max = P.rud_max; X.rud_bias = P.rud_bias;
    X.kp_kd = P.kp_h + P.kd_h * P.inv_dt;
    X.kp_s = P.kp_s; X.ki_s = P.ki_s; X.kp_w = P.kp_w; X.ki_w = P.ki_w;
    X.hotel = P.hotel;
    for (int j = 0; j < 3; ++j) { X.sfc_me[j] = P.sfc_me[j] * P.dt; X.sfc_dg[j] = P.sfc_dg[j] * P.dt; }
    X.sh_me = P.sh_me; X.sh_hsg = P.sh_hsg; X.sh_lin = P.sh_lin; X.sh_quad = P.sh_quad;
    X.au_vr = P.au_vr; X.au_vrr = P.au_vrr; X.au_lin = P.au_lin; X.au_nl = P.au_nl; X.au_wind = P.au_wind;
    X.au_thr = P.au_thr;   // already carries D^4 k_t (derive_ship)
    X.av_ur = P.av_ur; X.av_urr = P.av_urr; X.av_lin = P.av_lin; X.av_nl = P.av_nl; X.av_wind = P.av_wind; X.av_rud = P.av_rud;
    X.ar_ca = P.ar_ca; X.ar_lin = P.ar_lin; X.ar_nl = P.ar_nl; X.ar_wind = P.ar_wind; X.ar_rud = P.ar_rud;
    X.n_modes = P.n_modes;
    for (int j = 0; j < P.n_modes && j < SIT_MAX_MODES; ++j) {
        const ModeP<float>& M = P.mode[j];
        x2::ModeX& Y = X.mode[j];
        Y.avail_total = M.avail_total; Y.avail_me = M.avail_me; Y.avail_el = M.avail_el;
        Y.cap_tq_me = M.cap_tq_me; Y.cap_tq_el = M.cap_tq_el;
        Y.inv_me_cap = M.inv_me_cap; Y.inv_el_cap = M.inv_el_cap; Y.me_cap_kw = M.me_cap_kw;
        Y.fme_a = (float)((double)X.sfc_me[0] * M.inv_me_cap * M.inv_me_cap); Y.fme_b = (float)((double)X.sfc_me[1] * M.inv_me_cap);
        Y.fel_a = (float)((double)X.sfc_dg[0] * M.inv_el_cap * M.inv_el_cap); Y.fel_b = (float)((double)X.sfc_dg[1] * M.inv_el_cap);
        if (M.sg == SIT_SG_MOTOR) { Y.add_me = 0.0f; Y.cap_me = M.me_cap; Y.a_el = 1.0f; Y.c_el = 0.0f; }
        else { Y.add_me = M.load_me_add; Y.cap_me = INFINITY; Y.a_el = 0.0f; Y.c_el = M.load_el_const; }
    }
}

template <typename R>
static void derive_env(const SitConfig& c, EnvKArgs<R>& A) {
    derive_ship(c.ship[0], A.P[0]);
    derive_ship(c.ship[1], A.P[1]);
    const SitMap& m = c.map;
    MapP<R>& M = A.M;
    std::memset(&M, 0, sizeof(M));
    M.n_edges = m.n_edges; M.n_poly = m.n_poly;
    double mn_n = 1e300, mx_n = -1e300, mn_e = 1e300, mx_e = -1e300;
    for (int i = 0; i < m.n_edges; ++i) {
        M.x1[i] = (R)m.x1[i]; M.y1[i] = (R)m.y1[i]; M.x2[i] = (R)m.x2[i]; M.y2[i] = (R)m.y2[i];
        const double dx = m.x2[i] - m.x1[i], dy = m.y2[i] - m.y1[i], l2 = dx * dx + dy * dy;
        M.dx[i] = (R)dx; M.dy[i] = (R)dy; M.len2[i] = (R)l2; M.inv_len2[i] = (R)(l2 > 0 ? 1.0 / l2 : 0.0);
        M.poly_end[i] = (i + 1 == m.n_edges || m.poly[i + 1] != m.poly[i]) ? 1 : 0;
        mn_e = std::fmin(mn_e, std::fmin(m.x1[i], m.x2[i])); mx_e = std::fmax(mx_e, std::fmax(m.x1[i], m.x2[i]));
        mn_n = std::fmin(mn_n, std::fmin(m.y1[i], m.y2[i])); mx_n = std::fmax(mx_n, std::fmax(m.y1[i], m.y2[i]));
    }
    EnvP<R>& E = A.EP;
    std::memset(&E, 0, sizeof(E));
    if (m.n_edges == 0) { mn_n = mn_e = 0; mx_n = mx_e = 1; }
    E.min_n = (R)mn_n; E.max_n = (R)mx_n; E.min_e = (R)mn_e; E.max_e = (R)mx_e;
    double hl_max = 0;
    for (int r = 0; r < 2; ++r) {
        const double margin = c.ship[r].length_of_ship / 2;
        E.lo_n[r] = (R)(mn_n + margin); E.hi_n[r] = (R)(mx_n - margin);
        E.lo_e[r] = (R)(mn_e + margin); E.hi_e[r] = (R)(mx_e - margin);
        E.half_len[r] = (R)margin;
        hl_max = std::fmax(hl_max, margin);
    }
    E.arrive2 = largest_with_sqrt_le<double>(200.0);
    E.inv_max_north = (R)(1.0 / mx_n);
    E.e_tol = (R)1000; E.inv_e_tol = (R)(1.0 / 1000);
    E.rpm_max = (R)2000;
    E.collide2 = 50.0 * 50.0;
    const double nr = hl_max * std::sqrt(2.0) + 4.0;
    E.near2 = (R)(nr * nr);
    E.theta = c.theta; E.sampling_frequency = c.sampling_frequency;
}

static size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

static void build_layout(SitHandle_* h) {
    size_t off = 0;
    const bool f32 = h->cfg.precision == SIT_F32;
    auto add = [&](const FieldDef& d) {
        SitField f;
        std::memset(&f, 0, sizeof(f));
        std::snprintf(f.name, sizeof(f.name), "%s", d.name);
        f.dtype = d.kind == 'D' ? SIT_DT_F64 : d.kind == 'I' ? SIT_DT_I32 : d.kind == 'F' ? SIT_DT_F32 : (f32 ? SIT_DT_F32 : SIT_DT_F64);
        f.rows = d.rows < 0 ? h->cfg.max_wp : d.rows;
        f.count = d.env ? h->n_envs : h->n_ships;
        const size_t esz = f.dtype == SIT_DT_F64 ? 8 : 4;
        f.offset = (int64_t)off;
        off = align_up(off + esz * (size_t)f.rows * (size_t)f.count, 256);
        h->fields.push_back(f);
    };
    for (const FieldDef& d : kShipFields) add(d);
    if (h->cfg.kind == SIT_KIND_ENV)
        for (const FieldDef& d : kEnvFields) add(d);
    h->state_bytes = off;
}

static const SitField* find_field(const SitHandle_* h, const char* name) {
    for (const SitField& f : h->fields)
        if (std::strcmp(f.name, name) == 0) return &f;
    return nullptr;
}

template <typename T>
static T* fptr(const SitHandle_* h, void* state, const char* name) {
    const SitField* f = find_field(h, name);
    return f ? reinterpret_cast<T*>(static_cast<char*>(state) + f->offset) : nullptr;
}

template <typename R>
static ShipPtrs<R> ship_ptrs(const SitHandle_* h, void* st) {
    ShipPtrs<R> g;
    g.north = fptr<double>(h, st, "north"); g.east = fptr<double>(h, st, "east");
    g.init_north = fptr<double>(h, st, "init_north"); g.init_east = fptr<double>(h, st, "init_east");
#define F_(n) g.n = fptr<R>(h, st, #n)
    F_(psi); F_(u); F_(v); F_(r); F_(omega); F_(ei_speed); F_(ei_shaft); F_(ei_heading); F_(e_prev); F_(e_ct); F_(e_ct_int);
    F_(fuel_me); F_(fuel_el); F_(fuel); F_(u_sp); F_(vc_n); F_(vc_e); F_(wind_speed); F_(wind_dir);
    F_(init_psi); F_(init_u); F_(init_v); F_(init_r); F_(init_vc_n); F_(init_vc_e); F_(init_wind_speed); F_(init_wind_dir);
    F_(init_u_sp); F_(init_omega); F_(init_ei_shaft);
    F_(wp_n); F_(wp_e); F_(wp0_n); F_(wp0_e);
#undef F_
    g.k = fptr<int32_t>(h, st, "k"); g.n_wp = fptr<int32_t>(h, st, "n_wp"); g.mode = fptr<int32_t>(h, st, "mode");
    g.n_wp0 = fptr<int32_t>(h, st, "n_wp0"); g.steps = fptr<int32_t>(h, st, "steps");
    g.mc_idx = fptr<int32_t>(h, st, "mc_idx"); g.mc_n = fptr<float>(h, st, "mc_n"); g.mc_e = fptr<float>(h, st, "mc_e");
    g.mc_b2 = fptr<float>(h, st, "mc_b2");
    g.rel_n = fptr<float>(h, st, "rel_n"); g.rel_cn = fptr<float>(h, st, "rel_cn");
    g.rel_e = fptr<float>(h, st, "rel_e"); g.rel_ce = fptr<float>(h, st, "rel_ce");
    g.n_ships = h->n_ships; g.max_wp = h->cfg.max_wp; g._pad = 0;
    return g;
}

template <typename R>
static EnvPtrs<R> env_ptrs(const SitHandle_* h, void* st) {
    EnvPtrs<R> e;
    std::memset(&e, 0, sizeof(e));
    e.n_envs = h->n_envs;
    if (h->cfg.kind != SIT_KIND_ENV) return e;
    e.prev_n = fptr<double>(h, st, "prev_n"); e.prev_e = fptr<double>(h, st, "prev_e");
    e.eps_dist = fptr<double>(h, st, "eps_dist"); e.samp_dist = fptr<double>(h, st, "samp_dist");
    e.ab_seg = fptr<R>(h, st, "ab_seg"); e.act_n = fptr<R>(h, st, "act_n"); e.act_e = fptr<R>(h, st, "act_e");
    e.ep_return = fptr<R>(h, st, "ep_return");
    e.flags = fptr<int32_t>(h, st, "flags"); e.ep_steps = fptr<int32_t>(h, st, "ep_steps");
    e.ab_alpha = fptr<R>(h, st, "ab_alpha"); e.cur_obs = fptr<R>(h, st, "cur_obs");
    e.n_dec = fptr<int32_t>(h, st, "n_dec"); e.dec_serial = fptr<int32_t>(h, st, "dec_serial");
    e.tr_slot = fptr<int32_t>(h, st, "tr_slot");
    e.draw_t = fptr<double>(h, st, "draw_t");
    return e;
}

static int check_params(const SitShipParams& p, const char* who) {
    if (!(p.dt > 0)) return fail(SIT_E_INVALID, std::string(who) + ": dt must be > 0");
    if (p.n_modes < 1 || p.n_modes > SIT_MAX_MODES) return fail(SIT_E_INVALID, std::string(who) + ": n_modes out of range");
    for (int j = 0; j < p.n_modes; ++j)
        if (p.mode_shaft_generator_state[j] < 0 || p.mode_shaft_generator_state[j] > 2)
            return fail(SIT_E_INVALID, std::string(who) + ": bad shaft generator state");
    if (p.machinery_kind != SIT_MACH_FULL && p.machinery_kind != SIT_MACH_SIMPLIFIED)
        return fail(SIT_E_INVALID, std::string(who) + ": machinery_kind must be SIT_MACH_FULL or SIT_MACH_SIMPLIFIED");
    if (p.machinery_kind == SIT_MACH_SIMPLIFIED) {
        if (!(p.thrust_force_dynamic_time_constant > 0))
            return fail(SIT_E_INVALID, std::string(who) + ": thrust_force_dynamic_time_constant must be > 0");
        if (!(p.lookahead_distance > 0)) return fail(SIT_E_INVALID, std::string(who) + ": bad parameter");
        return SIT_OK;                                   // no shaft: propeller parameters are not read
    }
    if (!(p.lookahead_distance > 0) || !(p.propeller_inertia > 0)) return fail(SIT_E_INVALID, std::string(who) + ": bad parameter");
    return SIT_OK;
}

extern "C" {

int sit_abi_version(void) { return SIT_ABI_VERSION; }
const char* sit_last_error(void) { return g_err.c_str(); }

const char* sit_status_string(int32_t b) {
    static const char* k[12] = {
        "|Test ship reaches endpoint|", "|Test ship hits map horizon|", "|Test ship collides with the terrain|",
        "|Test ship mechanical failure|", "|Test ship navigation failure|", "|Test ship blackout failure|",
        "|Obstacle ship reaches endpoint|", "|Obstacle ship hits map horizon|", "|Obstacle ship collides with the terrain|",
        "|Obstacle ship IW sampled in terminal state|", "|Obstacle ship navigation failure|", "|Ship collision|"};
    return (b >= 0 && b < 12) ? k[b] : "";
}

int sit_create(const SitConfig* cfg, int64_t n_envs, int device, SitHandle* out) {
    if (!cfg || !out) return fail(SIT_E_INVALID, "null argument");
    if (cfg->abi_version != SIT_ABI_VERSION) return fail(SIT_E_INVALID, "SitConfig.abi_version mismatch");
    if (cfg->kind != SIT_KIND_FLEET && cfg->kind != SIT_KIND_ENV) return fail(SIT_E_INVALID, "bad kind");
    if (cfg->precision != SIT_F64 && cfg->precision != SIT_F32) return fail(SIT_E_INVALID, "bad precision");
    if (cfg->max_wp < 2 || cfg->max_wp > SIT_MAX_WP) return fail(SIT_E_INVALID, "max_wp out of range");
    if (n_envs < 1) return fail(SIT_E_INVALID, "n_envs must be >= 1");
    int rc = check_params(cfg->ship[0], "ship[0]");
    if (rc) return rc;
    if (cfg->kind == SIT_KIND_ENV) {
        rc = check_params(cfg->ship[1], "ship[1]");
        if (rc) return rc;
        if (cfg->ship[0].machinery_kind != SIT_MACH_FULL || cfg->ship[1].machinery_kind != SIT_MACH_FULL)
            return fail(SIT_E_INVALID, "SimplifiedMachineryModel is a fleet-only variant: MultiShipRLEnv passes the shaft speed to "
                                       "the throttle controller and reads the rpm (MSRL_env_ex.py:233-237,554-558)");
        if (cfg->map.n_edges < 0 || cfg->map.n_edges > SIT_MAX_EDGES) return fail(SIT_E_INVALID, "map.n_edges out of range");
        if (!(cfg->sampling_frequency > 0)) return fail(SIT_E_INVALID, "sampling_frequency must be > 0");
        int np_ = 0;
        for (int i = 0; i < cfg->map.n_edges; ++i)
            if (i + 1 == cfg->map.n_edges || cfg->map.poly[i + 1] != cfg->map.poly[i]) ++np_;
        if (np_ > 32) return fail(SIT_E_INVALID, "map has more than 32 polygons");
    }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        return fail(SIT_E_NODEVICE, "no CUDA device (libshipsim has no CPU fallback)");
    }
    if (device < 0 || device >= ndev) return fail(SIT_E_INVALID, "device index out of range");
    CUDA_TRY(cudaSetDevice(device));
    SitHandle_* h = new (std::nothrow) SitHandle_();
    if (!h) return fail(SIT_E_INVALID, "out of host memory");
    h->cfg = *cfg;
    h->n_envs = n_envs;
    h->ships_per_env = cfg->kind == SIT_KIND_ENV ? 2 : 1;
    h->n_ships = n_envs * h->ships_per_env;
    h->device = device;
    h->same_params = cfg->kind == SIT_KIND_FLEET || params_equal(cfg->ship[0], cfg->ship[1]);
    cudaDeviceGetAttribute(&h->sm_count, cudaDevAttrMultiProcessorCount, device);
    derive_ship(cfg->ship[0], h->pd[0]); derive_ship(cfg->ship[1].dt > 0 ? cfg->ship[1] : cfg->ship[0], h->pd[1]);
    derive_ship(cfg->ship[0], h->pf[0]); derive_ship(cfg->ship[1].dt > 0 ? cfg->ship[1] : cfg->ship[0], h->pf[1]);
    derive_shipx(h->pf[0], h->px);
    h->kd = nullptr; h->kf = nullptr;
    if (cfg->kind == SIT_KIND_ENV) {
        if (cfg->precision == SIT_F64) { h->kd = new EnvKArgs<double>(); derive_env(*cfg, *h->kd); }
        else { h->kf = new EnvKArgs<float>(); derive_env(*cfg, *h->kf); }
    }
    h->d_scratch = nullptr; h->scratch_bytes = 0; h->h_pinned = nullptr; h->pinned_bytes = 0;
    h->d_pol = nullptr; h->pol_bytes = 0;
    h->roll_graph = nullptr; h->tc_smem_set = 0; h->pipe_ready = false;
    build_layout(h);
    *out = h;
    return SIT_OK;
}

int sit_destroy(SitHandle h) {
    if (!h) return SIT_OK;
    if (h->d_scratch) cudaFree(h->d_scratch);
    if (h->d_pol) cudaFree(h->d_pol);
    if (h->h_pinned) cudaFreeHost(h->h_pinned);
    if (h->roll_graph) cudaGraphExecDestroy(h->roll_graph);
    if (h->pipe_ready) {
        for (int j = 0; j < SIT_PIPE; ++j) { cudaStreamDestroy(h->pipe[j]); cudaEventDestroy(h->pipe_done[j]); }
        for (int j = 0; j <= SIT_PIPE; ++j) cudaEventDestroy(h->cnt_ev[j]);
        cudaEventDestroy(h->pipe_in);
    }
    delete h->kd; delete h->kf;
    delete h;
    return SIT_OK;
}

int sit_state_bytes(SitHandle h, size_t* nbytes) {
    if (!h || !nbytes) return fail(SIT_E_INVALID, "null argument");
    *nbytes = h->state_bytes;
    return SIT_OK;
}

int64_t sit_n_ships(SitHandle h) { return h ? h->n_ships : 0; }

int sit_layout(SitHandle h, SitField* fields, int32_t capacity, int32_t* n_fields) {
    if (!h || !n_fields) return fail(SIT_E_INVALID, "null argument");
    *n_fields = (int32_t)h->fields.size();
    if (!fields) return SIT_OK;
    if (capacity < (int32_t)h->fields.size()) return fail(SIT_E_INVALID, "field array too small");
    std::memcpy(fields, h->fields.data(), h->fields.size() * sizeof(SitField));
    return SIT_OK;
}

static int grid_for(int64_t n, int block) { return (int)((n + block - 1) / block); }

}  // extern "C"
// fp32 fleets: every launch without per-step inputs / logs runs the packed kernel.  Returns false when the launch is not
// its case (fp64, per-step disturbances, 27-channel log, command stream).
template <typename R>
static bool fleet_x2(SitHandle, const ShipPtrs<R>&, int, const FleetIO<R>&, bool, bool, bool, cudaStream_t) { return false; }
template <>
bool fleet_x2<float>(SitHandle h, const ShipPtrs<float>& g, int K, const FleetIO<float>& k, bool quiet, bool fast, bool single,
                     cudaStream_t st) {
    const bool traj_all = quiet && k.traj && k.traj_steps == K && K > 1;
    if (!(traj_all || fast) || (k.first & 1)) return false;
    static const bool off = std::getenv("SIT_NO_X2") != nullptr;   // A/B switch for profiling runs
    if (off) return false;
    // CTA size: 128 threads (256 ships) for fleets of many waves; a shard of a strong-scaled fleet (131 072 ships on 148 SMs)
    // is a single wave, where the SM holding one CTA more than the others sets the time -- smaller CTAs even it out
    // (per-SM load = ceil(CTAs / SMs) x CTA size: 512 ships at 128 threads, 448 at 64 for 131 072 ships; ideal 442.8)
    static const int env_block = std::getenv("SIT_X2_BLOCK_RT") ? std::atoi(std::getenv("SIT_X2_BLOCK_RT")) : 0;
    int block = SIT_X2_BLOCK;
    const int64_t pairs = (k.count + 1) / 2;
    if (env_block == 32 || env_block == 64 || env_block == 128) block = env_block;
    else {
        int64_t best = -1;
        for (int b = SIT_X2_BLOCK; b >= 32; b >>= 1) {
            const int64_t ctas = (pairs + b - 1) / b, load = (ctas + h->sm_count - 1) / h->sm_count * b;
            if (best < 0 || load < best) { best = load; block = b; }
        }
    }
    const int grid = (int)((pairs + block - 1) / block);
    const x2::ShipPX& P = h->px;
#define X2_(A, M, T) fleet_step_x2_kernel<A, M, T><<<grid, block, 0, st>>>(P, g, K, k)
    if (traj_all) {
        if (k.avoid) { if (single) X2_(true, true, true); else X2_(true, false, true); }
        else { if (single) X2_(false, true, true); else X2_(false, false, true); }
    } else {
        if (k.avoid) { if (single) X2_(true, true, false); else X2_(true, false, false); }
        else { if (single) X2_(false, true, false); else X2_(false, false, false); }
    }
#undef X2_
    return true;
}

template <typename R>
static int fleet_step_t(SitHandle h, void* state, int K, const SitFleetIO* io, cudaStream_t st, int64_t first = 0,
                        int64_t count = -1, int8_t* k8 = nullptr) {
    FleetIO<R> k;
    std::memset(&k, 0, sizeof(k));
    k.store = 1; k.k8 = k8;
    k.first = first; k.count = count < 0 ? h->n_ships : count;
    if (io) {
        k.dist = static_cast<const R*>(io->dist); k.dist_steps = io->dist_steps; k.avoid = io->avoid; k.store = io->store; k.traj_steps = io->traj_steps;
        k.traj = static_cast<R*>(io->traj); k.log27 = static_cast<R*>(io->log27); k.cmd = static_cast<R*>(io->cmd);
        if (k.dist && !(k.dist_steps == 1 || k.dist_steps == K)) return fail(SIT_E_INVALID, "dist_steps must be 1 or K");
        if (!(k.traj_steps == 0 || k.traj_steps == 1 || k.traj_steps == K)) return fail(SIT_E_INVALID, "traj_steps must be 0, 1 or K");
    }
    const ShipPtrs<R> g = ship_ptrs<R>(h, state);
    const ShipP<R>& P = HostSel<R>::ship(h);
    // fp64: 96 registers per thread -> 2 CTAs of 256 (16 warps / SM) but 5 CTAs of 128 (20 warps / SM)
    static const int env_block = std::getenv("SIT_FLEET_BLOCK") ? std::atoi(std::getenv("SIT_FLEET_BLOCK")) : 0;
    const int block = (env_block >= 32 && env_block <= 256 && env_block % 32 == 0) ? env_block : (sizeof(R) == 8 ? 128 : 256);
    const int grid = grid_for(k.count, block);
    const bool quiet = (!k.dist || k.dist_steps == 1) && !k.log27 && !k.cmd && k.store;
    const bool fast = quiet && (!k.traj || k.traj_steps == 1);
    const bool single = P.n_modes == 1;
    if (P.mach == SIT_MACH_SIMPLIFIED) {   // SimplifiedMachineryModel: the general kernel only (a model variant, not a bench path)
        if (k.log27) fleet_step_kernel<R, true, SIT_MACH_SIMPLIFIED><<<grid, block, 0, st>>>(P, g, K, k);
        else fleet_step_kernel<R, false, SIT_MACH_SIMPLIFIED><<<grid, block, 0, st>>>(P, g, K, k);
    } else if (fleet_x2<R>(h, g, K, k, quiet, fast, single, st)) {
        // fp32: packed two-ships-per-thread kernel (ship_x2.cuh) launched
    } else if (quiet && k.traj && k.traj_steps == K && K > 1) {   // every-step trajectory on the specialised compute path
        if (k.avoid) {
            if (single) fleet_step_fast_kernel<R, true, true, true><<<grid, block, 0, st>>>(P, g, K, k);
            else fleet_step_fast_kernel<R, true, false, true><<<grid, block, 0, st>>>(P, g, K, k);
        } else {
            if (single) fleet_step_fast_kernel<R, false, true, true><<<grid, block, 0, st>>>(P, g, K, k);
            else fleet_step_fast_kernel<R, false, false, true><<<grid, block, 0, st>>>(P, g, K, k);
        }
    } else if (fast) {
        if (k.avoid) {
            if (single) fleet_step_fast_kernel<R, true, true><<<grid, block, 0, st>>>(P, g, K, k);
            else fleet_step_fast_kernel<R, true, false><<<grid, block, 0, st>>>(P, g, K, k);
        } else {
            if (single) fleet_step_fast_kernel<R, false, true><<<grid, block, 0, st>>>(P, g, K, k);
            else fleet_step_fast_kernel<R, false, false><<<grid, block, 0, st>>>(P, g, K, k);
        }
    } else if (k.log27) fleet_step_kernel<R, true><<<grid, block, 0, st>>>(P, g, K, k);
    else fleet_step_kernel<R, false><<<grid, block, 0, st>>>(P, g, K, k);
    CUDA_TRY(cudaGetLastError());
    return SIT_OK;
}
extern "C" {

int sit_fleet_step_k(SitHandle h, void* state, int32_t K, const SitFleetIO* io, void* stream) {
    if (!h || !state) return fail(SIT_E_INVALID, "null argument");
    if (K < 1) return fail(SIT_E_INVALID, "K must be >= 1");
    if (h->cfg.kind != SIT_KIND_FLEET) return fail(SIT_E_KIND, "sit_fleet_step_k needs a SIT_KIND_FLEET handle");
    CUDA_TRY(cudaSetDevice(h->device));
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    return h->cfg.precision == SIT_F64 ? fleet_step_t<double>(h, state, K, io, st) : fleet_step_t<float>(h, state, K, io, st);
}

static int ensure_pipe(SitHandle h);
static int ensure_scratch(SitHandle h, size_t total) {
    if (h->scratch_bytes >= total) return SIT_OK;
    if (h->d_scratch) cudaFree(h->d_scratch);
    h->d_scratch = nullptr; h->scratch_bytes = 0;
    CUDA_TRY(cudaMalloc(&h->d_scratch, total));
    h->scratch_bytes = total;
    return SIT_OK;
}

static int ensure_pol_scratch(SitHandle h, size_t total) {
    if (h->pol_bytes >= total) return SIT_OK;
    if (h->d_pol) cudaFree(h->d_pol);
    h->d_pol = nullptr; h->pol_bytes = 0;
    CUDA_TRY(cudaMalloc(&h->d_pol, total));
    h->pol_bytes = total;
    return SIT_OK;
}

// rows of pose_h selected by channel_mask, copied as runs of consecutive channels (one 2-D copy per run and chunk)
static int copy_pose_rows(char* pose_h, const char* d_pose, int mask, size_t first, size_t cnt, size_t S, size_t rs,
                          cudaStream_t st) {
    int out_row = 0;
    for (int c = 0; c < SIT_N_TRAJ;) {
        if (!((mask >> c) & 1)) { ++c; continue; }
        int run = 1;
        while (c + run < SIT_N_TRAJ && ((mask >> (c + run)) & 1)) ++run;
        CUDA_TRY(cudaMemcpy2DAsync(pose_h + ((size_t)out_row * S + first) * rs, S * rs, d_pose + ((size_t)c * S + first) * rs, S * rs,
                                   cnt * rs, run, cudaMemcpyDeviceToHost, st));
        out_row += run; c += run;
    }
    return SIT_OK;
}

// join the internal pipe streams back into the caller's stream (also on the error paths: the caller's stream must never
// run ahead of work that was forked from it)
static void join_pipes(SitHandle h, cudaStream_t st) {
    for (int j = 0; j < SIT_PIPE; ++j) {
        if (cudaEventRecord(h->pipe_done[j], h->pipe[j]) == cudaSuccess) cudaStreamWaitEvent(st, h->pipe_done[j], 0);
    }
}

int sit_fleet_step_host_sel(SitHandle h, void* state, int32_t K, const void* dist_h, int32_t avoid, void* pose_h,
                            int32_t channel_mask, int8_t* k8_h, void* stream) {
    if (!h || !state) return fail(SIT_E_INVALID, "null argument");
    if (h->cfg.kind != SIT_KIND_FLEET) return fail(SIT_E_KIND, "sit_fleet_step_host needs a SIT_KIND_FLEET handle");
    if (K < 1) return fail(SIT_E_INVALID, "K must be >= 1");
    if (pose_h && (channel_mask & 0xff) == 0) return fail(SIT_E_INVALID, "channel_mask selects no channel");
    channel_mask &= 0xff;
    CUDA_TRY(cudaSetDevice(h->device));
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const size_t S = (size_t)h->n_ships, rs = h->cfg.precision == SIT_F64 ? 8 : 4;
    const size_t o_dist = 0, o_pose = align_up(4 * S * rs, 256), o_k8 = o_pose + align_up(SIT_N_TRAJ * S * rs, 256),
                 total = o_k8 + align_up(S, 256);
    int rc = ensure_scratch(h, total);
    if (rc) return rc;
    char* d = static_cast<char*>(h->d_scratch);
    SitFleetIO io;
    std::memset(&io, 0, sizeof(io));
    io.dist = dist_h ? d + o_dist : nullptr; io.dist_steps = 1; io.avoid = avoid; io.store = 1;
    io.traj = pose_h ? d + o_pose : nullptr; io.traj_steps = 1;
    int8_t* d_k8 = k8_h ? reinterpret_cast<int8_t*>(d + o_k8) : nullptr;
    // Large fleets: the ships are cut into chunks that travel through SIT_PIPE internal streams, so the H2D copy of
    // chunk c+1 and the D2H copy of chunk c-1 run on the two copy engines while chunk c computes (ships are independent,
    // so a chunk is a complete unit of work).  Small fleets: one copy-in, one launch, one copy-out on the caller's stream.
    static const int env_chunks = std::getenv("SIT_HOST_CHUNKS") ? std::atoi(std::getenv("SIT_HOST_CHUNKS")) : 0;
    const int64_t min_chunk = env_chunks > 0 ? 16384 : 65536;
    int n_chunks = (int)std::min<int64_t>(env_chunks > 0 ? env_chunks : 16, (int64_t)S / min_chunk);   // 4 / 8 / 16 / 32 chunks at 1 M ships: 1.40 / 1.49 / 1.54 / 1.39e11 ship-steps/s
    if (n_chunks < 2 || (!dist_h && !pose_h && !k8_h)) {
        if (dist_h) CUDA_TRY(cudaMemcpyAsync(d + o_dist, dist_h, 4 * S * rs, cudaMemcpyHostToDevice, st));
        rc = h->cfg.precision == SIT_F64 ? fleet_step_t<double>(h, state, K, &io, st, 0, -1, d_k8)
                                         : fleet_step_t<float>(h, state, K, &io, st, 0, -1, d_k8);
        if (rc) return rc;
        if (pose_h && (rc = copy_pose_rows(static_cast<char*>(pose_h), d + o_pose, channel_mask, 0, S, S, rs, st))) return rc;
        if (k8_h) CUDA_TRY(cudaMemcpyAsync(k8_h, d_k8, S, cudaMemcpyDeviceToHost, st));
        CUDA_TRY(cudaStreamSynchronize(st));
        return SIT_OK;
    }
    if ((rc = ensure_pipe(h))) return rc;
    CUDA_TRY(cudaEventRecord(h->pipe_in, st));
    for (int j = 0; j < SIT_PIPE; ++j) CUDA_TRY(cudaStreamWaitEvent(h->pipe[j], h->pipe_in, 0));
    const int64_t chunk = ((int64_t)S / n_chunks + 255) / 256 * 256;
    const size_t pitch = S * rs;
    int c = 0;
    cudaError_t ce = cudaSuccess;
    for (int64_t first = 0; first < (int64_t)S && !rc && ce == cudaSuccess; first += chunk, ++c) {
        const int64_t cnt = std::min<int64_t>(chunk, (int64_t)S - first);
        cudaStream_t ps = h->pipe[c % SIT_PIPE];
        if (dist_h)
            ce = cudaMemcpy2DAsync(d + o_dist + first * rs, pitch, static_cast<const char*>(dist_h) + first * rs, pitch,
                                   cnt * rs, 4, cudaMemcpyHostToDevice, ps);
        if (ce != cudaSuccess) break;
        rc = h->cfg.precision == SIT_F64 ? fleet_step_t<double>(h, state, K, &io, ps, first, cnt, d_k8)
                                         : fleet_step_t<float>(h, state, K, &io, ps, first, cnt, d_k8);
        if (rc) break;
        if (pose_h) rc = copy_pose_rows(static_cast<char*>(pose_h), d + o_pose, channel_mask, (size_t)first, (size_t)cnt, S, rs, ps);
        if (!rc && k8_h) ce = cudaMemcpyAsync(k8_h + first, d_k8 + first, cnt, cudaMemcpyDeviceToHost, ps);
    }
    join_pipes(h, st);   // also after an error: the forked work is joined before the call returns
    if (rc) return rc;
    if (ce != cudaSuccess) return fail(SIT_E_CUDA, std::string("sit_fleet_step_host: ") + cudaGetErrorString(ce));
    CUDA_TRY(cudaStreamSynchronize(st));
    return SIT_OK;
}

int sit_fleet_step_host(SitHandle h, void* state, int32_t K, const void* dist_h, int32_t avoid, void* pose_h,
                        void* stream) {
    return sit_fleet_step_host_sel(h, state, K, dist_h, avoid, pose_h, 0xff, nullptr, stream);
}

int sit_capture_init(SitHandle h, void* state, void* stream) {
    if (!h || !state) return fail(SIT_E_INVALID, "null argument");
    CUDA_TRY(cudaSetDevice(h->device));
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int block = 256, grid = grid_for(h->n_ships, block);
    if (h->cfg.precision == SIT_F64)
        capture_init_kernel<double><<<grid, block, 0, st>>>(ship_ptrs<double>(h, state), env_ptrs<double>(h, state),
                                                            h->cfg.sampling_frequency, h->ships_per_env);
    else
        capture_init_kernel<float><<<grid, block, 0, st>>>(ship_ptrs<float>(h, state), env_ptrs<float>(h, state),
                                                           h->cfg.sampling_frequency, h->ships_per_env);
    CUDA_TRY(cudaGetLastError());
    return SIT_OK;
}

int sit_reset(SitHandle h, void* state, const uint8_t* mask, int32_t hard, void* stream) {
    if (!h || !state) return fail(SIT_E_INVALID, "null argument");
    CUDA_TRY(cudaSetDevice(h->device));
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int block = 256, grid = grid_for(h->n_ships, block);
    if (h->cfg.precision == SIT_F64)
        reset_kernel<double><<<grid, block, 0, st>>>(ship_ptrs<double>(h, state), env_ptrs<double>(h, state), mask, hard, h->ships_per_env);
    else
        reset_kernel<float><<<grid, block, 0, st>>>(ship_ptrs<float>(h, state), env_ptrs<float>(h, state), mask, hard, h->ships_per_env);
    CUDA_TRY(cudaGetLastError());
    return SIT_OK;
}

}  // extern "C"
template <typename R>
static int env_init_step_t(SitHandle h, void* state, const uint8_t* mask, cudaStream_t st) {
    const ShipPtrs<R> g = ship_ptrs<R>(h, state);
    const EnvKArgs<R>& A = HostSel<R>::env(h);
    const int block = 256, grid = grid_for(h->n_ships, block);
    if (h->same_params) env_init_step_kernel<R, true><<<grid, block, 0, st>>>(A.P[0], A.P[1], g, mask);
    else env_init_step_kernel<R, false><<<grid, block, 0, st>>>(A.P[0], A.P[1], g, mask);
    CUDA_TRY(cudaGetLastError());
    return SIT_OK;
}
extern "C" {

int sit_env_init_step(SitHandle h, void* state, const uint8_t* mask, void* stream) {
    if (!h || !state) return fail(SIT_E_INVALID, "null argument");
    if (h->cfg.kind != SIT_KIND_ENV) return fail(SIT_E_KIND, "sit_env_init_step needs a SIT_KIND_ENV handle");
    CUDA_TRY(cudaSetDevice(h->device));
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    return h->cfg.precision == SIT_F64 ? env_init_step_t<double>(h, state, mask, st) : env_init_step_t<float>(h, state, mask, st);
}

}  // extern "C"
// fp32 envs whose two roles share one parameter set CAN run the packed one-thread-per-env kernel (env_x2.cuh) unless the
// reward_parts / snapshot outputs are asked for: SIT_ENV_X2=1 (or -DSIT_ENV_X2_DEFAULT=1) switches it on.  Off by default:
// 38 % fewer instructions but no faster than the lane-pair kernel on the bench legs (measurements in env_x2.cuh)
#ifndef SIT_ENV_X2_DEFAULT
#define SIT_ENV_X2_DEFAULT 0
#endif
template <typename R, bool ROLL>
static bool env_x2(SitHandle, const ShipPtrs<R>&, const EnvPtrs<R>&, int, const EnvIO<R>&, const RollArgs&, cudaStream_t) { return false; }
template <bool ROLL>
static bool env_x2_f(SitHandle h, const ShipPtrs<float>& g, const EnvPtrs<float>& ge, int K, const EnvIO<float>& k, const RollArgs& ra,
                     cudaStream_t st) {
    const char* sw = std::getenv("SIT_ENV_X2");      // read per call: tests switch it inside one process
    const bool on = sw ? std::atoi(sw) != 0 : SIT_ENV_X2_DEFAULT != 0;
    if (!on || !h->same_params || k.rew_parts || k.snap || h->cfg.ship[0].machinery_kind != SIT_MACH_FULL) return false;
    static const int env_block = std::getenv("SIT_ENVX2_BLOCK") ? std::atoi(std::getenv("SIT_ENVX2_BLOCK")) : 0;
    const int block = (env_block == 32 || env_block == 64 || env_block == 128) ? env_block : 128, grid = grid_for(ra.env_count, block);
    if (h->pf[0].n_modes > 1) env_step_x2_kernel<ROLL, true><<<grid, block, 0, st>>>(*h->kf, h->px, g, ge, K, k, ra);
    else env_step_x2_kernel<ROLL, false><<<grid, block, 0, st>>>(*h->kf, h->px, g, ge, K, k, ra);
    return true;
}
template <>
bool env_x2<float, false>(SitHandle h, const ShipPtrs<float>& g, const EnvPtrs<float>& ge, int K, const EnvIO<float>& k,
                          const RollArgs& ra, cudaStream_t st) { return env_x2_f<false>(h, g, ge, K, k, ra, st); }
#ifdef SIT_ROLL_X2   // the rollout variant of the packed kernel: parity-green, slower at rollout sizes (see env_x2.cuh)
template <>
bool env_x2<float, true>(SitHandle h, const ShipPtrs<float>& g, const EnvPtrs<float>& ge, int K, const EnvIO<float>& k,
                         const RollArgs& ra, cudaStream_t st) { return env_x2_f<true>(h, g, ge, K, k, ra, st); }
#endif

template <typename R>
static int env_step_t(SitHandle h, void* state, int K, const SitEnvIO* io, cudaStream_t st, int64_t env_first = 0,
                      int64_t env_count = -1) {
    EnvIO<R> k;
    std::memset(&k, 0, sizeof(k));
    k.init = -1; k.rec_steps = 1; k.act_steps = 1;
    if (io) {
        k.action = static_cast<const R*>(io->action); k.sac_update = io->sac_update;
        k.act_steps = io->act_steps; k.init = io->init; k.auto_reset = io->auto_reset; k.rec_steps = io->rec_steps;
        k.obs = static_cast<R*>(io->obs); k.reward = static_cast<R*>(io->reward); k.done = io->done; k.status = io->status;
        k.reward_sum = static_cast<R*>(io->reward_sum); k.done_any = io->done_any; k.stats = io->stats;
        k.rew_parts = static_cast<R*>(io->reward_parts);
        k.snap = io->snap; k.draw_dt = h->cfg.ship[0].dt;
        if (k.snap && io->rec_steps != K) return fail(SIT_E_INVALID, "snap needs rec_steps == K (a snapshot is tied to its step's pose)");
        if (!(k.act_steps == 1 || k.act_steps == K)) return fail(SIT_E_INVALID, "act_steps must be 1 or K");
        if (!(k.rec_steps == 1 || k.rec_steps == K)) return fail(SIT_E_INVALID, "rec_steps must be 1 or K");
    }
    const ShipPtrs<R> g = ship_ptrs<R>(h, state);
    const EnvPtrs<R> ge = env_ptrs<R>(h, state);
    const EnvKArgs<R>& A = HostSel<R>::env(h);
    RollArgs ra;
    std::memset(&ra, 0, sizeof(ra));
    ra.env_first = env_first; ra.env_count = env_count < 0 ? h->n_envs : env_count;
    const int block = 128, grid = grid_for(2 * ra.env_count, block);
    if (env_x2<R, false>(h, g, ge, K, k, ra, st)) {
        // packed kernel launched
    } else if (k.rew_parts || k.snap) {
        if (h->same_params) env_step_kernel<R, true, false, true><<<grid, block, 0, st>>>(A, g, ge, K, k, ra);
        else env_step_kernel<R, false, false, true><<<grid, block, 0, st>>>(A, g, ge, K, k, ra);
    } else if (h->same_params) env_step_kernel<R, true, false><<<grid, block, 0, st>>>(A, g, ge, K, k, ra);
    else env_step_kernel<R, false, false><<<grid, block, 0, st>>>(A, g, ge, K, k, ra);
    CUDA_TRY(cudaGetLastError());
    return SIT_OK;
}
static int roll_args(SitHandle h, const SitRollout* ro, RollArgs& ra) {
    if (!ro || !ro->replay || !ro->counters) return fail(SIT_E_INVALID, "SitRollout needs replay and counters");
    if (ro->capacity < h->n_envs) return fail(SIT_E_INVALID, "replay capacity must be >= n_envs");
    if (ro->capacity > 0x7fffffffLL) return fail(SIT_E_INVALID, "replay capacity too large");
    std::memset(&ra, 0, sizeof(ra));
    ra.replay = ro->replay; ra.capacity = ro->capacity; ra.counters = ro->counters;
    ra.seed = ro->seed; ra.env_id_offset = ro->env_id_offset;
    ra.max_episode_steps = ro->max_episode_steps; ra.max_route_resampling = ro->max_route_resampling;
    ra.truncate = ro->truncate_episodes != 0;
    ra.sampling_frequency = (int32_t)h->cfg.sampling_frequency;
    ra.env_first = 0; ra.env_count = h->n_envs;
    return SIT_OK;
}

template <typename R>
static int rollout_step_t(SitHandle h, void* state, int K, const RollArgs& ra, const SitEnvIO* io, cudaStream_t st) {
    EnvIO<R> k;
    std::memset(&k, 0, sizeof(k));
    k.init = -1; k.rec_steps = 1; k.act_steps = 1; k.auto_reset = 1;
    if (io) {
        if (io->rec_steps > 1) return fail(SIT_E_INVALID, "sit_rollout_step_k records the last step only (rec_steps <= 1)");
        k.auto_reset = io->auto_reset == 2 ? 2 : 1;
        k.obs = static_cast<R*>(io->obs); k.reward = static_cast<R*>(io->reward); k.done = io->done; k.status = io->status;
        k.reward_sum = static_cast<R*>(io->reward_sum); k.done_any = io->done_any; k.stats = io->stats;
    }
    const ShipPtrs<R> g = ship_ptrs<R>(h, state);
    const EnvPtrs<R> ge = env_ptrs<R>(h, state);
    const EnvKArgs<R>& A = HostSel<R>::env(h);
    const int block = 128, grid = grid_for(2 * ra.env_count, block);
    if (env_x2<R, true>(h, g, ge, K, k, ra, st)) {
        // packed kernel launched
    } else if (h->same_params) env_step_kernel<R, true, true><<<grid, block, 0, st>>>(A, g, ge, K, k, ra);
    else env_step_kernel<R, false, true><<<grid, block, 0, st>>>(A, g, ge, K, k, ra);
    CUDA_TRY(cudaGetLastError());
    return SIT_OK;
}

extern "C" {

int sit_rollout_begin(SitHandle h, void* state, void* stream) {
    if (!h || !state) return fail(SIT_E_INVALID, "null argument");
    if (h->cfg.kind != SIT_KIND_ENV) return fail(SIT_E_KIND, "sit_rollout_begin needs a SIT_KIND_ENV handle");
    CUDA_TRY(cudaSetDevice(h->device));
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int block = 256, grid = grid_for(h->n_envs, block);
    if (h->cfg.precision == SIT_F64) rollout_begin_kernel<double><<<grid, block, 0, st>>>(ship_ptrs<double>(h, state), env_ptrs<double>(h, state));
    else rollout_begin_kernel<float><<<grid, block, 0, st>>>(ship_ptrs<float>(h, state), env_ptrs<float>(h, state));
    CUDA_TRY(cudaGetLastError());
    return SIT_OK;
}

}  // extern "C"

#define SIT_ROLL_GROUPS_MAX 8
static int policy_args(const SitPolicy* pol, PolicyArgs& pa, bool& tensor_cores) {
    const int base_mode = pol->mode & 0xff;
    tensor_cores = (pol->mode & SIT_POLICY_TC_BF16) != 0;
    if ((pol->mode & ~(0xff | SIT_POLICY_TC_BF16)) || base_mode > SIT_POLICY_MEAN) return fail(SIT_E_INVALID, "bad policy mode");
    std::memset(&pa, 0, sizeof(pa));
    pa.mode = base_mode; pa.hidden = pol->hidden;
    if (base_mode != SIT_POLICY_UNIFORM) {
        if (pol->hidden < 32 || pol->hidden > 256 || pol->hidden % 32) return fail(SIT_E_INVALID, "policy hidden width must be a multiple of 32 in [32, 256]");
        if (!pol->w0 || !pol->b0 || !pol->w1 || !pol->b1 || !pol->w2 || !pol->b2) return fail(SIT_E_INVALID, "policy weights missing");
        pa.w0 = pol->w0; pa.b0 = pol->b0; pa.w1 = pol->w1; pa.b1 = pol->b1; pa.w2 = pol->w2; pa.b2 = pol->b2;
        pa.obs_scale = pol->obs_scale;
    } else {
        tensor_cores = false;
    }
    return SIT_OK;
}

// Policy scratch: per-group header [SIT_ROLL_GROUPS_MAX][8 words: n_pending, n_retry, ticket, -, slot_base (u64), -, -] | env list [B] |
// obs [B][10] | out [B][2] | retry list [B] | packed W1 | per-CTA counts / offsets [B / 256 + groups].
// The lists are sized for the whole handle; an env group [first, first + count) uses the slices that start at `first`,
// so groups that run concurrently never share a row.
static int pol_scratch(SitHandle h, int group, int64_t first, PolScratch& sc) {
    const size_t B = (size_t)h->n_envs;
    const size_t o_env = 256, o_obs = align_up(o_env + B * 4, 256), o_out = align_up(o_obs + B * SIT_N_OBS * 4, 256),
                 o_retry = align_up(o_out + B * 2 * 4, 256), o_w1p = align_up(o_retry + B * 4, 256),
                 o_cta = o_w1p + 256 * 256 * 2, total = o_cta + (B / 256 + SIT_ROLL_GROUPS_MAX + 1) * 4;
    int rc = ensure_pol_scratch(h, total);
    if (rc) return rc;
    char* d = static_cast<char*>(h->d_pol);
    sc.n_pending = reinterpret_cast<unsigned int*>(d) + 8 * group; sc.n_retry = sc.n_pending + 1; sc.ticket = sc.n_pending + 2;
    sc.slot_base = reinterpret_cast<unsigned long long*>(sc.n_pending + 4);
    sc.cta_cnt = reinterpret_cast<int32_t*>(d + o_cta) + first / 256 + group;   // group starts are multiples of 256 envs
    sc.env = reinterpret_cast<int32_t*>(d + o_env) + first; sc.retry = reinterpret_cast<int32_t*>(d + o_retry) + first;
    sc.obs = reinterpret_cast<float*>(d + o_obs) + first * SIT_N_OBS; sc.out = reinterpret_cast<float*>(d + o_out) + first * 2;
    sc.w1p = reinterpret_cast<unsigned char*>(d + o_w1p);
    return SIT_OK;
}

static int pack_w1(SitHandle h, const PolicyArgs& pa, cudaStream_t st) {
    PolScratch sc;
    int rc = pol_scratch(h, 0, 0, sc);
    if (rc) return rc;
    const int H = pa.hidden;
    rollout_pack_w1_kernel<<<(H * (H >> 3) + 255) / 256, 256, 0, st>>>(pa, sc);
    CUDA_TRY(cudaGetLastError());
    return SIT_OK;
}

// gather -> MLP -> sample -> resample for the env range of `ra`; the packed W1 image must already be in the scratch
template <typename R>
static int rollout_policy_t(SitHandle h, void* state, const PolicyArgs& pa, bool tensor_cores, const RollArgs& ra, int group,
                            cudaStream_t st, cudaEvent_t count_after = nullptr, cudaEvent_t count_done = nullptr) {
    PolScratch sc;
    int rc = pol_scratch(h, group, ra.env_first, sc);
    if (rc) return rc;
    CUDA_TRY(cudaMemsetAsync(sc.n_pending, 0, 4 * sizeof(unsigned int), st));
    const ShipPtrs<R> g = ship_ptrs<R>(h, state);
    const EnvPtrs<R> ge = env_ptrs<R>(h, state);
    const EnvKArgs<R>& A = HostSel<R>::env(h);
    const int gather_grid = grid_for(ra.env_count, 256);
    // persistent: tiles / warps stride over the pending list (a small share of the group's envs)
    const int work_grid = h->sm_count * 4;
    // chained collect: this group's rows are reserved after the previous group's (count_after) and before the next one's
    if (count_after) CUDA_TRY(cudaStreamWaitEvent(st, count_after, 0));
    rollout_count_kernel<R><<<gather_grid, 256, 0, st>>>(ge, ra, sc);
    if (count_done) CUDA_TRY(cudaEventRecord(count_done, st));
    rollout_gather_kernel<R><<<gather_grid, 256, 0, st>>>(ge, ra, sc, pa.obs_scale);
    if (pa.mode != SIT_POLICY_UNIFORM) {
        if (tensor_cores) {
            const int H = pa.hidden;
            const size_t smem = (size_t)TC_M * H * 2 + (size_t)H * H * 2 + (size_t)(3 * H + 2 * TC_M) * 4 + 32;
            int cols = 32;
            while (cols < H) cols <<= 1;
            if (h->tc_smem_set < smem) {
                CUDA_TRY(cudaFuncSetAttribute(rollout_mlp_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
                h->tc_smem_set = smem;
            }
            const int tiles = (int)std::min<int64_t>(h->sm_count, (ra.env_count + TC_M - 1) / TC_M);
            rollout_mlp_tc_kernel<<<tiles, 256, smem, st>>>(pa, sc, cols);
        } else {
            rollout_mlp_kernel<<<work_grid, 256, 0, st>>>(pa, sc);
        }
    }
    rollout_sample_kernel<R><<<work_grid, 256, 0, st>>>(A, g, ge, pa, ra, sc);
    if (pa.mode != SIT_POLICY_MEAN && ra.max_route_resampling >= 32)
        rollout_resample_kernel<R><<<work_grid, 256, 0, st>>>(A, g, ge, pa, ra, sc);
    CUDA_TRY(cudaGetLastError());
    return SIT_OK;
}

extern "C" {

int sit_rollout_policy(SitHandle h, void* state, const SitPolicy* pol, const SitRollout* ro, void* stream) {
    if (!h || !state || !pol) return fail(SIT_E_INVALID, "null argument");
    if (h->cfg.kind != SIT_KIND_ENV) return fail(SIT_E_KIND, "sit_rollout_policy needs a SIT_KIND_ENV handle");
    PolicyArgs pa;
    bool tensor_cores = false;
    int rc = policy_args(pol, pa, tensor_cores);
    if (rc) return rc;
    RollArgs ra;
    rc = roll_args(h, ro, ra);
    if (rc) return rc;
    CUDA_TRY(cudaSetDevice(h->device));
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (tensor_cores) {
        rc = pack_w1(h, pa, st);
        if (rc) return rc;
    }
    return h->cfg.precision == SIT_F64 ? rollout_policy_t<double>(h, state, pa, tensor_cores, ra, 0, st)
                                       : rollout_policy_t<float>(h, state, pa, tensor_cores, ra, 0, st);
}

int sit_rollout_step_k(SitHandle h, void* state, int32_t K, const SitRollout* ro, const SitEnvIO* io, void* stream) {
    if (!h || !state) return fail(SIT_E_INVALID, "null argument");
    if (K < 1) return fail(SIT_E_INVALID, "K must be >= 1");
    if (h->cfg.kind != SIT_KIND_ENV) return fail(SIT_E_KIND, "sit_rollout_step_k needs a SIT_KIND_ENV handle");
    RollArgs ra;
    int rc = roll_args(h, ro, ra);
    if (rc) return rc;
    CUDA_TRY(cudaSetDevice(h->device));
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    return h->cfg.precision == SIT_F64 ? rollout_step_t<double>(h, state, K, ra, io, st) : rollout_step_t<float>(h, state, K, ra, io, st);
}

}  // extern "C"

static int ensure_pipe(SitHandle h) {
    if (h->pipe_ready) return SIT_OK;
    for (int j = 0; j < SIT_PIPE; ++j) {
        CUDA_TRY(cudaStreamCreateWithFlags(&h->pipe[j], cudaStreamNonBlocking));
        CUDA_TRY(cudaEventCreateWithFlags(&h->pipe_done[j], cudaEventDisableTiming));
    }
    CUDA_TRY(cudaEventCreateWithFlags(&h->pipe_in, cudaEventDisableTiming));
    for (int j = 0; j <= SIT_PIPE; ++j) CUDA_TRY(cudaEventCreateWithFlags(&h->cnt_ev[j], cudaEventDisableTiming));
    h->pipe_ready = true;
    return SIT_OK;
}

#ifndef SIT_ROLLOUT_GROUPS_DEFAULT
#define SIT_ROLLOUT_GROUPS_DEFAULT 0   /* 0: by size (below) */
#endif
// Number of env groups one rollout iteration is cut into (SIT_ROLLOUT_GROUPS overrides).  The policy chain of a group
// (gather -> MLP -> sample -> resample) is short, dependent and touches only the few envs at a decision event, so on its
// own it leaves the GPU almost idle; with several groups in flight the chain of one group runs under the env-step kernel of
// the others.  Envs are independent and every draw is keyed by (global env id, decision serial, try), so the grouping
// does not change any env's trajectory.
static int rollout_groups(SitHandle h) {
    // measured on B200, 65 536 envs, K = 32 (chained collect, tools/exp_groups.sh): 1 group 287 us per iteration, 2: 267,
    // 3: 261, 4: 263, 5: 267; a per-kernel launch priority for the policy kernels changed nothing
    int g = SIT_ROLLOUT_GROUPS_DEFAULT > 0 ? SIT_ROLLOUT_GROUPS_DEFAULT : (h->n_envs >= 49152 ? 3 : (h->n_envs >= 32768 ? 2 : 1));
    if (const char* e = std::getenv("SIT_ROLLOUT_GROUPS")) g = std::atoi(e);
    g = std::max(1, std::min(g, std::min(SIT_ROLL_GROUPS_MAX, SIT_PIPE + 1)));
    while (g > 1 && h->n_envs / g < 256) --g;
    return g;
}

// one iteration = policy + K steps for every env as ONE group on the caller's stream (several groups: rollout_chains)
static int rollout_iteration(SitHandle h, void* state, int K, const PolicyArgs& pa, bool tensor_cores, const RollArgs& ra,
                             const SitEnvIO* io, cudaStream_t st) {
    int rc = SIT_OK;
    if (tensor_cores && (rc = pack_w1(h, pa, st))) return rc;
    if (h->cfg.precision == SIT_F64) {
        rc = rollout_policy_t<double>(h, state, pa, tensor_cores, ra, 0, st);
        if (!rc) rc = rollout_step_t<double>(h, state, K, ra, io, st);
    } else {
        rc = rollout_policy_t<float>(h, state, pa, tensor_cores, ra, 0, st);
        if (!rc) rc = rollout_step_t<float>(h, state, K, ra, io, st);
    }
    return rc;
}

// `iterations` x (policy + K steps) as `groups` INDEPENDENT CHAINS, joined only at the end of the call.  Inside one iteration
// the groups cannot overlap much -- every group runs its short policy chain at the same time, then every group wants the
// whole GPU for its env-step kernel -- so joining after each iteration (rollout_iteration) gains nothing.  Left to run
// freely the chains fall out of phase after the first iteration: while one group's env-step kernel fills the SMs, the
// other group's policy chain (few small CTAs: <= 64 registers each, they fit beside three resident env-step CTAs) runs in
// the slots that are left, and the partial last wave of one step kernel is filled by the next one.  Per env the
// sequence policy, K steps, policy, ... is unchanged.  Replay rows: the slot reservations (rollout_count_kernel) are
// ordered group 0, 1, .., G-1, group 0 of the next iteration, ... by events, so the ring is filled in one fixed order.
static int rollout_chains(SitHandle h, void* state, int iterations, int K, const PolicyArgs& pa, bool tensor_cores,
                          const RollArgs& ra0, const SitEnvIO* io, int groups, cudaStream_t st) {
    int rc = SIT_OK;
    if (tensor_cores && (rc = pack_w1(h, pa, st))) return rc;      // the weights do not change inside one call
    const int64_t B = h->n_envs, per = ((B + groups - 1) / groups + 255) / 256 * 256;
    CUDA_TRY(cudaEventRecord(h->pipe_in, st));
    for (int gi = 1; gi < groups; ++gi) CUDA_TRY(cudaStreamWaitEvent(h->pipe[gi - 1], h->pipe_in, 0));
    cudaEvent_t prev = nullptr;
    // SIT_ROLLOUT_TRACE=1 (plain launches only): timed events around every policy chain and step kernel, printed as a
    // per-group timeline [us since the start of the call] on stderr after a synchronise -- a debugging aid, not a product path
    cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
    cudaStreamIsCapturing(st, &cap);
    const bool trace = cap == cudaStreamCaptureStatusNone && std::getenv("SIT_ROLLOUT_TRACE") != nullptr;
    std::vector<cudaEvent_t> tev;
    auto mark = [&](cudaStream_t s_) {
        if (!trace) return;
        cudaEvent_t e; cudaEventCreate(&e); cudaEventRecord(e, s_); tev.push_back(e);
    };
    mark(st);
    for (int it = 0; it < iterations; ++it) {
        for (int gi = 0; gi < groups; ++gi) {
            RollArgs ra = ra0;
            ra.env_first = std::min<int64_t>(B, gi * per);
            ra.env_count = std::min<int64_t>(B, (gi + 1) * per) - ra.env_first;
            if (ra.env_count <= 0) continue;
            cudaStream_t gs = gi == 0 ? st : h->pipe[gi - 1];
            if (h->cfg.precision == SIT_F64) {
                rc = rollout_policy_t<double>(h, state, pa, tensor_cores, ra, gi, gs, prev, h->cnt_ev[gi]);
                if (!rc) rc = rollout_step_t<double>(h, state, K, ra, io, gs);
            } else {
                mark(gs);
                rc = rollout_policy_t<float>(h, state, pa, tensor_cores, ra, gi, gs, prev, h->cnt_ev[gi]);
                mark(gs);
                if (!rc) rc = rollout_step_t<float>(h, state, K, ra, io, gs);
                mark(gs);
            }
            if (rc) break;
            prev = h->cnt_ev[gi];
        }
        if (rc) break;
    }
    for (int gi = 1; gi < groups; ++gi) {   // also after an error: forked work is joined before the call returns
        cudaError_t e1 = cudaEventRecord(h->pipe_done[gi - 1], h->pipe[gi - 1]);
        cudaError_t e2 = e1 == cudaSuccess ? cudaStreamWaitEvent(st, h->pipe_done[gi - 1], 0) : e1;
        if (e2 != cudaSuccess && !rc) rc = fail(SIT_E_CUDA, std::string("rollout_chains join: ") + cudaGetErrorString(e2));
    }
    if (trace && !tev.empty()) {
        mark(st);
        cudaStreamSynchronize(st);
        float total = 0.f;
        cudaEventElapsedTime(&total, tev.front(), tev.back());
        std::fprintf(stderr, "[sit trace] %d iterations x %d groups, K = %d: %.1f us\n", iterations, groups, K, total * 1e3f);
        if (h->cfg.precision != SIT_F64 && tev.size() == (size_t)(2 + 3 * iterations * groups))
            for (int it = 0; it < iterations; ++it)
                for (int gi = 0; gi < groups; ++gi) {
                    float a = 0.f, b = 0.f, c = 0.f;
                    const size_t j = 1 + 3 * ((size_t)it * groups + gi);
                    cudaEventElapsedTime(&a, tev.front(), tev[j]); cudaEventElapsedTime(&b, tev.front(), tev[j + 1]);
                    cudaEventElapsedTime(&c, tev.front(), tev[j + 2]);
                    std::fprintf(stderr, "[sit trace] it %2d group %d: policy %7.1f .. %7.1f  step .. %7.1f\n", it, gi, a * 1e3f, b * 1e3f, c * 1e3f);
                }
        for (cudaEvent_t e : tev) cudaEventDestroy(e);
    }
    return rc;
}

#define SIT_CHAIN_MAX_ITERS 64   /* iterations per graph of a chained sit_rollout_collect call */
extern "C" {

int sit_rollout_group_count(SitHandle h, int32_t* groups) {
    if (!h || !groups) return fail(SIT_E_INVALID, "null argument");
    if (h->cfg.kind != SIT_KIND_ENV) return fail(SIT_E_KIND, "sit_rollout_group_count needs a SIT_KIND_ENV handle");
    *groups = rollout_groups(h);
    return SIT_OK;
}

// `iterations` x (sit_rollout_policy, sit_rollout_step_k(K)).  One env group (rollout_groups above): one iteration is captured
// once into a CUDA graph and replayed, one graph launch per iteration instead of ~7 launches through the caller's language
// binding.  Several groups: the whole call is one graph of independent per-group chains (rollout_chains), captured on the
// second use of a set of arguments (the first use runs as plain launches and sizes the scratch).  Either graph is
// re-captured when any argument changes.  use_graph == 0: plain launches.
int sit_rollout_collect(SitHandle h, void* state, int32_t iterations, int32_t K, const SitPolicy* pol, const SitRollout* ro,
                        const SitEnvIO* io, int32_t use_graph, void* stream) {
    if (!h || !state || !pol || !ro) return fail(SIT_E_INVALID, "null argument");
    if (iterations < 0 || K < 1) return fail(SIT_E_INVALID, "iterations must be >= 0 and K >= 1");
    if (h->cfg.kind != SIT_KIND_ENV) return fail(SIT_E_KIND, "sit_rollout_collect needs a SIT_KIND_ENV handle");
    if (io && io->rec_steps > 1) return fail(SIT_E_INVALID, "sit_rollout_collect records the last step only (rec_steps <= 1)");
    if (iterations == 0) return SIT_OK;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    CUDA_TRY(cudaSetDevice(h->device));
    PolicyArgs pa;
    bool tensor_cores = false;
    int rc = policy_args(pol, pa, tensor_cores);
    if (rc) return rc;
    RollArgs ra;
    rc = roll_args(h, ro, ra);
    if (rc) return rc;
    const int groups = rollout_groups(h);
    if (groups > 1 && (rc = ensure_pipe(h))) return rc;
    cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
    if (st != nullptr) CUDA_TRY(cudaStreamIsCapturing(st, &cs));
    if (!use_graph || st == nullptr || cs != cudaStreamCaptureStatusNone) {   // legacy stream cannot be captured
        if (groups > 1) return rollout_chains(h, state, iterations, K, pa, tensor_cores, ra, io, groups, st);
        for (int it = 0; it < iterations; ++it)
            if ((rc = rollout_iteration(h, state, K, pa, tensor_cores, ra, io, st))) return rc;
        return SIT_OK;
    }
    if (groups > 1 && iterations > SIT_CHAIN_MAX_ITERS) {
        // a long chained call: whole chunks of SIT_CHAIN_MAX_ITERS iterations (one graph, reused), the rest as plain launches --
        // the graph of a call holds ~7 nodes per group and iteration
        for (int left = iterations; left > 0;) {
            const int n = std::min<int>(left, SIT_CHAIN_MAX_ITERS);
            rc = n == SIT_CHAIN_MAX_ITERS ? sit_rollout_collect(h, state, n, K, pol, ro, io, use_graph, stream)
                                          : rollout_chains(h, state, n, K, pa, tensor_cores, ra, io, groups, st);
            if (rc) return rc;
            left -= n;
        }
        return SIT_OK;
    }
    std::string key;
    const int32_t chain_iters = groups > 1 ? iterations : 0;   // a chained call is one graph of all its iterations
    key.append(reinterpret_cast<const char*>(&chain_iters), sizeof(chain_iters));
    key.append(reinterpret_cast<const char*>(&state), sizeof(state));
    key.append(reinterpret_cast<const char*>(&K), sizeof(K));
    key.append(reinterpret_cast<const char*>(&groups), sizeof(groups));
    key.append(reinterpret_cast<const char*>(pol), sizeof(*pol));
    key.append(reinterpret_cast<const char*>(ro), sizeof(*ro));
    SitEnvIO io0;
    std::memset(&io0, 0, sizeof(io0));
    if (io) io0 = *io;
    key.append(reinterpret_cast<const char*>(&io0), sizeof(io0));
    if (groups > 1) {
        if (!h->roll_graph || key != h->roll_key) {
            if (h->roll_graph) { cudaGraphExecDestroy(h->roll_graph); h->roll_graph = nullptr; h->roll_key.clear(); }
            // first use of these arguments: plain launches (validates, sizes the scratch); second use: captured
            if (key != h->roll_key_seen) {
                h->roll_key_seen = key;
                return rollout_chains(h, state, iterations, K, pa, tensor_cores, ra, io, groups, st);
            }
            CUDA_TRY(cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
            rc = rollout_chains(h, state, iterations, K, pa, tensor_cores, ra, io, groups, st);
            cudaGraph_t graph = nullptr;
            cudaError_t ce = cudaStreamEndCapture(st, &graph);
            if (rc || ce != cudaSuccess) {
                if (graph) cudaGraphDestroy(graph);
                return rc ? rc : fail(SIT_E_CUDA, std::string("graph capture: ") + cudaGetErrorString(ce));
            }
            ce = cudaGraphInstantiate(&h->roll_graph, graph, 0);
            cudaGraphDestroy(graph);
            if (ce != cudaSuccess) { h->roll_graph = nullptr; return fail(SIT_E_CUDA, std::string("graph instantiate: ") + cudaGetErrorString(ce)); }
            h->roll_key = key;
        }
        CUDA_TRY(cudaGraphLaunch(h->roll_graph, st));
        return SIT_OK;
    }
    if (!h->roll_graph || key != h->roll_key) {
        if (h->roll_graph) { cudaGraphExecDestroy(h->roll_graph); h->roll_graph = nullptr; }
        // one eager iteration first: validates the arguments and sizes the scratch outside the capture
        if ((rc = rollout_iteration(h, state, K, pa, tensor_cores, ra, io, st))) return rc;
        if (--iterations == 0) return SIT_OK;
        CUDA_TRY(cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
        rc = rollout_iteration(h, state, K, pa, tensor_cores, ra, io, st);
        cudaGraph_t graph = nullptr;
        cudaError_t ce = cudaStreamEndCapture(st, &graph);
        if (rc || ce != cudaSuccess) {
            if (graph) cudaGraphDestroy(graph);
            return rc ? rc : fail(SIT_E_CUDA, std::string("graph capture: ") + cudaGetErrorString(ce));
        }
        ce = cudaGraphInstantiate(&h->roll_graph, graph, 0);
        cudaGraphDestroy(graph);
        if (ce != cudaSuccess) { h->roll_graph = nullptr; return fail(SIT_E_CUDA, std::string("graph instantiate: ") + cudaGetErrorString(ce)); }
        h->roll_key = key;
    }
    for (int it = 0; it < iterations; ++it) CUDA_TRY(cudaGraphLaunch(h->roll_graph, st));
    return SIT_OK;
}

}  // extern "C"
// =================================================================================================
// Replay exchange: the rows of a rollout segment leave the sampler ring as ONE fixed-capacity block per rank (header row +
// block_rows rows) and are appended, rank block after rank block, to a replay memory.  Sizes are host constants, counts
// live in the block headers, so neither side synchronises with the host or exchanges counts.
// =================================================================================================
__global__ void __launch_bounds__(256) replay_pack_kernel(const float4* __restrict__ ring, const int64_t capacity,
                                                          const unsigned long long* __restrict__ closed,
                                                          const long long* __restrict__ seen, float4* __restrict__ block,
                                                          const int64_t block_rows) {
    const long long n_now = (long long)*closed, fresh = n_now - *seen;
    const long long n = fresh < block_rows ? fresh : block_rows;          // a too-small block keeps the newest rows
    const long long first = n_now - n;
    const int64_t items = n * (SIT_TR_FLOATS / 4);
    for (int64_t it = (int64_t)blockIdx.x * 256 + threadIdx.x; it < items; it += (int64_t)gridDim.x * 256) {
        const int64_t j = it / (SIT_TR_FLOATS / 4), q = it % (SIT_TR_FLOATS / 4);
        block[(1 + j) * (SIT_TR_FLOATS / 4) + q] = ring[((first + j) % capacity) * (SIT_TR_FLOATS / 4) + q];
    }
}
__global__ void replay_pack_finish_kernel(const unsigned long long* closed, long long* seen, long long* dropped, float* block,
                                          const int64_t block_rows) {
    const long long n_now = (long long)*closed, fresh = n_now - *seen;
    const long long n = fresh < block_rows ? fresh : block_rows;
    reinterpret_cast<long long*>(block)[0] = n;                           // header row: valid rows of this block
    if (fresh > block_rows && dropped) *dropped += fresh - block_rows;
    *seen = n_now;
}
// blocks: [n_blocks][1 + block_rows][24].  Row j of block r lands at ring position (count + sum of the earlier blocks' counts
// + j) mod capacity: rank order, row order, deterministic.  If more rows arrive than the ring holds only the newest land.
__global__ void __launch_bounds__(256) replay_append_kernel(float4* __restrict__ dst, const int64_t capacity,
                                                            const unsigned long long* __restrict__ count,
                                                            const float4* __restrict__ blocks, const int n_blocks,
                                                            const int64_t block_rows) {
    __shared__ long long s_pre[65];
    if (threadIdx.x == 0) {
        long long run = 0;
        for (int r = 0; r < n_blocks; ++r) {
            s_pre[r] = run;
            long long c = reinterpret_cast<const long long*>(blocks + (int64_t)r * (1 + block_rows) * (SIT_TR_FLOATS / 4))[0];
            run += c < 0 ? 0 : (c > block_rows ? block_rows : c);
        }
        s_pre[n_blocks] = run;
    }
    __syncthreads();
    const long long total = s_pre[n_blocks], skip = total > capacity ? total - capacity : 0;
    const long long base = (long long)*count;
    const int64_t items = total * (SIT_TR_FLOATS / 4);
    for (int64_t it = (int64_t)blockIdx.x * 256 + threadIdx.x; it < items; it += (int64_t)gridDim.x * 256) {
        const long long g = it / (SIT_TR_FLOATS / 4);
        const int q = (int)(it % (SIT_TR_FLOATS / 4));
        if (g < skip) continue;
        int r = 0;
        while (r + 1 < n_blocks && s_pre[r + 1] <= g) ++r;
        const long long j = g - s_pre[r];
        dst[((base + g) % capacity) * (SIT_TR_FLOATS / 4) + q] =
            blocks[((int64_t)r * (1 + block_rows) + 1 + j) * (SIT_TR_FLOATS / 4) + q];
    }
}
__global__ void replay_append_finish_kernel(unsigned long long* count, const float* blocks, const int n_blocks, const int64_t block_rows) {
    long long run = 0;
    for (int r = 0; r < n_blocks; ++r) {
        long long c = reinterpret_cast<const long long*>(blocks + (int64_t)r * (1 + block_rows) * SIT_TR_FLOATS)[0];
        run += c < 0 ? 0 : (c > block_rows ? block_rows : c);
    }
    *count += (unsigned long long)run;
}

extern "C" {

int sit_replay_pack(const SitRollout* ro, int64_t* seen, int64_t* dropped, float* block, int64_t block_rows, void* stream) {
    if (!ro || !ro->replay || !ro->counters || !seen || !block) return fail(SIT_E_INVALID, "null argument");
    if (block_rows < 1 || ro->capacity < 1) return fail(SIT_E_INVALID, "block_rows and capacity must be >= 1");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int grid = (int)std::min<int64_t>((block_rows * (SIT_TR_FLOATS / 4) + 255) / 256, 148 * 8);
    replay_pack_kernel<<<grid, 256, 0, st>>>(reinterpret_cast<const float4*>(ro->replay), ro->capacity, ro->counters + SIT_RC_CLOSED,
                                             reinterpret_cast<const long long*>(seen), reinterpret_cast<float4*>(block), block_rows);
    replay_pack_finish_kernel<<<1, 1, 0, st>>>(ro->counters + SIT_RC_CLOSED, reinterpret_cast<long long*>(seen),
                                               reinterpret_cast<long long*>(dropped), block, block_rows);
    CUDA_TRY(cudaGetLastError());
    return SIT_OK;
}

int sit_replay_append(float* dst, int64_t dst_capacity, unsigned long long* dst_count, const float* blocks, int32_t n_blocks,
                      int64_t block_rows, void* stream) {
    if (!dst || !dst_count || !blocks) return fail(SIT_E_INVALID, "null argument");
    if (n_blocks < 1 || n_blocks > 64 || block_rows < 1 || dst_capacity < 1) return fail(SIT_E_INVALID, "bad block geometry (1..64 blocks)");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int grid = (int)std::min<int64_t>(((int64_t)n_blocks * block_rows * (SIT_TR_FLOATS / 4) + 255) / 256, 148 * 8);
    replay_append_kernel<<<grid, 256, 0, st>>>(reinterpret_cast<float4*>(dst), dst_capacity, dst_count,
                                               reinterpret_cast<const float4*>(blocks), n_blocks, block_rows);
    replay_append_finish_kernel<<<1, 1, 0, st>>>(dst_count, blocks, n_blocks, block_rows);
    CUDA_TRY(cudaGetLastError());
    return SIT_OK;
}

}  // extern "C"

extern "C" {

int sit_env_step_k(SitHandle h, void* state, int32_t K, const SitEnvIO* io, void* stream) {
    if (!h || !state) return fail(SIT_E_INVALID, "null argument");
    if (K < 1) return fail(SIT_E_INVALID, "K must be >= 1");
    if (h->cfg.kind != SIT_KIND_ENV) return fail(SIT_E_KIND, "sit_env_step_k needs a SIT_KIND_ENV handle");
    CUDA_TRY(cudaSetDevice(h->device));
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    return h->cfg.precision == SIT_F64 ? env_step_t<double>(h, state, K, io, st) : env_step_t<float>(h, state, K, io, st);
}

int sit_env_step_host(SitHandle h, void* state, int32_t K, const void* action_h, const uint8_t* sac_update_h,
                      int32_t init, void* obs_h, void* reward_h, uint8_t* done_h, int32_t* status_h, void* stream) {
    if (!h || !state || !action_h) return fail(SIT_E_INVALID, "null argument");
    if (h->cfg.kind != SIT_KIND_ENV) return fail(SIT_E_KIND, "sit_env_step_host needs a SIT_KIND_ENV handle");
    CUDA_TRY(cudaSetDevice(h->device));
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const size_t B = (size_t)h->n_envs, rs = h->cfg.precision == SIT_F64 ? 8 : 4;
    // device scratch: action [2][B] | obs [10][B] | reward [B] | status [B] i32 | sac [B] u8 | done [B] u8
    const size_t o_act = 0, o_obs = align_up(o_act + 2 * B * rs, 256), o_rew = align_up(o_obs + SIT_N_OBS * B * rs, 256),
                 o_st = align_up(o_rew + B * rs, 256), o_sac = align_up(o_st + B * 4, 256), o_done = align_up(o_sac + B, 256),
                 total = align_up(o_done + B, 256);
    {
        int rc0 = ensure_scratch(h, total);
        if (rc0) return rc0;
    }
    char* d = static_cast<char*>(h->d_scratch);
    if (K < 1) return fail(SIT_E_INVALID, "K must be >= 1");
    SitEnvIO io;
    std::memset(&io, 0, sizeof(io));
    io.action = d + o_act; io.sac_update = sac_update_h ? reinterpret_cast<uint8_t*>(d + o_sac) : nullptr;
    io.act_steps = 1; io.init = init; io.auto_reset = 0; io.rec_steps = 1;
    io.obs = d + o_obs; io.reward = d + o_rew; io.done = reinterpret_cast<uint8_t*>(d + o_done);
    io.status = reinterpret_cast<int32_t*>(d + o_st);
    // envs are independent: large batches travel as chunks through the handle's internal streams, like sit_fleet_step_host
    const int64_t min_chunk = 32768;
    const int n_chunks = (int)std::min<int64_t>(8, (int64_t)B / min_chunk);
    if (n_chunks < 2) {
        CUDA_TRY(cudaMemcpyAsync(d + o_act, action_h, 2 * B * rs, cudaMemcpyHostToDevice, st));
        if (sac_update_h) CUDA_TRY(cudaMemcpyAsync(d + o_sac, sac_update_h, B, cudaMemcpyHostToDevice, st));
        int rc = sit_env_step_k(h, state, K, &io, stream);
        if (rc) return rc;
        if (obs_h) CUDA_TRY(cudaMemcpyAsync(obs_h, d + o_obs, SIT_N_OBS * B * rs, cudaMemcpyDeviceToHost, st));
        if (reward_h) CUDA_TRY(cudaMemcpyAsync(reward_h, d + o_rew, B * rs, cudaMemcpyDeviceToHost, st));
        if (done_h) CUDA_TRY(cudaMemcpyAsync(done_h, d + o_done, B, cudaMemcpyDeviceToHost, st));
        if (status_h) CUDA_TRY(cudaMemcpyAsync(status_h, d + o_st, B * 4, cudaMemcpyDeviceToHost, st));
        CUDA_TRY(cudaStreamSynchronize(st));
        return SIT_OK;
    }
    int rc = ensure_pipe(h);
    if (rc) return rc;
    CUDA_TRY(cudaEventRecord(h->pipe_in, st));
    for (int j = 0; j < SIT_PIPE; ++j) CUDA_TRY(cudaStreamWaitEvent(h->pipe[j], h->pipe_in, 0));
    const int64_t chunk = ((int64_t)B / n_chunks + 255) / 256 * 256;
    int c = 0;
    cudaError_t ce = cudaSuccess;
    auto ok = [&](cudaError_t e) { if (ce == cudaSuccess) ce = e; return e == cudaSuccess; };
    for (int64_t first = 0; first < (int64_t)B && !rc && ce == cudaSuccess; first += chunk, ++c) {
        const int64_t cnt = std::min<int64_t>(chunk, (int64_t)B - first);
        cudaStream_t ps = h->pipe[c % SIT_PIPE];
        if (!ok(cudaMemcpy2DAsync(d + o_act + first * rs, B * rs, static_cast<const char*>(action_h) + first * rs, B * rs,
                                  cnt * rs, 2, cudaMemcpyHostToDevice, ps))) break;
        if (sac_update_h && !ok(cudaMemcpyAsync(d + o_sac + first, sac_update_h + first, cnt, cudaMemcpyHostToDevice, ps))) break;
        rc = h->cfg.precision == SIT_F64 ? env_step_t<double>(h, state, K, &io, ps, first, cnt)
                                         : env_step_t<float>(h, state, K, &io, ps, first, cnt);
        if (rc) break;
        if (obs_h && !ok(cudaMemcpy2DAsync(static_cast<char*>(obs_h) + first * rs, B * rs, d + o_obs + first * rs, B * rs, cnt * rs,
                                           SIT_N_OBS, cudaMemcpyDeviceToHost, ps))) break;
        if (reward_h && !ok(cudaMemcpyAsync(static_cast<char*>(reward_h) + first * rs, d + o_rew + first * rs, cnt * rs,
                                            cudaMemcpyDeviceToHost, ps))) break;
        if (done_h && !ok(cudaMemcpyAsync(done_h + first, d + o_done + first, cnt, cudaMemcpyDeviceToHost, ps))) break;
        if (status_h && !ok(cudaMemcpyAsync(status_h + first, d + o_st + first * 4, cnt * 4, cudaMemcpyDeviceToHost, ps))) break;
    }
    join_pipes(h, st);   // also after an error: the forked work is joined before the call returns
    if (rc) return rc;
    if (ce != cudaSuccess) return fail(SIT_E_CUDA, std::string("sit_env_step_host: ") + cudaGetErrorString(ce));
    CUDA_TRY(cudaStreamSynchronize(st));
    return SIT_OK;
}

}  // extern "C"

// ---- sit_env_step_host_noalloc: the staging buffers of sit_env_step_host in a fixed device array ----------------------------
// No counterpart in the reference; added for callers that may not allocate (CUDA
// graph capture, tight serving loops, harnesses that forbid cudaMalloc).  The array below is part of the module image: every
// GPU that loads the library holds SIT_NOALLOC_SCRATCH_BYTES of it until the process exits.  In HOST code the name
// `g_noalloc_scratch` is only a stand-in for the symbol -- its device address comes from cudaGetSymbolAddress, once per device,
// with that device current (noalloc_scratch_ptr below); kernels are handed that address, never the name.
#define SIT_NOALLOC_SCRATCH_BYTES (2u << 20)
static __device__ __align__(256) unsigned char g_noalloc_scratch[SIT_NOALLOC_SCRATCH_BYTES];
#define SIT_NOALLOC_MAX_DEVICES 64
static void* g_noalloc_ptr[SIT_NOALLOC_MAX_DEVICES];
static std::mutex g_noalloc_mu;

static int noalloc_scratch_ptr(int device, char** out) {   // the caller has made `device` current
    if (device < 0 || device >= SIT_NOALLOC_MAX_DEVICES) return fail(SIT_E_INVALID, "device index beyond the no-alloc table (64)");
    std::lock_guard<std::mutex> lk(g_noalloc_mu);
    if (!g_noalloc_ptr[device]) {
        void* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess || !p) {
            cudaGetLastError();
            return fail(SIT_E_CUDA, std::string("cudaGetSymbolAddress(g_noalloc_scratch): ") + cudaGetErrorString(e));
        }
        g_noalloc_ptr[device] = p;
    }
    *out = static_cast<char*>(g_noalloc_ptr[device]);
    return SIT_OK;
}

extern "C" {

int sit_env_step_host_noalloc(SitHandle h, void* state, int32_t K, const void* action_h, const uint8_t* sac_update_h,
                              int32_t init, void* obs_h, void* reward_h, uint8_t* done_h, int32_t* status_h, void* stream) {
    if (!h || !state || !action_h) return fail(SIT_E_INVALID, "null argument");
    if (h->cfg.kind != SIT_KIND_ENV) return fail(SIT_E_KIND, "sit_env_step_host_noalloc needs a SIT_KIND_ENV handle");
    if (K < 1) return fail(SIT_E_INVALID, "K must be >= 1");
    if (h->n_envs > SIT_NOALLOC_MAX_ENVS)
        return fail(SIT_E_INVALID, "sit_env_step_host_noalloc: n_envs = " + std::to_string((long long)h->n_envs) + " exceeds the fixed scratch (SIT_NOALLOC_MAX_ENVS = " +
                                       std::to_string(SIT_NOALLOC_MAX_ENVS) + "); use sit_env_step_host");
    CUDA_TRY(cudaSetDevice(h->device));
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const size_t B = (size_t)h->n_envs, rs = h->cfg.precision == SIT_F64 ? 8 : 4;
    // same staging layout as sit_env_step_host: action [2][B] | obs [10][B] | reward [B] | status [B] i32 | sac [B] u8 | done [B] u8
    const size_t o_act = 0, o_obs = align_up(o_act + 2 * B * rs, 256), o_rew = align_up(o_obs + SIT_N_OBS * B * rs, 256),
                 o_st = align_up(o_rew + B * rs, 256), o_sac = align_up(o_st + B * 4, 256), o_done = align_up(o_sac + B, 256),
                 total = align_up(o_done + B, 256);
    if (total > SIT_NOALLOC_SCRATCH_BYTES) return fail(SIT_E_INVALID, "sit_env_step_host_noalloc: staging exceeds the fixed scratch");
    char* d = nullptr;
    {
        int rc0 = noalloc_scratch_ptr(h->device, &d);
        if (rc0) return rc0;
    }
    // the array keeps the previous call's contents: outputs start from zero, inputs are overwritten in full below
    CUDA_TRY(cudaMemsetAsync(d + o_obs, 0, total - o_obs, st));
    SitEnvIO io;
    std::memset(&io, 0, sizeof(io));
    io.action = d + o_act; io.sac_update = sac_update_h ? reinterpret_cast<uint8_t*>(d + o_sac) : nullptr;
    io.act_steps = 1; io.init = init; io.auto_reset = 0; io.rec_steps = 1;
    io.obs = d + o_obs; io.reward = d + o_rew; io.done = reinterpret_cast<uint8_t*>(d + o_done);
    io.status = reinterpret_cast<int32_t*>(d + o_st);
    CUDA_TRY(cudaMemcpyAsync(d + o_act, action_h, 2 * B * rs, cudaMemcpyHostToDevice, st));
    if (sac_update_h) CUDA_TRY(cudaMemcpyAsync(d + o_sac, sac_update_h, B, cudaMemcpyHostToDevice, st));
    int rc = h->cfg.precision == SIT_F64 ? env_step_t<double>(h, state, K, &io, st) : env_step_t<float>(h, state, K, &io, st);
    if (rc) return rc;
    if (obs_h) CUDA_TRY(cudaMemcpyAsync(obs_h, d + o_obs, SIT_N_OBS * B * rs, cudaMemcpyDeviceToHost, st));
    if (reward_h) CUDA_TRY(cudaMemcpyAsync(reward_h, d + o_rew, B * rs, cudaMemcpyDeviceToHost, st));
    if (done_h) CUDA_TRY(cudaMemcpyAsync(done_h, d + o_done, B, cudaMemcpyDeviceToHost, st));
    if (status_h) CUDA_TRY(cudaMemcpyAsync(status_h, d + o_st, B * 4, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaStreamSynchronize(st));
    return SIT_OK;
}

int sit_map_query(SitHandle h, int64_t n, const void* north, const void* east, uint8_t* flags, void* dist, void* stream) {
    if (!h || !north || !east) return fail(SIT_E_INVALID, "null argument");
    if (h->cfg.kind != SIT_KIND_ENV) return fail(SIT_E_KIND, "sit_map_query needs a SIT_KIND_ENV handle");
    if (n < 0) return fail(SIT_E_INVALID, "n must be >= 0");
    if (n == 0) return SIT_OK;
    CUDA_TRY(cudaSetDevice(h->device));
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int block = 256, grid = grid_for(n, block);
    if (h->cfg.precision == SIT_F64)
        map_query_kernel<double><<<grid, block, 0, st>>>(*h->kd, n, static_cast<const double*>(north), static_cast<const double*>(east),
                                                         flags, static_cast<double*>(dist));
    else
        map_query_kernel<float><<<grid, block, 0, st>>>(*h->kf, n, static_cast<const float*>(north), static_cast<const float*>(east),
                                                        flags, static_cast<float*>(dist));
    CUDA_TRY(cudaGetLastError());
    return SIT_OK;
}

int sit_circle_query(int32_t precision, int32_t n_circles, const double* circles, int64_t n, const void* north, const void* east,
                     uint8_t* flags, void* dist, void* stream) {
    if (!circles || !north || !east) return fail(SIT_E_INVALID, "null argument");
    if (n_circles < 1 || n_circles > SIT_MAX_CIRCLES) return fail(SIT_E_INVALID, "n_circles must be in [1, 64]");
    if (precision != SIT_F64 && precision != SIT_F32) return fail(SIT_E_INVALID, "bad precision");
    if (n < 0) return fail(SIT_E_INVALID, "n must be >= 0");
    if (n == 0) return SIT_OK;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { cudaGetLastError(); return fail(SIT_E_NODEVICE, "no CUDA device: there is no CPU fallback"); }
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int grid = grid_for(n, 256);
    if (precision == SIT_F64) {
        CircleSet<double> cs; std::memset(&cs, 0, sizeof(cs)); cs.n = n_circles;
        for (int c = 0; c < n_circles; ++c) { cs.cn[c] = circles[3 * c]; cs.ce[c] = circles[3 * c + 1]; cs.cr[c] = circles[3 * c + 2]; }
        circle_query_kernel<double><<<grid, 256, 0, st>>>(cs, n, static_cast<const double*>(north), static_cast<const double*>(east),
                                                          flags, static_cast<double*>(dist));
    } else {
        CircleSet<float> cs; std::memset(&cs, 0, sizeof(cs)); cs.n = n_circles;
        for (int c = 0; c < n_circles; ++c) { cs.cn[c] = (float)circles[3 * c]; cs.ce[c] = (float)circles[3 * c + 1]; cs.cr[c] = (float)circles[3 * c + 2]; }
        circle_query_kernel<float><<<grid, 256, 0, st>>>(cs, n, static_cast<const float*>(north), static_cast<const float*>(east),
                                                         flags, static_cast<float*>(dist));
    }
    CUDA_TRY(cudaGetLastError());
    return SIT_OK;
}

int sit_ship_outline(int32_t precision, int64_t n, const void* north, const void* east, const void* psi, void* x_ned, void* y_ned) {
    if (!north || !east || !psi || !x_ned || !y_ned) return fail(SIT_E_INVALID, "null argument");
    if (precision != SIT_F64 && precision != SIT_F32) return fail(SIT_E_INVALID, "bad precision");
    if (n < 0) return fail(SIT_E_INVALID, "n must be >= 0");
    if (n == 0) return SIT_OK;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { cudaGetLastError(); return fail(SIT_E_NODEVICE, "no CUDA device: there is no CPU fallback"); }
    const int grid = grid_for(n, 256);
    if (precision == SIT_F64)
        ship_outline_kernel<double><<<grid, 256>>>(n, static_cast<const double*>(north), static_cast<const double*>(east),
                                                   static_cast<const double*>(psi), static_cast<double*>(x_ned), static_cast<double*>(y_ned));
    else
        ship_outline_kernel<float><<<grid, 256>>>(n, static_cast<const float*>(north), static_cast<const float*>(east),
                                                  static_cast<const float*>(psi), static_cast<float*>(x_ned), static_cast<float*>(y_ned));
    CUDA_TRY(cudaGetLastError());
    return SIT_OK;
}

int sit_measure_fma_peak(int device, int32_t precision, double* tflops, double* ms_out) {
    if (!tflops) return fail(SIT_E_INVALID, "null argument");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { cudaGetLastError(); return fail(SIT_E_NODEVICE, "no CUDA device"); }
    CUDA_TRY(cudaSetDevice(device));
    int sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
    void* out = nullptr;
    CUDA_TRY(cudaMalloc(&out, 64));
    cudaEvent_t e0, e1;
    CUDA_TRY(cudaEventCreate(&e0)); CUDA_TRY(cudaEventCreate(&e1));
    const int block = 256, grid = sms * 8, iters = precision == SIT_F64 ? 2000 : 4000;
    float best = 1e30f;
    for (int rep = 0; rep < 5; ++rep) {
        cudaEventRecord(e0, 0);
        if (precision == SIT_F64) fma_peak_kernel<double><<<grid, block>>>(static_cast<double*>(out), iters, 1.0000001, 1e-9);
        else if (precision == SIT_PEAK_F32X2) fma_peak_x2_kernel<<<grid, block>>>(static_cast<float*>(out), iters, 1.0000001f, 1e-9f);
        else fma_peak_kernel<float><<<grid, block>>>(static_cast<float*>(out), iters, 1.0000001f, 1e-9f);
        cudaEventRecord(e1, 0);
        CUDA_TRY(cudaEventSynchronize(e1));
        float ms = 0;
        cudaEventElapsedTime(&ms, e0, e1);
        if (rep > 0 && ms < best) best = ms;
    }
    CUDA_TRY(cudaGetLastError());
    const double flops = (precision == SIT_PEAK_F32X2 ? 2.0 : 1.0) * 2.0 * 8 * 16 * (double)iters * block * (double)grid;
    *tflops = flops / (best * 1e-3) / 1e12;
    if (ms_out) *ms_out = best;
    cudaEventDestroy(e0); cudaEventDestroy(e1); cudaFree(out);
    return SIT_OK;
}

}  // extern "C"
#include "sac_learner.cuh"
