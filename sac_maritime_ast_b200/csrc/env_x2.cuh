// Packed fp32 form of the two-ship AST env step (MultiShipRLEnv.step, RLEnv/MSRL_env_ex.py:402-440): ONE THREAD OWNS ONE ENV,
// its two ships ride the two lanes of the Blackwell packed f32x2 instructions (lane 0 = test ship, lane 1 = obstacle ship) through
// ship_x2.cuh::ship_step2, exactly like the fleet kernel packs two independent ships.  Compared with env_step_kernel (one
// lane per ship, a lane pair per env) the FP32 work of the two ship steps is issued once instead of twice, the pair
// exchanges (__shfl_xor of positions, partial rewards, flags) disappear because both ships live in one thread, and rewards /
// terminal logic run once per env.  Everything that DECIDES is unchanged: waypoint switch, arrival, ship-ship collision and
// the travelled distances are evaluated in double on the float64 positions (reconstructed from the segment start and the
// compensated relative pair) with the reference's operation order; the map queries use the same nearest-ring cache and the
// same warp-cooperative rescans, once per ship.
//
// Used for fp32 env handles whose two roles share one parameter set (the reference's default, main_ast.py:207-237) when no
// reward_parts / snapshot output is asked for; every other case runs env_step_kernel.  Same state buffer, same outputs.
//
// MEASURED (B200, 262 144 envs x 64 steps per launch; rollout = 65 536 envs, K = 32), against env_step_kernel:
//   * warp instructions per env-step: 38 % fewer (73 M against 117 M per rollout launch);
//   * first version (two inlined step instances, every rare path inlined per ship; 4 700 SASS instructions): 1.39e10 -> 1.67e10
//     env-steps/s while episodes are young, but 1.27e10 -> 1.08e10 once resets, rescans and hull checks are frequent (the
//     bench's env leg) and 1.06e10 -> 0.86e10 over the 10 000-step horizon: issue slots 33 % busy with 5.3 instruction-fetch
//     stalls per issue (0.34 for the lane-pair kernel) -- half the warps, diverging into a 75 KB instruction footprint;
//   * this version (ONE step instance that also serves init_step through per-thread step counters, the per-ship rare paths
//     as two-trip loops around one copy, segment geometry as a real call; 3 400 instructions): young 1.49e10, bench leg
//     1.20e10 (lane-pair: 1.26e10), rollout 2.44e7 samples/s (lane-pair: 2.45e7) at (128, 2); (128, 3) and (128, 4) spill
//     and are slower in the rollout.
// The warp-cooperative rare paths (one warp rescans / hull-checks for one ship at a time) cost the same per event in both
// kernels, and with 64 ships per warp and half the resident warps there is less left to overlap them with.  The kernel
// is therefore OPT-IN (SIT_ENV_X2=1; the rollout variant additionally needs -DSIT_ROLL_X2 at build time), parity-tested on
// all ten env goldens (tests/test_gpu_env.py) -- a measured experiment, not the default path.
#pragma once

namespace sit {

// element l of a register pair (a run-time l compiles to selects: no local-memory array)
template <typename T>
__device__ __forceinline__ T& pick(T& a, T& b, const int l) { return l == 0 ? a : b; }

// MULTI: per-ship machinery modes (run-time mode switching); otherwise the single mode's constants are uniform operands.
#ifndef SIT_ENVX2_MINB
#define SIT_ENVX2_MINB 2   /* 243 registers, no spills; (128, 3) = 168 and (128, 4) = 128 registers spill 200 .. 570 B */
#endif
template <bool ROLL, bool MULTI>
__global__ void __launch_bounds__(128, SIT_ENVX2_MINB)
env_step_x2_kernel(const __grid_constant__ EnvKArgs<float> A, const __grid_constant__ x2::ShipPX P, const ShipPtrs<float> g,
                   const EnvPtrs<float> ge, const int K, const EnvIO<float> io, const RollArgs ro) {
    using namespace x2;
    typedef float R;
    const unsigned FULL = 0xffffffffu;
    __shared__ EdgeRec<R> s_edges[SIT_MAX_EDGES];
    stage_edges(A.M, s_edges);
    const int n_edges = A.M.n_edges;
    const int64_t e_raw = (int64_t)blockIdx.x * blockDim.x + threadIdx.x + ro.env_first;   // env sub-range of this launch
    const int64_t S = g.n_ships, B = ge.n_envs;
    const bool active = e_raw < ro.env_first + ro.env_count;
    const int64_t env = active ? e_raw : 0;        // idle lanes of the last warp shadow env 0 and never store
    const int64_t i0 = 2 * env;                    // test ship slot; the obstacle ship sits at i0 + 1
    const EnvP<R>& EP = A.EP;
    const int wl = threadIdx.x & 31;            // lane of the warp (`lane(v, l)` below selects a ship of the packed pair)
    auto idx = [&](int l) -> int64_t { return 2 * (active ? e_raw : 0) + l; };
    auto ld2 = [&](const float* p) -> V { return *reinterpret_cast<const float2*>(p + i0); };
    auto st2 = [&](float* p, const V v) { *reinterpret_cast<float2*>(p + i0) = v; };

    // ---- state: both ships of the env ------------------------------------------------------------------------------
    Ship2 s;
    s.psi = ld2(g.psi); s.u = ld2(g.u); s.v = ld2(g.v); s.r = ld2(g.r); s.om = ld2(g.omega);
    s.ei_s = ld2(g.ei_speed); s.ei_w = ld2(g.ei_shaft); s.ei_h = ld2(g.ei_heading); s.e_prev = ld2(g.e_prev);
    s.e_ct = ld2(g.e_ct); s.e_int = ld2(g.e_ct_int);
    s.f_me = ld2(g.fuel_me); s.f_el = ld2(g.fuel_el);
    V f_off = sub(ld2(g.fuel), add(s.f_me, s.f_el));    // fuel_cons beside its two parts (ship_engine.py:288-292)
    s.u_sp = ld2(g.u_sp); s.vcn = ld2(g.vc_n); s.vce = ld2(g.vc_e);
    auto set_wind2 = [&](const V w_speed, const V w_dir) {
        float sb, cb;
        sincos_(w_dir.x, sb, cb); s.wcb.x = w_speed.x * cb; s.wsb.x = w_speed.x * sb;
        sincos_(w_dir.y, sb, cb); s.wcb.y = w_speed.y * cb; s.wsb.y = w_speed.y * sb;
    };
    set_wind2(ld2(g.wind_speed), ld2(g.wind_dir));
    // per ship (0 = test, 1 = obstacle), as register pairs: step counter, segment index the float64 base belongs to, segment
    // start and absolute position (float64), route end point (constant: insertions go before it, controllers.py:298-303)
    int32_t steps0, steps1, kk0, kk1;
    double pN0, pN1, pE0, pE1, aN0, aN1, aE0, aE1;
    float endN0, endN1, endE0, endE1;
#pragma unroll 1
    for (int l = 0; l < 2; ++l) {
        const int64_t i = i0 + l;
        const int k = g.k[i], nw = g.n_wp[i];
        pick(s.k[0], s.k[1], l) = k; pick(s.n_wp[0], s.n_wp[1], l) = nw; pick(steps0, steps1, l) = g.steps[i];
        pick(kk0, kk1, l) = k;
        const double pN = g.wp_n[(int64_t)(k - 1) * S + i], pE = g.wp_e[(int64_t)(k - 1) * S + i];
        const double cN = g.wp_n[(int64_t)k * S + i], cE = g.wp_e[(int64_t)k * S + i];
        const double N = g.north[i], E = g.east[i];
        // the relative pair written by the previous launch is kept while it still describes the float64 position
        float x = g.rel_n[i], cx = g.rel_cn[i], y = g.rel_e[i], cy = g.rel_ce[i];
        if (!(abs_pos(pN, x, cx) == N)) split_rel(N - pN, x, cx);
        if (!(abs_pos(pE, y, cy) == E)) split_rel(E - pE, y, cy);
        lane(s.X, l) = x; lane(s.CX, l) = cx; lane(s.Y, l) = y; lane(s.CY, l) = cy;
        pick(pN0, pN1, l) = pN; pick(pE0, pE1, l) = pE; pick(aN0, aN1, l) = N; pick(aE0, aE1, l) = E;
        set_segment2<true>(s, P, pN, pE, cN, cE, l);
        pick(endN0, endN1, l) = g.wp_n[(int64_t)(nw - 1) * S + i]; pick(endE0, endE1, l) = g.wp_e[(int64_t)(nw - 1) * S + i];
    }
    ModeV MV;
    auto load_modes = [&]() {
        if (MULTI) {
            const ModeX &a = P.mode[g.mode[i0]], &b = P.mode[g.mode[i0 + 1]];
#define MV_(f) MV.f = make_float2(a.f, b.f)
            MV_(avail_total); MV_(avail_me); MV_(avail_el); MV_(cap_tq_me); MV_(cap_tq_el); MV_(add_me); MV_(cap_me);
            MV_(a_el); MV_(c_el); MV_(inv_me_cap); MV_(inv_el_cap); MV_(me_cap_kw);
            MV_(fme_a); MV_(fme_b); MV_(fel_a); MV_(fel_b);
#undef MV_
        }
    };
    load_modes();

    int32_t flags = ge.flags[env];
    bool stop_t = (flags & F_STOP_TEST) != 0, stop_o = (flags & F_STOP_OBS) != 0;
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
    unsigned my_cnt = 0, n_live = 0, n_eps = 0, n_closed = 0;
    MapCache mc0, mc1;
    mc0.idx = (uint32_t)g.mc_idx[i0]; mc0.p0n = g.mc_n[i0]; mc0.p0e = g.mc_e[i0]; mc0.b2 = g.mc_b2[i0];
    mc1.idx = (uint32_t)g.mc_idx[i0 + 1]; mc1.p0n = g.mc_n[i0 + 1]; mc1.p0e = g.mc_e[i0 + 1]; mc1.b2 = g.mc_b2[i0 + 1];
    const R rpm_per_rads = A.P[0].rpm_per_rads;

    // absolute float64 positions from the (possibly new) segment starts
    auto refresh_abs = [&]() {
        if (s.k[0] != kk0 || s.k[1] != kk1) {     // a ship switched segments inside the step: new float64 base
#pragma unroll 1
            for (int l = 0; l < 2; ++l) {
                const int k = pick(s.k[0], s.k[1], l);
                pick(kk0, kk1, l) = k;
                pick(pN0, pN1, l) = g.wp_n[(int64_t)(k - 1) * S + idx(l)]; pick(pE0, pE1, l) = g.wp_e[(int64_t)(k - 1) * S + idx(l)];
            }
        }
        aN0 = abs_pos(pN0, s.X.x, s.CX.x); aE0 = abs_pos(pE0, s.Y.x, s.CY.x);
        aN1 = abs_pos(pN1, s.X.y, s.CX.y); aE1 = abs_pos(pE1, s.Y.y, s.CY.y);
    };

    // One trip of the loop = ONE packed ship step.  A thread whose episode ended reloads the initial state and spends its next
    // trip on init_step (MSRL_env_ex.py:188-215: no avoidance, nothing stored, no rewards) WITHOUT advancing its step counter
    // t, so the kernel holds a single instance of the step code; the other threads of the warp take an ordinary step in that
    // trip.  t is therefore per thread, and the loop runs until every thread of the warp has done its K steps.
    int t = 0;
    bool init_mode = false;
    while (true) {
        const bool todo = t < K || init_mode, was_init = init_mode;
        if (!__any_sync(FULL, todo)) break;
        // ---- inputs: converted_action and SAC_update (MSRL_env_ex.py:402-405) -----------------------------------------
        bool su = false, live = todo;
        const bool step_mode = live && !init_mode;
        if (ROLL) {
            if (step_mode && pending) {
                if (have_act) { su = true; pending = false; have_act = false; tr_open = true; }   // action from sit_rollout_policy
                else live = false;
            }
            if (!__any_sync(FULL, live && active)) break;
        } else if (step_mode && io.action) {
            const int64_t ta = io.act_steps > 1 ? t : 0;
            const R a_n = io.action[(ta * 2 + 0) * B + env], a_e = io.action[(ta * 2 + 1) * B + env];
            if (!iw_valid || a_n != act_n || a_e != act_e) {
                // is_route_outside_horizon :517-542 or is_route_inside_obstacles :544-552 of the action, cached until it changes
                act_n = a_n; act_e = a_e;
                iw_bad = (a_n < EP.min_n || a_n > EP.max_n || a_e < EP.min_e || a_e > EP.max_e) ||
                         map_contains(s_edges, n_edges, a_n, a_e);
                iw_valid = true;
            }
            if (io.sac_update && (io.act_steps > 1 || t == 0)) su = io.sac_update[ta * B + env] != 0;
        }
        int status = 0;
        bool done = false;
        StepOut2 so;
        so.om_pre = bc(0.0f); so.load_me = bc(0.0f);
        double ship_d2 = 0.0;
        const bool stepping = live && !init_mode;   // an ordinary env step in this trip (false: idle, pending, or init_step)
        if (live) {   // ---- phase A: the two ship steps ------------------------------------------------------------------
            if (stepping) n_live += 1;
            const bool init = (ROLL || io.init < 0) ? fresh : (t == 0 && io.init != 0);
            const bool frozen = stepping && stop_o;                      // obs_step :285-310: a stopped obstacle ship is not advanced
            const double pre_n = aN1, pre_e = aE1;
            if (stepping && !frozen && su) {
                // obs_step :324-338 -> update_route controllers.py:298-303: list.insert(-1, wp)
                const int64_t i = i0 + 1;
                const int nw = s.n_wp[1];
                if (nw < g.max_wp) {
                    if (active) {
                        g.wp_n[(int64_t)nw * S + i] = endN1; g.wp_e[(int64_t)nw * S + i] = endE1;
                        g.wp_n[(int64_t)(nw - 1) * S + i] = act_n; g.wp_e[(int64_t)(nw - 1) * S + i] = act_e;
                    }
                    s.n_wp[1] = nw + 1;
                    if (s.k[1] == nw - 1) set_segment2<true>(s, P, pN1, pE1, (double)act_n, (double)act_e, 1);   // the segment in use ended at the old last point
                }
                samp_dist = 0.0;
            }
            // the frozen obstacle ship keeps its lane: save, step both lanes, restore
            float kp[24];
            int kpk = 0, kpn = 0;
            if (frozen) {
#define KP_(j, f) kp[j] = s.f.y
                KP_(0, X); KP_(1, Y); KP_(2, CX); KP_(3, CY); KP_(4, SX); KP_(5, SY); KP_(6, thr2); KP_(7, psi); KP_(8, u); KP_(9, v);
                KP_(10, r); KP_(11, om); KP_(12, ei_s); KP_(13, ei_w); KP_(14, ei_h); KP_(15, e_prev); KP_(16, e_ct); KP_(17, e_int);
                KP_(18, f_me); KP_(19, f_el); KP_(20, alpha); KP_(21, sa); KP_(22, ca);
#undef KP_
                kpk = s.k[1]; kpn = s.n_wp[1];
            }
            // test_step :217-283 (avoidance branch always taken) | obs_step :346-373 | init_step :188-215 (init_mode)
            if (MULTI) ship_step2<AVOID_LANE0_RT, true, true>(P, MV, s, g, idx, so, stepping, stepping);
            else ship_step2<AVOID_LANE0_RT, true, true>(P, P.mode[0], s, g, idx, so, stepping, stepping);
            if (stepping) steps0 += 1;
            if (frozen) {
#define KP_(j, f) s.f.y = kp[j]
                KP_(0, X); KP_(1, Y); KP_(2, CX); KP_(3, CY); KP_(4, SX); KP_(5, SY); KP_(6, thr2); KP_(7, psi); KP_(8, u); KP_(9, v);
                KP_(10, r); KP_(11, om); KP_(12, ei_s); KP_(13, ei_w); KP_(14, ei_h); KP_(15, e_prev); KP_(16, e_ct); KP_(17, e_int);
                KP_(18, f_me); KP_(19, f_el); KP_(20, alpha); KP_(21, sa); KP_(22, ca);
#undef KP_
                s.k[1] = kpk; s.n_wp[1] = kpn;
                steps1 += 2;                                             // frozen ship: next_time twice, :294,307
            } else if (stepping) {
                steps1 += 1;
            }
            refresh_abs();
            if (frozen) {
                if (nstored >= 1) nstored = 2;                           // store_last_simulation_data repeats the last row :292
            } else if (stepping) {
                if (!init && nstored >= 1) {                             // :391-395, last two *stored* (pre-step) positions
                    const double dn = pre_n - prev_n, de = pre_e - prev_e;
                    const double d = sqrt(__dadd_rn(__dmul_rn(dn, dn), __dmul_rn(de, de)));
                    eps_dist += d; samp_dist += d;
                }
                prev_n = pre_n; prev_e = pre_e;
                nstored = nstored < 2 ? nstored + 1 : 2;
            }
            const double ddn = aN0 - aN1, dde = aE0 - aE1;
            ship_d2 = __dadd_rn(__dmul_rn(ddn, ddn), __dmul_rn(dde, dde));
        }

        // ---- phase B (whole warp): distance to the nearest ring + hull containment, nearest-ring cache, once per ship, as a
        //      two-trip loop around ONE copy of the cooperative scans (instruction footprint) --------------------------------
        const R pn0 = (R)aN0, pe0 = (R)aE0, pn1 = (R)aN1, pe1 = (R)aE1;
        R d2map0 = (R)3.0e38, d2map1 = (R)3.0e38;
        bool inside0 = false, inside1 = false;
        if (n_edges > 0) {
#pragma unroll 1
            for (int l = 0; l < 2; ++l) {
                const R pnl = l == 0 ? pn0 : pn1, pel = l == 0 ? pe0 : pe1;
                MapCache m = l == 0 ? mc0 : mc1;
                const float ddn_ = pnl - m.p0n, dde_ = pel - m.p0e;
                unsigned nb = __ballot_sync(FULL, stepping && !(ddn_ * ddn_ + dde_ * dde_ < fabsf(m.b2)));
                while (nb) {   // budget spent: the whole warp rescans that lane's position
                    const int L = __ffs(nb) - 1;
                    nb &= nb - 1;
                    const MapCache c = coop_scan(s_edges, n_edges, __shfl_sync(FULL, pnl, L), __shfl_sync(FULL, pel, L), wl);
                    if (wl == L) m = c;
                }
                const R d2 = cached_dist2(s_edges, m.idx, pnl, pel);
                bool ins = (__float_as_uint(m.b2) >> 31) != 0u;   // far from every ring all four corners share the centre's side
                unsigned hb = __ballot_sync(FULL, stepping && d2 <= EP.near2);
                while (hb) {   // is_pos_inside_obstacles, MSRL_env_ex.py:490-515: the four corners of the +-L/2 square
                    const int L = __ffs(hb) - 1;
                    hb &= hb - 1;
                    const bool in = coop_hull_inside(s_edges, n_edges, __shfl_sync(FULL, pnl, L), __shfl_sync(FULL, pel, L),
                                                     l == 0 ? EP.half_len[0] : EP.half_len[1], wl);
                    if (wl == L) ins = in;
                }
                if (l == 0) { mc0 = m; d2map0 = d2; inside0 = ins; } else { mc1 = m; d2map1 = d2; inside1 = ins; }
            }
        }

        if (live && init_mode) {
            init_mode = false;                       // init_step done: the new episode starts with the next trip
        } else if (stepping) {   // ---- phase C: rewards, terminal checks, outputs, episode restart --------------------------------
            const R rpm_t = so.om_pre.x * rpm_per_rads;                  // stored (pre-step) observables of the test ship
            const R power_me_kw = so.load_me.x * 0.001f;
            const R avail_me_kw = MULTI ? MV.me_cap_kw.x : P.mode[0].me_cap_kw;
            const R ect_t = fabsf(s.e_ct.x), ect_o = fabsf(s.e_ct.y);
            const R dmap_t = sqrt_(d2map0), dmap_o = sqrt_(d2map1);
            bool dn_t = false, dn_o = false;
            int st = 0;
            // ---- test ship: test_ship_non_terminal_state_reward :628-664, test_ship_terminal_state_reward :734-809 --------
            R part_t;
            {
                const bool outside = pn0 < EP.lo_n[0] || pn0 > EP.hi_n[0] || pe0 < EP.lo_e[0] || pe0 > EP.hi_e[0];
                const double dnE = aN0 - (double)endN0, deE = aE0 - (double)endE0;
                const bool arrived = __dadd_rn(__dmul_rn(dnE, dnE), __dmul_rn(deE, deE)) <= EP.arrive2;
                const R ntt = ect_t * EP.inv_e_tol + ((R)1 - dmap_t * EP.inv_max_north) * (R)0.01;
                R rt = (R)0;
                if (arrived) { stop_t = true; dn_t = true; st |= SIT_ST_TEST_ENDPOINT; }
                if (outside) { stop_t = true; dn_t = true; st |= SIT_ST_TEST_HORIZON; }
                if (inside0) { if (!stop_t) rt += (R)1000; stop_t = true; dn_t = true; st |= SIT_ST_TEST_TERRAIN; }
                if (fabsf(rpm_t) > EP.rpm_max) { if (!stop_t) rt += (R)1000; stop_t = true; dn_t = true; st |= SIT_ST_TEST_MECH; }
                if (ect_t > EP.e_tol) { if (!stop_t) rt += (R)1000; stop_t = true; dn_t = true; st |= SIT_ST_TEST_NAV; }
                if (power_me_kw > avail_me_kw) { if (!stop_t) rt += (R)1000; stop_t = true; dn_t = true; st |= SIT_ST_TEST_BLACKOUT; }
                part_t = ntt + rt;
            }
            // ---- obstacle ship: obs_ship_non_terminal_state_reward :666-710 (stop flag from before this step's checks),
            //      obs_ship_terminal_state_reward :811-881, shared_non_terminal_state_reward :712-731 (flag after them) ------
            R nto = (R)0, ro_ = (R)0, snt = (R)0;
            {
                const bool outside = pn1 < EP.lo_n[1] || pn1 > EP.hi_n[1] || pe1 < EP.lo_e[1] || pe1 > EP.hi_e[1];
                const double dnE = aN1 - (double)endN1, deE = aE1 - (double)endE1;
                const bool arrived = __dadd_rn(__dmul_rn(dnE, dnE), __dmul_rn(deE, deE)) <= EP.arrive2;
                if (!stop_o) nto = (R)0.1 + (-(ect_o * EP.inv_e_tol) * (R)0.01) + (-((R)1 - dmap_o * EP.inv_max_north) * (R)0.01);
                if (arrived) { stop_o = true; st |= SIT_ST_OBS_ENDPOINT; }                       // no done
                if (outside) { stop_o = true; dn_o = true; st |= SIT_ST_OBS_HORIZON; }
                if (inside1) { if (!stop_o) ro_ -= (R)1000; dn_o = true; st |= SIT_ST_OBS_TERRAIN; }   // no stop flag
                if (iw_bad) { if (!stop_o) ro_ -= (R)1000; stop_o = true; dn_o = true; st |= SIT_ST_OBS_IW; }
                if (ect_o > EP.e_tol || samp_dist > (double)ab_seg * EP.theta) {
                    if (!stop_o) ro_ -= (R)1000; stop_o = true; dn_o = true; st |= SIT_ST_OBS_NAV;
                }
                if (!stop_o) snt = ((R)1 - (R)sqrt(ship_d2) * EP.inv_max_north) * (R)0.001;
            }
            // shared_terminal_state_reward :883-904, is_ship_collision :584-603
            const bool collide = ship_d2 < EP.collide2;
            if (collide) { stop_t = true; stop_o = true; dn_t = true; st |= SIT_ST_COLLISION; }
            done = dn_t || dn_o;
            status = st;
            const R reward = ((((part_t + nto) + ro_) + snt)) + (collide ? (R)2000 : (R)0);   // :970-972
            rsum += reward; ep_return += reward; ep_steps += 1;
            any_done |= done;
            fresh = false;

            // ---- outputs: coalesced SoA rows [step][channel][env] -------------------------------------------------------
            if (active && (io.rec_steps > 1 || t == K - 1)) {
                const int64_t rec = io.rec_steps > 1 ? t : 0;
                if (io.obs) {   // :424-435: post-step pose, pre-step stored rpm / e_ct / power
                    R* ob = io.obs + rec * SIT_N_OBS * B + env;
                    ob[0] = pn0; ob[B] = pe0; ob[2 * B] = s.psi.x; ob[3 * B] = rpm_t; ob[4 * B] = ect_t; ob[5 * B] = power_me_kw;
                    ob[6 * B] = pn1; ob[7 * B] = pe1; ob[8 * B] = s.psi.y; ob[9 * B] = ect_o;
                }
                if (io.reward) io.reward[rec * B + env] = reward;
                if (io.done) io.done[rec * B + env] = done ? 1 : 0;
                if (io.status) io.status[rec * B + env] = status;
            }
            bool episode_end = done && io.auto_reset != 0;
            if (ROLL) {
                // main_ast.py:387: mask = 1 if episode_steps == num_steps_episode else float(not done); the reference's
                // episode_steps counts init_step as 1, so it equals ep_steps + 1 here
                const bool timeup = ro.max_episode_steps > 0 && ep_steps + 1 == ro.max_episode_steps;
                if (tr_open) {   // close the replay row opened by sit_rollout_policy: (reward, next_state, mask), :395-396
                    if (active) {
                        float* row = ro.replay + (int64_t)ge.tr_slot[env] * SIT_TR_FLOATS;
                        row[12] = pn0; row[13] = pe0; row[14] = s.psi.x; row[15] = rpm_t; row[16] = ect_t; row[17] = power_me_kw;
                        row[18] = pn1; row[19] = pe1; row[20] = s.psi.y; row[21] = ect_o;
                        row[11] = reward; row[22] = timeup ? 1.0f : (done ? 0.0f : 1.0f);
                        n_closed += 1;
                    }
                    tr_open = false;
                }
                episode_end = done || (timeup && ro.truncate);
            }
            // ---- in-kernel reset() right after a terminal step; init_step() is this thread's next trip ----------------------
            if (episode_end) {
                if (ROLL && active) n_eps += 1;
                const bool hard = io.auto_reset == 2;   // MSRL_env_ex.py:145-186 keeps omega, integrators, fuel unless hard
                V w_speed, w_dir;
#pragma unroll 1
                for (int l = 0; l < 2; ++l) {
                    const int64_t i = i0 + l;
                    const double N = g.init_north[i], E = g.init_east[i];
                    lane(s.psi, l) = g.init_psi[i]; lane(s.u, l) = g.init_u[i]; lane(s.v, l) = g.init_v[i]; lane(s.r, l) = g.init_r[i];
                    const R vcn = g.init_vc_n[i], vce = g.init_vc_e[i], ws = g.init_wind_speed[i], wd = g.init_wind_dir[i], usp = g.init_u_sp[i];
                    lane(s.vcn, l) = vcn; lane(s.vce, l) = vce; lane(w_speed, l) = ws; lane(w_dir, l) = wd; lane(s.u_sp, l) = usp;
                    if (active) { g.vc_n[i] = vcn; g.vc_e[i] = vce; g.wind_speed[i] = ws; g.wind_dir[i] = wd; g.u_sp[i] = usp; }
                    pick(steps0, steps1, l) = 0; pick(s.k[0], s.k[1], l) = 1; pick(kk0, kk1, l) = 1;
                    lane(s.e_ct, l) = 0.0f; lane(s.e_int, l) = 0.0f;
                    const int nw = g.n_wp0[i];
                    pick(s.n_wp[0], s.n_wp[1], l) = nw;
                    if (active)
                        for (int j = 0; j < g.max_wp; ++j) {
                            g.wp_n[(int64_t)j * S + i] = g.wp0_n[(int64_t)j * S + i]; g.wp_e[(int64_t)j * S + i] = g.wp0_e[(int64_t)j * S + i];
                        }
                    const double pN = g.wp0_n[i], pE = g.wp0_e[i], cN = g.wp0_n[S + i], cE = g.wp0_e[S + i];
                    pick(pN0, pN1, l) = pN; pick(pE0, pE1, l) = pE; pick(aN0, aN1, l) = N; pick(aE0, aE1, l) = E;
                    split_rel(N - pN, lane(s.X, l), lane(s.CX, l));
                    split_rel(E - pE, lane(s.Y, l), lane(s.CY, l));
                    set_segment2<true>(s, P, pN, pE, cN, cE, l);
                    pick(endN0, endN1, l) = g.wp0_n[(int64_t)(nw - 1) * S + i]; pick(endE0, endE1, l) = g.wp0_e[(int64_t)(nw - 1) * S + i];
                    if (hard) {   // what the reference's reset() leaves alone (quirk Q3): a fresh construction
                        lane(s.om, l) = g.init_omega[i];
                        lane(s.ei_s, l) = 0.0f; lane(s.ei_w, l) = g.init_ei_shaft[i]; lane(s.ei_h, l) = 0.0f; lane(s.e_prev, l) = 0.0f;
                        lane(s.f_me, l) = 0.0f; lane(s.f_el, l) = 0.0f; lane(f_off, l) = 0.0f;
                    }
                }
                set_wind2(w_speed, w_dir);
                stop_t = false; stop_o = false; nstored = 0; fresh = true;
                eps_dist = 0.0; samp_dist = 0.0; prev_n = 0.0; prev_e = 0.0;
                ep_return = (R)0; ep_steps = 0;
                init_mode = true;                    // init_step :188-215 = the next trip of this thread
                if (ROLL) {
                    // the new episode's first decision acts on reset()'s float32 initial state (MSRL_env_ex.py:86-90,186)
                    pending = true; n_dec = 0;
                    if (active) {
                        R* co = ge.cur_obs + env;
                        co[0] = (R)g.init_north[i0]; co[B] = (R)g.init_east[i0]; co[2 * B] = g.init_psi[i0];
                        co[3 * B] = (R)0; co[4 * B] = (R)0; co[5 * B] = (R)0;
                        co[6 * B] = (R)g.init_north[i0 + 1]; co[7 * B] = (R)g.init_east[i0 + 1]; co[8 * B] = g.init_psi[i0 + 1];
                        co[9 * B] = (R)0;
                    }
                }
            } else if (ROLL) {
                // decision event: the obstacle ship is inside the radius of acceptance of its current intermediate
                // waypoint (controllers.py:317-325; the same squared test as the waypoint switch, LOS_guidance.py:96)
                bool dec = false;
                if (!stop_o && n_dec < ro.sampling_frequency && s.k[1] < s.n_wp[1] - 1) {
                    const float dnf = s.SX.y - s.X.y, def_ = s.SY.y - s.Y.y;
                    if (dnf * dnf + def_ * def_ <= P.ra2_band) {          // float pre-test, exact test in double
                        const int64_t i = i0 + 1;
                        const double dn = (double)g.wp_n[(int64_t)s.k[1] * S + i] - aN1, de = (double)g.wp_e[(int64_t)s.k[1] * S + i] - aE1;
                        dec = __dadd_rn(__dmul_rn(dn, dn), __dmul_rn(de, de)) <= P.ra2;
                    }
                }
                if (dec) {
                    pending = true;
                    if (active) {
                        R* co = ge.cur_obs + env;
                        co[0] = pn0; co[B] = pe0; co[2 * B] = s.psi.x; co[3 * B] = rpm_t; co[4 * B] = ect_t; co[5 * B] = power_me_kw;
                        co[6 * B] = pn1; co[7 * B] = pe1; co[8 * B] = s.psi.y; co[9 * B] = ect_o;
                    }
                }
            }
        }   // stepping
        // ---- warp-level reduction of termination / collision flags ---------------------------------------------------------
        if (io.stats) {
            const bool cnt_lane = active && stepping;
            if (__any_sync(FULL, cnt_lane && status != 0)) {
#pragma unroll
                for (int b = 0; b < 12; ++b) {
                    const unsigned bal = __ballot_sync(FULL, cnt_lane && ((status >> b) & 1));
                    if (wl == b) my_cnt += __popc(bal);
                }
                const unsigned bal = __ballot_sync(FULL, cnt_lane && done);
                if (wl == 12) my_cnt += __popc(bal);
            }
        }
        if (todo && !was_init) t += 1;   // a step slot is used whether the env stepped or waited for an action; init_step is free
    }

    // ---- write back ------------------------------------------------------------------------------------------------------
    if (active) {
        st2(g.psi, s.psi); st2(g.u, s.u); st2(g.v, s.v); st2(g.r, s.r); st2(g.omega, s.om);
        st2(g.ei_speed, s.ei_s); st2(g.ei_shaft, s.ei_w); st2(g.ei_heading, s.ei_h); st2(g.e_prev, s.e_prev);
        st2(g.e_ct, vabs(s.e_ct)); st2(g.e_ct_int, s.e_int);
        st2(g.fuel_me, s.f_me); st2(g.fuel_el, s.f_el); st2(g.fuel, add(add(s.f_me, s.f_el), f_off));
        st2(g.rel_n, s.X); st2(g.rel_cn, s.CX); st2(g.rel_e, s.Y); st2(g.rel_ce, s.CY);
        g.north[i0] = aN0; g.east[i0] = aE0; g.north[i0 + 1] = aN1; g.east[i0 + 1] = aE1;
        g.k[i0] = s.k[0]; g.n_wp[i0] = s.n_wp[0]; g.steps[i0] = steps0;
        g.k[i0 + 1] = s.k[1]; g.n_wp[i0 + 1] = s.n_wp[1]; g.steps[i0 + 1] = steps1;
        g.mc_idx[i0] = (int32_t)mc0.idx; g.mc_n[i0] = mc0.p0n; g.mc_e[i0] = mc0.p0e; g.mc_b2[i0] = mc0.b2;
        g.mc_idx[i0 + 1] = (int32_t)mc1.idx; g.mc_n[i0 + 1] = mc1.p0n; g.mc_e[i0 + 1] = mc1.p0e; g.mc_b2[i0 + 1] = mc1.b2;
        const int32_t f = (stop_t ? F_STOP_TEST : 0) | (stop_o ? F_STOP_OBS : 0) | (nstored << F_NSTORED_SHIFT) |
                          (iw_bad ? F_IW_BAD : 0) | (iw_valid ? F_IW_VALID : 0) | (fresh ? F_FRESH : 0) |
                          (pending ? F_PENDING : 0) | (have_act ? F_HAVE_ACT : 0);
        ge.flags[env] = f;
        ge.prev_n[env] = prev_n; ge.prev_e[env] = prev_e;
        ge.eps_dist[env] = eps_dist; ge.samp_dist[env] = samp_dist;
        ge.act_n[env] = act_n; ge.act_e[env] = act_e;
        ge.ep_return[env] = ep_return; ge.ep_steps[env] = ep_steps;
        if (ROLL) ge.n_dec[env] = n_dec;
        if (io.reward_sum) io.reward_sum[env] = rsum;
        if (io.done_any) io.done_any[env] = any_done ? 1 : 0;
    }
    auto warp_sum = [&](unsigned v) {
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(FULL, v, d);
        return v;
    };
    if (io.stats) {
        if (wl <= 12 && my_cnt) atomicAdd(&io.stats[wl], (unsigned long long)my_cnt);
        const unsigned steps_w = warp_sum(active ? n_live : 0u);       // env-steps executed by this warp
        if (wl == 13 && steps_w) atomicAdd(&io.stats[13], (unsigned long long)steps_w);
    }
    if (ROLL) {
        const unsigned eps = warp_sum(n_eps), cl = warp_sum(n_closed);
        if (wl == 14 && eps) atomicAdd(&ro.counters[SIT_RC_EPISODES], (unsigned long long)eps);
        if (wl == 15 && cl) atomicAdd(&ro.counters[SIT_RC_CLOSED], (unsigned long long)cl);
    }
}

}  // namespace sit
