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
// Measured (B200): 1.39e10 -> 1.67e10 env-steps/s on the bench's env leg (38 % fewer warp instructions per env-step).  The
// ROLLOUT variant (ROLL = true) is written and parity-green but NOT used by default: at 65 536 envs the kernel is one to two
// waves of 8 .. 16 warps per SM, its issue rate drops to 29 % (instruction-fetch stalls 1.8 per issue against 0.46 for the
// lane-pair kernel, which keeps twice the warps in flight) and the rollout runs 6 % slower than with env_step_kernel
// (2.32e7 against 2.46e7 samples/s); compile with -DSIT_ROLL_X2 to try it.
// Reference citations are relative to /root/reference/.
#pragma once

namespace sit {

// MULTI: per-ship machinery modes (run-time mode switching); otherwise the single mode's constants are uniform operands.
// 128 registers / 4 CTAs per SM (16 warps): the step loop spills ~300 B but the doubled occupancy wins (measured on B200,
// 262 144 envs x 64 steps: 1.60e10 env-steps/s at (128, 2) = 211 registers, 1.58e10 at (128, 3), 1.67e10 at (128, 4);
// 64-thread CTAs are 1 % slower at every setting)
#ifndef SIT_ENVX2_MINB
#define SIT_ENVX2_MINB 4
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
    int32_t steps[2], kk[2];
    double pN[2], pE[2], aN[2], aE[2];                 // segment start and absolute position of each ship (float64)
    float endN[2], endE[2];                            // route end point (constant: insertions go before it, controllers.py:298-303)
#pragma unroll
    for (int l = 0; l < 2; ++l) {
        const int64_t i = i0 + l;
        s.k[l] = g.k[i]; s.n_wp[l] = g.n_wp[i]; steps[l] = g.steps[i];
        const int k = s.k[l];
        kk[l] = k;
        pN[l] = g.wp_n[(int64_t)(k - 1) * S + i]; pE[l] = g.wp_e[(int64_t)(k - 1) * S + i];
        const double cN = g.wp_n[(int64_t)k * S + i], cE = g.wp_e[(int64_t)k * S + i];
        aN[l] = g.north[i]; aE[l] = g.east[i];
        // the relative pair written by the previous launch is kept while it still describes the float64 position
        float x = g.rel_n[i], cx = g.rel_cn[i], y = g.rel_e[i], cy = g.rel_ce[i];
        if (!(abs_pos(pN[l], x, cx) == aN[l])) split_rel(aN[l] - pN[l], x, cx);
        if (!(abs_pos(pE[l], y, cy) == aE[l])) split_rel(aE[l] - pE[l], y, cy);
        lane(s.X, l) = x; lane(s.CX, l) = cx; lane(s.Y, l) = y; lane(s.CY, l) = cy;
        set_segment2(s, P, pN[l], pE[l], cN, cE, l);
        endN[l] = g.wp_n[(int64_t)(s.n_wp[l] - 1) * S + i]; endE[l] = g.wp_e[(int64_t)(s.n_wp[l] - 1) * S + i];
    }
    ModeV MV;
    auto load_modes = [&]() {
        if (MULTI) {
            const ModeX &a = P.mode[g.mode[i0]], &b = P.mode[g.mode[i0 + 1]];
#define MV_(f) MV.f = make_float2(a.f, b.f)
            MV_(avail_total); MV_(avail_me); MV_(avail_el); MV_(cap_tq_me); MV_(cap_tq_el); MV_(add_me); MV_(cap_me);
            MV_(a_el); MV_(c_el); MV_(inv_me_cap); MV_(inv_el_cap); MV_(me_cap_kw);
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
    MapCache mc[2];
#pragma unroll
    for (int l = 0; l < 2; ++l) {
        mc[l].idx = (uint32_t)g.mc_idx[i0 + l]; mc[l].p0n = g.mc_n[i0 + l]; mc[l].p0e = g.mc_e[i0 + l]; mc[l].b2 = g.mc_b2[i0 + l];
    }
    const R rpm_per_rads = A.P[0].rpm_per_rads;

    // absolute float64 positions from the (possibly new) segment starts
    auto refresh_abs = [&]() {
#pragma unroll
        for (int l = 0; l < 2; ++l) {
            if (s.k[l] != kk[l]) {     // the ship switched segments inside the step: new float64 base
                kk[l] = s.k[l];
                pN[l] = g.wp_n[(int64_t)(kk[l] - 1) * S + idx(l)]; pE[l] = g.wp_e[(int64_t)(kk[l] - 1) * S + idx(l)];
            }
            aN[l] = abs_pos(pN[l], lane(s.X, l), lane(s.CX, l)); aE[l] = abs_pos(pE[l], lane(s.Y, l), lane(s.CY, l));
        }
    };

    for (int t = 0; t < K; ++t) {
        // ---- inputs: converted_action and SAC_update (MSRL_env_ex.py:402-405) -----------------------------------------
        bool su = false, live = true;
        if (ROLL) {
            if (pending) {
                if (have_act) { su = true; pending = false; have_act = false; tr_open = true; }   // action from sit_rollout_policy
                else live = false;
            }
            if (!__any_sync(FULL, live && active)) break;
        } else if (io.action) {
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
        if (live) {   // ---- phase A: the two ship steps ------------------------------------------------------------------
            n_live += 1;
            const bool init = (ROLL || io.init < 0) ? fresh : (t == 0 && io.init != 0);
            const bool frozen = stop_o;                                  // obs_step :285-310: a stopped obstacle ship is not advanced
            const double pre_n = aN[1], pre_e = aE[1];
            if (!frozen && su) {
                // obs_step :324-338 -> update_route controllers.py:298-303: list.insert(-1, wp)
                const int64_t i = i0 + 1;
                const int nw = s.n_wp[1];
                if (nw < g.max_wp) {
                    if (active) {
                        g.wp_n[(int64_t)nw * S + i] = endN[1]; g.wp_e[(int64_t)nw * S + i] = endE[1];
                        g.wp_n[(int64_t)(nw - 1) * S + i] = act_n; g.wp_e[(int64_t)(nw - 1) * S + i] = act_e;
                    }
                    s.n_wp[1] = nw + 1;
                    if (s.k[1] == nw - 1) set_segment2(s, P, pN[1], pE[1], (double)act_n, (double)act_e, 1);   // the segment in use ended at the old last point
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
            // test_step :217-283 (avoidance branch always taken) | obs_step :346-373
            if (MULTI) ship_step2<AVOID_LANE0, true>(P, MV, s, g, idx, so);
            else ship_step2<AVOID_LANE0, true>(P, P.mode[0], s, g, idx, so);
            steps[0] += 1;
            if (frozen) {
#define KP_(j, f) s.f.y = kp[j]
                KP_(0, X); KP_(1, Y); KP_(2, CX); KP_(3, CY); KP_(4, SX); KP_(5, SY); KP_(6, thr2); KP_(7, psi); KP_(8, u); KP_(9, v);
                KP_(10, r); KP_(11, om); KP_(12, ei_s); KP_(13, ei_w); KP_(14, ei_h); KP_(15, e_prev); KP_(16, e_ct); KP_(17, e_int);
                KP_(18, f_me); KP_(19, f_el); KP_(20, alpha); KP_(21, sa); KP_(22, ca);
#undef KP_
                s.k[1] = kpk; s.n_wp[1] = kpn;
                steps[1] += 2;                                           // frozen ship: next_time twice, :294,307
            } else {
                steps[1] += 1;
            }
            refresh_abs();
            if (frozen) {
                if (nstored >= 1) nstored = 2;                           // store_last_simulation_data repeats the last row :292
            } else {
                if (!init && nstored >= 1) {                             // :391-395, last two *stored* (pre-step) positions
                    const double dn = pre_n - prev_n, de = pre_e - prev_e;
                    const double d = sqrt(__dadd_rn(__dmul_rn(dn, dn), __dmul_rn(de, de)));
                    eps_dist += d; samp_dist += d;
                }
                prev_n = pre_n; prev_e = pre_e;
                nstored = nstored < 2 ? nstored + 1 : 2;
            }
            const double ddn = aN[0] - aN[1], dde = aE[0] - aE[1];
            ship_d2 = __dadd_rn(__dmul_rn(ddn, ddn), __dmul_rn(dde, dde));
        }

        // ---- phase B (whole warp): distance to the nearest ring + hull containment, nearest-ring cache, once per ship -------
        R pn[2], pe[2], d2map[2];
        bool inside[2];
#pragma unroll
        for (int l = 0; l < 2; ++l) {
            pn[l] = (R)aN[l]; pe[l] = (R)aE[l];
            d2map[l] = (R)3.0e38; inside[l] = false;
            if (n_edges > 0) {
                const float ddn_ = pn[l] - mc[l].p0n, dde_ = pe[l] - mc[l].p0e;
                unsigned nb = __ballot_sync(FULL, live && !(ddn_ * ddn_ + dde_ * dde_ < fabsf(mc[l].b2)));
                while (nb) {   // budget spent: the whole warp rescans that lane's position
                    const int L = __ffs(nb) - 1;
                    nb &= nb - 1;
                    const MapCache c = coop_scan(s_edges, n_edges, __shfl_sync(FULL, pn[l], L), __shfl_sync(FULL, pe[l], L), wl);
                    if (wl == L) mc[l] = c;
                }
                d2map[l] = cached_dist2(s_edges, mc[l].idx, pn[l], pe[l]);
                inside[l] = (__float_as_uint(mc[l].b2) >> 31) != 0u;   // far from every ring all four corners share the centre's side
                unsigned hb = __ballot_sync(FULL, live && d2map[l] <= EP.near2);
                while (hb) {   // is_pos_inside_obstacles, MSRL_env_ex.py:490-515: the four corners of the +-L/2 square
                    const int L = __ffs(hb) - 1;
                    hb &= hb - 1;
                    const bool in = coop_hull_inside(s_edges, n_edges, __shfl_sync(FULL, pn[l], L), __shfl_sync(FULL, pe[l], L),
                                                     EP.half_len[l], wl);
                    if (wl == L) inside[l] = in;
                }
            }
        }

        if (live) {   // ---- phase C: rewards, terminal checks, outputs, episode restart ---------------------------------------
            const R rpm_t = so.om_pre.x * rpm_per_rads;                  // stored (pre-step) observables of the test ship
            const R power_me_kw = so.load_me.x * 0.001f;
            const R avail_me_kw = MULTI ? MV.me_cap_kw.x : P.mode[0].me_cap_kw;
            const R ect_t = fabsf(s.e_ct.x), ect_o = fabsf(s.e_ct.y);
            const R dmap_t = sqrt_(d2map[0]), dmap_o = sqrt_(d2map[1]);
            bool dn_t = false, dn_o = false;
            int st = 0;
            // ---- test ship: test_ship_non_terminal_state_reward :628-664, test_ship_terminal_state_reward :734-809 --------
            R part_t;
            {
                const bool outside = pn[0] < EP.lo_n[0] || pn[0] > EP.hi_n[0] || pe[0] < EP.lo_e[0] || pe[0] > EP.hi_e[0];
                const double dnE = aN[0] - (double)endN[0], deE = aE[0] - (double)endE[0];
                const bool arrived = __dadd_rn(__dmul_rn(dnE, dnE), __dmul_rn(deE, deE)) <= EP.arrive2;
                const R ntt = ect_t * EP.inv_e_tol + ((R)1 - dmap_t * EP.inv_max_north) * (R)0.01;
                R rt = (R)0;
                if (arrived) { stop_t = true; dn_t = true; st |= SIT_ST_TEST_ENDPOINT; }
                if (outside) { stop_t = true; dn_t = true; st |= SIT_ST_TEST_HORIZON; }
                if (inside[0]) { if (!stop_t) rt += (R)1000; stop_t = true; dn_t = true; st |= SIT_ST_TEST_TERRAIN; }
                if (fabsf(rpm_t) > EP.rpm_max) { if (!stop_t) rt += (R)1000; stop_t = true; dn_t = true; st |= SIT_ST_TEST_MECH; }
                if (ect_t > EP.e_tol) { if (!stop_t) rt += (R)1000; stop_t = true; dn_t = true; st |= SIT_ST_TEST_NAV; }
                if (power_me_kw > avail_me_kw) { if (!stop_t) rt += (R)1000; stop_t = true; dn_t = true; st |= SIT_ST_TEST_BLACKOUT; }
                part_t = ntt + rt;
            }
            // ---- obstacle ship: obs_ship_non_terminal_state_reward :666-710 (stop flag from before this step's checks),
            //      obs_ship_terminal_state_reward :811-881, shared_non_terminal_state_reward :712-731 (flag after them) ------
            R nto = (R)0, ro_ = (R)0, snt = (R)0;
            {
                const bool outside = pn[1] < EP.lo_n[1] || pn[1] > EP.hi_n[1] || pe[1] < EP.lo_e[1] || pe[1] > EP.hi_e[1];
                const double dnE = aN[1] - (double)endN[1], deE = aE[1] - (double)endE[1];
                const bool arrived = __dadd_rn(__dmul_rn(dnE, dnE), __dmul_rn(deE, deE)) <= EP.arrive2;
                if (!stop_o) nto = (R)0.1 + (-(ect_o * EP.inv_e_tol) * (R)0.01) + (-((R)1 - dmap_o * EP.inv_max_north) * (R)0.01);
                if (arrived) { stop_o = true; st |= SIT_ST_OBS_ENDPOINT; }                       // no done
                if (outside) { stop_o = true; dn_o = true; st |= SIT_ST_OBS_HORIZON; }
                if (inside[1]) { if (!stop_o) ro_ -= (R)1000; dn_o = true; st |= SIT_ST_OBS_TERRAIN; }   // no stop flag
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
                    ob[0] = pn[0]; ob[B] = pe[0]; ob[2 * B] = s.psi.x; ob[3 * B] = rpm_t; ob[4 * B] = ect_t; ob[5 * B] = power_me_kw;
                    ob[6 * B] = pn[1]; ob[7 * B] = pe[1]; ob[8 * B] = s.psi.y; ob[9 * B] = ect_o;
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
                        row[12] = pn[0]; row[13] = pe[0]; row[14] = s.psi.x; row[15] = rpm_t; row[16] = ect_t; row[17] = power_me_kw;
                        row[18] = pn[1]; row[19] = pe[1]; row[20] = s.psi.y; row[21] = ect_o;
                        row[11] = reward; row[22] = timeup ? 1.0f : (done ? 0.0f : 1.0f);
                        n_closed += 1;
                    }
                    tr_open = false;
                }
                episode_end = done || (timeup && ro.truncate);
            }
            // ---- in-kernel reset() + init_step() right after a terminal step ---------------------------------------------
            if (episode_end) {
                if (ROLL && active) n_eps += 1;
                const bool hard = io.auto_reset == 2;   // MSRL_env_ex.py:145-186 keeps omega, integrators, fuel unless hard
                V w_speed, w_dir;
#pragma unroll
                for (int l = 0; l < 2; ++l) {
                    const int64_t i = i0 + l;
                    aN[l] = g.init_north[i]; aE[l] = g.init_east[i];
                    lane(s.psi, l) = g.init_psi[i]; lane(s.u, l) = g.init_u[i]; lane(s.v, l) = g.init_v[i]; lane(s.r, l) = g.init_r[i];
                    lane(s.vcn, l) = g.init_vc_n[i]; lane(s.vce, l) = g.init_vc_e[i];
                    lane(w_speed, l) = g.init_wind_speed[i]; lane(w_dir, l) = g.init_wind_dir[i];
                    lane(s.u_sp, l) = g.init_u_sp[i];
                    if (active) {
                        g.vc_n[i] = lane(s.vcn, l); g.vc_e[i] = lane(s.vce, l);
                        g.wind_speed[i] = lane(w_speed, l); g.wind_dir[i] = lane(w_dir, l); g.u_sp[i] = lane(s.u_sp, l);
                    }
                    steps[l] = 0; s.k[l] = 1; kk[l] = 1;
                    lane(s.e_ct, l) = 0.0f; lane(s.e_int, l) = 0.0f;
                    s.n_wp[l] = g.n_wp0[i];
                    for (int j = 0; j < g.max_wp; ++j) {
                        const R wn = g.wp0_n[(int64_t)j * S + i], we = g.wp0_e[(int64_t)j * S + i];
                        if (active) { g.wp_n[(int64_t)j * S + i] = wn; g.wp_e[(int64_t)j * S + i] = we; }
                    }
                    pN[l] = g.wp0_n[i]; pE[l] = g.wp0_e[i];
                    const double cN = g.wp0_n[S + i], cE = g.wp0_e[S + i];
                    split_rel(aN[l] - pN[l], lane(s.X, l), lane(s.CX, l));
                    split_rel(aE[l] - pE[l], lane(s.Y, l), lane(s.CY, l));
                    set_segment2(s, P, pN[l], pE[l], cN, cE, l);
                    endN[l] = g.wp0_n[(int64_t)(s.n_wp[l] - 1) * S + i]; endE[l] = g.wp0_e[(int64_t)(s.n_wp[l] - 1) * S + i];
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
                StepOut2 o2;
                if (MULTI) ship_step2<AVOID_NONE, false>(P, MV, s, g, idx, o2);      // init_step :188-215
                else ship_step2<AVOID_NONE, false>(P, P.mode[0], s, g, idx, o2);
                refresh_abs();
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
                        const double dn = (double)g.wp_n[(int64_t)s.k[1] * S + i] - aN[1], de = (double)g.wp_e[(int64_t)s.k[1] * S + i] - aE[1];
                        dec = __dadd_rn(__dmul_rn(dn, dn), __dmul_rn(de, de)) <= P.ra2;
                    }
                }
                if (dec) {
                    pending = true;
                    if (active) {
                        R* co = ge.cur_obs + env;
                        co[0] = pn[0]; co[B] = pe[0]; co[2 * B] = s.psi.x; co[3 * B] = rpm_t; co[4 * B] = ect_t; co[5 * B] = power_me_kw;
                        co[6 * B] = pn[1]; co[7 * B] = pe[1]; co[8 * B] = s.psi.y; co[9 * B] = ect_o;
                    }
                }
            }
        }   // live
        // ---- warp-level reduction of termination / collision flags ---------------------------------------------------------
        if (io.stats) {
            const bool cnt_lane = active && live;
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
    }

    // ---- write back ------------------------------------------------------------------------------------------------------
    if (active) {
        st2(g.psi, s.psi); st2(g.u, s.u); st2(g.v, s.v); st2(g.r, s.r); st2(g.omega, s.om);
        st2(g.ei_speed, s.ei_s); st2(g.ei_shaft, s.ei_w); st2(g.ei_heading, s.ei_h); st2(g.e_prev, s.e_prev);
        st2(g.e_ct, vabs(s.e_ct)); st2(g.e_ct_int, s.e_int);
        st2(g.fuel_me, s.f_me); st2(g.fuel_el, s.f_el); st2(g.fuel, add(add(s.f_me, s.f_el), f_off));
        st2(g.rel_n, s.X); st2(g.rel_cn, s.CX); st2(g.rel_e, s.Y); st2(g.rel_ce, s.CY);
#pragma unroll
        for (int l = 0; l < 2; ++l) {
            const int64_t i = i0 + l;
            g.north[i] = aN[l]; g.east[i] = aE[l];
            g.k[i] = s.k[l]; g.n_wp[i] = s.n_wp[l]; g.steps[i] = steps[l];
            g.mc_idx[i] = (int32_t)mc[l].idx; g.mc_n[i] = mc[l].p0n; g.mc_e[i] = mc[l].p0e; g.mc_b2[i] = mc[l].b2;
        }
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
