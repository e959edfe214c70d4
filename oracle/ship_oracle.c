/* CPU oracle (float64) for the ship-in-transit hot path -- TEST INFRASTRUCTURE, see ship_oracle.h.
 * Restatement in C of the reference's scalar Python step; literal formulation on purpose.
 * All file:line citations are relative to /root/reference/.
 */
#include "ship_oracle.h"

#include <math.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#ifndef M_PI
#define M_PI 3.14159265358979323846
#endif

static double sat(double val, double low, double hi)
{ /* controllers.py:64-69,95-100: max(low, min(val, hi)) */
    double m = val < hi ? val : hi;
    return low > m ? low : m;
}

static double clip(double x, double lo, double hi)
{ /* np.clip as used at MSRL_env_ex.py:245,249 */
    double m = x > lo ? x : lo;
    return m < hi ? m : hi;
}

void orc_derive(OrcParams* p)
{
    /* ship_model.py:71-80: payload, light-ship weight, mass, yaw inertia */
    double payload = 0.9 * (p->dead_weight_tonnage - p->bunkers);
    double lsw = p->dead_weight_tonnage / p->coefficient_of_deadweight_to_displacement - p->dead_weight_tonnage;
    p->mass = lsw + payload + p->bunkers + p->ballast;
    p->i_z = p->mass * (p->length_of_ship * p->length_of_ship + p->width_of_ship * p->width_of_ship) / 12;
    /* ship_model.py:193-209 set_added_mass */
    p->x_du = p->mass * p->added_mass_coeff_surge;
    p->y_dv = p->mass * p->added_mass_coeff_sway;
    p->n_dr = p->i_z * p->added_mass_coeff_yaw;
    /* ship_model.py:123-130 wind set-up */
    p->rho_a = 1.2;
    p->proj_area_f = p->width_of_ship * 8.0;
    p->proj_area_l = p->length_of_ship * 8.0;
    p->cx = 0.5; p->cy = 0.7; p->cn = 0.08;
    /* ship_engine.py:203 */
    p->rudder_max = p->max_rudder_angle_deg * M_PI / 180;
    /* ship_engine.py:32-44 update_available_propulsion_power */
    for (int i = 0; i < p->n_modes; ++i) {
        if (p->mode_sg_state[i] == ORC_SG_MOTOR) {
            p->avail_total[i] = p->mode_me_cap[i] + p->mode_el_cap[i] - p->hotel_load;
            p->avail_me[i] = p->mode_me_cap[i];
            p->avail_el[i] = p->mode_el_cap[i] - p->hotel_load;
        } else if (p->mode_sg_state[i] == ORC_SG_GEN) {
            p->avail_total[i] = p->mode_me_cap[i] - p->hotel_load;
            p->avail_me[i] = p->mode_me_cap[i] - p->hotel_load;
            p->avail_el[i] = 0;
        } else {
            p->avail_total[i] = p->mode_me_cap[i];
            p->avail_me[i] = p->mode_me_cap[i];
            p->avail_el[i] = 0;
        }
    }
}

/* ship_engine.py:46-76 MachineryMode.distribute_load */
static void distribute_load(const OrcParams* p, int mode, double load_perc,
                            double* load_me, double* load_el, double* pct_me, double* pct_el)
{
    double me_cap = p->mode_me_cap[mode], el_cap = p->mode_el_cap[mode], hotel = p->hotel_load;
    double total = load_perc * p->avail_total[mode];
    if (p->mode_sg_state[mode] == ORC_SG_MOTOR) {
        *load_me = total < me_cap ? total : me_cap;
        *load_el = total + hotel - *load_me;
        *pct_el = *load_el / el_cap;
        *pct_me = (me_cap == 0) ? 0.0 : *load_me / me_cap;
    } else if (p->mode_sg_state[mode] == ORC_SG_GEN) {
        *load_el = hotel < el_cap ? hotel : el_cap;
        *load_me = total + hotel - *load_el;
        *pct_me = *load_me / me_cap;
        *pct_el = (el_cap == 0) ? 0.0 : *load_el / el_cap;
    } else {
        *load_me = total;
        *load_el = hotel;
        *pct_me = *load_me / me_cap;
        *pct_el = *load_el / el_cap;
    }
}

/* ship_engine.py:256-261 spec_fuel_cons */
static double spec_fuel(double load_perc, const double c[3])
{
    double rate = c[0] * (load_perc * load_perc) + c[1] * load_perc + c[2];
    return rate / 3.6e9;
}

/* ship_engine.py:368-384 main_engine_torque / hsg_torque share this shape */
static double torque_limited(double load_perc, double avail, double omega)
{
    double a = load_perc * avail / (omega + 0.1);
    double b = avail / 5 * M_PI / 30;
    return a < b ? a : b;
}

void orc_ship_step(const OrcParams* p, OrcShip* s, int avoid, int store, double* log27,
                   double* rud_out, double* thr_out)
{
    const double dt = p->dt;
    const double N = s->north, E = s->east, psi = s->psi, u = s->u;

    /* --- LOS_guidance.py:88-103 next_wpt (squared distance, <=, end-of-route hold) --- */
    int k = s->k;
    {
        double dn = s->wp_n[k] - N, de = s->wp_e[k] - E;
        if (dn * dn + de * de <= p->ra * p->ra) {
            if (s->n_wp > k + 1) k = k + 1;
        }
    }
    s->k = k;

    /* --- LOS_guidance.py:105-121 los_guidance --- */
    double dx = s->wp_n[k] - s->wp_n[k - 1];
    double dy = s->wp_e[k] - s->wp_e[k - 1];
    double alpha = atan2(dy, dx);
    double e_ct = -(N - s->wp_n[k - 1]) * sin(alpha) + (E - s->wp_e[k - 1]) * cos(alpha);
    s->e_ct = fabs(e_ct);
    if (e_ct * e_ct >= p->lookahead * p->lookahead) e_ct = 0.99 * p->lookahead;  /* sign lost (quirk Q8) */
    double delta = sqrt(p->lookahead * p->lookahead - e_ct * e_ct);
    if (fabs(s->e_ct_int + e_ct / delta) <= p->windup_limit) s->e_ct_int += e_ct / delta;
    double chi = atan(-e_ct / delta - s->e_ct_int * p->ki_los);
    double heading_ref = alpha + chi;
    s->heading_ref = heading_ref;   /* controllers.py:312-313 */
    s->heading_mea = psi;

    /* --- controllers.py:81-93 pid_ctrl (no wrap, no anti-windup) + :188-189 --- */
    double err = heading_ref - psi;
    double d_err = (err - s->e_prev) / dt;
    double ei = s->ei_heading + err * dt;
    s->e_prev = err;
    s->ei_heading = ei;
    double rud = -(err * p->kp_heading + d_err * p->kd_heading + ei * p->ki_heading);
    rud = sat(rud, -p->rudder_max, p->rudder_max);

    /* --- controllers.py:138-143 throttle cascade; shaft PI measured input is u (quirk Q1,
     *     MSRL_env_ex.py:233-237,358-362) --- */
    double e1 = s->u_sp - u;
    s->ei_speed = s->ei_speed + e1 * dt;
    double thr;
    if (p->mach_kind == 1) {
        /* controllers.py:170-172 ThrottleFromSpeedSetPointSimplifiedPropulsion.throttle: pi_ctrl (:51-61) then
         * sat(val, 0, 1.1) = max(low, min(val, hi)) (:63-69) */
        thr = e1 * p->kp_simplified + s->ei_speed * p->ki_simplified;
        thr = sat(thr, 0.0, 1.1);
    } else {
        double omega_des = e1 * p->kp_speed + s->ei_speed * p->ki_speed;
        double e2 = omega_des - u;
        s->ei_shaft = s->ei_shaft + e2 * dt;
        thr = e2 * p->kp_shaft + s->ei_shaft * p->ki_shaft;
    }

    /* --- MSRL_env_ex.py:240-249 collision-avoidance branch (always taken for the test ship, Q2) --- */
    if (avoid) {
        thr *= 0.5;
        thr = clip(thr, 0.0, 1.1);
        rud += 3.0 * M_PI / 180.0;
        rud = clip(rud, -p->rudder_max, p->rudder_max);
    }
    if (rud_out) *rud_out = rud;
    if (thr_out) *thr_out = thr;

    const int mode = s->mode;
    const double omega = s->omega;

    /* --- ship_model.py:645-684 store_simulation_data (pre-step observables, fuel accumulators) --- */
    if (store) {
        double load_me, load_el, pct_me, pct_el;
        distribute_load(p, mode, thr, &load_me, &load_el, &pct_me, &pct_el);
        /* ship_engine.py:263-292 fuel_consumption */
        double rate_me = (load_me == 0) ? 0.0 : load_me * spec_fuel(pct_me, p->sfc_me);
        double rate_el = (pct_el == 0) ? 0.0 : load_el * spec_fuel(pct_el, p->sfc_dg);
        s->fuel_me = s->fuel_me + rate_me * dt;
        s->fuel_el = s->fuel_el + rate_el * dt;
        s->fuel = s->fuel + (rate_me + rate_el) * dt;
        if (log27) {
            log27[0] = s->time;
            log27[1] = N;
            log27[2] = E;
            log27[3] = psi * 180 / M_PI;
            log27[4] = rud * 180 / M_PI;
            log27[5] = u;
            log27[6] = s->v;
            log27[7] = s->r * 180 / M_PI;
            log27[8] = p->mach_kind == 1 ? 0.0 : omega * 30 / M_PI;   /* the simplified model has no shaft */
            log27[9] = pct_me;
            log27[10] = pct_el;
            log27[11] = load_me / 1000;
            log27[12] = p->mode_me_cap[mode] / 1000;
            log27[13] = load_el / 1000;
            log27[14] = p->mode_el_cap[mode] / 1000;
            log27[15] = (load_el + load_me) / 1000;
            log27[16] = (thr * p->avail_total[mode]) / 1000;
            log27[17] = rate_me;
            log27[18] = rate_el;
            log27[19] = rate_me + rate_el;
            log27[20] = s->fuel_me;
            log27[21] = s->fuel_el;
            log27[22] = s->fuel;
            log27[23] = p->mach_kind == 1 ? 0.0 : torque_limited(thr, p->avail_me[mode], omega);
            log27[24] = p->mach_kind == 1 ? omega / 1000 : pow(p->dp, 4) * p->kt * omega * fabs(omega) / 1000;
            log27[25] = s->e_ct;
            log27[26] = fabs(s->heading_mea - s->heading_ref);  /* radians, despite the key (Q12) */
        }
    }

    /* --- ship_model.py:233-250 three_dof_kinematics --- */
    const double c = cos(psi), sn = sin(psi);
    double d_north = c * u + (-sn) * s->v;
    double d_east = sn * u + c * s->v;
    double d_yaw = s->r;

    /* --- ship_engine.py:391-395,355-361 shaft equation --- */
    double d_omega, thrust;
    if (p->mach_kind == 1) {
        /* ship_engine.py:423-428 update_thrust_force: power = load_perc * (P_me + P_el);
         * d_thrust = (-k_thrust * thrust + power) / T, k_thrust = 2160 / 790 (:420); the kinetics get the PRE-update thrust
         * (the order of ShipModel.update_differentials, ship_model.py:484-490) */
        double power = thr * (p->avail_me[mode] + p->avail_el[mode]);
        d_omega = (-(2160.0 / 790.0) * omega + power) / p->thrust_time_constant;
        thrust = omega;
    } else {
        double t_me = torque_limited(thr, p->avail_me[mode], omega);
        double t_hsg = torque_limited(thr, p->avail_el[mode], omega);
        double eq_me = (t_me - p->d_me * omega) / p->r_me;
        double eq_hsg = (t_hsg - p->d_hsg * omega) / p->r_hsg;
        d_omega = (eq_me + eq_hsg - p->kp_torque * (omega * omega)) / p->jp;
        /* --- ship_engine.py:363-366 thrust (pre-update omega) --- */
        thrust = pow(p->dp, 4) * p->kt * omega * fabs(omega);
    }

    /* --- ship_model.py:576-622 kinetics --- */
    /* current to body frame: inv(R) * v_c */
    double uc = c * s->vc_n + sn * s->vc_e;
    double vc = -sn * s->vc_n + c * s->vc_e;
    double ur = u - uc, vr = s->v - vc;
    /* ship_model.py:608-622 rudder */
    double f_rud_v = -p->c_rudder_v * rud * (u - uc);
    double f_rud_r = -p->c_rudder_r * rud * (u - uc);
    /* ship_model.py:211-231 wind */
    double uw = s->wind_speed * cos(s->wind_dir - psi);
    double vw = s->wind_speed * sin(s->wind_dir - psi);
    double u_rw = uw - u, v_rw = vw - s->v;
    double gamma = -atan2(v_rw, u_rw);
    double wind_rw2 = u_rw * u_rw + v_rw * v_rw;
    double c_x = -p->cx * cos(gamma), c_y = p->cy * sin(gamma), c_n = p->cn * sin(2 * gamma);
    double tau_coeff = 0.5 * p->rho_a * wind_rw2;
    double tau_u = tau_coeff * c_x * p->proj_area_f;
    double tau_v = tau_coeff * c_y * p->proj_area_l;
    double tau_n = tau_coeff * c_n * p->proj_area_l * p->length_of_ship;
    /* ship_model.py:252-276 matrices with x_g = 0 (ship_model.py:79) */
    double m = p->mass, v = s->v, r = s->r;
    double cor0 = -m * v * r, cor1 = m * u * r, cor2 = m * v * u - m * u * v;
    double ca0 = p->y_dv * vr * r, ca1 = -p->x_du * ur * r, ca2 = -p->y_dv * vr * ur + p->x_du * ur * vr;
    double dm0 = (m / p->t_surge + p->ku * u) * ur;
    double dm1 = (m / p->t_sway + p->kv * v) * vr;
    double dm2 = (p->i_z / p->t_yaw + p->kr * r) * r;
    double d_u = (-cor0 - ca0 - dm0 + tau_u + thrust) / (m + p->x_du);
    double d_v = (-cor1 - ca1 - dm1 + tau_v + f_rud_v) / (m + p->y_dv);
    double d_r = (-cor2 - ca2 - dm2 + tau_n + f_rud_r) / (p->i_z + p->n_dr);

    /* --- ship_model.py:632-643, utils.py:50-53 explicit Euler (all derivatives from the old state) --- */
    s->north = N + d_north * dt;
    s->east = E + d_east * dt;
    s->psi = psi + d_yaw * dt;
    s->u = u + d_u * dt;
    s->v = v + d_v * dt;
    s->r = r + d_r * dt;
    s->omega = omega + d_omega * dt;
    /* utils.py:42-48 next_time (not called by init_step, MSRL_env_ex.py:188-215) */
    if (store) s->time = s->time + dt;
}

/* ---------------------------------------------------------------------------------------------
 * reset: ship_model.py:340-400,708-712 (pose, velocities, environment, time; machinery NOT
 * reset), controllers.py:145-151 (throttle reset is a no-op), controllers.py:333-350 +
 * LOS_guidance.py:123-136 (waypoint index, e_ct, e_ct_int, route reloaded), MSRL_env_ex.py:165
 * (desired_forward_speed).  Heading-PID integrator/prev_error, both PI integrators, omega and the
 * fuel accumulators keep their values (quirk Q3).
 * ------------------------------------------------------------------------------------------- */
void orc_ship_capture_init(OrcShip* s)
{
    s->init_north = s->north; s->init_east = s->east; s->init_psi = s->psi;
    s->init_u = s->u; s->init_v = s->v; s->init_r = s->r;
    s->init_vc_n = s->vc_n; s->init_vc_e = s->vc_e;
    s->init_wind_speed = s->wind_speed; s->init_wind_dir = s->wind_dir;
    s->init_u_sp = s->u_sp;
    s->n_wp0 = s->n_wp;
    memcpy(s->wp0_n, s->wp_n, sizeof(s->wp_n));
    memcpy(s->wp0_e, s->wp_e, sizeof(s->wp_e));
}

void orc_ship_reset(OrcShip* s)
{
    s->north = s->init_north; s->east = s->init_east; s->psi = s->init_psi;
    s->u = s->init_u; s->v = s->init_v; s->r = s->init_r;
    s->vc_n = s->init_vc_n; s->vc_e = s->init_vc_e;
    s->wind_speed = s->init_wind_speed; s->wind_dir = s->init_wind_dir;
    s->time = 0.0;
    s->k = 1;
    s->heading_ref = 0.0; s->heading_mea = 0.0;
    s->e_ct = 0.0; s->e_ct_int = 0.0;
    s->n_wp = s->n_wp0;
    memcpy(s->wp_n, s->wp0_n, sizeof(s->wp_n));
    memcpy(s->wp_e, s->wp0_e, sizeof(s->wp_e));
    s->u_sp = s->init_u_sp;
}

void orc_run_ships(const OrcParams* p, OrcShip* ships, long n, int nsteps, int avoid,
                   const double* dist, double* traj, int threads)
{
#ifdef _OPENMP
    if (threads <= 0) threads = omp_get_max_threads();
#pragma omp parallel for schedule(static) num_threads(threads)
#endif
    for (long i = 0; i < n; ++i) {
        OrcShip* s = &ships[i];
        for (int t = 0; t < nsteps; ++t) {
            if (dist) {
                const double* d = dist + (size_t)t * 4 * n;
                s->vc_n = d[i]; s->vc_e = d[n + i]; s->wind_speed = d[2 * n + i]; s->wind_dir = d[3 * n + i];
            }
            orc_ship_step(p, s, avoid, 1, 0, 0, 0);
            if (traj) {
                double* o = traj + (size_t)t * 8 * n;
                o[i] = s->north; o[n + i] = s->east; o[2 * n + i] = s->psi; o[3 * n + i] = s->u;
                o[4 * n + i] = s->v; o[5 * n + i] = s->r; o[6 * n + i] = s->omega; o[7 * n + i] = (double)s->k;
            }
        }
    }
}

/* ---------------------------------------------------------------------------------------------
 * Polygon map: restatement of the two shapely/GEOS calls at obstacle.py:126-129,138-141.
 * PARITY UNPINNED (see oracle/geometry.py): GEOS RayCrossingCounter + Distance::pointToSegment.
 * ------------------------------------------------------------------------------------------- */
static int orient_sign(double ax, double ay, double bx, double by, double px, double py)
{
    double d = (bx - ax) * (py - ay) - (by - ay) * (px - ax);
    return (d > 0) - (d < 0);
}

/* obstacle.py:126-129: any(poly.contains(Point(e_pos, n_pos))) -- interior only */
int orc_map_contains(const OrcMap* m, double n_pos, double e_pos)
{
    const double px = e_pos, py = n_pos;
    int i = 0;
    while (i < m->n_edges) {
        const int pid = m->poly[i];
        int crossings = 0, boundary = 0;
        for (; i < m->n_edges && m->poly[i] == pid; ++i) {
            if (boundary) continue;
            double x1 = m->x1[i], y1 = m->y1[i], x2 = m->x2[i], y2 = m->y2[i];
            if (x1 == x2 && y1 == y2) continue;
            if (x1 < px && x2 < px) continue;
            if (px == x2 && py == y2) { boundary = 1; continue; }
            if (y1 == py && y2 == py) {
                double lo = x1 <= x2 ? x1 : x2, hi = x1 <= x2 ? x2 : x1;
                if (lo <= px && px <= hi) boundary = 1;
                continue;
            }
            if ((y1 > py && y2 <= py) || (y2 > py && y1 <= py)) {
                int s = orient_sign(x1, y1, x2, y2, px, py);
                if (s == 0) { boundary = 1; continue; }
                if (y2 < y1) s = -s;
                if (s > 0) crossings++;
            }
        }
        if (!boundary && (crossings & 1)) return 1;
    }
    return 0;
}

static double seg_distance(double px, double py, double ax, double ay, double bx, double by)
{
    if (ax == bx && ay == by) return hypot(px - ax, py - ay);
    double len2 = (bx - ax) * (bx - ax) + (by - ay) * (by - ay);
    double r = ((px - ax) * (bx - ax) + (py - ay) * (by - ay)) / len2;
    if (r <= 0.0) return hypot(px - ax, py - ay);
    if (r >= 1.0) return hypot(px - bx, py - by);
    double s = ((ay - py) * (bx - ax) - (ax - px) * (by - ay)) / len2;
    return fabs(s) * sqrt(len2);
}

/* obstacle.py:138-141: min(poly.exterior.distance(Point(e_ship, n_ship))) */
double orc_map_distance(const OrcMap* m, double n_ship, double e_ship)
{
    double best = INFINITY;
    for (int i = 0; i < m->n_edges; ++i) {
        double d = seg_distance(e_ship, n_ship, m->x1[i], m->y1[i], m->x2[i], m->y2[i]);
        if (d < best) best = d;
    }
    return best;
}

/* ---------------------------------------------------------------------------------------------
 * Two-ship env: RLEnv/MSRL_env_ex.py
 * ------------------------------------------------------------------------------------------- */

/* MSRL_env_ex.py:117-141 reward_function_params (uses the obstacle ship's CURRENT route) */
void orc_env_reset_params(OrcEnv* e)
{
    const OrcShip* o = &e->obs;
    double dn = o->wp_n[o->n_wp - 1] - o->wp_n[0];
    double de = o->wp_e[o->n_wp - 1] - o->wp_e[0];
    double ab = sqrt(dn * dn + de * de);
    e->e_tolerance = 1000;
    e->ab_segment_length = ab / e->sampling_frequency;
    e->ab_alpha = atan2(de, dn);
}

/* MSRL_env_ex.py:145-186 */
void orc_env_reset(OrcEnv* e)
{
    orc_ship_reset(&e->test);
    orc_ship_reset(&e->obs);
    e->stop_test = 0; e->stop_obs = 0;
    e->n_stored_obs = 0;                 /* ship_model.py:708-712 clears simulation_results */
    e->eps_distance_travelled = 0; e->sampling_distance_travelled = 0;
    orc_env_reset_params(e);
}

/* MSRL_env_ex.py:188-215: both ships advance one dt without storing, without next_time and
 * without the collision-avoidance branch (quirk Q7) */
void orc_env_init_step(const OrcParams* pt, const OrcParams* po, OrcEnv* e)
{
    orc_ship_step(pt, &e->test, 0, 0, 0, 0, 0);
    orc_ship_step(po, &e->obs, 0, 0, 0, 0, 0);
}

/* MSRL_env_ex.py:460-488 */
static int pos_outside_horizon(const OrcMap* m, double n_pos, double e_pos, double ship_length)
{
    double margin = ship_length / 2;
    int outside_n = n_pos < m->min_north + margin || n_pos > m->max_north - margin;
    int outside_e = e_pos < m->min_east + margin || e_pos > m->max_east - margin;
    return outside_n || outside_e;
}

/* MSRL_env_ex.py:490-515: 4 corners of the +-L/2 square */
static int pos_inside_obstacles(const OrcMap* m, double n_pos, double e_pos, double ship_length)
{
    double margin = ship_length / 2;
    double n0 = n_pos - margin, e0 = e_pos - margin, n1 = n_pos + margin, e1 = e_pos + margin;
    int inside = 0;
    if (orc_map_contains(m, n0, e0)) inside = 1;
    if (orc_map_contains(m, n0, e1)) inside = 1;
    if (orc_map_contains(m, n1, e0)) inside = 1;
    if (orc_map_contains(m, n1, e1)) inside = 1;
    return inside;
}

/* MSRL_env_ex.py:517-542 */
static int route_outside_horizon(const OrcMap* m, double n_route, double e_route)
{
    int outside_n = n_route < m->min_north || n_route > m->max_north;
    int outside_e = e_route < m->min_east || e_route > m->max_east;
    return outside_n || outside_e;
}

void orc_env_step(const OrcParams* pt, const OrcParams* po, const OrcMap* m, OrcEnv* e,
                  double act_n, double act_e, int sac_update, int init,
                  double obs10[10], double* reward, int* done, int* status)
{
    double lg[ORC_N_LOG];

    /* ---- test_step, MSRL_env_ex.py:217-283 (avoidance branch always taken, quirk Q2) ---- */
    orc_ship_step(pt, &e->test, 1, 1, lg, 0, 0);
    e->last_rpm[0] = lg[8]; e->last_ect[0] = lg[25]; e->last_power_me[0] = lg[11];
    e->last_avail_me = lg[12];

    /* ---- obs_step, MSRL_env_ex.py:285-400 ---- */
    OrcShip* o = &e->obs;
    if (e->stop_obs) {
        /* :291-322 frozen ship: store_last_simulation_data + next_time twice (quirk Q5) */
        if (e->n_stored_obs >= 1) {
            e->stored_n[0] = e->stored_n[1]; e->stored_e[0] = e->stored_e[1];
            if (e->n_stored_obs < 2) e->n_stored_obs = 2;
        }
        o->time = o->time + po->dt;
        o->time = o->time + po->dt;
    } else {
        if (sac_update) {
            /* :324-338 -> controllers.py:298-303: list.insert(-1, .) = before the end point */
            if (o->n_wp < ORC_MAX_WP) {
                o->wp_n[o->n_wp] = o->wp_n[o->n_wp - 1];
                o->wp_e[o->n_wp] = o->wp_e[o->n_wp - 1];
                o->wp_n[o->n_wp - 1] = act_n;
                o->wp_e[o->n_wp - 1] = act_e;
                o->n_wp += 1;
            }
            e->sampling_distance_travelled = 0;
        }
        orc_ship_step(po, o, 0, 1, lg, 0, 0);
        e->last_rpm[1] = lg[8]; e->last_ect[1] = lg[25]; e->last_power_me[1] = lg[11];
        /* history of stored (pre-step) positions, last two kept */
        if (e->n_stored_obs >= 1) { e->stored_n[0] = e->stored_n[1]; e->stored_e[0] = e->stored_e[1]; }
        e->stored_n[1] = lg[1]; e->stored_e[1] = lg[2];
        if (e->n_stored_obs < 2) e->n_stored_obs += 1;
        if (!init && e->n_stored_obs >= 2) {
            /* :391-395 */
            double dn = e->stored_n[1] - e->stored_n[0], de = e->stored_e[1] - e->stored_e[0];
            double d = sqrt(dn * dn + de * de);
            e->eps_distance_travelled += d;
            e->sampling_distance_travelled += d;
        }
    }

    /* ---- :424-435 observation: post-step pose, pre-step stored rpm / e_ct / power (Q4) ---- */
    obs10[0] = e->test.north; obs10[1] = e->test.east; obs10[2] = e->test.psi;
    obs10[3] = e->last_rpm[0]; obs10[4] = e->last_ect[0]; obs10[5] = e->last_power_me[0];
    obs10[6] = o->north; obs10[7] = o->east; obs10[8] = o->psi; obs10[9] = e->last_ect[1];

    /* ---- reward_function, MSRL_env_ex.py:906-980 ---- */
    const double n_t = obs10[0], e_t = obs10[1], rpm_t = obs10[3], ect_t = obs10[4], pw_t = obs10[5];
    const double n_o = obs10[6], e_o = obs10[7], ect_o = obs10[9];
    const double avail = e->last_avail_me;   /* :920 */
    int st = 0;

    /* :628-664 test ship non-terminal */
    double r_ect_t = fabs(ect_t) / e->e_tolerance;
    double r_col_t = (1 - orc_map_distance(m, n_t, e_t) / m->max_north) / 100;
    double reward_ntt = r_ect_t + r_col_t;

    /* :734-809 test ship terminal (sequential stop_flag gating, quirk Q10) */
    double rt = 0; int done_t = 0;
    {
        const OrcShip* t = &e->test;
        double dn = n_t - t->wp_n[t->n_wp - 1], de = e_t - t->wp_e[t->n_wp - 1];
        double rel = sqrt(dn * dn + de * de);
        if (rel <= 200) { e->stop_test = 1; done_t = 1; st |= ORC_ST_TEST_ENDPOINT; }
        if (pos_outside_horizon(m, n_t, e_t, pt->length_of_ship)) { e->stop_test = 1; done_t = 1; st |= ORC_ST_TEST_HORIZON; }
        if (pos_inside_obstacles(m, n_t, e_t, pt->length_of_ship)) {
            if (!e->stop_test) rt += 1000;
            e->stop_test = 1; done_t = 1; st |= ORC_ST_TEST_TERRAIN;
        }
        if (fabs(rpm_t) > 2000) {                       /* :554-558 */
            if (!e->stop_test) rt += 1000;
            e->stop_test = 1; done_t = 1; st |= ORC_ST_TEST_MECH;
        }
        if (fabs(ect_t) > e->e_tolerance) {             /* :560-564 */
            if (!e->stop_test) rt += 1000;
            e->stop_test = 1; done_t = 1; st |= ORC_ST_TEST_NAV;
        }
        if (pw_t > avail) {                             /* :578-582 */
            if (!e->stop_test) rt += 1000;
            e->stop_test = 1; done_t = 1; st |= ORC_ST_TEST_BLACKOUT;
        }
    }

    /* :666-710 obstacle ship non-terminal (sees stop_flag from before its own terminal check) */
    double reward_nto = 0;
    if (!e->stop_obs) {
        double r_base = 0.1;
        double r_ect_o = -(fabs(ect_o) / e->e_tolerance) / 100;
        double r_col_o = -(1 - orc_map_distance(m, n_o, e_o) / m->max_north) / 100;
        reward_nto = r_base + r_ect_o + r_col_o;
    }

    /* :811-881 obstacle ship terminal */
    double ro = 0; int done_o = 0;
    {
        double dn = n_o - o->wp_n[o->n_wp - 1], de = e_o - o->wp_e[o->n_wp - 1];
        double rel = sqrt(dn * dn + de * de);
        if (rel <= 200) { e->stop_obs = 1; st |= ORC_ST_OBS_ENDPOINT; }        /* no done */
        if (pos_outside_horizon(m, n_o, e_o, po->length_of_ship)) { e->stop_obs = 1; done_o = 1; st |= ORC_ST_OBS_HORIZON; }
        if (pos_inside_obstacles(m, n_o, e_o, po->length_of_ship)) {
            if (!e->stop_obs) ro += -1000;
            done_o = 1; st |= ORC_ST_OBS_TERRAIN;                             /* no stop_flag */
        }
        if (route_outside_horizon(m, act_n, act_e) || orc_map_contains(m, act_n, act_e)) {
            if (!e->stop_obs) ro += -1000;
            e->stop_obs = 1; done_o = 1; st |= ORC_ST_OBS_IW;
        }
        /* :566-576 */
        if (fabs(ect_o) > e->e_tolerance || e->sampling_distance_travelled > e->ab_segment_length * e->theta) {
            if (!e->stop_obs) ro += -1000;
            e->stop_obs = 1; done_o = 1; st |= ORC_ST_OBS_NAV;
        }
    }

    /* :712-731 shared non-terminal (after the obstacle terminal check, quirk Q11) */
    double reward_snt = 0;
    double ddn = n_t - n_o, dde = e_t - e_o;
    if (!e->stop_obs) {
        double ship_distance = sqrt(ddn * ddn + dde * dde);
        reward_snt = (1 - ship_distance / m->max_north) / 1000;
    }
    /* :883-904, 584-603 shared terminal */
    double rs = 0; int done_s = 0;
    if (ddn * ddn + dde * dde < 50.0 * 50.0) {
        rs += 2000; st |= ORC_ST_COLLISION; e->stop_test = 1; e->stop_obs = 1; done_s = 1;
    }

    *reward = reward_ntt + rt + reward_nto + ro + reward_snt + rs;   /* :970-972, left to right */
    *done = done_t || done_o || done_s;
    *status = st;
}

void orc_run_envs(const OrcParams* pt, const OrcParams* po, const OrcMap* m, OrcEnv* envs, long n,
                  int nsteps, const double* act, const unsigned char* sac, int first_is_init,
                  double* obs, double* reward, unsigned char* done, int* status, int threads)
{
#ifdef _OPENMP
    if (threads <= 0) threads = omp_get_max_threads();
#pragma omp parallel for schedule(static) num_threads(threads)
#endif
    for (long i = 0; i < n; ++i) {
        for (int t = 0; t < nsteps; ++t) {
            double a_n = act ? act[((size_t)t * 2 + 0) * n + i] : 0.0;
            double a_e = act ? act[((size_t)t * 2 + 1) * n + i] : 0.0;
            int su = sac ? sac[(size_t)t * n + i] : 0;
            double o10[10], r; int d, s;
            orc_env_step(pt, po, m, &envs[i], a_n, a_e, su, first_is_init && t == 0, o10, &r, &d, &s);
            if (obs) for (int c = 0; c < 10; ++c) obs[((size_t)t * 10 + c) * n + i] = o10[c];
            if (reward) reward[(size_t)t * n + i] = r;
            if (done) done[(size_t)t * n + i] = (unsigned char)d;
            if (status) status[(size_t)t * n + i] = s;
        }
    }
}

int orc_sizeof_params(void) { return (int)sizeof(OrcParams); }
int orc_sizeof_ship(void) { return (int)sizeof(OrcShip); }
int orc_sizeof_env(void) { return (int)sizeof(OrcEnv); }
int orc_sizeof_map(void) { return (int)sizeof(OrcMap); }
int orc_max_threads(void)
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
