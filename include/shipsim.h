/* libshipsim -- C ABI of the B200-native batched ship-in-transit simulator.
 *
 * This header is the drop-in boundary for the hot path of AndreasKing-Goks/sac-maritime-ast:
 * the scalar Python ship simulator (simulators/ship_in_transit/) and the two-ship RL env step
 * (RLEnv/MSRL_env_ex.py).  The reference has no FFI of its own (pure Python); every entry point
 * below cites the reference interface it replaces.  Conventions:
 *   - extern "C", plain pointers and sizes, no torch / C++ types in any signature;
 *   - returns 0 on success or a negative SitError; never throws; sit_last_error() gives the
 *     thread-local message of the last failure;
 *   - a handle owns no SIMULATION memory: the caller allocates ONE device buffer of
 *     sit_state_bytes() bytes and passes it to every call (layout: sit_layout());
 *   - device entry points only enqueue work on the caller's CUDA stream (a cudaStream_t passed
 *     as void*; NULL = legacy default stream) and never synchronise;
 *   - what a handle does own is SCRATCH, allocated lazily and kept until sit_destroy: the staging
 *     buffers + 4 internal streams of the *_host entry points (first *_host call), and the policy
 *     scratch of sit_rollout_policy / sit_rollout_collect (compact env list, MLP inputs / outputs,
 *     the packed bf16 W1 image: ~60 B per env + 128 KiB, first rollout call; cudaMalloc is a
 *     device-wide synchronisation, so that FIRST call is not "enqueue only").  sit_rollout_collect
 *     runs plain launches (one iteration, or with several env groups its whole first call) before it
 *     captures its CUDA graph for exactly this reason;
 *   - the *_host entry points take HOST buffers, copy in, run, copy out and synchronise;
 *   - there is no CPU fallback: without a CUDA device every compute entry point fails.
 */
#ifndef SHIPSIM_H
#define SHIPSIM_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SIT_ABI_VERSION 3
#define SIT_MAX_WP 16     /* route capacity per ship: 2 + sampling_frequency (main_ast.py:47) fits */
#define SIT_MAX_MODES 4   /* machinery modes per ship (PTO / PTI / MEC, main_ast.py:136-158) */
#define SIT_MAX_EDGES 128 /* polygon ring segments of the map (default map: 39, main_ast.py:200-205) */
#define SIT_N_OBS 10      /* observation vector, MSRL_env_ex.py:424-435 */
#define SIT_N_REWARD_PARTS 8
#define SIT_N_TRAJ 8      /* trajectory channels N, E, psi, u, v, r, omega, k */
#define SIT_N_LOG 27      /* simulation_results channels, ship_model.py:645-684 */
#define SIT_N_STATS 16    /* 12 status-bit counters + done, steps, decisions, reserved */

typedef enum SitError {
    SIT_OK = 0,
    SIT_E_INVALID = -1,   /* bad argument / config */
    SIT_E_CUDA = -2,      /* CUDA runtime error (message in sit_last_error) */
    SIT_E_NODEVICE = -3,  /* no CUDA device: there is no CPU fallback */
    SIT_E_KIND = -4       /* call not valid for this handle kind */
} SitError;

enum { SIT_F64 = 0, SIT_F32 = 1 };            /* arithmetic of the step kernels */
enum { SIT_KIND_FLEET = 1, SIT_KIND_ENV = 2 };/* bare ships | two-ship AST envs */
enum { SIT_SG_MOTOR = 0, SIT_SG_GEN = 1, SIT_SG_OFF = 2 }; /* ship_engine.py:32-44 */
enum { SIT_ROLE_TEST = 0, SIT_ROLE_OBS = 1 };
/* machinery model of a ship: ShipMachineryModel (ship_engine.py:298-395: shaft-speed state, torque caps) or
 * SimplifiedMachineryModel (ship_engine.py:398-433: thrust force as a first-order lag of the commanded power, driven by
 * ThrottleFromSpeedSetPointSimplifiedPropulsion controllers.py:154-172).  The simplified model is a FLEET-only variant: the
 * reference's MultiShipRLEnv calls throttle(speed_set_point, measured_speed, measured_shaft_speed) and reads the shaft
 * rpm, which only the full model has. */
enum { SIT_MACH_FULL = 0, SIT_MACH_SIMPLIFIED = 1 };

/* status bits; strings at MSRL_env_ex.py:760-899 (sit_status_string) */
enum {
    SIT_ST_TEST_ENDPOINT = 1 << 0, SIT_ST_TEST_HORIZON = 1 << 1, SIT_ST_TEST_TERRAIN = 1 << 2,
    SIT_ST_TEST_MECH = 1 << 3, SIT_ST_TEST_NAV = 1 << 4, SIT_ST_TEST_BLACKOUT = 1 << 5,
    SIT_ST_OBS_ENDPOINT = 1 << 6, SIT_ST_OBS_HORIZON = 1 << 7, SIT_ST_OBS_TERRAIN = 1 << 8,
    SIT_ST_OBS_IW = 1 << 9, SIT_ST_OBS_NAV = 1 << 10, SIT_ST_COLLISION = 1 << 11
};

/* Uniform parameters of one ship role: the reference's config NamedTuples flattened.
 * ShipConfiguration ship_model.py:20-35; MachinerySystemConfiguration + MachineryModeParams
 * ship_engine.py:17-20,121-138; ThrottleControllerGains / HeadingControllerGains
 * controllers.py:16-31; LosParameters LOS_guidance.py:15-19; SimulationConfiguration
 * .integration_step ship_model.py:52.  Derived constants (mass, inertia, added mass, wind areas,
 * available power per mode: ship_model.py:71-86,123-130,193-209; ship_engine.py:32-44) are
 * computed inside sit_create. */
typedef struct SitShipParams {
    double dt;
    double dead_weight_tonnage, coefficient_of_deadweight_to_displacement, bunkers, ballast;
    double length_of_ship, width_of_ship;
    double added_mass_coefficient_in_surge, added_mass_coefficient_in_sway, added_mass_coefficient_in_yaw;
    double mass_over_linear_friction_coefficient_in_surge, mass_over_linear_friction_coefficient_in_sway,
           mass_over_linear_friction_coefficient_in_yaw;
    double nonlinear_friction_coefficient__in_surge, nonlinear_friction_coefficient__in_sway,
           nonlinear_friction_coefficient__in_yaw;
    int32_t n_modes, machinery_kind;   /* SIT_MACH_* */
    double mode_main_engine_capacity[SIT_MAX_MODES], mode_electrical_capacity[SIT_MAX_MODES];
    int32_t mode_shaft_generator_state[SIT_MAX_MODES];
    double hotel_load;
    double linear_friction_main_engine, linear_friction_hybrid_shaft_generator;
    double gear_ratio_between_main_engine_and_propeller, gear_ratio_between_hybrid_shaft_generator_and_propeller;
    double propeller_inertia, propeller_speed_to_torque_coefficient, propeller_diameter,
           propeller_speed_to_thrust_force_coefficient;
    double rudder_angle_to_sway_force_coefficient, rudder_angle_to_yaw_force_coefficient, max_rudder_angle_degrees;
    double sfc_me[3], sfc_dg[3];                       /* FuelConsumptionCoefficients a, b, c */
    double kp_ship_speed, ki_ship_speed, kp_shaft_speed, ki_shaft_speed;
    double kp_heading, kd_heading, ki_heading;
    double radius_of_acceptance, lookahead_distance, los_integral_gain, los_integrator_windup_limit;
    /* SIT_MACH_SIMPLIFIED only: SimplifiedPropulsionMachinerySystemConfiguration.thrust_force_dynamic_time_constant
     * (ship_engine.py:148-157) and the kp, ki of ThrottleFromSpeedSetPointSimplifiedPropulsion (controllers.py:160-168).
     * The thrust force [N] is the machinery state: it lives in the state field "omega" (initial_thrust_force at reset). */
    double thrust_force_dynamic_time_constant, kp_simplified, ki_simplified;
} SitShipParams;

/* PolygonObstacle (obstacle.py:92-148): ring segments in (x, y) = (east, north), one poly id per
 * segment; segments of one polygon are contiguous. */
typedef struct SitMap {
    int32_t n_edges, n_poly;
    double x1[SIT_MAX_EDGES], y1[SIT_MAX_EDGES], x2[SIT_MAX_EDGES], y2[SIT_MAX_EDGES];
    int32_t poly[SIT_MAX_EDGES];
} SitMap;

typedef struct SitConfig {
    int32_t abi_version;      /* SIT_ABI_VERSION */
    int32_t kind;             /* SIT_KIND_FLEET or SIT_KIND_ENV */
    int32_t precision;        /* SIT_F64 or SIT_F32 (positions are carried in f64 in both) */
    int32_t max_wp;           /* route capacity actually allocated, 2..SIT_MAX_WP */
    SitShipParams ship[2];    /* [SIT_ROLE_TEST], [SIT_ROLE_OBS]; a fleet uses ship[0] */
    SitMap map;               /* envs only */
    double sampling_frequency;/* args.sampling_frequency, MSRL_env_ex.py:125 */
    double theta;             /* args.theta, MSRL_env_ex.py:130 */
} SitConfig;

typedef struct SitHandle_* SitHandle;

/* one named SoA array inside the state buffer */
enum { SIT_DT_F64 = 0, SIT_DT_F32 = 1, SIT_DT_I32 = 2 };
typedef struct SitField {
    char name[32];
    int64_t offset;     /* bytes from the start of the state buffer */
    int32_t dtype;      /* SIT_DT_* */
    int32_t rows;       /* 1, or max_wp for route arrays: array is [rows][count] */
    int64_t count;      /* elements per row: n_ships for ship fields, n_envs for env fields */
} SitField;

/* ---- lifetime ------------------------------------------------------------------------------ */
/* Replaces constructing ShipModelAST / controllers / MultiShipRLEnv (ship_model.py:545-574,
 * controllers.py:108-136,253-296, MSRL_env_ex.py:40-114) for n_envs independent copies. */
int sit_create(const SitConfig* cfg, int64_t n_envs, int device, SitHandle* out);
int sit_destroy(SitHandle h);
int sit_state_bytes(SitHandle h, size_t* nbytes);
/* fields == NULL: only *n_fields is written */
int sit_layout(SitHandle h, SitField* fields, int32_t capacity, int32_t* n_fields);
int64_t sit_n_ships(SitHandle h);

/* ---- bare ships (simulators/ship_in_transit) ------------------------------------------------- */
typedef struct SitFleetIO {
    /* inputs, device pointers, may be NULL */
    const void* dist;        /* [dist_steps][4][n_ships] reals: current north, current east, wind speed,
                                wind dir assigned before each step (ship_model.py:154-156); dist_steps is
                                1 (held for the launch) or K */
    int32_t dist_steps;
    int32_t avoid;           /* 1: collision-avoidance branch MSRL_env_ex.py:240-249 on every ship */
    int32_t store;           /* 1: store_simulation_data + next_time (normal step); 0: init_step (MSRL_env_ex.py:188-215) */
    int32_t traj_steps;      /* traj holds K rows (0 or K: every step) or 1 row (the last step of the launch) */
    /* outputs, device pointers, may be NULL */
    void* traj;              /* [traj_steps][SIT_N_TRAJ][n_ships] reals, post-step */
    void* log27;             /* [K][SIT_N_LOG][n_ships] reals, the 27 stored channels of each step */
    void* cmd;               /* [K][2][n_ships] reals: rudder, throttle commanded in each step */
} SitFleetIO;

/* K ship-steps for every ship: guidance (LOS_guidance.py:88-121) -> heading PID
 * (controllers.py:72-100,175-189,306-314) -> throttle cascade (controllers.py:138-143) -> store
 * (ship_model.py:645-684, ship_engine.py:239-292) -> update_differentials / integrate
 * (ship_model.py:576-643, ship_engine.py:355-395) -> next_time (utils.py:42-48). */
int sit_fleet_step_k(SitHandle h, void* state, int32_t K, const SitFleetIO* io, void* stream);
/* Host-buffer form of sit_fleet_step_k for callers that keep their arrays on the host (the reference's
 * callers do): dist_h [4][n_ships] reals (may be NULL) is copied in and held for the K steps, the post-step
 * (N, E, psi, u, v, r, omega, k) of the LAST step is copied out to pose_h [SIT_N_TRAJ][n_ships]; synchronises. */
int sit_fleet_step_host(SitHandle h, void* state, int32_t K, const void* dist_h, int32_t avoid, void* pose_h,
                        void* stream);
/* Same, copying out only what the caller reads: pose_h receives the channels whose bit is set in channel_mask
 * (bit c = channel c of N, E, psi, u, v, r, omega, k), packed as [popcount(mask)][n_ships] rows in channel order;
 * k8_h (may be NULL) receives the waypoint index as int8 [n_ships] -- one byte per ship instead of a real.  The
 * reference's callers read the pose and speed of a ship between steps (ship_model.north / east / yaw_angle /
 * forward_speed, MSRL_env_ex.py:255-278): mask 0x0f + k8_h is 17 bytes per fp32 ship instead of 32. */
int sit_fleet_step_host_sel(SitHandle h, void* state, int32_t K, const void* dist_h, int32_t avoid, void* pose_h,
                            int32_t channel_mask, int8_t* k8_h, void* stream);
/* snapshot pose / environment / route as the values reset() restores (the deepcopy at
 * MSRL_env_ex.py:55-56 and the init_* attributes set in ship_model.py:132-157) */
int sit_capture_init(SitHandle h, void* state, void* stream);
/* ShipModelAST.reset + controller resets (ship_model.py:340-400,708-712; controllers.py:145-151,
 * 333-350; LOS_guidance.py:123-136).  mask: NULL = all, else uint8[n_envs].  hard != 0 also
 * restores what the reference leaves alone (omega, integrators, fuel): a fresh construction. */
int sit_reset(SitHandle h, void* state, const uint8_t* mask, int32_t hard, void* stream);

/* ---- two-ship AST env (RLEnv/MSRL_env_ex.py MultiShipRLEnv) ---------------------------------- */
typedef struct SitEnvIO {
    /* inputs (device) */
    const void* action;      /* [act_steps][2][n_envs] reals: converted_action (north, east) */
    const uint8_t* sac_update;/* [act_steps][n_envs]: insert the waypoint at that step (MSRL_env_ex.py:324-338) */
    int32_t act_steps;       /* 1: action held for the launch, sac_update applies to its first step; or K */
    int32_t init;            /* `init` argument of step(): 1/0 for the first step of the launch, -1 = per-env
                                "first step since reset" flag */
    int32_t auto_reset;      /* 1: reset() + init_step() inside the kernel right after a done step (reference
                                semantics: omega, integrators, fuel survive, quirk Q3); 2: same with a hard reset */
    int32_t rec_steps;       /* outputs hold rec_steps = K (every step) or 1 (last step of the launch) */
    /* outputs (device), may be NULL */
    void* obs;               /* [rec_steps][SIT_N_OBS][n_envs] reals */
    void* reward;            /* [rec_steps][n_envs] reals */
    uint8_t* done;           /* [rec_steps][n_envs] */
    int32_t* status;         /* [rec_steps][n_envs] bitmask SIT_ST_* */
    void* reward_sum;        /* [n_envs] reals: sum of the K step rewards of this launch */
    uint8_t* done_any;       /* [n_envs]: any step of this launch was done */
    unsigned long long* stats; /* [SIT_N_STATS] accumulated with warp-reduced atomics: counts of each status
                                bit (0..11), done steps (12), env-steps (13) */
    void* reward_parts;      /* [rec_steps][SIT_N_REWARD_PARTS][n_envs] reals, may be NULL: the per-step increments of the
                                reference's cumulative `reward_results` lists (MSRL_env_ex.py:133-141, appended in
                                reward_function :924-964): test_ship reward_e_ct, reward_near_col, total_non_terminal;
                                obs_ship reward_base, reward_e_ct, reward_near_col, total_non_terminal; shared
                                total_non_terminal */
    uint8_t* snap;           /* [rec_steps][n_envs], may be NULL; needs rec_steps == K.  1 where the reference would have called
                                ship_snap_shot() on both ships in that step (`ship_draw`, MSRL_env_ex.py:416-421: timer > 30 s ->
                                snapshot, timer = 0; then timer += dt).  The timer is the env state field "draw_t" (seconds, f64;
                                the constructor argument time_since_last_ship_drawing, never reset by reset()).  The outline
                                rows are sit_ship_outline of that step's obs pose (obs 0,1,2 / 6,7,8). */
} SitEnvIO;

/* MultiShipRLEnv.reset (MSRL_env_ex.py:145-186): same as sit_reset on an env handle */
/* MultiShipRLEnv.init_step (MSRL_env_ex.py:188-215) */
int sit_env_init_step(SitHandle h, void* state, const uint8_t* mask, void* stream);
/* K x MultiShipRLEnv.step (MSRL_env_ex.py:402-440): test_step :217-283, obs_step :285-400,
 * observation :424-435, reward_function :906-980 with the failure predicates :453-603 */
int sit_env_step_k(SitHandle h, void* state, int32_t K, const SitEnvIO* io, void* stream);

/* Host-buffer drop-in for one MultiShipRLEnv.step over all envs: action [2][n_envs] and
 * sac_update [n_envs] are HOST arrays, obs [10][n_envs], reward, done, status are HOST outputs
 * (reals in the handle's precision).  Copies, runs K steps (outputs of the last), synchronises. */
int sit_env_step_host(SitHandle h, void* state, int32_t K, const void* action_h, const uint8_t* sac_update_h,
                      int32_t init, void* obs_h, void* reward_h, uint8_t* done_h, int32_t* status_h, void* stream);

/* sit_env_step_host without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid it).  Same arguments, same results as sit_env_step_host; the device staging (action, sac_update, obs,
 * reward, done, status) lives in a fixed 2 MiB `static __device__` array inside the library instead of the handle's
 * lazily allocated scratch, so the call performs no cudaMalloc / cudaFree (it does cudaMemsetAsync, cudaMemcpyAsync,
 * one kernel launch and a stream synchronise).  Limits and rules:
 *   - handles with more than SIT_NOALLOC_MAX_ENVS envs are rejected with SIT_E_INVALID (use sit_env_step_host);
 *   - every GPU that loads the library holds the 2 MiB array until the process exits;
 *   - ALL calls on a device share the array (it keeps the previous call's contents; outputs are cleared at the
 *     start of each call): two calls that overlap on the GPU -- different handles or streams -- race.  The call
 *     synchronises its stream before returning, so calls issued one after another from one thread are safe. */
#define SIT_NOALLOC_MAX_ENVS 16384
int sit_env_step_host_noalloc(SitHandle h, void* state, int32_t K, const void* action_h, const uint8_t* sac_update_h,
                              int32_t init, void* obs_h, void* reward_h, uint8_t* done_h, int32_t* status_h, void* stream);

/* ---- device-resident rollout (the role of the empty ast_core/samplers/intermediate_waypoint_sampler.py; loop
 * contract of test_beds/main_ast.py:310-412).  The agent-side pieces the reference imports from the absent `ast_sac`
 * package (select_action, convert_action_to_simu_input) are DEFINED here -- parity unpinned, see DESIGN.md:
 *   decision event  = first step of an episode, or the obstacle ship is within the radius of acceptance of its
 *                     current intermediate waypoint (controllers.py:317-325) and fewer than sampling_frequency
 *                     decisions were taken this episode (main_ast.py:47);
 *   action          = a in [-1, 1]: UniformPolicy (policies/uniform_policy.py:21-22) or the tanh-Gaussian policy
 *                     (nn_models/mlp.py:95-148, distributions/normal.py:14-15,78-98, policies/gaussian_policy.py:116-117);
 *   conversion      = scoping angle a * pi/6 (action_space, MSRL_env_ex.py:79-83) around AB_alpha, one
 *                     AB_segment_length (MSRL_env_ex.py:122-127) ahead of the previous intermediate waypoint
 *                     (of the route start for the first), redrawn up to max_route_resampling times while it falls
 *                     inside an obstacle or off the map (main_ast.py:49-51);
 *   transition      = (state[10], action, reward, next_state[10], mask) of the decision step only
 *                     (main_ast.py:387,395-396), mask = 1 if episode_steps == num_steps_episode else !done.
 * An env that reaches a decision event stops stepping ("pending") until sit_rollout_policy has given it an action,
 * so the policy runs batched over the pending subset between step launches and nothing returns to the host. */
#define SIT_TR_FLOATS 24   /* replay row: s[0..9], a[10], r[11], s'[12..21], mask[22], global env id as int32 bits [23] */
enum { SIT_POLICY_UNIFORM = 0, SIT_POLICY_GAUSSIAN = 1, SIT_POLICY_MEAN = 2 };
/* OR into SitPolicy.mode: run the H x H hidden layer on the tensor cores (tcgen05.mma, bf16 operands -- layer-0 activations
 * and w1 rounded to nearest-even bf16 -- fp32 accumulate in TMEM); everything else stays fp32.  Without it the whole MLP
 * is fp32 on the CUDA cores. */
enum { SIT_POLICY_TC_BF16 = 0x100 };
/* TRANSITIONS counts replay rows OPENED by sit_rollout_policy (state, action written); CLOSED counts rows whose reward,
 * next_state and mask the following step has written.  Rows [0, CLOSED) (ring order) are complete: a learner reads CLOSED. */
enum { SIT_RC_TRANSITIONS = 0, SIT_RC_DECISIONS = 1, SIT_RC_RESAMPLES = 2, SIT_RC_EPISODES = 3, SIT_RC_CLOSED = 4, SIT_RC_N = 8 };

typedef struct SitPolicy {
    int32_t mode;              /* SIT_POLICY_* */
    int32_t hidden;            /* width H of both hidden layers (multiple of 32, <= 256); ignored for UNIFORM */
    const float* w0;           /* [10][H] row-major: layer0 = relu(b0 + obs . w0)   (device pointers, fp32) */
    const float* b0;           /* [H] */
    const float* w1;           /* [H][H]:  layer1 = relu(b1 + layer0 . w1) */
    const float* b1;           /* [H] */
    const float* w2;           /* [H][2]:  (mu, log_sigma) = b2 + layer1 . w2, log_sigma clipped to [-20, 2] */
    const float* b2;           /* [2] */
    const float* obs_scale;    /* [10] or NULL: the network sees obs[c] * obs_scale[c] (e.g. 1 / observation_space.high,
                                  MSRL_env_ex.py:67-78); the replay rows keep the raw observation */
} SitPolicy;

typedef struct SitRollout {
    uint64_t seed;             /* Philox key; draws are keyed by (global env id, decision serial, resample index) */
    int64_t env_id_offset;     /* global id of env 0 of this handle: results do not depend on the sharding */
    int32_t max_episode_steps; /* args.num_steps_episode (0 = none).  Reference semantics (main_ast.py:322,387): the row that
                                  closes at episode_steps == num_steps_episode gets mask = 1 and the episode CONTINUES until
                                  a terminal state (`while not done`); see truncate_episodes */
    int32_t max_route_resampling;
    float* replay;             /* [capacity][SIT_TR_FLOATS] device ring, capacity >= n_envs */
    int64_t capacity;
    unsigned long long* counters; /* [SIT_RC_N] device counters, accumulated */
    int32_t truncate_episodes; /* opt-in extension: != 0 also ENDS the episode (reset + init_step) at num_steps_episode */
    int32_t _pad;
} SitRollout;

/* give every pending env an action: policy forward on its current observation, sample, convert, validate, open
 * the replay row.  Enqueues only. */
int sit_rollout_policy(SitHandle h, void* state, const SitPolicy* pol, const SitRollout* ro, void* stream);
/* K x step for every env that is not pending (pending envs with a fresh action consume it first); closes replay
 * rows, auto-resets finished episodes (io->auto_reset: 1 reference reset, 2 hard) and marks new decision events.
 * io may carry the usual outputs (rec_steps must be 1) and stats; io->action / sac_update / init are ignored. */
int sit_rollout_step_k(SitHandle h, void* state, int32_t K, const SitRollout* ro, const SitEnvIO* io, void* stream);
/* `iterations` x (sit_rollout_policy, sit_rollout_step_k(K)): the inner loop of main_ast.py:320-412 between two learner
 * updates.  use_graph != 0 (and a non-default stream): replayed from a CUDA graph -- with one env group a graph of one
 * iteration, launched `iterations` times; with several groups (sit_rollout_group_count) ONE graph of the whole call, the
 * groups as independent chains that are joined at the end of the call only (captured on the second call with the same
 * arguments; the first runs as plain launches; a call of more than 64 iterations runs as 64-iteration graphs plus a
 * plain-launch remainder).  The graph is re-captured when any argument changes.  Enqueues only. */
int sit_rollout_collect(SitHandle h, void* state, int32_t iterations, int32_t K, const SitPolicy* pol, const SitRollout* ro,
                        const SitEnvIO* io, int32_t use_graph, void* stream);
/* mark every env pending at the start of a rollout (after sit_reset + sit_env_init_step) */
int sit_rollout_begin(SitHandle h, void* state, void* stream);
/* number of env groups of sit_rollout_collect (1 = a single chain on the caller's stream; more: every group runs its own
 * chain policy -> K steps -> policy -> ... on an internal stream forked from the caller's at the start of the call and
 * joined at its end, so one group's policy kernels run beside another group's step kernel).  Envs are independent and every
 * random draw is keyed by (global env id, decision serial, try), so no env's trajectory depends on it; replay rows are
 * reserved group after group (ascending env ranges) inside an iteration and iteration after iteration, so the ring is
 * filled row for row as with one group.  Chosen from n_envs (1 below 32 768 envs, 2, 3 from 49 152); the environment variable SIT_ROLLOUT_GROUPS overrides. */
int sit_rollout_group_count(SitHandle h, int32_t* groups);

/* ---- replay exchange after a rollout segment (the only multi-GPU exchange of the loop: main_ast.py:396 memory.push, batched).
 * A block is [1 + block_rows][SIT_TR_FLOATS] floats: row 0 is a header whose first 8 bytes hold the int64 count of valid rows,
 * rows 1 .. count are transitions.  block_rows is a HOST constant (e.g. policy calls per segment x n_envs), so every rank
 * ships the same number of bytes and no counts are exchanged; nothing here synchronises with the host.
 *   sit_replay_pack:   rows closed since *seen (device int64, SIT_RC_CLOSED at the previous pack; updated) -> block; rows
 *                      that do not fit are counted in *dropped (device int64, may be NULL) and the newest are kept;
 *   sit_replay_append: n_blocks blocks laid out back to back (the output of an all-gather of every rank's block) are
 *                      appended to the ring dst [dst_capacity][SIT_TR_FLOATS] in block order, row order; *dst_count
 *                      (device, rows ever appended) advances by the number of valid rows. */
int sit_replay_pack(const SitRollout* ro, int64_t* seen, int64_t* dropped, float* block, int64_t block_rows, void* stream);
int sit_replay_append(float* dst, int64_t dst_capacity, unsigned long long* dst_count, const float* blocks, int32_t n_blocks,
                      int64_t block_rows, void* stream);

/* Map predicates for n candidate waypoints (device arrays of reals): flags[i] bit 0 = inside an obstacle polygon
 * (is_route_inside_obstacles MSRL_env_ex.py:544-552 -> obstacle.py:131-136), bit 1 = outside the map AABB
 * (is_route_outside_horizon :517-542); dist[i] = obstacles_distance (obstacle.py:138-141).  flags / dist may be NULL. */
int sit_map_query(SitHandle h, int64_t n, const void* north, const void* east, uint8_t* flags, void* dist, void* stream);

/* ---- model variants and snapshot data ------------------------------------------------------------------
 * StaticObstacle (obstacle.py:8-90): n_circles rows (north, east, radius) of HOST doubles; for n points (device reals of
 * `precision`): flags[i] = 1 if inside or on any circle (if_ship_inside_obstacles / if_route_inside_obstacles :69-85),
 * dist[i] = obstacles_distance :55-67 (negative inside).  flags / dist may be NULL.  Needs no handle: circles are not part
 * of MultiShipRLEnv (it requires PolygonObstacle.map_boundaries). */
int sit_circle_query(int32_t precision, int32_t n_circles, const double* circles, int64_t n, const void* north, const void* east,
                     uint8_t* flags, void* dist, void* stream);
/* ship_snap_shot (ship_model.py:326-338, ShipDraw utils.py:56-101) for n poses (device reals): x_ned, y_ned [n][6], the
 * closed outline of the 80 m x 20 m hull rotated by psi and translated to (north, east).  Legacy default stream. */
int sit_ship_outline(int32_t precision, int64_t n, const void* north, const void* east, const void* psi, void* x_ned, void* y_ned);

/* ---- SAC learner (the step after the path: the empty ast_core/algorithms/sac.py; call contract
 * `critic_1_loss, critic_2_loss, policy_loss, ent_loss, alpha = agent.update_parameters(memory, batch_size, updates)`
 * main_ast.py:350-357; hyper-parameters main_ast.py:41-68).  The reference's `ast_sac` agent is absent, so the update
 * rule is DEFINED here as standard soft actor-critic with twin Q networks -- parity unpinned, checked against a plain
 * PyTorch autograd restatement in tests/:
 *   (s, a, r, s', mask) ~ uniform rows of the replay ring (SIT_TR_FLOATS layout above);
 *   a' = tanh(mu(s') + sigma(s') eps'), log pi' = log N(eps') - log sigma - log(1 - a'^2 + 1e-6);
 *   y = r + mask * gamma * (min(Q1t, Q2t)(s', a') - alpha log pi');     critic loss = mse(Q1(s,a), y) + mse(Q2(s,a), y);
 *   policy loss = mean(alpha log pi(a~|s) - min(Q1, Q2)(s, a~)) with the updated critics, reparameterised;
 *   alpha loss  = -mean(log_alpha * (log pi + target_entropy)), target_entropy = -1 (one action); alpha = exp(log_alpha)
 *                 takes effect from the next update (the first update uses SitSacConfig.alpha);
 *   Adam (beta 0.9 / 0.999, eps 1e-8) on policy, critics and log_alpha;  target <- (1 - tau) target + tau critic every
 *   target_update_interval updates.
 * Networks: policy 10 -> H -> H -> (mu, log sigma), Q(s, a) 11 -> H -> H -> 1, ReLU, Glorot-uniform weights, zero biases.
 * The policy block has the SitPolicy layout, so the rollout kernels read the weights the learner is updating.
 * All device memory is ONE caller-owned workspace (sit_sac_workspace_bytes, sit_sac_field); an update is 37 small
 * kernels captured once into a CUDA graph and replayed; every random draw is Philox keyed by (seed, update index). */
typedef struct SitSacConfig {
    int32_t hidden;                    /* H: multiple of 32 in [32, 256] (main_ast.py:67: 256) */
    int32_t batch_size;                /* main_ast.py:63: 64; any 1..65536 */
    float gamma, tau, lr, alpha;       /* main_ast.py:41-58: 0.99, 0.005, 3e-4, 0.2 */
    int32_t automatic_entropy_tuning;  /* main_ast.py:59 */
    int32_t target_update_interval;    /* main_ast.py:84 */
    uint64_t seed;                     /* main_ast.py:61 */
    float obs_scale[SIT_N_OBS];        /* per-channel input scale of all five networks (state part of s and s'); an entry of 0
                                          means 1.  The raw observation is metres / rpm / kW (up to 2e4): unscaled it saturates
                                          tanh(mu) and zeroes every policy gradient */
    int32_t tensor_cores;              /* 1: H x H products of the update on tcgen05 (bf16 operands, fp32 accumulate) when
                                          batch_size >= 1024 (below that the update stays in fp32, within fp32 rounding of a
                                          torch autograd restatement); 2: at every batch size (the reference's batch 64 runs
                                          as 22 launches instead of 37); 0: fp32 CUDA cores everywhere.  With hidden a multiple
                                          of 128 a network's forward and its input-gradient chain are one kernel each */
    int32_t _pad;
} SitSacConfig;
typedef struct SitSac_* SitSac;
enum { SIT_SAC_LOSS_Q1 = 0, SIT_SAC_LOSS_Q2 = 1, SIT_SAC_LOSS_POLICY = 2, SIT_SAC_LOSS_ALPHA = 3, SIT_SAC_ALPHA = 4,
       SIT_SAC_MEAN_LOGPI = 5, SIT_SAC_MEAN_Q = 6, SIT_SAC_N_LOSS = 8 };

int sit_sac_create(const SitSacConfig* cfg, int device, SitSac* out);
int sit_sac_destroy(SitSac h);
int sit_sac_workspace_bytes(SitSac h, size_t* nbytes);
/* named float32 / int32 region of the workspace: "policy", "q1", "q2", "q1_target", "q2_target", "log_alpha", "alpha",
 * "grad", "adam_m", "adam_v", "idx", "eps_next", "eps_cur", "losses", "updates", "obs_scale".  offset in bytes, count in
 * elements. */
int sit_sac_field(SitSac h, const char* name, int64_t* offset, int64_t* count);
/* Glorot-uniform weights (Philox, cfg.seed), zero biases, targets = critics, log_alpha = 0, Adam state and counter 0 */
int sit_sac_init(SitSac h, void* workspace, void* stream);
/* fill a SitPolicy (mode GAUSSIAN) with pointers into the workspace's policy block (and its obs_scale) */
int sit_sac_policy(SitSac h, void* workspace, SitPolicy* out);
/* n_updates x update_parameters on batches drawn from replay[0 .. min(*n_rows, capacity)) (n_rows: device counter, e.g.
 * SitRollout.counters + SIT_RC_TRANSITIONS; NULL = all `capacity` rows are valid).  losses (device, may be NULL)
 * receives [SIT_SAC_N_LOSS] floats of the LAST update.  Enqueues only; fails if fewer rows than one batch is NOT checked
 * on the host (the counter lives on the device): an empty replay draws row 0. */
int sit_sac_update(SitSac h, void* workspace, const float* replay, int64_t capacity, const unsigned long long* n_rows,
                   int32_t n_updates, float* losses, int32_t use_graph, void* stream);

/* ---- misc ------------------------------------------------------------------------------------ */
const char* sit_last_error(void);
const char* sit_status_string(int32_t bit_index);  /* exact reference strings, MSRL_env_ex.py:760-899 */
int sit_abi_version(void);
/* CUDA-core FMA peak of this device measured with a register-resident FMA chain (roofline
 * denominator for the compute-bound step kernels).  precision SIT_F32 / SIT_F64, or SIT_PEAK_F32X2 for the same
 * chain issued as packed f32x2 FMAs (SASS FFMA2, the instruction the fp32 fleet kernel uses); returns TFLOP/s. */
enum { SIT_PEAK_F32X2 = 2 };
int sit_measure_fma_peak(int device, int32_t precision, double* tflops, double* ms);

#ifdef __cplusplus
}
#endif
#endif /* SHIPSIM_H */
