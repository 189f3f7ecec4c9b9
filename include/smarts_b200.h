/*
 * smarts_b200 C-ABI: the batched per-tick multi-agent step (controllers + vehicle dynamics +
 * observation sensors) for E independent environments on one B200.
 *
 * The reference (fahmyadan/SMARTS, 100% Python) has no FFI boundary; the entry points below are
 * what a binding of its hot path would call.  Each one names the reference interface it replaces
 * (paths relative to /root/reference).  All pointers are plain host or device addresses; there
 * are no torch / C++ types in any signature.  INTEGRATION.md shows the ctypes stub.
 *
 * Conventions
 *   - E envs, A agent slots per env, M social slots per env, V = A + M vehicle slots per env.
 *     Vehicle arrays are indexed [e*V + slot]; slots 0..A-1 are agents.  Agent arrays are
 *     indexed [e*A + a].
 *   - Frames/angles follow smarts/core/coordinates.py:169-185: heading in radians, 0 = +y,
 *     counter-clockwise, wrapped to (-pi, pi].
 *   - Every function returns 0 on success, a negative SB_E* code otherwise; sb_last_error()
 *     gives the message.  CUDA errors are sticky per handle.  No exceptions cross the ABI.
 *   - Calls on one handle are not re-entrant.  sb_step/sb_observe only enqueue work on the
 *     given stream (0 = legacy default stream) and return; the caller synchronises.
 */
#ifndef SMARTS_B200_H
#define SMARTS_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SB_ABI_VERSION 7

/* error codes */
#define SB_OK 0
#define SB_EINVAL (-1)   /* bad argument / config */
#define SB_ECUDA (-2)    /* CUDA runtime error (sticky) */
#define SB_ESTATE (-3)   /* call order: map / rays not uploaded, handle destroyed */
#define SB_ENOMEM (-4)

/* ActionSpaceType values, smarts/core/controllers/__init__.py:42-54 */
#define SB_ACTION_CONTINUOUS 0
#define SB_ACTION_LANE 1
#define SB_ACTION_ACTUATOR_DYNAMIC 2
#define SB_ACTION_LANE_WITH_CONTINUOUS_SPEED 3
#define SB_ACTION_TRAJECTORY 5

/* Lane action command codes (strings in the reference, controllers/__init__.py:136-143) */
#define SB_LANE_KEEP 0
#define SB_LANE_SLOW_DOWN 1
#define SB_LANE_CHANGE_LEFT 2
#define SB_LANE_CHANGE_RIGHT 3

/* internal caps (status bit SB_STATUS_PATH_OVERFLOW is raised when exceeded) */
#define SB_MAX_PATHS 16
#define SB_MAX_LOOKAHEAD 64
#define SB_MAX_TRAJ 32
#define SB_MAX_POLY_VERTS 8 /* drivable-area polygons: convex, counter-clockwise, 3..8 vertices */

#define SB_STATUS_PATH_OVERFLOW 1u
#define SB_STATUS_NO_WAYPOINTS 2u
#define SB_STATUS_LIDAR_OVERFLOW 4u  /* more candidate boxes / edges near one agent than the kernel stages */
#define SB_STATUS_TTC_TRUNCATED 8u   /* lane_ttc saw a capped neighbour / path / waypoint block (or a lane index
                                        beyond the path list, an IndexError in the reference) */
#define SB_STATUS_RW_TRUNCATED 16u   /* RoadWaypoints: more lanes / paths / nearby roads than the caps below, or the
                                        walk back through incoming lanes went deeper than SB_RW_MAX_DEPTH */
#define SB_RW_MAX_DEPTH 8            /* nesting of RoadWaypointsSensor.paths_for_lane (sensors.py:1020-1027) */
#define SB_RW_MAX_ROADS 64           /* the road + its parallel roads + oncoming roads (sensors.py:1008-1010) */

/* Events bits (smarts/core/events.py:23-34) in SbObsPtrs.ev_events and SbConfig.done_mask
 * (DoneCriteria, smarts/core/agent_interface.py:185-206). */
#define SB_EV_COLLISION 1u
#define SB_EV_OFF_ROAD 2u
#define SB_EV_OFF_ROUTE 4u
#define SB_EV_ON_SHOULDER 8u
#define SB_EV_WRONG_WAY 16u
#define SB_EV_NOT_MOVING 32u
#define SB_EV_REACHED_GOAL 64u          /* always ends the episode (sensors.py:464-474) */
#define SB_EV_REACHED_MAX_STEPS 128u    /* always ends the episode */
#define SB_EV_AGENTS_ALIVE_DONE 256u    /* multi-agent alive criteria: evaluated by the host layer */
#define SB_MAX_VIAS 16                  /* via points per agent mission */
#define SB_DRIVEN_PATH_MAX 500          /* DrivenPathSensor deque length, sensors.py:838-840 */

typedef struct SbHandle_ *SbHandle;

/* Compiled from AgentInterface (smarts/core/agent_interface.py:209-277) + HiWayEnv kwargs
 * (smarts/env/hiway_env.py:43-65).  One interface for all agents of a batch. */
typedef struct SbConfig {
  int32_t abi_version;     /* SB_ABI_VERSION */
  int32_t device;          /* CUDA device ordinal */
  int32_t n_envs;          /* E */
  int32_t n_agents;        /* A */
  int32_t n_social;        /* M */
  double dt;               /* fixed_timestep_sec, hiway_env.py:112-113 */
  int32_t substeps;        /* dynamics substeps per tick; 0 -> int(dt*240) (smarts.py:607-615) */
  int32_t action_space;    /* SB_ACTION_* */
  int32_t traj_len;        /* N points per Trajectory action (<= SB_MAX_TRAJ) */
  int32_t route_cap;       /* max roads per agent route */
  /* NeighborhoodVehicles (agent_interface.py:88-93, sensors.py:950-969) */
  int32_t nbr_enable;
  int32_t nbr_cap;         /* neighbours written per agent (StdObs: 10, format_obs.py:41) */
  double nbr_radius;       /* < 0  ==  radius=None (all vehicles) */
  /* Waypoints (agent_interface.py:74-79, sensors.py:972-988) */
  int32_t wp_enable;
  int32_t wp_lookahead;    /* sensor lookahead (default 32) */
  int32_t wp_out_paths;    /* paths written per agent (StdObs: 4) */
  int32_t wp_out_len;      /* waypoints written per path (StdObs: 20; raw: lookahead+1) */
  /* Lidar (agent_interface.py:64-71, lidar.py, sensors.py:797-827) */
  int32_t lidar_enable;
  int32_t lidar_rays;      /* R, must equal the uploaded table */
  int32_t lidar_miss_inf;  /* 1: miss -> (inf,inf,inf) as lidar.py:131; 0: zeros as format_obs.py:437-443 */
  double lidar_wall_height;/* road-edge wall height in metres; 0 disables edges (reference has none) */
  /* OGM / DrivableAreaGridMap (agent_interface.py:36-62, sensors.py:675-758) */
  int32_t ogm_enable, ogm_width, ogm_height;
  double ogm_resolution;
  int32_t dagm_enable, dagm_width, dagm_height;
  double dagm_resolution;
  /* Events / done criteria / trip-meter reward (sensors.py:443-594, :830-947; agent_manager.py:233-234) */
  int32_t events_enable;
  uint32_t done_mask;          /* SB_EV_* bits that end an agent's episode (DoneCriteria) */
  int32_t max_episode_steps;   /* AgentInterface.max_episode_steps; <= 0: None */
  double not_moving_time;      /* EventConfiguration, agent_interface.py:174-182 (60 s) */
  double not_moving_distance;  /* (1 m) */
  /* lane_ttc (smarts/env/custom_observations.py:148-290), the `ttc` member of StdObs (format_obs.py:546-556);
   * needs nbr_enable and wp_enable */
  int32_t ttc_enable;
  /* AccelerometerSensor (sensors.py:1046-1084): linear / angular acceleration and jerk of the ego */
  int32_t accel_enable;
  /* ViaSensor (sensors.py:1090-1150): via points per agent mission (Mission.via, plan.py:180-188);
   * via_cap = 0 disables.  Defaults of vehicle.py:552-558: range 40 m, speed accuracy 1.5 m/s */
  int32_t via_cap;              /* <= SB_MAX_VIAS */
  double via_acquisition_range;
  double via_speed_accuracy;
  /* ego_centric_observation_adapter (smarts/core/utils/adapters/ego_centric_adapters.py:62-178): 1 = sb_observe
   * ends by rewriting the positional buffers in the ego frame (origin at the ego, +x along its heading):
   * nbr_pos / nbr_heading, wp_pos / wp_heading, lidar_points.  Only filled entries are touched; a lidar miss
   * stays (0,0,0) with lidar_miss_inf = 0 and becomes (NaN,NaN,NaN) with lidar_miss_inf = 1 (inf - x through
   * the rotation, as in the reference). */
  int32_t ego_frame;
  /* RoadWaypoints (agent_interface.py:81-91, sensors.py:991-1043): waypoint paths along every lane of the road
   * the ego is on, its parallel roads and the oncoming roads next to it, from `horizon` metres behind the ego over
   * 2 * horizon lane points.  rw_lanes_cap lanes and rw_paths_cap paths (over all lanes) are written per agent;
   * 2 * rw_horizon <= SB_MAX_LOOKAHEAD. */
  int32_t rw_enable;
  int32_t rw_horizon;      /* RoadWaypoints.horizon (default 20) */
  int32_t rw_lanes_cap;
  int32_t rw_paths_cap;
  /* AgentInterface.vehicle_type (agent_interface.py:246-250; vehicle.py:393-406): SB_VEHICLE_SEDAN (0) or
   * SB_VEHICLE_BUS (1) of include/sb_vehicle.h -- chassis box, masses, torques and controller gains of the agents */
  int32_t vehicle_type;
} SbConfig;

/* Per-scenario constant road map (host pointers, copied once).  Built by
 * smarts_b200/road_map.py from a SUMO net.xml; restates SumoRoadNetwork + LanePoints
 * (smarts/core/sumo_road_network.py, smarts/core/lanepoints.py). */
typedef struct SbMap {
  int32_t n_roads, n_lanes, n_lanepoints, n_shape_pts, n_polys, n_poly_verts, n_edges;
  int32_t n_lane_out, n_lp_next;
  int32_t n_lane_in, n_road_par;
  double lanepoint_spacing;
  /* roads */
  const int32_t *road_lane_start, *road_lane_count;   /* [n_roads] lanes of a road are contiguous */
  const uint8_t *road_is_junction;                     /* [n_roads] Edge.isSpecial() */
  const int32_t *road_lp_start, *road_lp_count;        /* [n_roads] lane-point range */
  const int32_t *road_first_out;                       /* [n_roads] first outgoing road or -1 (Road.outgoing_roads[0]) */
  /* lanes */
  const int32_t *lane_road, *lane_idx_in_road;         /* [n_lanes] */
  const double *lane_width, *lane_speed, *lane_length; /* [n_lanes] */
  const int32_t *lane_out_start;                       /* [n_lanes+1] CSR into lane_out */
  const int32_t *lane_out;                             /* [n_lane_out] outgoing lane (via or to) */
  const int32_t *lane_lp_start, *lane_lp_count;        /* [n_lanes] */
  const int32_t *lane_shape_start;                     /* [n_lanes+1] CSR into shape_x/y */
  /* lane points */
  const double *lp_x, *lp_y, *lp_heading;              /* [n_lanepoints] */
  const uint8_t *lp_inferred;                          /* [n_lanepoints] */
  const int32_t *lp_lane;                              /* [n_lanepoints] */
  const int32_t *lp_next_start;                        /* [n_lanepoints+1] CSR into lp_next */
  const int32_t *lp_next;                              /* [n_lp_next] */
  /* lane centre polylines */
  const double *shape_x, *shape_y;                     /* [n_shape_pts] */
  /* drivable-area convex CCW polygons (3..SB_MAX_POLY_VERTS vertices each) and road boundary segments */
  const int32_t *poly_start;                           /* [n_polys+1] CSR into poly_x/y */
  const double *poly_x, *poly_y;                       /* [n_poly_verts] */
  const double *road_edges;                            /* [n_edges,4] ax,ay,bx,by */
  /* RoadWaypoints tables */
  const int32_t *lane_in_start;                        /* [n_lanes+1] CSR into lane_in */
  const int32_t *lane_in;                              /* [n_lane_in] Lane.incoming_lanes, sumo_road_network.py:342-348 */
  const int32_t *road_par_start;                       /* [n_roads+1] CSR into road_par */
  const int32_t *road_par;                             /* [n_road_par] Road.parallel_roads, sumo_road_network.py:607-618 */
} SbMap;

/* Device pointers owned by the library (valid until sb_destroy). */
typedef struct SbStatePtrs {
  /* vehicles [E*V]  (VehicleState, smarts/core/vehicle.py:53-80) */
  double *pos_x, *pos_y, *pos_z, *heading, *speed, *yaw_rate;
  double *length, *width, *height;
  double *cos_h, *sin_h;     /* of heading, by the deterministic sincos of include/sb_detmath.h */
  uint8_t *alive;
  /* agents [E*A]  (AckermannChassis read-backs, smarts/core/chassis.py:493-566) */
  double *steering;          /* chassis.steering (rad, clockwise positive) */
  double *v_long, *v_lat;    /* longitudinal_lateral_speed */
  double *ctrl;              /* [E*A, SB_CTRL_STATE] controller state */
  double *last_control;      /* [E*A, 3] throttle, brake, steering applied this tick */
  int32_t *route;            /* [E*A, route_cap] road indices, -1 padded */
  int32_t *route_len;        /* [E*A] */
  uint32_t *status;          /* [E*A] SB_STATUS_* bits, sticky until sb_reset */
} SbStatePtrs;

#define SB_CTRL_STATE 12

/* Observation buffers (device, owned by the library).  Layouts are StdObs
 * (smarts/env/wrappers/format_obs.py:45-197) so a Python view is zero-copy. */
typedef struct SbObsPtrs {
  /* ego (EgoVehicleObservation, sensors.py:65-101) */
  int32_t *ego_lane;           /* [E*A] nearest lane index into the map or -1 (sensors.py:277-285) */
  /* neighbours */
  int32_t *nbr_count;          /* [E*A] total vehicles within radius (uncapped) */
  int32_t *nbr_id;             /* [E*A, cap] vehicle slot within env, -1 padded */
  double *nbr_pos;             /* [E*A, cap, 3] */
  float *nbr_box;              /* [E*A, cap, 3] length,width,height */
  float *nbr_heading;          /* [E*A, cap] */
  float *nbr_speed;            /* [E*A, cap] */
  int32_t *nbr_lane;           /* [E*A, cap] map lane index or -1 (sensors.py:244-254) */
  int8_t *nbr_lane_index;      /* [E*A, cap] */
  /* waypoints */
  int32_t *wp_n_paths;         /* [E*A] total paths found (uncapped) */
  int32_t *wp_len;             /* [E*A, P] */
  double *wp_pos;              /* [E*A, P, L, 3] */
  float *wp_heading;           /* [E*A, P, L] */
  float *wp_lane_width;        /* [E*A, P, L] */
  float *wp_speed_limit;       /* [E*A, P, L] */
  int32_t *wp_lane;            /* [E*A, P, L] map lane index (lane_id) */
  int8_t *wp_lane_index;       /* [E*A, P, L] */
  /* lidar */
  double *lidar_points;        /* [E*A, R, 3] */
  int8_t *lidar_hits;          /* [E*A, R] */
  /* grids */
  uint8_t *ogm;                /* [E*A, H, W] */
  uint8_t *dagm;               /* [E*A, H, W] */
  /* events / reward (events_enable) */
  uint32_t *ev_events;         /* [E*A] SB_EV_* bits (Events, events.py:23-34) */
  uint8_t *ev_done;            /* [E*A] Sensors._is_done_with_events, sensors.py:464-474 */
  int32_t *ev_n_collisions;    /* [E*A] vehicles in contact (smarts.py:1270-1291) */
  int32_t *ev_collidee;        /* [E*A] lowest colliding vehicle slot or -1 */
  double *ev_distance;         /* [E*A] TripMeterSensor(), sensors.py:880-947 (Observation.distance_travelled) */
  double *ev_reward;           /* [E*A] TripMeterSensor(increment=True), agent_manager.py:233-234 */
  int32_t *ev_steps;           /* [E*A] SensorState.steps_completed, sensors.py:609-637 */
  /* lane_ttc (ttc_enable): StdObs `ttc` (format_obs.py:546-556) */
  float *ttc_angle_error;      /* [E*A] */
  float *ttc_distance_from_center; /* [E*A] */
  float *ttc_dtc;              /* [E*A, 3] right / current / left lane */
  float *ttc_ttc;              /* [E*A, 3] */
  uint8_t *ttc_valid;          /* [E*A] 0: no neighbours or no waypoint paths (the reference returns None) */
  /* accelerometer (accel_enable): EgoVehicleObservation.linear_acceleration ... angular_jerk, sensors.py:94-101 */
  double *acc_linear;          /* [E*A, 3] */
  double *acc_angular;         /* [E*A, 3] */
  double *acc_linear_jerk;     /* [E*A, 3] */
  double *acc_angular_jerk;    /* [E*A, 3] */
  /* vias (via_cap > 0): Observation.via_data (Vias, sensors.py:168-175) as indices into the agent's via list */
  int32_t *via_near_count;     /* [E*A] */
  int32_t *via_near_idx;       /* [E*A, via_cap] nearest first (stable), -1 padded */
  int32_t *via_hit_count;      /* [E*A] points collected by this observation */
  int32_t *via_hit_idx;        /* [E*A, via_cap] in mission order, -1 padded */
  /* road waypoints (rw_enable): Observation.road_waypoints (RoadWaypoints.lanes, sensors.py:104-107) as a lane
   * list in the reference's dict insertion order plus one flat path list; path p belongs to lane slot
   * rw_path_slot[p]; the paths of one lane are consecutive and in the reference's list order (the lane groups
   * themselves follow the order in which the sensor last computed each lane, which is the slot order unless a
   * road is listed twice).  Lc = rw_lanes_cap, Pc = rw_paths_cap, Hz = 2 * rw_horizon + 1. */
  int32_t *rw_n_lanes;         /* [E*A] lanes found (uncapped) */
  int32_t *rw_lane;            /* [E*A, Lc] map lane index (the dict key), -1 padded */
  int32_t *rw_n_paths;         /* [E*A] paths found (uncapped) */
  int32_t *rw_path_slot;       /* [E*A, Pc] index into rw_lane, -1 padded */
  int32_t *rw_len;             /* [E*A, Pc] waypoints of the path */
  double *rw_pos;              /* [E*A, Pc, Hz, 2] */
  double *rw_heading;          /* [E*A, Pc, Hz] */
  float *rw_lane_width;        /* [E*A, Pc, Hz] */
  float *rw_speed_limit;       /* [E*A, Pc, Hz] */
  int32_t *rw_wp_lane;         /* [E*A, Pc, Hz] Waypoint.lane_id as a map lane index */
  int8_t *rw_lane_index;       /* [E*A, Pc, Hz] */
} SbObsPtrs;

/* Initial / reset state for a set of envs (host pointers).  Replaces SMARTS.reset ->
 * setup -> vehicle creation (smarts/core/smarts.py:365-460, vehicle.py:344-457). */
typedef struct SbResetState {
  const double *agent_x, *agent_y, *agent_heading, *agent_speed; /* [n, A] */
  const int32_t *route;      /* [n, A, route_cap] or NULL (no route) */
  const int32_t *route_len;  /* [n, A] or NULL */
  /* PositionalGoal (plan.py:95-124): goal_radius < 0 or NULL pointers = EndlessGoal */
  const double *goal_x, *goal_y, *goal_radius; /* [n, A] or NULL */
  /* Mission.via (plan.py:180-188), used when via_cap > 0; NULL via_count = no vias */
  const int32_t *via_count;  /* [n, A] 0..via_cap */
  const int32_t *via_lane;   /* [n, A, via_cap] map lane index (Via.lane_id) */
  const double *via_x, *via_y, *via_hit_distance, *via_required_speed; /* [n, A, via_cap] */
  /* 0: the envs' social slots are emptied (a fresh SMARTS.setup, smarts.py:440-460);  1: they are left as the last
   * sb_set_social wrote them -- the auto-reset of a batched env whose injected traffic keeps running */
  int32_t keep_social;
} SbResetState;

/* HiWayEnv.__init__ -> SMARTS.__init__ (smarts/env/hiway_env.py:166-173, smarts/core/smarts.py:96-185) */
int sb_create(const SbConfig *cfg, SbHandle *out);
/* HiWayEnv.close / SMARTS.destroy (hiway_env.py:287-291, smarts.py:688-703) */
void sb_destroy(SbHandle h);
/* SMARTS.setup -> road map load (smarts.py:440-460; SumoRoadNetwork.from_spec, sumo_road_network.py:145-176).
 * Calling it again replaces the map (the scenario iterator of HiWayEnv.reset, hiway_env.py:265-285, hands
 * SMARTS.reset the next scenario and SMARTS.setup reloads the road map when it differs, smarts.py:411-460): the
 * device is synchronised, the old map's arrays are freed, and every later call except a full sb_reset
 * (env_ids = NULL) answers SB_ESTATE until that reset has placed the agents on the new map. */
int sb_upload_map(SbHandle h, const SbMap *map);
/* Lidar._compute_rays base table (smarts/core/lidar.py:89-106); host [R,3] direction*max_distance */
int sb_upload_lidar_rays(SbHandle h, const double *base_rays, int32_t n_rays);
/* SMARTS.reset for a subset of envs (smarts.py:365-438); env_ids host [n] or NULL = all envs */
int sb_reset(SbHandle h, const int32_t *env_ids, int32_t n, const SbResetState *init);
/* Provider state injection (ExternalProvider.state_update, smarts/core/external_provider.py:44-51 ->
 * BoxChassis.control, chassis.py:211-217).  dev_states: device [E, M, 8] doubles
 * (x, y, heading, speed, length, width, height, alive) for this tick. */
int sb_set_social(SbHandle h, const double *dev_states, void *stream);
/* SMARTS._step_providers: Controllers.perform_action + chassis step (smarts.py:1048-1110,
 * controllers/__init__.py:61-152).  dev_actions layout by action space:
 *   LANE: int32 [E*A] command;  LANE_WITH_CONTINUOUS_SPEED: double [E*A,2] (speed, lane_change);
 *   CONTINUOUS / ACTUATOR_DYNAMIC: double [E*A,3];  TRAJECTORY: double [E*A,4,N] (x,y,heading,speed). */
int sb_step(SbHandle h, const void *dev_actions, void *stream);
/* AgentManager.observe -> Sensors.observe for every agent (agent_manager.py:161-240, sensors.py:238-396) */
int sb_observe(SbHandle h, void *stream);
/* The same for the agents of the listed envs only (host env_ids [n]; NULL = every env): the reset observation of
 * an auto-reset.  Every env of the reference is its own SMARTS instance (ParallelEnv, parallel_env.py:96-127), so
 * observing the envs that were just reset must not touch the sensor state (driven path, trip meter,
 * accelerometer history, vias, step counters) or the observation buffers of the others.  Whether an observation
 * follows a tick (step counter + 1, dones may be raised) or is the `sim.resetting` kind is kept per agent: sb_step
 * marks every agent, the events pass of sb_observe / sb_observe_envs uses the mark up for the agents it covers, and
 * sb_reset clears it for the envs it resets -- so any split of the envs over several sb_observe_envs calls after a
 * sb_step gives what one sb_observe gives. */
int sb_observe_envs(SbHandle h, const int32_t *env_ids, int32_t n, void *stream);
/* Convenience: the reference-facing call with HOST buffers (hiway_env.py:216-263 step()):
 * copies actions and social states in, runs sb_set_social + sb_step + sb_observe, synchronises. */
int sb_step_host(SbHandle h, const void *host_actions, int64_t action_bytes,
                 const double *host_social);
/* The other half of the host-buffer call: copy every enabled observation buffer to the HOST
 * pointers in *host_dst (NULL members are skipped) and synchronise.  Layouts as SbObsPtrs.
 * The grids (0 / 255 pixels) cross PCIe as 1 bit per pixel and are widened to uint8 by a pool of
 * host threads while later chunks are still in flight; every other buffer is a plain copy. */
int sb_copy_obs_to_host(SbHandle h, const SbObsPtrs *host_dst);
/* Opt-in for callers that reuse their host grid buffers: on = 1 promises that the ogm / dagm pointers given to
 * sb_copy_obs_to_host are the ones of the previous call and that nothing else wrote to them in between.  The
 * library then rewrites only the 64-pixel words that changed since that call (an ego-centred road image changes
 * along its edges only), which takes the host-memory write volume -- the bound of the call -- down by an order of
 * magnitude.  The buffers hold the complete uint8 images after every call either way.  A new pointer, a new size
 * or on = 0 -> 1 makes the next call a full write.  (hiway_env.py:216-263: the reference allocates fresh arrays
 * per step; a vectorised env that keeps its pinned staging buffers is the use case.) */
int sb_set_incremental_host_grids(SbHandle h, int on);
/* sb_copy_obs_to_host with the grids delivered as the 1-bit image instead of uint8: host_dst->ogm / ->dagm point to
 * W*H/8 bytes per agent, row-major, bit (i & 7) of byte (i >> 3) = pixel i of the [H, W] image (set = 255), i.e.
 * numpy.unpackbits(buf.reshape(n, H, W // 8), axis=-1, bitorder="little") * 255 is the uint8 image
 * sb_copy_obs_to_host delivers (widths are multiples of 32).  The uint8 widening is 8x the host-memory write
 * volume of everything else in the call and is its bound (profiles/README.md); a consumer that keeps grids
 * packed (a frame stack, a replay buffer, a policy whose first layer unpacks on its own device) should not
 * pay for it.  Every other member as in sb_copy_obs_to_host.  (Replaces the np.array(...) conversions of
 * smarts/core/sensors.py:706-716, :748-758 on the delivery side.) */
int sb_copy_obs_to_host_packed(SbHandle h, const SbObsPtrs *host_dst);
/* The grid half of sb_copy_obs_to_host_packed for callers that cannot allocate.  (No counterpart in the
 * reference: it is Python and never meets the problem.)
 *
 * sb_copy_obs_to_host_packed sizes its device bit image with cudaMalloc and its pinned staging with
 * cudaHostAlloc on first use, and again whenever the grids grow.  That is no use to
 *   - a trainer whose memory pool was handed all of HBM at start-up (cudaMalloc then answers out-of-memory),
 *   - a control loop that must not stall: cudaMalloc / cudaFree synchronise the whole device,
 *   - a thread that runs while another thread of the process has a stream capture open in the global capture mode,
 *     where cudaMalloc is an error,
 *   - an embedding that audits allocations and forbids them after initialisation.
 * This call performs no cudaMalloc, cudaFree, cudaHostAlloc or any other allocating CUDA / library routine, on its
 * first call or later.  Its scratch is a 256-byte-aligned `static __device__` array of SB_NOALLOC_BITS_MAX_BYTES
 * (16 MiB) that is part of the library image: every GPU that loads libsmartsb200.so holds those 16 MiB until the
 * process exits, whether or not this entry point is ever called.
 *
 * Inputs: the handle's ogm / dagm device images as the last sb_observe left them.  Outputs: host_ogm_bits /
 * host_dagm_bits (HOST pointers, pinned or pageable; NULL skips that grid), E*A*H*W/8 bytes each, in exactly the bit
 * layout sb_copy_obs_to_host_packed documents.  The call waits for the stream of the last sb_set_social / sb_step /
 * sb_observe, works on the legacy default stream and returns after the host buffers are complete.
 *
 * Limit: the bit images of the requested grids together must fit the array: E*A*(ogm W*H + dagm W*H)/8 <=
 * SB_NOALLOC_BITS_MAX_BYTES (e.g. 1024 agents with two 256 x 256 grids, or 16 384 agents with two 64 x 64 grids).
 * Anything larger returns SB_EINVAL with the two numbers in sb_last_error; nothing is launched.  W*H must be a
 * multiple of 32 as for the packed call.
 *
 * First call on a device: the library asks the runtime once for the device address of the array
 * (cudaGetSymbolAddress with that device current; the host-side name of a __device__ array is not a device
 * pointer).  Make that first call before opening any stream capture.
 *
 * Sharing: every handle on the same GPU uses the same array, and it keeps what the last call wrote.  Nothing has to
 * start at zero -- the kernel overwrites every word that is copied out afterwards -- so nothing is cleared.  Two
 * calls whose GPU work overlaps (two host threads, two handles on one GPU) race on the array and deliver garbage;
 * serialise them.  sb_set_timing does not see this call's kernel (timing events are created on demand, which
 * allocates); sb_launch_count does. */
#define SB_NOALLOC_BITS_MAX_BYTES (16u << 20)
int sb_copy_grids_to_host_packed_noalloc(SbHandle h, uint8_t *host_ogm_bits, uint8_t *host_dagm_bits);
int sb_get_obs(SbHandle h, SbObsPtrs *out);
int sb_get_state(SbHandle h, SbStatePtrs *out);
/* number of kernel launches issued by this handle so far (bench gpu_launches) */
int64_t sb_launch_count(SbHandle h);
/* Per-kernel device timing (the tracing hook; the reference wraps SMARTS.step in a `timeit`
 * logger, smarts/core/utils/logging.py:29-43).  With timing on, every kernel launch is bracketed
 * by CUDA events on its stream; sb_collect_timing synchronises and returns, per kernel kind,
 * the summed milliseconds and the number of launches since the last collect. */
#define SB_K_SOCIAL 0
#define SB_K_CONTROL 1
#define SB_K_OBSERVE 2
#define SB_K_LIDAR 3
#define SB_K_RASTER 4
#define SB_K_EVENTS 5
#define SB_K_PACK 6    /* k_pack_bits: grid transport of sb_copy_obs_to_host */
#define SB_K_TTC 7
#define SB_K_EXTRAS 8   /* accelerometer + vias */
#define SB_K_EGO_FRAME 9
#define SB_K_ROAD_WP 10  /* k_road_waypoints */
#define SB_K_COUNT 11
int sb_set_timing(SbHandle h, int on);
int sb_collect_timing(SbHandle h, double *ms_sum /*[SB_K_COUNT]*/, int64_t *count /*[SB_K_COUNT]*/);
const char *sb_last_error(SbHandle h);
int sb_abi_version(void);

#ifdef __cplusplus
}
#endif
#endif /* SMARTS_B200_H */
