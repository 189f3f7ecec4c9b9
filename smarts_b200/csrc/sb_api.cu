// sb_api.cu -- the C-ABI of include/smarts_b200.h over the kernels of sb_kernels.cuh.
// Compiled for sm_100a only (-gencode arch=compute_100a,code=sm_100a -fmad=false).  There is no
// CPU path in this library: every entry point that computes launches CUDA kernels.
#include <cstdio>
#include <cstdlib>
#include <algorithm>
#include <cmath>
#include <cstring>
#include <string>
#include <vector>

#include <chrono>
#include <thread>

#include "sb_kernels.cuh"

// sb_hostxfer.cpp (g++): host threads that widen bit-packed grids back to uint8
extern "C" {
void* sbx_pool_create(int threads, int first_cpu);
void sbx_pool_destroy(void* p);
void sbx_pool_submit(void* p, const uint8_t* bits, uint8_t* out, size_t n_bytes_in, size_t piece);
void sbx_pool_submit_diff(void* p, const uint8_t* bits, uint8_t* prev, uint8_t* out, size_t n_bytes_in, size_t piece);
void sbx_pool_wait(void* p);
int sbx_pool_threads(void* p);
}

struct SbHandle_ {
  SbConfig cfg;
  DCfg dc;
  DMap dmap;
  DState st;
  DObs ob;
  DEvents ev{};
  DEvOut evo{};
  DTtcOut ttc{};
  DExtras ex{};
  DRoadWp rw{};
  int32_t* d_veh_lane = nullptr;  // k_vehicle_lanes -> k_observe (only when V <= A * nbr_cap)
  int32_t *h_env_list = nullptr, *d_env_list = nullptr;  // sb_observe_envs: pinned staging + device copy of the env list
  // Ordering against the caller's streams is kept with events the handle owns, never with the caller's stream
  // handles (a stream may be destroyed, or be under capture, by the time the next call looks at it).
  cudaEvent_t work_ev = nullptr;      // recorded behind the last sb_set_social / sb_step / sb_observe on a non-legacy stream
  bool work_pending = false;          // ... and not yet waited for by a host-buffer call
  cudaEvent_t env_list_ev = nullptr;  // recorded behind the H2D copy of the sb_observe_envs list
  bool env_list_pending = false;
  bool have_map = false, have_rays = false;
  bool stale_agents = false;          // a new map was uploaded: routes / lanes of the old one are meaningless until a full reset
  std::vector<void*> map_allocs;      // device arrays of the current map (freed when another one replaces it)
  LTable ray_table{};   // factored ray table (include/sb_lidar_table.h), sorted by azimuth group
  float ray_reach = 0.f;
  std::vector<void*> allocs;
  CtrlPick* d_picks = nullptr;  // k_control_paths -> k_control_law
  int lidar_tile_min_agents = 2;      // agents per env from which k_lidar_tiles is used (SB_LIDAR_TILE_MIN_A: A/B runs)
  int lidar_list_cap = SB_LIDAR_LIST;  // intervals a k_lidar_tiles block holds (SB_LIDAR_LIST_CAP lowers it: tests of the fallback path)
  double* d_edge_inv = nullptr;  // LTable::edge_inv (rebuilt when the map or the ray table changes)
  void* d_actions = nullptr;  // staging for sb_step_host
  size_t actions_cap = 0;
  double* d_social = nullptr;
  int64_t launches = 0;
  // optional per-kernel CUDA-event timing (sb_set_timing / sb_collect_timing)
  bool timing = false;
  std::vector<cudaEvent_t> ev_pool;
  size_t ev_used = 0;
  struct Span { int kind; cudaEvent_t a, b; };
  std::vector<Span> spans;
  int cuda_err = 0;
  std::string err;
  size_t raster_smem = 0;
  RasterDims raster_dims{};
  // grid transport of sb_copy_obs_to_host: device bit image, pinned staging, expansion pool
  uint32_t* d_bits = nullptr;
  uint8_t* h_bits = nullptr;
  size_t bits_bytes = 0;
  void* pool = nullptr;
  std::vector<cudaEvent_t> xfer_ev;
  double xfer_frac = 1.0;  // share of the agents whose grids take the bit-packed path (SB_XFER_PACKED_SHARE)
  // incremental grid writes (sb_set_incremental_host_grids): the bit image the destination was last written from
  bool incremental = false, prev_valid = false;
  uint8_t* h_prev = nullptr;
  size_t prev_bytes = 0;
  const void* prev_dst[2] = {nullptr, nullptr};
  size_t prev_len[2] = {0, 0};
  bool xfer_diff = false;
  size_t xfer_bits = 0;
};

namespace {

int fail(SbHandle h, int code, const std::string& msg) {
  if (h) h->err = msg;
  return code;
}

bool cuda_ok(SbHandle h, cudaError_t e, const char* what) {
  if (e == cudaSuccess) return true;
  h->cuda_err = (int)e;
  h->err = std::string(what) + ": " + cudaGetErrorString(e);
  return false;
}

template <class T>
bool dev_alloc(SbHandle h, T** p, size_t n, bool zero = true) {
  void* q = nullptr;
  size_t bytes = (n ? n : 1) * sizeof(T);
  if (!cuda_ok(h, cudaMalloc(&q, bytes), "cudaMalloc")) return false;
  if (zero && !cuda_ok(h, cudaMemset(q, 0, bytes), "cudaMemset")) return false;
  h->allocs.push_back(q);
  *p = static_cast<T*>(q);
  return true;
}

template <class T>
bool dev_upload(SbHandle h, const T** dst, const T* src, size_t n) {
  T* q = nullptr;
  if (!dev_alloc(h, &q, n, false)) return false;
  if (n && !cuda_ok(h, cudaMemcpy(q, src, n * sizeof(T), cudaMemcpyHostToDevice), "cudaMemcpy(map)")) return false;
  *dst = q;
  return true;
}

// One bit per grid pixel (pixels are 0 / 255): thread t packs bytes [32 t, 32 t + 32) into word t,
// bit i = byte 32 t + i.  Pure streaming: 256-bit loads, 1/8 of that written.
__global__ void __launch_bounds__(256) k_pack_bits(const uint8_t* __restrict__ src, uint32_t* __restrict__ dst,
                                                   size_t n_words) {
  size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n_words) return;
  const uint4* p = reinterpret_cast<const uint4*>(src) + t * 2;
  const uint4 a = __ldcs(p), b = __ldcs(p + 1);
  auto nib = [](uint32_t w) { return ((w & 0x01010101u) * 0x01020408u) >> 24; };
  dst[t] = nib(a.x) | (nib(a.y) << 4) | (nib(a.z) << 8) | (nib(a.w) << 12) | (nib(b.x) << 16) | (nib(b.y) << 20) |
           (nib(b.z) << 24) | (nib(b.w) << 28);
}

__global__ void k_fill(double* p, size_t n, double v) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) p[i] = v;
}

// SMARTS.reset for a list of envs: agents at their mission start with the sedan chassis box
// (models/vehicle.urdf:12-16), social slots cleared.
__global__ void k_reset(DCfg c, DState s, DEvents ev, const int32_t* env_ids, int n, const double* ax,
                        const double* ay, const double* ah, const double* av, const int32_t* route,
                        const int32_t* route_len, const double* gx, const double* gy, const double* gr,
                        int keep_social) {
  int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n * c.V) return;
  int i = t / c.V, slot = t - i * c.V;
  int e = env_ids ? env_ids[i] : i;
  int vi = e * c.V + slot;
  if (slot < c.A) {
    int ai = e * c.A + slot;
    size_t src = (size_t)i * c.A + slot;
    double h = d_wrap_heading(ah[src]);
    s.pos_x[vi] = ax[src];
    s.pos_y[vi] = ay[src];
    s.pos_z[vi] = 0.0;
    s.heading[vi] = h;
    s.speed[vi] = av[src];
    s.yaw_rate[vi] = 0.0;
    s.length[vi] = c.veh.length;
    s.width[vi] = c.veh.width;
    s.height[vi] = c.veh.height;
    sb_sincos(h, &s.sin_h[vi], &s.cos_h[vi]);
    s.alive[vi] = 1;
    s.steering[ai] = 0.0;
    s.v_long[ai] = av[src];
    s.v_lat[ai] = 0.0;
    for (int k = 0; k < SB_CTRL_STATE; ++k) s.ctrl[(size_t)ai * SB_CTRL_STATE + k] = 0.0;
    for (int k = 0; k < 3; ++k) s.last_control[(size_t)ai * 3 + k] = 0.0;
    s.status[ai] = 0;
    s.ticked[ai] = 0;  // the first observation of a new episode is the `sim.resetting` one
    for (int k = 0; k < c.route_cap; ++k)
      s.route[(size_t)ai * c.route_cap + k] = route ? route[src * c.route_cap + k] : -1;
    s.route_len[ai] = route_len ? route_len[src] : 0;
    if (c.events_enable) {  // new SensorState / TripMeterSensor / DrivenPathSensor (vehicle.py:344-457)
      ev.trip_init[ai] = 0;
      ev.trip_has[ai] = 0;
      ev.trip_x[ai] = ev.trip_y[ai] = ev.trip_h[ai] = ev.trip_dist[ai] = ev.trip_last[ai] = 0.0;
      ev.trip_wp[(size_t)ai * 8 + 3] = -1.0;
      ev.trip_wp[(size_t)ai * 8 + 7] = -1.0;
      ev.steps[ai] = 0;
      ev.path_n[ai] = 0;
      ev.last_px[ai] = ev.last_py[ai] = 0.0;
      ev.goal_x[ai] = gx ? gx[src] : 0.0;
      ev.goal_y[ai] = gy ? gy[src] : 0.0;
      ev.goal_r[ai] = gr ? gr[src] : -1.0;
    }
  } else if (!keep_social) {
    s.alive[vi] = 0;
    s.pos_x[vi] = s.pos_y[vi] = s.pos_z[vi] = s.heading[vi] = s.speed[vi] = s.yaw_rate[vi] = 0.0;
    s.length[vi] = s.width[vi] = s.height[vi] = 0.0;
    s.cos_h[vi] = 1.0;
    s.sin_h[vi] = 0.0;
  }
}

cudaEvent_t next_event(SbHandle h) {
  if (h->ev_used == h->ev_pool.size()) {
    cudaEvent_t e;
    cudaEventCreate(&e);
    h->ev_pool.push_back(e);
  }
  return h->ev_pool[h->ev_used++];
}

// brackets one kernel launch with events on its stream when timing is enabled
struct LaunchScope {
  SbHandle h;
  cudaStream_t st;
  cudaEvent_t a = nullptr, b = nullptr;
  int kind;
  LaunchScope(SbHandle h_, int kind_, cudaStream_t st_) : h(h_), st(st_), kind(kind_) {
    h->launches++;
    if (h->timing) {
      a = next_event(h);
      b = next_event(h);
      cudaEventRecord(a, st);
    }
  }
  ~LaunchScope() {
    if (h->timing) {
      cudaEventRecord(b, st);
      h->spans.push_back({kind, a, b});
    }
  }
};

// Every entry point runs with the handle's device current and hands the caller's back (a process may hold handles on
// several GPUs, or call from a thread whose current device is another one).
struct DeviceGuard {
  int prev = -1;
  bool ok = true;
  explicit DeviceGuard(int dev) {
    if (cudaGetDevice(&prev) != cudaSuccess) prev = -1;
    if (prev != dev) ok = cudaSetDevice(dev) == cudaSuccess;
  }
  ~DeviceGuard() {
    int cur = -1;
    if (prev >= 0 && cudaGetDevice(&cur) == cudaSuccess && cur != prev) cudaSetDevice(prev);
  }
};

// Called at the end of sb_set_social / sb_step / sb_observe: remembers "everything queued on st so far" as an event
// the host-buffer calls can wait for.  Work on the legacy stream needs no mark (those calls run on it), and a stream
// under capture cannot be marked (the record would become a graph node and the event could not be waited on): after
// replaying a captured tick on a non-blocking stream the caller synchronises it before a host-buffer call.
void mark_work(SbHandle h, cudaStream_t st) {
  if (st == nullptr || st == cudaStreamLegacy) return;  // (an earlier mark on another stream stays pending)
  cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
  if (cudaStreamIsCapturing(st, &cs) != cudaSuccess) {
    cudaGetLastError();
    return;
  }
  if (cs != cudaStreamCaptureStatusNone) return;
  if (!h->work_ev && cudaEventCreateWithFlags(&h->work_ev, cudaEventDisableTiming) != cudaSuccess) {
    cudaGetLastError();
    h->work_ev = nullptr;
    return;
  }
  h->work_pending = cudaEventRecord(h->work_ev, st) == cudaSuccess;
}

// The host-buffer calls work on the legacy stream: make it wait (on the device, no host stall) for the marked work.
void wait_marked_work(SbHandle h) {
  if (h->work_pending && cudaStreamWaitEvent(nullptr, h->work_ev, 0) != cudaSuccess) cudaGetLastError();
  h->work_pending = false;
}

size_t action_bytes(const SbConfig& c) {
  size_t na = (size_t)c.n_envs * c.n_agents;
  switch (c.action_space) {
    case SB_ACTION_LANE: return na * sizeof(int32_t);
    case SB_ACTION_LANE_WITH_CONTINUOUS_SPEED: return na * 2 * sizeof(double);
    case SB_ACTION_TRAJECTORY: return na * 4 * (size_t)c.traj_len * sizeof(double);
    default: return na * 3 * sizeof(double);
  }
}

}  // namespace

extern "C" {

int sb_abi_version(void) { return SB_ABI_VERSION; }

const char* sb_last_error(SbHandle h) { return h ? h->err.c_str() : "null handle"; }

int64_t sb_launch_count(SbHandle h) { return h ? h->launches : 0; }

int sb_create(const SbConfig* cfg, SbHandle* out) {
  if (!cfg || !out) return SB_EINVAL;
  *out = nullptr;
  SbHandle h = new SbHandle_();
  h->cfg = *cfg;
  *out = h;  // returned even on failure so sb_last_error works; caller must sb_destroy
  const SbConfig& c = h->cfg;
  if (c.abi_version != SB_ABI_VERSION) return fail(h, SB_EINVAL, "abi_version mismatch");
  if (c.n_envs < 1 || c.n_agents < 1 || c.n_social < 0) return fail(h, SB_EINVAL, "n_envs/n_agents/n_social");
  if (!(c.dt > 0)) return fail(h, SB_EINVAL, "dt must be > 0");
  switch (c.action_space) {
    case SB_ACTION_CONTINUOUS: case SB_ACTION_LANE: case SB_ACTION_ACTUATOR_DYNAMIC:
    case SB_ACTION_LANE_WITH_CONTINUOUS_SPEED: break;
    case SB_ACTION_TRAJECTORY:
      if (c.traj_len < 2 || c.traj_len > SB_MAX_TRAJ) return fail(h, SB_EINVAL, "traj_len out of range");
      break;
    default: return fail(h, SB_EINVAL, "unsupported action_space");
  }
  if (c.route_cap < 1) return fail(h, SB_EINVAL, "route_cap must be >= 1");
  if (c.nbr_enable && c.nbr_cap < 1) return fail(h, SB_EINVAL, "nbr_cap");
  if (c.wp_enable && (c.wp_lookahead < 1 || c.wp_lookahead > SB_MAX_LOOKAHEAD || c.wp_out_paths < 1 ||
                      c.wp_out_paths > SB_MAX_PATHS || c.wp_out_len < 1 || c.wp_out_len > SB_MAX_LOOKAHEAD + 1))
    return fail(h, SB_EINVAL, "waypoint sensor limits");
  if (c.lidar_enable && (c.lidar_rays < 1 || c.lidar_rays > SB_LIDAR_MAX_RAYS)) return fail(h, SB_EINVAL, "lidar_rays");
  auto grid_ok = [](int w, int hh, double r) { return w >= 1 && hh >= 1 && r > 0 && w <= 1024 && hh <= 1024; };
  if (c.ogm_enable && !grid_ok(c.ogm_width, c.ogm_height, c.ogm_resolution))
    return fail(h, SB_EINVAL, "ogm: width and height must be in 1..1024, resolution > 0");
  if (c.dagm_enable && !grid_ok(c.dagm_width, c.dagm_height, c.dagm_resolution))
    return fail(h, SB_EINVAL, "dagm: width and height must be in 1..1024, resolution > 0");
  if (c.n_agents + c.n_social > SB_LIDAR_MAX_BOXES + 1 && c.lidar_enable)
    return fail(h, SB_EINVAL, "lidar: too many vehicles per env");

  DeviceGuard dg(c.device);  // (the caller's current device is restored on return)
  if (!cuda_ok(h, dg.ok ? cudaSuccess : cudaErrorInvalidDevice, "cudaSetDevice")) return SB_ECUDA;
  DCfg& d = h->dc;
  d.E = c.n_envs; d.A = c.n_agents; d.M = c.n_social; d.V = d.A + d.M;
  d.dt = c.dt;
  d.substeps = c.substeps > 0 ? c.substeps : ((int)(c.dt * 240) > 1 ? (int)(c.dt * 240) : 1);
  d.action_space = c.action_space; d.traj_len = c.traj_len; d.route_cap = c.route_cap;
  d.nbr_enable = c.nbr_enable; d.nbr_cap = c.nbr_cap; d.nbr_radius = c.nbr_radius;
  d.wp_enable = c.wp_enable; d.wp_lookahead = c.wp_lookahead; d.wp_out_paths = c.wp_out_paths;
  d.wp_out_len = c.wp_out_len;
  d.lidar_enable = c.lidar_enable; d.lidar_rays = c.lidar_rays; d.lidar_miss_inf = c.lidar_miss_inf;
  d.lidar_wall_height = c.lidar_wall_height;
  d.ogm_enable = c.ogm_enable; d.ogm_w = c.ogm_width; d.ogm_h = c.ogm_height; d.ogm_res = c.ogm_resolution;
  d.dagm_enable = c.dagm_enable; d.dagm_w = c.dagm_width; d.dagm_h = c.dagm_height; d.dagm_res = c.dagm_resolution;
  d.events_enable = c.events_enable; d.max_steps = c.max_episode_steps; d.done_mask = c.done_mask;
  d.nm_time = c.not_moving_time; d.nm_dist = c.not_moving_distance;
  d.ttc_enable = c.ttc_enable;
  d.accel_enable = c.accel_enable; d.via_cap = c.via_cap; d.via_range = c.via_acquisition_range;
  d.via_speed_acc = c.via_speed_accuracy;
  d.ego_frame = c.ego_frame;
  if (c.vehicle_type != SB_VEHICLE_SEDAN && c.vehicle_type != SB_VEHICLE_BUS) return fail(h, SB_EINVAL, "vehicle_type must be SB_VEHICLE_SEDAN or SB_VEHICLE_BUS");
  sb_vehicle(c.vehicle_type, &d.veh);
  d.rw_enable = c.rw_enable; d.rw_horizon = c.rw_horizon; d.rw_lanes_cap = c.rw_lanes_cap; d.rw_paths_cap = c.rw_paths_cap;
  d.env_list = nullptr;
  d.n_list = 0;
  // (A = 1: the magic number would be 2^32 + 1 and wrap to 1 -- every work item would land in env 0; plain division there)
  d.a_magic = (c.n_agents >= 2 && (uint64_t)c.n_envs * c.n_agents * c.n_agents < (1ull << 32))
                  ? (uint32_t)((1ull << 32) / c.n_agents) + 1u : 0u;
  if (c.via_cap < 0 || c.via_cap > SB_MAX_VIAS) return fail(h, SB_EINVAL, "via_cap out of range");
  if (c.via_cap > 0 && !(c.via_acquisition_range >= 0 && c.via_speed_accuracy >= 0))
    return fail(h, SB_EINVAL, "via range / speed accuracy must be >= 0");
  if (c.ttc_enable && !(c.nbr_enable && c.wp_enable)) return fail(h, SB_EINVAL, "ttc needs the neighbour and waypoint sensors");
  if (c.rw_enable && !(c.rw_horizon >= 1 && 2 * c.rw_horizon <= SB_MAX_LOOKAHEAD))
    return fail(h, SB_EINVAL, "rw_horizon must be in 1 .. SB_MAX_LOOKAHEAD / 2");
  if (c.rw_enable && !(c.rw_lanes_cap >= 1 && c.rw_paths_cap >= 1)) return fail(h, SB_EINVAL, "rw_lanes_cap / rw_paths_cap must be >= 1");
  if (c.events_enable && !(c.not_moving_time >= 0)) return fail(h, SB_EINVAL, "not_moving_time must be >= 0");

  const size_t nv = (size_t)d.E * d.V, na = (size_t)d.E * d.A;
  DState& s = h->st;
  bool ok = dev_alloc(h, &s.pos_x, nv) && dev_alloc(h, &s.pos_y, nv) && dev_alloc(h, &s.pos_z, nv) &&
            dev_alloc(h, &s.heading, nv) && dev_alloc(h, &s.speed, nv) && dev_alloc(h, &s.yaw_rate, nv) &&
            dev_alloc(h, &s.length, nv) && dev_alloc(h, &s.width, nv) && dev_alloc(h, &s.height, nv) &&
            dev_alloc(h, &s.cos_h, nv) && dev_alloc(h, &s.sin_h, nv) && dev_alloc(h, &s.alive, nv) &&
            dev_alloc(h, &s.steering, na) && dev_alloc(h, &s.v_long, na) && dev_alloc(h, &s.v_lat, na) &&
            dev_alloc(h, &s.ctrl, na * SB_CTRL_STATE) && dev_alloc(h, &s.last_control, na * 3) &&
            dev_alloc(h, &s.route, na * c.route_cap) && dev_alloc(h, &s.route_len, na) &&
            dev_alloc(h, &s.status, na) && dev_alloc(h, &s.ticked, na);
  if (!ok) return SB_ENOMEM;
  k_fill<<<(unsigned)((nv + 255) / 256), 256>>>(s.cos_h, nv, 1.0);
  cudaMemset(s.route, 0xff, na * c.route_cap * sizeof(int32_t));
  DObs& o = h->ob;
  std::memset(&o, 0, sizeof(o));
  ok = dev_alloc(h, &o.ego_lane, na);
  if (ok && c.nbr_enable) {
    size_t n = na * c.nbr_cap;
    ok = dev_alloc(h, &o.nbr_count, na) && dev_alloc(h, &o.nbr_id, n) && dev_alloc(h, &o.nbr_pos, n * 3) &&
         dev_alloc(h, &o.nbr_box, n * 3) && dev_alloc(h, &o.nbr_heading, n) && dev_alloc(h, &o.nbr_speed, n) &&
         dev_alloc(h, &o.nbr_lane, n) && dev_alloc(h, &o.nbr_lane_index, n);
  }
  if (ok && c.wp_enable) {
    size_t np = na * c.wp_out_paths, n = np * c.wp_out_len;
    ok = dev_alloc(h, &o.wp_n_paths, na) && dev_alloc(h, &o.wp_len, np) && dev_alloc(h, &o.wp_pos, n * 3) &&
         dev_alloc(h, &o.wp_heading, n) && dev_alloc(h, &o.wp_lane_width, n) && dev_alloc(h, &o.wp_speed_limit, n) &&
         dev_alloc(h, &o.wp_lane, n) && dev_alloc(h, &o.wp_lane_index, n);
  }
  if (ok && c.lidar_enable)
    ok = dev_alloc(h, &o.lidar_points, na * c.lidar_rays * 3) && dev_alloc(h, &o.lidar_hits, na * c.lidar_rays);
  // the grids are the bulk of the memory (E*A*W*H bytes each); written in full every tick
  if (ok && c.ogm_enable) ok = dev_alloc(h, &o.ogm, na * c.ogm_width * c.ogm_height, false);
  if (ok && c.dagm_enable) ok = dev_alloc(h, &o.dagm, na * c.dagm_width * c.dagm_height, false);
  if (!ok) return SB_ENOMEM;
  if (c.events_enable) {
    DEvents& v = h->ev;
    DEvOut& q = h->evo;
    // DrivenPathSensor keeps 500 entries (sensors.py:838-840); the not-moving window needs
    // floor(not_moving_time / dt) intervals of them
    double win = floor(c.not_moving_time / c.dt + 1e-9);
    v.ring = (int)(win + 1 < SB_DRIVEN_PATH_MAX - 1 ? win + 1 : SB_DRIVEN_PATH_MAX - 1) + 1;  // one spare interval
    sb_clock_decimals(c.dt, &v.dt_num, &v.dt_den);
    ok = dev_alloc(h, &v.trip_init, na) && dev_alloc(h, &v.trip_has, na) && dev_alloc(h, &v.trip_x, na) &&
         dev_alloc(h, &v.trip_y, na) && dev_alloc(h, &v.trip_h, na) && dev_alloc(h, &v.trip_dist, na) &&
         dev_alloc(h, &v.trip_last, na) && dev_alloc(h, &v.trip_wp, na * 8) && dev_alloc(h, &v.steps, na) &&
         dev_alloc(h, &v.path_n, na) && dev_alloc(h, &v.path_cum, na * v.ring) && dev_alloc(h, &v.last_px, na) &&
         dev_alloc(h, &v.last_py, na) && dev_alloc(h, &v.goal_x, na) && dev_alloc(h, &v.goal_y, na) &&
         dev_alloc(h, &v.goal_r, na) && dev_alloc(h, &q.events, na) && dev_alloc(h, &q.done, na) &&
         dev_alloc(h, &q.n_collisions, na) && dev_alloc(h, &q.collidee, na) && dev_alloc(h, &q.distance, na) &&
         dev_alloc(h, &q.reward, na) && dev_alloc(h, &q.steps, na);
    if (!ok) return SB_ENOMEM;
    k_fill<<<(unsigned)((na + 255) / 256), 256>>>(v.goal_r, na, -1.0);
  }
  if (c.ttc_enable) {
    DTtcOut& q = h->ttc;
    if (!(dev_alloc(h, &q.angle_error, na) && dev_alloc(h, &q.dist_center, na) && dev_alloc(h, &q.dtc, na * 3) &&
          dev_alloc(h, &q.ttc, na * 3) && dev_alloc(h, &q.valid, na)))
      return SB_ENOMEM;
  }
  if (c.accel_enable) {
    DExtras& q = h->ex;
    if (!(dev_alloc(h, &q.lin_hist, na * 4) && dev_alloc(h, &q.ang_hist, na * 2) && dev_alloc(h, &q.hist_n, na) &&
          dev_alloc(h, &q.acc_lin, na * 3) && dev_alloc(h, &q.acc_ang, na * 3) && dev_alloc(h, &q.jerk_lin, na * 3) &&
          dev_alloc(h, &q.jerk_ang, na * 3)))
      return SB_ENOMEM;
  }
  if (c.via_cap > 0) {
    DExtras& q = h->ex;
    const size_t nc = na * c.via_cap;
    if (!(dev_alloc(h, &q.via_count, na) && dev_alloc(h, &q.via_lane, nc) && dev_alloc(h, &q.via_x, nc) &&
          dev_alloc(h, &q.via_y, nc) && dev_alloc(h, &q.via_hit, nc) && dev_alloc(h, &q.via_speed, nc) &&
          dev_alloc(h, &q.via_consumed, na) && dev_alloc(h, &q.near_count, na) && dev_alloc(h, &q.near_idx, nc) &&
          dev_alloc(h, &q.hit_count, na) && dev_alloc(h, &q.hit_idx, nc)))
      return SB_ENOMEM;
  }
  if (c.rw_enable) {
    DRoadWp& q = h->rw;
    const size_t np = na * c.rw_paths_cap, n = np * (2 * c.rw_horizon + 1);
    if (!(dev_alloc(h, &q.n_lanes, na) && dev_alloc(h, &q.lane, na * c.rw_lanes_cap) && dev_alloc(h, &q.n_paths, na) &&
          dev_alloc(h, &q.path_slot, np) && dev_alloc(h, &q.len, np) && dev_alloc(h, &q.pos, n * 2) &&
          dev_alloc(h, &q.heading, n) && dev_alloc(h, &q.lane_width, n) && dev_alloc(h, &q.speed_limit, n) &&
          dev_alloc(h, &q.wp_lane, n) && dev_alloc(h, &q.lane_index, n)))
      return SB_ENOMEM;
  }
  if (c.nbr_enable && d.V <= d.A * d.nbr_cap && !dev_alloc(h, &h->d_veh_lane, nv)) return SB_ENOMEM;
  if (!dev_alloc(h, &h->d_picks, na)) return SB_ENOMEM;
  if (!dev_alloc(h, &h->d_social, (size_t)d.E * (d.M ? d.M : 1) * 8)) return SB_ENOMEM;
  h->actions_cap = action_bytes(c);
  void* da = nullptr;
  if (!cuda_ok(h, cudaMalloc(&da, h->actions_cap), "cudaMalloc(actions)")) return SB_ENOMEM;
  h->allocs.push_back(da);
  h->d_actions = da;
  {  // k_raster shared memory: table | staging area | bit image 0 | bit image 1
    // grids of at most 128 rows: 256 / T agents per block, T threads each (k_raster_slices)
    int hmax = 0;
    if (c.ogm_enable) hmax = c.ogm_height;
    if (c.dagm_enable && c.dagm_height > hmax) hmax = c.dagm_height;
    const int T = hmax <= 32 ? 32 : hmax <= 64 ? 64 : hmax <= 128 ? 128 : SB_RASTER_THREADS;
    auto rows256 = [T](int hh) { return (size_t)(hh < T ? hh : T); };
    auto row_words = [](int ww) { return (size_t)(((ww + 31) / 32 + 7) / 8 * 8 + 1); };
    size_t b0 = c.ogm_enable ? row_words(c.ogm_width) * 4 * rows256(c.ogm_height) : 0;
    size_t b1 = c.dagm_enable ? row_words(c.dagm_width) * 4 * rows256(c.dagm_height) : 0;
    int wmax = 0;
    if (c.ogm_enable) wmax = c.ogm_width;
    if (c.dagm_enable && c.dagm_width > wmax) wmax = c.dagm_width;
    size_t stage = T == 32 ? sizeof(RasterStageT<32, SB_RASTER_SLICE_ROUND>) : T == 64 ? sizeof(RasterStageT<64, SB_RASTER_SLICE_ROUND>)
                   : T == 128 ? sizeof(RasterStageT<128, SB_RASTER_SLICE_ROUND>) : sizeof(RasterStage);
    stage = (stage + 15) / 16 * 16;
    (void)wmax;
    const size_t slice_bytes = stage + (b0 + b1 + 15) / 16 * 16;
    h->raster_smem = SB_RASTER_LUT_BYTES + (SB_RASTER_THREADS / T) * slice_bytes;
    RasterDims& g = h->raster_dims;
    g.W0 = c.ogm_enable ? c.ogm_width : 0;
    g.H0 = c.ogm_enable ? c.ogm_height : 0;
    g.W1 = c.dagm_enable ? c.dagm_width : 0;
    g.H1 = c.dagm_enable ? c.dagm_height : 0;
    g.wpr0 = (g.W0 + 31) / 32;
    g.wpr1 = (g.W1 + 31) / 32;
    g.rows0 = (int)rows256(g.H0);
    g.rows1 = (int)rows256(g.H1);
    g.n_veh = c.ogm_enable ? d.V : 0;
    g.n_src = g.n_veh;  // + map polygons once the map is uploaded (sb_upload_map)
    g.bits_off = (int)stage;
    g.slice_threads = T;
    g.slice_bytes = (int)slice_bytes;
    g.stride0 = g.W0 ? (int)row_words(g.W0) : 0;
    g.stride1 = g.W1 ? (int)row_words(g.W1) : 0;
    g.n_bit_words = g.rows0 * g.stride0 + g.rows1 * g.stride1;
    g.rad0 = 0.5f * (float)c.ogm_resolution * sqrtf((float)g.W0 * g.W0 + (float)g.H0 * g.H0) + 1.0f;
    g.rad1 = 0.5f * (float)c.dagm_resolution * sqrtf((float)g.W1 * g.W1 + (float)g.H1 * g.H1) + 1.0f;
    if (h->raster_smem > 200 * 1024) return fail(h, SB_EINVAL, "grid too large for the shared-memory bit image");
    const void* fn = T == 32 ? (const void*)k_raster_slices<32> : T == 64 ? (const void*)k_raster_slices<64>
                     : T == 128 ? (const void*)k_raster_slices<128> : (const void*)k_raster;
    if (!cuda_ok(h, cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->raster_smem),
                 "cudaFuncSetAttribute"))
      return SB_ECUDA;
  }
  // Test / A-B hooks for the LiDAR kernel choice.  They exist for tests/ and profiles/ only, so they are dead unless
  // SB_DEBUG_HOOKS=1 is set as well (a stray SB_LIDAR_* in a production environment changes nothing), and a value that
  // is not a whole number in range fails sb_create instead of silently selecting the slow path.
  if (const char* on = getenv("SB_DEBUG_HOOKS"); on && on[0] == '1' && on[1] == '\0') {
    auto hook = [&](const char* name, long lo, long hi, int* dst) {
      const char* v = getenv(name);
      if (!v) return true;
      char* end = nullptr;
      const long x = strtol(v, &end, 10);
      if (end == v || *end != '\0' || x < lo || x > hi) {
        fail(h, SB_EINVAL, std::string(name) + "=\"" + v + "\": expected a whole number in [" + std::to_string(lo) + ", " +
                               std::to_string(hi) + "]");
        return false;
      }
      *dst = (int)x;
      return true;
    };
    // agents per env from which k_lidar_tiles is used (1: always; A + 1 or more: never)
    if (!hook("SB_LIDAR_TILE_MIN_A", 1, 1 << 20, &h->lidar_tile_min_agents)) return SB_EINVAL;
    // intervals a k_lidar_tiles block may hold before it falls back to the one-agent cast (0: always fall back)
    if (!hook("SB_LIDAR_LIST_CAP", 0, SB_LIDAR_LIST, &h->lidar_list_cap)) return SB_EINVAL;
  }
  if (c.lidar_enable &&
      !cuda_ok(h, cudaFuncSetAttribute((const void*)k_lidar_tiles, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)sizeof(LidarShared)), "cudaFuncSetAttribute(k_lidar_tiles)"))
    return SB_ECUDA;
  if (!cuda_ok(h, cudaDeviceSynchronize(), "sb_create sync")) return SB_ECUDA;
  return SB_OK;
}

void sb_destroy(SbHandle h) {
  if (!h) return;
  DeviceGuard dg(h->cfg.device);
  cudaDeviceSynchronize();
  for (void* p : h->allocs) cudaFree(p);
  for (void* p : h->map_allocs) cudaFree(p);
  if (h->pool) sbx_pool_destroy(h->pool);
  if (h->h_bits) cudaFreeHost(h->h_bits);
  if (h->h_env_list) cudaFreeHost(h->h_env_list);
  std::free(h->h_prev);
  for (auto e : h->xfer_ev) cudaEventDestroy(e);
  for (auto e : h->ev_pool) cudaEventDestroy(e);
  if (h->work_ev) cudaEventDestroy(h->work_ev);
  if (h->env_list_ev) cudaEventDestroy(h->env_list_ev);
  if (h->d_edge_inv) cudaFree(h->d_edge_inv);
  delete h;
}

static bool build_lidar_edge_inv(SbHandle h);

int sb_upload_map(SbHandle h, const SbMap* mp) {
  if (!h || !mp) return SB_EINVAL;
  if (h->cuda_err) return SB_ECUDA;
  if (mp->n_lanes < 1 || mp->n_lanepoints < 1) return fail(h, SB_EINVAL, "empty map");
  DeviceGuard dg(h->cfg.device);
  if (h->have_map) {  // scenario cycling (hiway_env.py:265-285 draws the next scenario on reset): replace the map
    if (!cuda_ok(h, cudaDeviceSynchronize(), "sync before map replacement")) return SB_ECUDA;
    for (void* q : h->map_allocs) cudaFree(q);
    h->map_allocs.clear();
    h->dmap = DMap{};
    h->have_map = false;
    h->stale_agents = true;
    build_lidar_edge_inv(h);  // (no map: drops the table built on the old one)
  }
  const size_t alloc_mark = h->allocs.size();
  struct Claim {  // whatever this call allocated belongs to the map, also when it fails half-way
    SbHandle h;
    size_t mark;
    ~Claim() {
      h->map_allocs.insert(h->map_allocs.end(), h->allocs.begin() + (long)mark, h->allocs.end());
      h->allocs.resize(mark);
    }
  } claim{h, alloc_mark};
  DMap& m = h->dmap;
  m.n_roads = mp->n_roads; m.n_lanes = mp->n_lanes; m.n_lanepoints = mp->n_lanepoints;
  m.n_shape_pts = mp->n_shape_pts; m.n_polys = mp->n_polys; m.n_poly_verts = mp->n_poly_verts;
  m.n_edges = mp->n_edges; m.lanepoint_spacing = mp->lanepoint_spacing;
  {  // chained lane points?  (enables the lane-level path walk of sb_waypoints.cuh)
    bool chained = true;
    for (int l = 0; l < mp->n_lanes && chained; ++l) {
      const int s0 = mp->lane_lp_start[l], cnt = mp->lane_lp_count[l];
      for (int i = s0; i < s0 + cnt; ++i) {
        if (mp->lp_lane[i] != l) chained = false;
        if (i + 1 < s0 + cnt && !(mp->lp_next_start[i + 1] - mp->lp_next_start[i] == 1 &&
                                  mp->lp_next[mp->lp_next_start[i]] == i + 1))
          chained = false;
      }
    }
    m.lp_chained = chained ? 1 : 0;
  }
  {  // lane ranges tile the road range?  (lets one per-lane scan of a road serve as its road-level scan)
    bool tiled = true;
    for (int r = 0; r < mp->n_roads && tiled; ++r) {
      int expect = mp->road_lp_start[r];
      for (int l = mp->road_lane_start[r]; l < mp->road_lane_start[r] + mp->road_lane_count[r]; ++l) {
        if (mp->lane_lp_count[l] <= 0) continue;
        if (mp->lane_lp_start[l] != expect) tiled = false;
        expect += mp->lane_lp_count[l];
      }
      if (mp->road_lp_count[r] > 0 && expect != mp->road_lp_start[r] + mp->road_lp_count[r]) tiled = false;
      if (mp->road_lane_count[r] > SB_WP_CAND) tiled = false;
    }
    m.road_lp_tiled = tiled ? 1 : 0;
  }
  // derived host tables: lane of every shape vertex; polygon AABBs (float, rounded outwards)
  std::vector<int32_t> shape_lane(mp->n_shape_pts);
  for (int l = 0; l < mp->n_lanes; ++l)
    for (int i = mp->lane_shape_start[l]; i < mp->lane_shape_start[l + 1]; ++i) shape_lane[i] = l;
  std::vector<float> bbox((size_t)mp->n_polys * 4);
  for (int p = 0; p < mp->n_polys; ++p) {
    int s = mp->poly_start[p], n = mp->poly_start[p + 1] - s;
    if (n > SB_MAX_POLY_VERTS) return fail(h, SB_EINVAL, "drivable-area polygons must have <= SB_MAX_POLY_VERTS vertices");
    double x0 = 1e300, y0 = 1e300, x1 = -1e300, y1 = -1e300;
    for (int k = 0; k < n; ++k) {
      x0 = fmin(x0, mp->poly_x[s + k]); x1 = fmax(x1, mp->poly_x[s + k]);
      y0 = fmin(y0, mp->poly_y[s + k]); y1 = fmax(y1, mp->poly_y[s + k]);
    }
    bbox[p * 4 + 0] = nextafterf((float)x0, -INFINITY);
    bbox[p * 4 + 1] = nextafterf((float)y0, -INFINITY);
    bbox[p * 4 + 2] = nextafterf((float)x1, INFINITY);
    bbox[p * 4 + 3] = nextafterf((float)y1, INFINITY);
  }
  const size_t R = mp->n_roads, L = mp->n_lanes, P = mp->n_lanepoints;
  bool ok = dev_upload(h, &m.road_lane_start, mp->road_lane_start, R) &&
            dev_upload(h, &m.road_lane_count, mp->road_lane_count, R) &&
            dev_upload(h, &m.road_is_junction, mp->road_is_junction, R) &&
            dev_upload(h, &m.road_lp_start, mp->road_lp_start, R) &&
            dev_upload(h, &m.road_lp_count, mp->road_lp_count, R) &&
            dev_upload(h, &m.road_first_out, mp->road_first_out, R) &&
            dev_upload(h, &m.lane_road, mp->lane_road, L) &&
            dev_upload(h, &m.lane_idx_in_road, mp->lane_idx_in_road, L) &&
            dev_upload(h, &m.lane_width, mp->lane_width, L) && dev_upload(h, &m.lane_speed, mp->lane_speed, L) &&
            dev_upload(h, &m.lane_length, mp->lane_length, L) &&
            dev_upload(h, &m.lane_out_start, mp->lane_out_start, L + 1) &&
            dev_upload(h, &m.lane_out, mp->lane_out, (size_t)(mp->n_lane_out > 0 ? mp->n_lane_out : 1)) &&
            dev_upload(h, &m.lane_lp_start, mp->lane_lp_start, L) &&
            dev_upload(h, &m.lane_lp_count, mp->lane_lp_count, L) &&
            dev_upload(h, &m.lane_shape_start, mp->lane_shape_start, L + 1) &&
            dev_upload(h, &m.lp_x, mp->lp_x, P) && dev_upload(h, &m.lp_y, mp->lp_y, P) &&
            dev_upload(h, &m.lp_heading, mp->lp_heading, P) && dev_upload(h, &m.lp_inferred, mp->lp_inferred, P) &&
            dev_upload(h, &m.lp_lane, mp->lp_lane, P) && dev_upload(h, &m.lp_next_start, mp->lp_next_start, P + 1) &&
            dev_upload(h, &m.lp_next, mp->lp_next, (size_t)(mp->n_lp_next > 0 ? mp->n_lp_next : 1)) &&
            dev_upload(h, &m.shape_x, mp->shape_x, (size_t)mp->n_shape_pts) &&
            dev_upload(h, &m.shape_y, mp->shape_y, (size_t)mp->n_shape_pts) &&
            dev_upload(h, &m.shape_lane, shape_lane.data(), shape_lane.size()) &&
            dev_upload(h, &m.poly_start, mp->poly_start, (size_t)mp->n_polys + 1) &&
            dev_upload(h, &m.poly_x, mp->poly_x, (size_t)mp->n_poly_verts) &&
            dev_upload(h, &m.poly_y, mp->poly_y, (size_t)mp->n_poly_verts) &&
            dev_upload(h, &m.poly_bbox, bbox.data(), bbox.size()) &&
            dev_upload(h, &m.road_edges, mp->road_edges, (size_t)mp->n_edges * 4) &&
            dev_upload(h, &m.lane_in_start, mp->lane_in_start, L + 1) &&
            dev_upload(h, &m.lane_in, mp->lane_in, (size_t)(mp->n_lane_in > 0 ? mp->n_lane_in : 1)) &&
            dev_upload(h, &m.road_par_start, mp->road_par_start, R + 1) &&
            dev_upload(h, &m.road_par, mp->road_par, (size_t)(mp->n_road_par > 0 ? mp->n_road_par : 1));
  if (!ok) return SB_ECUDA;
  {  // bounding box of every road's lane points: lower bound for the "closest lane point per route road" search
    std::vector<double> rb((size_t)mp->n_roads * 4);
    for (int r = 0; r < mp->n_roads; ++r) {
      double x0 = 1e300, y0 = 1e300, x1 = -1e300, y1 = -1e300;
      for (int i = mp->road_lp_start[r]; i < mp->road_lp_start[r] + mp->road_lp_count[r]; ++i) {
        x0 = fmin(x0, mp->lp_x[i]); x1 = fmax(x1, mp->lp_x[i]);
        y0 = fmin(y0, mp->lp_y[i]); y1 = fmax(y1, mp->lp_y[i]);
      }
      rb[r * 4 + 0] = x0; rb[r * 4 + 1] = y0; rb[r * 4 + 2] = x1; rb[r * 4 + 3] = y1;
    }
    if (!dev_upload(h, &m.road_lp_bbox, rb.data(), rb.size())) return SB_ECUDA;
  }
  m.seg_nx = m.seg_ny = 0;
  m.seg_x0 = m.seg_y0 = m.seg_inv_cell = m.seg_rmax = 0.0;
  m.seg_cell_start = m.seg_items = nullptr;
  if (mp->n_shape_pts > 32) {  // uniform grid over the lane-centre segments (small maps: the full scan is one pass)
    const double rmax = 12.0, cell = 16.0, eps = 1e-6;
    double x0 = 1e300, y0 = 1e300, x1 = -1e300, y1 = -1e300;
    for (int i = 0; i < mp->n_shape_pts; ++i) {
      x0 = fmin(x0, mp->shape_x[i]); x1 = fmax(x1, mp->shape_x[i]);
      y0 = fmin(y0, mp->shape_y[i]); y1 = fmax(y1, mp->shape_y[i]);
    }
    x0 -= rmax + eps; y0 -= rmax + eps; x1 += rmax + eps; y1 += rmax + eps;
    const int nx = (int)floor((x1 - x0) / cell) + 1, ny = (int)floor((y1 - y0) / cell) + 1;
    if ((long long)nx * ny <= (1 << 22)) {
      const double inv = 1.0 / cell;
      auto span = [&](double lo, double hi, double o, int n, int* c0, int* c1) {
        *c0 = std::max(0, (int)floor((lo - rmax - eps - o) * inv));
        *c1 = std::min(n - 1, (int)floor((hi + rmax + eps - o) * inv));
      };
      std::vector<int32_t> cnt((size_t)nx * ny + 1, 0);
      for (int pass = 0; pass < 2; ++pass) {
        std::vector<int32_t> fill;
        std::vector<int32_t> items;
        if (pass == 1) {
          for (size_t c2 = 0, acc = 0; c2 <= (size_t)nx * ny; ++c2) {  // exclusive prefix
            int32_t v = cnt[c2];
            cnt[c2] = (int32_t)acc;
            acc += v;
          }
          fill.assign(cnt.begin(), cnt.end() - 1);
          items.resize(cnt.back());
        }
        for (int i = 0; i + 1 < mp->n_shape_pts; ++i) {
          if (shape_lane[i] != shape_lane[i + 1]) continue;
          int cx0, cx1, cy0, cy1;
          span(fmin(mp->shape_x[i], mp->shape_x[i + 1]), fmax(mp->shape_x[i], mp->shape_x[i + 1]), x0, nx, &cx0, &cx1);
          span(fmin(mp->shape_y[i], mp->shape_y[i + 1]), fmax(mp->shape_y[i], mp->shape_y[i + 1]), y0, ny, &cy0, &cy1);
          for (int cy = cy0; cy <= cy1; ++cy)
            for (int cx = cx0; cx <= cx1; ++cx) {
              if (pass == 0) cnt[(size_t)cy * nx + cx]++;
              else items[fill[(size_t)cy * nx + cx]++] = i;  // ascending segment index within a cell
            }
        }
        if (pass == 1) {
          if (!dev_upload(h, &m.seg_cell_start, cnt.data(), cnt.size()) ||
              !dev_upload(h, &m.seg_items, items.data(), items.size()))
            return SB_ECUDA;
          m.seg_nx = nx; m.seg_ny = ny; m.seg_x0 = x0; m.seg_y0 = y0; m.seg_inv_cell = inv; m.seg_rmax = rmax;
        }
      }
    }
  }
  m.lpg_nx = m.lpg_ny = 0;
  m.lpg_x0 = m.lpg_y0 = m.lpg_cell = m.lpg_inv_cell = 0.0;
  m.lpg_cell_start = m.lpg_items = nullptr;
  m.lpg_px = m.lpg_py = nullptr;
  if (mp->n_lanepoints > 64) {  // uniform grid over the lane points (unrouted closest-lane-point queries)
    const double cell = 8.0, inv = 1.0 / cell;
    double x0 = 1e300, y0 = 1e300, x1 = -1e300, y1 = -1e300;
    for (int i = 0; i < mp->n_lanepoints; ++i) {
      x0 = fmin(x0, mp->lp_x[i]); x1 = fmax(x1, mp->lp_x[i]);
      y0 = fmin(y0, mp->lp_y[i]); y1 = fmax(y1, mp->lp_y[i]);
    }
    const int nx = (int)floor((x1 - x0) * inv) + 1, ny = (int)floor((y1 - y0) * inv) + 1;
    if (std::isfinite(x0) && std::isfinite(y0) && (long long)nx * ny <= (1 << 22)) {
      auto cell_of = [&](int i) {
        const int cx = std::min(nx - 1, std::max(0, (int)floor((mp->lp_x[i] - x0) * inv)));
        const int cy = std::min(ny - 1, std::max(0, (int)floor((mp->lp_y[i] - y0) * inv)));
        return (size_t)cy * nx + cx;
      };
      std::vector<int32_t> start((size_t)nx * ny + 1, 0), items(mp->n_lanepoints);
      for (int i = 0; i < mp->n_lanepoints; ++i) start[cell_of(i) + 1]++;
      for (size_t c2 = 0; c2 < (size_t)nx * ny; ++c2) start[c2 + 1] += start[c2];
      std::vector<int32_t> fill(start.begin(), start.end() - 1);
      for (int i = 0; i < mp->n_lanepoints; ++i) items[fill[cell_of(i)]++] = i;
      std::vector<double> px(items.size()), py(items.size());
      for (size_t t = 0; t < items.size(); ++t) {
        px[t] = mp->lp_x[items[t]];
        py[t] = mp->lp_y[items[t]];
      }
      if (!dev_upload(h, &m.lpg_cell_start, start.data(), start.size()) ||
          !dev_upload(h, &m.lpg_items, items.data(), items.size()) || !dev_upload(h, &m.lpg_px, px.data(), px.size()) ||
          !dev_upload(h, &m.lpg_py, py.data(), py.size()))
        return SB_ECUDA;
      m.lpg_nx = nx; m.lpg_ny = ny; m.lpg_x0 = x0; m.lpg_y0 = y0; m.lpg_cell = cell; m.lpg_inv_cell = inv;
    }
  }
  h->raster_dims.n_src = h->raster_dims.n_veh + (h->cfg.dagm_enable ? mp->n_polys : 0);
  h->have_map = true;
  if (!build_lidar_edge_inv(h)) return SB_ECUDA;
  return SB_OK;
}

// LTable::edge_inv: needs the map's road edges and the ray table's groups, whichever arrives second builds it
static bool build_lidar_edge_inv(SbHandle h) {
  if (h->d_edge_inv) {
    cudaFree(h->d_edge_inv);
    h->d_edge_inv = nullptr;
  }
  h->ray_table.edge_inv = nullptr;
  if (!(h->have_map && h->have_rays && h->cfg.lidar_enable && h->cfg.lidar_wall_height > 0.0)) return true;
  const int n = h->ray_table.n_groups * h->dmap.n_edges;
  if (n <= 0) return true;
  if (!cuda_ok(h, cudaMalloc(&h->d_edge_inv, (size_t)n * sizeof(double)), "cudaMalloc(edge_inv)")) return false;
  k_lidar_edge_inv<<<(n + 255) / 256, 256>>>(h->dmap.road_edges, h->dmap.n_edges, h->ray_table.group_u,
                                             h->ray_table.n_groups, h->d_edge_inv);
  if (!cuda_ok(h, cudaDeviceSynchronize(), "k_lidar_edge_inv")) return false;
  h->ray_table.edge_inv = h->d_edge_inv;
  return true;
}

int sb_upload_lidar_rays(SbHandle h, const double* base_rays, int32_t n_rays) {
  if (!h || !base_rays) return SB_EINVAL;
  if (h->cuda_err) return SB_ECUDA;
  if (n_rays != h->cfg.lidar_rays) return fail(h, SB_EINVAL, "n_rays != cfg.lidar_rays");
  DeviceGuard dg(h->cfg.device);
  if (n_rays > SB_LIDAR_MAX_RAYS) return fail(h, SB_EINVAL, "lidar: too many rays");
  // Factor the table into azimuth groups (include/sb_lidar_table.h; the oracle runs the same function) and
  // sort the records by group, vertical rays last: a warp's 32 rays then walk the same interval list.
  std::vector<SbRayRec> rec(n_rays);
  std::vector<double> gu((size_t)n_rays * 2, 0.0);
  const int n_groups = sb_lidar_factor(base_rays, n_rays, rec.data(), gu.data());
  std::stable_sort(rec.begin(), rec.end(), [&](const SbRayRec& p, const SbRayRec& q) {
    const int gp = p.group < 0 ? n_groups : p.group, gq = q.group < 0 ? n_groups : q.group;
    return gp < gq;
  });
  std::vector<int32_t> gstart(n_groups + 1, n_rays);
  for (int i = n_rays - 1; i >= 0; --i)
    if (rec[i].group >= 0) gstart[rec[i].group] = i;
  {  // first vertical ray (or R)
    int nv = n_rays;
    for (int i = n_rays - 1; i >= 0 && rec[i].group < 0; --i) nv = i;
    gstart[n_groups] = nv;
  }
  double reach = 0.0;
  for (int i = 0; i < n_rays; ++i) reach = fmax(reach, rec[i].hlen);
  h->ray_reach = (float)(reach * 1.0001 + 0.01);
  const SbRayRec* d_rec = nullptr;
  const double* d_gu = nullptr;
  const int32_t* d_gs = nullptr;
  if (!dev_upload(h, &d_rec, rec.data(), rec.size()) || !dev_upload(h, &d_gu, gu.data(), (size_t)(n_groups > 0 ? n_groups : 1) * 2) ||
      !dev_upload(h, &d_gs, gstart.data(), gstart.size()))
    return SB_ECUDA;
  h->ray_table.rec = d_rec;
  h->ray_table.group_u = reinterpret_cast<const double2*>(d_gu);
  h->ray_table.group_start = d_gs;
  h->ray_table.n_groups = n_groups;
  h->ray_table.max_hlen = reach;
  h->have_rays = true;
  if (!build_lidar_edge_inv(h)) return SB_ECUDA;
  return SB_OK;
}

// second half of sb_reset: fresh AccelerometerSensor / ViaSensor state and the missions' via points
static bool reset_extras(SbHandle h, const int32_t* env_ids, int n, const SbResetState* init) {
  const DCfg& c = h->dc;
  const size_t na = (size_t)n * c.A, nc = na * (c.via_cap > 0 ? c.via_cap : 0);
  const bool vias = c.via_cap > 0 && init->via_count;
  if (vias) {
    if (!(init->via_lane && init->via_x && init->via_y && init->via_hit_distance && init->via_required_speed)) {
      fail(h, SB_EINVAL, "via_count given without the via arrays");
      return false;
    }
    for (size_t t = 0; t < na; ++t) {
      const int cnt = init->via_count[t];
      if (cnt < 0 || cnt > c.via_cap) {
        fail(h, SB_EINVAL, "via_count out of range");
        return false;
      }
      for (int k = 0; k < cnt; ++k) {
        const int l = init->via_lane[t * c.via_cap + k];
        if (l < 0 || l >= h->dmap.n_lanes) {
          fail(h, SB_EINVAL, "via lane out of range");
          return false;
        }
      }
    }
  }
  const size_t off_ids = 0, off_cnt = ((size_t)n * 4 + 7) / 8 * 8, off_lane = off_cnt + (na * 4 + 7) / 8 * 8,
               off_x = off_lane + (nc * 4 + 7) / 8 * 8, off_y = off_x + nc * 8, off_hit = off_y + nc * 8,
               off_sp = off_hit + nc * 8, total = off_sp + nc * 8 + 8;
  std::vector<unsigned char> host(total, 0);
  if (env_ids) std::memcpy(&host[off_ids], env_ids, (size_t)n * 4);
  if (vias) {
    std::memcpy(&host[off_cnt], init->via_count, na * 4);
    std::memcpy(&host[off_lane], init->via_lane, nc * 4);
    std::memcpy(&host[off_x], init->via_x, nc * 8);
    std::memcpy(&host[off_y], init->via_y, nc * 8);
    std::memcpy(&host[off_hit], init->via_hit_distance, nc * 8);
    std::memcpy(&host[off_sp], init->via_required_speed, nc * 8);
  }
  unsigned char* dev = nullptr;
  if (!cuda_ok(h, cudaMalloc(&dev, total), "cudaMalloc(reset extras)")) return false;
  bool ok = cuda_ok(h, cudaMemcpy(dev, host.data(), total, cudaMemcpyHostToDevice), "cudaMemcpy(reset extras)");
  if (ok) {
    const int threads = n * c.A;
    k_reset_extras<<<(threads + 127) / 128, 128>>>(
        c, h->ex, env_ids ? reinterpret_cast<const int32_t*>(dev + off_ids) : nullptr, n,
        vias ? reinterpret_cast<const int32_t*>(dev + off_cnt) : nullptr,
        reinterpret_cast<const int32_t*>(dev + off_lane), reinterpret_cast<const double*>(dev + off_x),
        reinterpret_cast<const double*>(dev + off_y), reinterpret_cast<const double*>(dev + off_hit),
        reinterpret_cast<const double*>(dev + off_sp));
    h->launches++;
    ok = cuda_ok(h, cudaDeviceSynchronize(), "k_reset_extras");
  }
  cudaFree(dev);
  return ok;
}

int sb_reset(SbHandle h, const int32_t* env_ids, int32_t n, const SbResetState* init) {
  if (!h || !init || !init->agent_x || !init->agent_y || !init->agent_heading || !init->agent_speed) return SB_EINVAL;
  if (h->cuda_err) return SB_ECUDA;
  if (!h->have_map) return fail(h, SB_ESTATE, "sb_reset before sb_upload_map");
  const DCfg& c = h->dc;
  if (!env_ids) n = c.E;
  if (h->stale_agents && env_ids) return fail(h, SB_ESTATE, "the map was replaced: the first reset must cover every env");
  if (n < 1 || n > c.E) return fail(h, SB_EINVAL, "n out of range");
  if (env_ids)
    for (int i = 0; i < n; ++i)
      if (env_ids[i] < 0 || env_ids[i] >= c.E) return fail(h, SB_EINVAL, "env id out of range");
  DeviceGuard dg(h->cfg.device);
  size_t na = (size_t)n * c.A;
  // one staging buffer: ids | x | y | h | v | route | route_len
  size_t off_ids = 0, off_x = ((size_t)n * 4 + 7) / 8 * 8, off_y = off_x + na * 8, off_h = off_y + na * 8,
         off_v = off_h + na * 8, off_gx = off_v + na * 8, off_gy = off_gx + na * 8, off_gr = off_gy + na * 8,
         off_r = off_gr + na * 8, off_rl = off_r + na * c.route_cap * 4, total = off_rl + na * 4;
  const bool goals = init->goal_x && init->goal_y && init->goal_radius;
  std::vector<unsigned char> host(total, 0);
  if (env_ids) std::memcpy(&host[off_ids], env_ids, (size_t)n * 4);
  std::memcpy(&host[off_x], init->agent_x, na * 8);
  std::memcpy(&host[off_y], init->agent_y, na * 8);
  std::memcpy(&host[off_h], init->agent_heading, na * 8);
  std::memcpy(&host[off_v], init->agent_speed, na * 8);
  if (goals) {
    std::memcpy(&host[off_gx], init->goal_x, na * 8);
    std::memcpy(&host[off_gy], init->goal_y, na * 8);
    std::memcpy(&host[off_gr], init->goal_radius, na * 8);
  }
  if (init->route) std::memcpy(&host[off_r], init->route, na * c.route_cap * 4);
  if (init->route_len) std::memcpy(&host[off_rl], init->route_len, na * 4);
  unsigned char* dev = nullptr;
  if (!cuda_ok(h, cudaMalloc(&dev, total), "cudaMalloc(reset)")) return SB_ECUDA;
  bool ok = cuda_ok(h, cudaMemcpy(dev, host.data(), total, cudaMemcpyHostToDevice), "cudaMemcpy(reset)");
  if (ok) {
    int threads = n * c.V;
    k_reset<<<(threads + 127) / 128, 128>>>(
        c, h->st, h->ev, env_ids ? reinterpret_cast<const int32_t*>(dev + off_ids) : nullptr, n,
        reinterpret_cast<const double*>(dev + off_x), reinterpret_cast<const double*>(dev + off_y),
        reinterpret_cast<const double*>(dev + off_h), reinterpret_cast<const double*>(dev + off_v),
        init->route ? reinterpret_cast<const int32_t*>(dev + off_r) : nullptr,
        init->route_len ? reinterpret_cast<const int32_t*>(dev + off_rl) : nullptr,
        goals ? reinterpret_cast<const double*>(dev + off_gx) : nullptr,
        goals ? reinterpret_cast<const double*>(dev + off_gy) : nullptr,
        goals ? reinterpret_cast<const double*>(dev + off_gr) : nullptr, init->keep_social);
    h->launches++;
    ok = cuda_ok(h, cudaDeviceSynchronize(), "k_reset");
  }
  cudaFree(dev);
  if (ok && (c.accel_enable || c.via_cap > 0)) ok = reset_extras(h, env_ids, n, init);
  if (!ok && !h->cuda_err) return SB_EINVAL;
  if (ok && !env_ids) h->stale_agents = false;
  return ok ? SB_OK : SB_ECUDA;
}

int sb_set_social(SbHandle h, const double* dev_states, void* stream) {
  if (!h) return SB_EINVAL;
  if (h->cuda_err) return SB_ECUDA;
  const DCfg& c = h->dc;
  if (c.M == 0) return SB_OK;
  if (!dev_states) return fail(h, SB_EINVAL, "dev_states is null");
  DeviceGuard dg(h->cfg.device);
  int n = c.E * c.M;
  {
    LaunchScope ls(h, SB_K_SOCIAL, (cudaStream_t)stream);
    k_set_social<<<(n + 255) / 256, 256, 0, (cudaStream_t)stream>>>(c, h->st, dev_states);
  }
  if (!cuda_ok(h, cudaGetLastError(), "k_set_social")) return SB_ECUDA;
  mark_work(h, (cudaStream_t)stream);
  return SB_OK;
}

int sb_step(SbHandle h, const void* dev_actions, void* stream) {
  if (!h || !dev_actions) return SB_EINVAL;
  if (h->cuda_err) return SB_ECUDA;
  if (!h->have_map) return fail(h, SB_ESTATE, "sb_step before sb_upload_map");
  if (h->stale_agents) return fail(h, SB_ESTATE, "the map was replaced: reset every env (sb_reset with env_ids = NULL) first");
  const DCfg& c = h->dc;
  int na = c.E * c.A;
  cudaStream_t st = (cudaStream_t)stream;
  DeviceGuard dg(h->cfg.device);
  {
    LaunchScope ls(h, SB_K_CONTROL, st);
    if (c.action_space == SB_ACTION_LANE || c.action_space == SB_ACTION_LANE_WITH_CONTINUOUS_SPEED) {
      k_control_paths<<<(na + SB_CTRL_WARPS - 1) / SB_CTRL_WARPS, SB_CTRL_WARPS * 32, 0, st>>>(c, h->dmap, h->st,
                                                                                             dev_actions, h->d_picks);
      h->launches++;
      k_control_law<<<(na + SB_CTRL_LAW_THREADS - 1) / SB_CTRL_LAW_THREADS, SB_CTRL_LAW_THREADS, 0, st>>>(
          c, h->dmap, h->st, dev_actions, h->d_picks);
    } else {
      k_control_simple<<<(na + 127) / 128, 128, 0, st>>>(c, h->st, static_cast<const double*>(dev_actions));
    }
  }
  if (!cuda_ok(h, cudaGetLastError(), "k_control")) return SB_ECUDA;
  mark_work(h, st);
  return SB_OK;
}

int sb_observe(SbHandle h, void* stream) { return sb_observe_envs(h, nullptr, 0, stream); }

int sb_observe_envs(SbHandle h, const int32_t* env_ids, int32_t n, void* stream) {
  if (!h) return SB_EINVAL;
  if (h->cuda_err) return SB_ECUDA;
  if (!h->have_map) return fail(h, SB_ESTATE, "sb_observe before sb_upload_map");
  if (h->stale_agents) return fail(h, SB_ESTATE, "the map was replaced: reset every env (sb_reset with env_ids = NULL) first");
  if (h->dc.lidar_enable && !h->have_rays) return fail(h, SB_ESTATE, "lidar enabled but no ray table uploaded");
  cudaStream_t st = (cudaStream_t)stream;
  DeviceGuard dg(h->cfg.device);
  DCfg c = h->dc;  // this call's copy: the env list rides in it
  if (env_ids) {
    if (n < 1 || n > c.E) return fail(h, SB_EINVAL, "n out of range");
    for (int i = 0; i < n; ++i)
      if (env_ids[i] < 0 || env_ids[i] >= c.E) return fail(h, SB_EINVAL, "env id out of range");
    // staged through a pinned buffer owned by the handle so that the copy is stream-ordered like the kernels
    if (!h->h_env_list) {
      if (!cuda_ok(h, cudaHostAlloc(&h->h_env_list, (size_t)c.E * sizeof(int32_t), cudaHostAllocDefault), "cudaHostAlloc(env list)") ||
          !dev_alloc(h, &h->d_env_list, (size_t)c.E))
        return SB_ENOMEM;
    } else if (h->env_list_pending) {
      cudaEventSynchronize(h->env_list_ev);  // the previous list's copy has left the pinned buffer
      h->env_list_pending = false;
    }
    std::memcpy(h->h_env_list, env_ids, (size_t)n * sizeof(int32_t));
    if (!cuda_ok(h, cudaMemcpyAsync(h->d_env_list, h->h_env_list, (size_t)n * sizeof(int32_t), cudaMemcpyHostToDevice, st),
                 "H2D env list"))
      return SB_ECUDA;
    {  // (a stream under capture is not marked: the copy is then a graph node that reads the buffer at replay time)
      cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
      if (cudaStreamIsCapturing(st, &cs) != cudaSuccess) cudaGetLastError();
      else if (cs == cudaStreamCaptureStatusNone &&
               (h->env_list_ev || cudaEventCreateWithFlags(&h->env_list_ev, cudaEventDisableTiming) == cudaSuccess))
        h->env_list_pending = cudaEventRecord(h->env_list_ev, st) == cudaSuccess;
    }
    c.env_list = h->d_env_list;
    c.n_list = n;
  }
  const int na = (env_ids ? n : c.E) * c.A;
  {
    LaunchScope ls(h, SB_K_OBSERVE, st);
    if (h->d_veh_lane) {  // fewer vehicles than neighbour slots: one lane lookup per vehicle, shared by the agents
      const int nvs = (env_ids ? n : c.E) * c.V;
      k_vehicle_lanes<<<(nvs + SB_OBS_WARPS - 1) / SB_OBS_WARPS, SB_OBS_WARPS * 32, 0, st>>>(c, h->dmap, h->st,
                                                                                          h->d_veh_lane);
      h->launches++;
    }
    if (!c.wp_enable || c.wp_lookahead <= 32) {
      k_observe<32><<<(na + SB_OBS_WARPS - 1) / SB_OBS_WARPS, SB_OBS_WARPS * 32, 0, st>>>(c, h->dmap, h->st, h->ob,
                                                                                      h->ev.trip_wp, h->ev.trip_init, h->d_veh_lane);
    } else {
      k_observe<SB_MAX_LOOKAHEAD><<<(na + SB_OBS_WARPS - 1) / SB_OBS_WARPS, SB_OBS_WARPS * 32, 0, st>>>(c, h->dmap, h->st, h->ob,
                                                                                      h->ev.trip_wp, h->ev.trip_init, h->d_veh_lane);
    }
  }
  if (c.ttc_enable) {
    LaunchScope ls(h, SB_K_TTC, st);
    k_ttc<<<(na + SB_TTC_WARPS - 1) / SB_TTC_WARPS, SB_TTC_WARPS * 32, 0, st>>>(c, h->st, h->ob, h->ttc);
  }
  if (c.accel_enable || c.via_cap > 0) {
    LaunchScope ls(h, SB_K_EXTRAS, st);
    k_extras<<<(na + 127) / 128, 128, 0, st>>>(c, h->dmap, h->st, h->ex);
  }
  if (c.rw_enable) {
    LaunchScope ls(h, SB_K_ROAD_WP, st);
    k_road_waypoints<<<(na + SB_RW_WARPS - 1) / SB_RW_WARPS, SB_RW_WARPS * 32, 0, st>>>(c, h->dmap, h->st, h->rw);
  }
  if (c.lidar_enable) {
    LaunchScope ls(h, SB_K_LIDAR, st);
    const int G = h->ray_table.n_groups;
    if (c.A >= h->lidar_tile_min_agents && c.V <= SB_LIDAR_MAX_BOXES && G <= SB_LIDAR_MAX_GROUPS && G * c.V <= SB_LIDAR_RECIP) {
      // several agents of an env per block: vehicles staged and (group, vehicle) reciprocals computed once
      const int tiles = (c.A + SB_LIDAR_TILE - 1) / SB_LIDAR_TILE;
      k_lidar_tiles<<<(env_ids ? n : c.E) * tiles, SB_LIDAR_THREADS, sizeof(LidarShared), st>>>(
          c, h->dmap, h->st, h->ob, h->ray_table, h->ray_reach, tiles, h->lidar_list_cap);
    } else {
      k_lidar<<<na, SB_LIDAR_THREADS, 0, st>>>(c, h->dmap, h->st, h->ob, h->ray_table, h->ray_reach);
    }
  }
  if (c.ogm_enable || c.dagm_enable) {
    LaunchScope ls(h, SB_K_RASTER, st);
    const RasterDims& g = h->raster_dims;
    const int per = SB_RASTER_THREADS / g.slice_threads, nb = (na + per - 1) / per;
    if (g.slice_threads == 32)
      k_raster_slices<32><<<nb, SB_RASTER_THREADS, h->raster_smem, st>>>(c, h->dmap, h->st, h->ob, g, na);
    else if (g.slice_threads == 64)
      k_raster_slices<64><<<nb, SB_RASTER_THREADS, h->raster_smem, st>>>(c, h->dmap, h->st, h->ob, g, na);
    else if (g.slice_threads == 128)
      k_raster_slices<128><<<nb, SB_RASTER_THREADS, h->raster_smem, st>>>(c, h->dmap, h->st, h->ob, g, na);
    else
      k_raster<<<na, SB_RASTER_THREADS, h->raster_smem, st>>>(c, h->dmap, h->st, h->ob, g);
  }
  if (c.events_enable) {
    LaunchScope ls(h, SB_K_EVENTS, st);
    k_events<<<(na + SB_EV_WARPS - 1) / SB_EV_WARPS, SB_EV_WARPS * 32, 0, st>>>(c, h->dmap, h->st, h->ev, h->evo);
  }
  if (c.ego_frame && (c.lidar_enable || c.nbr_enable || c.wp_enable || c.rw_enable)) {  // last: k_ttc / k_events read world-frame buffers
    LaunchScope ls(h, SB_K_EGO_FRAME, st);
    k_ego_frame<<<na, SB_EGO_THREADS, 0, st>>>(c, h->st, h->ob, h->rw);
  }
  if (!cuda_ok(h, cudaGetLastError(), "k_observe")) return SB_ECUDA;
  mark_work(h, st);
  return SB_OK;
}

int sb_step_host(SbHandle h, const void* host_actions, int64_t n_bytes, const double* host_social) {
  if (!h || !host_actions) return SB_EINVAL;
  if (h->cuda_err) return SB_ECUDA;
  if ((size_t)n_bytes != h->actions_cap) return fail(h, SB_EINVAL, "action_bytes does not match the action space");
  const DCfg& c = h->dc;
  DeviceGuard dg(h->cfg.device);
  // this call works on the legacy stream; work queued by sb_step / sb_observe on another (non-blocking) stream comes first
  wait_marked_work(h);
  if (!cuda_ok(h, cudaMemcpyAsync(h->d_actions, host_actions, h->actions_cap, cudaMemcpyHostToDevice, 0), "H2D actions"))
    return SB_ECUDA;
  if (c.M > 0) {
    if (!host_social) return fail(h, SB_EINVAL, "host_social is null");
    if (!cuda_ok(h, cudaMemcpyAsync(h->d_social, host_social, (size_t)c.E * c.M * 8 * sizeof(double),
                                    cudaMemcpyHostToDevice, 0), "H2D social"))
      return SB_ECUDA;
    int rc = sb_set_social(h, h->d_social, nullptr);
    if (rc) return rc;
  }
  int rc = sb_step(h, h->d_actions, nullptr);
  if (rc) return rc;
  rc = sb_observe(h, nullptr);
  if (rc) return rc;
  return cuda_ok(h, cudaStreamSynchronize(0), "sb_step_host sync") ? SB_OK : SB_ECUDA;
}

static int copy_obs_to_host(SbHandle h, const SbObsPtrs* dst, bool packed_dst);

int sb_copy_obs_to_host(SbHandle h, const SbObsPtrs* dst) { return copy_obs_to_host(h, dst, false); }

int sb_copy_obs_to_host_packed(SbHandle h, const SbObsPtrs* dst) { return copy_obs_to_host(h, dst, true); }

static int copy_obs_to_host(SbHandle h, const SbObsPtrs* dst, bool packed_dst) {
  if (!h || !dst) return SB_EINVAL;
  if (h->cuda_err) return SB_ECUDA;
  const SbConfig& c = h->cfg;
  const DObs& o = h->ob;
  size_t na = (size_t)c.n_envs * c.n_agents;
  DeviceGuard dg(c.device);
  // copies run on the legacy stream: observations written on another (non-blocking) stream are waited for first
  wait_marked_work(h);
  bool ok = true;
  size_t link_total = 0;  // bytes queued on the PCIe link by this call
  auto cpy = [&](void* d, const void* s, size_t bytes) {
    if (ok && d && s && bytes) {
      ok = cuda_ok(h, cudaMemcpyAsync(d, s, bytes, cudaMemcpyDeviceToHost, 0), "D2H obs");
      link_total += bytes;
    }
  };
  // Grids: packed to 1 bit / pixel on the device, copied in chunks, widened by host threads while
  // the next chunk is in flight (8x fewer PCIe bytes than the byte image).
  struct Sent { size_t off, len; uint8_t* host; };
  std::vector<Sent> sent;
  struct GridJob { const uint8_t* dev; uint8_t* host; size_t bytes; };
  GridJob jobs[2];
  int nj = 0;
  // The first n_packed agents' grids go bit-packed + widened by host threads, the rest as plain byte
  // copies by the DMA engine, both at once.  Default: everything packed -- on the B200 hosts measured the
  // call is bound by host-memory write bandwidth either way (profiles/README.md), and packing keeps the
  // PCIe link free; SB_XFER_PACKED_SHARE=0..1 moves the split for hosts with few cores per GPU.
  if (const char* sh = getenv("SB_XFER_PACKED_SHARE")) {
    const double v = atof(sh);
    h->xfer_frac = v < 0.0 ? 0.0 : (v > 1.0 ? 1.0 : v);
  }
  const size_t n_packed = packed_dst ? na : std::min(na, (size_t)llround(h->xfer_frac * (double)na));
  // (the bit image is 32 pixels per word over the flat [n, H, W] byte array: a grid whose W * H is not a multiple of
  // 32 travels as plain bytes, and cannot be delivered packed)
  const bool pk0 = ((size_t)c.ogm_width * c.ogm_height) % 32 == 0, pk1 = ((size_t)c.dagm_width * c.dagm_height) % 32 == 0;
  if (packed_dst && ((c.ogm_enable && dst->ogm && !pk0) || (c.dagm_enable && dst->dagm && !pk1)))
    return fail(h, SB_EINVAL, "packed delivery needs width * height to be a multiple of 32");
  if (c.ogm_enable && dst->ogm && pk0) jobs[nj++] = {o.ogm, dst->ogm, n_packed * c.ogm_width * c.ogm_height};
  if (c.dagm_enable && dst->dagm && pk1) jobs[nj++] = {o.dagm, dst->dagm, n_packed * c.dagm_width * c.dagm_height};
  const auto t0 = std::chrono::steady_clock::now();
  if (ok && nj && n_packed) {
    size_t need = 0;
    for (int j = 0; j < nj; ++j) need += (jobs[j].bytes / n_packed) * na / 8;  // room for every agent; widths are multiples of 32
    if (h->bits_bytes < need) {
      if (h->h_bits) cudaFreeHost(h->h_bits);
      h->h_bits = nullptr;
      void* q = nullptr;
      if (!cuda_ok(h, cudaMalloc(&q, need), "cudaMalloc(bits)")) return SB_ENOMEM;
      h->allocs.push_back(q);
      h->d_bits = static_cast<uint32_t*>(q);
      if (!cuda_ok(h, cudaHostAlloc(&q, need, cudaHostAllocDefault), "cudaHostAlloc(bits)")) return SB_ENOMEM;
      h->h_bits = static_cast<uint8_t*>(q);
      h->bits_bytes = need;
    }
    bool diff = false;  // incremental: same destinations and sizes as the previous call, written by this library
    if (h->incremental && !packed_dst) {
      if (h->prev_bytes < need) {
        std::free(h->h_prev);
        h->h_prev = static_cast<uint8_t*>(std::malloc(need));
        if (!h->h_prev) return SB_ENOMEM;
        h->prev_bytes = need;
        h->prev_valid = false;
      }
      diff = h->prev_valid;
      for (int j = 0; j < 2; ++j) {
        const void* d = j < nj ? jobs[j].host : nullptr;
        const size_t len = j < nj ? jobs[j].bytes : 0;
        if (h->prev_dst[j] != d || h->prev_len[j] != len) diff = false;
        h->prev_dst[j] = d;
        h->prev_len[j] = len;
      }
    }
    h->xfer_diff = diff;
    if (!h->pool && !packed_dst) {  // host threads: the cores of this box shared by the ranks of a torchrun launch
      unsigned hc = std::thread::hardware_concurrency();
      const char* lws = getenv("LOCAL_WORLD_SIZE");
      const int ranks = lws && atoi(lws) > 0 ? atoi(lws) : 1;
      // every hardware thread of this rank's share: measured best on the B200 host (16 threads 15.0 ms,
      // 8 threads 17.0 ms, 4 threads 19.7 ms per 2.15 GB of grids); SB_XFER_THREADS overrides
      int nt = (int)(hc == 0 ? 4 : hc) / ranks;
      const char* forced = getenv("SB_XFER_THREADS");
      if (forced && atoi(forced) > 0) nt = atoi(forced);
      nt = nt < 2 ? 2 : (nt > 32 ? 32 : nt);
      // several ranks on the box: rank r's workers on its own nt cores (SB_XFER_PIN=0: leave it to the scheduler)
      const char* lr = getenv("LOCAL_RANK");
      const char* pin = getenv("SB_XFER_PIN");
      const bool do_pin = ranks > 1 && lr && !(pin && pin[0] == '0');
      h->pool = sbx_pool_create(nt, do_pin ? atoi(lr) * nt : -1);
    }
    const size_t chunk = (size_t)8 << 20;  // bytes of bits per D2H chunk (64 MB of pixels)
    size_t off = 0, n_ev = 0;
    for (int j = 0; j < nj && ok; ++j) {
      const size_t words = jobs[j].bytes / 32;
      {
        LaunchScope ls(h, SB_K_PACK, 0);
        k_pack_bits<<<(unsigned)((words + 255) / 256), 256>>>(jobs[j].dev, h->d_bits + off / 4, words);
      }
      if (packed_dst) {  // the bit image is the delivery: one copy into the caller's buffer, nothing to widen
        cpy(jobs[j].host, reinterpret_cast<const uint8_t*>(h->d_bits) + off, words * 4);
        off += words * 4;
        continue;
      }
      for (size_t c0 = 0; c0 < words * 4 && ok; c0 += chunk) {
        const size_t len = words * 4 - c0 < chunk ? words * 4 - c0 : chunk;
        ok = cuda_ok(h, cudaMemcpyAsync(h->h_bits + off + c0, reinterpret_cast<const uint8_t*>(h->d_bits) + off + c0,
                                        len, cudaMemcpyDeviceToHost, 0), "D2H grid bits");
        if (n_ev == h->xfer_ev.size()) {
          cudaEvent_t e;
          cudaEventCreateWithFlags(&e, cudaEventDisableTiming);
          h->xfer_ev.push_back(e);
        }
        cudaEventRecord(h->xfer_ev[n_ev++], 0);
        link_total += len;
        sent.push_back({off + c0, len, jobs[j].host + c0 * 8});
      }
      off += words * 4;
    }
  }
  // everything else: plain copies, queued behind the grid chunks so that widening starts at once
  cpy(dst->ego_lane, o.ego_lane, na * 4);
  if (c.nbr_enable) {
    size_t n = na * c.nbr_cap;
    cpy(dst->nbr_count, o.nbr_count, na * 4); cpy(dst->nbr_id, o.nbr_id, n * 4); cpy(dst->nbr_pos, o.nbr_pos, n * 24);
    cpy(dst->nbr_box, o.nbr_box, n * 12); cpy(dst->nbr_heading, o.nbr_heading, n * 4);
    cpy(dst->nbr_speed, o.nbr_speed, n * 4); cpy(dst->nbr_lane, o.nbr_lane, n * 4);
    cpy(dst->nbr_lane_index, o.nbr_lane_index, n);
  }
  if (c.wp_enable) {
    size_t np = na * c.wp_out_paths, n = np * c.wp_out_len;
    cpy(dst->wp_n_paths, o.wp_n_paths, na * 4); cpy(dst->wp_len, o.wp_len, np * 4); cpy(dst->wp_pos, o.wp_pos, n * 24);
    cpy(dst->wp_heading, o.wp_heading, n * 4); cpy(dst->wp_lane_width, o.wp_lane_width, n * 4);
    cpy(dst->wp_speed_limit, o.wp_speed_limit, n * 4); cpy(dst->wp_lane, o.wp_lane, n * 4);
    cpy(dst->wp_lane_index, o.wp_lane_index, n);
  }
  if (c.lidar_enable) {
    cpy(dst->lidar_points, o.lidar_points, na * c.lidar_rays * 24);
    cpy(dst->lidar_hits, o.lidar_hits, na * c.lidar_rays);
  }
  if (c.events_enable) {
    const DEvOut& q = h->evo;
    cpy(dst->ev_events, q.events, na * 4); cpy(dst->ev_done, q.done, na); cpy(dst->ev_n_collisions, q.n_collisions, na * 4);
    cpy(dst->ev_collidee, q.collidee, na * 4); cpy(dst->ev_distance, q.distance, na * 8);
    cpy(dst->ev_reward, q.reward, na * 8); cpy(dst->ev_steps, q.steps, na * 4);
  }
  if (c.accel_enable) {
    const DExtras& q = h->ex;
    cpy(dst->acc_linear, q.acc_lin, na * 24); cpy(dst->acc_angular, q.acc_ang, na * 24);
    cpy(dst->acc_linear_jerk, q.jerk_lin, na * 24); cpy(dst->acc_angular_jerk, q.jerk_ang, na * 24);
  }
  if (c.via_cap > 0) {
    const DExtras& q = h->ex;
    cpy(dst->via_near_count, q.near_count, na * 4); cpy(dst->via_near_idx, q.near_idx, na * 4 * c.via_cap);
    cpy(dst->via_hit_count, q.hit_count, na * 4); cpy(dst->via_hit_idx, q.hit_idx, na * 4 * c.via_cap);
  }
  if (c.rw_enable) {
    const DRoadWp& q = h->rw;
    const size_t np = na * c.rw_paths_cap, n = np * (2 * c.rw_horizon + 1);
    cpy(dst->rw_n_lanes, q.n_lanes, na * 4); cpy(dst->rw_lane, q.lane, na * c.rw_lanes_cap * 4);
    cpy(dst->rw_n_paths, q.n_paths, na * 4); cpy(dst->rw_path_slot, q.path_slot, np * 4); cpy(dst->rw_len, q.len, np * 4);
    cpy(dst->rw_pos, q.pos, n * 16); cpy(dst->rw_heading, q.heading, n * 8); cpy(dst->rw_lane_width, q.lane_width, n * 4);
    cpy(dst->rw_speed_limit, q.speed_limit, n * 4); cpy(dst->rw_wp_lane, q.wp_lane, n * 4);
    cpy(dst->rw_lane_index, q.lane_index, n);
  }
  if (c.ttc_enable) {
    const DTtcOut& q = h->ttc;
    cpy(dst->ttc_angle_error, q.angle_error, na * 4); cpy(dst->ttc_distance_from_center, q.dist_center, na * 4);
    cpy(dst->ttc_dtc, q.dtc, na * 12); cpy(dst->ttc_ttc, q.ttc, na * 12); cpy(dst->ttc_valid, q.valid, na);
  }
  if (n_packed < na) {  // the byte-image share, behind everything else on the link
    const size_t g0 = (size_t)c.ogm_width * c.ogm_height, g1 = (size_t)c.dagm_width * c.dagm_height;
    if (c.ogm_enable && dst->ogm && pk0) { cpy(dst->ogm + n_packed * g0, o.ogm + n_packed * g0, (na - n_packed) * g0); }
    if (c.dagm_enable && dst->dagm && pk1) { cpy(dst->dagm + n_packed * g1, o.dagm + n_packed * g1, (na - n_packed) * g1); }
  }
  if (c.ogm_enable && dst->ogm && !pk0) cpy(dst->ogm, o.ogm, na * c.ogm_width * c.ogm_height);
  if (c.dagm_enable && dst->dagm && !pk1) cpy(dst->dagm, o.dagm, na * c.dagm_width * c.dagm_height);
  static const bool debug = getenv("SB_DEBUG_XFER") != nullptr;
  auto ms_since = [&] { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count(); };
  double t_first = 0, t_last = 0, t_pool = 0;
  if (ok && !sent.empty()) {
    for (size_t k = 0; k < sent.size() && ok; ++k) {
      ok = cuda_ok(h, cudaEventSynchronize(h->xfer_ev[k]), "D2H grid bits wait");
      if (k == 0) t_first = ms_since();
      if (!ok) break;
      if (h->xfer_diff)
        sbx_pool_submit_diff(h->pool, h->h_bits + sent[k].off, h->h_prev + sent[k].off, sent[k].host, sent[k].len,
                             (size_t)128 << 10);
      else
        sbx_pool_submit(h->pool, h->h_bits + sent[k].off, sent[k].host, sent[k].len, (size_t)128 << 10);
    }
    t_last = ms_since();
    sbx_pool_wait(h->pool);
    if (ok && h->incremental && !h->xfer_diff) {  // a full write: remember what the destination now holds
      size_t total = 0;
      for (const Sent& q : sent) total = std::max(total, q.off + q.len);
      std::memcpy(h->h_prev, h->h_bits, total);
      h->prev_valid = true;
    }
    if (!ok) h->prev_valid = false;
    t_pool = ms_since();
  }
  if (!ok) return SB_ECUDA;
  ok = cuda_ok(h, cudaStreamSynchronize(0), "D2H obs sync");
  if (debug)
    fprintf(stderr, "[sb xfer] first chunk %.2f ms, last chunk %.2f ms, widened %.2f ms, all copies %.2f ms "
            "(%d threads, %.1f MB over the link, packed share %.2f)\n",
            t_first, t_last, t_pool, ms_since(), h->pool ? sbx_pool_threads(h->pool) : 0, link_total / 1e6, h->xfer_frac);
  return ok ? SB_OK : SB_ECUDA;
}

int sb_set_incremental_host_grids(SbHandle h, int on) {
  if (!h) return SB_EINVAL;
  h->incremental = on != 0;
  h->prev_valid = false;  // the next call writes everything
  return SB_OK;
}

int sb_set_timing(SbHandle h, int on) {
  if (!h) return SB_EINVAL;
  h->timing = on != 0;
  h->spans.clear();
  h->ev_used = 0;
  return SB_OK;
}

int sb_collect_timing(SbHandle h, double* ms_sum, int64_t* count) {
  if (!h || !ms_sum || !count) return SB_EINVAL;
  DeviceGuard dg(h->cfg.device);
  if (!cuda_ok(h, cudaDeviceSynchronize(), "sb_collect_timing sync")) return SB_ECUDA;
  for (int k = 0; k < SB_K_COUNT; ++k) {
    ms_sum[k] = 0.0;
    count[k] = 0;
  }
  for (auto& s : h->spans) {
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, s.a, s.b) == cudaSuccess) {
      ms_sum[s.kind] += ms;
      count[s.kind]++;
    }
  }
  h->spans.clear();
  h->ev_used = 0;
  return SB_OK;
}

int sb_get_obs(SbHandle h, SbObsPtrs* out) {
  if (!h || !out) return SB_EINVAL;
  const DObs& o = h->ob;
  out->ego_lane = o.ego_lane;
  out->nbr_count = o.nbr_count; out->nbr_id = o.nbr_id; out->nbr_pos = o.nbr_pos; out->nbr_box = o.nbr_box;
  out->nbr_heading = o.nbr_heading; out->nbr_speed = o.nbr_speed; out->nbr_lane = o.nbr_lane;
  out->nbr_lane_index = o.nbr_lane_index;
  out->wp_n_paths = o.wp_n_paths; out->wp_len = o.wp_len; out->wp_pos = o.wp_pos; out->wp_heading = o.wp_heading;
  out->wp_lane_width = o.wp_lane_width; out->wp_speed_limit = o.wp_speed_limit; out->wp_lane = o.wp_lane;
  out->wp_lane_index = o.wp_lane_index;
  out->lidar_points = o.lidar_points; out->lidar_hits = o.lidar_hits;
  out->ogm = o.ogm; out->dagm = o.dagm;
  const DEvOut& q = h->evo;
  out->ev_events = q.events; out->ev_done = q.done; out->ev_n_collisions = q.n_collisions;
  out->ev_collidee = q.collidee; out->ev_distance = q.distance; out->ev_reward = q.reward; out->ev_steps = q.steps;
  out->ttc_angle_error = h->ttc.angle_error; out->ttc_distance_from_center = h->ttc.dist_center;
  out->ttc_dtc = h->ttc.dtc; out->ttc_ttc = h->ttc.ttc; out->ttc_valid = h->ttc.valid;
  out->acc_linear = h->ex.acc_lin; out->acc_angular = h->ex.acc_ang;
  out->acc_linear_jerk = h->ex.jerk_lin; out->acc_angular_jerk = h->ex.jerk_ang;
  out->via_near_count = h->ex.near_count; out->via_near_idx = h->ex.near_idx;
  out->via_hit_count = h->ex.hit_count; out->via_hit_idx = h->ex.hit_idx;
  out->rw_n_lanes = h->rw.n_lanes; out->rw_lane = h->rw.lane; out->rw_n_paths = h->rw.n_paths;
  out->rw_path_slot = h->rw.path_slot; out->rw_len = h->rw.len; out->rw_pos = h->rw.pos; out->rw_heading = h->rw.heading;
  out->rw_lane_width = h->rw.lane_width; out->rw_speed_limit = h->rw.speed_limit; out->rw_wp_lane = h->rw.wp_lane;
  out->rw_lane_index = h->rw.lane_index;
  return SB_OK;
}

int sb_get_state(SbHandle h, SbStatePtrs* out) {
  if (!h || !out) return SB_EINVAL;
  const DState& s = h->st;
  out->pos_x = s.pos_x; out->pos_y = s.pos_y; out->pos_z = s.pos_z; out->heading = s.heading; out->speed = s.speed;
  out->yaw_rate = s.yaw_rate; out->length = s.length; out->width = s.width; out->height = s.height;
  out->cos_h = s.cos_h; out->sin_h = s.sin_h; out->alive = s.alive;
  out->steering = s.steering; out->v_long = s.v_long; out->v_lat = s.v_lat; out->ctrl = s.ctrl;
  out->last_control = s.last_control; out->route = s.route; out->route_len = s.route_len; out->status = s.status;
  return SB_OK;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------------------------
// sb_copy_grids_to_host_packed_noalloc: packed grid delivery for callers that must not allocate (smarts_b200.h).
// The bit image lives in a file-scope device array instead of the cudaMalloc'ed SbHandle_::d_bits.
// ---------------------------------------------------------------------------------------------------------------
#include <mutex>

// 16 MiB of every GPU that loads this library, for the life of the process.  In host code this NAME is only a handle
// for the runtime's symbol table: a kernel launched with it as an argument would read or write host memory (or
// fault).  The device address comes from cudaGetSymbolAddress, below.
static __device__ __align__(256) uint32_t g_noalloc_bits[SB_NOALLOC_BITS_MAX_BYTES / sizeof(uint32_t)];

namespace {

constexpr int kNoallocMaxDevices = 64;
std::mutex g_noalloc_mu;
uint32_t* g_noalloc_dev_ptr[kNoallocMaxDevices] = {};  // device address of g_noalloc_bits on each GPU, once known

// `device` must be the current device: a __device__ array exists once per GPU and the runtime resolves the symbol
// for whichever one is current.
cudaError_t noalloc_bits_address(int device, uint32_t** out) {
  if (device < 0 || device >= kNoallocMaxDevices) return cudaErrorInvalidDevice;
  std::lock_guard<std::mutex> lock(g_noalloc_mu);
  if (!g_noalloc_dev_ptr[device]) {
    void* p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_bits);
    if (e != cudaSuccess) return e;
    g_noalloc_dev_ptr[device] = static_cast<uint32_t*>(p);
  }
  *out = g_noalloc_dev_ptr[device];
  return cudaSuccess;
}

// k_pack_bits for up to two grids in one launch: words [0, n0) pack src0, words [n0, n0 + n1) pack src1, each word
// written exactly once (so whatever an earlier call left in `bits` never reaches the host).
__global__ void __launch_bounds__(256) k_pack_bits_noalloc(const uint8_t* __restrict__ src0, size_t n0,
                                                           const uint8_t* __restrict__ src1, size_t n1,
                                                           uint32_t* __restrict__ bits) {
  const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n0 + n1) return;
  const uint4* p = t < n0 ? reinterpret_cast<const uint4*>(src0) + t * 2 : reinterpret_cast<const uint4*>(src1) + (t - n0) * 2;
  const uint4 a = __ldcs(p), b = __ldcs(p + 1);
  auto nib = [](uint32_t w) { return ((w & 0x01010101u) * 0x01020408u) >> 24; };
  bits[t] = nib(a.x) | (nib(a.y) << 4) | (nib(a.z) << 8) | (nib(a.w) << 12) | (nib(b.x) << 16) | (nib(b.y) << 20) |
            (nib(b.z) << 24) | (nib(b.w) << 28);
}

}  // namespace

extern "C" int sb_copy_grids_to_host_packed_noalloc(SbHandle h, uint8_t* host_ogm_bits, uint8_t* host_dagm_bits) {
  if (!h) return SB_EINVAL;
  if (h->cuda_err) return SB_ECUDA;
  const SbConfig& c = h->cfg;
  const size_t na = (size_t)c.n_envs * c.n_agents;
  const bool want0 = host_ogm_bits != nullptr, want1 = host_dagm_bits != nullptr;
  if (!want0 && !want1) return fail(h, SB_EINVAL, "sb_copy_grids_to_host_packed_noalloc: both destinations are NULL");
  if ((want0 && !c.ogm_enable) || (want1 && !c.dagm_enable))
    return fail(h, SB_EINVAL, "sb_copy_grids_to_host_packed_noalloc: a destination was given for a grid the config does not enable");
  const size_t px0 = want0 ? (size_t)c.ogm_width * c.ogm_height : 0, px1 = want1 ? (size_t)c.dagm_width * c.dagm_height : 0;
  if (px0 % 32 || px1 % 32)
    return fail(h, SB_EINVAL, "packed delivery needs width * height to be a multiple of 32");
  const size_t words0 = na * px0 / 32, words1 = na * px1 / 32, need = (words0 + words1) * 4;
  if (need > (size_t)SB_NOALLOC_BITS_MAX_BYTES)
    return fail(h, SB_EINVAL, "sb_copy_grids_to_host_packed_noalloc: the bit images need " + std::to_string(need) +
                              " bytes, the static staging array holds " + std::to_string((size_t)SB_NOALLOC_BITS_MAX_BYTES) +
                              " (SB_NOALLOC_BITS_MAX_BYTES); use sb_copy_obs_to_host_packed, which allocates");
  DeviceGuard dg(c.device);
  if (!dg.ok) return fail(h, SB_ECUDA, "cudaSetDevice failed");
  uint32_t* bits = nullptr;  // first call on this device: resolves the array's device address (before any capture)
  if (!cuda_ok(h, noalloc_bits_address(c.device, &bits), "cudaGetSymbolAddress(g_noalloc_bits)")) return SB_ECUDA;
  // the grids may have been written on another (non-blocking) stream: wait for it, as sb_copy_obs_to_host does
  wait_marked_work(h);
  h->launches++;
  k_pack_bits_noalloc<<<(unsigned)((words0 + words1 + 255) / 256), 256>>>(h->ob.ogm, words0, h->ob.dagm, words1, bits);
  if (!cuda_ok(h, cudaGetLastError(), "k_pack_bits_noalloc")) return SB_ECUDA;
  const uint8_t* src = reinterpret_cast<const uint8_t*>(bits);
  if (want0 && !cuda_ok(h, cudaMemcpyAsync(host_ogm_bits, src, words0 * 4, cudaMemcpyDeviceToHost, 0), "D2H ogm bits"))
    return SB_ECUDA;
  if (want1 && !cuda_ok(h, cudaMemcpyAsync(host_dagm_bits, src + words0 * 4, words1 * 4, cudaMemcpyDeviceToHost, 0), "D2H dagm bits"))
    return SB_ECUDA;
  return cuda_ok(h, cudaStreamSynchronize(0), "sb_copy_grids_to_host_packed_noalloc sync") ? SB_OK : SB_ECUDA;
}
