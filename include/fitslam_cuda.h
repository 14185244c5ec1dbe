/*
 * fitslam_cuda.h -- C ABI of the B200-native FIT-SLAM exploration hot path.
 *
 * This is the drop-in boundary.  Every entry point below
 * replaces one public method of the reference's `frontier_exploration` C++
 * classes; the header-only C++ classes under fit-slam_b200/cpp/frontier_exploration_cuda.hpp wrap these
 * calls behind the reference's own signatures.  Plain pointers and sizes only:
 * no STL, no torch types, no exceptions across the boundary.
 *
 * Reference paths are relative to /root/reference:
 *   DEP/  = dev_ws/src/DEPRECATED/frontier_exploration/frontier_exploration/
 *   ORBW/ = ORB-SLAM3-ROS2-Docker/orb_slam3_ros2_wrapper/
 *
 * Threading: one context = one CUDA stream.  Calls on one context are not
 * re-entrant (the reference is single-threaded per behaviour-tree tick,
 * DEP/src/ExplorationBT.cpp:166-197); different contexts are independent.
 *
 * There is NO CPU fallback: every compute entry point launches sm_100a kernels
 * and returns FITSLAM_E_CUDA if the device is unavailable.
 */
#ifndef FITSLAM_CUDA_H_
#define FITSLAM_CUDA_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(_WIN32)
#define FITSLAM_API
#else
#define FITSLAM_API __attribute__((visibility("default")))
#endif

#define FITSLAM_ABI_VERSION 2

typedef struct fitslam_ctx fitslam_ctx;

typedef enum fitslam_status {
  FITSLAM_OK = 0,
  FITSLAM_E_INVALID = 1,             /* bad argument / parameter out of the supported range     */
  FITSLAM_E_CUDA = 2,                /* CUDA runtime error; text in fitslam_last_error()         */
  FITSLAM_E_CAPACITY = 3,            /* caller buffer too small; required count written back     */
  FITSLAM_E_NO_COSTMAP = 4,          /* compute call before fitslam_set_costmap                  */
  FITSLAM_E_ROBOT_OUT_OF_BOUNDS = 5, /* DEP/src/FrontierSearch.cpp:28-32 (empty list + CRITICAL) */
  FITSLAM_E_NO_FRONTIERS = 6,        /* DEP/src/FrontierCostsManager.cpp:55-59 (returns false)   */
  FITSLAM_E_NO_POLYGON = 7,          /* DEP/src/FrontierCostsManager.cpp:61-65 (returns false)   */
  FITSLAM_E_ARRIVAL_UTILITY = 8,     /* "ARRIVAL UTILITY ERROR! MORE THAN 1" :149 (throws)       */
  FITSLAM_E_COST_OUT_OF_BOUNDS = 9,  /* "Cost out of bounds" :174 (throws)                       */
  FITSLAM_E_NO_LOOKUP_TABLE = 10,    /* "Cannot load lookup table..." FisherInfoManager.cpp:238  */
  FITSLAM_E_NO_MAP_POINTS = 11,      /* Fisher call before fitslam_set_map_points                */
  FITSLAM_E_IO = 12
} fitslam_status;

/* ---- parameters (defaults = DEP/params/exploration_params.yaml:50-87) ---- */
typedef struct fitslam_params {
  /* frontierSearch/ (DEP/src/FrontierSearch.cpp:12-18) */
  int32_t min_frontier_cluster_size; /* 1   */
  int32_t max_frontier_cluster_size; /* 20 (supported: 1..63) */
  double max_frontier_distance;      /* 50.0 m */
  int32_t lethal_threshold;          /* 160 */
  int32_t _pad0;
  /* costCalculator/ (DEP/src/CostCalculator.cpp:7-9,19) */
  double max_camera_depth; /* 2.0 m  */
  double delta_theta;      /* 0.10 rad */
  double camera_fov;       /* 1.04 rad */
  double robot_radius;     /* 0.60 m (Costmap2DROS::getRobotRadius) */
  double max_arrival_factor; /* 1.2  (DEP/src/CostCalculator.cpp:186) */
  double min_arrival_factor; /* 0.70 (:188) */
  /* frontierCostsManager/ (DEP/src/FrontierCostsManager.cpp:10-16) */
  double alpha;  /* 0.25 */
  double beta;   /* 1.0  */
  double vx_max; /* 0.5  */
  double wz_max; /* 0.5  */
  /* FullPathOptimizer constants (DEP/include/.../FullPathOptimizer.hpp:13,16) */
  int32_t num_frontiers_in_local_area; /* 5 */
  int32_t _pad1;
  double local_frontier_search_radius; /* 12.0 m */
  /* Fisher information (DEP/include/.../FisherInfoManager.hpp:23-30, FisherInfoManager.cpp:41,89) */
  float voxel_step;      /* step_max = 0.3f */
  float crowd_exponent;  /* 0.8f */
  double fisher_information_threshold; /* 550.0 */
  /* lookup-table bounds (DEP/src/fisher_information/GenerateLookupMain.cpp:8) */
  float lut_min_x, lut_max_x; /* 0, 21 */
  float lut_min_y, lut_max_y; /* -+ 8.5*1.732 */
  float lut_min_z, lut_max_z; /* -+ 8.5*1.732 */
  /* pinhole intrinsics + image bounds (ORBW/params/gazebo_rgbd.yaml:10-13, ORBW/src/orb_slam3_interface.cpp:218-221) */
  float fx, fy, cx, cy;       /* 432.496042035043 x2, 320, 240 */
  float img_max_u, img_max_v; /* 640, 480 */
  /* z of the candidate poses built from frontier goals (0.3) */
  double candidate_pose_z;
} fitslam_params;

/* One frontier as produced by FrontierSearch::searchFrom (DEP/src/FrontierSearch.cpp:21-96,
 * Frontier fields DEP/include/frontier_exploration/Frontier.hpp:39-136). */
typedef struct fitslam_frontier {
  double goal_x, goal_y;    /* goal point = centre of the median cell (FrontierSearch.cpp:161-165) */
  uint32_t goal_mx, goal_my;/* the same cell in map coordinates */
  int32_t size;             /* Frontier::getSize() */
  uint32_t component;       /* ordinal of the 8-connected frontier component, in discovery order */
  uint32_t chunk;           /* ordinal of this chunk inside its component */
  uint32_t first_cell;      /* offset of this chunk's first cell in the every_frontier cell list */
} fitslam_frontier;

/* Output of FrontierCostCalculator::setArrivalInformationForFrontier (DEP/src/CostCalculator.cpp:23-121). */
typedef struct fitslam_arrival {
  int32_t information;    /* maxValue: best window sum of unknown cells (:93-102,112)        */
  int32_t max_index;      /* maxIndex: first window attaining it                              */
  double theta_s_star;    /* maxIndex*delta_theta + camera_fov/2 (:119); 0.0 on a failed ray  */
  uint8_t achievable;     /* Frontier::isAchievable() after the call                          */
  uint8_t footprint_in_lethal; /* isRobotFootprintInLethal (:77)                              */
  uint8_t ray_failed;     /* getTracedCells returned false (:50-55)                           */
  uint8_t _pad[5];
} fitslam_arrival;

/* A pose in the ROS map frame (geometry_msgs::msg::Pose). */
typedef struct fitslam_pose {
  double x, y, z;
  double qx, qy, qz, qw;
} fitslam_pose;

/* Result of FullPathOptimizer::getFilteredFrontiersN (DEP/src/FullPathOptimizer.cpp:157-227).
 * `sorted` (caller array, capacity >= F) receives the indices of the eligible frontiers in
 * ascending path_length_m order.  local_frontiers = sorted[0 .. n_local),
 * global_frontiers = sorted[global_begin .. global_begin + n_global). */
typedef struct fitslam_selection {
  uint32_t n_sorted;
  uint32_t n_local;
  uint32_t global_begin;
  uint32_t n_global;
  int32_t closest_global; /* frontier index of closest_global_frontier, -1 if none */
  int32_t _pad;
} fitslam_selection;

/* Whole-cycle summary (fitslam_cycle). */
typedef struct fitslam_cycle_result {
  uint32_t n_frontiers;
  uint32_t n_frontier_cells;
  uint32_t n_components;
  uint32_t n_candidates;   /* frontier x yaw-window candidates scored */
  uint32_t n_fisher_poses; /* poses scored by the Fisher stage */
  uint32_t bfs_levels;
  uint32_t ambiguous_sorts;/* chunks whose median sort was redone on the host with libm's atan2 (an angle within 0.03 ulp of a
                              rounding boundary and 2 ulp of another angle or of a comparator threshold) */
  uint32_t _pad;
  fitslam_selection selection;
  float stage_ms[8];       /* extract, arrival, distance, fisher, normalise+select, cycle_begin total, cycle_end total, outer-wavefront kernel alone */
  uint32_t widest_level;   /* cells in the widest level of the outer wavefront */
  uint32_t wavefront_ctas; /* CTAs the outer wavefront ran on: 1, or the size of the thread-block cluster */
} fitslam_cycle_result;

enum { FITSLAM_DIST_GIVEN = 0, FITSLAM_DIST_EUCLIDEAN = 1 };
enum { FITSLAM_FISHER_OFF = 0, FITSLAM_FISHER_BEST_YAW = 1, FITSLAM_FISHER_ALL_YAWS = 2 };

/* ------------------------------------------------------------------ lifecycle */
FITSLAM_API int fitslam_abi_version(void);
FITSLAM_API void fitslam_default_params(fitslam_params *p);
FITSLAM_API fitslam_status fitslam_create(const fitslam_params *p, int device, fitslam_ctx **out);
FITSLAM_API void fitslam_destroy(fitslam_ctx *ctx);
FITSLAM_API const char *fitslam_last_error(const fitslam_ctx *ctx);
FITSLAM_API const char *fitslam_status_string(fitslam_status s);
/* Candidate sharding (8e).  In a cycle this rank scores the frontiers f with f % nranks == rank (known the moment a
 * frontier exists, so the scoring overlaps the search); the stand-alone scoring calls on a caller's list take the
 * contiguous block [rank * B, min(F, (rank + 1) * B)), B = ceil(F / nranks).  Records of the other frontiers are left
 * zero in the output buffers.  fitslam_comm_init sets the shard itself. */
FITSLAM_API fitslam_status fitslam_set_shard(fitslam_ctx *ctx, int rank, int nranks);
/* The CUDA stream all work of this context is enqueued on (a cudaStream_t). */
FITSLAM_API void *fitslam_stream(fitslam_ctx *ctx);
/* Number of kernels this context has launched so far (bench.py's gpu_launches). */
FITSLAM_API uint64_t fitslam_launch_count(const fitslam_ctx *ctx);

/* ------------------------------------------------------------------ inputs */
/* nav2_costmap_2d::Costmap2D snapshot (getCharMap/getSizeInCellsX/getResolution/getOriginX; T1).
 * `data` may be a host or a device pointer (detected).  dirty_xywh = {x0,y0,w,h} restricts the
 * upload to a sub-rectangle of an already-resident map of the same geometry (rapid mode), NULL = full. */
FITSLAM_API fitslam_status fitslam_set_costmap(fitslam_ctx *ctx, const uint8_t *data, uint32_t size_x,
                                               uint32_t size_y, double resolution, double origin_x,
                                               double origin_y, const uint32_t *dirty_xywh);
/* ORB-SLAM3 map points, ROS map frame, xyz float32 triples (slam_msgs/MapPoint.position; T7). */
FITSLAM_API fitslam_status fitslam_set_map_points(fitslam_ctx *ctx, const float *xyz, uint64_t n_points);

/* ------------------------------------------------------------------ frontier extraction */
/* FrontierSearch::reset(); searchFrom(position); getAllFrontiers()
 * (DEP/src/FrontierSearch.cpp:21-96,251-254; FrontierSearch.hpp:39-52).
 * *n_frontiers in: capacity of `out`; out: number found.  every_cells (optional) receives the map index
 * of every frontier cell in discovery order (getAllFrontiers); *n_cells in/out likewise.  */
FITSLAM_API fitslam_status fitslam_search_frontiers(fitslam_ctx *ctx, double robot_x, double robot_y,
                                                    fitslam_frontier *out, uint32_t *n_frontiers,
                                                    uint32_t *every_cells, uint32_t *n_cells);
FITSLAM_API fitslam_status fitslam_set_search_distance(fitslam_ctx *ctx, double max_frontier_distance);

/* ------------------------------------------------------------------ costmap ingest */
/* nav_msgs/OccupancyGrid (int8: -1 unknown, 0..100) -> the uint8 costmap the search reads, on the device:
 * (1) TraversabilityThresholdNode::callback (ORB-SLAM3-ROS2-Docker/traversability_mapping/traversability_ros_interface/
 *     traversability_mapping_ros/src/threshold_occupancy.cpp:25-49): value >= lethal_obstacle_threshold -> 100,
 *     -1 -> -1, else 0 (skipped when lethal_obstacle_threshold < 0);
 * (2) nav2_costmap_2d::StaticLayer::interpretValue (ROS 2 Humble binary package, not under the reference tree; the
 *     reference configures it with track_unknown_space: true and the defaults lethal_cost_threshold 100,
 *     unknown_cost_value -1, trinary_costmap true: DEP/params/exploration_params.yaml:25-38): unknown -> 255 (or 0
 *     without track_unknown_space), >= lethal_cost_threshold -> 254, trinary -> 0, else
 *     (unsigned char)((double)value / lethal_cost_threshold * 254).
 * Same geometry / dirty-rectangle rules as fitslam_set_costmap.  `data`: host or device, size_x * size_y int8. */
typedef struct fitslam_occupancy_params {
  int32_t lethal_obstacle_threshold; /* threshold node: 85 (threshold_occupancy.cpp:17); < 0 = no threshold node */
  int32_t lethal_cost_threshold;     /* static layer: 100 */
  int32_t unknown_cost_value;        /* static layer: -1 */
  int32_t trinary_costmap;           /* static layer: 1 */
  int32_t track_unknown_space;       /* costmap: 1 (exploration_params.yaml:25) */
} fitslam_occupancy_params;
FITSLAM_API void fitslam_default_occupancy_params(fitslam_occupancy_params *p);
FITSLAM_API fitslam_status fitslam_set_occupancy_grid(fitslam_ctx *ctx, const int8_t *data, uint32_t size_x,
                                                      uint32_t size_y, double resolution, double origin_x,
                                                      double origin_y, const fitslam_occupancy_params *p,
                                                      const uint32_t *dirty_xywh);

/* Copy-out of the resident costmap (Costmap2D::getCharMap: size_x * size_y bytes, row-major, no padding); `out`: host or
 * device, `capacity` bytes. */
FITSLAM_API fitslam_status fitslam_get_costmap(fitslam_ctx *ctx, uint8_t *out, uint64_t capacity);

/* ------------------------------------------------------------------ lethal marker (stamping half) */
/* LethalMarker (DEP/src/nav2_plugins/lethal_marker.cpp; the same layer in FS2/fit_slam2_nav2_plugins/plugins/
 * keepout_layer.cpp): each marked area is a fan of Bresenham rays (rayTraceGeneric, lethal_marker.cpp:12-41) from the
 * centre / apex cell; LethalMarker::updateCosts -> markCells writes 253 into every traced cell of the master grid
 * (:242-248, :303-327).
 *   kind 0 = addNewMarkedArea (:219-228): 360 rays to the circle of (unsigned)(extent / resolution) cells
 *            (getPointsInSemiCircle with robot_yaw 0, :52-73);
 *   kind 1 = addNewMarkedAreaFOV (:230-239): 20 rays from the apex to samples along the base of an isosceles
 *            triangle of height (unsigned)(extent / resolution) cells, apex angle 45 degrees, pointing along `yaw`
 *            (getPointsInIsoscelesTriangle, :75-127).
 * Ray end points are clamped to the map (cellEnforceBoundaries, :5-11).  An area whose centre is off the map is
 * skipped, as in the reference.  The stamps go into the resident costmap (see fitslam_get_costmap); n_marked
 * (optional) = areas that were on the map. */
typedef struct fitslam_lethal_area {
  double wx, wy;  /* centre (kind 0) / apex (kind 1), world frame */
  double extent;  /* radius (kind 0) / triangle height (kind 1), metres */
  double yaw;     /* kind 1: direction from the apex toward the base centre, radians */
  int32_t kind;
  int32_t reserved;
} fitslam_lethal_area;
FITSLAM_API fitslam_status fitslam_mark_lethal(fitslam_ctx *ctx, const fitslam_lethal_area *areas, uint32_t n,
                                               uint32_t *n_marked);

/* ------------------------------------------------------------------ sensor simulator */
/* SensorSimulator::timerCallback ray loop + markRay (RE/src/SensorSimulator.cpp:193-202, 211-246): for
 * a = min_angle; a <= max_angle; a += angular_resolution, a ray from the robot position along base_yaw + a in steps
 * of `resolution` up to max_range copies latest[idx] into explored[idx] cell by cell and stops behind the first cell
 * whose value is > 80 or at the map border.  latest / explored: size_x * size_y int8 (nav_msgs/OccupancyGrid data),
 * both host or both device; explored is updated in place.  n_rays (optional) = rays cast.  `resolution` is narrowed to float32
 * first: that is what the message carries (nav_msgs/MapMetaData.resolution) and what the reference widens again (:217). */
typedef struct fitslam_sensor_params {
  double sensor_min_angle;   /* -pi/4 (RE/src/Parameters.cpp:233-234) */
  double sensor_max_angle;   /*  pi/4 (:236-237) */
  double angular_resolution; /* 0.013 (:227-228) */
  double sensor_max_range;   /* 3.0   (:239-240) */
} fitslam_sensor_params;
FITSLAM_API void fitslam_default_sensor_params(fitslam_sensor_params *p);
FITSLAM_API fitslam_status fitslam_sensor_mark_rays(fitslam_ctx *ctx, const int8_t *latest, int8_t *explored,
                                                    uint32_t size_x, uint32_t size_y, double resolution,
                                                    double origin_x, double origin_y, double robot_x, double robot_y,
                                                    double base_yaw, const fitslam_sensor_params *p,
                                                    uint32_t *n_rays);

/* ------------------------------------------------------------------ roadmap edges */
/* FrontierRoadMap::isConnectable for a batch of node pairs (RE/src/planners/FrontierRoadmap.cpp:671-694): one ray
 * from (x0, y0) to (x1, y1) per pair through getTracedCells (RE/src/Helpers.cpp:30-91; at most
 * (unsigned)(1.5 * radius_to_decide_edges / resolution) steps), obstacle band 253..254, unknown = 255.
 * connectable[i] = 1 iff both ends are on the map, no cell of the ray is in the obstacle band, and the ray crosses at
 * most radius_to_decide_edges / resolution * 0.3 unknown cells.  segments: n x 4 doubles {x0, y0, x1, y1} (world
 * frame, host or device); num_unknown (optional, n int32) exposes getNumUnknown() for stage-wise parity (-1 = an end
 * is off the map). */
FITSLAM_API fitslam_status fitslam_roadmap_connectable(fitslam_ctx *ctx, const double *segments, uint32_t n,
                                                       double radius_to_decide_edges, uint8_t *connectable,
                                                       int32_t *num_unknown);

/* ------------------------------------------------------------------ arrival information */
/* FrontierCostCalculator::setMaxArrivalInformation (DEP/src/CostCalculator.cpp:123-191). */
FITSLAM_API fitslam_status fitslam_arrival_limits(fitslam_ctx *ctx, double *max_arrival_info_gt,
                                                  double *min_arrival_info_gt, int32_t *max_value);
/* setArrivalInformationForFrontier for a list of frontiers (DEP/src/CostCalculator.cpp:23-121).
 * frontiers == NULL scores the list left on the device by the last fitslam_search_frontiers.
 * info_along_ray (optional, F x n_rays uint16) and window_sums (optional, F x n_windows int32) expose
 * the intermediate arrays `information_along_ray` / `result` of the reference for stage-wise parity. */
FITSLAM_API fitslam_status fitslam_score_arrival(fitslam_ctx *ctx, const fitslam_frontier *frontiers,
                                                 uint32_t n_frontiers, const double poly_xy_min_max[4],
                                                 fitslam_arrival *out, uint16_t *info_along_ray,
                                                 int32_t *window_sums);
/* Ray fan geometry implied by the parameters: number of rays (theta loop :36), window k (:87), windows. */
FITSLAM_API fitslam_status fitslam_ray_shape(const fitslam_ctx *ctx, uint32_t *n_rays, uint32_t *k,
                                             uint32_t *n_windows, uint32_t *max_length);

/* ------------------------------------------------------------------ Fisher information */
/* FisherInformationManager::generateLookupTable / loadLookupTable
 * (DEP/src/fisher_information/FisherInfoManager.cpp:117-262).  path may be NULL (memory only). */
FITSLAM_API fitslam_status fitslam_fisher_generate_lookup(fitslam_ctx *ctx, const char *save_path);
FITSLAM_API fitslam_status fitslam_fisher_load_lookup(fitslam_ctx *ctx, const char *path);
/* Dense copy of the table: out[(ix*ny + iy)*nz + iz] (NaN = no entry); dims and index origin returned. */
FITSLAM_API fitslam_status fitslam_fisher_lookup_dims(const fitslam_ctx *ctx, int32_t dims[3], int32_t origin[3]);
FITSLAM_API fitslam_status fitslam_fisher_lookup_copy(fitslam_ctx *ctx, float *out, uint64_t capacity);
/* FisherInformationManager::isPoseSafe for a batch of poses (FisherInfoManager.cpp:39-115) with the
 * landmark service replaced by the resident map points culled as ORBW/src/orb_slam3_interface.cpp:188-317
 * (exhaustive search) and transformed as ORBW/src/rgbd/rgbd-slam-node.cpp:414-429.
 * n_visible[i] = landmarks in view of pose i that hit a lookup-table entry (the ones that contribute; the
 * reference has no far plane, landmarks beyond the table domain return NaN and are skipped, :90-94). */
FITSLAM_API fitslam_status fitslam_score_fisher(fitslam_ctx *ctx, const fitslam_pose *poses, uint32_t n_poses,
                                                float *information, uint8_t *safe, uint32_t *n_visible);

/* The landmark service behind isPoseSafe (exhaustive mode): the `map_points` of a
 * slam_msgs/srv/GetLandmarksInView response for one pose, i.e. ORBSLAM3Interface::mapPointsVisibleFromPose
 * (ORBW/src/orb_slam3_interface.cpp:188-317, exhaustive_search = true) over the resident map points, each visible
 * one transformed into the pose frame as ORBW/src/rgbd/rgbd-slam-node.cpp:406-429 does, in map-point order.
 * out_xyz: capacity x 3 floats (host or device; the wire widens them to float64); *count = visible landmarks, also
 * when it exceeds capacity (FITSLAM_E_CAPACITY, the first `capacity` are written).  The keyframe-structured
 * non-exhaustive mode (:228-272) is not built.
 * NOT the service's response on a multi-keyframe atlas: the service stops adding keyframes once more than maxLandmarks = 1000
 * landmarks are collected (:270-271, also in exhaustive mode; rgbd-slam-node.cpp:413).  That cutoff needs the atlas's
 * keyframe-to-point lists: this call treats the resident cloud as ONE keyframe, for which the reference's loop returns every
 * visible point too.  The same holds for fitslam_score_fisher and the cycle's Fisher stage: their totals are upper bounds of
 * what isPoseSafe would sum behind the real service.  fitslam_landmarks_in_view_keyframes / fitslam_score_fisher_keyframes
 * below take the keyframe lists and reproduce the cutoff. */
FITSLAM_API fitslam_status fitslam_landmarks_in_view(fitslam_ctx *ctx, const fitslam_pose *pose, float *out_xyz,
                                                     uint32_t capacity, uint32_t *count);

/* The keyframe-structured query: ORBSLAM3Interface::mapPointsVisibleFromPose as the service really runs it
 * (ORBW/src/orb_slam3_interface.cpp:188-317), over an atlas given as keyframes with their map-point lists.
 *   fitslam_set_keyframes   the atlas: n_kf keyframes in the order the service iterates them (the reference walks a
 *                           std::vector made from a std::set<KeyFrame*>, i.e. pointer order, which no caller can know: the order
 *                           given here IS the iteration order), each with GetPoseInverse() (Twc, ORB world frame) and the range
 *                           [first, first + count) of `point_index` that lists its map points (indices into the points of
 *                           fitslam_set_map_points; a point may appear in several keyframes and is taken once, :263-265).
 *   max_landmarks           the service passes 1000 (ORBW/src/rgbd/rgbd-slam-node.cpp:413).  A keyframe is skipped with everything
 *                           after it once MORE than max_landmarks landmarks are collected (:270-271; the check sits outside the
 *                           exhaustive_search test, so it applies in both modes).
 *   exhaustive_search = 0   the per-keyframe prefilter of :240-267 (distance to the keyframe against max_distance, behind /
 *                           in front, relative yaw) runs first.  max_angle is accepted and unused, as in the reference.
 * out_xyz / count as fitslam_landmarks_in_view (pose frame, service order: keyframe by keyframe, list order inside a keyframe);
 * n_keyframes_checked = numKFsChecked.  The prefilter multiplies 3x3 float matrices where Sophus multiplies unit quaternions:
 * a keyframe within a few ulp of one of its thresholds may fall on the other side. */
typedef struct fitslam_keyframe {
  float Rwc[9];   /* KeyFrame::GetPoseInverse() rotation, row-major */
  float twc[3];   /* ... translation */
  uint32_t first; /* its map points: point_index[first .. first + count) */
  uint32_t count;
} fitslam_keyframe;
FITSLAM_API fitslam_status fitslam_set_keyframes(fitslam_ctx *ctx, const fitslam_keyframe *keyframes, uint32_t n_keyframes,
                                                 const uint32_t *point_index, uint64_t n_entries);
FITSLAM_API fitslam_status fitslam_landmarks_in_view_keyframes(fitslam_ctx *ctx, const fitslam_pose *pose, int32_t max_landmarks,
                                                               float max_distance, float max_angle, int exhaustive_search,
                                                               float *out_xyz, uint32_t capacity, uint32_t *count,
                                                               uint32_t *n_keyframes_checked);
/* FisherInformationManager::isPoseSafe (DEP/src/fisher_information/FisherInfoManager.cpp:39-115) behind that service: the
 * landmarks of fitslam_landmarks_in_view_keyframes (max_landmarks = 1000 and the request's max_dist_pose_observation /
 * exhaustive_search in the reference) summed through the lookup table with the per-voxel crowding factor.
 * n_landmarks = landmarks the service returned, n_used = those that hit a lookup-table entry. */
FITSLAM_API fitslam_status fitslam_score_fisher_keyframes(fitslam_ctx *ctx, const fitslam_pose *pose, int32_t max_landmarks,
                                                          float max_distance, int exhaustive_search, float *information,
                                                          uint8_t *safe, uint32_t *n_landmarks, uint32_t *n_used);

/* The same query without any allocation.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * It calls no cudaMalloc / cudaFree / allocating routine, stages nothing and does not synchronise: three kernels and one
 * 4-byte device-to-device copy are enqueued on the context stream (fitslam_stream) and the call returns.
 *   d_xyz     n_points x 3 floats, DEVICE memory, the caller's map points (the resident points of fitslam_set_map_points are
 *             not used; n_points <= FITSLAM_LIV_STATIC_MAX_POINTS, else FITSLAM_E_CAPACITY and nothing is enqueued)
 *   d_pose    one fitslam_pose in DEVICE memory
 *   d_out_xyz capacity x 3 floats, DEVICE memory: the first min(count, capacity) visible landmarks, as fitslam_landmarks_in_view
 *   d_count   one uint32 in DEVICE memory: the number of visible landmarks (may exceed capacity; the caller compares)
 * The kernels' scratch (one count per 1024 points + the total, 16 KB) is a fixed device array inside the library, one per GPU
 * that loads it, held until the process exits.  Every call on a GPU shares it and finds the last call's contents there; a
 * call overwrites every word before it reads it, so nothing needs clearing, but two calls that overlap on the GPU (different
 * contexts / streams of one device) race: keep them ordered. */
#define FITSLAM_LIV_STATIC_MAX_POINTS 4194304ull
FITSLAM_API fitslam_status fitslam_landmarks_in_view_static(fitslam_ctx *ctx, const float *d_xyz, uint64_t n_points,
                                                            const fitslam_pose *d_pose, float *d_out_xyz, uint32_t capacity,
                                                            uint32_t *d_count);

/* ------------------------------------------------------------------ normalisation + selection */
/* FrontierCostsManager::assignCosts U1 block (DEP/src/FrontierCostsManager.cpp:121-205) followed by
 * FullPathOptimizer::getFilteredFrontiersN (DEP/src/FullPathOptimizer.cpp:157-227).
 * arrival/achievable come from the last fitslam_score_arrival.  path_length_m / path_heading: one per
 * frontier (inputs produced by the planner in the reference; DBL_MAX = the planner found the frontier not achievable,
 * DEP/src/CostCalculator.cpp:449-466); blacklisted may be NULL. */
FITSLAM_API fitslam_status fitslam_normalise_select(fitslam_ctx *ctx, const double *path_length_m,
                                                    const double *path_heading, const uint8_t *blacklisted,
                                                    uint32_t n_frontiers, double *weighted_cost,
                                                    double *arrival_utility, double *distance_utility,
                                                    uint8_t *achievable, uint32_t *sorted,
                                                    fitslam_selection *sel);

/* ------------------------------------------------------------------ one whole exploration cycle */
/* searchFrom -> setArrivalInformationForFrontier -> setPlanForFrontierEuclidean -> (Fisher) -> U1 ->
 * getFilteredFrontiersN, enqueued back to back on the context stream with a single host sync at the end.
 * robot_yaw feeds path_heading (DEP/src/CostCalculator.cpp:467-477). */
FITSLAM_API fitslam_status fitslam_cycle(fitslam_ctx *ctx, double robot_x, double robot_y, double robot_yaw,
                                         const double poly_xy_min_max[4], int fisher_mode,
                                         fitslam_cycle_result *result);
/* The same cycle in two halves, for candidate sharding across GPUs (8e).  cycle_begin: searchFrom (replicated on
 * every rank), with arrival information and Fisher information for this rank's share of the frontiers (frontier f
 * belongs to rank f % nranks, fitslam_set_shard) scored while the search runs; the 16-byte score records are left on the
 * device (fitslam_score_records): one row per frontier, zero where another rank scores.  The caller sums them over the
 * ranks as int32 (NCCL all-reduce), then cycle_end unpacks the merged records and runs the global steps:
 * setPlanForFrontierEuclidean, U1, getFilteredFrontiersN.  With one rank, cycle == begin + end. */
FITSLAM_API fitslam_status fitslam_cycle_begin(fitslam_ctx *ctx, double robot_x, double robot_y, double robot_yaw,
                                               const double poly_xy_min_max[4], int fisher_mode, uint32_t *n_frontiers);
FITSLAM_API fitslam_status fitslam_cycle_end(fitslam_ctx *ctx, fitslam_cycle_result *result);
/* Copy-out of the per-frontier records of the last cycle (any pointer may be NULL). */
FITSLAM_API fitslam_status fitslam_cycle_fetch(fitslam_ctx *ctx, fitslam_frontier *frontiers,
                                               fitslam_arrival *arrival, double *path_length_m,
                                               double *path_heading, double *weighted_cost, float *fisher_info,
                                               uint32_t *sorted, uint32_t capacity);
/* Copy-out of the per-CANDIDATE Fisher results of the last cycle: one entry per candidate pose, pose index =
 * frontier * poses_per_frontier + window (FITSLAM_FISHER_ALL_YAWS: every yaw window, poses_per_frontier = n_windows;
 * FITSLAM_FISHER_BEST_YAW: one pose per frontier).  Each entry is what FisherInformationManager::isPoseSafe returns for
 * that pose (DEP/src/fisher_information/FisherInfoManager.cpp:39-115): `information`, the bool, and the number of
 * landmarks that contributed.  Any pointer may be NULL; capacity in poses; *n_poses receives the count.  With several
 * ranks the entries of frontiers another rank scored are zero (the merged records only carry the best yaw). */
FITSLAM_API fitslam_status fitslam_cycle_fetch_candidates(fitslam_ctx *ctx, float *information, uint8_t *safe,
                                                          uint32_t *n_visible, uint32_t capacity, uint32_t *n_poses);

/* ------------------------------------------------------------------ multi-GPU plumbing (8e) */
/* Device address + byte size of the packed per-frontier score records of the last scoring call, laid out
 * for an in-place all-gather: nranks equal blocks of B = ceil(F / nranks) records, rank r's frontiers
 * [r * B, (r + 1) * B) in block r, so record f lives at index f.  16 bytes per frontier:
 * {int32 information, int32 max_index, float fisher, uint32 flags (bit0 achievable, bit1 footprint in lethal,
 * bit2 ray failed, bit3 fisher valid, bit4 fisher safe)}. */
FITSLAM_API fitslam_status fitslam_score_records(fitslam_ctx *ctx, void **device_ptr, uint64_t *bytes);
/* One NCCL communicator per context, one process per GPU (the reference is one CPU thread and has no
 * analogue).  fitslam_comm_unique_id: 128 bytes (an ncclUniqueId) made on rank 0 and carried to the other ranks by the
 * host's own means.  fitslam_comm_init joins the communicator and sets the shard (rank, nranks).  From then on
 *   - fitslam_set_costmap / fitslam_set_map_points are collective: rank 0 passes the data, the other ranks pass NULL (with
 *     the same geometry / count) and receive it by ncclBroadcast over NVLink (a dirty rectangle broadcasts its rows only);
 *   - fitslam_cycle merges the ranks' score records with an ncclAllReduce (int32 sum: one owner per row, zeros elsewhere)
 *     on the context stream between its two halves; a caller that drives cycle_begin / cycle_end itself calls
 *     fitslam_cycle_merge in between.
 * libnccl.so.2 is loaded at run time by fitslam_comm_unique_id / fitslam_comm_init; a single-GPU host never needs it. */
FITSLAM_API fitslam_status fitslam_comm_unique_id(void *id, uint64_t bytes);
FITSLAM_API fitslam_status fitslam_comm_init(fitslam_ctx *ctx, const void *nccl_unique_id, int rank, int nranks);
FITSLAM_API fitslam_status fitslam_comm_destroy(fitslam_ctx *ctx);
FITSLAM_API fitslam_status fitslam_cycle_merge(fitslam_ctx *ctx);

#ifdef __cplusplus
}
#endif
#endif /* FITSLAM_CUDA_H_ */
