// fisher.cu -- stage F: Fisher information of the ORB-SLAM3 map points for a batch of candidate poses.
//
// Replaces FisherInformationManager::{generateLookupTable, loadLookupTable, getVoxelCoordinate,
// getInformationFromLookup, isPoseSafe} (DEP/src/fisher_information/FisherInfoManager.cpp:39-324,
// DEP/include/frontier_exploration/fisher_information/FisherInfoManager.hpp:23-123), the Fisher helpers
// computeJacobianForPointLocal / computeFIM / computeInformationOfPointLocal (DEP/src/Helpers.cpp:397-449) and the
// server side of the landmark service: ORBSLAM3Interface::mapPointsVisibleFromPose
// (ORBW/src/orb_slam3_interface.cpp:188-317, exhaustive search), RgbdSlamNode::getMapPointsInViewServer
// (ORBW/src/rgbd/rgbd-slam-node.cpp:406-429), se3ROSToORB (ORBW/src/type_conversion.cpp:74-100) and
// Pinhole::project (ORB/src/CameraModels/Pinhole.cpp:43-49).
//
// Layout.  The lookup table is a dense float grid indexed by the integer voxel index round(x / step) per axis
// (NaN = no entry).  The map points are bucketed once per map update into a uniform xy bin grid (counting sort,
// float4 records).  One CTA scores one pose: it culls the bins around the pose against the frustum and the lookup
// domain (conservative plane tests), runs the reference's per-point float arithmetic on the surviving points,
// counts the points per camera-frame voxel in a shared-memory hash table (the crowding rank of
// getInformationFromLookup is order-free per voxel: information * S(count), S a host-tabulated prefix sum of
// exp(1 - j^0.8)) and reduces the voxel terms with warp shuffles.  Float arithmetic follows Eigen's evaluation
// order (x0*y0 + (x1*y1 + x2*y2)); this file is compiled with -fmad=false so it rounds like the host.
#include "ctx.h"

#include <algorithm>
#include <climits>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <limits>
#include <mutex>
#include <string>
#include <vector>

namespace fitslam {

static constexpr int CROWD_MAX = 512;       // exp(1 - j^0.8) underflows to 0 in float long before j = 512
// Per-CTA voxel bookkeeping, both in dynamic shared memory.  A bitmap over the part of the lookup domain a visible point can
// fall in (per x-slab of voxels, the z-rows the image height admits: `XRow`) says whether a voxel has its first landmark; the
// second and later landmarks of a voxel (about 1 in 8 at C2) and the points outside the bitmap's rows are counted in a
// small hash table of packed words, (voxel + 1 | outside << 21) << CNT_BITS | count.
static constexpr int BM_WORDS = 10240;      // 327 680 bits, 40 KB
static constexpr int SEC_SLOTS = 8192;      // 32 KB
static constexpr int CNT_BITS = 10;         // low bits of a slot: landmarks counted in it, saturating (>= CROWD_MAX is enough)
static constexpr uint32_t CNT_MASK = (1u << CNT_BITS) - 1u;
static constexpr uint32_t OUTSIDE_BIT = 1u << 21;  // key space of the voxels the bitmap has no bit for
static constexpr int MAX_XROWS = 256;       // x-slabs of the lookup domain the per-slab table holds (more: no bitmap at all)
static constexpr int FISHER_DYN_SMEM = (BM_WORDS + SEC_SLOTS) * 4;
static constexpr int MAX_BINS = 1024;       // bins one pose can touch (bounding square of the lookup domain)
// 768 threads x 2 CTAs per SM = 48 warps (40 registers; the only spills are in thread 0's build_pose_frame).  The point
// loop is latency-bound (one L2 load and one shared-memory CAS chain per iteration), so warps in flight are what pays:
// measured on C2, 256 x 3: 6.5 ms, 512 x 2: 5.4 ms, 1024 x 1: 6.4 ms, 768 x 2: 5.1 ms, 1024 x 2 (32 registers): 5.1 ms.
#ifndef FS_FISHER_THREADS
#define FS_FISHER_THREADS 768
#endif
#ifndef FS_FISHER_MINB
#define FS_FISHER_MINB 2
#endif
#ifndef FS_FISHER_BOUND   // the register budget is that of this many threads x FS_FISHER_MINB CTAs, whatever the CTA launched
#define FS_FISHER_BOUND (FS_FISHER_THREADS > 768 ? FS_FISHER_THREADS : 768)
#endif
static constexpr int FISHER_THREADS = FS_FISHER_THREADS;
#ifndef FS_CHUNK_PTS
#define FS_CHUNK_PTS 64
#endif
static constexpr int CHUNK_PTS = FS_CHUNK_PTS;  // points per unit of work
static constexpr int MAX_CHUNKS = 768 * 64 / CHUNK_PTS;  // 6 KB of static shared memory at 64 points per chunk

struct M3f { float m[3][3]; };
struct V3f { float v[3]; };

__host__ __device__ __forceinline__ float dot3(float a0, float b0, float a1, float b1, float a2, float b2) {
  return a0 * b0 + (a1 * b1 + a2 * b2);
}
__host__ __device__ inline V3f mulv(const M3f &A, const V3f &x) {
  V3f r;
  for (int i = 0; i < 3; ++i) r.v[i] = dot3(A.m[i][0], x.v[0], A.m[i][1], x.v[1], A.m[i][2], x.v[2]);
  return r;
}
__host__ __device__ inline M3f mulm(const M3f &A, const M3f &B) {
  M3f r;
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) r.m[i][j] = dot3(A.m[i][0], B.m[0][j], A.m[i][1], B.m[1][j], A.m[i][2], B.m[2][j]);
  return r;
}
__host__ __device__ inline M3f transp(const M3f &A) {
  M3f r;
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) r.m[i][j] = A.m[j][i];
  return r;
}
// Eigen::Quaternionf::toRotationMatrix
__host__ __device__ inline M3f quat_to_matrix(float w, float x, float y, float z) {
  M3f r;
  const float tx = 2.0f * x, ty = 2.0f * y, tz = 2.0f * z;
  const float twx = tx * w, twy = ty * w, twz = tz * w;
  const float txx = tx * x, txy = ty * x, txz = tz * x;
  const float tyy = ty * y, tyz = tz * y, tzz = tz * z;
  r.m[0][0] = 1.0f - (tyy + tzz); r.m[0][1] = txy - twz;          r.m[0][2] = txz + twy;
  r.m[1][0] = txy + twz;          r.m[1][1] = 1.0f - (txx + tzz); r.m[1][2] = tyz - twx;
  r.m[2][0] = txz - twy;          r.m[2][1] = tyz + twx;          r.m[2][2] = 1.0f - (txx + tyy);
  return r;
}
__host__ __device__ inline float cofac(const M3f &m, int i, int j) {
  const int i1 = (i + 1) % 3, i2 = (i + 2) % 3, j1 = (j + 1) % 3, j2 = (j + 2) % 3;
  return m.m[i1][j1] * m.m[i2][j2] - m.m[i1][j2] * m.m[i2][j1];
}
// Eigen compute_inverse<Matrix3f>: cofactors, determinant from column 0, scaled by 1/det
__host__ __device__ inline M3f inverse3(const M3f &m) {
  const float c0 = cofac(m, 0, 0), c1 = cofac(m, 1, 0), c2 = cofac(m, 2, 0);
  const float det = dot3(c0, m.m[0][0], c1, m.m[1][0], c2, m.m[2][0]);
  const float invdet = 1.0f / det;
  M3f r;
  r.m[0][0] = c0 * invdet; r.m[0][1] = c1 * invdet; r.m[0][2] = c2 * invdet;
  r.m[1][0] = cofac(m, 0, 1) * invdet; r.m[1][1] = cofac(m, 1, 1) * invdet; r.m[1][2] = cofac(m, 2, 1) * invdet;
  r.m[2][0] = cofac(m, 0, 2) * invdet; r.m[2][1] = cofac(m, 1, 2) * invdet; r.m[2][2] = cofac(m, 2, 2) * invdet;
  return r;
}

// computeInformationOfPointLocal(p, Q = I) = trace(J^T Q^-1 J), J = df_dpc * [-I | skew(p)]  (DEP/src/Helpers.cpp:397-449)
__host__ __device__ inline float information_of_point_local(const V3f &p) {
  const float n = sqrtf(dot3(p.v[0], p.v[0], p.v[1], p.v[1], p.v[2], p.v[2]));
  const float a = 1 / n;
  const float s = 1 / (n * n * n);
  M3f df;
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) df.m[i][j] = a * (i == j ? 1.0f : 0.0f) - (s * p.v[i]) * p.v[j];
  float right[3][6];
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) right[i][j] = (float)((-1.0) * (i == j ? 1.0f : 0.0f));
  right[0][3] = 0;        right[0][4] = -p.v[2]; right[0][5] = p.v[1];
  right[1][3] = p.v[2];   right[1][4] = 0;       right[1][5] = -p.v[0];
  right[2][3] = -p.v[1];  right[2][4] = p.v[0];  right[2][5] = 0;
  float J[3][6];
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 6; ++j) J[i][j] = dot3(df.m[i][0], right[0][j], df.m[i][1], right[1][j], df.m[i][2], right[2][j]);
  M3f Q;
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) Q.m[i][j] = i == j ? 1.0f : 0.0f;
  const M3f Qi = inverse3(Q);
  float d[6];
  for (int i = 0; i < 6; ++i) {
    float q[3];
    for (int k = 0; k < 3; ++k) q[k] = dot3(J[0][i], Qi.m[0][k], J[1][i], Qi.m[1][k], J[2][i], Qi.m[2][k]);
    d[i] = dot3(q[0], J[0][i], q[1], J[1][i], q[2], J[2][i]);
  }
  return (d[0] + (d[1] + d[2])) + (d[3] + (d[4] + d[5]));
}

struct FisherState {
  // lookup table: lut[((ix - org0) * dims1 + (iy - org1)) * dims2 + (iz - org2)], NaN = absent
  DevBuf<float> lut;
  int32_t dims[3] = {0, 0, 0}, org[3] = {0, 0, 0};
  bool have_lut = false;
  bool attr_set = false;
  bool lut_dense = false;       // no NaN inside dims (true for a generated table)
  std::vector<int> axis[3];     // voxel indices per axis in generation order (file order of the .dat records)
  DevBuf<int32_t> d_axis;
  DevBuf<double> crowd;         // S[k] = sum_{j<=k} (float)exp(1 - pow(j, s)), k = 0..CROWD_MAX
  DevBuf<uint2> xrow;           // per x-slab of the table: {first bit of the slab in the CTA bitmap, z_lo | rows << 16} (build_xrows)
  // map points, bucketed
  DevBuf<float4> pts;           // sorted by bin
  DevBuf<float> raw;            // staging copy of the caller's xyz
  DevBuf<uint32_t> bin_start;   // nbx*nby + 1
  DevBuf<uint32_t> bin_cursor;
  DevBuf<float> bbox;           // device reduction scratch: 6 ordered-int floats
  uint64_t n_points = 0;
  bool have_points = false;
  float bx0 = 0, by0 = 0, bin_size = 3.0f, zmin = 0, zmax = 0;
  int nbx = 0, nby = 0;
  // per-call buffers
  DevBuf<fitslam_pose> poses;
  DevBuf<float> info;
  DevBuf<uint8_t> safe;
  DevBuf<uint32_t> nvis;
  // keyframe-structured atlas (fitslam_set_keyframes) and the scratch of its query
  DevBuf<fitslam_keyframe> kf;
  DevBuf<uint32_t> kf_entry_pt, kf_entry_kf;  // per entry: map point, keyframe
  uint32_t n_kf = 0;
  uint64_t n_entries = 0;
  bool have_kf = false;
  DevBuf<uint32_t> kf_first;   // per map point: its first entry among the keyframes that pass the prefilter
  DevBuf<uint32_t> kf_state;   // 4 words per keyframe: passes the prefilter, visible landmarks, landmarks before it, processed
  DevBuf<uint8_t> kf_vis;      // per entry: a visible landmark taken for the first time
  DevBuf<uint32_t> kf_tiles;   // ordered compaction: per-tile counts + total + keyframes checked
  DevBuf<float> kf_out;        // the response, pose frame
  DevBuf<uint32_t> kf_vox;     // Fisher over the response: landmarks per voxel, per landmark {voxel, rank}
  DevBuf<int32_t> kf_lm;
  DevBuf<double> kf_part;
};

static FisherState *fstate(fitslam_ctx *c) {
  if (!c->fi) c->fi = new FisherState();
  return c->fi;
}

void destroy_fisher_state(fitslam_ctx *c) {
  if (!c->fi) return;
  FisherState *f = c->fi;
  f->lut.release(); f->d_axis.release(); f->crowd.release(); f->pts.release(); f->raw.release();
  f->bin_start.release(); f->bin_cursor.release(); f->bbox.release(); f->poses.release(); f->info.release();
  f->safe.release(); f->nvis.release(); f->xrow.release();
  f->kf.release(); f->kf_entry_pt.release(); f->kf_entry_kf.release(); f->kf_first.release(); f->kf_state.release();
  f->kf_vis.release(); f->kf_tiles.release(); f->kf_out.release(); f->kf_vox.release(); f->kf_lm.release(); f->kf_part.release();
  delete f;
  c->fi = nullptr;
}

// ------------------------------------------------------------------------------------------------ lookup table
// One thread per (ix, iy, iz) of the axis lists.  key = (float)(index * (double)step) is what getVoxelCoordinate
// returns for every landmark that rounds to this index (FisherInfoManager.hpp:108-123).
__global__ void lut_fill_kernel(const int32_t *__restrict__ ax, int n0, int n1, int n2, int o0, int o1, int o2, int d1, int d2,
                                double step, float *__restrict__ lut) {
  const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (long)n0 * n1 * n2) return;
  const int a = (int)(i / ((long)n1 * n2)), b = (int)((i / n2) % n1), cc = (int)(i % n2);
  const int ix = ax[a], iy = ax[n0 + b], iz = ax[n0 + n1 + cc];
  V3f p;
  p.v[0] = (float)((double)ix * step);
  p.v[1] = (float)((double)iy * step);
  p.v[2] = (float)((double)iz * step);
  lut[((long)(ix - o0) * d1 + (iy - o1)) * d2 + (iz - o2)] = information_of_point_local(p);
}
__global__ void fill_nan_kernel(float *p, long n) {
  const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) p[i] = __int_as_float(0x7fc00000);
}
// key (0,0,0) := max over the table (FisherInfoManager.cpp:190-197); one block
__global__ void lut_max_kernel(float *lut, long n, long zero_index) {
  __shared__ float sh[32];
  float m = -3.402823466e+38f;
  for (long i = threadIdx.x; i < n; i += blockDim.x) {
    const float v = lut[i];
    if (v == v) m = fmaxf(m, v);
  }
  for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
  if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = m;
  __syncthreads();
  if (threadIdx.x == 0) {
    for (unsigned w = 1; w < blockDim.x / 32; ++w) m = fmaxf(m, sh[w]);
    if (zero_index >= 0) lut[zero_index] = m;
  }
}

// Voxel indices visited by one axis of the generateLookupTable loop nest (FisherInfoManager.cpp:130-150): float bounds
// snapped with floor/ceil, a float counter advanced by step_max (the step_min branch is dead: subSampleVoxelUntil_m =
// -1, so increment_value becomes step_max before the first inner iteration), index = round(counter * (1 / (double)step)).
static void axis_indices(float lo, float hi, float step_max, std::vector<int> &out) {
  out.clear();
  lo = std::floor(lo * (1 / step_max)) * step_max;
  hi = std::ceil(hi * (1 / step_max)) * step_max;
  const double corrected_step = step_max;
  for (float counter = lo; counter <= hi; counter += step_max) {
    const int idx = (int)std::round(counter * (1 / corrected_step));
    if (std::find(out.begin(), out.end(), idx) == out.end()) out.push_back(idx);
    if (out.size() > 4096) break;
  }
}

static fitslam_status upload_crowd_table(fitslam_ctx *c) {
  FisherState *f = fstate(c);
  std::vector<double> S(CROWD_MAX + 1, 0.0);
  for (int j = 1; j <= CROWD_MAX; ++j) {
    // getFactorFromNum (FisherInfoManager.hpp:102-106): float = std::exp(1 - std::pow(int, float))
    const float factor = (float)std::exp(1 - std::pow((double)j, (double)c->prm.crowd_exponent));
    S[j] = S[j - 1] + (double)factor;
  }
  FS_CUDA(c, f->crowd.reserve(S.size()));
  FS_CUDA(c, cudaMemcpyAsync(f->crowd.p, S.data(), S.size() * sizeof(double), cudaMemcpyHostToDevice, c->stream));
  FS_CUDA(c, cudaStreamSynchronize(c->stream));
  return FITSLAM_OK;
}

static fitslam_status build_xrows(fitslam_ctx *c);

static fitslam_status alloc_lut(fitslam_ctx *c, const int lo[3], const int hi[3]) {
  FisherState *f = fstate(c);
  long n = 1;
  for (int a = 0; a < 3; ++a) {
    f->org[a] = std::min(lo[a], 0);  // key (0,0,0) is always an entry
    f->dims[a] = std::max(hi[a], 0) - f->org[a] + 1;
    n *= f->dims[a];
  }
  if (n <= 0 || n >= (long)OUTSIDE_BIT - 1) return fail(c, FITSLAM_E_INVALID, "lookup table too large (voxel index must fit 21 bits)");
  FS_CUDA(c, f->lut.reserve((size_t)n));
  fill_nan_kernel<<<(unsigned)((n + 255) / 256), 256, 0, c->stream>>>(f->lut.p, n);
  ++c->launches;
  return build_xrows(c);
}

// Rows of the CTA bitmap (fisher_kernel).  A visible point has |z| <= t * x in the pose frame, t = the larger half-height of
// the image over fy, so the voxels of x-slab ix it can fall in have |iz| <= t * (ix + 0.5) + 1.5: per slab, dims[1] rows of
// that many bits.  If that does not fit BM_WORDS (a wide image, a long table) the rows are narrowed until it does; landmarks
// outside the rows are counted in the hash table instead, so the rows only decide speed, never the result.
static fitslam_status build_xrows(fitslam_ctx *c) {
  FisherState *f = fstate(c);
  const int nx = f->dims[0], ny = f->dims[1], nz = f->dims[2];
  if (nx > MAX_XROWS || nz > 0xffff) { f->xrow.release(); return FITSLAM_OK; }
  const double t = std::max((double)c->prm.cy, (double)c->prm.img_max_v - (double)c->prm.cy) / (double)c->prm.fy;
  std::vector<uint2> rows((size_t)nx);
  for (double scale = 1.0; scale > 1e-3; scale *= 0.95) {
    uint64_t bits = 0;
    for (int i = 0; i < nx; ++i) {
      const int ix = f->org[0] + i;  // voxel index along the viewing axis
      int z_lo = 0, z_hi = -1;
      if (ix >= 0) {
        const int hz = (int)std::ceil(scale * (t * (ix + 0.5) + 1.5));
        z_lo = std::max(0, -f->org[2] - hz);
        z_hi = std::min(nz - 1, -f->org[2] + hz);
      }
      const int n = std::max(0, z_hi - z_lo + 1);
      rows[(size_t)i] = make_uint2((uint32_t)bits, (uint32_t)z_lo | ((uint32_t)n << 16));
      bits += (uint64_t)ny * n;
    }
    if (bits <= (uint64_t)BM_WORDS * 32) {
      FS_CUDA(c, f->xrow.reserve((size_t)nx));
      FS_CUDA(c, cudaMemcpyAsync(f->xrow.p, rows.data(), (size_t)nx * sizeof(uint2), cudaMemcpyHostToDevice, c->stream));
      FS_CUDA(c, cudaStreamSynchronize(c->stream));
      return FITSLAM_OK;
    }
  }
  f->xrow.release();
  return FITSLAM_OK;
}

static long lut_size(const FisherState *f) { return (long)f->dims[0] * f->dims[1] * f->dims[2]; }

}  // namespace fitslam

using namespace fitslam;

extern "C" {

fitslam_status fitslam_fisher_generate_lookup(fitslam_ctx *c, const char *save_path) {
  if (!c) return FITSLAM_E_INVALID;
  cudaSetDevice(c->device);
  FisherState *f = fstate(c);
  const fitslam_params &p = c->prm;
  if (!(p.voxel_step > 0.0f)) return fail(c, FITSLAM_E_INVALID, "voxel_step must be positive");
  axis_indices(p.lut_min_x, p.lut_max_x, p.voxel_step, f->axis[0]);
  axis_indices(p.lut_min_y, p.lut_max_y, p.voxel_step, f->axis[1]);
  axis_indices(p.lut_min_z, p.lut_max_z, p.voxel_step, f->axis[2]);
  int lo[3], hi[3];
  std::vector<int32_t> flat;
  for (int a = 0; a < 3; ++a) {
    if (f->axis[a].empty()) return fail(c, FITSLAM_E_INVALID, "empty lookup-table axis");
    lo[a] = *std::min_element(f->axis[a].begin(), f->axis[a].end());
    hi[a] = *std::max_element(f->axis[a].begin(), f->axis[a].end());
    flat.insert(flat.end(), f->axis[a].begin(), f->axis[a].end());
  }
  fitslam_status st = alloc_lut(c, lo, hi);
  if (st != FITSLAM_OK) return st;
  FS_CUDA(c, f->d_axis.reserve(flat.size()));
  FS_CUDA(c, cudaMemcpyAsync(f->d_axis.p, flat.data(), flat.size() * 4, cudaMemcpyHostToDevice, c->stream));
  const int n0 = (int)f->axis[0].size(), n1 = (int)f->axis[1].size(), n2 = (int)f->axis[2].size();
  const long n = (long)n0 * n1 * n2;
  lut_fill_kernel<<<(unsigned)((n + 255) / 256), 256, 0, c->stream>>>(f->d_axis.p, n0, n1, n2, f->org[0], f->org[1], f->org[2],
                                                                     f->dims[1], f->dims[2], (double)p.voxel_step, f->lut.p);
  const long zero = ((long)(0 - f->org[0]) * f->dims[1] + (0 - f->org[1])) * f->dims[2] + (0 - f->org[2]);
  lut_max_kernel<<<1, 1024, 0, c->stream>>>(f->lut.p, lut_size(f), zero);
  c->launches += 2;
  FS_CUDA(c, cudaGetLastError());
  FS_CUDA(c, cudaStreamSynchronize(c->stream));
  f->have_lut = true;
  {  // dense iff the table has no hole: one host pass over a copy (once per table)
    std::vector<float> chk((size_t)lut_size(f));
    FS_CUDA(c, cudaMemcpy(chk.data(), f->lut.p, chk.size() * 4, cudaMemcpyDeviceToHost));
    f->lut_dense = true;
    for (float v : chk)
      if (v != v) { f->lut_dense = false; break; }
  }
  st = upload_crowd_table(c);
  if (st != FITSLAM_OK) return st;
  if (save_path) {
    // .dat layout (FisherInfoManager.cpp:185-186): {float key[3], float value} per entry, loop order, (0,0,0) last
    std::vector<float> host((size_t)lut_size(f));
    FS_CUDA(c, cudaMemcpy(host.data(), f->lut.p, host.size() * 4, cudaMemcpyDeviceToHost));
    FILE *fp = fopen(save_path, "wb");
    if (!fp) return fail(c, FITSLAM_E_IO, std::string("cannot open ") + save_path);
    const double step = (double)p.voxel_step;
    auto at = [&](int ix, int iy, int iz) {
      return host[((size_t)(ix - f->org[0]) * f->dims[1] + (iy - f->org[1])) * f->dims[2] + (iz - f->org[2])];
    };
    for (int ix : f->axis[0])
      for (int iy : f->axis[1])
        for (int iz : f->axis[2]) {
          if (ix == 0 && iy == 0 && iz == 0) continue;  // NaN at generation time: skipped, appended below
          const float v = at(ix, iy, iz);
          if (v != v) continue;
          const float rec[4] = {(float)(ix * step), (float)(iy * step), (float)(iz * step), v};
          fwrite(rec, sizeof(float), 4, fp);
        }
    const float last[4] = {0.0f, 0.0f, 0.0f, at(0, 0, 0)};
    fwrite(last, sizeof(float), 4, fp);
    fclose(fp);
  }
  return FITSLAM_OK;
}

fitslam_status fitslam_fisher_load_lookup(fitslam_ctx *c, const char *path) {
  if (!c || !path) return fail(c, FITSLAM_E_INVALID, "load_lookup: null argument");
  cudaSetDevice(c->device);
  FisherState *f = fstate(c);
  FILE *fp = fopen(path, "rb");
  if (!fp) return fail(c, FITSLAM_E_NO_LOOKUP_TABLE, "Cannot load lookup table. Please generate it first.");
  std::vector<float> rec;
  float r[4];
  while (fread(r, sizeof(float), 4, fp) == 4) rec.insert(rec.end(), r, r + 4);
  fclose(fp);
  const size_t n = rec.size() / 4;
  if (n == 0) return fail(c, FITSLAM_E_NO_LOOKUP_TABLE, "Cannot load lookup table. Please generate it first.");
  const double inv = 1 / (double)c->prm.voxel_step;
  int lo[3] = {0, 0, 0}, hi[3] = {0, 0, 0};
  std::vector<int> idx(3 * n);
  for (size_t i = 0; i < n; ++i)
    for (int a = 0; a < 3; ++a) {
      const int v = (int)std::round((double)rec[4 * i + a] * inv);  // the key IS index * step: the same index comes back
      idx[3 * i + a] = v;
      lo[a] = std::min(lo[a], v);
      hi[a] = std::max(hi[a], v);
    }
  fitslam_status st = alloc_lut(c, lo, hi);
  if (st != FITSLAM_OK) return st;
  std::vector<float> host((size_t)lut_size(f), std::numeric_limits<float>::quiet_NaN());
  for (size_t i = 0; i < n; ++i)  // later records overwrite earlier ones, like unordered_map::operator[]
    host[((size_t)(idx[3 * i] - f->org[0]) * f->dims[1] + (idx[3 * i + 1] - f->org[1])) * f->dims[2] + (idx[3 * i + 2] - f->org[2])] = rec[4 * i + 3];
  FS_CUDA(c, cudaMemcpyAsync(f->lut.p, host.data(), host.size() * 4, cudaMemcpyHostToDevice, c->stream));
  FS_CUDA(c, cudaStreamSynchronize(c->stream));
  for (int a = 0; a < 3; ++a) f->axis[a].clear();
  f->lut_dense = true;
  for (float v : host)
    if (v != v) { f->lut_dense = false; break; }
  f->have_lut = true;
  return upload_crowd_table(c);
}

fitslam_status fitslam_fisher_lookup_dims(const fitslam_ctx *c, int32_t dims[3], int32_t origin[3]) {
  if (!c || !c->fi || !c->fi->have_lut) return FITSLAM_E_NO_LOOKUP_TABLE;
  for (int a = 0; a < 3; ++a) {
    if (dims) dims[a] = c->fi->dims[a];
    if (origin) origin[a] = c->fi->org[a];
  }
  return FITSLAM_OK;
}

fitslam_status fitslam_fisher_lookup_copy(fitslam_ctx *c, float *out, uint64_t capacity) {
  if (!c || !out) return fail(c, FITSLAM_E_INVALID, "lookup_copy: null argument");
  if (!c->fi || !c->fi->have_lut) return fail(c, FITSLAM_E_NO_LOOKUP_TABLE, "Cannot load lookup table. Please generate it first.");
  const uint64_t n = (uint64_t)lut_size(c->fi);
  if (capacity < n) return fail(c, FITSLAM_E_CAPACITY, "lookup_copy: buffer too small");
  cudaSetDevice(c->device);
  FS_CUDA(c, cudaMemcpyAsync(out, c->fi->lut.p, n * 4, cudaMemcpyDefault, c->stream));
  FS_CUDA(c, cudaStreamSynchronize(c->stream));
  return FITSLAM_OK;
}

}  // extern "C"

// ================================================================================================ map points
namespace fitslam {

__device__ __forceinline__ uint32_t f2ord(float f) {
  const uint32_t u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
static inline float ord2f(uint32_t o) {
  const uint32_t u = (o & 0x80000000u) ? (o & 0x7fffffffu) : ~o;
  float f;
  memcpy(&f, &u, 4);
  return f;
}
__device__ __forceinline__ bool finite3(float x, float y, float z) { return isfinite(x) && isfinite(y) && isfinite(z); }

// ord[0..2] = min xyz, ord[3..5] = max xyz over the finite points
__global__ void points_bbox_kernel(const float *__restrict__ xyz, uint64_t n, uint32_t *ord) {
  uint32_t lo[3] = {0xffffffffu, 0xffffffffu, 0xffffffffu}, hi[3] = {0, 0, 0};
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
    const float x = xyz[3 * i], y = xyz[3 * i + 1], z = xyz[3 * i + 2];
    if (!finite3(x, y, z)) continue;
    const uint32_t o[3] = {f2ord(x), f2ord(y), f2ord(z)};
    for (int a = 0; a < 3; ++a) {
      lo[a] = min(lo[a], o[a]);
      hi[a] = max(hi[a], o[a]);
    }
  }
  for (int a = 0; a < 3; ++a) {
    for (int s = 16; s > 0; s >>= 1) {
      lo[a] = min(lo[a], __shfl_xor_sync(0xffffffffu, lo[a], s));
      hi[a] = max(hi[a], __shfl_xor_sync(0xffffffffu, hi[a], s));
    }
    if ((threadIdx.x & 31) == 0) {
      atomicMin(&ord[a], lo[a]);
      atomicMax(&ord[3 + a], hi[a]);
    }
  }
}

struct BinGrid {
  float x0, y0, inv, size;
  int nbx, nby;
};
__device__ __forceinline__ int bin_of(const BinGrid &g, float x, float y) {
  int bx = (int)floorf((x - g.x0) * g.inv), by = (int)floorf((y - g.y0) * g.inv);
  bx = max(0, min(g.nbx - 1, bx));
  by = max(0, min(g.nby - 1, by));
  return by * g.nbx + bx;
}
__global__ void points_count_kernel(const float *__restrict__ xyz, uint64_t n, BinGrid g, uint32_t *count) {
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
    const float x = xyz[3 * i], y = xyz[3 * i + 1], z = xyz[3 * i + 2];
    if (finite3(x, y, z)) atomicAdd(&count[bin_of(g, x, y)], 1u);
  }
}
// exclusive scan of count[0..n) in place, count[n] = total; start[] = cursor[] copies.  One block.
__global__ void bins_scan_kernel(uint32_t *count, uint32_t *cursor, uint32_t n) {
  __shared__ uint32_t wsum[32];
  __shared__ uint32_t carry;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  const uint32_t lane = threadIdx.x & 31u, wid = threadIdx.x >> 5;
  for (uint32_t base = 0; base < n; base += blockDim.x) {
    const uint32_t i = base + threadIdx.x;
    const uint32_t v = i < n ? count[i] : 0;
    uint32_t incl = v;
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
      if ((int)lane >= o) incl += t;
    }
    if (lane == 31) wsum[wid] = incl;
    __syncthreads();
    if (wid == 0) {
      uint32_t w = lane < blockDim.x / 32 ? wsum[lane] : 0, wi = w;
      for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, wi, o);
        if ((int)lane >= o) wi += t;
      }
      wsum[lane] = wi - w;
    }
    __syncthreads();
    const uint32_t excl = carry + wsum[wid] + incl - v;
    if (i < n) {
      count[i] = excl;
      cursor[i] = excl;
    }
    __syncthreads();
    if (threadIdx.x == blockDim.x - 1) carry = excl + v;
    __syncthreads();
  }
  if (threadIdx.x == 0) count[n] = carry;
}
__global__ void points_scatter_kernel(const float *__restrict__ xyz, uint64_t n, BinGrid g, uint32_t *cursor, float4 *out) {
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
    const float x = xyz[3 * i], y = xyz[3 * i + 1], z = xyz[3 * i + 2];
    if (!finite3(x, y, z)) continue;
    const uint32_t pos = atomicAdd(&cursor[bin_of(g, x, y)], 1u);
    out[pos] = make_float4(x, y, z, __uint_as_float((uint32_t)i));
  }
}

// radius (map frame, metres) that contains the whole lookup domain seen from the pose origin, any orientation
static double lut_reach(const fitslam_ctx *c) {
  const FisherState *f = c->fi;
  const double step = (double)c->prm.voxel_step;
  double r2 = 0;
  for (int a = 0; a < 3; ++a) {
    const double lo = (f->org[a] - 0.5) * step, hi = (f->org[a] + f->dims[a] - 0.5) * step;
    const double m = std::max(std::fabs(lo), std::fabs(hi));
    r2 += m * m;
  }
  return std::sqrt(r2) + 0.1;
}

}  // namespace fitslam

extern "C" fitslam_status fitslam_set_map_points(fitslam_ctx *c, const float *xyz, uint64_t n) {
  // with a communicator the call is collective: rank 0 uploads, the others pass xyz = NULL and receive
  const bool receive = c && comm_active(c) && c->rank != 0;
  if (!c || (!xyz && n && !receive)) return fail(c, FITSLAM_E_INVALID, "set_map_points: null argument");
  if (n >= (1ull << 31)) return fail(c, FITSLAM_E_INVALID, "set_map_points: more than 2^31 points");
  cudaSetDevice(c->device);
  FisherState *f = fstate(c);
  f->have_points = false;
  f->have_kf = false;  // the keyframes' point indices meant the previous points
  f->n_points = n;
  FS_CUDA(c, f->raw.reserve(3 * n + 3));
  FS_CUDA(c, f->pts.reserve(n + 1));
  FS_CUDA(c, f->bbox.reserve(8));
  if (n && !receive) FS_CUDA(c, cudaMemcpyAsync(f->raw.p, xyz, 3 * n * sizeof(float), cudaMemcpyDefault, c->stream));
  {
    const fitslam_status bst = comm_broadcast(c, f->raw.p, 3 * n * sizeof(float));
    if (bst != FITSLAM_OK) return bst;
  }
  const uint32_t init[6] = {0xffffffffu, 0xffffffffu, 0xffffffffu, 0, 0, 0};
  FS_CUDA(c, cudaMemcpyAsync(f->bbox.p, init, sizeof(init), cudaMemcpyHostToDevice, c->stream));
  const unsigned grid = (unsigned)std::min<uint64_t>(148 * 8, (n + 255) / 256 + 1);
  points_bbox_kernel<<<grid, 256, 0, c->stream>>>(f->raw.p, n, reinterpret_cast<uint32_t *>(f->bbox.p));
  ++c->launches;
  uint32_t ord[6];
  FS_CUDA(c, cudaMemcpyAsync(ord, f->bbox.p, sizeof(ord), cudaMemcpyDeviceToHost, c->stream));
  FS_CUDA(c, cudaStreamSynchronize(c->stream));
  float lo[3] = {0, 0, 0}, hi[3] = {0, 0, 0};
  if (ord[0] <= ord[3])
    for (int a = 0; a < 3; ++a) {
      lo[a] = ord2f(ord[a]);
      hi[a] = ord2f(ord[3 + a]);
    }
  // bin size: >= 3 m, and large enough that (a) the grid has at most 2^20 bins, (b) a pose touches <= 30 x 30 bins
  double bs = 3.0;
  const double ex = (double)hi[0] - lo[0], ey = (double)hi[1] - lo[1];
  while ((std::floor(ex / bs) + 1) * (std::floor(ey / bs) + 1) > (double)(1 << 20)) bs *= 2;
  f->bin_size = (float)bs;
  f->bx0 = lo[0];
  f->by0 = lo[1];
  f->zmin = lo[2];
  f->zmax = hi[2];
  f->nbx = (int)std::floor(ex / bs) + 1;
  f->nby = (int)std::floor(ey / bs) + 1;
  const uint32_t nb = (uint32_t)f->nbx * f->nby;
  FS_CUDA(c, f->bin_start.reserve(nb + 1));
  FS_CUDA(c, f->bin_cursor.reserve(nb + 1));
  FS_CUDA(c, cudaMemsetAsync(f->bin_start.p, 0, (nb + 1) * 4, c->stream));
  BinGrid g{f->bx0, f->by0, (float)(1.0 / bs), f->bin_size, f->nbx, f->nby};
  points_count_kernel<<<grid, 256, 0, c->stream>>>(f->raw.p, n, g, f->bin_start.p);
  bins_scan_kernel<<<1, 1024, 0, c->stream>>>(f->bin_start.p, f->bin_cursor.p, nb);
  points_scatter_kernel<<<grid, 256, 0, c->stream>>>(f->raw.p, n, g, f->bin_cursor.p, f->pts.p);
  c->launches += 3;
  FS_CUDA(c, cudaGetLastError());
  FS_CUDA(c, cudaStreamSynchronize(c->stream));
  f->have_points = true;
  return FITSLAM_OK;
}

// ================================================================================================ scoring
namespace fitslam {

struct FisherParams {
  const float4 *pts;
  const uint32_t *bin_start;
  BinGrid grid;
  float zmin, zmax;
  const float *lut;
  int dims[3], org[3];
  const double *crowd;
  double inv_step;      // 1 / (double)step_max, the factor getVoxelCoordinate multiplies by
  float inv_step_f;     // the same in float: prefilter of the rounding, see voxel_of_point
  double step;
  double reach;         // lut_reach()
  double threshold;
  float fx, fy, cx, cy, max_u, max_v;
  uint32_t n_poses;
  uint32_t first;       // this launch scores poses [first, first + gridDim.x)
  uint32_t poses_per_frontier;  // record f = pose / poses_per_frontier (cycle mode)
  int lut_dense;        // every voxel inside dims has an entry: the visibility test needs no table read
  const uint2 *xrow;    // dims[0] entries (build_xrows); null: no bitmap, every landmark is counted in the hash table
  int use_slice;        // poses are frontier * poses_per_frontier + window for the frontiers of `slice` (cycle.cu)
  ScoreSlice slice;
};

struct PoseFrame {      // per-CTA, built by thread 0
  M3f Rcw;  V3f tcw;    // se3ROSToORB(poseToAffine(pose))
  M3f Rinv; V3f tinv;   // inverse(poseToAffine(pose)): map -> pose frame
  double plane[11][4];  // n.p + d >= -margin inside (map frame, unit normals): 5 frustum + 6 lookup-domain planes
  float px, py, pz;
  float guard;          // metres: bound on |one-transform local coordinates - either of the reference's float evaluations|
  float gu, gv, gvox;   // the guard in the units of the image tests (times Zc) and of the voxel rounding
  int exact_only;       // Rinv is not the transpose of R to rounding (a quaternion far from unit length): no shortcut
};

__device__ void build_pose_frame(const fitslam_pose &pose, const FisherParams &P, PoseFrame &F) {
  // poseToAffine (ORBW/src/type_conversion.cpp:141-153): doubles narrowed to float
  V3f t;
  t.v[0] = (float)pose.x; t.v[1] = (float)pose.y; t.v[2] = (float)pose.z;
  const M3f R = quat_to_matrix((float)pose.qw, (float)pose.qx, (float)pose.qy, (float)pose.qz);
  M3f ros2orb, orb2ros;
  const float a[9] = {0, -1, 0, 0, 0, -1, 1, 0, 0}, b[9] = {0, 0, 1, -1, 0, 0, 0, -1, 0};
  for (int i = 0; i < 9; ++i) { ros2orb.m[i / 3][i % 3] = a[i]; orb2ros.m[i / 3][i % 3] = b[i]; }
  // se3ROSToORB (type_conversion.cpp:74-100)
  const M3f Rtemp = mulm(ros2orb, R);
  const V3f ttemp = mulv(ros2orb, t);
  const M3f Rti = transp(Rtemp);
  const V3f ti = mulv(Rti, ttemp);
  V3f tneg; tneg.v[0] = -ti.v[0]; tneg.v[1] = -ti.v[1]; tneg.v[2] = -ti.v[2];
  F.Rcw = mulm(ros2orb, Rti);
  F.tcw = mulv(ros2orb, tneg);
  // Eigen Affine3f::inverse(): linear().inverse(), translation = -linear_inv * t
  F.Rinv = inverse3(R);
  M3f neg;
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) neg.m[i][j] = -F.Rinv.m[i][j];
  F.tinv = mulv(neg, t);
  F.px = t.v[0]; F.py = t.v[1]; F.pz = t.v[2];
  // The point loop evaluates l = Rinv * (p - t) once and reads both the camera coordinates (-l.y, -l.z, l.x) and the pose-frame
  // coordinates off it; the reference evaluates them separately as Rcw * Pw + tcw and Rinv * p + tinv.  All three round the
  // same real vector (when Rinv = R^T, which holds to rounding for a unit quaternion).  Error of a reference evaluation:
  // three products and two sums per dot product, each 2^-24 relative to at most |p|_1 = |t|_1 + |d|_1, the same for the
  // translation term, one more rounding for the final sum: under 2^-22 * (2 |t|_1 + |d|_1).  Of the one-transform form:
  // 2^-22 * |d|_1.  |d|_1 <= 3 * reach < 80 m.  The guard doubles that sum; a decision closer than the guard to its boundary
  // is taken again with the reference's own arithmetic.
  float asym = 0.f;
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) asym = fmaxf(asym, fabsf(F.Rinv.m[i][j] - R.m[j][i]));
  F.exact_only = !(asym <= 1e-5f) ? 1 : 0;  // (also NaN)
  F.guard = 1e-6f * (fabsf(t.v[0]) + fabsf(t.v[1]) + fabsf(t.v[2]) + 96.f) + 3.f * asym * 96.f;
  F.gu = (P.fx + fmaxf(P.cx, P.max_u - P.cx)) * F.guard;
  F.gv = (P.fy + fmaxf(P.cy, P.max_v - P.cy)) * F.guard;
  F.gvox = F.guard * P.inv_step_f + 2e-4f;  // + the rounding of the float product
}

// ---- conservative culling planes (double; only used to skip whole bins): plane k of 11, one thread each
__device__ void build_pose_plane(const FisherParams &P, PoseFrame &F, int k) {
  if (k < 5) {
    // camera coords = Rcw * Perm * p + tcw with Perm = ros2orb: column j of M = Rcw * Perm
    double M[3][3];
    for (int i = 0; i < 3; ++i) {
      M[i][0] = F.Rcw.m[i][2]; M[i][1] = -(double)F.Rcw.m[i][0]; M[i][2] = -(double)F.Rcw.m[i][1];
    }
    double cam[3];
    if (k == 0) { cam[0] = 0; cam[1] = 0; cam[2] = 1; }
    else if (k == 1) { cam[0] = P.fx; cam[1] = 0; cam[2] = P.cx; }
    else if (k == 2) { cam[0] = -(double)P.fx; cam[1] = 0; cam[2] = (double)P.max_u - P.cx; }
    else if (k == 3) { cam[0] = 0; cam[1] = P.fy; cam[2] = P.cy; }
    else { cam[0] = 0; cam[1] = -(double)P.fy; cam[2] = (double)P.max_v - P.cy; }
    const double nn = sqrt(cam[0] * cam[0] + cam[1] * cam[1] + cam[2] * cam[2]);
    double d = 0;
    for (int j = 0; j < 3; ++j) {
      F.plane[k][j] = (cam[0] * M[0][j] + cam[1] * M[1][j] + cam[2] * M[2][j]) / nn;
      d += cam[j] * (double)F.tcw.v[j];
    }
    F.plane[k][3] = d / nn;
  } else {
    const int ax = (k - 5) >> 1;
    const bool upper = (k - 5) & 1;
    const double lo = (P.org[ax] - 0.5) * P.step, hi = (P.org[ax] + P.dims[ax] - 0.5) * P.step;
    for (int j = 0; j < 3; ++j) F.plane[k][j] = upper ? -(double)F.Rinv.m[ax][j] : (double)F.Rinv.m[ax][j];
    F.plane[k][3] = upper ? hi - (double)F.tinv.v[ax] : (double)F.tinv.v[ax] - lo;
  }
}

__device__ __forceinline__ uint32_t mix32(uint32_t x) {
  x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
  return x;
}

// ORBSLAM3Interface::mapPointsVisibleFromPose for one map point: in front of the camera and inside the image
// (orb_slam3_interface.cpp:286-300, Pinhole.cpp:43-49)
__device__ __forceinline__ bool point_visible(const float4 q, const PoseFrame &F, const FisherParams &P) {
  V3f Pw; Pw.v[0] = -q.y; Pw.v[1] = -q.z; Pw.v[2] = q.x;  // ROS map axes -> ORB world axes
  const V3f RP = mulv(F.Rcw, Pw);
  const float Xc = RP.v[0] + F.tcw.v[0], Yc = RP.v[1] + F.tcw.v[1], Zc = RP.v[2] + F.tcw.v[2];
  if (Zc < 0.0f) return false;
  const float u = P.fx * Xc / Zc + P.cx;
  const float v = P.fy * Yc / Zc + P.cy;
  return !(u < 0.0f || u > P.max_u) && !(v < 0.0f || v > P.max_v);
}

// voxel of a visible map point in the pose frame: -1 if outside the table or a NaN entry
__device__ __forceinline__ int voxel_of_visible_point(const float4 q, const PoseFrame &F, const FisherParams &P) {
  // rgbd-slam-node.cpp:419-425: vector3fORBToROS, then inverse(pose) * position.  vector3fORBToROS multiplies by a signed
  // permutation matrix, 0*a + (0*b + 1*c): for finite inputs that is the original ROS triple again (up to the sign of a
  // zero, which cannot change a voxel index)
  V3f pos; pos.v[0] = q.x; pos.v[1] = q.y; pos.v[2] = q.z;
  const V3f lp = mulv(F.Rinv, pos);
  int idx[3];
  bool ok = true;
#pragma unroll
  for (int a = 0; a < 3; ++a) {
    const float c = lp.v[a] + F.tinv.v[a];
    // getVoxelCoordinate (FisherInfoManager.hpp:108-123): round((double)c * (1 / (double)step)).  The float product is
    // within 2^-23 |c / step| < 6.2e-5 of the double one while |c / step| < 512, so farther than 1e-4 from a
    // half-integer both round to the same index; only the remaining cases (2 in 10^4) take the double path.
    const float rf = c * P.inv_step_f;
    const float ri = rintf(rf);
    int r;
    if (fabsf(rf - ri) < 0.4999f && fabsf(rf) < 512.0f) {
      r = (int)ri;
    } else {
      const double rd = round((double)c * P.inv_step);
      r = (rd >= -2147483000.0 && rd <= 2147483000.0) ? (int)rd : INT_MIN;  // also rejects NaN
    }
    ok = ok && r >= P.org[a] && r < P.org[a] + P.dims[a];
    idx[a] = r - P.org[a];
  }
  if (!ok) return -1;
  const int lin = (idx[0] * P.dims[1] + idx[1]) * P.dims[2] + idx[2];
  if (P.lut_dense) return lin;
  const float val = __ldg(P.lut + lin);
  return (val == val) ? lin : -1;
}

// The same decision as point_visible with one approximate reciprocal instead of two IEEE divisions: u and v are within 5e-4
// of the exact ones (|u|, |v| < 1000 px), so a point at least 0.01 px inside all four image borders is visible and one a pixel
// outside any of them is not; the others (3 in 1000) take the reference's arithmetic.
__device__ __forceinline__ bool point_visible_fast(const float4 q, const PoseFrame &F, const FisherParams &P) {
  V3f Pw; Pw.v[0] = -q.y; Pw.v[1] = -q.z; Pw.v[2] = q.x;
  const V3f RP = mulv(F.Rcw, Pw);
  const float Xc = RP.v[0] + F.tcw.v[0], Yc = RP.v[1] + F.tcw.v[1], Zc = RP.v[2] + F.tcw.v[2];
  if (Zc < 0.0f) return false;
  const float tx = P.fx * Xc, ty = P.fy * Yc;
  if (Zc >= 1e-30f) {
    const float rz = __frcp_rn(Zc);
    const float ua = tx * rz + P.cx, va = ty * rz + P.cy;
    if (ua >= 0.01f && ua <= P.max_u - 0.01f && va >= 0.01f && va <= P.max_v - 0.01f) return true;
    if (ua < -1.0f || ua > P.max_u + 1.0f || va < -1.0f || va > P.max_v + 1.0f) return false;
  }
  const float u = tx / Zc + P.cx;
  const float v = ty / Zc + P.cy;
  return !(u < 0.0f || u > P.max_u) && !(v < 0.0f || v > P.max_v);
}

// voxel of a visible map point in the pose frame: the index into the lookup table (-1 if outside the table or a NaN entry)
// and its bit in the CTA bitmap (-1 if the bitmap has no row for it)
__device__ __forceinline__ int voxel_and_bit(const float4 q, const PoseFrame &F, const FisherParams &P, const uint2 *xrow, int *bit) {
  V3f pos; pos.v[0] = q.x; pos.v[1] = q.y; pos.v[2] = q.z;
  const V3f lp = mulv(F.Rinv, pos);
  int idx[3];
  bool ok = true;
#pragma unroll
  for (int a = 0; a < 3; ++a) {  // as voxel_of_visible_point
    const float c = lp.v[a] + F.tinv.v[a];
    const float rf = c * P.inv_step_f;
    const float ri = rintf(rf);
    int r;
    if (fabsf(rf - ri) < 0.4999f && fabsf(rf) < 512.0f) {
      r = (int)ri;
    } else {
      const double rd = round((double)c * P.inv_step);
      r = (rd >= -2147483000.0 && rd <= 2147483000.0) ? (int)rd : INT_MIN;
    }
    ok = ok && r >= P.org[a] && r < P.org[a] + P.dims[a];
    idx[a] = r - P.org[a];
  }
  *bit = -1;
  if (!ok) return -1;
  const int lin = (idx[0] * P.dims[1] + idx[1]) * P.dims[2] + idx[2];
  if (!P.lut_dense) {
    const float val = __ldg(P.lut + lin);
    if (!(val == val)) return -1;
  }
  if (xrow) {
    const uint2 xr = xrow[idx[0]];
    const int rows = (int)(xr.y >> 16), zr = idx[2] - (int)(xr.y & 0xffffu);
    if ((unsigned)zr < (unsigned)rows) *bit = (int)xr.x + idx[1] * rows + zr;
  }
  return lin;
}

// count one landmark in the hash table: packed slot key << CNT_BITS | count; plain loads walk the probe sequence and only
// the claiming / counting step is an atomic (shared-memory atomics cost 2-4 cycles per LANE on this part).
// Returns the count including this landmark (1 = it claimed the slot), or 0 if the table is full.
__device__ __forceinline__ uint32_t table_count(uint32_t *tab, uint32_t h, uint32_t key22) {
  uint32_t slot = h & (SEC_SLOTS - 1);
  const uint32_t key = key22 << CNT_BITS;
  for (int probe = 0; probe < 256; ++probe) {
    uint32_t prev = ((volatile uint32_t *)tab)[slot];
    if (prev == 0u) prev = atomicCAS(&tab[slot], 0u, key | 1u);
    if (prev == 0u) return 1u;
    if ((prev & ~CNT_MASK) == key) {
      // saturates far above CROWD_MAX: ranks beyond it carry a zero crowding factor anyway
      if ((prev & CNT_MASK) >= CNT_MASK - 256u) return CROWD_MAX + 1u;
      return min((atomicAdd(&tab[slot], 1u) & CNT_MASK) + 1u, (uint32_t)CROWD_MAX + 1u);
    }
    slot = (slot + 1) & (SEC_SLOTS - 1);
  }
  return 0u;
}

__global__ void __launch_bounds__(FS_FISHER_BOUND, FS_FISHER_MINB)
fisher_kernel(const FisherParams P, const fitslam_pose *__restrict__ poses, float *__restrict__ info_out,
              uint8_t *__restrict__ safe_out, uint32_t *__restrict__ nvis_out, ScoreRecord *__restrict__ rec_out,
              const fitslam_arrival *__restrict__ arrival) {
  __shared__ PoseFrame F;
  extern __shared__ uint32_t dyn[];
  uint32_t *bm = dyn;               // BM_WORDS: bit = the voxel has its first landmark
  uint32_t *tab = dyn + BM_WORDS;   // SEC_SLOTS packed slots, 0 = empty
  __shared__ uint32_t s_used, s_coll, s_overflow, s_nchunk, s_next, s_chunk_ovf;
  __shared__ uint2 s_chunk[MAX_CHUNKS];  // [first, last) point ranges of the bins that survive the cull
  __shared__ uint2 s_cq[FISHER_THREADS / 32][64];      // per warp: {table index | outside << 31, information} of landmarks
                                                       // waiting for the hash-table step
  __shared__ uint32_t s_uq[FISHER_THREADS / 32][64];   // per warp: points waiting for the reference's own arithmetic
  __shared__ uint2 s_xrow[MAX_XROWS];
  __shared__ double s_red[FISHER_THREADS / 32];
  uint32_t pi;
  if (P.use_slice) {
    const uint32_t slot = blockIdx.x / P.poses_per_frontier;
    uint32_t f = 0;
    if (P.slice.work_count ? !slice_work_frontier(P.slice, slot, f) : !slice_frontier(P.slice, slot, f)) return;
    pi = f * P.poses_per_frontier + (blockIdx.x - slot * P.poses_per_frontier);
  } else {
    pi = P.first + blockIdx.x;
    if (pi >= P.n_poses) return;
  }
  if (threadIdx.x == 0) build_pose_frame(poses[pi], P, F);
  else if (P.xrow && threadIdx.x >= 32 && threadIdx.x - 32 < (uint32_t)P.dims[0]) s_xrow[threadIdx.x - 32] = __ldg(P.xrow + threadIdx.x - 32);
  const uint2 *xrow = P.xrow ? s_xrow : nullptr;  // (the exact path's argument)
  const bool have_rows = P.xrow != nullptr;
  __syncthreads();
  if (threadIdx.x < 11) build_pose_plane(P, F, (int)threadIdx.x);
  __syncthreads();
  const uint32_t lane = threadIdx.x & 31u, wid = threadIdx.x >> 5, nw = FISHER_THREADS / 32;
  // bins of the bounding square of the lookup domain around the pose
  const BinGrid g = P.grid;
  const int bx_lo = max(0, (int)floor(((double)F.px - P.reach - g.x0) * g.inv)), bx_hi = min(g.nbx - 1, (int)floor(((double)F.px + P.reach - g.x0) * g.inv));
  const int by_lo = max(0, (int)floor(((double)F.py - P.reach - g.y0) * g.inv)), by_hi = min(g.nby - 1, (int)floor(((double)F.py + P.reach - g.y0) * g.inv));
  const int sq_w = bx_hi - bx_lo + 1, sq_h = by_hi - by_lo + 1;
  const int n_sq = (sq_w > 0 && sq_h > 0) ? sq_w * sq_h : 0;
  const double margin = 0.05 + 1e-4 * (fabs((double)F.px) + fabs((double)F.py));

  // ---- cull the bins once, and cut the surviving ones into chunks of at most CHUNK_PTS points: the chunk list is
  // the unit of work the warps pull from a shared counter, so a warp that met empty or culled bins takes more chunks
  // (a static bins-per-warp split left half of the warps waiting at the barrier)
  for (uint32_t chunk_pts = CHUNK_PTS;; chunk_pts <<= 1) {  // a very dense cloud overflows the list: coarser chunks then
  if (threadIdx.x == 0) { s_nchunk = 0; s_chunk_ovf = 0; }
  __syncthreads();
  for (int b0 = (int)wid * 32; b0 < n_sq; b0 += (int)nw * 32) {
    const int b = b0 + (int)lane;
    uint32_t start = 0, end = 0;
    bool keep = false;
    if (b < n_sq) {
      const int bx = bx_lo + b % sq_w, by = by_lo + b / sq_w;
      const uint32_t bi = (uint32_t)by * g.nbx + bx;
      start = __ldg(P.bin_start + bi);
      end = __ldg(P.bin_start + bi + 1);
      if (end > start) {
        // clamped edge bins also hold points outside their nominal square: open the box on that side
        const double x0 = bx == 0 ? -1e30 : (double)g.x0 + bx * (double)g.size, x1 = bx == g.nbx - 1 ? 1e30 : (double)g.x0 + (bx + 1) * (double)g.size;
        const double y0 = by == 0 ? -1e30 : (double)g.y0 + by * (double)g.size, y1 = by == g.nby - 1 ? 1e30 : (double)g.y0 + (by + 1) * (double)g.size;
        keep = true;
#pragma unroll 1
        for (int k = 0; k < 11 && keep; ++k) {
          const double nx = F.plane[k][0], ny = F.plane[k][1], nz = F.plane[k][2];
          const double mx = fmax(nx * x0, nx * x1) + fmax(ny * y0, ny * y1) + fmax(nz * (double)P.zmin, nz * (double)P.zmax) + F.plane[k][3];
          if (mx < -margin) keep = false;
        }
      }
    }
    const uint32_t nch = keep ? (end - start + chunk_pts - 1) / chunk_pts : 0u;
    uint32_t incl = nch;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
      if ((int)lane >= o) incl += t;
    }
    const uint32_t tot = __shfl_sync(0xffffffffu, incl, 31);
    uint32_t base = 0;
    if (lane == 31 && tot) base = atomicAdd(&s_nchunk, tot);
    base = __shfl_sync(0xffffffffu, base, 31) + incl - nch;
    for (uint32_t k = 0; k < nch; ++k) {
      if (base + k < MAX_CHUNKS) s_chunk[base + k] = make_uint2(start + k * chunk_pts, min(end, start + (k + 1) * chunk_pts));
      else s_chunk_ovf = 1;
    }
  }
  __syncthreads();
  const bool again = s_chunk_ovf != 0 && chunk_pts < (1u << 30);
  __syncthreads();
  if (!again) break;
  }
  const uint32_t n_chunk = min(s_nchunk, (uint32_t)MAX_CHUNKS);

  double total = 0.0;
  uint32_t used = 0;
  // Attempt 1 takes every contributing landmark in one pass.  If the hash table fills up (more voxels with a second landmark
  // than slots), the rest of the attempt only counts; the count then sizes a multi-pass run in which pass p owns the voxels
  // whose hash falls in partition p, so that every pass fits with a load factor <= 1/2.
  volatile uint32_t *v_overflow = &s_overflow;
  uint2 *cq = s_cq[wid];
  uint32_t *uq = s_uq[wid];
  for (uint32_t npass = 1; npass <= 65536;) {
    total = 0.0;
    if (threadIdx.x == 0) { s_used = 0; s_coll = 0; s_overflow = 0; }
    for (uint32_t pass = 0; pass < npass; ++pass) {
      for (int i = threadIdx.x; i < BM_WORDS + SEC_SLOTS; i += FISHER_THREADS) dyn[i] = 0;
      if (threadIdx.x == 0) s_next = 0;
      __syncthreads();
      uint32_t my_used = 0, my_coll = 0;
      // Two steps per point.  The first runs on every point of the chunk: one transform into the pose frame, the visibility
      // test and the voxel read off it, the voxel's bit in the CTA bitmap.  A point closer than the guard to one of those
      // decisions' boundaries (1 or 2 in 100) waits in a per-warp stack for the reference's own arithmetic; a landmark whose
      // voxel already had one (1 in 8) waits in another for the hash-table step.  Both stacks are taken 32 entries at a time.
      uint32_t n_c = 0, n_u = 0;
      // hash-table step on the top `take` entries of the collision stack
      auto drain_collisions = [&]() {
        const uint32_t take = min(n_c, 32u), at0 = n_c - take;
        if (lane < take && !*v_overflow) {
          const uint2 e = cq[at0 + lane];
          const uint32_t outside = e.x >> 31, vox = e.x & 0x7fffffffu;
          // the landmark that raises a voxel's count to k adds information(voxel) * exp(1 - k^s), as a difference of
          // the host-tabulated prefix sums; inside the bitmap the table counts from the second landmark on
          const uint32_t cnt = table_count(tab, mix32(vox), (vox + 1u) | (outside ? OUTSIDE_BIT : 0u));
          const uint32_t k = cnt + 1u - outside;
          if (cnt == 0u) *v_overflow = 1;
          else if (k <= (uint32_t)CROWD_MAX) total += (double)__uint_as_float(e.y) * (__ldg(P.crowd + k) - __ldg(P.crowd + k - 1u));
        }
        n_c = at0;
        __syncwarp();
      };
      // one landmark per lane (vox < 0: none): bitmap, or onto the collision stack
      auto commit = [&](int vox, int bit) {
        bool collide = false;
        uint32_t c_key = 0;
        float inf_v = 0.f;
        bool mine = vox >= 0;
        if (npass != 1) mine = mine && ((mix32((uint32_t)vox) >> 12) & (npass - 1)) == pass;  // (uniform branch: one pass is the rule)
        if (mine) {
          ++my_used;
          if (!*v_overflow) {  // a lost attempt keeps counting only
            inf_v = __ldg(P.lut + vox);
            if (bit >= 0) {
              const uint32_t w = (uint32_t)bit >> 5, m = 1u << (bit & 31);
              uint32_t old = ((volatile uint32_t *)bm)[w];
              if (!(old & m)) old = atomicOr(&bm[w], m);
              // first landmark of its voxel: factor exp(1 - 1^s) = 1.  (In double: which thread meets which landmark changes from
              // run to run, and only a double sum rounds to the same float whatever the order; a float partial sum was tried.)
              if (!(old & m)) total += (double)inf_v;
              else { collide = true; c_key = (uint32_t)vox; }
            } else {
              collide = true;
              c_key = (uint32_t)vox | 0x80000000u;
            }
          }
        }
        const uint32_t cbal = __ballot_sync(0xffffffffu, collide);
        if (cbal) {
          if (collide) cq[n_c + __popc(cbal & ((1u << lane) - 1u))] = make_uint2(c_key, __float_as_uint(inf_v));
          n_c += __popc(cbal);
          my_coll += collide ? 1u : 0u;
          __syncwarp();
          if (n_c >= 32) drain_collisions();  // (the stack holds 64)
        }
      };
      // the reference's own arithmetic on the top entries of the stack of uncertain points
      auto drain_unsure = [&]() {
        const uint32_t take = min(n_u, 32u), at0 = n_u - take;
        int vox = -1, bit = -1;
        if (lane < take) {
          const float4 q = __ldg(P.pts + uq[at0 + lane]);
          if (point_visible(q, F, P)) vox = voxel_and_bit(q, F, P, xrow, &bit);
        }
        n_u = at0;
        __syncwarp();
        commit(vox, bit);
      };
      while (true) {
        uint32_t ci = 0;
        if (lane == 0) ci = atomicAdd(&s_next, 1u);
        ci = __shfl_sync(0xffffffffu, ci, 0);
        const bool last = ci >= n_chunk;
        const uint2 ch = last ? make_uint2(0, 0) : s_chunk[ci];
        for (uint32_t i0 = ch.x; i0 < ch.y; i0 += 32) {
          const uint32_t i = i0 + lane;
          int vox = -1, bit = -1;
          bool unsure = false;
          if (i < ch.y) {
            const float4 q = __ldg(P.pts + i);
            const float dx = q.x - F.px, dy = q.y - F.py, dz = q.z - F.pz;
            const float lx = __fmaf_rn(F.Rinv.m[0][0], dx, __fmaf_rn(F.Rinv.m[0][1], dy, F.Rinv.m[0][2] * dz));
            const float ly = __fmaf_rn(F.Rinv.m[1][0], dx, __fmaf_rn(F.Rinv.m[1][1], dy, F.Rinv.m[1][2] * dz));
            const float lz = __fmaf_rn(F.Rinv.m[2][0], dx, __fmaf_rn(F.Rinv.m[2][1], dy, F.Rinv.m[2][2] * dz));
            // camera coordinates (Xc, Yc, Zc) = (-ly, -lz, lx); u = fx Xc / Zc + cx in [0, max_u] <=> -cx Zc <= fx Xc <= (max_u - cx) Zc
            // (fminf / fmaxf drop NaNs: non-finite points must not reach them)
            const bool finite = fabsf(dx) + fabsf(dy) + fabsf(dz) < 1e6f;
            bool sure = !F.exact_only && finite && lx >= 0.05f, vis = false;
            if (sure) {
              // (fused by hand: the library is built with -fmad=false, and these feed a guarded decision only)
              const float au = -P.fx * ly, av = -P.fy * lz;
              const float eu = __fmaf_rn(4e-4f, lx, F.gu), ev = __fmaf_rn(4e-4f, lx, F.gv);  // + the roundings of u, v themselves (3 ulp of 10^3 px) times Zc
              const float u0 = __fmaf_rn(P.cx, lx, au), u1 = __fmaf_rn(P.max_u - P.cx, lx, -au);
              const float v0 = __fmaf_rn(P.cy, lx, av), v1 = __fmaf_rn(P.max_v - P.cy, lx, -av);
              const float inner = fminf(fminf(u0 - eu, u1 - eu), fminf(v0 - ev, v1 - ev));
              const float outer = fminf(fminf(u0 + eu, u1 + eu), fminf(v0 + ev, v1 + ev));
              vis = inner >= 0.f;
              sure = vis || outer < 0.f;
            } else if (!F.exact_only && finite && lx < -F.guard) {
              sure = true;  // behind the camera whichever evaluation is asked
            }
            if (sure && vis) {
              // voxel = round(l / step) per axis, outside the guard band of a half-integer
              const float rx = lx * P.inv_step_f, ry = ly * P.inv_step_f, rz = lz * P.inv_step_f;
              const float fx_ = rintf(rx), fy_ = rintf(ry), fz_ = rintf(rz);
              const float edge = fmaxf(fmaxf(fabsf(rx - fx_), fabsf(ry - fy_)), fabsf(rz - fz_));
              // (an index inside the table implies |r| < 512, which is what the guard's bound on the float product assumes; the
              // conversion saturates for huge values and `finite` has kept NaNs out)
              if (edge < 0.5f - F.gvox) {
                const int i0_ = (int)fx_ - P.org[0], i1_ = (int)fy_ - P.org[1], i2_ = (int)fz_ - P.org[2];
                if ((unsigned)i0_ < (unsigned)P.dims[0] && (unsigned)i1_ < (unsigned)P.dims[1] && (unsigned)i2_ < (unsigned)P.dims[2]) {
                  vox = (i0_ * P.dims[1] + i1_) * P.dims[2] + i2_;
                  if (!P.lut_dense) {
                    const float val = __ldg(P.lut + vox);
                    if (!(val == val)) vox = -1;
                  }
                  if (vox >= 0 && have_rows) {
                    const uint2 xr = s_xrow[i0_];
                    const int rows = (int)(xr.y >> 16), zr = i2_ - (int)(xr.y & 0xffffu);
                    if ((unsigned)zr < (unsigned)rows) bit = (int)xr.x + i1_ * rows + zr;
                  }
                }
              } else {
                sure = false;
              }
            }
            unsure = !sure;
          }
          const uint32_t ubal = __ballot_sync(0xffffffffu, unsure);
          if (ubal) {
            if (unsure) uq[n_u + __popc(ubal & ((1u << lane) - 1u))] = i;
            n_u += __popc(ubal);
            __syncwarp();
            if (n_u >= 32) drain_unsure();  // (the stack holds 64)
          }
          commit(vox, bit);
        }
        if (last) {
          while (n_u > 0) drain_unsure();
          while (n_c > 0) drain_collisions();
          break;
        }
      }
      if (my_used) atomicAdd(&s_used, my_used);
      if (my_coll) atomicAdd(&s_coll, my_coll);
      __syncthreads();
      if (s_overflow && npass > 1) break;  // npass == 1: finish the (only) pass so that the counts are complete
      __syncthreads();
    }
    used = s_used;
    const uint32_t coll = s_coll;
    const bool ovf = s_overflow != 0;
    __syncthreads();
    if (!ovf) break;
    if (npass == 1) {  // size the partitioning from the count of table entries at most needed: <= SEC_SLOTS / 2 per pass on average
      uint32_t want = 2;
      while ((uint64_t)want * (SEC_SLOTS / 2) < coll) want <<= 1;
      npass = want;
    } else {
      npass <<= 1;  // skewed partitions: halve them again
    }
  }
  for (int o = 16; o > 0; o >>= 1) total += __shfl_xor_sync(0xffffffffu, total, o);
  if (lane == 0) s_red[wid] = total;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0;
    for (uint32_t w = 0; w < nw; ++w) t += s_red[w];
    const float information = (float)t;
    const bool is_safe = (double)information > P.threshold;  // FisherInfoManager.cpp:112-114
    if (info_out) info_out[pi] = information;
    if (safe_out) safe_out[pi] = is_safe ? 1 : 0;
    if (nvis_out) nvis_out[pi] = used;
    if (rec_out) {  // the frontier's record carries the Fisher information of its best-yaw pose
      const uint32_t f = pi / P.poses_per_frontier, w = pi - f * P.poses_per_frontier;
      if (P.poses_per_frontier == 1 || (int32_t)w == arrival[f].max_index) {
        rec_out[f].fisher = information;
        rec_out[f].flags |= 8u | (is_safe ? 16u : 0u);
      }
    }
  }
}

bool fisher_ready(const fitslam_ctx *c) { return c->fi && c->fi->have_lut && c->fi->have_points; }

static fitslam_status fisher_params(fitslam_ctx *c, FisherParams &P) {
  FisherState *f = c->fi;
  if (!f || !f->have_lut) return fail(c, FITSLAM_E_NO_LOOKUP_TABLE, "Cannot load lookup table. Please generate it first.");
  if (!f->have_points) return fail(c, FITSLAM_E_NO_MAP_POINTS, "no map points set");
  memset(&P, 0, sizeof(P));
  P.pts = f->pts.p;
  P.bin_start = f->bin_start.p;
  P.grid = BinGrid{f->bx0, f->by0, 1.0f / f->bin_size, f->bin_size, f->nbx, f->nby};
  P.zmin = f->zmin;
  P.zmax = f->zmax;
  P.lut = f->lut.p;
  for (int a = 0; a < 3; ++a) { P.dims[a] = f->dims[a]; P.org[a] = f->org[a]; }
  P.crowd = f->crowd.p;
  P.step = (double)c->prm.voxel_step;
  P.inv_step = 1 / P.step;
  P.inv_step_f = (float)P.inv_step;
  P.reach = lut_reach(c);
  P.threshold = c->prm.fisher_information_threshold;
  P.fx = c->prm.fx; P.fy = c->prm.fy; P.cx = c->prm.cx; P.cy = c->prm.cy;
  P.max_u = c->prm.img_max_u; P.max_v = c->prm.img_max_v;
  P.lut_dense = f->lut_dense ? 1 : 0;
  P.xrow = f->xrow.p;
  if (!f->attr_set) {  // (per device)
    FS_CUDA(c, cudaFuncSetAttribute(fisher_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, FISHER_DYN_SMEM));
    f->attr_set = true;
  }
  return FITSLAM_OK;
}

// Scores poses [first, first + count) of the n resident poses.
fitslam_status fisher_score_device(fitslam_ctx *c, const fitslam_pose *d_poses, uint32_t n, uint32_t first, uint32_t count,
                                   float *d_info, uint8_t *d_safe, uint32_t *d_nvis, ScoreRecord *d_rec,
                                   const fitslam_arrival *d_arrival, uint32_t poses_per_frontier) {
  FisherParams P;
  const fitslam_status st = fisher_params(c, P);
  if (st != FITSLAM_OK) return st;
  if (n == 0 || count == 0) return FITSLAM_OK;
  P.n_poses = n;
  P.first = first;
  P.poses_per_frontier = poses_per_frontier ? poses_per_frontier : 1;
  fisher_kernel<<<count, FISHER_THREADS, FISHER_DYN_SMEM, c->stream>>>(P, d_poses, d_info, d_safe, d_nvis, d_rec, d_arrival);
  ++c->launches;
  FS_CUDA(c, cudaGetLastError());
  return FITSLAM_OK;
}

// Scores the candidate poses (frontier * poses_per_frontier + window) of the frontiers of a slice; CTAs of slots outside
// the slice, or of frontiers another launch scores, return at once.
fitslam_status fisher_score_slice(fitslam_ctx *c, cudaStream_t stream, const fitslam_pose *d_poses, uint32_t poses_per_frontier,
                                  const ScoreSlice &slice, float *d_info, uint8_t *d_safe, uint32_t *d_nvis, ScoreRecord *d_rec,
                                  const fitslam_arrival *d_arrival) {
  FisherParams P;
  const fitslam_status st = fisher_params(c, P);
  if (st != FITSLAM_OK) return st;
  const uint64_t slots = slice.work_count ? slice.work_cap : (slice.range ? slice.cap : slice.count);
  if (slots == 0) return FITSLAM_OK;
  P.poses_per_frontier = poses_per_frontier ? poses_per_frontier : 1;
  P.use_slice = 1;
  P.slice = slice;
  fisher_kernel<<<(unsigned)(slots * P.poses_per_frontier), FISHER_THREADS, FISHER_DYN_SMEM, stream>>>(P, d_poses, d_info, d_safe, d_nvis,
                                                                                                     d_rec, d_arrival);
  ++c->launches;
  FS_CUDA(c, cudaGetLastError());
  return FITSLAM_OK;
}

}  // namespace fitslam

extern "C" fitslam_status fitslam_score_fisher(fitslam_ctx *c, const fitslam_pose *poses, uint32_t n, float *information,
                                               uint8_t *safe, uint32_t *n_visible) {
  if (!c || (!poses && n)) return fail(c, FITSLAM_E_INVALID, "score_fisher: null argument");
  cudaSetDevice(c->device);
  FisherState *f = fstate(c);
  if (!f->have_lut) return fail(c, FITSLAM_E_NO_LOOKUP_TABLE, "Cannot load lookup table. Please generate it first.");
  if (!f->have_points) return fail(c, FITSLAM_E_NO_MAP_POINTS, "no map points set");
  if (n == 0) return FITSLAM_OK;
  FS_CUDA(c, f->poses.reserve(n));
  FS_CUDA(c, f->info.reserve(n));
  FS_CUDA(c, f->safe.reserve(n));
  FS_CUDA(c, f->nvis.reserve(n));
  FS_CUDA(c, cudaMemcpyAsync(f->poses.p, poses, (size_t)n * sizeof(fitslam_pose), cudaMemcpyDefault, c->stream));
  if (c->nranks > 1) {
    FS_CUDA(c, cudaMemsetAsync(f->info.p, 0, (size_t)n * 4, c->stream));
    FS_CUDA(c, cudaMemsetAsync(f->safe.p, 0, n, c->stream));
    FS_CUDA(c, cudaMemsetAsync(f->nvis.p, 0, (size_t)n * 4, c->stream));
  }
  uint32_t first = 0, count = n;
  shard_range(c, n, &first, &count);
  fitslam_status st = fisher_score_device(c, f->poses.p, n, first, count, f->info.p, f->safe.p, f->nvis.p, nullptr, nullptr, 1);
  if (st != FITSLAM_OK) return st;
  if (information) FS_CUDA(c, cudaMemcpyAsync(information, f->info.p, (size_t)n * 4, cudaMemcpyDefault, c->stream));
  if (safe) FS_CUDA(c, cudaMemcpyAsync(safe, f->safe.p, n, cudaMemcpyDefault, c->stream));
  if (n_visible) FS_CUDA(c, cudaMemcpyAsync(n_visible, f->nvis.p, (size_t)n * 4, cudaMemcpyDefault, c->stream));
  FS_CUDA(c, cudaStreamSynchronize(c->stream));
  return FITSLAM_OK;
}

// ================================================================================================ GetLandmarksInView
// The service the Fisher plugin calls (exhaustive mode): ORBSLAM3Interface::mapPointsVisibleFromPose
// (ORBW/src/orb_slam3_interface.cpp:188-317) over every map point, each visible one converted as the node does before it
// fills the response (ORBW/src/rgbd/rgbd-slam-node.cpp:406-429), in map-point order.  Ordered compaction in three launches
// over the caller's (unsorted) points: per-tile counts, a scan of the tile counts, and the writes.
namespace fitslam {

static constexpr int LIV_THREADS = 256, LIV_PER_THREAD = 4, LIV_TILE = LIV_THREADS * LIV_PER_THREAD;

__device__ __forceinline__ uint32_t liv_visible_mask(const float *__restrict__ xyz, uint64_t n, uint64_t p0, const PoseFrame &F,
                                                     const FisherParams &P, float4 q[LIV_PER_THREAD]) {
  uint32_t m = 0;
  if (p0 + LIV_PER_THREAD <= n) {  // 12 consecutive floats, 48-byte aligned
    const float4 a = __ldg(reinterpret_cast<const float4 *>(xyz + 3 * p0));
    const float4 b = __ldg(reinterpret_cast<const float4 *>(xyz + 3 * p0) + 1);
    const float4 c = __ldg(reinterpret_cast<const float4 *>(xyz + 3 * p0) + 2);
    q[0] = make_float4(a.x, a.y, a.z, 0.f);
    q[1] = make_float4(a.w, b.x, b.y, 0.f);
    q[2] = make_float4(b.z, b.w, c.x, 0.f);
    q[3] = make_float4(c.y, c.z, c.w, 0.f);
#pragma unroll
    for (int k = 0; k < LIV_PER_THREAD; ++k) m |= point_visible(q[k], F, P) ? (1u << k) : 0u;
  } else {
    for (int k = 0; k < LIV_PER_THREAD; ++k) {
      q[k] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (p0 + k < n) {
        q[k] = make_float4(xyz[3 * (p0 + k)], xyz[3 * (p0 + k) + 1], xyz[3 * (p0 + k) + 2], 0.f);
        m |= point_visible(q[k], F, P) ? (1u << k) : 0u;
      }
    }
  }
  return m;
}

// exclusive offset of this thread's count within the CTA, and the CTA total
__device__ __forceinline__ uint32_t liv_block_scan(uint32_t cnt, uint32_t *warp_sums, uint32_t *total) {
  const uint32_t lane = threadIdx.x & 31u, wid = threadIdx.x >> 5;
  uint32_t incl = cnt;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const uint32_t v = __shfl_up_sync(0xffffffffu, incl, d);
    if (lane >= (uint32_t)d) incl += v;
  }
  if (lane == 31) warp_sums[wid] = incl;
  __syncthreads();
  uint32_t base = 0, tot = 0;
  for (uint32_t w = 0; w < LIV_THREADS / 32; ++w) {
    const uint32_t s = warp_sums[w];
    if (w < wid) base += s;
    tot += s;
  }
  *total = tot;
  return base + incl - cnt;
}

// pass 0: tile_count[b]; pass 1: writes at tile_offset[b]
template <int PASS>
__global__ void __launch_bounds__(LIV_THREADS) landmarks_in_view_kernel(const FisherParams P, const fitslam_pose *__restrict__ pose,
                                                                         const float *__restrict__ xyz, uint64_t n, uint32_t *tile_count,
                                                                         const uint32_t *__restrict__ tile_offset, float *__restrict__ out,
                                                                         uint32_t capacity) {
  __shared__ PoseFrame F;
  __shared__ uint32_t warp_sums[LIV_THREADS / 32];
  if (threadIdx.x == 0) build_pose_frame(*pose, P, F);
  __syncthreads();
  const uint64_t p0 = ((uint64_t)blockIdx.x * LIV_THREADS + threadIdx.x) * LIV_PER_THREAD;
  float4 q[LIV_PER_THREAD];
  const uint32_t m = liv_visible_mask(xyz, n, p0, F, P, q);
  uint32_t total;
  const uint32_t excl = liv_block_scan((uint32_t)__popc(m), warp_sums, &total);
  if (PASS == 0) {
    if (threadIdx.x == 0) tile_count[blockIdx.x] = total;
    return;
  }
  uint32_t dst = tile_offset[blockIdx.x] + excl;
#pragma unroll
  for (int k = 0; k < LIV_PER_THREAD; ++k) {
    if (!((m >> k) & 1u)) continue;
    if (dst < capacity) {
      // vector3fORBToROS of the ORB-frame point, then affinePosToMap * position (rgbd-slam-node.cpp:419-425).  The signed
      // permutation is multiplied out as the reference does: for a finite point it returns the ROS triple, for a
      // non-finite one (which passes the cull: every comparison with NaN is false) 0 * inf turns coordinates into NaN
      M3f orb2ros;
      const float b[9] = {0, 0, 1, -1, 0, 0, 0, -1, 0};
      for (int i = 0; i < 9; ++i) orb2ros.m[i / 3][i % 3] = b[i];
      V3f Pw; Pw.v[0] = -q[k].y; Pw.v[1] = -q[k].z; Pw.v[2] = q[k].x;
      const V3f pos = mulv(orb2ros, Pw);
      const V3f lp = mulv(F.Rinv, pos);
      out[3 * (size_t)dst] = lp.v[0] + F.tinv.v[0];
      out[3 * (size_t)dst + 1] = lp.v[1] + F.tinv.v[1];
      out[3 * (size_t)dst + 2] = lp.v[2] + F.tinv.v[2];
    }
    ++dst;
  }
}

// exclusive scan of the tile counts in place (one CTA, chunks of 1024 with a running carry); the total goes to *total
__global__ void __launch_bounds__(1024) liv_scan_kernel(uint32_t *tile, uint32_t n_tiles, uint32_t *total) {
  __shared__ uint32_t ws[32];
  __shared__ uint32_t carry;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  const uint32_t lane = threadIdx.x & 31u, wid = threadIdx.x >> 5;
  for (uint32_t base = 0; base < n_tiles; base += 1024) {
    const uint32_t i = base + threadIdx.x;
    const uint32_t v = i < n_tiles ? tile[i] : 0u;
    uint32_t incl = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const uint32_t u = __shfl_up_sync(0xffffffffu, incl, d);
      if (lane >= (uint32_t)d) incl += u;
    }
    if (lane == 31) ws[wid] = incl;
    __syncthreads();
    uint32_t wb = 0, tot = 0;
    for (uint32_t w = 0; w < 32; ++w) {
      const uint32_t s = ws[w];
      if (w < wid) wb += s;
      tot += s;
    }
    const uint32_t c0 = carry;
    if (i < n_tiles) tile[i] = c0 + wb + incl - v;
    __syncthreads();
    if (threadIdx.x == 0) carry = c0 + tot;
    __syncthreads();
  }
  if (threadIdx.x == 0) *total = carry;
}

}  // namespace fitslam

extern "C" fitslam_status fitslam_landmarks_in_view(fitslam_ctx *c, const fitslam_pose *pose, float *out_xyz, uint32_t capacity,
                                                    uint32_t *count) {
  if (!c || !pose || !count || (!out_xyz && capacity)) return fail(c, FITSLAM_E_INVALID, "landmarks_in_view: null argument");
  cudaSetDevice(c->device);
  FisherState *f = fstate(c);
  if (!f->have_points) return fail(c, FITSLAM_E_NO_MAP_POINTS, "no map points set");
  *count = 0;
  const uint64_t n = f->n_points;
  if (n == 0) return FITSLAM_OK;
  FisherParams P;
  memset(&P, 0, sizeof(P));
  P.fx = c->prm.fx; P.fy = c->prm.fy; P.cx = c->prm.cx; P.cy = c->prm.cy;
  P.max_u = c->prm.img_max_u; P.max_v = c->prm.img_max_v;
  const uint32_t n_tiles = (uint32_t)((n + LIV_TILE - 1) / LIV_TILE);
  FS_CUDA(c, f->poses.reserve(1));
  FS_CUDA(c, f->nvis.reserve((size_t)n_tiles + 1));
  FS_CUDA(c, cudaMemcpyAsync(f->poses.p, pose, sizeof(fitslam_pose), cudaMemcpyDefault, c->stream));
  uint32_t *tiles = f->nvis.p, *total = f->nvis.p + n_tiles;
  landmarks_in_view_kernel<0><<<n_tiles, LIV_THREADS, 0, c->stream>>>(P, f->poses.p, f->raw.p, n, tiles, nullptr, nullptr, 0);
  liv_scan_kernel<<<1, 1024, 0, c->stream>>>(tiles, n_tiles, total);
  c->launches += 2;
  const bool out_dev = [&] {
    cudaPointerAttributes a;
    const bool d = out_xyz && cudaPointerGetAttributes(&a, out_xyz) == cudaSuccess && (a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged);
    cudaGetLastError();
    return d;
  }();
  float *d_out = out_xyz;
  if (!out_dev && capacity) {
    FS_CUDA(c, f->info.reserve(3 * (size_t)std::min<uint64_t>(capacity, n)));
    d_out = f->info.p;
  }
  const uint32_t cap = (uint32_t)std::min<uint64_t>(capacity, n);
  if (cap) {
    landmarks_in_view_kernel<1><<<n_tiles, LIV_THREADS, 0, c->stream>>>(P, f->poses.p, f->raw.p, n, nullptr, tiles, d_out, cap);
    ++c->launches;
  }
  FS_CUDA(c, cudaGetLastError());
  FS_CUDA(c, cudaMemcpyAsync(count, total, sizeof(uint32_t), cudaMemcpyDeviceToHost, c->stream));
  FS_CUDA(c, cudaStreamSynchronize(c->stream));
  if (!out_dev && cap) {
    const uint32_t m = std::min(*count, cap);
    if (m) FS_CUDA(c, cudaMemcpyAsync(out_xyz, d_out, 3 * (size_t)m * sizeof(float), cudaMemcpyDeviceToHost, c->stream));
    FS_CUDA(c, cudaStreamSynchronize(c->stream));
  }
  if (*count > capacity) return fail(c, FITSLAM_E_CAPACITY, "landmarks_in_view: more visible landmarks than capacity (count holds the number)");
  return FITSLAM_OK;
}

// ================================================================================================ keyframe-structured query
// ORBSLAM3Interface::mapPointsVisibleFromPose as the service runs it (ORBW/src/orb_slam3_interface.cpp:188-317): keyframes in
// the caller's order, a per-keyframe prefilter unless exhaustive_search (:240-267), every map point once (:263-265), and the
// cutoff "more than maxLandmarks collected: stop" checked at the start of a keyframe (:270-271).  The loop is sequential in the
// reference; what it computes is order-free up to a prefix sum:
//   pass[k]      the prefilter of keyframe k (independent per keyframe)
//   first[p]     the first entry of map point p among the passing keyframes (atomicMin of entry indices): the entry that
//                processes p.  The cutoff removes a suffix of keyframes, and the first entry of p inside the kept prefix is
//                first[p] or nothing, so it can be taken before the cutoff is known
//   cnt[k]       visible fresh landmarks of keyframe k;  before[k] = sum of cnt over j < k
//   processed[k] = pass[k] and before[k] <= maxLandmarks (monotone, so this is the reference's `break`)
// and the response is the ordered compaction of the visible fresh entries of processed keyframes.
namespace fitslam {

static constexpr uint32_t KF_PASS = 0, KF_CNT = 1, KF_BEFORE = 2, KF_PROC = 3;

// Eigen 3.4 eulerAngles(2, 0, 1) (odd = 0, i = 2, j = 0, k = 1) + rotationORBToEulerROS (ORBW/src/type_conversion.cpp:143-153),
// component 2
__device__ float kf_relative_yaw(const M3f &R) {
  float res0 = atan2f(R.m[0][1], R.m[1][1]);
  if (res0 > 0.0f) res0 -= (float)M_PI;
  const float s1 = sinf(res0), c1 = cosf(res0);
  float res2 = atan2f(s1 * R.m[1][2] - c1 * R.m[0][2], c1 * R.m[0][0] - s1 * R.m[1][0]);
  res2 = -res2;
  float e2 = -res2;
  e2 = e2 < 0 ? (float)((2 * M_PI) + (double)e2) : e2;
  return e2;
}

__global__ void kf_prefilter_kernel(const FisherParams P, const fitslam_pose *__restrict__ pose, const fitslam_keyframe *__restrict__ kfs,
                                    uint32_t n_kf, float max_distance, int exhaustive, uint32_t *state) {
  __shared__ PoseFrame F;
  if (threadIdx.x == 0) build_pose_frame(*pose, P, F);
  __syncthreads();
  const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n_kf) return;
  bool pass = true;
  if (!exhaustive) {
    M3f Rwc;
    V3f twc;
    for (int a = 0; a < 3; ++a) {
      for (int b = 0; b < 3; ++b) Rwc.m[a][b] = kfs[k].Rwc[3 * a + b];
      twc.v[a] = kfs[k].twc[a];
    }
    const M3f relR = mulm(F.Rcw, Rwc);  // rel_T_CamToKF = Tcw * pKF->GetPoseInverse()
    const V3f rt = mulv(F.Rcw, twc);
    const float tx = rt.v[0] + F.tcw.v[0], ty = rt.v[1] + F.tcw.v[1], tz = rt.v[2] + F.tcw.v[2];
    const float yaw = kf_relative_yaw(relR);
    const float dist = sqrtf(dot3(tx, tx, ty, ty, tz, tz));
    if ((double)dist >= (double)max_distance * 2.0) {
      pass = false;
    } else {
      if (dist > max_distance) {
        if (tz < 0.0f) pass = false;
        else if (tz > 0.0f && (((double)yaw >= 0 && (double)yaw < M_PI / 2.0) || ((double)yaw > 3.0 * M_PI / 2.0 && (double)yaw < 2 * M_PI)))
          pass = false;
      }
      if (pass && dist <= max_distance && tz < 0.0f && ((double)yaw >= 2.0944 && (double)yaw < 4.18879)) pass = false;
    }
  }
  state[4 * k + KF_PASS] = pass ? 1u : 0u;
  state[4 * k + KF_CNT] = 0u;
}

__global__ void kf_first_kernel(const uint32_t *__restrict__ entry_pt, const uint32_t *__restrict__ entry_kf, uint64_t n_entries,
                                const uint32_t *__restrict__ state, uint32_t *first) {
  const uint64_t e = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= n_entries) return;
  if (state[4 * entry_kf[e] + KF_PASS]) atomicMin(&first[entry_pt[e]], (uint32_t)e);
}

__global__ void kf_visible_kernel(const FisherParams P, const fitslam_pose *__restrict__ pose, const float *__restrict__ xyz,
                                  const uint32_t *__restrict__ entry_pt, const uint32_t *__restrict__ entry_kf, uint64_t n_entries,
                                  const uint32_t *__restrict__ first, uint32_t *state, uint8_t *vis) {
  __shared__ PoseFrame F;
  if (threadIdx.x == 0) build_pose_frame(*pose, P, F);
  __syncthreads();
  const uint64_t e = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= n_entries) return;
  const uint32_t pt = entry_pt[e], k = entry_kf[e];
  bool v = false;
  if (state[4 * k + KF_PASS] && first[pt] == (uint32_t)e) {
    const float4 q = make_float4(xyz[3 * (size_t)pt], xyz[3 * (size_t)pt + 1], xyz[3 * (size_t)pt + 2], 0.f);
    v = point_visible(q, F, P);
  }
  vis[e] = v ? 1 : 0;
  if (v) atomicAdd(&state[4 * k + KF_CNT], 1u);
}

// one CTA: before[k] = landmarks collected when the loop reaches keyframe k; processed[k]; *checked = numKFsChecked
__global__ void __launch_bounds__(1024) kf_cutoff_kernel(uint32_t *state, uint32_t n_kf, long long max_landmarks, uint32_t *checked) {
  __shared__ uint32_t ws[32];
  __shared__ uint32_t carry, s_checked, s_break;
  if (threadIdx.x == 0) { carry = 0; s_checked = 0; s_break = 0; }
  __syncthreads();
  const uint32_t lane = threadIdx.x & 31u, wid = threadIdx.x >> 5;
  for (uint32_t base = 0; base < n_kf; base += 1024) {
    const uint32_t k = base + threadIdx.x;
    const bool pass = k < n_kf && state[4 * k + KF_PASS];
    const uint32_t v = pass ? state[4 * k + KF_CNT] : 0u;
    uint32_t incl = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const uint32_t u = __shfl_up_sync(0xffffffffu, incl, d);
      if (lane >= (uint32_t)d) incl += u;
    }
    if (lane == 31) ws[wid] = incl;
    __syncthreads();
    uint32_t wb = 0, tot = 0;
    for (uint32_t w = 0; w < 32; ++w) {
      const uint32_t x = ws[w];
      if (w < wid) wb += x;
      tot += x;
    }
    const uint32_t c0 = carry;
    if (k < n_kf) {
      const uint32_t before = c0 + wb + incl - v;
      const bool proc = pass && (long long)before <= max_landmarks;
      state[4 * k + KF_BEFORE] = before;
      state[4 * k + KF_PROC] = proc ? 1u : 0u;
      if (proc) atomicAdd(&s_checked, 1u);
      // numKFsChecked++ comes before the cutoff test (:269-271): the keyframe that breaks the loop is counted too.  It is the
      // first passing keyframe that is not processed (before[] is monotone: the passing keyframes after it are never reached)
      if (pass && !proc) s_break = 1u;
    }
    __syncthreads();
    if (threadIdx.x == 0) carry = c0 + tot;
    __syncthreads();
  }
  if (threadIdx.x == 0) *checked = s_checked + s_break;
}

// ordered compaction of the response; pass 0: tile counts, pass 1: writes (same transform as landmarks_in_view_kernel)
template <int PASS>
__global__ void __launch_bounds__(LIV_THREADS) kf_compact_kernel(const FisherParams P, const fitslam_pose *__restrict__ pose,
                                                                  const float *__restrict__ xyz, const uint32_t *__restrict__ entry_pt,
                                                                  const uint32_t *__restrict__ entry_kf, uint64_t n_entries,
                                                                  const uint32_t *__restrict__ state, const uint8_t *__restrict__ vis,
                                                                  uint32_t *tile_count, const uint32_t *__restrict__ tile_offset,
                                                                  float *__restrict__ out, uint32_t capacity) {
  __shared__ PoseFrame F;
  __shared__ uint32_t warp_sums[LIV_THREADS / 32];
  if (PASS == 1) {
    if (threadIdx.x == 0) build_pose_frame(*pose, P, F);
    __syncthreads();
  }
  const uint64_t e0 = ((uint64_t)blockIdx.x * LIV_THREADS + threadIdx.x) * LIV_PER_THREAD;
  uint32_t m = 0;
  for (int k = 0; k < LIV_PER_THREAD; ++k) {
    const uint64_t e = e0 + k;
    if (e < n_entries && vis[e] && state[4 * entry_kf[e] + KF_PROC]) m |= 1u << k;
  }
  uint32_t total;
  const uint32_t excl = liv_block_scan((uint32_t)__popc(m), warp_sums, &total);
  if (PASS == 0) {
    if (threadIdx.x == 0) tile_count[blockIdx.x] = total;
    return;
  }
  uint32_t dst = tile_offset[blockIdx.x] + excl;
  for (int k = 0; k < LIV_PER_THREAD; ++k) {
    if (!((m >> k) & 1u)) continue;
    if (dst < capacity) {
      const uint32_t pt = entry_pt[e0 + k];
      const float qx = xyz[3 * (size_t)pt], qy = xyz[3 * (size_t)pt + 1], qz = xyz[3 * (size_t)pt + 2];
      // vector3fORBToROS of the ORB-frame point, then affinePosToMap * position (rgbd-slam-node.cpp:419-425)
      M3f orb2ros;
      const float b[9] = {0, 0, 1, -1, 0, 0, 0, -1, 0};
      for (int i = 0; i < 9; ++i) orb2ros.m[i / 3][i % 3] = b[i];
      V3f Pw; Pw.v[0] = -qy; Pw.v[1] = -qz; Pw.v[2] = qx;
      const V3f pos = mulv(orb2ros, Pw);
      const V3f lp = mulv(F.Rinv, pos);
      out[3 * (size_t)dst] = lp.v[0] + F.tinv.v[0];
      out[3 * (size_t)dst + 1] = lp.v[1] + F.tinv.v[1];
      out[3 * (size_t)dst + 2] = lp.v[2] + F.tinv.v[2];
    }
    ++dst;
  }
}

// getVoxelCoordinate + table entry for a landmark given in the pose frame (the client side of the service response)
__device__ __forceinline__ int voxel_of_local_point(float cx, float cy, float cz, const FisherParams &P) {
  const float c3[3] = {cx, cy, cz};
  int idx[3];
  bool ok = true;
#pragma unroll
  for (int a = 0; a < 3; ++a) {
    const double rd = round((double)c3[a] * P.inv_step);  // FisherInfoManager.hpp:108-123
    const int r = (rd >= -2147483000.0 && rd <= 2147483000.0) ? (int)rd : INT_MIN;  // also rejects NaN
    ok = ok && r >= P.org[a] && r < P.org[a] + P.dims[a];
    idx[a] = r - P.org[a];
  }
  if (!ok) return -1;
  const int lin = (idx[0] * P.dims[1] + idx[1]) * P.dims[2] + idx[2];
  const float val = __ldg(P.lut + lin);
  return (val == val) ? lin : -1;
}

__global__ void kf_fisher_count_kernel(const FisherParams P, const float *__restrict__ lm, uint32_t n, uint32_t *vox_cnt, int32_t *lm_vox) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int lin = voxel_of_local_point(lm[3 * (size_t)i], lm[3 * (size_t)i + 1], lm[3 * (size_t)i + 2], P);
  int rank = -1;
  if (lin >= 0) rank = (int)atomicAdd(&vox_cnt[lin], 1u);
  lm_vox[2 * (size_t)i] = lin;
  lm_vox[2 * (size_t)i + 1] = rank;
}

// the first landmark of a voxel adds the whole voxel: information(v) * sum_{k <= n} exp(1 - k^s); per-block sums in a fixed order
__global__ void __launch_bounds__(256) kf_fisher_sum_kernel(const FisherParams P, uint32_t n, const uint32_t *__restrict__ vox_cnt,
                                                            const int32_t *__restrict__ lm_vox, double *part, uint32_t *used) {
  __shared__ double sd[256];
  __shared__ uint32_t su[256];
  const uint32_t i = blockIdx.x * 256 + threadIdx.x;
  double t = 0.0;
  uint32_t u = 0;
  if (i < n) {
    const int lin = lm_vox[2 * (size_t)i], rank = lm_vox[2 * (size_t)i + 1];
    if (lin >= 0) {
      u = 1;
      if (rank == 0) t = (double)__ldg(P.lut + lin) * __ldg(P.crowd + min(vox_cnt[lin], (uint32_t)CROWD_MAX));
    }
  }
  sd[threadIdx.x] = t;
  su[threadIdx.x] = u;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if ((int)threadIdx.x < o) { sd[threadIdx.x] += sd[threadIdx.x + o]; su[threadIdx.x] += su[threadIdx.x + o]; }
    __syncthreads();
  }
  if (threadIdx.x == 0) { part[blockIdx.x] = sd[0]; atomicAdd(used, su[0]); }
}

__global__ void kf_fisher_final_kernel(const double *__restrict__ part, uint32_t n_blocks, double threshold, float *info, uint8_t *safe) {
  double t = 0.0;
  for (uint32_t b = 0; b < n_blocks; ++b) t += part[b];
  const float information = (float)t;
  *info = information;
  *safe = (double)information > threshold ? 1 : 0;  // FisherInfoManager.cpp:112-114
}

// the query on the context stream; leaves the response in f->kf_out, *count / *checked on the host (one sync)
static fitslam_status kf_query(fitslam_ctx *c, const fitslam_pose *pose, long long max_landmarks, float max_distance, int exhaustive,
                               uint32_t *count, uint32_t *checked) {
  FisherState *f = fstate(c);
  *count = 0;
  *checked = 0;
  if (!f->have_points) return fail(c, FITSLAM_E_NO_MAP_POINTS, "no map points set");
  if (!f->have_kf) return fail(c, FITSLAM_E_NO_MAP_POINTS, "no keyframes set (fitslam_set_keyframes, after fitslam_set_map_points)");
  if (f->n_kf == 0) return FITSLAM_OK;
  FisherParams P;
  memset(&P, 0, sizeof(P));
  P.fx = c->prm.fx; P.fy = c->prm.fy; P.cx = c->prm.cx; P.cy = c->prm.cy;
  P.max_u = c->prm.img_max_u; P.max_v = c->prm.img_max_v;
  const uint64_t ne = f->n_entries;
  const uint32_t n_tiles = (uint32_t)((ne + LIV_TILE - 1) / LIV_TILE);
  const uint32_t out_cap = (uint32_t)std::min<uint64_t>(ne, f->n_points);  // every map point at most once
  FS_CUDA(c, f->poses.reserve(1));
  FS_CUDA(c, f->kf_first.reserve(f->n_points + 1));
  FS_CUDA(c, f->kf_state.reserve(4 * (size_t)f->n_kf));
  FS_CUDA(c, f->kf_vis.reserve(ne + 1));
  FS_CUDA(c, f->kf_tiles.reserve((size_t)n_tiles + 4));
  FS_CUDA(c, f->kf_out.reserve(3 * (size_t)out_cap + 3));
  FS_CUDA(c, cudaMemcpyAsync(f->poses.p, pose, sizeof(fitslam_pose), cudaMemcpyDefault, c->stream));
  FS_CUDA(c, cudaMemsetAsync(f->kf_first.p, 0xFF, (f->n_points + 1) * sizeof(uint32_t), c->stream));
  uint32_t *tiles = f->kf_tiles.p, *total = tiles + n_tiles, *d_checked = tiles + n_tiles + 1;
  kf_prefilter_kernel<<<(f->n_kf + 255) / 256, 256, 0, c->stream>>>(P, f->poses.p, f->kf.p, f->n_kf, max_distance, exhaustive, f->kf_state.p);
  ++c->launches;
  if (ne) {
    const unsigned eg = (unsigned)((ne + 255) / 256);
    kf_first_kernel<<<eg, 256, 0, c->stream>>>(f->kf_entry_pt.p, f->kf_entry_kf.p, ne, f->kf_state.p, f->kf_first.p);
    kf_visible_kernel<<<eg, 256, 0, c->stream>>>(P, f->poses.p, f->raw.p, f->kf_entry_pt.p, f->kf_entry_kf.p, ne, f->kf_first.p,
                                                 f->kf_state.p, f->kf_vis.p);
    c->launches += 2;
  }
  kf_cutoff_kernel<<<1, 1024, 0, c->stream>>>(f->kf_state.p, f->n_kf, max_landmarks, d_checked);
  ++c->launches;
  if (ne) {
    kf_compact_kernel<0><<<n_tiles, LIV_THREADS, 0, c->stream>>>(P, f->poses.p, f->raw.p, f->kf_entry_pt.p, f->kf_entry_kf.p, ne,
                                                                 f->kf_state.p, f->kf_vis.p, tiles, nullptr, nullptr, 0);
    liv_scan_kernel<<<1, 1024, 0, c->stream>>>(tiles, n_tiles, total);
    kf_compact_kernel<1><<<n_tiles, LIV_THREADS, 0, c->stream>>>(P, f->poses.p, f->raw.p, f->kf_entry_pt.p, f->kf_entry_kf.p, ne,
                                                                 f->kf_state.p, f->kf_vis.p, nullptr, tiles, f->kf_out.p, out_cap);
    c->launches += 3;
  } else {
    FS_CUDA(c, cudaMemsetAsync(total, 0, sizeof(uint32_t), c->stream));
  }
  FS_CUDA(c, cudaGetLastError());
  uint32_t h[2] = {0, 0};
  FS_CUDA(c, cudaMemcpyAsync(h, total, 2 * sizeof(uint32_t), cudaMemcpyDeviceToHost, c->stream));
  FS_CUDA(c, cudaStreamSynchronize(c->stream));
  *count = h[0];
  *checked = h[1];
  return FITSLAM_OK;
}

}  // namespace fitslam

extern "C" fitslam_status fitslam_set_keyframes(fitslam_ctx *c, const fitslam_keyframe *kfs, uint32_t n_kf, const uint32_t *point_index,
                                                uint64_t n_entries) {
  if (!c || (!kfs && n_kf) || (!point_index && n_entries)) return fail(c, FITSLAM_E_INVALID, "set_keyframes: null argument");
  if (n_entries >= (1ull << 32) - 1) return fail(c, FITSLAM_E_INVALID, "set_keyframes: more than 2^32 - 2 entries");
  cudaSetDevice(c->device);
  FisherState *f = fstate(c);
  if (!f->have_points) return fail(c, FITSLAM_E_NO_MAP_POINTS, "set_keyframes: set the map points first (the lists index them)");
  f->have_kf = false;
  // the lists must tile [0, n_entries) in keyframe order (the service order is keyframe by keyframe), indices inside the points
  std::vector<uint32_t> entry_kf((size_t)n_entries);
  uint64_t at = 0;
  for (uint32_t k = 0; k < n_kf; ++k) {
    if (kfs[k].first != at || (uint64_t)kfs[k].first + kfs[k].count > n_entries)
      return fail(c, FITSLAM_E_INVALID, "set_keyframes: keyframe " + std::to_string(k) + ": the lists must follow each other in point_index");
    for (uint32_t i = 0; i < kfs[k].count; ++i) entry_kf[(size_t)at + i] = k;
    at += kfs[k].count;
  }
  if (at != n_entries) return fail(c, FITSLAM_E_INVALID, "set_keyframes: the lists cover fewer entries than n_entries");
  for (uint64_t e = 0; e < n_entries; ++e)
    if (point_index[e] >= f->n_points) return fail(c, FITSLAM_E_INVALID, "set_keyframes: point index beyond the resident map points");
  FS_CUDA(c, f->kf.reserve((size_t)n_kf + 1));
  FS_CUDA(c, f->kf_entry_pt.reserve((size_t)n_entries + 1));
  FS_CUDA(c, f->kf_entry_kf.reserve((size_t)n_entries + 1));
  if (n_kf) FS_CUDA(c, cudaMemcpyAsync(f->kf.p, kfs, (size_t)n_kf * sizeof(fitslam_keyframe), cudaMemcpyHostToDevice, c->stream));
  if (n_entries) {
    FS_CUDA(c, cudaMemcpyAsync(f->kf_entry_pt.p, point_index, (size_t)n_entries * 4, cudaMemcpyHostToDevice, c->stream));
    FS_CUDA(c, cudaMemcpyAsync(f->kf_entry_kf.p, entry_kf.data(), (size_t)n_entries * 4, cudaMemcpyHostToDevice, c->stream));
  }
  FS_CUDA(c, cudaStreamSynchronize(c->stream));
  f->n_kf = n_kf;
  f->n_entries = n_entries;
  f->have_kf = true;
  return FITSLAM_OK;
}

extern "C" fitslam_status fitslam_landmarks_in_view_keyframes(fitslam_ctx *c, const fitslam_pose *pose, int32_t max_landmarks,
                                                              float max_distance, float max_angle, int exhaustive_search,
                                                              float *out_xyz, uint32_t capacity, uint32_t *count,
                                                              uint32_t *n_keyframes_checked) {
  (void)max_angle;  // accepted and unused, as in the reference
  if (!c || !pose || !count || (!out_xyz && capacity)) return fail(c, FITSLAM_E_INVALID, "landmarks_in_view_keyframes: null argument");
  cudaSetDevice(c->device);
  uint32_t checked = 0;
  const fitslam_status st = kf_query(c, pose, max_landmarks, max_distance, exhaustive_search, count, &checked);
  if (st != FITSLAM_OK) return st;
  if (n_keyframes_checked) *n_keyframes_checked = checked;
  const uint32_t m = std::min(*count, capacity);
  if (m) {
    FS_CUDA(c, cudaMemcpyAsync(out_xyz, c->fi->kf_out.p, 3 * (size_t)m * sizeof(float), cudaMemcpyDefault, c->stream));
    FS_CUDA(c, cudaStreamSynchronize(c->stream));
  }
  if (*count > capacity) return fail(c, FITSLAM_E_CAPACITY, "landmarks_in_view_keyframes: more landmarks than capacity (count holds the number)");
  return FITSLAM_OK;
}

extern "C" fitslam_status fitslam_score_fisher_keyframes(fitslam_ctx *c, const fitslam_pose *pose, int32_t max_landmarks, float max_distance,
                                                         int exhaustive_search, float *information, uint8_t *safe, uint32_t *n_landmarks,
                                                         uint32_t *n_used) {
  if (!c || !pose || !information) return fail(c, FITSLAM_E_INVALID, "score_fisher_keyframes: null argument");
  cudaSetDevice(c->device);
  FisherParams P;
  fitslam_status st = fisher_params(c, P);  // (lookup table and map points present)
  if (st != FITSLAM_OK) return st;
  FisherState *f = c->fi;
  uint32_t count = 0, checked = 0;
  st = kf_query(c, pose, max_landmarks, max_distance, exhaustive_search, &count, &checked);
  if (st != FITSLAM_OK) return st;
  if (n_landmarks) *n_landmarks = count;
  const size_t lut_n = (size_t)lut_size(f);
  const uint32_t nb = (count + 255) / 256;
  FS_CUDA(c, f->kf_vox.reserve(lut_n + 1));
  FS_CUDA(c, f->kf_lm.reserve(2 * (size_t)count + 2));
  FS_CUDA(c, f->kf_part.reserve((size_t)nb + 1));
  FS_CUDA(c, f->info.reserve(4));
  FS_CUDA(c, f->safe.reserve(4));
  FS_CUDA(c, f->nvis.reserve(4));
  FS_CUDA(c, cudaMemsetAsync(f->kf_vox.p, 0, lut_n * sizeof(uint32_t), c->stream));
  FS_CUDA(c, cudaMemsetAsync(f->nvis.p, 0, sizeof(uint32_t), c->stream));
  if (count) {
    kf_fisher_count_kernel<<<nb, 256, 0, c->stream>>>(P, f->kf_out.p, count, f->kf_vox.p, f->kf_lm.p);
    kf_fisher_sum_kernel<<<nb, 256, 0, c->stream>>>(P, count, f->kf_vox.p, f->kf_lm.p, f->kf_part.p, f->nvis.p);
    c->launches += 2;
  }
  kf_fisher_final_kernel<<<1, 1, 0, c->stream>>>(f->kf_part.p, nb, P.threshold, f->info.p, f->safe.p);
  ++c->launches;
  FS_CUDA(c, cudaGetLastError());
  uint8_t h_safe = 0;
  uint32_t h_used = 0;
  FS_CUDA(c, cudaMemcpyAsync(information, f->info.p, sizeof(float), cudaMemcpyDeviceToHost, c->stream));
  FS_CUDA(c, cudaMemcpyAsync(&h_safe, f->safe.p, 1, cudaMemcpyDeviceToHost, c->stream));
  FS_CUDA(c, cudaMemcpyAsync(&h_used, f->nvis.p, sizeof(uint32_t), cudaMemcpyDeviceToHost, c->stream));
  FS_CUDA(c, cudaStreamSynchronize(c->stream));
  if (safe) *safe = h_safe;
  if (n_used) *n_used = h_used;
  return FITSLAM_OK;
}

// ================================================================================================ no-allocation variant
// fitslam_landmarks_in_view_static: the same three launches as fitslam_landmarks_in_view for callers that cannot call
// cudaMalloc (CUDA graph capture, serving loops, harnesses that forbid it).  No counterpart in the reference.
// The scratch the kernels need (one count per tile of LIV_TILE points + the total) lives in
// a file-scope device array of fixed size instead of a context buffer; the caller passes every input and output as a device
// pointer, nothing is staged, nothing is synchronised.
namespace fitslam {

static constexpr uint64_t LIV_STATIC_MAX_POINTS = FITSLAM_LIV_STATIC_MAX_POINTS;
static constexpr uint32_t LIV_STATIC_MAX_TILES = (uint32_t)((LIV_STATIC_MAX_POINTS + LIV_TILE - 1) / LIV_TILE);
// 16 KB + one word per GPU that loads the library, held until the process exits
static __device__ __align__(256) uint32_t liv_static_scratch[LIV_STATIC_MAX_TILES + 1];

// In host code `liv_static_scratch` is a host-side stand-in, not a device address: the kernels get the address
// cudaGetSymbolAddress returns, looked up once per device with that device current.
static constexpr int LIV_STATIC_MAX_DEVICES = 64;
static uint32_t *liv_static_ptr[LIV_STATIC_MAX_DEVICES] = {};
static std::mutex liv_static_mutex;

static cudaError_t liv_static_address(int device, uint32_t **out) {
  if (device < 0 || device >= LIV_STATIC_MAX_DEVICES) return cudaErrorInvalidDevice;
  std::lock_guard<std::mutex> lock(liv_static_mutex);
  if (!liv_static_ptr[device]) {
    uint32_t *p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress((void **)&p, liv_static_scratch);  // (the caller has made `device` current)
    if (e != cudaSuccess) return e;
    liv_static_ptr[device] = p;
  }
  *out = liv_static_ptr[device];
  return cudaSuccess;
}

}  // namespace fitslam

extern "C" fitslam_status fitslam_landmarks_in_view_static(fitslam_ctx *c, const float *d_xyz, uint64_t n_points,
                                                           const fitslam_pose *d_pose, float *d_out_xyz, uint32_t capacity,
                                                           uint32_t *d_count) {
  if (!c || !d_pose || !d_count || (!d_xyz && n_points) || (!d_out_xyz && capacity))
    return fail(c, FITSLAM_E_INVALID, "landmarks_in_view_static: null argument");
  if (n_points > LIV_STATIC_MAX_POINTS)
    return fail(c, FITSLAM_E_CAPACITY, "landmarks_in_view_static: n_points = " + std::to_string(n_points) + " exceeds the static scratch (" +
                                           std::to_string(LIV_STATIC_MAX_POINTS) + " points, FITSLAM_LIV_STATIC_MAX_POINTS); use fitslam_landmarks_in_view");
  FS_CUDA(c, cudaSetDevice(c->device));
  if (n_points == 0) {
    FS_CUDA(c, cudaMemsetAsync(d_count, 0, sizeof(uint32_t), c->stream));
    return FITSLAM_OK;
  }
  uint32_t *scratch = nullptr;
  FS_CUDA(c, liv_static_address(c->device, &scratch));
  FisherParams P;
  memset(&P, 0, sizeof(P));
  P.fx = c->prm.fx; P.fy = c->prm.fy; P.cx = c->prm.cx; P.cy = c->prm.cy;
  P.max_u = c->prm.img_max_u; P.max_v = c->prm.img_max_v;
  const uint32_t n_tiles = (uint32_t)((n_points + LIV_TILE - 1) / LIV_TILE);
  // every scratch word this call reads is written by it first (pass 0 writes all n_tiles counts, the scan writes the total):
  // nothing has to start at zero, whatever the last call left behind
  uint32_t *tiles = scratch, *total = scratch + LIV_STATIC_MAX_TILES;
  landmarks_in_view_kernel<0><<<n_tiles, LIV_THREADS, 0, c->stream>>>(P, d_pose, d_xyz, n_points, tiles, nullptr, nullptr, 0);
  liv_scan_kernel<<<1, 1024, 0, c->stream>>>(tiles, n_tiles, total);
  c->launches += 2;
  const uint32_t cap = (uint32_t)std::min<uint64_t>(capacity, n_points);
  if (cap) {
    landmarks_in_view_kernel<1><<<n_tiles, LIV_THREADS, 0, c->stream>>>(P, d_pose, d_xyz, n_points, nullptr, tiles, d_out_xyz, cap);
    ++c->launches;
  }
  FS_CUDA(c, cudaGetLastError());
  FS_CUDA(c, cudaMemcpyAsync(d_count, total, sizeof(uint32_t), cudaMemcpyDeviceToDevice, c->stream));
  return FITSLAM_OK;
}
