// Synthetic ray-cast spinning-LiDAR generator (test/bench infrastructure).
//
// Produces motion-distorted scans of a closed box world in the LidarSlam point
// layout (32-byte LidarPoint, reference slam_lib/include/LidarSlam/LidarPoint.h:31-64)
// for the sensor geometries BASELINE.json names (VLP-16, HDL-64, OS-128, VLS-128,
// Pandar64).  The reference ships no dataset (ros_wrapping/tests/README.md:3-4);
// beam tables are uniform approximations of the public datasheets.
// Every ray returns (closed world), points are emitted in firing order
// (for col, for ring ascending), time = (col/C - 1)*period <= 0 relative to the
// frame stamp (= end of sweep), as the ParaView wrapper does.
//
// Build: g++ -O2 -fopenmp -shared -fPIC (see tools/synth/Makefile)
#include <cmath>
#include <cstdint>
#include <cstring>
#include <vector>
#include <algorithm>

extern "C" {

struct synth_point {  // bit-identical to LidarSlam::LidarPoint
  float x, y, z, w;
  double time;
  float intensity;
  uint16_t laser_id;
  uint8_t device_id;
  uint8_t label;
};

struct synth_sensor {
  int n_beams;
  int n_cols;
  double elev_min_deg;   // ring 0 (lowest)
  double elev_max_deg;   // ring n_beams-1 (highest)
  int device_id;
  double mount_xyz[3];   // sensor origin in BASE
  double mount_rpy[3];   // sensor orientation in BASE (Rz*Ry*Rx), radians
};

}  // extern "C"

namespace {

struct V3 { double x, y, z; };
inline V3 operator+(V3 a, V3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline V3 operator-(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline V3 operator*(double s, V3 a) { return {s * a.x, s * a.y, s * a.z}; }
inline double dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline V3 cross(V3 a, V3 b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
inline V3 unit(V3 a) { double n = std::sqrt(dot(a, a)); return {a.x / n, a.y / n, a.z / n}; }

struct M3 { double m[9]; };
inline V3 mul(const M3& R, V3 v) { return {R.m[0] * v.x + R.m[1] * v.y + R.m[2] * v.z, R.m[3] * v.x + R.m[4] * v.y + R.m[5] * v.z, R.m[6] * v.x + R.m[7] * v.y + R.m[8] * v.z}; }
inline V3 mulT(const M3& R, V3 v) { return {R.m[0] * v.x + R.m[3] * v.y + R.m[6] * v.z, R.m[1] * v.x + R.m[4] * v.y + R.m[7] * v.z, R.m[2] * v.x + R.m[5] * v.y + R.m[8] * v.z}; }
inline M3 mul(const M3& A, const M3& B)
{
  M3 C;
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) C.m[i * 3 + j] = A.m[i * 3] * B.m[j] + A.m[i * 3 + 1] * B.m[3 + j] + A.m[i * 3 + 2] * B.m[6 + j];
  return C;
}
inline M3 rpy(double rx, double ry, double rz)
{
  const double cx = std::cos(rx), sx = std::sin(rx), cy = std::cos(ry), sy = std::sin(ry), cz = std::cos(rz), sz = std::sin(rz);
  return {{cy * cz, sx * sy * cz - cx * sz, cx * sy * cz + sx * sz, cy * sz, sx * sy * sz + cx * cz, cx * sy * sz - sx * cz, -sy, sx * cy, cx * cy}};
}

inline uint64_t splitmix(uint64_t x)
{
  x += 0x9E3779B97F4A7C15ull;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
  return x ^ (x >> 31);
}
inline double u01(uint64_t h) { return ((h >> 11) + 0.5) * (1.0 / 9007199254740992.0); }

struct Rng {
  uint64_t s;
  explicit Rng(uint64_t seed) : s(seed) {}
  double uni() { s = splitmix(s); return u01(s); }
  double uni(double a, double b) { return a + (b - a) * uni(); }
};

struct Box { V3 lo, hi; int mat; };
struct Cyl { double cx, cy, r, h; int mat; };
struct Quad { V3 c, n, u, v; double hu, hv; int mat; double brad; };

struct Scene {
  double xmin, xmax, ymin, ymax, zmax;
  std::vector<Box> boxes;
  std::vector<Cyl> cyls;
  std::vector<Quad> quads;
};

struct Handle {
  Scene scene;
  uint64_t noise_seed;
  double speed, period, range_noise;
};

void build_scene(Scene& sc, uint64_t seed)
{
  Rng r(splitmix(seed));
  sc.xmin = -60; sc.xmax = 460; sc.ymin = -120; sc.ymax = 120; sc.zmax = 40;
  auto path_y = [](double x) { return 10.0 * std::sin(0.05 * x); };
  // keep a 6 m corridor around the trajectory free of obstacles
  auto clear_of_path = [&](double x, double y, double rad) { return std::fabs(y - path_y(x)) > rad + 6.0; };
  for (int i = 0; i < 130; ++i) {
    for (int tries = 0; tries < 50; ++tries) {
      const double sx = r.uni(5, 20), sy = r.uni(5, 20), h = r.uni(3, 25);
      const double cx = r.uni(sc.xmin + 15, sc.xmax - 15), cy = r.uni(sc.ymin + 15, sc.ymax - 15);
      if (!clear_of_path(cx, cy, 0.75 * std::max(sx, sy))) continue;
      sc.boxes.push_back({{cx - sx / 2, cy - sy / 2, 0}, {cx + sx / 2, cy + sy / 2, h}, (int)(r.uni() * 4)});
      break;
    }
  }
  for (int i = 0; i < 170; ++i) {
    for (int tries = 0; tries < 50; ++tries) {
      const double rad = r.uni(0.15, 0.4), h = r.uni(4, 12);
      const double cx = r.uni(sc.xmin + 5, sc.xmax - 5);
      // poles line the route (both sides, 7..30 m away) so edges are always in view
      const double side = (r.uni() < 0.5) ? -1.0 : 1.0;
      const double cy = path_y(cx) + side * r.uni(7, 30);
      if (cy < sc.ymin + 2 || cy > sc.ymax - 2) continue;
      sc.cyls.push_back({cx, cy, rad, h, (int)(r.uni() * 4)});
      break;
    }
  }
  for (int i = 0; i < 20; ++i) {
    for (int tries = 0; tries < 50; ++tries) {
      const double cx = r.uni(sc.xmin + 15, sc.xmax - 15), cy = r.uni(sc.ymin + 15, sc.ymax - 15);
      if (!clear_of_path(cx, cy, 8.0)) continue;
      const double yaw = r.uni(0, 2 * M_PI), tilt = r.uni(0.3, 1.0);
      V3 n = unit({std::sin(tilt) * std::cos(yaw), std::sin(tilt) * std::sin(yaw), std::cos(tilt)});
      V3 u = unit(cross({0, 0, 1}, n));
      V3 v = cross(n, u);
      const double hu = r.uni(3, 8), hv = r.uni(3, 8);
      sc.quads.push_back({{cx, cy, hv * std::sin(tilt) + 0.2}, n, u, v, hu, hv, (int)(r.uni() * 4), std::sqrt(hu * hu + hv * hv)});
      break;
    }
  }
}

struct Hit { double t; V3 n; int mat; };

inline void hit_plane(V3 o, V3 d, V3 n, double off, int mat, Hit& best)
{
  const double dn = dot(d, n);
  if (std::fabs(dn) < 1e-12) return;
  const double t = (off - dot(o, n)) / dn;
  if (t > 1e-6 && t < best.t) best = {t, n, mat};
}

inline void hit_box(V3 o, V3 d, const Box& b, Hit& best)
{
  double tmin = -1e300, tmax = 1e300;
  int axmin = 0;
  const double oo[3] = {o.x, o.y, o.z}, dd[3] = {d.x, d.y, d.z}, lo[3] = {b.lo.x, b.lo.y, b.lo.z}, hi[3] = {b.hi.x, b.hi.y, b.hi.z};
  for (int a = 0; a < 3; ++a) {
    if (std::fabs(dd[a]) < 1e-12) {
      if (oo[a] < lo[a] || oo[a] > hi[a]) return;
      continue;
    }
    double t1 = (lo[a] - oo[a]) / dd[a], t2 = (hi[a] - oo[a]) / dd[a];
    if (t1 > t2) std::swap(t1, t2);
    if (t1 > tmin) { tmin = t1; axmin = a; }
    if (t2 < tmax) tmax = t2;
    if (tmin > tmax) return;
  }
  if (tmin > 1e-6 && tmin < best.t) {
    V3 n = {0, 0, 0};
    const double s = (dd[axmin] > 0) ? -1.0 : 1.0;
    if (axmin == 0) n.x = s; else if (axmin == 1) n.y = s; else n.z = s;
    best = {tmin, n, b.mat};
  }
}

inline void hit_cyl(V3 o, V3 d, const Cyl& c, Hit& best)
{
  const double ox = o.x - c.cx, oy = o.y - c.cy;
  const double a = d.x * d.x + d.y * d.y;
  if (a < 1e-14) return;
  const double b = ox * d.x + oy * d.y;
  const double cc = ox * ox + oy * oy - c.r * c.r;
  const double disc = b * b - a * cc;
  if (disc < 0) return;
  const double t = (-b - std::sqrt(disc)) / a;
  if (t <= 1e-6 || t >= best.t) return;
  const double z = o.z + t * d.z;
  if (z < 0 || z > c.h) return;
  V3 n = {(ox + t * d.x) / c.r, (oy + t * d.y) / c.r, 0};
  best = {t, n, c.mat};
}

inline void hit_quad(V3 o, V3 d, const Quad& q, Hit& best)
{
  const double dn = dot(d, q.n);
  if (std::fabs(dn) < 1e-12) return;
  const double t = dot(q.c - o, q.n) / dn;
  if (t <= 1e-6 || t >= best.t) return;
  const V3 p = o + t * d - q.c;
  if (std::fabs(dot(p, q.u)) > q.hu || std::fabs(dot(p, q.v)) > q.hv) return;
  best = {t, q.n, q.mat};
}

// continuous-time trajectory of BASE in WORLD
void traj(const Handle& h, double t, V3& p, M3& R)
{
  // the vehicle starts at rest and accelerates smoothly to `speed` (time constant 1.5 s)
  const double tau = 1.5;
  const double x = h.speed * (t - tau * (1.0 - std::exp(-t / tau)));
  p = {x, 10.0 * std::sin(0.05 * x), 1.8};
  const double yaw = std::atan(0.5 * std::cos(0.05 * x));
  const double rp = (0.5 * M_PI / 180.0) * std::sin(0.7 * t);
  R = rpy(rp, rp, yaw);
}

}  // namespace

extern "C" {

void* synth_create(uint64_t scene_seed, uint64_t sequence_id)
{
  Handle* h = new Handle;
  build_scene(h->scene, scene_seed);
  h->noise_seed = splitmix(scene_seed * 1000003ull + sequence_id);
  h->speed = 8.0;
  h->period = 0.1;
  h->range_noise = 0.01;
  return h;
}

void synth_destroy(void* hp) { delete static_cast<Handle*>(hp); }

void synth_set_motion(void* hp, double speed, double period, double range_noise)
{
  Handle* h = static_cast<Handle*>(hp);
  h->speed = speed; h->period = period; h->range_noise = range_noise;
}

// Fills out[n_beams*n_cols]; gt_pose = row-major 4x4 BASE-in-WORLD at the frame stamp.
int synth_frame(void* hp, const synth_sensor* s, int frame, synth_point* out, double* gt_pose, uint64_t* stamp_us)
{
  const Handle& h = *static_cast<Handle*>(hp);
  const Scene& sc = h.scene;
  const int R = s->n_beams, C = s->n_cols;
  const double t_end = (frame + 1) * h.period;
  *stamp_us = (uint64_t)llround(t_end * 1e6);
  {
    V3 p; M3 Rm;
    traj(h, t_end, p, Rm);
    const double T[16] = {Rm.m[0], Rm.m[1], Rm.m[2], p.x, Rm.m[3], Rm.m[4], Rm.m[5], p.y, Rm.m[6], Rm.m[7], Rm.m[8], p.z, 0, 0, 0, 1};
    std::memcpy(gt_pose, T, sizeof(T));
  }
  const M3 Rmount = rpy(s->mount_rpy[0], s->mount_rpy[1], s->mount_rpy[2]);
  const V3 tmount = {s->mount_xyz[0], s->mount_xyz[1], s->mount_xyz[2]};

  // azimuth sectors (world frame, around the mid-sweep sensor position) to cull primitives
  const int NS = 128;
  std::vector<std::vector<int>> sect_box(NS), sect_cyl(NS), sect_quad(NS);
  {
    V3 pm; M3 Rm;
    traj(h, t_end - 0.5 * h.period, pm, Rm);
    const V3 om = pm + mul(Rm, tmount);
    const double slack = h.speed * h.period + 0.5;  // sensor displacement during the sweep
    auto add = [&](std::vector<std::vector<int>>& sect, int id, double cx, double cy, double rad) {
      const double dx = cx - om.x, dy = cy - om.y, dist = std::sqrt(dx * dx + dy * dy);
      const double rr = rad + slack;
      if (dist <= rr) { for (int k = 0; k < NS; ++k) sect[k].push_back(id); return; }
      const double a = std::atan2(dy, dx), w = std::asin(rr / dist) + 2 * M_PI / NS;
      const int k0 = (int)std::floor((a - w + M_PI) / (2 * M_PI) * NS), k1 = (int)std::floor((a + w + M_PI) / (2 * M_PI) * NS);
      for (int k = k0; k <= k1; ++k) sect[((k % NS) + NS) % NS].push_back(id);
    };
    for (size_t i = 0; i < sc.boxes.size(); ++i) {
      const Box& b = sc.boxes[i];
      add(sect_box, (int)i, 0.5 * (b.lo.x + b.hi.x), 0.5 * (b.lo.y + b.hi.y), 0.5 * std::hypot(b.hi.x - b.lo.x, b.hi.y - b.lo.y));
    }
    for (size_t i = 0; i < sc.cyls.size(); ++i) add(sect_cyl, (int)i, sc.cyls[i].cx, sc.cyls[i].cy, sc.cyls[i].r);
    for (size_t i = 0; i < sc.quads.size(); ++i) add(sect_quad, (int)i, sc.quads[i].c.x, sc.quads[i].c.y, sc.quads[i].brad);
  }

#pragma omp parallel for schedule(dynamic, 16)
  for (int col = 0; col < C; ++col) {
    const double trel = ((double)col / C - 1.0) * h.period;
    V3 pb; M3 Rb;
    traj(h, t_end + trel, pb, Rb);
    const M3 Rs = mul(Rb, Rmount);          // sensor orientation in world
    const V3 o = pb + mul(Rb, tmount);      // sensor origin in world
    const double az = -2.0 * M_PI * col / C;  // clockwise
    for (int ring = 0; ring < R; ++ring) {
      const double el = (R > 1 ? s->elev_min_deg + (s->elev_max_deg - s->elev_min_deg) * ring / (R - 1) : s->elev_min_deg) * M_PI / 180.0;
      const V3 ds = {std::cos(el) * std::cos(az), std::cos(el) * std::sin(az), std::sin(el)};
      const V3 d = mul(Rs, ds);
      Hit best = {1e300, {0, 0, 1}, 0};
      hit_plane(o, d, {0, 0, 1}, 0.0, 0, best);
      hit_plane(o, d, {0, 0, -1}, -sc.zmax, 1, best);
      hit_plane(o, d, {1, 0, 0}, sc.xmin, 2, best);
      hit_plane(o, d, {-1, 0, 0}, -sc.xmax, 2, best);
      hit_plane(o, d, {0, 1, 0}, sc.ymin, 3, best);
      hit_plane(o, d, {0, -1, 0}, -sc.ymax, 3, best);
      int k = (int)std::floor((std::atan2(d.y, d.x) + M_PI) / (2 * M_PI) * NS);
      k = ((k % NS) + NS) % NS;
      for (int id : sect_box[k]) hit_box(o, d, sc.boxes[id], best);
      for (int id : sect_cyl[k]) hit_cyl(o, d, sc.cyls[id], best);
      for (int id : sect_quad[k]) hit_quad(o, d, sc.quads[id], best);

      const uint64_t key = splitmix(h.noise_seed ^ splitmix(((uint64_t)frame << 40) ^ ((uint64_t)s->device_id << 32) ^ ((uint64_t)col * 1024 + ring)));
      const double u1 = u01(splitmix(key)), u2 = u01(splitmix(key + 1)), u3 = u01(splitmix(key + 2));
      const double gauss = std::sqrt(-2.0 * std::log(u1)) * std::cos(2 * M_PI * u2);
      const double range = best.t + h.range_noise * gauss;
      const double cosi = std::fabs(dot(d, best.n));
      double inten = 40.0 + 150.0 * cosi + 20.0 * best.mat + u3;
      if (inten < 0) inten = 0;
      if (inten >= 255.0) inten = 254.999;

      synth_point& q = out[(size_t)col * R + ring];
      q.x = (float)(range * ds.x); q.y = (float)(range * ds.y); q.z = (float)(range * ds.z); q.w = 1.0f;
      q.time = trel;
      q.intensity = (float)inten;
      q.laser_id = (uint16_t)ring;
      q.device_id = (uint8_t)s->device_id;
      q.label = 0;
    }
  }
  return R * C;
}

}  // extern "C"
