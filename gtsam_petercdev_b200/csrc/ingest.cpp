// ingest.cpp -- dataset files straight into the struct-of-arrays problem tables.
//
// Host C++, no CUDA.  The reference parses g2o / TORO / BAL files into hundreds of thousands of shared_ptr factors and
// std::map nodes (slam/dataset.cpp:505-570 load2D, :922-946 load3D, :621-633 readG2o; sfm/SfmData.cpp:189-247
// FromBalFile) which a device packer would flatten again at once; here every record goes directly into the flat arrays
// that gtb_set_variables / gtb_set_factors take.  Behaviour mirrored from the reference:
//   * a record is "tag fields..." on one line; unknown tags are skipped (parseLines, slam/dataset.cpp:129-140)
//   * 2D: VERTEX2 | VERTEX_SE2 | VERTEX  id x y yaw;  EDGE2 | EDGE | EDGE_SE2 | ODOMETRY  id1 id2 x y yaw v0..v5 with the
//     g2o information layout [v0 v1 v2; v1 v3 v4; v2 v4 v5] (createNoiseModel, :216-297, NoiseFormatG2O), model =
//     Gaussian::Information(M, smart); vertices referenced by an edge but absent are created by chaining the
//     measurement on the first key (:541-545)
//   * 3D: VERTEX3 id x y z roll pitch yaw (Rot3::Ypr = RzRyRx, geometry/Rot3M.cpp:84-107), VERTEX_SE3:QUAT id x y z qx
//     qy qz qw (normalised, Eigen's quaternion -> matrix), EDGE3 / EDGE_SE3:QUAT with the upper triangle of a 6 x 6
//     information matrix; the QUAT form stores it in (t, R) order and is permuted to GTSAM's (R, t) (:846-856);
//     unlike 2D, missing vertices are NOT created (:929-931)
//   * BAL: every number but the counts is read as a 32-bit float, v is negated, the OpenGL camera (R, t) becomes
//     wRc = R^T diag(1,-1,-1), wtc = -R^T t (openGL2gtsam, sfm/SfmData.cpp:79-86); factors follow the track order of
//     examples/SFMExample_bal.cpp:52-63 (for every point, its measurements in file order), cameras get the plain keys
//     0..n-1 and points P(j) as in tests/testGeneralSFMFactorB.cpp:44-58, Unit measurement noise
//   * writers: writeG2o (slam/dataset.cpp:636-735), default stream formatting (6 significant digits)
// 2D landmark records (VERTEX_XY, BR, LANDMARK) have no variable type on the device path and are rejected; 3D landmarks
// (VERTEX_TRACKXYZ) become Point3 variables keyed L(id), as in the reference.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <map>
#include <sstream>
#include <string>
#include <vector>

#include "../../include/gtsam_b200.h"

struct gtb_dataset {
  std::vector<uint64_t> vkeys[GTB_NUM_VTYPES];
  std::vector<double> vstate[GTB_NUM_VTYPES];
  std::vector<uint64_t> fkeys[GTB_NUM_BUCKETS];
  std::vector<double> fz[GTB_NUM_BUCKETS];
  std::vector<int32_t> fnk[GTB_NUM_BUCKETS];
  std::vector<double> fnoise[GTB_NUM_BUCKETS];
  int32_t loss = GTB_LOSS_NONE;   // kernelFunctionType of readG2o: Huber(1.345) / Tukey(4.6851) on every edge
  double loss_k = 0.0;
};

// the reference holds a 2-D rotation as (cos, sin) and prints Rot2::theta() = atan2(sin, cos) (geometry/Rot2.h:190-192):
// an angle read as 4.71 is written back as -1.57
static inline double wrap_theta(double th) { return std::atan2(std::sin(th), std::cos(th)); }

namespace {

thread_local std::string g_err;
const int kStateIn[GTB_NUM_VTYPES] = {3, 12, 17, 3};
const int kZ[GTB_NUM_BUCKETS] = {3, 12, 2, 3, 12, 17, 3};
const int kDim[GTB_NUM_BUCKETS] = {3, 6, 2, 3, 6, 9, 3};
const int kNKeys[GTB_NUM_BUCKETS] = {2, 2, 2, 1, 1, 1, 1};

struct ParseError { std::string msg; };

// whitespace-separated fields of one record
struct Fields {
  const char* p;
  explicit Fields(const char* s) : p(s) {}
  bool next_token(const char*& b, const char*& e) {
    while (*p == ' ' || *p == '\t' || *p == '\r') ++p;
    if (!*p || *p == '\n') return false;
    b = p;
    while (*p && *p != ' ' && *p != '\t' && *p != '\r' && *p != '\n') ++p;
    e = p;
    return true;
  }
  double f64(const char* what) {
    const char *b, *e;
    if (!next_token(b, e)) throw ParseError{std::string("malformed line: missing ") + what};
    char* end = nullptr;
    const double v = std::strtod(b, &end);
    if (end == b) throw ParseError{std::string("malformed line: bad number for ") + what};
    return v;
  }
  double f32(const char* what) {   // BAL: read as float, then widened
    const char *b, *e;
    if (!next_token(b, e)) throw ParseError{std::string("malformed file: missing ") + what};
    char* end = nullptr;
    const float v = std::strtof(b, &end);
    if (end == b) throw ParseError{std::string("malformed file: bad number for ") + what};
    return (double)v;
  }
  uint64_t u64(const char* what) {
    const char *b, *e;
    if (!next_token(b, e)) throw ParseError{std::string("malformed line: missing ") + what};
    char* end = nullptr;
    const unsigned long long v = std::strtoull(b, &end, 10);
    if (end == b) throw ParseError{std::string("malformed line: bad index for ") + what};
    return (uint64_t)v;
  }
};

// Rot3::RzRyRx(x = roll, y = pitch, z = yaw), row-major (geometry/Rot3M.cpp:84-107)
void rzryrx(double x, double y, double z, double* R) {
  const double cx = std::cos(x), sx = std::sin(x), cy = std::cos(y), sy = std::sin(y), cz = std::cos(z), sz = std::sin(z);
  const double ss_ = sx * sy, cs_ = cx * sy, sc_ = sx * cy, cc_ = cx * cy, c_s = cx * sz, s_s = sx * sz, _cs = cy * sz,
               _cc = cy * cz, s_c = sx * cz, c_c = cx * cz;
  const double ssc = ss_ * cz, csc = cs_ * cz, sss = ss_ * sz, css = cs_ * sz;
  R[0] = _cc; R[1] = -c_s + ssc; R[2] = s_s + csc;
  R[3] = _cs; R[4] = c_c + sss;  R[5] = -s_c + css;
  R[6] = -sy; R[7] = sc_;        R[8] = cc_;
}

// file quaternion (x, y, z, w) -> normalised -> rotation matrix, Eigen's QuaternionBase::toRotationMatrix order of
// operations (slam/dataset.cpp:738-745 then Rot3(const Quaternion&), geometry/Rot3M.cpp)
void quat_to_R(double x, double y, double z, double w, double* R) {
  const double norm = std::sqrt(w * w + x * x + y * y + z * z), f = 1.0 / norm;
  w = f * w; x = f * x; y = f * y; z = f * z;
  const double tx = 2.0 * x, ty = 2.0 * y, tz = 2.0 * z;
  const double twx = tx * w, twy = ty * w, twz = tz * w, txx = tx * x, txy = ty * x, txz = tz * x, tyy = ty * y, tyz = tz * y,
               tzz = tz * z;
  R[0] = 1.0 - (tyy + tzz); R[1] = txy - twz;         R[2] = txz + twy;
  R[3] = txy + twz;         R[4] = 1.0 - (txx + tzz); R[5] = tyz - twx;
  R[6] = txz - twy;         R[7] = tyz + twx;         R[8] = 1.0 - (txx + tyy);
}

// rotation matrix (row-major) -> quaternion (w, x, y, z), Eigen's trace-based conversion (Rot3::toQuaternion)
void R_to_quat(const double* R, double* q) {
  const double t = R[0] + R[4] + R[8];
  if (t > 0.0) {
    double s = std::sqrt(t + 1.0);
    q[0] = 0.5 * s;
    s = 0.5 / s;
    q[1] = (R[7] - R[5]) * s; q[2] = (R[2] - R[6]) * s; q[3] = (R[3] - R[1]) * s;
  } else {
    int i = 0;
    if (R[4] > R[0]) i = 1;
    if (R[8] > R[4 * i]) i = 2;
    const int j = (i + 1) % 3, k = (j + 1) % 3;
    double s = std::sqrt(R[4 * i] - R[4 * j] - R[4 * k] + 1.0);
    q[1 + i] = 0.5 * s;
    s = 0.5 / s;
    q[0] = (R[3 * k + j] - R[3 * j + k]) * s;
    q[1 + j] = (R[3 * j + i] + R[3 * i + j]) * s;
    q[1 + k] = (R[3 * k + i] + R[3 * i + k]) * s;
  }
}

void read_lines(const char* path, std::vector<std::string>& lines) {
  std::ifstream is(path);
  if (!is) throw ParseError{std::string("can not find file ") + path};
  std::string l;
  while (std::getline(is, l)) lines.push_back(l);
}

bool split_tag(const std::string& line, std::string& tag, const char*& rest) {
  size_t a = 0;
  while (a < line.size() && (line[a] == ' ' || line[a] == '\t' || line[a] == '\r')) ++a;
  size_t b = a;
  while (b < line.size() && line[b] != ' ' && line[b] != '\t' && line[b] != '\r') ++b;
  if (b == a) return false;
  tag.assign(line, a, b - a);
  rest = line.c_str() + b;
  return true;
}

void from_map(const std::map<uint64_t, std::vector<double>>& m, std::vector<uint64_t>& keys, std::vector<double>& state) {
  for (const auto& kv : m) {   // Values iterates in ascending key order
    keys.push_back(kv.first);
    state.insert(state.end(), kv.second.begin(), kv.second.end());
  }
}

void load_2d(const char* path, gtb_dataset* d) {
  std::vector<std::string> lines;
  read_lines(path, lines);
  std::map<uint64_t, std::vector<double>> poses;   // (x, y, theta)
  std::map<uint64_t, std::pair<double, double>> chain_cs;   // (cos, sin) of the poses created by chaining odometry
  std::string tag;
  const char* rest;
  // first pass: vertices (slam/dataset.cpp:511-523)
  for (const std::string& l : lines) {
    if (!split_tag(l, tag, rest)) continue;
    if (tag == "VERTEX2" || tag == "VERTEX_SE2" || tag == "VERTEX") {
      Fields f(rest);
      const uint64_t id = f.u64("id");
      const double x = f.f64("x"), y = f.f64("y"), yaw = f.f64("yaw");
      if (poses.count(id)) throw ParseError{"vertex " + std::to_string(id) + " appears twice (ValuesKeyAlreadyExists)"};
      poses[id] = {x, y, yaw};
    } else if (tag == "VERTEX_XY") {
      throw ParseError{"2D landmarks (VERTEX_XY) have no variable type on the device path"};
    }
  }
  // second pass: edges (:525-566)
  for (const std::string& l : lines) {
    if (!split_tag(l, tag, rest)) continue;
    if (tag == "BR" || tag == "LANDMARK") throw ParseError{"bearing-range measurements have no factor bucket on the device path"};
    if (!(tag == "EDGE2" || tag == "EDGE" || tag == "EDGE_SE2" || tag == "ODOMETRY")) continue;
    Fields f(rest);
    const uint64_t id1 = f.u64("id1"), id2 = f.u64("id2");
    const double x = f.f64("x"), y = f.f64("y"), yaw = f.f64("yaw");
    double v[6];
    for (int i = 0; i < 6; ++i) v[i] = f.f64("information");
    if (v[0] == 0.0 || v[3] == 0.0 || v[5] == 0.0) throw ParseError{"load2D::readNoiseModel looks like this is not G2O matrix order"};
    const double M[9] = {v[0], v[1], v[2], v[1], v[3], v[4], v[2], v[4], v[5]};
    d->fkeys[GTB_BETWEEN2].push_back(id1);
    d->fkeys[GTB_BETWEEN2].push_back(id2);
    d->fz[GTB_BETWEEN2].insert(d->fz[GTB_BETWEEN2].end(), {x, y, yaw});
    d->fnk[GTB_BETWEEN2].push_back(GTB_NOISE_INFORMATION);
    d->fnoise[GTB_BETWEEN2].insert(d->fnoise[GTB_BETWEEN2].end(), M, M + 9);
    // pure odometry file: chain the measurement (Pose2::operator*, geometry/Pose2.h:138, Rot2::operator*, Rot2.h:120)
    // The reference keeps a rotation as (cos, sin) (geometry/Rot2.h:44), so a chained pose carries the PRODUCT of the
    // cosines and sines along its chain, not cos / sin of the accumulated angle: kept here next to the pose so that x and
    // y come out bit-identical along a 20,000-pose chain; theta = atan2(s, c) (Rot2::theta, Rot2.h:190) is what is stored.
    if (!poses.count(id1)) { poses[id1] = {0.0, 0.0, 0.0}; chain_cs[id1] = {1.0, 0.0}; }
    if (!poses.count(id2)) {
      const std::vector<double>& a = poses[id1];
      const auto it = chain_cs.find(id1);
      const double c1 = it != chain_cs.end() ? it->second.first : std::cos(a[2]);
      const double s1 = it != chain_cs.end() ? it->second.second : std::sin(a[2]);
      const double c2 = std::cos(yaw), s2 = std::sin(yaw);
      double c = c1 * c2 - s1 * s2, s = s1 * c2 + c1 * s2;
      double n2 = c * c + s * s;   // Rot2::fromCosSin -> normalize (geometry/Rot2.cpp:56-64)
      if (std::abs(n2 - 1.0) > 1e-10) { n2 = 1 / std::sqrt(n2); c *= n2; s *= n2; }
      poses[id2] = {a[0] + (c1 * x + -s1 * y), a[1] + (s1 * x + c1 * y), std::atan2(s, c)};
      chain_cs[id2] = {c, s};
    }
  }
  from_map(poses, d->vkeys[GTB_POSE2], d->vstate[GTB_POSE2]);
}

void load_3d(const char* path, gtb_dataset* d) {
  std::vector<std::string> lines;
  read_lines(path, lines);
  std::map<uint64_t, std::vector<double>> poses;   // (R 9 row-major, t 3)
  std::map<uint64_t, std::vector<double>> points;
  std::string tag;
  const char* rest;
  for (const std::string& l : lines) {
    if (!split_tag(l, tag, rest)) continue;
    if (tag == "VERTEX3" || tag == "VERTEX_SE3:QUAT") {
      Fields f(rest);
      const uint64_t id = f.u64("id");
      const double x = f.f64("x"), y = f.f64("y"), z = f.f64("z");
      std::vector<double> st(12);
      if (tag == "VERTEX3") {
        const double roll = f.f64("roll"), pitch = f.f64("pitch"), yaw = f.f64("yaw");
        rzryrx(roll, pitch, yaw, st.data());
      } else {
        const double qx = f.f64("qx"), qy = f.f64("qy"), qz = f.f64("qz"), qw = f.f64("qw");
        quat_to_R(qx, qy, qz, qw, st.data());
      }
      st[9] = x; st[10] = y; st[11] = z;
      if (poses.count(id)) throw ParseError{"vertex " + std::to_string(id) + " appears twice (ValuesKeyAlreadyExists)"};
      poses[id] = st;
    } else if (tag == "VERTEX_TRACKXYZ") {   // Point3 landmark, key L(id) (slam/dataset.cpp:781-790, :933-934)
      Fields f(rest);
      const uint64_t id = f.u64("id");
      const double x = f.f64("x"), y = f.f64("y"), z = f.f64("z");
      const uint64_t key = ((uint64_t)'l' << 56) | id;
      if (points.count(key)) throw ParseError{"landmark " + std::to_string(id) + " appears twice (ValuesKeyAlreadyExists)"};
      points[key] = {x, y, z};
    } else if (tag == "EDGE3" || tag == "EDGE_SE3:QUAT") {
      Fields f(rest);
      const uint64_t id1 = f.u64("id1"), id2 = f.u64("id2");
      const double x = f.f64("x"), y = f.f64("y"), z = f.f64("z");
      double zz[12];
      const bool quat = tag == "EDGE_SE3:QUAT";
      if (!quat) {
        const double roll = f.f64("roll"), pitch = f.f64("pitch"), yaw = f.f64("yaw");
        rzryrx(roll, pitch, yaw, zz);
      } else {
        const double qx = f.f64("qx"), qy = f.f64("qy"), qz = f.f64("qz"), qw = f.f64("qw");
        quat_to_R(qx, qy, qz, qw, zz);
      }
      zz[9] = x; zz[10] = y; zz[11] = z;
      double m[36];
      for (int i = 0; i < 6; ++i)
        for (int j = i; j < 6; ++j) m[i * 6 + j] = m[j * 6 + i] = f.f64("information");
      double M[36];
      if (quat) {   // (t, R) -> (R, t) block permutation
        for (int i = 0; i < 6; ++i)
          for (int j = 0; j < 6; ++j) M[i * 6 + j] = m[((i + 3) % 6) * 6 + (j + 3) % 6];
      } else {
        std::memcpy(M, m, sizeof(M));
      }
      d->fkeys[GTB_BETWEEN3].push_back(id1);
      d->fkeys[GTB_BETWEEN3].push_back(id2);
      d->fz[GTB_BETWEEN3].insert(d->fz[GTB_BETWEEN3].end(), zz, zz + 12);
      d->fnk[GTB_BETWEEN3].push_back(GTB_NOISE_INFORMATION);
      d->fnoise[GTB_BETWEEN3].insert(d->fnoise[GTB_BETWEEN3].end(), M, M + 36);
    }
  }
  from_map(poses, d->vkeys[GTB_POSE3], d->vstate[GTB_POSE3]);
  from_map(points, d->vkeys[GTB_POINT3], d->vstate[GTB_POINT3]);
}

// Rot3::Rodrigues(w) = Expmap (full exponential map, geometry/SO3.cpp ExpmapFunctor): R = I + A W + B W^2
void rodrigues(double wx, double wy, double wz, double* R) {
  const double t2 = wx * wx + wy * wy + wz * wz;
  double A, B;
  if (t2 > 1e-10) {   // nearZero threshold of so3::ExpmapFunctor
    const double t = std::sqrt(t2), s2 = std::sin(t / 2.0);
    A = std::sin(t) / t;
    B = 2.0 * s2 * s2 / t2;
  } else {
    A = 1.0 - t2 / 6.0;
    B = 0.5 - t2 / 24.0;
  }
  const double W[9] = {0, -wz, wy, wz, 0, -wx, -wy, wx, 0};
  double WW[9];
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) WW[i * 3 + j] = W[i * 3] * W[j] + W[i * 3 + 1] * W[3 + j] + W[i * 3 + 2] * W[6 + j];
  for (int i = 0; i < 9; ++i) R[i] = (i % 4 == 0 ? 1.0 : 0.0) + A * W[i] + B * WW[i];
}

void load_bal(const char* path, gtb_dataset* d) {
  std::ifstream is(path, std::ios::binary);
  if (!is) throw ParseError{"Error in FromBalFile: can not find the file!!"};
  std::stringstream ss;
  ss << is.rdbuf();
  const std::string text = ss.str();
  // the whole file is one whitespace-separated token stream
  std::string flat = text;
  for (char& c : flat)
    if (c == '\n') c = ' ';
  Fields f(flat.c_str());
  const uint64_t n_cams = f.u64("nrPoses"), n_points = f.u64("nrPoints"), n_obs = f.u64("nrObservations");
  std::vector<std::vector<std::pair<uint64_t, std::pair<double, double>>>> tracks(n_points);
  for (uint64_t k = 0; k < n_obs; ++k) {
    const uint64_t i = f.u64("camera index"), j = f.u64("point index");
    const double u = f.f32("u"), v = f.f32("v");
    if (i >= n_cams || j >= n_points) throw ParseError{"observation refers to a camera or point out of range"};
    tracks[j].push_back({i, {u, -v}});
  }
  std::vector<double>& cs = d->vstate[GTB_CAMERA];
  for (uint64_t i = 0; i < n_cams; ++i) {
    const double wx = f.f32("wx"), wy = f.f32("wy"), wz = f.f32("wz");
    const double tx = f.f32("tx"), ty = f.f32("ty"), tz = f.f32("tz");
    const double fl = f.f32("f"), k1 = f.f32("k1"), k2 = f.f32("k2");
    double R[9];
    rodrigues(wx, wy, wz, R);
    // wRc = R^T diag(1, -1, -1); wtc = R^T (-t)
    double st[17];
    for (int r = 0; r < 3; ++r) {
      st[r * 3 + 0] = R[0 * 3 + r];
      st[r * 3 + 1] = -R[1 * 3 + r];
      st[r * 3 + 2] = -R[2 * 3 + r];
      st[9 + r] = R[0 * 3 + r] * (-tx) + R[1 * 3 + r] * (-ty) + R[2 * 3 + r] * (-tz);
    }
    st[12] = fl; st[13] = k1; st[14] = k2; st[15] = 0.0; st[16] = 0.0;
    d->vkeys[GTB_CAMERA].push_back(i);
    cs.insert(cs.end(), st, st + 17);
  }
  for (uint64_t j = 0; j < n_points; ++j) {
    const double x = f.f32("x"), y = f.f32("y"), z = f.f32("z");
    d->vkeys[GTB_POINT3].push_back(((uint64_t)'p' << 56) | j);
    d->vstate[GTB_POINT3].insert(d->vstate[GTB_POINT3].end(), {x, y, z});
  }
  for (uint64_t j = 0; j < n_points; ++j)
    for (const auto& m : tracks[j]) {
      d->fkeys[GTB_SFM].push_back(m.first);
      d->fkeys[GTB_SFM].push_back(((uint64_t)'p' << 56) | j);
      d->fz[GTB_SFM].push_back(m.second.first);
      d->fz[GTB_SFM].push_back(m.second.second);
      d->fnk[GTB_SFM].push_back(GTB_NOISE_UNIT);
      d->fnoise[GTB_SFM].insert(d->fnoise[GTB_SFM].end(), 4, 0.0);
    }
}

template <class Fn> gtb_status guarded_read(gtb_dataset** out, Fn fn) {
  if (!out) return GTB_ERR_INVALID;
  *out = nullptr;
  gtb_dataset* d = new gtb_dataset();
  try {
    fn(d);
  } catch (const ParseError& e) {
    g_err = e.msg;
    delete d;
    return GTB_ERR_INVALID;
  } catch (const std::exception& e) {
    g_err = e.what();
    delete d;
    return GTB_ERR_INVALID;
  }
  *out = d;
  return GTB_OK;
}

}  // namespace

extern "C" {

gtb_status gtb_dataset_read_g2o(const char* path, int is3d, int kernel, gtb_dataset** out) {
  return guarded_read(out, [&](gtb_dataset* d) {
    if (!path) throw ParseError{"null path"};
    if (is3d) {
      // load3D ignores kernelFunctionType (slam/dataset.cpp:621-624)
      load_3d(path, d);
    } else {
      load_2d(path, d);
      if (kernel == GTB_LOSS_HUBER) { d->loss = GTB_LOSS_HUBER; d->loss_k = 1.345; }
      else if (kernel == GTB_LOSS_TUKEY) { d->loss = GTB_LOSS_TUKEY; d->loss_k = 4.6851; }
      else if (kernel != GTB_LOSS_NONE) throw ParseError{"load2D: invalid kernel function type"};
    }
  });
}

gtb_status gtb_dataset_read_bal(const char* path, gtb_dataset** out) {
  return guarded_read(out, [&](gtb_dataset* d) {
    if (!path) throw ParseError{"null path"};
    load_bal(path, d);
  });
}

void gtb_dataset_free(gtb_dataset* d) { delete d; }
const char* gtb_dataset_last_error(void) { return g_err.c_str(); }

int64_t gtb_dataset_num_variables(const gtb_dataset* d, int vtype) {
  return (d && vtype >= 0 && vtype < GTB_NUM_VTYPES) ? (int64_t)d->vkeys[vtype].size() : -1;
}
const uint64_t* gtb_dataset_variable_keys(const gtb_dataset* d, int vtype) {
  return (d && vtype >= 0 && vtype < GTB_NUM_VTYPES) ? d->vkeys[vtype].data() : nullptr;
}
const double* gtb_dataset_variable_state(const gtb_dataset* d, int vtype) {
  return (d && vtype >= 0 && vtype < GTB_NUM_VTYPES) ? d->vstate[vtype].data() : nullptr;
}
int64_t gtb_dataset_num_factors(const gtb_dataset* d, int bucket) {
  return (d && bucket >= 0 && bucket < GTB_NUM_BUCKETS) ? (int64_t)d->fnk[bucket].size() : -1;
}
const uint64_t* gtb_dataset_factor_keys(const gtb_dataset* d, int bucket) {
  return (d && bucket >= 0 && bucket < GTB_NUM_BUCKETS) ? d->fkeys[bucket].data() : nullptr;
}
const double* gtb_dataset_factor_z(const gtb_dataset* d, int bucket) {
  return (d && bucket >= 0 && bucket < GTB_NUM_BUCKETS) ? d->fz[bucket].data() : nullptr;
}
const int32_t* gtb_dataset_factor_noise_kind(const gtb_dataset* d, int bucket) {
  return (d && bucket >= 0 && bucket < GTB_NUM_BUCKETS) ? d->fnk[bucket].data() : nullptr;
}
const double* gtb_dataset_factor_noise(const gtb_dataset* d, int bucket) {
  return (d && bucket >= 0 && bucket < GTB_NUM_BUCKETS) ? d->fnoise[bucket].data() : nullptr;
}
int gtb_dataset_robust(const gtb_dataset* d, double* k) {
  if (!d) return GTB_LOSS_NONE;
  if (k) *k = d->loss_k;
  return d->loss;
}

gtb_status gtb_dataset_apply(const gtb_dataset* d, gtb_engine* e) {
  if (!d || !e) return GTB_ERR_INVALID;
  for (int t = 0; t < GTB_NUM_VTYPES; ++t)
    if (!d->vkeys[t].empty()) {
      const gtb_status st = gtb_set_variables(e, t, (int64_t)d->vkeys[t].size(), d->vkeys[t].data(), d->vstate[t].data());
      if (st != GTB_OK) return st;
    }
  for (int b = 0; b < GTB_NUM_BUCKETS; ++b)
    if (!d->fnk[b].empty()) {
      gtb_status st = gtb_set_factors(e, b, (int64_t)d->fnk[b].size(), d->fkeys[b].data(), d->fz[b].data(), d->fnk[b].data(),
                                      d->fnoise[b].data());
      if (st != GTB_OK) return st;
      if (d->loss != GTB_LOSS_NONE && (b == GTB_BETWEEN2)) {
        st = gtb_set_robust(e, b, d->loss, d->loss_k);
        if (st != GTB_OK) return st;
      }
    }
  return GTB_OK;
}

// writeG2o (slam/dataset.cpp:636-735): poses of both kinds, then the between factors in graph order.  `information`
// per factor is the d x d matrix R^T R of its Gaussian model (what the reference recomputes from the model), row-major.
gtb_status gtb_write_g2o(const char* path, int64_t n_pose2, const uint64_t* pose2_keys, const double* pose2,
                         int64_t n_pose3, const uint64_t* pose3_keys, const double* pose3,
                         int64_t n_between2, const uint64_t* between2_keys, const double* between2_z, const double* between2_information,
                         int64_t n_between3, const uint64_t* between3_keys, const double* between3_z, const double* between3_information) {
  if (!path) return GTB_ERR_INVALID;
  std::ofstream os(path);
  if (!os) { g_err = std::string("can not open ") + path; return GTB_ERR_INVALID; }
  auto index = [](uint64_t key) { return key & 0x00ffffffffffffffull; };   // Symbol(key).index()
  for (int64_t i = 0; i < n_pose2; ++i)
    os << "VERTEX_SE2 " << index(pose2_keys[i]) << " " << pose2[3 * i] << " " << pose2[3 * i + 1] << " " << wrap_theta(pose2[3 * i + 2]) << std::endl;
  for (int64_t i = 0; i < n_pose3; ++i) {
    const double* s = pose3 + 12 * i;
    double q[4];
    R_to_quat(s, q);
    os << "VERTEX_SE3:QUAT " << index(pose3_keys[i]) << " " << s[9] << " " << s[10] << " " << s[11] << " " << q[1] << " " << q[2] << " "
       << q[3] << " " << q[0] << std::endl;
  }
  for (int64_t i = 0; i < n_between2; ++i) {
    const double* z = between2_z + 3 * i;
    const double* I = between2_information + 9 * i;
    os << "EDGE_SE2 " << index(between2_keys[2 * i]) << " " << index(between2_keys[2 * i + 1]) << " " << z[0] << " " << z[1] << " " << wrap_theta(z[2]);
    for (int r = 0; r < 3; ++r)
      for (int c = r; c < 3; ++c) os << " " << I[r * 3 + c];
    os << std::endl;
  }
  for (int64_t i = 0; i < n_between3; ++i) {
    const double* z = between3_z + 12 * i;
    const double* I = between3_information + 36 * i;
    double q[4];
    R_to_quat(z, q);
    os << "EDGE_SE3:QUAT " << index(between3_keys[2 * i]) << " " << index(between3_keys[2 * i + 1]) << " " << z[9] << " " << z[10] << " "
       << z[11] << " " << q[1] << " " << q[2] << " " << q[3] << " " << q[0];
    for (int r = 0; r < 6; ++r)
      for (int c = r; c < 6; ++c) os << " " << I[((r + 3) % 6) * 6 + (c + 3) % 6];   // (R, t) -> g2o's (t, R) order
    os << std::endl;
  }
  return os ? GTB_OK : GTB_ERR_INVALID;
}

}  // extern "C"
