// Stand-ins for the value types of the 3D scan-matching boundary (see
// cartographer_compat.h for why they exist).  In a full Cartographer tree the real
// headers replace this file:
//   transform::Rigid3f / Eigen::Quaternionf   cartographer/transform/rigid_transform.h:116-207
//   transform::GetYaw, AngleAxisVectorToRotationQuaternion
//                                              cartographer/transform/transform.h:57-61,94-108
//   mapping::HybridGrid                        cartographer/mapping/3d/hybrid_grid.h:68-545
//   proto::FastCorrelativeScanMatcherOptions3D
//       cartographer/mapping/proto/scan_matching/fast_correlative_scan_matcher_options_3d.proto
// All float arithmetic is written operation by operation in the order Eigen 3.3
// evaluates it; the host layer is compiled with -ffp-contract=off.
#ifndef CSM_HOST_CARTOGRAPHER_COMPAT_3D_H_
#define CSM_HOST_CARTOGRAPHER_COMPAT_3D_H_

#include <unordered_map>

#include "cartographer_compat.h"

namespace Eigen {
using VectorXf = std::vector<float>;
using Quaterniond = cartographer::transform::Quaterniond;
struct Array3i {
  int v[3] = {0, 0, 0};
  Array3i() = default;
  Array3i(int x, int y, int z) : v{x, y, z} {}
  int x() const { return v[0]; }
  int y() const { return v[1]; }
  int z() const { return v[2]; }
  int operator[](int i) const { return v[i]; }
};
}  // namespace Eigen

namespace cartographer {
namespace transform {

using Vector3f = sensor::Vector3f;

struct Quaternionf {
  float w_ = 1.f, x_ = 0.f, y_ = 0.f, z_ = 0.f;
  Quaternionf() = default;
  Quaternionf(float w, float x, float y, float z) : w_(w), x_(x), y_(y), z_(z) {}
  float w() const { return w_; }
  float x() const { return x_; }
  float y() const { return y_; }
  float z() const { return z_; }
  // Quaternion product (Eigen/src/Geometry/Quaternion.h, generic quat_product).
  Quaternionf operator*(const Quaternionf& b) const {
    const Quaternionf& a = *this;
    return Quaternionf(a.w_ * b.w_ - a.x_ * b.x_ - a.y_ * b.y_ - a.z_ * b.z_,
                       a.w_ * b.x_ + a.x_ * b.w_ + a.y_ * b.z_ - a.z_ * b.y_,
                       a.w_ * b.y_ + a.y_ * b.w_ + a.z_ * b.x_ - a.x_ * b.z_,
                       a.w_ * b.z_ + a.z_ * b.w_ + a.x_ * b.y_ - a.y_ * b.x_);
  }
  // QuaternionBase::_transformVector
  Vector3f operator*(const Vector3f& v) const {
    float ux = y_ * v.z() - z_ * v.y();
    float uy = z_ * v.x() - x_ * v.z();
    float uz = x_ * v.y() - y_ * v.x();
    ux += ux;
    uy += uy;
    uz += uz;
    const float cx = y_ * uz - z_ * uy;
    const float cy = z_ * ux - x_ * uz;
    const float cz = x_ * uy - y_ * ux;
    return Vector3f((v.x() + w_ * ux) + cx, (v.y() + w_ * uy) + cy, (v.z() + w_ * uz) + cz);
  }
  float squaredNorm() const { return ((x_ * x_ + y_ * y_) + z_ * z_) + w_ * w_; }
  Quaternionf conjugate() const { return Quaternionf(w_, -x_, -y_, -z_); }
  Quaternionf inverse() const {
    const float n2 = squaredNorm();
    return Quaternionf(w_ / n2, -x_ / n2, -y_ / n2, -z_ / n2);
  }
  Quaternionf normalized() const {
    const float n = std::sqrt(squaredNorm());
    return Quaternionf(w_ / n, x_ / n, y_ / n, z_ / n);
  }
};

// Quaterniond::inverse().cast<float>() as one step (conjugate / squaredNorm in double).
inline Quaternionf InverseAsFloat(const Quaterniond& q) {
  const double n2 = ((q.x() * q.x() + q.y() * q.y()) + q.z() * q.z()) + q.w() * q.w();
  return Quaternionf(static_cast<float>(q.w() / n2), static_cast<float>(-q.x() / n2),
                     static_cast<float>(-q.y() / n2), static_cast<float>(-q.z() / n2));
}
inline Quaternionf CastToFloat(const Quaterniond& q) {
  return Quaternionf(static_cast<float>(q.w()), static_cast<float>(q.x()), static_cast<float>(q.y()),
                     static_cast<float>(q.z()));
}

class Rigid3f {
 public:
  Rigid3f() = default;
  Rigid3f(const Vector3f& t, const Quaternionf& r) : translation_(t), rotation_(r) {}
  static Rigid3f Rotation(const Quaternionf& r) { return Rigid3f(Vector3f(), r); }
  static Rigid3f Translation(const Vector3f& t) { return Rigid3f(t, Quaternionf()); }
  static Rigid3f Identity() { return Rigid3f(); }
  const Vector3f& translation() const { return translation_; }
  const Quaternionf& rotation() const { return rotation_; }
  // rigid_transform.h:160-164
  Rigid3f inverse() const {
    const Quaternionf r = rotation_.conjugate();
    const Vector3f t = r * translation_;
    return Rigid3f(Vector3f(-t.x(), -t.y(), -t.z()), r);
  }
  Rigid3d cast_double() const {
    return Rigid3d(Vector3d(translation_.x(), translation_.y(), translation_.z()),
                   Quaterniond(rotation_.w(), rotation_.x(), rotation_.y(), rotation_.z()));
  }

 private:
  Vector3f translation_;
  Quaternionf rotation_;
};
// rigid_transform.h:186-190: rotation product is normalised.
inline Rigid3f operator*(const Rigid3f& l, const Rigid3f& r) {
  const Vector3f rt = l.rotation() * r.translation();
  return Rigid3f(Vector3f(rt.x() + l.translation().x(), rt.y() + l.translation().y(),
                          rt.z() + l.translation().z()),
                 (l.rotation() * r.rotation()).normalized());
}
inline Vector3f operator*(const Rigid3f& t, const Vector3f& p) {
  const Vector3f r = t.rotation() * p;
  return Vector3f(r.x() + t.translation().x(), r.y() + t.translation().y(), r.z() + t.translation().z());
}
inline Rigid3f CastToFloat(const Rigid3d& t) {
  return Rigid3f(Vector3f(static_cast<float>(t.translation().x()), static_cast<float>(t.translation().y()),
                          static_cast<float>(t.translation().z())),
                 CastToFloat(t.rotation()));
}
// transform.h:41-47 with T = float.  The reference's unqualified atan2(float, float) resolves to the
// C library's double atan2 (its translation units include <cmath>, not <math.h>), so the yaw is
// computed in double and narrowed; atan2f would differ in the last bit for ~16 % of inputs.
inline float GetYaw(const Quaternionf& q) {
  const Vector3f d = q * Vector3f(1.f, 0.f, 0.f);
  return static_cast<float>(std::atan2(static_cast<double>(d.y()), static_cast<double>(d.x())));
}
// transform.h:94-108 with T = float; the literals 2. and 1e-8 are doubles.
inline Quaternionf AngleAxisVectorToRotationQuaternion(const Vector3f& a) {
  float scale = 0.5f, w = 1.f;
  const float squared = (a.x() * a.x() + a.y() * a.y()) + a.z() * a.z();
  if (squared > 1e-8) {
    const float norm = std::sqrt(squared);
    scale = static_cast<float>(std::sin(norm / 2.) / norm);
    w = static_cast<float>(std::cos(norm / 2.));
  }
  return Quaternionf(w, scale * a.x(), scale * a.y(), scale * a.z());
}
inline float Norm(const Vector3f& v) { return std::sqrt((v.x() * v.x() + v.y() * v.y()) + v.z() * v.z()); }

}  // namespace transform

namespace mapping {

// What the matcher observes of a HybridGrid: value(), GetCellIndex(), resolution(),
// grid_size() (starts at 128 voxels and doubles until an inserted index fits,
// hybrid_grid.h:255-301) and the iteration over non-default voxels.  The tree of
// 8^3 bricks behind it is replaced by a hash map.
class HybridGrid {
 public:
  explicit HybridGrid(const float resolution) : resolution_(resolution) {}
  float resolution() const { return resolution_; }
  int grid_size() const { return grid_size_; }
  Eigen::Array3i GetCellIndex(const transform::Vector3f& p) const {
    return Eigen::Array3i(common::RoundToInt(p.x() / resolution_), common::RoundToInt(p.y() / resolution_),
                          common::RoundToInt(p.z() / resolution_));
  }
  uint16 value(const Eigen::Array3i& i) const {
    const auto it = cells_.find(Key(i));
    return it == cells_.end() ? 0 : it->second;
  }
  uint16* mutable_value(const Eigen::Array3i& i) {
    for (int a = 0; a < 3; ++a) CSM_CHECK(i[a] >= -(1 << 20) && i[a] < (1 << 20));
    const auto fits = [&]() {
      const int h = grid_size_ >> 1;
      return i.x() >= -h && i.x() < h && i.y() >= -h && i.y() < h && i.z() >= -h && i.z() < h;
    };
    while (!fits()) grid_size_ <<= 1;
    return &cells_[Key(i)];
  }
  // probability_values.h:32-46,88-91
  void SetProbability(const Eigen::Array3i& i, const float probability) {
    const float clamped = probability < 0.1f ? 0.1f : (probability > 0.9f ? 0.9f : probability);
    *mutable_value(i) =
        static_cast<uint16>(common::RoundToInt((clamped - 0.1f) * (32766.f / (0.9f - 0.1f))) + 1);
  }
  // Non-default voxels as flat lists (what HybridGrid::Iterator visits).
  void Export(std::vector<int32_t>* xyz, std::vector<uint16>* values) const {
    xyz->clear();
    values->clear();
    for (const auto& kv : cells_) {
      if (kv.second == 0) continue;
      xyz->push_back(static_cast<int32_t>((kv.first >> 42) & 0x1FFFFF) - (1 << 20));
      xyz->push_back(static_cast<int32_t>((kv.first >> 21) & 0x1FFFFF) - (1 << 20));
      xyz->push_back(static_cast<int32_t>(kv.first & 0x1FFFFF) - (1 << 20));
      values->push_back(kv.second);
    }
  }

 private:
  static uint64_t Key(const Eigen::Array3i& i) {
    return (static_cast<uint64_t>(i.x() + (1 << 20)) << 42) | (static_cast<uint64_t>(i.y() + (1 << 20)) << 21) |
           static_cast<uint64_t>(i.z() + (1 << 20));
  }
  float resolution_;
  int grid_size_ = 128;
  std::unordered_map<uint64_t, uint16> cells_;
};

// The parts of Submap3D the constraint builder reads (mapping/3d/submap_3d.h:43-95):
// the two grids and the rotational histogram are borrowed, not owned.
class Submap3D {
 public:
  Submap3D(const transform::Rigid3d& local_pose, const HybridGrid* high_resolution_hybrid_grid,
           const HybridGrid* low_resolution_hybrid_grid,
           const Eigen::VectorXf* rotational_scan_matcher_histogram)
      : local_pose_(local_pose),
        high_(high_resolution_hybrid_grid),
        low_(low_resolution_hybrid_grid),
        histogram_(rotational_scan_matcher_histogram) {}
  transform::Rigid3d local_pose() const { return local_pose_; }
  const HybridGrid& high_resolution_hybrid_grid() const { return *high_; }
  const HybridGrid& low_resolution_hybrid_grid() const { return *low_; }
  const Eigen::VectorXf& rotational_scan_matcher_histogram() const { return *histogram_; }

 private:
  transform::Rigid3d local_pose_;
  const HybridGrid* high_;
  const HybridGrid* low_;
  const Eigen::VectorXf* histogram_;
};

}  // namespace mapping
}  // namespace cartographer

#endif  // CSM_HOST_CARTOGRAPHER_COMPAT_3D_H_
