// See object.hpp. Reference file:line cited per function (paths relative to /root/reference/).
#include "object.hpp"

#include <cassert>
#include <cmath>
#include <functional>

namespace optimo {

// ---- Curve ---------------------------------------------------------------------------------------------

/// OptiMo/curve.cpp:100-128 -- first segment, in list order, that accepts t; a segment without an end
/// accepts anything.
size_t Curve::FindSegmentIndex(double t) const
{
    for (size_t k = 0; k < segments_.size(); ++k) {
        const Segment& s = segments_[k];
        if (s.end_point_index < 0) return k;
        const double t_end = control_points_[s.end_point_index].time;
        if (s.start_point_index < 0) {
            if (t_end > t) return k;
            continue;
        }
        if (control_points_[s.start_point_index].time <= t && t_end > t) return k;
    }
    return segments_.size() - 1;
}

/// OptiMo/curve.cpp:130-148 -- append the point and split the segment that holds its time.
void Curve::AddControlPoint(const ControlPoint& cp)
{
    control_points_.push_back(cp);
    const size_t k = FindSegmentIndex(cp.time);
    const Segment hit = segments_[k];
    const int fresh = static_cast<int>(control_points_.size()) - 1;
    segments_[k] = Segment{fresh, hit.end_point_index};
    segments_.insert(segments_.begin() + static_cast<long>(k), Segment{hit.start_point_index, fresh});
}

std::vector<int> Curve::ChainOrder() const
{
    std::vector<int> order;
    for (const Segment& s : segments_)
        if (s.end_point_index >= 0) order.push_back(s.end_point_index);
    return order;
}

// ---- Joint ---------------------------------------------------------------------------------------------

Joint::Joint(std::shared_ptr<Joint> parent, const Vec3& offset, const Vec3& rotation, const std::string& name)
    : parent_(parent), offset_(offset), rot_x_(rotation[0]), rot_y_(rotation[1]), rot_z_(rotation[2])
{
    name_ = name;
    SetPhysicalParameters();
}

Joint::Joint(std::shared_ptr<Joint> parent, const Vec3& offset, const Variable& rx, const Variable& ry, const Variable& rz,
             const std::string& name)
    : parent_(parent), offset_(offset), rot_x_(rx), rot_y_(ry), rot_z_(rz)
{
    name_ = name;
    SetPhysicalParameters();
}

std::shared_ptr<Joint> Joint::AddChild(const Vec3& offset, const Vec3& rotation, const std::string& name)
{
    auto child = std::make_shared<Joint>(shared_from_this(), offset, rotation, name);
    children_.push_back(child);
    return child;
}

std::shared_ptr<Joint> Joint::AddChild(std::shared_ptr<Joint> child)
{
    children_.push_back(child);
    return child;
}

/// OptiMo/joint.cpp:90-119 -- a 0.1 x |offset| x 0.1 m box of density 1000 along the bone; COM at half the
/// offset; inertia about the COM = R I_body R^T with R = FromTwoVectors(offset^, y^) (Eigen's quaternion
/// formula: c = a.b, axis = a x b, s = sqrt(2(1+c)), q = (s/2, axis/s)).
void Joint::SetPhysicalParameters()
{
    const double x_body = 0.1, z_body = 0.1;
    const double len = std::sqrt(offset_[0] * offset_[0] + offset_[1] * offset_[1] + offset_[2] * offset_[2]);
    const double y_body = len;
    if (parent_.expired()) {
        mass_ = 0.0;
        center_of_mass_position_parent_ = Vec3{{0, 0, 0}};
        inertia_at_center_parent_.fill(0.0);
        return;
    }
    mass_ = 1000.0 * x_body * z_body * len;
    center_of_mass_position_parent_ = Vec3{{0.5 * offset_[0], 0.5 * offset_[1], 0.5 * offset_[2]}};
    const double Ib[3] = {(mass_ / 12.0) * (y_body * y_body + z_body * z_body), (mass_ / 12.0) * (x_body * x_body + z_body * z_body),
                          (mass_ / 12.0) * (x_body * x_body + y_body * y_body)};
    // a = offset normalised (Eigen normalized(): only when the squared norm is positive), b = unit y
    double a[3] = {offset_[0], offset_[1], offset_[2]};
    const double z2 = a[0] * a[0] + a[1] * a[1] + a[2] * a[2];
    if (z2 > 0.0) { const double n = std::sqrt(z2); a[0] /= n; a[1] /= n; a[2] /= n; }
    // FromTwoVectors normalises its arguments again (Eigen/src/Geometry/Quaternion.h), which moves a by an ulp
    const double z2b = a[0] * a[0] + a[1] * a[1] + a[2] * a[2];
    if (z2b > 0.0) { const double n = std::sqrt(z2b); a[0] /= n; a[1] /= n; a[2] /= n; }
    const double c = a[1];  // dot(b, a)
    double R[3][3];
    if (c < -1.0 + 1e-12) {
        // Antiparallel: Eigen picks (by SVD) some axis perpendicular to y and rotates by pi. I_body is
        // diag(A, B, A), which every such rotation leaves unchanged, so any choice gives the same inertia.
        R[0][0] = 1; R[0][1] = 0; R[0][2] = 0; R[1][0] = 0; R[1][1] = -1; R[1][2] = 0; R[2][0] = 0; R[2][1] = 0; R[2][2] = -1;
    } else {
        const double ax = a[1] * 0.0 - a[2] * 1.0, ay = a[2] * 0.0 - a[0] * 0.0, az = a[0] * 1.0 - a[1] * 0.0;  // a x b
        const double s = std::sqrt((1.0 + c) * 2.0), invs = 1.0 / s;
        const double x = ax * invs, y = ay * invs, z = az * invs, w = s * 0.5;
        const double tx = 2.0 * x, ty = 2.0 * y, tz = 2.0 * z;
        const double twx = tx * w, twy = ty * w, twz = tz * w, txx = tx * x, txy = ty * x, txz = tz * x, tyy = ty * y,
                     tyz = tz * y, tzz = tz * z;
        R[0][0] = 1.0 - (tyy + tzz); R[0][1] = txy - twz;         R[0][2] = txz + twy;
        R[1][0] = txy + twz;         R[1][1] = 1.0 - (txx + tzz); R[1][2] = tyz - twx;
        R[2][0] = txz - twy;         R[2][1] = tyz + twx;         R[2][2] = 1.0 - (txx + tyy);
    }
    // (R * diag(Ib)) * R^T
    double RI[3][3];
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) RI[i][j] = R[i][j] * Ib[j];
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j)
            inertia_at_center_parent_[3 * i + j] = RI[i][0] * R[j][0] + RI[i][1] * R[j][1] + RI[i][2] * R[j][2];
}

// ---- IkHandler -----------------------------------------------------------------------------------------

/// OptiMo/ikhandler.cpp:45-67,77-101 -- the chain is the path root .. end effector; every joint of it except
/// the end effector becomes IK-driven and loses its own rotation keys.
void IkHandler::Initialize(std::shared_ptr<Joint> ik_root, std::shared_ptr<Joint> end_effector)
{
    ik_root_ = ik_root;
    end_effector_ = end_effector;
    target_joints_.clear();
    for (std::shared_ptr<Joint> j = end_effector;; j = j->parent_.lock()) {
        assert(j);
        target_joints_.insert(target_joints_.begin(), j);
        if (j == ik_root) break;
    }
    for (auto& j : target_joints_) {
        if (j == end_effector_) continue;
        j->rot_x_.curve_ = Curve();
        j->rot_y_.curve_ = Curve();
        j->rot_z_.curve_ = Curve();
    }
}

// ---- Object --------------------------------------------------------------------------------------------

Object::Object() { root_joint_ = std::make_shared<Joint>(nullptr, Vec3{{0, 0, 0}}, Vec3{{0, 0, 0}}, "root"); }

std::vector<std::shared_ptr<Joint>> Object::GetJoints() const
{
    std::vector<std::shared_ptr<Joint>> out;
    std::function<void(const std::shared_ptr<Joint>&)> visit = [&](const std::shared_ptr<Joint>& j) {
        out.push_back(j);
        for (const auto& c : j->children_) visit(c);
    };
    if (root_joint_) visit(root_joint_);
    return out;
}

std::vector<std::shared_ptr<Item>> Object::GetItems() const
{
    std::vector<std::shared_ptr<Item>> out;
    for (const auto& j : GetJoints()) out.push_back(j);
    for (const auto& h : ik_handlers_) out.push_back(h);
    return out;
}

std::shared_ptr<Item> Object::GetItemByName(const std::string& name) const
{
    for (const auto& it : GetItems()) if (it->name_ == name) return it;
    return nullptr;
}

std::shared_ptr<Joint> Object::GetJointByName(const std::string& name) const
{
    for (const auto& j : GetJoints()) if (j->name_ == name) return j;
    return nullptr;
}

int Object::GetItemIndexByName(const std::string& name) const
{
    int k = 0;
    for (const auto& it : GetItems()) { if (it->name_ == name) return k; ++k; }
    return -1;
}

VecN Object::GetDescriptor() const
{
    VecN x;
    for (const auto& it : GetItems())
        for (Variable* v : it->GetVariablePointers()) {
            if (!v->IsKeyframed()) continue;
            for (const ControlPoint& p : v->curve_.GetControlPoints()) {
                x.push_back(p.time); x.push_back(p.value); x.push_back(p.theta);
                x.push_back(p.time_backward); x.push_back(p.time_forward);
            }
        }
    return x;
}

void Object::SetDescriptor(const VecN& x)
{
    size_t k = 0;
    for (const auto& it : GetItems())
        for (Variable* v : it->GetVariablePointers()) {
            if (!v->IsKeyframed()) continue;
            for (ControlPoint& p : v->curve_.GetControlPoints()) {
                assert(k + 5 <= x.size());
                p.time = x[k]; p.value = x[k + 1]; p.theta = x[k + 2]; p.time_backward = x[k + 3]; p.time_forward = x[k + 4];
                k += 5;
            }
        }
    if (is_cyclic_) MakeMotionCyclic();
}

void Object::MakeMotionCyclic()
{
    for (const auto& it : GetItems())
        for (Variable* v : it->GetVariablePointers()) {
            if (!v->IsKeyframed()) continue;
            auto& pts = v->curve_.GetControlPoints();
            const double th = 0.5 * (pts.front().theta + pts.back().theta);  // curve.cpp:179-188
            pts.front().theta = th;
            pts.back().theta = th;
        }
}

}  // namespace optimo
