// Headless scene model of the B200 OptiMo objective path: the host-side mirror of the reference's
// Item / Variable / Curve / Joint / IkHandler / Object classes, reduced to what the objective path needs
// (structure, descriptor packing, physical parameters). All evaluation (curves, IK, FK, dynamics) happens
// on the GPU behind include/omo.h; nothing here samples a curve.
//
// Reference (paths relative to /root/reference/):
//   ControlPoint / Curve      OptiMo/curve.h:21-106, curve.cpp:95-148,179-188
//   Variable                  OptiMo/variable.h
//   Joint                     OptiMo/joint.h, joint.cpp:90-119 (SetPhysicalParameters)
//   IkHandler                 OptiMo/ikhandler.h, ikhandler.cpp:45-101
//   Object                    OptiMo/object.h, object.cpp:27-186
#ifndef OPTIMO_B200_OBJECT_HPP
#define OPTIMO_B200_OBJECT_HPP

#include <array>
#include <memory>
#include <string>
#include <vector>

namespace optimo {

using VecN = std::vector<double>;
using Vec3 = std::array<double, 3>;

struct ControlPoint {
    double time, value, theta, time_backward, time_forward;
    ControlPoint(double t, double v, double th = 0.0, double tb = -1.0, double tf = 1.0)  // curve.h:56-63 defaults
        : time(t), value(v), theta(th), time_backward(tb), time_forward(tf) {}
};

/// Control points in insertion order plus the segment chain (time order at insertion).
class Curve {
public:
    struct Segment { int start_point_index, end_point_index; };  // -1 == open end
    Curve() { segments_.push_back(Segment{-1, -1}); }
    void AddControlPoint(const ControlPoint& cp);
    void AddControlPoint(double t, double v) { AddControlPoint(ControlPoint(t, v)); }
    const std::vector<ControlPoint>& GetControlPoints() const { return control_points_; }
    std::vector<ControlPoint>& GetControlPoints() { return control_points_; }
    const std::vector<Segment>& GetSegments() const { return segments_; }
    /// Control-point indices in chain order (the end point of every segment but the last).
    std::vector<int> ChainOrder() const;

private:
    size_t FindSegmentIndex(double t) const;
    std::vector<ControlPoint> control_points_;
    std::vector<Segment> segments_;
};

struct Variable {
    explicit Variable(double v = 0.0) : var_(v) {}
    double var_;
    Curve curve_;
    bool IsKeyframed() const { return !curve_.GetControlPoints().empty(); }
    int GetNumOfControlPoints() const { return static_cast<int>(curve_.GetControlPoints().size()); }
};

struct Item {
    virtual ~Item() {}
    std::string name_;
    bool is_selected_ = false;
    virtual std::vector<Variable*> GetVariablePointers() = 0;
};

class IkHandler;

class Joint final : public Item, public std::enable_shared_from_this<Joint> {
public:
    Joint(std::shared_ptr<Joint> parent, const Vec3& offset, const Vec3& rotation, const std::string& name);
    Joint(std::shared_ptr<Joint> parent, const Vec3& offset, const Variable& rx, const Variable& ry, const Variable& rz,
          const std::string& name);

    std::vector<std::shared_ptr<Joint>> children_;
    std::weak_ptr<Joint> parent_;
    std::shared_ptr<Joint> AddChild(const Vec3& offset, const Vec3& rotation, const std::string& name);
    std::shared_ptr<Joint> AddChild(std::shared_ptr<Joint> child);

    Vec3 offset_;
    Variable rot_x_, rot_y_, rot_z_;
    std::vector<Variable*> GetVariablePointers() override { return {&rot_x_, &rot_y_, &rot_z_}; }

    bool IsControlledByIk() const { return !associated_ik_handler_.expired(); }
    std::weak_ptr<IkHandler> associated_ik_handler_;

    Vec3 center_of_mass_position_parent_{{0, 0, 0}};
    std::array<double, 9> inertia_at_center_parent_{{0, 0, 0, 0, 0, 0, 0, 0, 0}};  // row-major
    double mass_ = 0.0;
    /// joint.cpp:90-119. Call again after changing offset_.
    void SetPhysicalParameters();
};

class IkHandler final : public Item, public std::enable_shared_from_this<IkHandler> {
public:
    explicit IkHandler(const std::string& name) { name_ = name; }
    void Initialize(std::shared_ptr<Joint> ik_root, std::shared_ptr<Joint> end_effector);  // ikhandler.cpp:77-101
    Variable pos_x_, pos_y_, pos_z_;
    std::vector<Variable*> GetVariablePointers() override { return {&pos_x_, &pos_y_, &pos_z_}; }
    std::shared_ptr<Joint> ik_root_, end_effector_;
    std::vector<std::shared_ptr<Joint>> target_joints_;  // root .. end effector
};

class Object {
public:
    Object();
    std::vector<std::shared_ptr<Joint>> GetJoints() const;        // DFS pre-order, object.cpp:27-40
    std::vector<std::shared_ptr<IkHandler>> GetIkHandlers() const { return ik_handlers_; }
    std::vector<std::shared_ptr<Item>> GetItems() const;          // joints then IK handlers, object.cpp:47-61
    void AddIkHandler(std::shared_ptr<IkHandler> h) { ik_handlers_.push_back(h); }
    std::shared_ptr<Item> GetItemByName(const std::string& name) const;
    std::shared_ptr<Joint> GetJointByName(const std::string& name) const;
    int GetItemIndexByName(const std::string& name) const;

    VecN GetDescriptor() const;          // object.cpp:110-145
    void SetDescriptor(const VecN& x);   // object.cpp:147-172 (+ MakeMotionCyclic when is_cyclic_)
    void ClearIkHandlers() { ik_handlers_.clear(); }
    std::shared_ptr<Joint> GetRootJoint() const { return root_joint_; }
    void SetRootJoint(std::shared_ptr<Joint> r) { root_joint_ = r; }
    bool is_cyclic_ = false;

private:
    std::shared_ptr<Joint> root_joint_;
    std::vector<std::shared_ptr<IkHandler>> ik_handlers_;
    void MakeMotionCyclic();  // object.cpp:174-186
};

}  // namespace optimo
#endif
