// See core.hpp. Reference file:line cited per function (paths relative to /root/reference/).
#include "core.hpp"

#include <algorithm>
#include <cstring>
#include <cassert>
#include <cstdlib>
#include <fstream>
#include <iterator>

#include "json.hpp"

namespace optimo {

namespace {

constexpr double kPi = 3.14159265358979323846;

// ---- scene JSON (OptiMo/core_io.cpp:100-151) ------------------------------------------------------------
Variable InterpretJson2Variable(const Json& json)
{
    Variable var(json["var"].number_value());
    for (const Json& j : json["curve"].array_items())
        var.curve_.AddControlPoint(ControlPoint(j[0].number_value(), j[1].number_value(), j[2].number_value(), j[3].number_value(),
                                                j[4].number_value()));
    return var;
}

std::shared_ptr<Joint> InterpretJson2Joint(const Json& json, std::shared_ptr<Joint> parent = nullptr)
{
    const Vec3 offset{{json["offset"][0].number_value(), json["offset"][1].number_value(), json["offset"][2].number_value()}};
    auto joint = std::make_shared<Joint>(parent, offset, InterpretJson2Variable(json["rot_x"]), InterpretJson2Variable(json["rot_y"]),
                                         InterpretJson2Variable(json["rot_z"]), json["name"].string_value());
    for (const Json& c : json["children"].array_items()) joint->AddChild(InterpretJson2Joint(c, joint));
    return joint;
}

void InterpretJson2IkHandlerAndAddToObject(const Json& json, const std::shared_ptr<Object>& object)
{
    auto root = object->GetJointByName(json["root_name"].string_value());
    auto end = object->GetJointByName(json["end_effector_name"].string_value());
    if (!root || !end) throw std::runtime_error("IK handler refers to an unknown joint");
    auto handler = std::make_shared<IkHandler>(json["name"].string_value());
    handler->Initialize(root, end);
    for (auto& j : handler->target_joints_)
        if (j != end) j->associated_ik_handler_ = handler;
    handler->pos_x_ = InterpretJson2Variable(json["pos_x"]);
    handler->pos_y_ = InterpretJson2Variable(json["pos_y"]);
    handler->pos_z_ = InterpretJson2Variable(json["pos_z"]);
    object->AddIkHandler(handler);
}

Json Variable2Json(const Variable& var)
{
    Json curve = Json::array();
    for (const ControlPoint& p : var.curve_.GetControlPoints()) {
        Json cp = Json::array();
        for (double v : {p.time, p.value, p.theta, p.time_backward, p.time_forward}) cp.push_back(Json::number(v));
        curve.push_back(cp);
    }
    Json o = Json::object();
    o.set("var", Json::number(var.var_));
    o.set("curve", curve);
    return o;
}

Json Joint2Json(const std::shared_ptr<Joint>& joint)
{
    Json children = Json::array();
    for (const auto& c : joint->children_) children.push_back(Joint2Json(c));
    Json off = Json::array();
    for (double v : joint->offset_) off.push_back(Json::number(v));
    Json o = Json::object();
    o.set("name", Json::string(joint->name_));
    o.set("offset", off);
    o.set("rot_x", Variable2Json(joint->rot_x_));
    o.set("rot_y", Variable2Json(joint->rot_y_));
    o.set("rot_z", Variable2Json(joint->rot_z_));
    o.set("children", children);
    return o;
}

}  // namespace

// ---- construction -------------------------------------------------------------------------------------------

/// OptiMo/core.cpp:23-103 -- the default humanoid, default frozen dofs (t, v) and weights. Unlike the
/// reference this constructor evaluates nothing: the engine is created on first use (never before main).
Core::Core()
{
    object_ = std::make_shared<Object>();
    auto V = [](double x, double y, double z) { return Vec3{{x, y, z}}; };
    const Vec3 Z = V(0, 0, 0);
    auto root = object_->GetRootJoint();
    auto spine = root->AddChild(V(0.0, 0.2, 0.0), Z, "spine");
    auto neck = spine->AddChild(V(0.0, 0.5, 0.0), Z, "neck");
    neck->AddChild(V(0.0, 0.5, 0.0), Z, "head");
    auto r_shld = spine->AddChild(V(-0.2, 0.5, 0.0), V(0.0, 0.0, +kPi / 6.0), "right_shoulder");
    auto r_arm = r_shld->AddChild(V(-0.5, 0.0, 0.0), Z, "right_arm");
    r_arm->AddChild(V(-0.5, 0.0, 0.0), Z, "right_hand");
    auto l_shld = spine->AddChild(V(+0.2, 0.5, 0.0), V(0.0, 0.0, -kPi / 6.0), "left_shoulder");
    auto l_arm = l_shld->AddChild(V(+0.5, 0.0, 0.0), V(0.0, -0.1 * kPi, 0.0), "left_arm");
    l_arm->AddChild(V(+0.5, 0.0, 0.0), Z, "left_hand");
    auto l_uleg = root->AddChild(V(+0.15, -0.10, 0.00), Z, "left_upleg");
    auto l_leg = l_uleg->AddChild(V(0.03, -0.50, 0.05), Z, "left_leg");
    auto l_foot = l_leg->AddChild(V(0.00, -0.50, -0.05), Z, "left_foot");
    l_foot->AddChild(V(0.00, -0.05, 0.15), Z, "left_toe");
    auto r_uleg = root->AddChild(V(-0.15, -0.10, 0.00), Z, "right_upleg");
    auto r_leg = r_uleg->AddChild(V(-0.03, -0.50, 0.05), Z, "right_leg");
    auto r_foot = r_leg->AddChild(V(0.00, -0.50, -0.05), Z, "right_foot");
    r_foot->AddChild(V(0.00, -0.05, 0.15), Z, "right_toe");

    struct Key { double t, v; };
    auto keys = [](Variable& var, std::initializer_list<Key> ks) { for (const Key& k : ks) var.curve_.AddControlPoint(k.t, k.v); };
    keys(spine->rot_z_, {{4, 0.0 * kPi}, {10, 0.1 * kPi}, {28, 0.1 * kPi}, {36, 0.0 * kPi}});
    keys(r_shld->rot_z_, {{4, kPi / 4.0}, {10, kPi / 3.0}, {28, kPi / 3.0}, {36, kPi / 4.0}});
    keys(l_arm->rot_z_, {{4, 0.0 * kPi}, {12, kPi / 6.0}, {20, kPi / 8.0}, {28, kPi / 6.0}, {36, 0.0 * kPi}});
    keys(l_shld->rot_z_, {{4, -kPi / 4.0}, {12, 0.4 * kPi}, {20, 0.0 * kPi}, {26, 0.4 * kPi}, {36, -kPi / 4.0}});
    keys(r_shld->rot_y_, {{4, 0.0 * kPi}, {10, -0.1 * kPi}, {28, -0.1 * kPi}, {36, 0.0 * kPi}});
    keys(r_arm->rot_y_, {{4, 0.1 * kPi}, {10, 0.2 * kPi}, {28, 0.2 * kPi}, {36, 0.1 * kPi}});
    root->offset_ = V(0.0, -0.6, 0.0);

    const int n = static_cast<int>(object_->GetDescriptor().size()) / 5;
    freezed_dofs_.assign(static_cast<size_t>(n) * 5, false);
    for (int i = 0; i < n; ++i) { freezed_dofs_[i * 5 + 0] = true; freezed_dofs_[i * 5 + 1] = true; }
    per_item_weights_.assign(object_->GetItems().size(), 0.0);
    weight_trajectory_ = 0.0;
    x_original_ = object_->GetDescriptor();

    const char* dev = std::getenv("OPTIMO_DEVICE");
    if (!dev) dev = std::getenv("LOCAL_RANK");
    if (dev) device_ = std::atoi(dev);
}

Core::~Core()
{
    if (multi_) { omo_destroy_multi(multi_); engine_ = nullptr; }
    if (engine_) omo_destroy(engine_);
}

void Core::SetDevice(int device)
{
    if (device != device_) { device_ = device; InitializeDynamicsModel(); }
}

// ---- IO ---------------------------------------------------------------------------------------------------------

void Core::ReadJson(const std::string& file_path)
{
    std::ifstream in(file_path);
    if (!in) throw std::runtime_error("cannot open " + file_path);
    ReadJsonString(std::string(std::istreambuf_iterator<char>(in), std::istreambuf_iterator<char>()));
}

/// core_io.cpp:228-250 (ApplyJsonToScene) + :154-163 (rebuild the dynamics model). The runtime state that the
/// reference keeps only in UI-driven members (frozen dofs, weights, kernels, x_original) is
/// reset to the Core() defaults for the new scene.
void Core::ApplyJsonToScene(const std::string& json_text)
{
    const Json json = Json::parse(json_text);
    object_->SetRootJoint(nullptr);
    object_->ClearIkHandlers();
    min_frame_ = json["min_frame"].int_value();
    max_frame_ = json["max_frame"].int_value();
    frames_per_second_ = json["frames_per_second"].int_value();
    if (json["joints"].type() != Json::Object) throw std::runtime_error("scene JSON has no \"joints\" object");
    object_->SetRootJoint(InterpretJson2Joint(json["joints"]));
    for (const Json& h : json["ik_handlers"].array_items()) InterpretJson2IkHandlerAndAddToObject(h, object_);
    ClearOptimizationCache();
}

void Core::ReadJsonString(const std::string& json_text)
{
    ApplyJsonToScene(json_text);
    const int n = static_cast<int>(object_->GetDescriptor().size()) / 5;
    freezed_dofs_.assign(static_cast<size_t>(n) * 5, false);
    for (int i = 0; i < n; ++i) { freezed_dofs_[i * 5 + 0] = true; freezed_dofs_[i * 5 + 1] = true; }
    per_item_weights_.assign(object_->GetItems().size(), 0.0);
    weight_trajectory_ = 0.0;
    kernels_.clear();
    x_original_ = object_->GetDescriptor();
    InitializeDynamicsModel();
}

std::string Core::ConvertCurrentSceneToJson() const
{
    Json iks = Json::array();
    for (const auto& h : object_->GetIkHandlers()) {
        Json o = Json::object();
        o.set("name", Json::string(h->name_));
        o.set("pos_x", Variable2Json(h->pos_x_));
        o.set("pos_y", Variable2Json(h->pos_y_));
        o.set("pos_z", Variable2Json(h->pos_z_));
        o.set("root_name", Json::string(h->ik_root_->name_));
        o.set("end_effector_name", Json::string(h->end_effector_->name_));
        iks.push_back(o);
    }
    Json obj = Json::object();
    obj.set("min_frame", Json::number(min_frame_));
    obj.set("max_frame", Json::number(max_frame_));
    obj.set("frames_per_second", Json::number(frames_per_second_));
    obj.set("joints", Joint2Json(object_->GetRootJoint()));
    obj.set("ik_handlers", iks);
    return obj.dump();
}

void Core::WriteJson(const std::string& file_path) const
{
    std::ofstream out(file_path);
    out << ConvertCurrentSceneToJson();
}

// ---- history / motion export (next row N3) ------------------------------------------------------------------------

void Core::WriteHistory(const std::string& file_path) const
{
    std::ofstream out(file_path);
    out << "[";
    bool first = true;
    for (const std::string& stamp : history_) {
        if (!first) out << ",";
        first = false;
        out << stamp;
    }
    out << "]";
}

void Core::PushHistory() { history_.push_back(ConvertCurrentSceneToJson()); }

void Core::PopHistory()
{
    if (history_.size() <= 1) return;   // "Cannot undo any more." (core_io.cpp:198-202)
    history_.pop_back();
    ApplyJsonToScene(history_.back());
    // The reference leaves model_ as it is (an undo never changes the skeleton in its UI); here the engine snapshots the
    // scene, so it is rebuilt whenever the stamp differs structurally -- InitializeDynamicsModel just drops it.
    const size_t n = object_->GetDescriptor().size();
    if (x_original_.size() != n) {
        x_original_ = object_->GetDescriptor();
        freezed_dofs_.assign(n, false);
        for (size_t i = 0; i < n / 5; ++i) { freezed_dofs_[i * 5] = true; freezed_dofs_[i * 5 + 1] = true; }
        per_item_weights_.assign(object_->GetItems().size(), 0.0);
    }
    InitializeDynamicsModel();
}

std::string Core::FrameJson(int frame, const double* euler) const
{
    Json joints = Json::array();
    int i = 0;
    for (const auto& joint : object_->GetJoints()) {
        Json j = Json::object();
        j.set("name", Json::string(joint->name_));
        j.set("r_x", Json::number(euler[3 * i + 0]));
        j.set("r_y", Json::number(euler[3 * i + 1]));
        j.set("r_z", Json::number(euler[3 * i + 2]));
        joints.push_back(j);
        ++i;
    }
    Json f = Json::object();
    f.set("frame", Json::number(frame));
    f.set("joints", joints);
    return f.dump();
}

std::string Core::ConvertSpecificFrameToJson(int t)
{
    const double tt = t;
    const std::vector<double> e = CalculateEulerAngles(&tt, 1);
    return FrameJson(t, e.data());
}

std::string Core::ConvertMotionToJson()
{
    // every frame's angles in ONE launch (the reference evaluates curve by curve, frame by frame: core_io.cpp:256-262)
    std::vector<double> times;
    for (int t = min_frame_; t <= max_frame_; ++t) times.push_back(t);
    const std::vector<double> e = CalculateEulerAngles(times.data(), static_cast<int>(times.size()));
    const size_t w = 3 * object_->GetJoints().size();
    std::string out = "[";
    for (int t = min_frame_; t <= max_frame_; ++t) {
        if (t != min_frame_) out += ", ";
        out += FrameJson(t, e.data() + (size_t)(t - min_frame_) * w);
    }
    out += "]";
    return out;
}

void Core::ExportMotion(const std::string& file_path)
{
    std::ofstream out(file_path);
    out << ConvertMotionToJson();
}

// ---- weights / frozen dofs ------------------------------------------------------------------------------------------

void Core::UpdateFreezedDofsFromSelection()
{
    const int n = static_cast<int>(object_->GetDescriptor().size()) / 5;
    freezed_dofs_.assign(static_cast<size_t>(n) * 5, true);
    int cp = 0;
    for (const auto& item : object_->GetItems()) {
        const int begin = cp;
        for (Variable* v : item->GetVariablePointers()) cp += v->GetNumOfControlPoints();
        if (!item->is_selected_) continue;
        for (int i = begin; i < cp; ++i) { freezed_dofs_[i * 5 + 2] = false; freezed_dofs_[i * 5 + 3] = false; freezed_dofs_[i * 5 + 4] = false; }
    }
}

void Core::UpdateGlobalWeight(double weight) { weight_trajectory_ = weight; }

void Core::UpdatePerItemWeight(double weight, const std::string& item_name)
{
    const int index = object_->GetItemIndexByName(item_name);
    if (index < 0) throw std::invalid_argument("no item named " + item_name);
    per_item_weights_[index] = weight;
}

VecN Core::CalculateParameterWeights() const
{
    const int n = static_cast<int>(object_->GetDescriptor().size()) / 5;
    VecN w(static_cast<size_t>(n) * 5, weight_trajectory_);
    int item_index = 0, cp = 0;
    for (const auto& item : object_->GetItems()) {
        for (Variable* v : item->GetVariablePointers()) {
            const int np = v->GetNumOfControlPoints();
            for (int k = cp * 5; k < (cp + np) * 5; ++k) w[k] += per_item_weights_[item_index];
            cp += np;
        }
        ++item_index;
    }
    return w;
}

void Core::UpdateOriginalParameters(const VecN& x_original) { x_original_ = x_original; }

double Core::GetCostWeight(double t) const
{
    double w = 1.0;
    for (const GaussianKernel& k : kernels_) w -= k.GetValue(t);
    return std::max(w, 0.0);
}

// ---- engine ---------------------------------------------------------------------------------------------------------------

void Core::Check(omo_status st, const char* what)
{
    if (st == OMO_OK) return;
    throw engine_error(std::string(what) + ": " + omo_last_error(engine_));
}

void Core::InitializeDynamicsModel()
{
    std::lock_guard<std::recursive_mutex> eg(engine_mutex_);
    if (multi_) { omo_destroy_multi(multi_); multi_ = nullptr; engine_ = nullptr; }   // the multi engine owned engine_
    if (engine_) { omo_destroy(engine_); engine_ = nullptr; }
    ClearOptimizationCache();
    have_last_ = false;
}

void Core::SetDevices(const std::vector<int>& devices)
{
    std::lock_guard<std::recursive_mutex> eg(engine_mutex_);
    if (devices == devices_) return;
    devices_ = devices;
    if (!devices_.empty()) device_ = devices_[0];
    InitializeDynamicsModel();
}

omo_multi* Core::MultiEngine()
{
    std::lock_guard<std::recursive_mutex> eg(engine_mutex_);
    if (devices_.size() < 2) return nullptr;
    Engine();
    return multi_;
}

/// Flattens the object graph into the tables of omo_scene_desc: what Core::InitializeDynamicsModel
/// (core_rbdl.cpp:200-262) feeds to RBDL plus the channel / control-point layout of Object::GetDescriptor.
omo_ctx* Core::Engine()
{
    std::lock_guard<std::recursive_mutex> eg(engine_mutex_);
    if (engine_) return engine_;
    const auto joints = object_->GetJoints();
    const auto iks = object_->GetIkHandlers();
    const int J = static_cast<int>(joints.size());
    std::vector<int> parent(J, -1), ik_root, ik_end, first_cp, n_cp_ch, chain;
    std::vector<double> offset(3 * J), mass(J), com(3 * J), inertia(9 * J), chan_const;
    auto index_of = [&](const std::shared_ptr<Joint>& j) { return static_cast<int>(std::find(joints.begin(), joints.end(), j) - joints.begin()); };
    for (int k = 0; k < J; ++k) {
        const auto& j = joints[k];
        if (auto p = j->parent_.lock()) parent[k] = index_of(p);
        for (int a = 0; a < 3; ++a) { offset[3 * k + a] = j->offset_[a]; com[3 * k + a] = j->center_of_mass_position_parent_[a]; }
        mass[k] = j->mass_;
        for (int a = 0; a < 9; ++a) inertia[9 * k + a] = j->inertia_at_center_parent_[a];
    }
    int cp_total = 0;
    auto add_channel = [&](const Variable& v) {
        chan_const.push_back(v.var_);
        const int n = v.GetNumOfControlPoints();
        first_cp.push_back(n > 0 ? cp_total : -1);
        n_cp_ch.push_back(n);
        for (int idx : v.curve_.ChainOrder()) chain.push_back(cp_total + idx);
        cp_total += n;
    };
    for (const auto& j : joints) { add_channel(j->rot_x_); add_channel(j->rot_y_); add_channel(j->rot_z_); }
    for (const auto& h : iks) {
        add_channel(h->pos_x_); add_channel(h->pos_y_); add_channel(h->pos_z_);
        ik_root.push_back(index_of(h->ik_root_));
        ik_end.push_back(index_of(h->end_effector_));
    }
    omo_scene_desc d;
    d.n_joints = J; d.parent = parent.data(); d.offset = offset.data(); d.mass = mass.data(); d.com = com.data(); d.inertia = inertia.data();
    d.n_ik = static_cast<int>(iks.size()); d.ik_root = ik_root.data(); d.ik_end = ik_end.data();
    d.n_channels = static_cast<int>(chan_const.size()); d.chan_const = chan_const.data(); d.chan_first_cp = first_cp.data();
    d.chan_n_cp = n_cp_ch.data(); d.n_cp = cp_total; d.chain = chain.data();
    if (probe_only_ && debug_table_out_) {
        // host-only: the flattened device scene table (include/omo.h: omo_debug_scene_table); no device is touched
        probe_status_ = static_cast<int>(omo_debug_scene_table(&d, debug_table_out_, debug_table_cap_, debug_sub_end_out_));
        return nullptr;
    }
    if (probe_only_) {
        // host-only check of the scene-specialised kernel (include/omo.h: omo_jit_probe); no device is touched
        std::vector<char> log(1 << 16, 0);
        const omo_status ps = omo_jit_probe(&d, probe_fk_, probe_count_, probe_prefix_.empty() ? nullptr : probe_prefix_.c_str(), log.data(),
                                            static_cast<int>(log.size()));
        probe_log_ = log.data();
        probe_status_ = static_cast<int>(ps);
        return nullptr;
    }
    if (devices_.size() >= 2) {
        const omo_status st = omo_create_multi(&d, devices_.data(), static_cast<int>(devices_.size()), &multi_);
        if (st != OMO_OK) { multi_ = nullptr; throw engine_error(std::string("omo_create_multi: ") + omo_multi_last_error(nullptr)); }
        engine_ = omo_multi_ctx(multi_, 0);
        return engine_;
    }
    const omo_status st = omo_create(&d, device_, &engine_);
    if (st != OMO_OK) { engine_ = nullptr; throw engine_error(std::string("omo_create: ") + omo_last_error(nullptr)); }
    return engine_;
}

int Core::ProbeSpecialisedKernel(bool fk, int flags, const std::string& dump_prefix, std::string* log)
{
    std::lock_guard<std::recursive_mutex> eg(engine_mutex_);
    omo_ctx* keep = engine_;
    engine_ = nullptr;
    probe_only_ = true; probe_fk_ = fk ? 1 : 0; probe_count_ = flags; probe_prefix_ = dump_prefix;
    try { Engine(); } catch (...) { probe_only_ = false; engine_ = keep; throw; }
    probe_only_ = false;
    engine_ = keep;
    if (log) *log = probe_log_;
    return probe_status_;
}

int Core::DebugSceneTable(void* out, unsigned long long cap, int* sub_end_out)
{
    std::lock_guard<std::recursive_mutex> eg(engine_mutex_);
    omo_ctx* keep = engine_;
    engine_ = nullptr;
    probe_only_ = true; debug_table_out_ = out; debug_table_cap_ = cap; debug_sub_end_out_ = sub_end_out;
    try { Engine(); } catch (...) { probe_only_ = false; debug_table_out_ = nullptr; engine_ = keep; throw; }
    probe_only_ = false; debug_table_out_ = nullptr;
    engine_ = keep;
    return probe_status_;
}

/// Snapshots the state the objective reads (weights, kernels, cyclic flag, window, x_original) into the engine.
/// Called at the top of every evaluation so that mid-solve slider changes behave as in the reference.
void Core::SyncEngine(double min_t, double max_t, const VecN& x_original)
{
    Check(omo_set_window(Engine(), min_t, max_t, 0.20), "omo_set_window");   // dt = 0.20, core_rbdl.cpp:351
    SyncEngineState(x_original);
}

void Core::SyncEngineState(const VecN& x_original)
{
    omo_ctx* e = Engine();
    const VecN w = CalculateParameterWeights();
    if (x_original.size() != w.size()) throw std::invalid_argument("x_original has the wrong length");
    std::vector<omo_gauss> ks;
    for (const GaussianKernel& k : kernels_) ks.push_back(omo_gauss{k.amp, k.mu, k.sigma});
    Check(omo_set_weights(e, w.data(), static_cast<int>(w.size()), weight_trajectory_, ks.data(), static_cast<int>(ks.size()),
                          object_->is_cyclic_ ? 1 : 0), "omo_set_weights");
    Check(omo_set_original(e, x_original.data(), static_cast<int>(x_original.size())), "omo_set_original");
}

/// The same snapshot pushed to every device of the multi engine (batches and gradients only).
void Core::SyncMulti(double min_t, double max_t, const VecN& x_original)
{
    const VecN w = CalculateParameterWeights();
    if (x_original.size() != w.size()) throw std::invalid_argument("x_original has the wrong length");
    std::vector<omo_gauss> ks;
    for (const GaussianKernel& k : kernels_) ks.push_back(omo_gauss{k.amp, k.mu, k.sigma});
    auto check = [&](omo_status st, const char* what) { if (st != OMO_OK) throw engine_error(std::string(what) + ": " + omo_multi_last_error(multi_)); };
    check(omo_multi_set_window(multi_, min_t, max_t, 0.20), "omo_multi_set_window");
    check(omo_multi_set_weights(multi_, w.data(), static_cast<int>(w.size()), weight_trajectory_, ks.data(), static_cast<int>(ks.size()),
                                object_->is_cyclic_ ? 1 : 0), "omo_multi_set_weights");
    check(omo_multi_set_original(multi_, x_original.data(), static_cast<int>(x_original.size())), "omo_multi_set_original");
}

double Core::CalculateFullObjective(double min_t, double max_t, const VecN& x, const VecN& x_original)
{
    std::lock_guard<std::recursive_mutex> eg(engine_mutex_);
    SyncEngine(min_t, max_t, x_original);
    if (x.size() != x_original.size()) throw std::invalid_argument("x has the wrong length");
    double f = 0.0;
    Check(omo_eval(engine_, x.data(), 1, &f, &last_terms_), "omo_eval");
    NoteEvaluated(x, min_t, max_t);
    return f;
}

/// What every evaluation leaves behind: the object sits at x (core_opt.cpp:40 "modifies the object and does not restore it",
/// core_rbdl.cpp:371) and the OptimizationCache belongs to x from now on.
void Core::NoteEvaluated(const VecN& x, double min_t, double max_t)
{
    object_->SetDescriptor(x);
    std::lock_guard<std::mutex> g(optimization_cache_mutex_);
    last_x_ = x; last_min_t_ = min_t; last_max_t_ = max_t; have_last_ = true; cache_stale_ = true;
}

void Core::NoteEvaluatedFromCache(const VecN& x, double min_t, double max_t)
{
    std::lock_guard<std::recursive_mutex> eg(engine_mutex_);
    NoteEvaluated(x, min_t, max_t);
}

void Core::CalculateFullObjectiveBatch(double min_t, double max_t, const double* X, int B, const VecN& x_original, double* f_out,
                                       omo_terms* terms_out)
{
    std::lock_guard<std::recursive_mutex> eg(engine_mutex_);
    Engine();
    if (multi_) {
        SyncMulti(min_t, max_t, x_original);
        if (omo_multi_eval(multi_, X, B, f_out, terms_out) != OMO_OK) throw engine_error(std::string("omo_multi_eval: ") + omo_multi_last_error(multi_));
        return;
    }
    SyncEngine(min_t, max_t, x_original);
    Check(omo_eval(engine_, X, B, f_out, terms_out), "omo_eval");
}

double Core::CalculateFullObjectiveAndGradient(double min_t, double max_t, const VecN& x, const VecN& x_original,
                                               const std::vector<int>& free_indices, double h, VecN* grad, VecN* f_plus_minus)
{
    std::lock_guard<std::recursive_mutex> eg(engine_mutex_);
    Engine();
    if (x.size() != x_original.size()) throw std::invalid_argument("x has the wrong length");
    const int F = static_cast<int>(free_indices.size());
    grad->assign(F, 0.0);
    if (f_plus_minus) f_plus_minus->assign(2 * static_cast<size_t>(F), 0.0);
    double f0 = 0.0;
    if (multi_) {
        SyncMulti(min_t, max_t, x_original);
        if (omo_multi_eval_grad(multi_, x.data(), free_indices.data(), F, h, &f0, grad->data(), f_plus_minus ? f_plus_minus->data() : nullptr) != OMO_OK)
            throw engine_error(std::string("omo_multi_eval_grad: ") + omo_multi_last_error(multi_));
    } else {
        SyncEngine(min_t, max_t, x_original);
        Check(omo_eval_grad(engine_, x.data(), free_indices.data(), F, h, &f0, grad->data(), f_plus_minus ? f_plus_minus->data() : nullptr),
              "omo_eval_grad");
    }
    NoteEvaluated(x, min_t, max_t);
    return f0;
}

double Core::CalculateFullObjectiveAndAnalyticGradient(double min_t, double max_t, const VecN& x, const VecN& x_original,
                                                       const std::vector<int>& free_indices, VecN* grad, bool* analytic_used)
{
    std::lock_guard<std::recursive_mutex> eg(engine_mutex_);
    SyncEngine(min_t, max_t, x_original);
    if (x.size() != x_original.size()) throw std::invalid_argument("x has the wrong length");
    const int F = static_cast<int>(free_indices.size());
    grad->assign(F, 0.0);
    double f0 = 0.0;
    const omo_status st = omo_eval_grad_analytic(engine_, x.data(), free_indices.data(), F, &f0, grad->data());
    if (analytic_used) *analytic_used = st == OMO_OK;
    if (st == OMO_UNSUPPORTED)   // IK handles / unsorted keys: the finite-difference gradient covers them
        return CalculateFullObjectiveAndGradient(min_t, max_t, x, x_original, free_indices, 1e-4, grad);
    Check(st, "omo_eval_grad_analytic");
    NoteEvaluated(x, min_t, max_t);
    return f0;
}

std::vector<double> Core::CalculatePositions(double min_t, double max_t, const VecN& x)
{
    std::lock_guard<std::recursive_mutex> eg(engine_mutex_);
    SyncEngine(min_t, max_t, x_original_);
    std::vector<double> p(static_cast<size_t>(omo_num_samples(engine_)) * omo_dof(engine_));
    Check(omo_positions(engine_, x.data(), p.data()), "omo_positions");
    return p;
}

// ---- next row N2: single-t torques (core_rbdl.cpp:264-289) and their UI-side consumers -----------------------------

void Core::CalculateTorquesBatch(const double* times, int M, double* tau, double* q, double* qd, double* qdd)
{
    std::lock_guard<std::recursive_mutex> eg(engine_mutex_);
    const VecN x = object_->GetDescriptor();
    SyncEngineState(x_original_.size() == x.size() ? x_original_ : x);
    Check(omo_eval_states_at(engine_, x.data(), times, M, 0.20, q, qd, qdd, tau, nullptr, nullptr), "omo_eval_states_at");
}

VecN Core::CalculateTorques(double t)
{
    VecN tau(3 * object_->GetJoints().size(), 0.0);
    CalculateTorquesBatch(&t, 1, tau.data());
    return tau;
}

std::vector<double> Core::CalculateEulerAngles(const double* times, int M)
{
    std::lock_guard<std::recursive_mutex> eg(engine_mutex_);
    const VecN x = object_->GetDescriptor();
    SyncEngineState(x_original_.size() == x.size() ? x_original_ : x);
    const size_t dof = 3 * object_->GetJoints().size();
    std::vector<double> e(static_cast<size_t>(M) * dof), root(static_cast<size_t>(M) * 3);
    Check(omo_eval_states_at(engine_, x.data(), times, M, 0.20, e.data(), nullptr, nullptr, nullptr, root.data(), nullptr), "omo_eval_states_at");
    // q[0:3] is the root OFFSET (core_rbdl.cpp:178-182); ExportMotion wants the root joint's own angles there
    for (int i = 0; i < M; ++i) for (int k = 0; k < 3; ++k) e[i * dof + k] = root[3 * static_cast<size_t>(i) + k];
    return e;
}

void Core::UpdateMaximumTorqueAndCost()
{
    const OptimizationCache cache = GetOptimizationCache();
    max_torque_ = 0.0; max_cost_ = 0.0;
    for (int i = 0; i < cache.GetSize(); ++i) {
        double n2 = 0.0;
        for (double v : cache.tau[i]) n2 += v * v;
        max_torque_ = std::max(max_torque_, std::sqrt(n2));
        max_cost_ = std::max(max_cost_, cache.local_cost_sequence[i]);
    }
}

Core::OptimizationCache Core::GetOptimizationCache()
{
    // lock order: engine, then cache (CalculateFullObjective holds the engine and takes the cache mutex at its end)
    std::lock_guard<std::recursive_mutex> eg(engine_mutex_);
    std::lock_guard<std::mutex> g(optimization_cache_mutex_);
    if (cache_stale_ || optimization_cache_.time_sequence.empty()) {
        // The reference computes CalculateCost(min_frame-1, max_frame+1, 0.2) of the current object when no cache
        // exists (core.h:136-139); otherwise the cache belongs to the last evaluated x.
        VecN x = have_last_ ? last_x_ : object_->GetDescriptor();
        const double lo = have_last_ ? last_min_t_ : min_frame_ - 1.0, hi = have_last_ ? last_max_t_ : max_frame_ + 1.0;
        SyncEngine(lo, hi, x_original_.size() == x.size() ? x_original_ : x);
        const int n = omo_num_samples(engine_), dof = omo_dof(engine_);
        std::vector<double> q((size_t)n * dof), qd((size_t)n * dof), qdd((size_t)n * dof), tau((size_t)n * dof), lc(n), ts(n);
        Check(omo_eval_cached(engine_, x.data(), nullptr, nullptr, q.data(), qd.data(), qdd.data(), tau.data(), lc.data()), "omo_eval_cached");
        Check(omo_get_time_grid(engine_, ts.data()), "omo_get_time_grid");
        OptimizationCache c;
        c.time_sequence = ts; c.local_cost_sequence = lc;
        auto rows = [&](const std::vector<double>& a) {
            std::vector<VecN> out(n);
            for (int i = 0; i < n; ++i) out[i].assign(a.begin() + (size_t)i * dof, a.begin() + (size_t)(i + 1) * dof);
            return out;
        };
        c.q = rows(q); c.q_dot = rows(qd); c.q_ddot = rows(qdd); c.tau = rows(tau);
        optimization_cache_ = c;
        cache_stale_ = false;
    }
    return optimization_cache_;
}

void Core::ClearOptimizationCache()
{
    std::lock_guard<std::mutex> g(optimization_cache_mutex_);
    optimization_cache_ = OptimizationCache();
    cache_stale_ = true;
}

// ---- the NLopt callback (core_opt.cpp:42-57) ------------------------------------------------------------------------------------

double EvaluateObjective(const std::vector<double>& x, std::vector<double>& grad, void* data)
{
    EvaluateObjectiveData* d = static_cast<EvaluateObjectiveData*>(data);
    Core& core = *d->core;
    // revert: X^T x + Y x_initial
    VecN full = d->x_initial;
    if (x.size() != d->free_indices.size()) throw std::invalid_argument("EvaluateObjective: wrong number of free dofs");
    for (size_t k = 0; k < x.size(); ++k) full[d->free_indices[k]] = x[k];
    if (core.stop_optimization_) {
        core.stop_optimization_ = false;
        throw forced_stop();
    }
    d->evaluations += 1;
    if (!d->start_cache_x.empty() && !(d->want_gradient && !grad.empty())) {
        // batched start: the solver's x, bit for bit, among the points evaluated before it began?
        for (size_t b = 0; b < d->start_cache_x.size(); ++b) {
            if (d->start_cache_x[b].size() == x.size() && std::memcmp(d->start_cache_x[b].data(), x.data(), x.size() * sizeof(double)) == 0) {
                d->start_cache_hits += 1;
                core.NoteEvaluatedFromCache(full, d->min_t, d->max_t);
                return d->start_cache_f[b];
            }
        }
    }
    if (d->want_gradient && !grad.empty()) {
        VecN g;
        const double f = d->analytic_gradient
                             ? core.CalculateFullObjectiveAndAnalyticGradient(d->min_t, d->max_t, full, d->x_original, d->free_indices, &g)
                             : core.CalculateFullObjectiveAndGradient(d->min_t, d->max_t, full, d->x_original, d->free_indices, d->gradient_step, &g);
        std::copy(g.begin(), g.end(), grad.begin());
        return f;
    }
    return core.CalculateFullObjective(d->min_t, d->max_t, full, d->x_original);
}

}  // namespace optimo
