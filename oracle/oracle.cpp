// =====================================================================================================
// TEST INFRASTRUCTURE -- CPU ORACLE. Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
// `--impl reference` legs may load this. The product (optimo_b200/, include/) never includes, links
// or calls anything in oracle/.
// =====================================================================================================
//
// CPU restatement of OptiMo's objective path, written without Eigen / Qt.
// Two builds share this file (oracle/Makefile):
//
//   liboracle.so            -- "port": everything restated in this repo (omath.hpp, own_rbdl.hpp and the
//                              curve code below). Travels as source, builds anywhere.
//   _ref/liboracle_ref.so   -- "reference-linked" (-DORACLE_USE_REF): the OptiMo glue below is the same
//                              restatement, but the numerics underneath are the reference's OWN objects,
//                              compiled from /root/reference where they lie: RBDL v2.5.0 (SimpleMath backend:
//                              InverseDynamics, InverseKinematics, jcalc, AddBody), OptiMo/curve.cpp +
//                              variable.cpp (verbatim, against oracle/shim/Eigen/Core), and parallel-util.
//                              OptiMo's own core*.cpp/joint.cpp/object.cpp/ikhandler.cpp cannot be compiled
//                              here (they need Eigen3 + Qt5/OpenGL headers), hence the restatement.
//
// PARITY PINNING: the reference ships no OptiMo-level golden vectors. This oracle
// is pinned (tests/test_oracle_pin.py) by (i) port == reference-linked build on all three shipped scenes,
// (ii) the RBDL/curve known answers recorded from real reference objects (SURVEY.md Appendix A.2/A.3/A.5,
// tests/golden/), (iii) RBDL's own test identities (EulerXYZ == chain of revolutes, MultiDofTests.cc:655-702).
// Eigen itself is absent, so the Eigen-vs-SimpleMath last-ulp differences are "parity unpinned".
//
// Each function cites the reference file:line it follows (paths relative to /root/reference/).

#include <algorithm>
#include <atomic>
#include <cassert>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <functional>
#include <list>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "omath.hpp"

#ifdef ORACLE_USE_REF
#include <rbdl/rbdl.h>         // the reference's RBDL v2.5.0 (SimpleMath backend)
#include <parallel-util.hpp>   // the reference's parallel-util
#include "variable.h"          // the reference's OptiMo/variable.h + curve.h (verbatim)
namespace RBD = RigidBodyDynamics;
#else
#include "own_rbdl.hpp"
#endif

namespace orc {

// ---------------------------------------------------------------------------------------------------
// a16  parallel_for          external/parallel-util/include/parallel-util.hpp:28-63
// ---------------------------------------------------------------------------------------------------
#ifdef ORACLE_USE_REF
template <typename F> void parallel_for(int n, F f, int conc = 0) { parallelutil::parallel_for(n, f, conc); }
#else
/// Static block partition over freshly spawned std::threads; the first threads take
/// ceil(n/T) tasks, the last `lacking` threads one fewer. conc == 0 => hardware_concurrency (or 4).
template <typename F> void parallel_for(int n, F f, int conc = 0)
{
    if (n <= 0) return;
    const int hint = (conc == 0) ? static_cast<int>(std::thread::hardware_concurrency()) : conc;
    const int T = std::min(n, hint == 0 ? 4 : hint);
    const int per = n / T + (n % T == 0 ? 0 : 1);
    const int lacking = per * T - n;
    auto work = [&](int ti) {
        const int short_before = std::max(ti - T + lacking, 0);
        const int begin = ti * per - short_before;
        const int end = begin + per - (ti - T + lacking >= 0 ? 1 : 0);
        for (int k = begin; k < end; ++k) f(k);
    };
    std::vector<std::thread> th;
    for (int j = 0; j < T; ++j) th.emplace_back(work, j);
    for (auto& t : th) t.join();
}
#endif

// ---------------------------------------------------------------------------------------------------
// a6  curves                 OptiMo/curve.h:21-76, curve.cpp:20-54,57-93,100-148,179-188, variable.h:24
// ---------------------------------------------------------------------------------------------------
#ifdef ORACLE_USE_REF
using ::ControlPoint;
using ::Curve;
using ::Variable;
static std::atomic<long> g_newton_iters{0}, g_bezier_evals{0};  // not counted in this build
#else
static std::atomic<long> g_newton_iters{0}, g_bezier_evals{0};

struct ControlPoint {
    double time, value, theta, time_backward, time_forward;
    ControlPoint(double t, double v, double th, double tb, double tf)
        : time(t), value(v), theta(th), time_backward(tb), time_forward(tf) {}
    double GetValueBackward() const { return time_backward * std::tan(theta); }  // curve.h:29-32
    double GetValueForward() const { return time_forward * std::tan(theta); }    // curve.h:34-37
    double GetT() const { return time; }
    double GetV() const { return value; }
};

class Curve {
public:
    struct Segment { int start_point_index, end_point_index; };
    Curve() { chain_.push_back(Segment{-1, -1}); }  // curve.cpp:95-98

    std::vector<ControlPoint>& GetControlPoints() { return pts_; }
    const std::vector<ControlPoint>& GetControlPoints() const { return pts_; }
    ControlPoint& GetControlPoint(int i) { return pts_[i]; }
    const std::vector<Segment>& GetCurveSegments() const { return chain_; }

    /// curve.cpp:100-128 -- first segment, in list order, that accepts t.
    size_t FindSegmentIndex(double t) const
    {
        for (size_t k = 0; k < chain_.size(); ++k) {
            const Segment& s = chain_[k];
            if (s.end_point_index < 0) return k;
            const double te = pts_[s.end_point_index].GetT();
            if (s.start_point_index < 0) { if (te > t) return k; continue; }
            if (pts_[s.start_point_index].GetT() <= t && te > t) return k;
        }
        return chain_.size() - 1;
    }
    /// curve.cpp:130-148 -- append the point, split the segment that contains its time.
    void AddControlPoint(const ControlPoint& cp)
    {
        pts_.push_back(cp);
        const size_t k = FindSegmentIndex(cp.GetT());
        const Segment old = chain_[k];
        const int idx = static_cast<int>(pts_.size()) - 1;
        chain_[k] = Segment{idx, old.end_point_index};
        chain_.insert(chain_.begin() + k, Segment{old.start_point_index, idx});
    }
    /// curve.cpp:150-154 + 57-93
    double GetValue(double t) const
    {
        const Segment& s = chain_[FindSegmentIndex(t)];
        if (s.start_point_index < 0 && s.end_point_index < 0) return 0.0;
        if (s.start_point_index < 0) return pts_[s.end_point_index].GetV();
        if (s.end_point_index < 0) return pts_[s.start_point_index].GetV();
        const ControlPoint& sp = pts_[s.start_point_index];
        const ControlPoint& ep = pts_[s.end_point_index];
        const double p0 = sp.GetT(), p1 = sp.GetT() + sp.time_forward, p2 = ep.GetT() + ep.time_backward,
                     p3 = ep.GetT();
        const double a = -p0 + 3.0 * p1 - 3.0 * p2 + p3;
        const double b = 3.0 * (p0 - 2.0 * p1 + p2);
        const double c = -3.0 * p0 + 3.0 * p1;
        const double d = p0 - t;
        const double u = SolveCubic(a, b, c, d);
        return Bezier(u, sp.GetV(), sp.GetV() + sp.GetValueForward(), ep.GetV() + ep.GetValueBackward(), ep.GetV());
    }
    /// curve.cpp:179-188
    void MakeMotionCyclic()
    {
        const int n = static_cast<int>(pts_.size());
        const double th = 0.5 * (pts_[0].theta + pts_[n - 1].theta);
        pts_[0].theta = th;
        pts_[n - 1].theta = th;
    }

private:
    /// curve.cpp:43-54 -- Newton from 0.5, at most 50 updates, |f| tested before each update.
    static double SolveCubic(double a, double b, double c, double d)
    {
        double x = 0.5;
        int k = 0;
        for (; k < 50; ++k) {
            const double f = x * (x * (a * x + b) + c) + d;
            const double fp = x * (3.0 * a * x + 2.0 * b) + c;
            if (std::abs(f) < 1e-09) break;
            x = x - f / fp;
        }
        g_newton_iters += k;
        g_bezier_evals += 1;
        return x;
    }
    /// curve.cpp:20-40 -- Bernstein form, each basis polynomial evaluated on its own.
    static double Bezier(double u, double p0, double p1, double p2, double p3)
    {
        const double w = 1.0 - u;
        return p0 * (w * w * w) + p1 * (3.0 * w * w * u) + p2 * (3.0 * w * u * u) + p3 * (u * u * u);
    }
    std::vector<ControlPoint> pts_;   // insertion order
    std::vector<Segment> chain_;      // time order at insertion
};

struct Variable {
    explicit Variable(double v = 0.0) : var_(v) {}
    double var_;
    Curve curve_;
    bool IsKeyframed() const { return !curve_.GetControlPoints().empty(); }
    double GetValue(double t) const { return IsKeyframed() ? curve_.GetValue(t) : var_; }
    int GetNumOfControlPoints() const { return static_cast<int>(curve_.GetControlPoints().size()); }
};
#endif

// ---------------------------------------------------------------------------------------------------
// Backend wrappers over RBDL (real objects or own_rbdl.hpp)
// ---------------------------------------------------------------------------------------------------
enum JointKind { kTranslationXYZ = 0, kEulerXYZ = 1, kEulerZYX = 2 };

#ifdef ORACLE_USE_REF
struct RbdModel {
    RBD::Model m;
    static RBD::Math::Vector3d v3(const V3& a) { return RBD::Math::Vector3d(a[0], a[1], a[2]); }
    unsigned add_body_gyration(unsigned parent, const V3& xtrans, int kind, double mass, const V3& com, const V3& gyr)
    {
        return m.AddBody(parent, RBD::Math::Xtrans(v3(xtrans)), RBD::Joint(jt(kind)), RBD::Body(mass, v3(com), v3(gyr)));
    }
    unsigned add_body(unsigned parent, const V3& xtrans, int kind, double mass, const V3& com, const M3& inertia)
    {
        RBD::Math::Matrix3d I(inertia(0, 0), inertia(0, 1), inertia(0, 2), inertia(1, 0), inertia(1, 1), inertia(1, 2),
                              inertia(2, 0), inertia(2, 1), inertia(2, 2));
        return m.AddBody(parent, RBD::Math::Xtrans(v3(xtrans)), RBD::Joint(jt(kind)), RBD::Body(mass, v3(com), I));
    }
    void set_gravity(const V3& g) { m.gravity = v3(g); }
    unsigned dof_count() const { return m.dof_count; }
    void inverse_dynamics(const VecN& q, const VecN& qd, const VecN& qdd, VecN& tau)
    {
        const int n = static_cast<int>(q.size());
        RBD::Math::VectorNd Q(n), QD(n), QDD(n), T(n);
        for (int i = 0; i < n; ++i) { Q[i] = q[i]; QD[i] = qd[i]; QDD[i] = qdd[i]; T[i] = 0.0; }
        RBD::InverseDynamics(m, Q, QD, QDD, T);
        for (int i = 0; i < n; ++i) tau[i] = T[i];
    }
    bool inverse_kinematics(const VecN& qinit, unsigned body, const V3& point, const V3& target, VecN& qres, double tol,
                            double lambda, unsigned max_iter, int* iters)
    {
        const int n = static_cast<int>(qinit.size());
        RBD::Math::VectorNd Q0(n), QR(n);
        for (int i = 0; i < n; ++i) { Q0[i] = qinit[i]; QR[i] = 0.0; }
        const bool ok = RBD::InverseKinematics(m, Q0, std::vector<unsigned>{body}, std::vector<RBD::Math::Vector3d>{v3(point)},
                                               std::vector<RBD::Math::Vector3d>{v3(target)}, QR, tol, lambda, max_iter);
        qres.resize(n);
        for (int i = 0; i < n; ++i) qres[i] = QR[i];
        if (iters) *iters = -1;
        return ok;
    }
private:
    static RBD::JointType jt(int kind)
    {
        return kind == kTranslationXYZ ? RBD::JointTypeTranslationXYZ
                                       : (kind == kEulerXYZ ? RBD::JointTypeEulerXYZ : RBD::JointTypeEulerZYX);
    }
};
#else
struct RbdModel {
    own::Model m;
    unsigned add_body_gyration(unsigned parent, const V3& xtrans, int kind, double mass, const V3& com, const V3& gyr)
    {
        M3 I; I(0, 0) = gyr[0]; I(1, 1) = gyr[1]; I(2, 2) = gyr[2];  // Body.h:61-72
        return add_body(parent, xtrans, kind, mass, com, I);
    }
    unsigned add_body(unsigned parent, const V3& xtrans, int kind, double mass, const V3& com, const M3& inertia)
    {
        return m.add_body(parent, own::Xform(M3::identity(), xtrans), kind, mass, com, inertia);
    }
    void set_gravity(const V3& g) { m.gravity = g; }
    unsigned dof_count() const { return m.dof_count; }
    void inverse_dynamics(const VecN& q, const VecN& qd, const VecN& qdd, VecN& tau) { m.inverse_dynamics(q, qd, qdd, tau); }
    bool inverse_kinematics(const VecN& qinit, unsigned body, const V3& point, const V3& target, VecN& qres, double tol,
                            double lambda, unsigned max_iter, int* iters)
    {
        return m.inverse_kinematics(qinit, body, point, target, qres, tol, lambda, max_iter, iters);
    }
};
#endif

static std::atomic<long> g_ik_iters{0}, g_ik_solves{0};

// ---------------------------------------------------------------------------------------------------
// a7 / a10  Joint             OptiMo/joint.h, joint.cpp:36-52,64-83,90-119
// a8        IkHandler         OptiMo/ikhandler.cpp:77-188
// a4        Object            OptiMo/object.cpp:27-61,110-186
// ---------------------------------------------------------------------------------------------------
struct IkHandler;

struct Item {
    std::string name;
    bool is_selected = false;
    virtual ~Item() {}
    virtual std::vector<Variable*> vars() = 0;
};

struct Joint : Item {
    Joint* parent = nullptr;
    std::vector<Joint*> children;
    V3 offset;
    Variable rot[3];
    IkHandler* ik = nullptr;   // associated_ik_handler_
    double mass = 0.0; V3 com; M3 inertia;
    bool degenerate_inertia = false;

    std::vector<Variable*> vars() override { return {&rot[0], &rot[1], &rot[2]}; }

    /// joint.cpp:90-119 -- box of 0.1 x |offset| x 0.1 m, density 1000, COM at half the offset,
    /// inertia rotated by R = FromTwoVectors(offset^, y^).
    void SetPhysicalParameters()
    {
        const double x_body = 0.1, y_body = norm(offset), z_body = 0.1;
        if (!parent) { mass = 0.0; com = V3(); inertia = M3(); return; }
        mass = 1000.0 * x_body * z_body * norm(offset);
        com = 0.5 * offset;
        const M3 R = from_two_vectors(normalized(offset), V3(0.0, 1.0, 0.0), &degenerate_inertia);
        M3 Ib;
        Ib(0, 0) = (mass / 12.0) * (y_body * y_body + z_body * z_body);
        Ib(1, 1) = (mass / 12.0) * (x_body * x_body + z_body * z_body);
        Ib(2, 2) = (mass / 12.0) * (x_body * x_body + y_body * y_body);
        inertia = (R * Ib) * transpose(R);
    }

    V3 GetEulerAngle(double t);            // joint.cpp:36-52 (defined after IkHandler)
    Affine GetAffineRelativeToParent(double t)   // joint.cpp:64-74: T(offset) Rz Ry Rx
    {
        const V3 e = GetEulerAngle(t);
        Affine a;
        a.translate(offset);
        a.rotate(angle_axis_unit(e[2], 2));
        a.rotate(angle_axis_unit(e[1], 1));
        a.rotate(angle_axis_unit(e[0], 0));
        return a;
    }
    Affine GetAffineRelativeToWorld(double t)    // joint.cpp:76-83 (recursion to the root)
    {
        if (!parent) return GetAffineRelativeToParent(t);
        return parent->GetAffineRelativeToWorld(t) * GetAffineRelativeToParent(t);
    }
};

struct IkHandler : Item {
    Variable pos[3];
    Joint* root = nullptr;
    Joint* end = nullptr;
    std::vector<Joint*> targets;   // root .. end effector
    RbdModel model;
    unsigned end_body_id = 0;
    // 1-entry cache shared by all threads (ikhandler.cpp:107-122,153-161)
    std::vector<V3> cached_result;
    double cached_parameter = 0.0;
    std::mutex* mutex = nullptr;   // the reference uses one function-static mutex for all handlers

    std::vector<Variable*> vars() override { return {&pos[0], &pos[1], &pos[2]}; }

    /// ikhandler.cpp:77-101 + 45-67
    void Initialize(Joint* ik_root, Joint* end_effector)
    {
        root = ik_root; end = end_effector;
        targets.clear();
        for (Joint* j = end_effector;; j = j->parent) {
            targets.insert(targets.begin(), j);
            if (j == ik_root) break;
            assert(j->parent);
        }
        for (Joint* j : targets) {
            if (j == end) continue;
            j->ik = this;
            for (auto& r : j->rot) r.curve_ = Curve();   // reset FK keys
        }
        // ikhandler.cpp:166-188 -- chain of EulerZYX bodies, body i+1 framed at targets[i].offset
        model = RbdModel();
        unsigned parent_id = 0;
        for (size_t i = 0; i + 1 < targets.size(); ++i) {
            Joint* j = targets[i + 1];
            parent_id = model.add_body(parent_id, j->parent->offset, kEulerZYX, j->mass, j->com, j->inertia);
        }
        end_body_id = parent_id;
    }

    int GetIndex(const Joint* j) const
    {
        for (size_t i = 0; i < targets.size(); ++i) if (targets[i] == j) return static_cast<int>(i);
        assert(false); return 0;
    }

    /// ikhandler.cpp:103-164
    std::vector<V3> ComputeInverseKinematics(double t)
    {
        const size_t n_ctl = model.dof_count() / 3;
        {
            std::lock_guard<std::mutex> g(*mutex);
            const std::vector<V3> copy = cached_result;
            const bool hit = (cached_result.size() == n_ctl) && (std::abs(t - cached_parameter) < 1e-12);
            if (hit) return copy;
        }
        const Affine affine = root->parent ? root->parent->GetAffineRelativeToWorld(t) : Affine();
        const V3 target_world(pos[0].GetValue(t), pos[1].GetValue(t), pos[2].GetValue(t));
        const V3 target_local = inverse(affine) * target_world;

        const VecN q_init(model.dof_count(), 0.0);
        VecN q_res(model.dof_count(), 0.0);
        int iters = 0;
        RbdModel copied = model;   // THREAD_SAFE copy (ikhandler.cpp:139-141)
        copied.inverse_kinematics(q_init, end_body_id, end->offset, target_local, q_res, 1e-06, 0.40, 20, &iters);
        if (iters > 0) { g_ik_iters += iters; g_ik_solves += 1; }

        std::vector<V3> result(n_ctl);
        for (size_t i = 0; i < n_ctl; ++i) result[i] = V3(q_res[3 * i + 2], q_res[3 * i + 1], q_res[3 * i]);  // SwapXZ
        {
            std::lock_guard<std::mutex> g(*mutex);
            cached_result = result;
            cached_parameter = t;
        }
        return result;
    }
};

inline V3 Joint::GetEulerAngle(double t)
{
    if (ik) return ik->ComputeInverseKinematics(t)[ik->GetIndex(this)];
    return V3(rot[0].GetValue(t), rot[1].GetValue(t), rot[2].GetValue(t));
}

struct GaussianKernel {   // core.h:27-33
    double amp, mu, sigma;
    double GetValue(double x) const { return amp * std::exp(-(x - mu) * (x - mu) / (2.0 * sigma * sigma)); }
};

// ---------------------------------------------------------------------------------------------------
// Scene == headless Core + Object
// ---------------------------------------------------------------------------------------------------
struct BuildOp {   // recorded so that a scene can be cloned for candidate-parallel runs
    int kind; int a, b; double d[5]; std::string s;
};

struct Scene {
    std::vector<std::unique_ptr<Joint>> joint_store;   // creation order == DFS pre-order of the JSON
    std::vector<std::unique_ptr<IkHandler>> ik_store;
    Joint* root = nullptr;
    bool is_cyclic = false;
    std::mutex ik_mutex;
    std::vector<BuildOp> ops;

    int min_frame = 1, max_frame = 40;
    std::vector<char> freezed_dofs;
    std::vector<double> per_item_weights;
    double weight_trajectory = 0.0;
    std::vector<GaussianKernel> kernels;
    VecN x_original;
    std::unique_ptr<RbdModel> model;
    int sampling_concurrency = 0;   // 0 => hardware_concurrency (MULTI_THREAD, core_rbdl.cpp:157-161)

    struct OriginalMotionCache { std::vector<VecN> p_original; VecN x_original; };
    std::unique_ptr<OriginalMotionCache> original_motion_cache;

    struct OptimizationCache { std::vector<double> time_sequence; std::vector<VecN> q, qd, qdd, tau; std::vector<double> local_cost; };
    OptimizationCache optimization_cache;
    double last_terms[3] = {0.0, 0.0, 0.0};

    // ---- Object ----------------------------------------------------------------------------------
    std::vector<Joint*> GetJoints() const   // object.cpp:27-40 (DFS pre-order)
    {
        std::vector<Joint*> out;
        std::function<void(Joint*)> rec = [&](Joint* j) { out.push_back(j); for (Joint* c : j->children) rec(c); };
        if (root) rec(root);
        return out;
    }
    std::vector<Item*> GetItems() const     // object.cpp:47-61 (joints, then IK handlers)
    {
        std::vector<Item*> out;
        for (Joint* j : GetJoints()) out.push_back(j);
        for (auto& h : ik_store) out.push_back(h.get());
        return out;
    }
    VecN GetDescriptor() const              // object.cpp:110-145
    {
        VecN x;
        for (Item* it : GetItems())
            for (Variable* v : it->vars()) {
                if (!v->IsKeyframed()) continue;
                for (const ControlPoint& p : v->curve_.GetControlPoints()) {
                    x.push_back(p.time); x.push_back(p.value); x.push_back(p.theta);
                    x.push_back(p.time_backward); x.push_back(p.time_forward);
                }
            }
        return x;
    }
    void SetDescriptor(const VecN& x)       // object.cpp:147-186
    {
        size_t k = 0;
        for (Item* it : GetItems())
            for (Variable* v : it->vars()) {
                if (!v->IsKeyframed()) continue;
                for (ControlPoint& p : v->curve_.GetControlPoints()) {
                    p.time = x[k]; p.value = x[k + 1]; p.theta = x[k + 2];
                    p.time_backward = x[k + 3]; p.time_forward = x[k + 4];
                    k += 5;
                }
            }
        if (is_cyclic)
            for (Item* it : GetItems())
                for (Variable* v : it->vars()) if (v->IsKeyframed()) v->curve_.MakeMotionCyclic();
    }

    // ---- a10: Core::InitializeDynamicsModel  core_rbdl.cpp:200-262 ------------------------------------
    void InitializeDynamicsModel()
    {
        const std::vector<Joint*> joints = GetJoints();
        model.reset(new RbdModel());
        std::vector<unsigned> id_of(joints.size(), 0);
        id_of[0] = model->add_body_gyration(0, root->offset, kTranslationXYZ, 0.0, V3(), V3(1.0, 1.0, 1.0));
        for (size_t k = 1; k < joints.size(); ++k) {
            Joint* j = joints[k];
            const size_t pk = std::find(joints.begin(), joints.end(), j->parent) - joints.begin();
            id_of[k] = model->add_body(id_of[pk], j->parent->offset, kEulerXYZ, j->mass, j->com, j->inertia);
        }
        model->set_gravity(V3(0.0 * 0.0, 0.0 * -9.8, 0.0 * 0.0));
    }

    // ---- Core() defaults for weights / frozen DoFs   core.cpp:80-94 -------------------------------
    void ResetRuntimeDefaults()
    {
        const int n = static_cast<int>(GetDescriptor().size()) / 5;
        freezed_dofs.assign(n * 5, 0);
        for (int i = 0; i < n; ++i) { freezed_dofs[i * 5 + 0] = 1; freezed_dofs[i * 5 + 1] = 1; }
        per_item_weights.assign(GetItems().size(), 0.0);
        weight_trajectory = 0.0;
        kernels.clear();
        UpdateOriginalParameters(GetDescriptor());
    }
    void UpdateOriginalParameters(const VecN& x) { x_original = x; original_motion_cache.reset(); }  // core.cpp:184-189

    /// core.cpp:118-148
    void UpdateFreezedDofsFromSelection()
    {
        const int n = static_cast<int>(GetDescriptor().size()) / 5;
        freezed_dofs.assign(n * 5, 1);
        int cp = 0;
        for (Item* it : GetItems()) {
            const int begin = cp;
            for (Variable* v : it->vars()) cp += v->GetNumOfControlPoints();
            if (it->is_selected)
                for (int i = begin; i < cp; ++i) { freezed_dofs[i * 5 + 2] = 0; freezed_dofs[i * 5 + 3] = 0; freezed_dofs[i * 5 + 4] = 0; }
        }
    }

    /// a13: core.cpp:161-182
    VecN CalculateParameterWeights() const
    {
        const int n = static_cast<int>(GetDescriptor().size()) / 5;
        VecN w(5 * n, weight_trajectory);
        int item_index = 0, cp = 0;
        for (Item* it : GetItems()) {
            for (Variable* v : it->vars()) {
                const int np = v->GetNumOfControlPoints();
                for (int k = cp * 5; k < (cp + np) * 5; ++k) w[k] += per_item_weights[item_index];
                cp += np;
            }
            ++item_index;
        }
        return w;
    }

    /// a11: core.cpp:191-199
    double GetCostWeight(double t) const
    {
        double w = 1.0;
        for (const GaussianKernel& k : kernels) w -= k.GetValue(t);
        return std::max(w, 0.0);
    }

    // ---- a9: stencils   core_rbdl.cpp:28-95 (SECOND_ORDER_ACCURATE off) ----------------------------
    static void FirstCentral(double dx, const std::vector<VecN>& v, std::vector<VecN>* r)
    {
        const int n = static_cast<int>(v.size());
        const size_t m = v[0].size();
        for (int i = 2; i < n - 2; ++i)
            for (size_t k = 0; k < m; ++k)
                (*r)[i][k] = (-v[i + 2][k] + 8.0 * v[i + 1][k] - 8.0 * v[i - 1][k] + v[i - 2][k]) / (12.0 * dx);
        for (int i : {1, n - 2})
            for (size_t k = 0; k < m; ++k) (*r)[i][k] = (v[i + 1][k] - v[i - 1][k]) / (2.0 * dx);
        for (size_t k = 0; k < m; ++k) {
            (*r)[0][k] = (v[1][k] - v[0][k]) / dx;
            (*r)[n - 1][k] = (v[n - 1][k] - v[n - 2][k]) / dx;
        }
    }
    static void SecondCentral(double dx, const std::vector<VecN>& v, std::vector<VecN>* r)
    {
        const int n = static_cast<int>(v.size());
        const size_t m = v[0].size();
        for (int i = 2; i < n - 2; ++i)
            for (size_t k = 0; k < m; ++k)
                (*r)[i][k] = (-v[i + 2][k] + 16.0 * v[i + 1][k] - 30.0 * v[i][k] + 16.0 * v[i - 1][k] - v[i - 2][k]) / (12.0 * dx * dx);
        for (int i : {1, n - 2})
            for (size_t k = 0; k < m; ++k) (*r)[i][k] = (v[i + 1][k] - 2.0 * v[i][k] + v[i - 1][k]) / (dx * dx);
        for (size_t k = 0; k < m; ++k) {
            (*r)[0][k] = (v[2][k] - 2.0 * v[1][k] + v[0][k]) / (2.0 * dx * dx);
            (*r)[n - 1][k] = (v[n - 1][k] - 2.0 * v[n - 2][k] + v[n - 3][k]) / (2.0 * dx * dx);
        }
    }

    // ---- a5: Core::SetStates   core_rbdl.cpp:108-166 ------------------------------------------------
    void SetStates(double min_t, double max_t, double dt, int n_dof, std::vector<double>* ts, std::vector<VecN>* q,
                   std::vector<VecN>* qd, std::vector<VecN>* qdd, std::vector<VecN>* p)
    {
        const std::vector<Joint*> joints = GetJoints();
        ts->clear();
        for (double t = min_t; t <= max_t; t += dt) ts->push_back(t);
        const int n_samples = static_cast<int>(ts->size());
        q->assign(n_samples, VecN(n_dof, 0.0));
        qd->assign(n_samples, VecN(n_dof, 0.0));
        qdd->assign(n_samples, VecN(n_dof, 0.0));
        if (p) p->assign(n_samples, VecN(joints.size() * 3, 0.0));

        auto body = [&](int ti) {
            int ji = 0;
            const double t = (*ts)[ti];
            for (Joint* j : joints) {
                if (!j->parent) {
                    for (int k = 0; k < 3; ++k) (*q)[ti][k] = j->offset[k];
                } else {
                    const V3 e = j->GetEulerAngle(t);
                    for (int k = 0; k < 3; ++k) (*q)[ti][ji * 3 + k] = e[k];
                }
                if (p) {
                    const V3 pos = j->GetAffineRelativeToWorld(t).t;
                    for (int k = 0; k < 3; ++k) (*p)[ti][3 * ji + k] = pos[k];
                }
                ++ji;
            }
        };
        parallel_for(n_samples, body, sampling_concurrency);
        FirstCentral(dt, *q, qd);
        SecondCentral(dt, *q, qdd);
    }

    // ---- a11: Core::CalculateCost   core_rbdl.cpp:307-346 -------------------------------------------
    double CalculateCost(double dt, const std::vector<double>& ts, const std::vector<VecN>& q, const std::vector<VecN>& qd,
                         const std::vector<VecN>& qdd, bool keep_cache)
    {
        const int n = static_cast<int>(q.size());
        const int dof = static_cast<int>(model->dof_count());
        double cost = 0.0;
        std::vector<VecN> tau(n, VecN(dof, 0.0));
        std::vector<double> local_cost(n), local_weight(n);
        for (int i = 0; i < n; ++i) local_weight[i] = GetCostWeight(ts[i]);
        for (int i = 0; i < n; ++i) {
            model->inverse_dynamics(q[i], qd[i], qdd[i], tau[i]);
            double s = 0.0;
            for (int k = 3; k < dof; ++k) s += tau[i][k] * tau[i][k];
            local_cost[i] = s * dt;
            cost += local_weight[i] * local_cost[i];
        }
        if (keep_cache) optimization_cache = OptimizationCache{ts, q, qd, qdd, tau, local_cost};
        return cost;
    }

    // ---- a3: Core::CalculateFullObjective   core_rbdl.cpp:348-401 -----------------------------------
    double CalculateFullObjective(double min_t, double max_t, const VecN& x, const VecN& x_orig, bool keep_cache = true)
    {
        const double dt = 0.20;
        const VecN w = CalculateParameterWeights();
        const int dof = static_cast<int>(model->dof_count());
        if (!original_motion_cache || !is_approx(original_motion_cache->x_original, x_orig)) {
            SetDescriptor(x_orig);
            std::vector<double> ts; std::vector<VecN> q, qd, qdd, p;
            SetStates(min_t, max_t, dt, dof, &ts, &q, &qd, &qdd, &p);
            original_motion_cache.reset(new OriginalMotionCache{p, x_orig});
        }
        SetDescriptor(x);
        std::vector<double> ts; std::vector<VecN> q, qd, qdd, p;
        SetStates(min_t, max_t, dt, dof, &ts, &q, &qd, &qdd, &p);
        const double C = CalculateCost(dt, ts, q, qd, qdd, keep_cache);
        double D_param = 0.0;
        for (size_t k = 0; k < x.size(); ++k) D_param += ((x[k] - x_orig[k]) * w[k]) * (x[k] - x_orig[k]);
        double D_traj = 0.0;
        for (size_t i = 0; i < p.size(); ++i) D_traj += weight_trajectory * sqdist(p[i], original_motion_cache->p_original[i]) * dt;
        last_terms[0] = C; last_terms[1] = D_param; last_terms[2] = D_traj;
        return C + D_param + D_traj;
    }

    // ---- N2: Core::SetCurrentState + Core::CalculateTorques   core_rbdl.cpp:168-198, 264-289 -------------------
    // Scalar differences of core_rbdl.cpp:97-106: (succ - prev) / (2 eps), (succ - 2 v + prev) / (eps eps).
    void SetCurrentState(double t, double dt, VecN* q, VecN* qd, VecN* qdd)
    {
        int i = 0;
        for (Joint* j : GetJoints()) {
            if (!j->parent) {
                for (int k = 0; k < 3; ++k) (*q)[k] = j->offset[k];
            } else {
                const V3 e = j->GetEulerAngle(t), ep = j->GetEulerAngle(t - dt), es = j->GetEulerAngle(t + dt);
                for (int k = 0; k < 3; ++k) {
                    (*q)[i * 3 + k] = e[k];
                    (*qd)[i * 3 + k] = (es[k] - ep[k]) / (2.0 * dt);
                    (*qdd)[i * 3 + k] = (es[k] - 2.0 * e[k] + ep[k]) / (dt * dt);
                }
            }
            ++i;
        }
    }
    VecN CalculateTorques(double t, VecN* q_out = nullptr, VecN* qd_out = nullptr, VecN* qdd_out = nullptr)
    {
        const double dt = 0.20;
        const int dof = static_cast<int>(model->dof_count());
        VecN q(dof, 0.0), qd(dof, 0.0), qdd(dof, 0.0), tau(dof, 0.0);
        SetCurrentState(t, dt, &q, &qd, &qdd);
        model->inverse_dynamics(q, qd, qdd, tau);
        if (q_out) *q_out = q;
        if (qd_out) *qd_out = qd;
        if (qdd_out) *qdd_out = qdd;
        return tau;
    }

    // ---- a2: select / revert   core_opt.cpp:99-121 ---------------------------------------------------
    std::vector<int> FreeIndices() const
    {
        std::vector<int> idx;
        for (size_t i = 0; i < freezed_dofs.size(); ++i) if (!freezed_dofs[i]) idx.push_back(static_cast<int>(i));
        return idx;
    }
};

static Scene* clone_scene(const Scene& s);

}  // namespace orc

// =====================================================================================================
// C API (ctypes)
// =====================================================================================================
using namespace orc;

extern "C" {

const char* orc_backend()
{
#ifdef ORACLE_USE_REF
    return "reference-linked: RBDL v2.5.0 (SimpleMath) + OptiMo/curve.cpp + parallel-util compiled from /root/reference";
#else
    return "port: restated in oracle/ (own_rbdl.hpp, omath.hpp, oracle.cpp)";
#endif
}

void* orc_scene_new() { return new Scene(); }
void orc_scene_free(void* h) { delete static_cast<Scene*>(h); }

static void record(Scene* s, int kind, int a, int b, const double* d, int nd, const char* str)
{
    BuildOp op; op.kind = kind; op.a = a; op.b = b;
    for (int i = 0; i < 5; ++i) op.d[i] = (i < nd) ? d[i] : 0.0;
    op.s = str ? str : "";
    s->ops.push_back(op);
}

/// core_io.cpp:112-130 -- joints must be added in the JSON's recursive (pre-order) sequence.
int orc_add_joint(void* h, int parent, const double* offset, const double* var3, const char* name)
{
    Scene* s = static_cast<Scene*>(h);
    const double d[5] = {offset[0], offset[1], offset[2], 0, 0};
    record(s, 0, parent, 0, d, 3, name);
    record(s, 5, 0, 0, var3, 3, nullptr);
    std::unique_ptr<Joint> j(new Joint());
    j->name = name;
    j->offset = V3(offset[0], offset[1], offset[2]);
    for (int k = 0; k < 3; ++k) j->rot[k] = Variable(var3[k]);
    if (parent >= 0) { j->parent = s->joint_store[parent].get(); j->parent->children.push_back(j.get()); }
    else s->root = j.get();
    j->SetPhysicalParameters();
    s->joint_store.push_back(std::move(j));
    return static_cast<int>(s->joint_store.size()) - 1;
}

/// core_io.cpp:100-110 (InterpretJson2Variable -> Curve::AddControlPoint)
void orc_joint_add_cp(void* h, int joint, int axis, double t, double v, double th, double tb, double tf)
{
    Scene* s = static_cast<Scene*>(h);
    const double d[5] = {t, v, th, tb, tf};
    record(s, 1, joint, axis, d, 5, nullptr);
    s->joint_store[joint]->rot[axis].curve_.AddControlPoint(ControlPoint(t, v, th, tb, tf));
}

/// core_io.cpp:132-151
int orc_add_ik(void* h, const char* name, int root_joint, int end_joint, const double* var3)
{
    Scene* s = static_cast<Scene*>(h);
    record(s, 2, root_joint, end_joint, var3, 3, name);
    std::unique_ptr<IkHandler> ik(new IkHandler());
    ik->name = name;
    ik->mutex = &s->ik_mutex;
    ik->Initialize(s->joint_store[root_joint].get(), s->joint_store[end_joint].get());
    for (int k = 0; k < 3; ++k) ik->pos[k] = Variable(var3[k]);
    s->ik_store.push_back(std::move(ik));
    return static_cast<int>(s->ik_store.size()) - 1;
}

void orc_ik_add_cp(void* h, int ik, int axis, double t, double v, double th, double tb, double tf)
{
    Scene* s = static_cast<Scene*>(h);
    const double d[5] = {t, v, th, tb, tf};
    record(s, 3, ik, axis, d, 5, nullptr);
    s->ik_store[ik]->pos[axis].curve_.AddControlPoint(ControlPoint(t, v, th, tb, tf));
}

/// Core::ReadJson tail (core_io.cpp:154-163) + the Core() defaults (core.cpp:80-94).
void orc_finalize(void* h, int min_frame, int max_frame)
{
    Scene* s = static_cast<Scene*>(h);
    const double d[5] = {0, 0, 0, 0, 0};
    record(s, 4, min_frame, max_frame, d, 0, nullptr);
    s->min_frame = min_frame; s->max_frame = max_frame;
    s->InitializeDynamicsModel();
    s->ResetRuntimeDefaults();
}

int orc_n_joints(void* h) { return static_cast<int>(static_cast<Scene*>(h)->joint_store.size()); }
int orc_n_items(void* h) { return static_cast<int>(static_cast<Scene*>(h)->GetItems().size()); }
int orc_descriptor_size(void* h) { return static_cast<int>(static_cast<Scene*>(h)->GetDescriptor().size()); }
void orc_get_descriptor(void* h, double* x)
{
    const VecN d = static_cast<Scene*>(h)->GetDescriptor();
    std::copy(d.begin(), d.end(), x);
}
void orc_set_descriptor(void* h, const double* x, int n) { static_cast<Scene*>(h)->SetDescriptor(VecN(x, x + n)); }
int orc_has_degenerate_inertia(void* h)
{
    for (auto& j : static_cast<Scene*>(h)->joint_store) if (j->degenerate_inertia) return 1;
    return 0;
}

void orc_set_cyclic(void* h, int c) { static_cast<Scene*>(h)->is_cyclic = (c != 0); }
void orc_set_weight_trajectory(void* h, double w) { static_cast<Scene*>(h)->weight_trajectory = w; }   // core.cpp:150-153
void orc_set_per_item_weight(void* h, int item, double w) { static_cast<Scene*>(h)->per_item_weights[item] = w; }
void orc_clear_kernels(void* h) { static_cast<Scene*>(h)->kernels.clear(); }
void orc_add_kernel(void* h, double amp, double mu, double sigma)
{
    static_cast<Scene*>(h)->kernels.push_back(GaussianKernel{amp, mu, sigma});
}
void orc_set_sampling_concurrency(void* h, int c) { static_cast<Scene*>(h)->sampling_concurrency = c; }
void orc_update_original(void* h, const double* x, int n) { static_cast<Scene*>(h)->UpdateOriginalParameters(VecN(x, x + n)); }
void orc_get_original(void* h, double* x)
{
    Scene* s = static_cast<Scene*>(h);
    std::copy(s->x_original.begin(), s->x_original.end(), x);
}
void orc_set_selected(void* h, int item, int sel) { static_cast<Scene*>(h)->GetItems()[item]->is_selected = (sel != 0); }
void orc_update_freezed_from_selection(void* h) { static_cast<Scene*>(h)->UpdateFreezedDofsFromSelection(); }
void orc_set_freezed(void* h, const int* f, int n)
{
    Scene* s = static_cast<Scene*>(h);
    s->freezed_dofs.assign(n, 0);
    for (int i = 0; i < n; ++i) s->freezed_dofs[i] = static_cast<char>(f[i] != 0);
}
void orc_get_freezed(void* h, int* f)
{
    Scene* s = static_cast<Scene*>(h);
    for (size_t i = 0; i < s->freezed_dofs.size(); ++i) f[i] = s->freezed_dofs[i];
}
double orc_cost_weight(void* h, double t) { return static_cast<Scene*>(h)->GetCostWeight(t); }
void orc_parameter_weights(void* h, double* w)
{
    const VecN v = static_cast<Scene*>(h)->CalculateParameterWeights();
    std::copy(v.begin(), v.end(), w);
}

/// Number of samples of the accumulated `t += dt` grid (core_rbdl.cpp:120-123) and the grid itself.
int orc_time_grid(double min_t, double max_t, double dt, double* out, int cap)
{
    int n = 0;
    for (double t = min_t; t <= max_t; t += dt) { if (out && n < cap) out[n] = t; ++n; }
    return n;
}

/// a3. terms = {C, D_param, D_traj}.
double orc_full_objective(void* h, double min_t, double max_t, const double* x, const double* x_orig, int n, double* terms)
{
    Scene* s = static_cast<Scene*>(h);
    const double f = s->CalculateFullObjective(min_t, max_t, VecN(x, x + n), VecN(x_orig, x_orig + n));
    if (terms) for (int k = 0; k < 3; ++k) terms[k] = s->last_terms[k];
    return f;
}

int orc_cache_samples(void* h) { return static_cast<int>(static_cast<Scene*>(h)->optimization_cache.time_sequence.size()); }
int orc_dof(void* h) { return static_cast<int>(static_cast<Scene*>(h)->model->dof_count()); }
/// which: 0 time, 1 q, 2 qd, 3 qdd, 4 tau, 5 local_cost, 6 p_original (N_s x 3J)
void orc_get_cache(void* h, int which, double* out)
{
    Scene* s = static_cast<Scene*>(h);
    const auto& c = s->optimization_cache;
    auto dump = [&](const std::vector<VecN>& a) { for (const VecN& r : a) { std::copy(r.begin(), r.end(), out); out += r.size(); } };
    switch (which) {
        case 0: std::copy(c.time_sequence.begin(), c.time_sequence.end(), out); break;
        case 1: dump(c.q); break;
        case 2: dump(c.qd); break;
        case 3: dump(c.qdd); break;
        case 4: dump(c.tau); break;
        case 5: std::copy(c.local_cost.begin(), c.local_cost.end(), out); break;
        case 6: if (s->original_motion_cache) dump(s->original_motion_cache->p_original); break;
    }
}

/// Joint world positions p (N_s x 3J) of descriptor x, via SetStates (core_rbdl.cpp:150-153).
int orc_positions(void* h, double min_t, double max_t, const double* x, int n, double* p_out)
{
    Scene* s = static_cast<Scene*>(h);
    s->SetDescriptor(VecN(x, x + n));
    std::vector<double> ts; std::vector<VecN> q, qd, qdd, p;
    s->SetStates(min_t, max_t, 0.20, static_cast<int>(s->model->dof_count()), &ts, &q, &qd, &qdd, &p);
    for (const VecN& r : p) { std::copy(r.begin(), r.end(), p_out); p_out += r.size(); }
    return static_cast<int>(ts.size());
}

/// N2: Core::CalculateTorques(t) (core_rbdl.cpp:264-289) of descriptor x at M times; tau / q / qd / qdd are [M][dof] (any NULL).
void orc_torques_at(void* h, const double* x, int n, const double* times, int M, double* tau, double* q, double* qd, double* qdd)
{
    Scene* s = static_cast<Scene*>(h);
    s->SetDescriptor(VecN(x, x + n));
    const size_t dof = s->model->dof_count();
    for (int i = 0; i < M; ++i) {
        VecN a, b, c;
        const VecN T = s->CalculateTorques(times[i], &a, &b, &c);
        if (tau) std::copy(T.begin(), T.end(), tau + i * dof);
        if (q) std::copy(a.begin(), a.end(), q + i * dof);
        if (qd) std::copy(b.begin(), b.end(), qd + i * dof);
        if (qdd) std::copy(c.begin(), c.end(), qdd + i * dof);
    }
}

/// N3: what Core::ConvertSpecificFrameToJson reads (core_io.cpp:264-294): Joint::GetEulerAngle(t) of EVERY joint in
/// GetJoints() order, the root's own rot channels included. out is [M][3J].
void orc_euler_angles_at(void* h, const double* x, int n, const double* times, int M, double* out)
{
    Scene* s = static_cast<Scene*>(h);
    s->SetDescriptor(VecN(x, x + n));
    for (int i = 0; i < M; ++i)
        for (Joint* j : s->GetJoints()) {
            const V3 e = j->GetEulerAngle(times[i]);
            for (int k = 0; k < 3; ++k) *out++ = e[k];
        }
}

/// a1/a2: EvaluateObjective (core_opt.cpp:42-57) with revert(x) = X^T x + Y x_initial.
double orc_evaluate_objective(void* h, const double* x_free, int n_free, const double* x_initial, int n_full, double min_t,
                              double max_t)
{
    Scene* s = static_cast<Scene*>(h);
    const std::vector<int> idx = s->FreeIndices();
    if (static_cast<int>(idx.size()) != n_free) return NAN;
    VecN full(x_initial, x_initial + n_full);
    for (int k = 0; k < n_free; ++k) full[idx[k]] = x_free[k];
    return s->CalculateFullObjective(min_t, max_t, full, s->x_original);
}

/// Constructed oracle for the NEW gradient: central differences of the CPU
/// objective on the free parameters, g_k = (f(x + h e_k) - f(x - h e_k)) / 2h. fpm (2F) receives f+ / f-.
double orc_gradient(void* h, const double* x_full, int n_full, double min_t, double max_t, double step, double* grad, double* fpm)
{
    Scene* s = static_cast<Scene*>(h);
    const std::vector<int> idx = s->FreeIndices();
    VecN x(x_full, x_full + n_full);
    const double f0 = s->CalculateFullObjective(min_t, max_t, x, s->x_original);
    for (size_t k = 0; k < idx.size(); ++k) {
        VecN xp = x, xm = x;
        xp[idx[k]] += step; xm[idx[k]] -= step;
        const double fp = s->CalculateFullObjective(min_t, max_t, xp, s->x_original, false);
        const double fm = s->CalculateFullObjective(min_t, max_t, xm, s->x_original, false);
        if (fpm) { fpm[2 * k] = fp; fpm[2 * k + 1] = fm; }
        grad[k] = (fp - fm) / (2.0 * step);
    }
    return f0;
}

// ---- low-level probes used to pin the oracle against real-RBDL known answers ------------------------

/// Generic tree: body i (1-based) has parent[i-1] (0 = base), Xtrans(xt), joint kind, mass, com, inertia(9, row-major).
void orc_probe_inverse_dynamics(int n_bodies, const int* parent, const int* kind, const double* xt, const double* mass,
                                const double* com, const double* inertia, const double* gravity, const double* q,
                                const double* qd, const double* qdd, double* tau)
{
    RbdModel m;
    for (int i = 0; i < n_bodies; ++i) {
        M3 I; for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) I(r, c) = inertia[9 * i + 3 * r + c];
        m.add_body(parent[i], V3(xt[3 * i], xt[3 * i + 1], xt[3 * i + 2]), kind[i], mass[i], V3(com[3 * i], com[3 * i + 1], com[3 * i + 2]), I);
    }
    m.set_gravity(V3(gravity[0], gravity[1], gravity[2]));
    const int n = 3 * n_bodies;
    VecN T(n, 0.0);
    m.inverse_dynamics(VecN(q, q + n), VecN(qd, qd + n), VecN(qdd, qdd + n), T);
    std::copy(T.begin(), T.end(), tau);
}

/// Serial chain of EulerZYX bodies (as ikhandler.cpp:166-188 builds it); returns 1 when converged.
int orc_probe_inverse_kinematics(int n_bodies, const double* xt, const double* mass, const double* com, const double* inertia,
                                 const double* point, const double* target, double tol, double lambda, int max_iter, double* qres,
                                 int* iters)
{
    RbdModel m;
    unsigned parent = 0;
    for (int i = 0; i < n_bodies; ++i) {
        M3 I; for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) I(r, c) = inertia[9 * i + 3 * r + c];
        parent = m.add_body(parent, V3(xt[3 * i], xt[3 * i + 1], xt[3 * i + 2]), kEulerZYX, mass[i], V3(com[3 * i], com[3 * i + 1], com[3 * i + 2]), I);
    }
    const int n = 3 * n_bodies;
    VecN Q;
    int it = 0;
    const bool ok = m.inverse_kinematics(VecN(n, 0.0), parent, V3(point[0], point[1], point[2]), V3(target[0], target[1], target[2]), Q,
                                         tol, lambda, max_iter, &it);
    std::copy(Q.begin(), Q.end(), qres);
    if (iters) *iters = it;
    return ok ? 1 : 0;
}

/// One curve built with AddControlPoint(t, v, theta, tb, tf) in the given order, evaluated at ts.
void orc_probe_curve(int n_cp, const double* cps5, int n_t, const double* ts, double* out)
{
    Variable v(0.0);
    for (int i = 0; i < n_cp; ++i) v.curve_.AddControlPoint(ControlPoint(cps5[5 * i], cps5[5 * i + 1], cps5[5 * i + 2], cps5[5 * i + 3], cps5[5 * i + 4]));
    for (int i = 0; i < n_t; ++i) out[i] = v.GetValue(ts[i]);
}

/// Segment chain (time order at insertion) of a variable: writes n_cp control-point indices.
int orc_var_chain(void* h, int item, int axis, int* order)
{
    Scene* s = static_cast<Scene*>(h);
    Variable* v = s->GetItems()[item]->vars()[axis];
    int n = 0;
    for (const auto& seg : v->curve_.GetCurveSegments()) if (seg.end_point_index >= 0) order[n++] = seg.end_point_index;
    return n;
}

void orc_joint_physical(void* h, int joint, double* out14)
{
    Joint* j = static_cast<Scene*>(h)->joint_store[joint].get();
    out14[0] = j->mass;
    for (int k = 0; k < 3; ++k) out14[1 + k] = j->com[k];
    for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) out14[4 + 3 * r + c] = j->inertia(r, c);
    out14[13] = 0.0;
}

void orc_counters(long* out4, int reset)
{
    out4[0] = g_bezier_evals; out4[1] = g_newton_iters; out4[2] = g_ik_solves; out4[3] = g_ik_iters;
    if (reset) { g_bezier_evals = 0; g_newton_iters = 0; g_ik_solves = 0; g_ik_iters = 0; }
}

int orc_hardware_concurrency() { return static_cast<int>(std::thread::hardware_concurrency()); }

/// CPU baseline: evaluates B candidates (rows of X) and returns wall seconds.
///   mode 0 "ref-as-is": candidates one after another; inside each, parallel_for over samples with
///          hardware_concurrency threads for sampling and a serial RNEA loop (what OptiMo does today).
///   mode 1 "ref-candidate-parallel": parallel_for over candidates with `threads` workers, each owning a
///          private object graph + model (the THREAD_SAFE idiom, core_rbdl.cpp:281-284); inner loops serial.
double orc_bench_batch(void* h, const double* X, int B, int D, double min_t, double max_t, int mode, int threads, double* f_out)
{
    Scene* s = static_cast<Scene*>(h);
    const VecN x0 = s->x_original;
    const auto t0 = std::chrono::steady_clock::now();
    if (mode == 0) {
        s->sampling_concurrency = 0;
        for (int c = 0; c < B; ++c) f_out[c] = s->CalculateFullObjective(min_t, max_t, VecN(X + (size_t)c * D, X + (size_t)(c + 1) * D), x0, false);
    } else {
        const int T = std::max(1, std::min(threads == 0 ? (int)std::thread::hardware_concurrency() : threads, B));
        std::vector<std::unique_ptr<Scene>> workers;
        for (int t = 0; t < T; ++t) {
            workers.emplace_back(clone_scene(*s));
            workers.back()->sampling_concurrency = 1;
        }
        std::atomic<int> next{0};
        std::vector<std::thread> th;
        for (int t = 0; t < T; ++t)
            th.emplace_back([&, t]() {
                Scene* w = workers[t].get();
                for (int c = next++; c < B; c = next++)
                    f_out[c] = w->CalculateFullObjective(min_t, max_t, VecN(X + (size_t)c * D, X + (size_t)(c + 1) * D), x0, false);
            });
        for (auto& t : th) t.join();
    }
    return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
}

}  // extern "C"

namespace orc {
static Scene* clone_scene(const Scene& s)
{
    void* h = orc_scene_new();
    for (size_t i = 0; i < s.ops.size(); ++i) {
        const BuildOp& op = s.ops[i];
        switch (op.kind) {
            case 0: orc_add_joint(h, op.a, op.d, s.ops[i + 1].d, op.s.c_str()); ++i; break;
            case 1: orc_joint_add_cp(h, op.a, op.b, op.d[0], op.d[1], op.d[2], op.d[3], op.d[4]); break;
            case 2: orc_add_ik(h, op.s.c_str(), op.a, op.b, op.d); break;
            case 3: orc_ik_add_cp(h, op.a, op.b, op.d[0], op.d[1], op.d[2], op.d[3], op.d[4]); break;
            case 4: orc_finalize(h, op.a, op.b); break;
            default: break;
        }
    }
    Scene* c = static_cast<Scene*>(h);
    c->is_cyclic = s.is_cyclic;
    c->freezed_dofs = s.freezed_dofs;
    c->per_item_weights = s.per_item_weights;
    c->weight_trajectory = s.weight_trajectory;
    c->kernels = s.kernels;
    c->x_original = s.x_original;
    return c;
}
}  // namespace orc
