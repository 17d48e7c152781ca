// C wrapper over the headless Core (core.hpp) so that non-C++ hosts (the Python tests and bench harness via
// ctypes, or any FFI) can drive the reference-facing API. Declared in include/optimo_core.h.
#include <cstring>
#include <string>

#include "../../include/optimo_core.h"
#include "core.hpp"

using optimo::Core;
using optimo::VecN;

namespace {
thread_local std::string g_err;

template <typename F> int guarded(F f)
{
    try { f(); g_err.clear(); return 0; }
    catch (const optimo::forced_stop& e) { g_err = e.what(); return 2; }
    catch (const std::exception& e) { g_err = e.what(); return 1; }
    catch (...) { g_err = "unknown exception"; return 1; }
}
Core* C(void* h) { return static_cast<Core*>(h); }
}  // namespace

extern "C" {

const char* optimo_last_error(void) { return g_err.c_str(); }

void* optimo_core_new(void) { try { return new Core(); } catch (...) { return nullptr; } }
void* optimo_core_instance(void) { return &Core::GetInstance(); }
void optimo_core_free(void* h) { delete C(h); }

int optimo_read_json(void* h, const char* path) { return guarded([&] { C(h)->ReadJson(path); }); }
int optimo_read_json_string(void* h, const char* text) { return guarded([&] { C(h)->ReadJsonString(text); }); }
int optimo_write_json(void* h, const char* path) { return guarded([&] { C(h)->WriteJson(path); }); }
int optimo_scene_json(void* h, char* out, int cap)
{
    std::string s;
    const int rc = guarded([&] { s = C(h)->ConvertCurrentSceneToJson(); });
    if (rc) return -1;
    if (out && cap > 0) { std::strncpy(out, s.c_str(), cap - 1); out[cap - 1] = 0; }
    return static_cast<int>(s.size());
}
int optimo_set_device(void* h, int device) { return guarded([&] { C(h)->SetDevice(device); }); }
int optimo_set_devices(void* h, const int* devices, int n)
{
    return guarded([&] { C(h)->SetDevices(n > 0 && devices ? std::vector<int>(devices, devices + n) : std::vector<int>()); });
}

int optimo_min_frame(void* h) { return C(h)->min_frame_; }
int optimo_max_frame(void* h) { return C(h)->max_frame_; }
void optimo_set_frames(void* h, int min_frame, int max_frame) { C(h)->min_frame_ = min_frame; C(h)->max_frame_ = max_frame; }
int optimo_num_joints(void* h) { return static_cast<int>(C(h)->object_->GetJoints().size()); }
int optimo_num_items(void* h) { return static_cast<int>(C(h)->object_->GetItems().size()); }
int optimo_item_name(void* h, int index, char* out, int cap)
{
    const auto items = C(h)->object_->GetItems();
    if (index < 0 || index >= static_cast<int>(items.size())) return -1;
    std::strncpy(out, items[index]->name_.c_str(), cap - 1); out[cap - 1] = 0;
    return 0;
}
int optimo_descriptor_size(void* h) { return static_cast<int>(C(h)->object_->GetDescriptor().size()); }
int optimo_get_descriptor(void* h, double* x)
{
    return guarded([&] { const VecN d = C(h)->object_->GetDescriptor(); std::copy(d.begin(), d.end(), x); });
}
int optimo_set_descriptor(void* h, const double* x, int n) { return guarded([&] { C(h)->object_->SetDescriptor(VecN(x, x + n)); }); }
void optimo_set_cyclic(void* h, int cyclic) { C(h)->object_->is_cyclic_ = cyclic != 0; }

int optimo_update_original_parameters(void* h, const double* x, int n) { return guarded([&] { C(h)->UpdateOriginalParameters(VecN(x, x + n)); }); }
int optimo_retrieve_original_parameters(void* h, double* x)
{
    return guarded([&] { const VecN& d = C(h)->RetrieveOriginalParameters(); std::copy(d.begin(), d.end(), x); });
}
void optimo_update_global_weight(void* h, double w) { C(h)->UpdateGlobalWeight(w); }
int optimo_update_per_item_weight(void* h, double w, const char* item) { return guarded([&] { C(h)->UpdatePerItemWeight(w, item); }); }
void optimo_clear_kernels(void* h) { C(h)->ClearKernels(); }
void optimo_add_kernel(void* h, double amp, double mu, double sigma)
{
    C(h)->AddKernel(mu);
    optimo::GaussianKernel* k = C(h)->GetLastKernelPointer();
    k->amp = amp; k->mu = mu; k->sigma = sigma;
}
double optimo_get_cost_weight(void* h, double t) { return C(h)->GetCostWeight(t); }
int optimo_parameter_weights(void* h, double* w)
{
    return guarded([&] { const VecN v = C(h)->CalculateParameterWeights(); std::copy(v.begin(), v.end(), w); });
}
int optimo_set_selected(void* h, const char* item, int selected)
{
    return guarded([&] {
        auto it = C(h)->object_->GetItemByName(item);
        if (!it) throw std::invalid_argument(std::string("no item named ") + item);
        it->is_selected_ = selected != 0;
    });
}
void optimo_update_freezed_dofs_from_selection(void* h) { C(h)->UpdateFreezedDofsFromSelection(); }
int optimo_get_freezed_dofs(void* h, int* out)
{
    const auto& f = C(h)->freezed_dofs_;
    for (size_t i = 0; i < f.size(); ++i) out[i] = f[i] ? 1 : 0;
    return static_cast<int>(f.size());
}
int optimo_set_freezed_dofs(void* h, const int* in, int n)
{
    return guarded([&] {
        if (n != static_cast<int>(C(h)->object_->GetDescriptor().size())) throw std::invalid_argument("freezed_dofs length mismatch");
        C(h)->freezed_dofs_.assign(n, false);
        for (int i = 0; i < n; ++i) C(h)->freezed_dofs_[i] = in[i] != 0;
    });
}
int optimo_joint_physical(void* h, int joint, double* out13)
{
    return guarded([&] {
        const auto joints = C(h)->object_->GetJoints();
        const auto& j = joints.at(joint);
        out13[0] = j->mass_;
        for (int k = 0; k < 3; ++k) out13[1 + k] = j->center_of_mass_position_parent_[k];
        for (int k = 0; k < 9; ++k) out13[4 + k] = j->inertia_at_center_parent_[k];
    });
}

int optimo_calculate_full_objective(void* h, double min_t, double max_t, const double* x, const double* x_original, int n, double* f,
                                    double* terms3)
{
    return guarded([&] {
        *f = C(h)->CalculateFullObjective(min_t, max_t, VecN(x, x + n), VecN(x_original, x_original + n));
        if (terms3) { terms3[0] = C(h)->last_terms_.C; terms3[1] = C(h)->last_terms_.D_param; terms3[2] = C(h)->last_terms_.D_traj; }
    });
}
int optimo_calculate_full_objective_batch(void* h, double min_t, double max_t, const double* X, int B, const double* x_original, int n,
                                          double* f, double* terms3)
{
    return guarded([&] {
        C(h)->CalculateFullObjectiveBatch(min_t, max_t, X, B, VecN(x_original, x_original + n), f, reinterpret_cast<omo_terms*>(terms3));
    });
}
int optimo_calculate_gradient(void* h, double min_t, double max_t, const double* x, const double* x_original, int n, const int* free_idx,
                              int F, double step, double* f0, double* grad, double* fpm)
{
    return guarded([&] {
        VecN g, pm;
        *f0 = C(h)->CalculateFullObjectiveAndGradient(min_t, max_t, VecN(x, x + n), VecN(x_original, x_original + n),
                                                      std::vector<int>(free_idx, free_idx + F), step, &g, fpm ? &pm : nullptr);
        std::copy(g.begin(), g.end(), grad);
        if (fpm) std::copy(pm.begin(), pm.end(), fpm);
    });
}
int optimo_positions(void* h, double min_t, double max_t, const double* x, int n, double* p)
{
    return guarded([&] { const auto v = C(h)->CalculatePositions(min_t, max_t, VecN(x, x + n)); std::copy(v.begin(), v.end(), p); });
}
int optimo_num_samples(void* h, double min_t, double max_t)
{
    int n = 0;
    for (double t = min_t; t <= max_t; t += 0.20) ++n;
    (void)h;
    return n;
}

/// EvaluateObjective exactly as NLopt would call it: x over the free dofs (freezed_dofs_), frozen ones from x_initial.
int optimo_evaluate_objective(void* h, const double* x_free, int n_free, const double* x_initial, int n_full, double min_t, double max_t,
                              int want_gradient, double step, double* f, double* grad)
{
    return guarded([&] {
        optimo::EvaluateObjectiveData d;
        d.core = C(h); d.min_t = min_t; d.max_t = max_t;
        d.x_original = C(h)->RetrieveOriginalParameters();
        d.x_initial.assign(x_initial, x_initial + n_full);
        for (size_t i = 0; i < C(h)->freezed_dofs_.size(); ++i) if (!C(h)->freezed_dofs_[i]) d.free_indices.push_back(static_cast<int>(i));
        d.want_gradient = want_gradient != 0; d.gradient_step = step;
        std::vector<double> x(x_free, x_free + n_free), g(grad ? n_free : 0, 0.0);
        *f = optimo::EvaluateObjective(x, g, &d);
        if (grad) std::copy(g.begin(), g.end(), grad);
    });
}
void optimo_set_stop_optimization(void* h, int flag) { C(h)->stop_optimization_ = flag != 0; }
int optimo_get_stop_optimization(void* h) { return C(h)->stop_optimization_ ? 1 : 0; }

int optimo_perform_optimization(void* h, int explicit_termination, double* final_value, int* evals)
{
    return guarded([&] {
        C(h)->PerformOptimization(explicit_termination != 0);
        if (final_value) *final_value = C(h)->last_solve_value_;
        if (evals) *evals = C(h)->last_solve_evals_;
    });
}

void optimo_set_batched_start(void* h, int on) { C(h)->batched_start_ = on != 0; }
int optimo_last_solve_batched(void* h) { return C(h)->last_solve_batched_; }
int optimo_perform_optimization_gradient(void* h, int explicit_termination, const char* algorithm, double* final_value, int* evals)
{
    return guarded([&] {
        // "LD_SLSQP+analytic": the same algorithm over the analytic gradient (next row N4)
        std::string alg = algorithm ? algorithm : "LD_LBFGS";
        bool analytic = false;
        const size_t plus = alg.find("+analytic");
        if (plus != std::string::npos) { analytic = true; alg.erase(plus); }
        C(h)->PerformOptimizationWithGradient(explicit_termination != 0, alg, analytic);
        if (final_value) *final_value = C(h)->last_solve_value_;
        if (evals) *evals = C(h)->last_solve_evals_;
    });
}

int optimo_debug_scene_table(void* h, void* out, unsigned long long cap, int* sub_end_out)
{
    int status = 1;
    const int rc = guarded([&] { status = C(h)->DebugSceneTable(out, cap, sub_end_out); });
    return rc ? rc : status;
}
int optimo_calculate_analytic_gradient(void* h, double min_t, double max_t, const double* x, const double* x_original, int n, const int* free_idx, int F,
                                       double* f0, double* grad, int* analytic_used)
{
    return guarded([&] {
        VecN g;
        bool used = false;
        *f0 = C(h)->CalculateFullObjectiveAndAnalyticGradient(min_t, max_t, VecN(x, x + n), VecN(x_original, x_original + n), std::vector<int>(free_idx, free_idx + F), &g, &used);
        std::copy(g.begin(), g.end(), grad);
        if (analytic_used) *analytic_used = used ? 1 : 0;
    });
}
int optimo_probe_specialised_kernel(void* h, int fk, int count, const char* dump_prefix, char* log, int cap)
{
    int status = 0;
    const int rc = guarded([&] {
        std::string l;
        status = C(h)->ProbeSpecialisedKernel(fk != 0, count, dump_prefix ? dump_prefix : "", &l);
        if (log && cap > 0) { std::strncpy(log, l.c_str(), static_cast<size_t>(cap) - 1); log[cap - 1] = 0; }
    });
    return rc != 0 ? rc : (status == 0 ? 0 : 1);
}

void optimo_record_solve_trace(void* h, int enable) { C(h)->record_solve_trace_ = enable != 0; }
int optimo_solve_trace_size(void* h, int* rows, int* dim)
{
    return guarded([&] {
        *dim = C(h)->solve_trace_dim_;
        *rows = *dim > 0 ? static_cast<int>(C(h)->solve_trace_.size()) / (*dim + 1) : 0;
    });
}
int optimo_get_solve_trace(void* h, double* out)
{
    return guarded([&] { std::copy(C(h)->solve_trace_.begin(), C(h)->solve_trace_.end(), out); });
}

// optimo_cache_size takes ONE consistent copy of the cache (Core::GetOptimizationCache, thread-safe) and keeps it per calling
// thread; optimo_get_cache then reads that copy. A UI thread polling during a solve therefore never sees arrays of two
// different evaluations (or of two different windows) in one read-out.
namespace { thread_local Core::OptimizationCache t_cache_snapshot; }
int optimo_cache_size(void* h, int* n_samples, int* dof)
{
    return guarded([&] {
        t_cache_snapshot = C(h)->GetOptimizationCache();
        const auto& c = t_cache_snapshot;
        *n_samples = c.GetSize();
        *dof = c.q.empty() ? 0 : static_cast<int>(c.q[0].size());
    });
}
/// which: 0 time, 1 q, 2 q_dot, 3 q_ddot, 4 tau, 5 local_cost -- of the copy taken by this thread's last optimo_cache_size
int optimo_get_cache(void* h, int which, double* out)
{
    return guarded([&] {
        const auto& c = t_cache_snapshot;
        auto dump = [&](const std::vector<VecN>& a) { for (const VecN& r : a) { std::copy(r.begin(), r.end(), out); out += r.size(); } };
        switch (which) {
            case 0: std::copy(c.time_sequence.begin(), c.time_sequence.end(), out); break;
            case 1: dump(c.q); break;
            case 2: dump(c.q_dot); break;
            case 3: dump(c.q_ddot); break;
            case 4: dump(c.tau); break;
            case 5: std::copy(c.local_cost_sequence.begin(), c.local_cost_sequence.end(), out); break;
            default: throw std::invalid_argument("bad cache selector");
        }
    });
}
void optimo_clear_optimization_cache(void* h) { C(h)->ClearOptimizationCache(); }

// ---- next rows N2 / N3 --------------------------------------------------------------------------------------------------
int optimo_calculate_torques(void* h, const double* times, int M, double* tau, double* q, double* qd, double* qdd)
{
    return guarded([&] { C(h)->CalculateTorquesBatch(times, M, tau, q, qd, qdd); });
}
int optimo_euler_angles(void* h, const double* times, int M, double* out)
{
    return guarded([&] { const auto v = C(h)->CalculateEulerAngles(times, M); std::copy(v.begin(), v.end(), out); });
}
int optimo_update_maximum_torque_and_cost(void* h, double* max_torque, double* max_cost)
{
    return guarded([&] {
        C(h)->UpdateMaximumTorqueAndCost();
        if (max_torque) *max_torque = C(h)->max_torque_;
        if (max_cost) *max_cost = C(h)->max_cost_;
    });
}
int optimo_export_motion(void* h, const char* path) { return guarded([&] { C(h)->ExportMotion(path); }); }
int optimo_motion_json(void* h, char* out, int cap)
{
    std::string s;
    const int rc = guarded([&] { s = C(h)->ConvertMotionToJson(); });
    if (rc) return -1;
    if (out && cap > 0) { std::strncpy(out, s.c_str(), cap - 1); out[cap - 1] = 0; }
    return static_cast<int>(s.size());
}
int optimo_push_history(void* h) { return guarded([&] { C(h)->PushHistory(); }); }
int optimo_pop_history(void* h) { return guarded([&] { C(h)->PopHistory(); }); }
void optimo_clear_history(void* h) { C(h)->ClearHistory(); }
int optimo_history_size(void* h) { return C(h)->GetHistorySize(); }
int optimo_write_history(void* h, const char* path) { return guarded([&] { C(h)->WriteHistory(path); }); }

void* optimo_engine(void* h)
{
    void* e = nullptr;
    guarded([&] { e = C(h)->Engine(); });
    return e;
}
void* optimo_engine_multi(void* h)
{
    void* e = nullptr;
    guarded([&] { e = C(h)->MultiEngine(); });
    return e;
}
/// Pushes window / weights / x_original into the engine (what every evaluation does first) and returns it.
void* optimo_engine_synced(void* h, double min_t, double max_t)
{
    void* e = nullptr;
    guarded([&] {
        double f;  // one evaluation of the current descriptor is the simplest full sync
        const VecN x = C(h)->object_->GetDescriptor();
        f = C(h)->CalculateFullObjective(min_t, max_t, x, C(h)->RetrieveOriginalParameters());
        (void)f;
        e = C(h)->Engine();
    });
    return e;
}

}  // extern "C"
