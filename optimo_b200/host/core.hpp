// Headless `Core` of the B200 OptiMo objective path: same public names, argument meaning and error
// behaviour as the reference's Core for everything the objective path touches (OptiMo/core.h:35-245),
// minus camera / drawing / Qt. The arithmetic is NOT here: CalculateFullObjective and friends forward to the
// CUDA engine behind include/omo.h (libomo.so). There is no CPU implementation of the objective in this class.
//
// The Eigen::VectorXd arguments of the reference become std::vector<double> (Eigen is not a dependency).
#ifndef OPTIMO_B200_CORE_HPP
#define OPTIMO_B200_CORE_HPP

#include <cmath>
#include <list>
#include <memory>
#include <mutex>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/omo.h"
#include "object.hpp"

namespace optimo {

/// Gaussian kernel: amp * exp(- (x - mu)^2 / (2 sigma^2))   (OptiMo/core.h:27-33)
struct GaussianKernel {
    double amp, mu, sigma;
    double GetValue(double x) const { return amp * std::exp(-(x - mu) * (x - mu) / (2.0 * sigma * sigma)); }
};

/// Thrown (like nlopt::forced_stop) by EvaluateObjective when Core::stop_optimization_ is set (core_opt.cpp:50-54).
struct forced_stop : std::runtime_error { forced_stop() : std::runtime_error("nlopt forced stop") {} };
/// Any failure of the CUDA engine (there is no fallback path).
struct engine_error : std::runtime_error { explicit engine_error(const std::string& m) : std::runtime_error(m) {} };

class Core {
public:
    Core();   // default humanoid (core.cpp:23-103); touches no CUDA state
    ~Core();
    Core(const Core&) = delete;
    Core& operator=(const Core&) = delete;

    static Core& GetInstance() { static Core core; return core; }

    int min_frame_ = 1, max_frame_ = 40, current_frame_ = 1, frames_per_second_ = 24;
    /// If true, the next objective callback throws forced_stop (set from another thread; core.h:64-66).
    bool stop_optimization_ = false;
    std::shared_ptr<Object> object_;

    /// True: the dof is not optimised. Same order as the descriptor (core.h:73-76).
    std::vector<bool> freezed_dofs_;
    void UpdateFreezedDofsFromSelection();                                  // core.cpp:118-148
    std::vector<double> per_item_weights_;
    void UpdatePerItemWeight(double weight, const std::string& item_name);  // core.cpp:155-159
    double weight_trajectory_ = 0.0;
    void UpdateGlobalWeight(double weight);                                 // core.cpp:150-153

    /// core_opt.cpp:60-157 -- LN_BOBYQA over the free dofs through NLopt (loaded at run time, see core_opt.cpp).
    void PerformOptimization(bool explicit_termination = false);
    /// Next row N1: the same solve driven by a gradient-based NLopt algorithm, as the reference's helper
    /// nloptutil::unconstrained::grad_based::bounded::solve would set it up (nlopt-util.hpp:226-274) -- every callback
    /// returns f and the central-difference gradient over the free dofs from one sparse GPU batch (include/omo.h:
    /// omo_eval_grad). Bounds on t_b / t_f are pulled in by the difference step so that x +- h stays feasible.
    /// analytic = true: the gradient comes from omo_eval_grad_analytic (next row N4) -- one evaluation of the per-sample arrays
    /// instead of 2F+1 objective columns; scenes it does not cover (IK handles) silently keep the finite-difference gradient.
    void PerformOptimizationWithGradient(bool explicit_termination = false, const std::string& algorithm = "LD_LBFGS", bool analytic = false);

    /// core_rbdl.cpp:348-401. Like the reference it leaves the object at `x`.
    double CalculateFullObjective(double min_t, double max_t, const VecN& x, const VecN& x_original);
    /// Batch form: X is row-major [B][D] in host memory; f_out[B]; terms_out[B] optional.
    void CalculateFullObjectiveBatch(double min_t, double max_t, const double* X, int B, const VecN& x_original, double* f_out,
                                     omo_terms* terms_out = nullptr);
    /// NEW: f(x) and its central-difference gradient on the free dofs.
    double CalculateFullObjectiveAndGradient(double min_t, double max_t, const VecN& x, const VecN& x_original,
                                             const std::vector<int>& free_indices, double h, VecN* grad, VecN* f_plus_minus = nullptr);
    /// Next row N4: f(x) and its ANALYTIC gradient on the free dofs (include/omo.h: omo_eval_grad_analytic); scenes it does not
    /// cover (IK handles, unsorted key times) get the finite-difference gradient instead (*analytic_used tells which).
    double CalculateFullObjectiveAndAnalyticGradient(double min_t, double max_t, const VecN& x, const VecN& x_original,
                                                     const std::vector<int>& free_indices, VecN* grad, bool* analytic_used = nullptr);
    /// (C, D_param, D_traj) of the last CalculateFullObjective call.
    omo_terms last_terms_{0, 0, 0};
    /// What an evaluation leaves behind (object at x, cache stale) for a value that was served from the batched start.
    void NoteEvaluatedFromCache(const VecN& x, double min_t, double max_t);

    void ReadJson(const std::string& file_path);          // core_io.cpp:154-163
    void ReadJsonString(const std::string& json_text);    // ApplyJsonToScene + InitializeDynamicsModel
    void WriteJson(const std::string& file_path) const;   // core_io.cpp:165-169
    std::string ConvertCurrentSceneToJson() const;        // core_io.cpp:216-226

    // ---- next rows N2 / N3: the callers either side of the path ------------------------------------
    /// core_rbdl.cpp:264-289: joint torques of the CURRENT object at time t (3-point differences over t -+ 0.2,
    /// SetCurrentState core_rbdl.cpp:168-198). One GPU launch; see CalculateTorquesBatch for many t at once.
    VecN CalculateTorques(double t);
    /// The same for M times in one launch (the UI's heat strip asks for one t per pixel column, graphwidget.cpp:325-340).
    /// tau [M][dof] row-major; q / qd / qdd optional, same shape.
    void CalculateTorquesBatch(const double* times, int M, double* tau, double* q = nullptr, double* qd = nullptr, double* qdd = nullptr);
    /// Joint::GetEulerAngle(t) of every joint in GetJoints() order, root included (joint.cpp:36-52), M x 3J, on the GPU.
    std::vector<double> CalculateEulerAngles(const double* times, int M);
    double max_torque_ = 0.0, max_cost_ = 0.0;            // core.h:96-97
    void UpdateMaximumTorqueAndCost();                    // core.cpp:105-116
    void WriteHistory(const std::string& file_path) const;          // core_io.cpp:171-186
    void PushHistory();                                            // core_io.cpp:188-192
    void PopHistory();                                             // core_io.cpp:194-208 (undo: re-applies the previous stamp)
    void ClearHistory() { history_.clear(); }                      // core_io.cpp:210-213
    int GetHistorySize() const { return static_cast<int>(history_.size()); }
    void ExportMotion(const std::string& file_path);                // core_io.cpp:252-264
    std::string ConvertSpecificFrameToJson(int t);                  // core_io.cpp:266-294
    std::string ConvertMotionToJson();                              // the text ExportMotion writes

    struct OptimizationCache {  // core.h:120-131
        std::vector<double> time_sequence;
        std::vector<VecN> q, q_dot, q_ddot, tau;
        std::vector<double> local_cost_sequence;
        int GetSize() const { return static_cast<int>(time_sequence.size()); }
    };
    /// Thread-safe copy (the UI thread polls it at 30 Hz while the solver thread evaluates, mainwindow.cpp:228-235,
    /// core.h:132-143). The per-sample arrays of the last evaluated x are fetched from the GPU on demand (the reference
    /// fills them on every evaluation, core_rbdl.cpp:341-343; the observable content is the same). Like every method
    /// that drives the engine it takes engine_mutex_ for the whole omo_* sequence: an omo_ctx serves one host thread
    /// at a time (include/omo.h), so a cache / torque query waits for the evaluation in flight and vice versa.
    OptimizationCache GetOptimizationCache();
    void ClearOptimizationCache();

    void UpdateOriginalParameters(const VecN& x_original);                       // core.cpp:184-189
    const VecN& RetrieveOriginalParameters() const { return x_original_; }
    void AddKernel(double t) { kernels_.push_back(GaussianKernel{1.0, t, 1.0}); }  // core.h:163
    double GetCostWeight(double t) const;                                          // core.cpp:191-199
    GaussianKernel* GetKernelPointer(int index) { return &kernels_[index]; }
    GaussianKernel* GetLastKernelPointer() { return GetKernelPointer(static_cast<int>(kernels_.size()) - 1); }
    void ClearKernels() { kernels_.clear(); }
    int GetNumKernels() const { return static_cast<int>(kernels_.size()); }

    VecN CalculateParameterWeights() const;   // core.cpp:161-182
    /// Joint world positions (N_s x 3J) of descriptor x over [min_t, max_t] (SetStates' p, core_rbdl.cpp:150-153).
    std::vector<double> CalculatePositions(double min_t, double max_t, const VecN& x);

    /// CUDA device the engine is created on (default 0, or $OPTIMO_DEVICE / $LOCAL_RANK).
    void SetDevice(int device);
    int GetDevice() const { return device_; }
    /// Several GPUs of one box behind the SAME calls: batches are sharded over candidates and gradients
    /// over their 2F+1 finite-difference columns by an omo_multi (include/omo.h); single evaluations run on the first
    /// device. Results are bit-identical to the one-device engine. An empty / one-entry list goes back to SetDevice.
    void SetDevices(const std::vector<int>& devices);
    const std::vector<int>& GetDevices() const { return devices_; }
    omo_multi* MultiEngine();   // nullptr unless SetDevices named >= 2 devices
    /// The engine behind the objective; created lazily (never before main), rebuilt when the scene changes.
    omo_ctx* Engine();
    /// Generates and NVRTC-compiles the scene-specialised objective kernel of the current scene on the host (no GPU
    /// needed); returns the omo_status, `log` gets the placement summary + compiler log. See omo_jit_probe.
    int ProbeSpecialisedKernel(bool fk, int flags, const std::string& dump_prefix, std::string* log);   // flags: omo_jit_probe's `count`
    /// Host-only: the flattened device scene table of the current scene (omo_debug_scene_table), for tests/native.
    int DebugSceneTable(void* out, unsigned long long cap, int* sub_end_out);
    /// Called after any structural change of object_ (core_rbdl.cpp:200-262): drops the engine.
    void InitializeDynamicsModel();

    void Solve(bool explicit_termination, const char* algorithm, bool want_gradient, bool analytic = false);   // body of both PerformOptimization forms
    /// Result of the last PerformOptimization (function value, evaluations).
    double last_solve_value_ = 0.0;
    int last_solve_evals_ = 0;
    std::string last_solve_status_;
    /// Batched BOBYQA start: evaluate the solver's first 2M+1 points as one GPU batch (default on; results are bit-identical to
    /// the serial start). last_solve_batched_ = callbacks of the last solve that were served from that batch.
    bool batched_start_ = true;
    int last_solve_batched_ = 0;
    /// Optional record of every callback of the last solve, in call order: M free parameters then f per row
    /// (used to replay a solve against another implementation of the objective).
    bool record_solve_trace_ = false;
    int solve_trace_dim_ = 0;
    std::vector<double> solve_trace_;

private:
    void SyncEngine(double min_t, double max_t, const VecN& x_original);
    void SyncEngineState(const VecN& x_original);   // weights + x_original, no window
    void SyncMulti(double min_t, double max_t, const VecN& x_original);
    void NoteEvaluated(const VecN& x, double min_t, double max_t);
    void ApplyJsonToScene(const std::string& json_text);   // core_io.cpp:228-250
    std::string FrameJson(int frame, const double* euler) const;
    std::vector<std::string> history_;              // core.h: std::list<std::string> history_
    void Check(omo_status st, const char* what);

    std::vector<GaussianKernel> kernels_;
    VecN x_original_;
    omo_ctx* engine_ = nullptr;       // owned, or omo_multi_ctx(multi_, 0) when multi_ exists
    omo_multi* multi_ = nullptr;
    std::vector<int> devices_;
    /// Serialises every omo_* sequence (SyncEngine .. last copy-out) between the solver thread and UI-side queries.
    std::recursive_mutex engine_mutex_;
    bool probe_only_ = false;
    int probe_fk_ = 0, probe_count_ = 0, probe_status_ = 0;
    std::string probe_prefix_, probe_log_;
    void* debug_table_out_ = nullptr; unsigned long long debug_table_cap_ = 0; int* debug_sub_end_out_ = nullptr;
    int device_ = 0;

    std::mutex optimization_cache_mutex_;
    OptimizationCache optimization_cache_;
    bool cache_stale_ = true;
    VecN last_x_;
    double last_min_t_ = 0.0, last_max_t_ = 0.0;
    bool have_last_ = false;
};

/// nlopt::vfunc-compatible objective (core_opt.cpp:42-57): `data` is an EvaluateObjectiveData*.
/// `grad` is filled only when data->want_gradient is set AND grad is non-empty (SURVEY.md §8(b) trap).
struct EvaluateObjectiveData {
    Core* core;
    double min_t, max_t;
    VecN x_original;   // #{full dofs}
    VecN x_initial;    // #{full dofs}: values of the frozen dofs (Y * x_initial, core_opt.cpp:121)
    std::vector<int> free_indices;  // X: free -> full
    bool want_gradient = false;
    bool analytic_gradient = false;   // next row N4: omo_eval_grad_analytic instead of the 2F+1 finite-difference columns
    double gradient_step = 1e-4;
    int evaluations = 0;
    // batched BOBYQA start (core_opt.cpp): predicted points in the callback's parameter space and their values
    std::vector<VecN> start_cache_x;
    std::vector<double> start_cache_f;
    int start_cache_hits = 0;
};
double EvaluateObjective(const std::vector<double>& x, std::vector<double>& grad, void* data);

}  // namespace optimo
#endif
