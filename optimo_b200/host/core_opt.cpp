// Core::PerformOptimization -- the caller of the objective path (OptiMo/core_opt.cpp:60-157 over
// external/nlopt-util/include/nlopt-util.hpp:44-195). NLopt itself is not rebuilt here: it is the reference's
// own solver library, loaded at run time through its C API (nlopt.h) from
//   $OPTIMO_NLOPT_LIB, else <this library's dir>/../../third_party/_ref/libnlopt.so, else "libnlopt.so[.0]".
// Only the objective callback is ours, and it runs on the GPU.
#include <dlfcn.h>

#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <iostream>

#include "core.hpp"

namespace optimo {

namespace {

constexpr double kGradientStep = 1e-4;   // central-difference step of the gradient-based solve

typedef double (*nlopt_func)(unsigned n, const double* x, double* gradient, void* func_data);
typedef struct nlopt_opt_s* nlopt_opt;

struct NloptApi {
    void* handle = nullptr;
    int (*algorithm_from_string)(const char*) = nullptr;
    nlopt_opt (*create)(int, unsigned) = nullptr;
    void (*destroy)(nlopt_opt) = nullptr;
    int (*optimize)(nlopt_opt, double*, double*) = nullptr;
    int (*set_min_objective)(nlopt_opt, nlopt_func, void*) = nullptr;
    int (*set_lower_bounds)(nlopt_opt, const double*) = nullptr;
    int (*set_upper_bounds)(nlopt_opt, const double*) = nullptr;
    int (*set_maxeval)(nlopt_opt, int) = nullptr;
    int (*set_ftol_rel)(nlopt_opt, double) = nullptr;
    int (*set_xtol_rel)(nlopt_opt, double) = nullptr;
    int (*get_initial_step)(nlopt_opt, const double*, double*) = nullptr;
    int (*set_initial_step)(nlopt_opt, const double*) = nullptr;
    int (*force_stop)(nlopt_opt) = nullptr;
    int (*get_numevals)(nlopt_opt) = nullptr;
    std::string path;
};

std::string self_dir()
{
    Dl_info info;
    if (dladdr(reinterpret_cast<void*>(&self_dir), &info) && info.dli_fname) {
        std::string p(info.dli_fname);
        const size_t k = p.rfind('/');
        return k == std::string::npos ? "." : p.substr(0, k);
    }
    return ".";
}

NloptApi& nlopt_api()
{
    static NloptApi api;
    if (api.handle) return api;
    std::vector<std::string> candidates;
    if (const char* env = std::getenv("OPTIMO_NLOPT_LIB")) candidates.push_back(env);
    candidates.push_back(self_dir() + "/../../third_party/_ref/libnlopt.so");
    candidates.push_back("libnlopt.so");
    candidates.push_back("libnlopt.so.0");
    for (const std::string& c : candidates) {
        api.handle = dlopen(c.c_str(), RTLD_NOW | RTLD_LOCAL);
        if (api.handle) { api.path = c; break; }
    }
    if (!api.handle) throw std::runtime_error("PerformOptimization: cannot load NLopt (set OPTIMO_NLOPT_LIB or build third_party/_ref)");
#define OPTIMO_SYM(field, name)                                                             \
    api.field = reinterpret_cast<decltype(api.field)>(dlsym(api.handle, name));             \
    if (!api.field) throw std::runtime_error(std::string("NLopt symbol missing: ") + name)
    OPTIMO_SYM(algorithm_from_string, "nlopt_algorithm_from_string");
    OPTIMO_SYM(create, "nlopt_create");
    OPTIMO_SYM(destroy, "nlopt_destroy");
    OPTIMO_SYM(optimize, "nlopt_optimize");
    OPTIMO_SYM(set_min_objective, "nlopt_set_min_objective");
    OPTIMO_SYM(set_lower_bounds, "nlopt_set_lower_bounds");
    OPTIMO_SYM(set_upper_bounds, "nlopt_set_upper_bounds");
    OPTIMO_SYM(set_maxeval, "nlopt_set_maxeval");
    OPTIMO_SYM(set_ftol_rel, "nlopt_set_ftol_rel");
    OPTIMO_SYM(set_xtol_rel, "nlopt_set_xtol_rel");
    OPTIMO_SYM(get_initial_step, "nlopt_get_initial_step");
    OPTIMO_SYM(set_initial_step, "nlopt_set_initial_step");
    OPTIMO_SYM(force_stop, "nlopt_force_stop");
    OPTIMO_SYM(get_numevals, "nlopt_get_numevals");
#undef OPTIMO_SYM
    return api;
}

struct Trampoline {
    EvaluateObjectiveData* data;
    nlopt_opt opt;
    NloptApi* api;
    std::vector<double> xtmp, gradtmp;
    std::string error;
};

// What nlopt::opt::myvfunc does (external/nlopt/src/api/nlopt-in.hpp:170-193): copy x into a std::vector,
// call the vfunc, map exceptions to nlopt_force_stop.
double trampoline(unsigned n, const double* x, double* grad, void* p)
{
    Trampoline* t = static_cast<Trampoline*>(p);
    try {
        t->xtmp.assign(x, x + n);
        if (grad) t->gradtmp.assign(n, 0.0); else t->gradtmp.clear();
        const double val = EvaluateObjective(t->xtmp, t->gradtmp, t->data);
        Core* core = t->data->core;
        if (core->record_solve_trace_) {
            core->solve_trace_.insert(core->solve_trace_.end(), x, x + n);
            core->solve_trace_.push_back(val);
        }
        if (grad) std::copy(t->gradtmp.begin(), t->gradtmp.end(), grad);
        return val;
    } catch (const forced_stop&) {
        t->error = "forced_stop";
    } catch (const std::exception& e) {
        t->error = e.what();
    }
    t->api->force_stop(t->opt);
    return HUGE_VAL;
}

}  // namespace

// The first 2n+1 points BOBYQA evaluates do not depend on any function value: x0 and x0 +- rho e_i, shifted where a bound is
// nearer than rho (external/nlopt/src/algs/bobyqa/bobyqa.c: the bound preparation of bobyqa() :3205-3257, prelim_ :1822-1868,
// NLopt's rescaling to equal initial steps :3093-3117 with src/util/rescale.c:29-71; reached from optimize.c:692-706 with
// npt = 2n+1). This restates that arithmetic operation for operation -- same IEEE doubles, this file is built without FMA
// contraction -- so that the points can be evaluated as ONE batch on the GPU before the solver starts.
// A prediction that is off in a single bit simply never matches the solver's x and costs nothing but the unused evaluation.
std::vector<VecN> PredictBobyqaStart(const VecN& x0, const VecN& lb, const VecN& ub, const VecN& dx)
{
    const size_t n = x0.size();
    std::vector<VecN> pts;
    if (n < 2) return pts;
    VecN s(n, 1.0);
    {
        size_t i = 1;
        for (; i < n && dx[i] == dx[i - 1]; ++i) {}
        if (i < n) for (i = 1; i < n; ++i) s[i] = dx[i] / dx[0];
    }
    VecN x(n), xl(n), xu(n), sl(n), su(n);
    for (size_t j = 0; j < n; ++j) {
        if (s[j] == 0.0 || !std::isfinite(s[j])) return pts;
        x[j] = x0[j] / s[j]; xl[j] = lb[j] / s[j]; xu[j] = ub[j] / s[j];
        if (xl[j] > xu[j]) std::swap(xl[j], xu[j]);
    }
    const double rhobeg = std::fabs(dx[0] / s[0]);
    for (size_t j = 0; j < n; ++j) {
        const double temp = xu[j] - xl[j];
        if (temp < rhobeg + rhobeg) return pts;   // BOBYQA refuses this itself
        sl[j] = xl[j] - x[j];
        su[j] = xu[j] - x[j];
        if (sl[j] >= -rhobeg) {
            if (sl[j] >= 0.0) { x[j] = xl[j]; sl[j] = 0.0; su[j] = temp; }
            else { x[j] = xl[j] + rhobeg; sl[j] = -rhobeg; su[j] = std::max(xu[j] - x[j], rhobeg); }
        } else if (su[j] <= rhobeg) {
            if (su[j] <= 0.0) { x[j] = xu[j]; sl[j] = -temp; su[j] = 0.0; }
            else { x[j] = xu[j] - rhobeg; sl[j] = std::min(xl[j] - x[j], -rhobeg); su[j] = rhobeg; }
        }
    }
    const VecN xbase = x;
    VecN stepa_of(n, 0.0);
    for (size_t nf = 1; nf <= 2 * n + 1; ++nf) {
        VecN xpt(n, 0.0);
        const size_t nfm = nf - 1;
        if (nfm >= 1 && nfm <= n) {
            double stepa = rhobeg;
            if (su[nfm - 1] == 0.0) stepa = -stepa;
            xpt[nfm - 1] = stepa;
            stepa_of[nfm - 1] = stepa;
        } else if (nfm > n) {
            const size_t nfx = nfm - n;
            double stepb = -rhobeg;
            if (sl[nfx - 1] == 0.0) stepb = std::min(2.0 * rhobeg, su[nfx - 1]);
            if (su[nfx - 1] == 0.0) stepb = std::max(-2.0 * rhobeg, sl[nfx - 1]);
            xpt[nfx - 1] = stepb;
        }
        VecN p(n);
        for (size_t j = 0; j < n; ++j) {
            double v = std::min(std::max(xl[j], xbase[j] + xpt[j]), xu[j]);
            if (xpt[j] == sl[j]) v = xl[j];
            if (xpt[j] == su[j]) v = xu[j];
            p[j] = v * s[j];   // nlopt_unscale: what the callback receives
        }
        pts.push_back(p);
    }
    return pts;
}

void Core::PerformOptimization(bool explicit_termination) { Solve(explicit_termination, "LN_BOBYQA", false); }

void Core::PerformOptimizationWithGradient(bool explicit_termination, const std::string& algorithm, bool analytic)
{
    if (algorithm.size() < 2 || algorithm[1] != 'D') throw std::invalid_argument("PerformOptimizationWithGradient needs a gradient-based (LD_ / GD_) NLopt algorithm");
    Solve(explicit_termination, algorithm.c_str(), true, analytic);
}

void Core::Solve(bool explicit_termination, const char* algorithm, bool want_gradient, bool analytic)
{
    stop_optimization_ = false;
    constexpr int dim = 5;
    constexpr double epsilon = 1e-06;

    VecN x_initial;
    { std::lock_guard<std::recursive_mutex> eg(engine_mutex_); x_initial = object_->GetDescriptor(); }   // UI-side queries read the object too
    const int n = static_cast<int>(x_initial.size()) / dim;
    const VecN x_original = RetrieveOriginalParameters();
    if (x_initial.size() != x_original.size()) throw std::logic_error("x_original does not match the descriptor");
    if (freezed_dofs_.size() != x_initial.size()) throw std::logic_error("freezed_dofs_ does not match the descriptor");

    // Bounds (core_opt.cpp:82-91) and clamping of the start point (:94-97)
    VecN upper(x_initial), lower(x_initial);
    for (double& v : upper) v += 6.0;
    for (double& v : lower) v -= 6.0;
    for (int i = 0; i < n; ++i) {
        upper[i * dim + 2] = +M_PI * 0.4;
        lower[i * dim + 2] = -M_PI * 0.4;
        // a central difference evaluates x +- h e_k: keep the handles on their side of zero there too
        upper[i * dim + 3] = -(epsilon + (want_gradient ? kGradientStep : 0.0));
        lower[i * dim + 4] = +(epsilon + (want_gradient ? kGradientStep : 0.0));
    }
    for (size_t i = 0; i < x_initial.size(); ++i) x_initial[i] = std::max(std::min(x_initial[i], upper[i]), lower[i]);

    // select / revert (core_opt.cpp:99-121) as an index table
    EvaluateObjectiveData data;
    data.core = this;
    data.min_t = min_frame_ - 2.0;
    data.max_t = max_frame_ + 2.0;
    data.x_original = x_original;
    data.x_initial = x_initial;
    data.want_gradient = want_gradient;
    data.analytic_gradient = analytic;
    data.gradient_step = kGradientStep;
    for (size_t i = 0; i < freezed_dofs_.size(); ++i) if (!freezed_dofs_[i]) data.free_indices.push_back(static_cast<int>(i));
    const unsigned M = static_cast<unsigned>(data.free_indices.size());
    auto sel = [&](const VecN& v) { VecN r(M); for (unsigned k = 0; k < M; ++k) r[k] = v[data.free_indices[k]]; return r; };
    auto rev = [&](const VecN& v) { VecN r = x_initial; for (unsigned k = 0; k < M; ++k) r[data.free_indices[k]] = v[k]; return r; };

    const int max_evaluations = explicit_termination ? 20000 : 2000;
    const double relative_function_tolerance = explicit_termination ? 1e-24 : 1e-07;
    const double relative_parameter_tolerance = explicit_termination ? 1e-24 : 1e-07;
    constexpr double initial_step_scale = 0.06;

    last_solve_evals_ = 0;
    last_solve_status_ = "";
    solve_trace_.clear();
    solve_trace_dim_ = static_cast<int>(M);
    if (M == 0) {
        std::lock_guard<std::recursive_mutex> eg(engine_mutex_);
        object_->SetDescriptor(x_initial); last_solve_status_ = "no free dofs"; return;
    }

    // nloptutil::solve (nlopt-util.hpp:44-195)
    NloptApi& api = nlopt_api();
    const int alg = api.algorithm_from_string(algorithm);
    if (alg < 0) throw std::invalid_argument(std::string("unknown NLopt algorithm ") + algorithm);
    nlopt_opt solver = api.create(alg, M);
    if (!solver) throw std::runtime_error("nlopt_create failed");
    const VecN u = sel(upper), l = sel(lower);
    api.set_upper_bounds(solver, u.data());
    api.set_lower_bounds(solver, l.data());
    api.set_maxeval(solver, max_evaluations);
    api.set_ftol_rel(solver, relative_function_tolerance);
    api.set_xtol_rel(solver, relative_parameter_tolerance);
    Trampoline tr{&data, solver, &api, {}, {}, ""};
    api.set_min_objective(solver, trampoline, &tr);
    VecN x_star = sel(x_initial);

    // verbose banner: one direct evaluation with a dummy gradient (nlopt-util.hpp:102-106); want_gradient is
    // false for LN_BOBYQA so no finite-difference batch is triggered.
    double initial_cost_value = 0.0;
    bool cancelled = false;
    try {
        std::vector<double> dummy;   // empty: the banner wants f only, whatever the algorithm
        initial_cost_value = EvaluateObjective(x_star, dummy, &data);
    } catch (const forced_stop&) {
        cancelled = true;
    }
    VecN step(M, 0.0);
    api.get_initial_step(solver, x_star.data(), step.data());
    for (double& d : step) d *= initial_step_scale;
    api.set_initial_step(solver, step.data());

    // Batched start: BOBYQA's first 2M+1 interpolation points in one GPU batch; the callback then serves
    // them from the cache (bitwise match on x) while the solver walks through its serial initialisation.
    data.start_cache_x.clear(); data.start_cache_f.clear(); data.start_cache_hits = 0;
    if (!cancelled && !want_gradient && batched_start_ && std::string(algorithm) == "LN_BOBYQA") {
        const std::vector<VecN> pts = PredictBobyqaStart(x_star, l, u, step);
        if (!pts.empty()) {
            std::vector<double> X(pts.size() * x_initial.size());
            for (size_t b = 0; b < pts.size(); ++b) {
                const VecN full = rev(pts[b]);
                std::copy(full.begin(), full.end(), X.begin() + b * x_initial.size());
            }
            std::vector<double> fb(pts.size());
            {
                std::lock_guard<std::recursive_mutex> eg(engine_mutex_);
                // the scene-specialised kernel, as for the single evaluations it stands in for: bit-identical values
                omo_set_batch_hint(Engine(), 1 << 30);
                try { CalculateFullObjectiveBatch(data.min_t, data.max_t, X.data(), static_cast<int>(pts.size()), x_original, fb.data()); }
                catch (...) { omo_set_batch_hint(Engine(), 0); throw; }
                omo_set_batch_hint(Engine(), 0);
            }
            data.start_cache_x = pts;
            data.start_cache_f = fb;
        }
    }

    const auto t_start = std::chrono::steady_clock::now();
    double final_cost_value = initial_cost_value;
    int result = 0;
    if (!cancelled) result = api.optimize(solver, x_star.data(), &final_cost_value);
    const auto t_end = std::chrono::steady_clock::now();
    const int numevals = api.get_numevals(solver);
    api.destroy(solver);

    // result < 0: ROUNDOFF_LIMITED (-4) and FORCED_STOP (-5) keep the best x (nlopt-util.hpp:129-148); any other
    // failure returns x_initial (:149-156). An engine error inside the callback is re-raised.
    if (!tr.error.empty() && tr.error != "forced_stop") throw engine_error("objective callback failed: " + tr.error);
    if (result < 0 && result != -4 && result != -5) x_star = sel(x_initial);
    last_solve_value_ = final_cost_value;
    last_solve_evals_ = numevals;
    last_solve_batched_ = data.start_cache_hits;
    last_solve_status_ = cancelled ? "forced_stop" : std::to_string(result);

    const double ms = std::chrono::duration<double, std::milli>(t_end - t_start).count();
    std::cout << "---- nlopt-util ----" << std::endl;
    std::cout << "Dimensions     : " << M << std::endl;
    std::cout << "Function value : " << initial_cost_value << " => " << final_cost_value << std::endl;
    std::cout << "Elapsed time   : " << ms << " [ms]" << std::endl;
    std::cout << "Function evals : " << numevals << std::endl;
    std::cout << "--------------------" << std::endl;

    { std::lock_guard<std::recursive_mutex> eg(engine_mutex_); object_->SetDescriptor(rev(x_star)); }
}

}  // namespace optimo
