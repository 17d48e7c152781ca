/*
 * optimo_core.h -- C view of the headless `Core` (optimo_b200/host/core.hpp, liboptimo_core.so): the
 * reference-facing operator API of the objective path, for hosts that are not C++ (ctypes, FFI).
 * Each function mirrors the Core member of the same name in the reference (OptiMo/core.h:35-245,
 * OptiMo/core_opt.cpp:42-157); `h` is a Core*. Functions returning int give 0 on success, 1 on error
 * (message: optimo_last_error()), 2 when the evaluation was cancelled through stop_optimization_
 * (the reference throws nlopt::forced_stop there, core_opt.cpp:50-54).
 *
 * The arithmetic runs on the GPU behind include/omo.h; there is no CPU fallback.
 */
#ifndef OPTIMO_CORE_H
#define OPTIMO_CORE_H

#ifdef __cplusplus
extern "C" {
#endif

const char* optimo_last_error(void);

void* optimo_core_new(void);          /* Core::Core(): default humanoid (core.cpp:23-103) */
void* optimo_core_instance(void);     /* Core::GetInstance() (core.h:40-44) */
void optimo_core_free(void* h);

int optimo_read_json(void* h, const char* path);          /* Core::ReadJson, core_io.cpp:154-163 */
int optimo_read_json_string(void* h, const char* text);   /* Core::ApplyJsonToScene, core_io.cpp:228-250 */
int optimo_write_json(void* h, const char* path);         /* Core::WriteJson, core_io.cpp:165-169 */
int optimo_scene_json(void* h, char* out, int cap);       /* ConvertCurrentSceneToJson; returns the length */
int optimo_set_device(void* h, int device);
/* Several GPUs of one box behind the same calls: batches are sharded over candidates, gradients over their finite-difference
 * columns (include/omo.h: omo_multi). n <= 1 goes back to one device. */
int optimo_set_devices(void* h, const int* devices, int n);

int optimo_min_frame(void* h);
int optimo_max_frame(void* h);
void optimo_set_frames(void* h, int min_frame, int max_frame);
int optimo_num_joints(void* h);
int optimo_num_items(void* h);
int optimo_item_name(void* h, int index, char* out, int cap);
int optimo_descriptor_size(void* h);                      /* Object::GetDescriptor().rows(), object.cpp:110-145 */
int optimo_get_descriptor(void* h, double* x);
int optimo_set_descriptor(void* h, const double* x, int n);  /* Object::SetDescriptor, object.cpp:147-172 */
void optimo_set_cyclic(void* h, int cyclic);              /* Object::is_cyclic_ */

int optimo_update_original_parameters(void* h, const double* x, int n);  /* core.cpp:184-189 */
int optimo_retrieve_original_parameters(void* h, double* x);
void optimo_update_global_weight(void* h, double w);                     /* core.cpp:150-153 */
int optimo_update_per_item_weight(void* h, double w, const char* item);  /* core.cpp:155-159 */
void optimo_clear_kernels(void* h);
void optimo_add_kernel(void* h, double amp, double mu, double sigma);    /* AddKernel + GetLastKernelPointer, core.h:163-170 */
double optimo_get_cost_weight(void* h, double t);                        /* core.cpp:191-199 */
int optimo_parameter_weights(void* h, double* w);                        /* core.cpp:161-182 */
int optimo_set_selected(void* h, const char* item, int selected);        /* Item::is_selected_ */
void optimo_update_freezed_dofs_from_selection(void* h);                 /* core.cpp:118-148 */
int optimo_get_freezed_dofs(void* h, int* out);
int optimo_set_freezed_dofs(void* h, const int* in, int n);
int optimo_joint_physical(void* h, int joint, double* out13);            /* mass, com[3], inertia[9]; joint.cpp:90-119 */

/* Core::CalculateFullObjective, core_rbdl.cpp:348-401. terms3 (optional) = C, D_param, D_traj. */
int optimo_calculate_full_objective(void* h, double min_t, double max_t, const double* x, const double* x_original, int n, double* f,
                                    double* terms3);
/* Batch of B descriptors (row-major X[B][n]). terms3 optional, [B][3]. */
int optimo_calculate_full_objective_batch(void* h, double min_t, double max_t, const double* X, int B, const double* x_original, int n,
                                          double* f, double* terms3);
/* NEW: central-difference gradient on free_idx. fpm optional, [F][2] = f+, f-. */
int optimo_calculate_gradient(void* h, double min_t, double max_t, const double* x, const double* x_original, int n, const int* free_idx,
                              int F, double step, double* f0, double* grad, double* fpm);
int optimo_positions(void* h, double min_t, double max_t, const double* x, int n, double* p);  /* SetStates' p, core_rbdl.cpp:150-153 */
int optimo_num_samples(void* h, double min_t, double max_t);                                   /* core_rbdl.cpp:120-123 */

/* EvaluateObjective as NLopt calls it (core_opt.cpp:42-57): x over the free dofs, frozen ones from x_initial. */
int optimo_evaluate_objective(void* h, const double* x_free, int n_free, const double* x_initial, int n_full, double min_t, double max_t,
                              int want_gradient, double step, double* f, double* grad);
void optimo_set_stop_optimization(void* h, int flag);     /* Core::stop_optimization_ */
int optimo_get_stop_optimization(void* h);
int optimo_perform_optimization(void* h, int explicit_termination, double* final_value, int* evals);  /* core_opt.cpp:60-157 */
/* The same solve with a gradient-based NLopt algorithm ("LD_LBFGS", "LD_SLSQP", "LD_MMA", "LD_TNEWTON" ...; NULL = LD_LBFGS), as
 * nloptutil::unconstrained::grad_based::bounded::solve sets one up (external/nlopt-util/include/nlopt-util.hpp:226-274): every
 * callback returns f and the central-difference gradient over the free dofs from one GPU batch. */
/* Batched BOBYQA start: the solver's first 2M+1 interpolation points -- which do not depend on any function value
 * (external/nlopt/src/algs/bobyqa/bobyqa.c prelim_) -- are predicted, evaluated as ONE GPU batch and served to the callback from a cache
 * keyed on the bits of x. Default on; the iterates are identical either way. optimo_last_solve_batched: callbacks served that way. */
void optimo_set_batched_start(void* h, int on);
int optimo_last_solve_batched(void* h);
/* algorithm: any gradient-based NLopt algorithm name; a "+analytic" suffix ("LD_SLSQP+analytic") takes the analytic gradient (N4) */
int optimo_perform_optimization_gradient(void* h, int explicit_termination, const char* algorithm, double* final_value, int* evals);

/* Host-only (no GPU): generate + NVRTC-compile the scene-specialised objective kernel of the current scene
 * (include/omo.h: omo_jit_probe). Returns 0 when it compiled; `log` = stack placement summary and compiler log. */
/* Next row N4: f and its analytic gradient (include/omo.h: omo_eval_grad_analytic); *analytic_used = 0 when the scene / x is outside its
 * coverage and the finite-difference gradient was returned instead. */
int optimo_calculate_analytic_gradient(void* h, double min_t, double max_t, const double* x, const double* x_original, int n, const int* free_idx, int F,
                                       double* f0, double* grad, int* analytic_used);
int optimo_debug_scene_table(void* h, void* out, unsigned long long cap, int* sub_end_out);   /* host-only, see omo_debug_scene_table */
int optimo_probe_specialised_kernel(void* h, int fk, int count, const char* dump_prefix, char* log, int cap);

/* Record of the last solve's callbacks in call order (rows x (dim + 1): the free parameters, then f); off by default. */
void optimo_record_solve_trace(void* h, int enable);
int optimo_solve_trace_size(void* h, int* rows, int* dim);
int optimo_get_solve_trace(void* h, double* out);

/* Core::GetOptimizationCache, core.h:134-144 (thread-safe copy). optimo_cache_size takes the copy and keeps it for the calling
 * thread; optimo_get_cache reads arrays of THAT copy, so one read-out never mixes two evaluations. */
int optimo_cache_size(void* h, int* n_samples, int* dof);
int optimo_get_cache(void* h, int which, double* out);    /* 0 time, 1 q, 2 q_dot, 3 q_ddot, 4 tau, 5 local_cost */
void optimo_clear_optimization_cache(void* h);

/* Next rows N2 / N3. Core::CalculateTorques(t) (core_rbdl.cpp:264-289) of the current object for M times
 * in one launch: tau [M][dof]; q, qd, qdd optional (SetCurrentState, core_rbdl.cpp:168-198). */
int optimo_calculate_torques(void* h, const double* times, int M, double* tau, double* q, double* qd, double* qdd);
int optimo_euler_angles(void* h, const double* times, int M, double* out);   /* Joint::GetEulerAngle of every joint, [M][3J] */
int optimo_update_maximum_torque_and_cost(void* h, double* max_torque, double* max_cost);   /* core.cpp:105-116 */
int optimo_export_motion(void* h, const char* path);                  /* Core::ExportMotion, core_io.cpp:252-264 */
int optimo_motion_json(void* h, char* out, int cap);                  /* the text ExportMotion writes; returns the length */
int optimo_push_history(void* h);                                     /* core_io.cpp:188-192 */
int optimo_pop_history(void* h);                                      /* core_io.cpp:194-208 */
void optimo_clear_history(void* h);
int optimo_history_size(void* h);
int optimo_write_history(void* h, const char* path);                  /* core_io.cpp:171-186 */

void* optimo_engine(void* h);                                   /* the omo_ctx* behind this Core (include/omo.h) */
void* optimo_engine_synced(void* h, double min_t, double max_t); /* same, after pushing window/weights/x_original */
void* optimo_engine_multi(void* h);                             /* the omo_multi* behind this Core, NULL unless optimo_set_devices named >= 2 */

#ifdef __cplusplus
}
#endif
#endif /* OPTIMO_CORE_H */
