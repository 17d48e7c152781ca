// omo_multi -- one host thread, several B200s (include/omo.h, "multi-device engine").
//
// The reference's callback (EvaluateObjective, OptiMo/core_opt.cpp:42-57) runs on ONE thread of ONE process and its only
// parallel construct is parallelutil::parallel_for inside Core::SetStates (OptiMo/core_rbdl.cpp:157-161). An embedder that
// keeps that shape cannot drive eight GPUs without writing the plumbing itself, so the plumbing lives here, behind the C ABI:
// an omo_multi owns one omo_ctx per device, one persistent worker thread per device (a ctx is single-threaded by contract),
// the pinned gather buffer and -- when the gather goes through NVLink -- the NCCL communicators.
//
// Sharding: candidates of a batch / columns of a finite-difference gradient are split into
// contiguous slices by omo_shard_range; scene tables, weights and x_original are replicated; the only exchange is the gather
// of one double per candidate / column. Two gathers with identical results:
//   OMO_GATHER_HOST  every device copies its slice straight into the caller's (or the pinned) host array -- no inter-GPU
//                    traffic at all; the consumer (NLopt) lives on the host anyway. Default.
//   OMO_GATHER_NCCL  ncclAllGather of the (padded) slices over NVLink / NVSwitch, then one D2H from the first device.
// NCCL is dlopen'ed ("libnccl.so.2"; OMO_NCCL_LIB overrides): libomo.so does not link against it.
#include <cuda_runtime.h>
#include <dlfcn.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "../../include/omo.h"

namespace {

std::string g_multi_create_error;

// ---- NCCL entry points (only what the gather needs) ------------------------------------------------------------------
typedef struct ncclComm* ncclComm_t;
struct NcclApi {
    void* handle = nullptr;
    std::string error;
    int (*ncclCommInitAll)(ncclComm_t*, int, const int*) = nullptr;
    int (*ncclCommDestroy)(ncclComm_t) = nullptr;
    int (*ncclAllGather)(const void*, void*, size_t, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*ncclGroupStart)() = nullptr;
    int (*ncclGroupEnd)() = nullptr;
    const char* (*ncclGetErrorString)(int) = nullptr;
    int (*ncclGetVersion)(int*) = nullptr;
};
constexpr int kNcclFloat64 = 8;   // ncclDataType_t::ncclFloat64 (nccl.h)

NcclApi& nccl_api()
{
    static NcclApi a;
    static std::once_flag once;
    std::call_once(once, [] {
        std::vector<std::string> names;
        if (const char* e = std::getenv("OMO_NCCL_LIB")) names.push_back(e);
        names.push_back("libnccl.so.2");
        names.push_back("libnccl.so");
        for (const std::string& n : names) {
            a.handle = dlopen(n.c_str(), RTLD_NOW | RTLD_LOCAL);
            if (a.handle) break;
        }
        if (!a.handle) { a.error = "libnccl.so.2 is not loadable (set OMO_NCCL_LIB)"; return; }
#define OMO_NCCL_SYM(f)                                                       \
    a.f = reinterpret_cast<decltype(a.f)>(dlsym(a.handle, #f));               \
    if (!a.f) { a.error += std::string(" missing ") + #f; }
        OMO_NCCL_SYM(ncclCommInitAll) OMO_NCCL_SYM(ncclCommDestroy) OMO_NCCL_SYM(ncclAllGather) OMO_NCCL_SYM(ncclGroupStart)
        OMO_NCCL_SYM(ncclGroupEnd) OMO_NCCL_SYM(ncclGetErrorString) OMO_NCCL_SYM(ncclGetVersion)
#undef OMO_NCCL_SYM
        if (!a.error.empty()) a.handle = nullptr;
    });
    return a;
}

// ---- one worker thread per device -------------------------------------------------------------------------------------
// Jobs are posted by the (single) calling thread and joined before the entry point returns. A worker spins briefly on its
// job counter before it blocks: a gradient call is ~100 us of GPU work and a futex wake-up costs a fifth of that.
struct Worker {
    std::thread th;
    std::mutex mu;
    std::condition_variable cv;
    std::atomic<unsigned long long> posted{0}, done{0};
    std::function<omo_status()> job;
    omo_status status = OMO_OK;
    bool quit = false;
};

}  // namespace

struct omo_multi {
    int n = 0;
    std::vector<int> dev;
    std::vector<omo_ctx*> ctx;
    std::vector<Worker*> workers;
    std::string err;
    int gather = OMO_GATHER_HOST;
    std::vector<ncclComm_t> comms;
    // gradient scratch: per device the slice it evaluates (padded to the largest slice) and, for NCCL, the gathered columns
    std::vector<double*> d_slice, d_all;
    int slice_cap = 0;
    double* h_cols = nullptr;   // pinned [n * slice_cap]
    int h_cols_cap = 0;
    std::vector<double> cols;   // unpadded columns of the last gradient
    unsigned long long jobs_run = 0;
};

namespace {

void worker_main(Worker* w)
{
    unsigned long long seen = 0;
    for (;;) {
        // spin, then sleep
        const auto t0 = std::chrono::steady_clock::now();
        while (w->posted.load(std::memory_order_acquire) == seen) {
            if (std::chrono::steady_clock::now() - t0 > std::chrono::microseconds(200)) {
                std::unique_lock<std::mutex> lk(w->mu);
                w->cv.wait(lk, [&] { return w->posted.load(std::memory_order_acquire) != seen || w->quit; });
                break;
            }
        }
        if (w->quit && w->posted.load(std::memory_order_acquire) == seen) return;
        seen = w->posted.load(std::memory_order_acquire);
        w->status = w->job ? w->job() : OMO_OK;
        w->done.store(seen, std::memory_order_release);
    }
}

void post(Worker* w, std::function<omo_status()> job)
{
    {
        std::lock_guard<std::mutex> lk(w->mu);
        w->job = std::move(job);
        w->posted.fetch_add(1, std::memory_order_release);
    }
    w->cv.notify_one();
}

omo_status join(Worker* w)
{
    const unsigned long long want = w->posted.load(std::memory_order_acquire);
    while (w->done.load(std::memory_order_acquire) != want) std::this_thread::yield();
    return w->status;
}

// Runs job(i) on every device's worker; the first failure wins and its message is copied from that device's context.
omo_status run_all(omo_multi* m, const std::function<omo_status(int)>& job)
{
    for (int i = 0; i < m->n; ++i) post(m->workers[i], [i, &job] { return job(i); });
    omo_status st = OMO_OK;
    for (int i = 0; i < m->n; ++i) {
        const omo_status s = join(m->workers[i]);
        if (s != OMO_OK && st == OMO_OK) {
            st = s;
            m->err = "device " + std::to_string(m->dev[i]) + ": " + (m->ctx[i] ? omo_last_error(m->ctx[i]) : "no context");
        }
    }
    m->jobs_run += 1;
    return st;
}

omo_status mfail(omo_multi* m, omo_status st, const std::string& msg)
{
    if (m) m->err = msg; else g_multi_create_error = msg;
    return st;
}

omo_status ensure_grad_buffers(omo_multi* m, int max_count)
{
    if (max_count <= m->slice_cap) return OMO_OK;
    const int cap = std::max(max_count, 64);
    omo_status st = run_all(m, [&](int i) -> omo_status {
        if (m->d_slice[i]) omo_device_free(m->ctx[i], m->d_slice[i]);
        if (m->d_all[i]) omo_device_free(m->ctx[i], m->d_all[i]);
        m->d_slice[i] = m->d_all[i] = nullptr;
        void* p = nullptr;
        omo_status s = omo_device_alloc(m->ctx[i], (unsigned long long)cap * sizeof(double), &p);
        if (s != OMO_OK) return s;
        m->d_slice[i] = static_cast<double*>(p);
        s = omo_device_alloc(m->ctx[i], (unsigned long long)cap * m->n * sizeof(double), &p);
        if (s != OMO_OK) return s;
        m->d_all[i] = static_cast<double*>(p);
        return OMO_OK;
    });
    if (st != OMO_OK) return st;
    if (m->h_cols) { omo_host_free_pinned(m->ctx[0], m->h_cols); m->h_cols = nullptr; }
    void* hp = nullptr;
    if ((st = omo_host_alloc_pinned(m->ctx[0], (unsigned long long)cap * m->n * sizeof(double), &hp)) != OMO_OK)
        return mfail(m, st, std::string("pinned gather buffer: ") + omo_last_error(m->ctx[0]));
    m->h_cols = static_cast<double*>(hp);
    m->slice_cap = cap;
    return OMO_OK;
}

}  // namespace

extern "C" {

const char* omo_multi_last_error(const omo_multi* m) { return m ? m->err.c_str() : g_multi_create_error.c_str(); }

void omo_destroy_multi(omo_multi* m)
{
    if (!m) return;
    if (!m->workers.empty()) {
        for (int i = 0; i < m->n; ++i) {
            Worker* w = m->workers[i];
            if (!w) continue;
            post(w, [m, i]() -> omo_status {
                if (m->ctx[i]) {
                    if (m->d_slice[i]) omo_device_free(m->ctx[i], m->d_slice[i]);
                    if (m->d_all[i]) omo_device_free(m->ctx[i], m->d_all[i]);
                    if (i == 0 && m->h_cols) omo_host_free_pinned(m->ctx[0], m->h_cols);
                }
                return OMO_OK;
            });
            join(w);
        }
    }
    if (!m->comms.empty()) {
        NcclApi& a = nccl_api();
        for (ncclComm_t c : m->comms) if (c && a.ncclCommDestroy) a.ncclCommDestroy(c);
    }
    for (int i = 0; i < (int)m->workers.size(); ++i) {
        Worker* w = m->workers[i];
        if (!w) continue;
        { std::lock_guard<std::mutex> lk(w->mu); w->quit = true; }
        w->cv.notify_one();
        if (w->th.joinable()) w->th.join();
        delete w;
    }
    for (omo_ctx* c : m->ctx) if (c) omo_destroy(c);
    delete m;
}

omo_status omo_create_multi(const omo_scene_desc* scene, const int* device_ids, int n_dev, omo_multi** out)
{
    if (!scene || !out) return mfail(nullptr, OMO_BAD_ARG, "null argument");
    *out = nullptr;
    int visible = 0;
    const cudaError_t e = cudaGetDeviceCount(&visible);
    if (e != cudaSuccess || visible == 0)
        return mfail(nullptr, OMO_CUDA_ERROR, std::string("no usable CUDA device (there is no CPU fallback): ") + cudaGetErrorString(e));
    if (n_dev <= 0) { n_dev = visible; device_ids = nullptr; }
    if (n_dev > 64) return mfail(nullptr, OMO_BAD_ARG, "at most 64 devices");
    omo_multi* m = new omo_multi();
    m->n = n_dev;
    for (int i = 0; i < n_dev; ++i) {
        const int d = device_ids ? device_ids[i] : i;
        if (d < 0 || d >= visible || std::find(m->dev.begin(), m->dev.end(), d) != m->dev.end()) {
            delete m;
            return mfail(nullptr, OMO_BAD_ARG, "device_ids must be distinct indices of visible devices");
        }
        m->dev.push_back(d);
    }
    m->ctx.assign(n_dev, nullptr);
    m->d_slice.assign(n_dev, nullptr);
    m->d_all.assign(n_dev, nullptr);
    for (int i = 0; i < n_dev; ++i) {
        Worker* w = new Worker();
        m->workers.push_back(w);
        w->th = std::thread(worker_main, w);
    }
    std::vector<std::string> errs(n_dev);
    const omo_status st = run_all(m, [&](int i) -> omo_status {
        const omo_status s = omo_create(scene, m->dev[i], &m->ctx[i]);
        if (s != OMO_OK) errs[i] = omo_last_error(nullptr);
        return s;
    });
    if (st != OMO_OK) {
        for (int i = 0; i < n_dev; ++i) if (!errs[i].empty()) { g_multi_create_error = "device " + std::to_string(m->dev[i]) + ": " + errs[i]; break; }
        omo_destroy_multi(m);
        return st;
    }
    if (const char* g = std::getenv("OMO_GATHER")) {
        if (std::strcmp(g, "nccl") == 0 && omo_multi_set_gather(m, OMO_GATHER_NCCL) != OMO_OK) {
            g_multi_create_error = m->err;
            omo_destroy_multi(m);
            return OMO_NCCL_ERROR;
        }
    }
    *out = m;
    return OMO_OK;
}

int omo_multi_num_devices(const omo_multi* m) { return m ? m->n : 0; }
int omo_multi_device(const omo_multi* m, int i) { return (m && i >= 0 && i < m->n) ? m->dev[i] : -1; }
omo_ctx* omo_multi_ctx(omo_multi* m, int i) { return (m && i >= 0 && i < m->n) ? m->ctx[i] : nullptr; }

omo_status omo_multi_set_gather(omo_multi* m, int mode)
{
    if (!m || (mode != OMO_GATHER_HOST && mode != OMO_GATHER_NCCL)) return mfail(m, OMO_BAD_ARG, "gather mode must be OMO_GATHER_HOST or OMO_GATHER_NCCL");
    if (mode == OMO_GATHER_NCCL && m->comms.empty()) {
        NcclApi& a = nccl_api();
        if (!a.handle) return mfail(m, OMO_NCCL_ERROR, "NCCL unavailable: " + a.error);
        m->comms.assign(m->n, nullptr);
        const int rc = a.ncclCommInitAll(m->comms.data(), m->n, m->dev.data());
        if (rc != 0) {
            m->comms.clear();
            return mfail(m, OMO_NCCL_ERROR, std::string("ncclCommInitAll: ") + a.ncclGetErrorString(rc));
        }
    }
    m->gather = mode;
    return OMO_OK;
}

int omo_multi_get_gather(const omo_multi* m) { return m ? m->gather : -1; }

omo_status omo_multi_set_window(omo_multi* m, double min_t, double max_t, double dt)
{
    if (!m) return OMO_BAD_ARG;
    return run_all(m, [&](int i) { return omo_set_window(m->ctx[i], min_t, max_t, dt); });
}

omo_status omo_multi_set_weights(omo_multi* m, const double* w_param, int D, double w_traj, const omo_gauss* kernels, int n_kernels, int cyclic)
{
    if (!m) return OMO_BAD_ARG;
    return run_all(m, [&](int i) { return omo_set_weights(m->ctx[i], w_param, D, w_traj, kernels, n_kernels, cyclic); });
}

omo_status omo_multi_set_original(omo_multi* m, const double* x_original, int D)
{
    if (!m) return OMO_BAD_ARG;
    return run_all(m, [&](int i) { return omo_set_original(m->ctx[i], x_original, D); });
}

omo_status omo_multi_eval(omo_multi* m, const double* X, int B, double* f_out, omo_terms* terms_out)
{
    if (!m || !f_out || B < 0) return mfail(m, OMO_BAD_ARG, "bad argument");
    const int D = omo_descriptor_size(m->ctx[0]);
    if (B > 0 && D > 0 && !X) return mfail(m, OMO_BAD_ARG, "null X");
    return run_all(m, [&](int i) -> omo_status {
        int b0 = 0, nb = 0;
        omo_shard_range(B, i, m->n, &b0, &nb);
        if (nb == 0) return OMO_OK;
        omo_set_batch_hint(m->ctx[i], B);   // the kernel the whole batch would get, whatever this device's share
        const omo_status s = omo_eval(m->ctx[i], X + (size_t)b0 * D, nb, f_out + b0, terms_out ? terms_out + b0 : nullptr);
        omo_set_batch_hint(m->ctx[i], 0);
        return s;
    });
}

omo_status omo_multi_eval_free(omo_multi* m, const double* x_base, const int* free_idx, int F, const double* Xfree, int B, double* f_out,
                               omo_terms* terms_out)
{
    if (!m || !f_out || B < 0 || F < 0) return mfail(m, OMO_BAD_ARG, "bad argument");
    return run_all(m, [&](int i) -> omo_status {
        int b0 = 0, nb = 0;
        omo_shard_range(B, i, m->n, &b0, &nb);
        if (nb == 0) return OMO_OK;
        omo_set_batch_hint(m->ctx[i], B);
        const omo_status s = omo_eval_free(m->ctx[i], x_base, free_idx, F, Xfree + (size_t)b0 * F, nb, f_out + b0, terms_out ? terms_out + b0 : nullptr);
        omo_set_batch_hint(m->ctx[i], 0);
        return s;
    });
}

omo_status omo_multi_eval_grad(omo_multi* m, const double* x, const int* free_idx, int F, double h, double* f0_out, double* grad_out, double* fpm_out)
{
    if (!m || F < 0 || (F > 0 && (!free_idx || !grad_out))) return mfail(m, OMO_BAD_ARG, "bad argument");
    const int P = 2 * F + 1;
    const int max_count = (P + m->n - 1) / m->n;
    omo_status st = ensure_grad_buffers(m, max_count);
    if (st != OMO_OK) return st;
    const bool nccl = m->gather == OMO_GATHER_NCCL && m->n > 1;
    NcclApi& a = nccl_api();
    std::vector<int> nccl_rc(m->n, 0);
    st = run_all(m, [&](int i) -> omo_status {
        int c0 = 0, nc = 0;
        omo_shard_range(P, i, m->n, &c0, &nc);
        omo_status s = OMO_OK;
        if (nc > 0) s = omo_eval_grad_columns(m->ctx[i], x, free_idx, F, h, c0, nc, m->d_slice[i], 0);
        if (s != OMO_OK) return s;
        if (!nccl) {
            // every device writes its slice of the column array on the host itself: the gather is the copy
            if (nc > 0) s = omo_memcpy_d2h(m->ctx[i], m->h_cols + c0, m->d_slice[i], (unsigned long long)nc * sizeof(double));
            return s;
        }
        // equal counts per rank: slices are padded to the largest one (the padding carries whatever the buffer held)
        nccl_rc[i] = a.ncclAllGather(m->d_slice[i], m->d_all[i], (size_t)max_count, kNcclFloat64, m->comms[i], static_cast<cudaStream_t>(omo_stream(m->ctx[i])));
        if (nccl_rc[i] != 0) return OMO_NCCL_ERROR;
        if (i == 0) return omo_memcpy_d2h(m->ctx[0], m->h_cols, m->d_all[0], (unsigned long long)max_count * m->n * sizeof(double));
        return omo_synchronize(m->ctx[i]);
    });
    if (st == OMO_NCCL_ERROR) {
        for (int i = 0; i < m->n; ++i) if (nccl_rc[i] != 0) return mfail(m, OMO_NCCL_ERROR, std::string("ncclAllGather: ") + a.ncclGetErrorString(nccl_rc[i]));
    }
    if (st != OMO_OK) return st;
    m->cols.resize(P);
    if (!nccl) std::copy(m->h_cols, m->h_cols + P, m->cols.begin());
    else
        for (int i = 0; i < m->n; ++i) {
            int c0 = 0, nc = 0;
            omo_shard_range(P, i, m->n, &c0, &nc);
            std::copy(m->h_cols + (size_t)i * max_count, m->h_cols + (size_t)i * max_count + nc, m->cols.begin() + c0);
        }
    if (fpm_out) std::copy(m->cols.begin() + 1, m->cols.end(), fpm_out);
    return omo_grad_from_columns(m->cols.data(), F, h, f0_out, grad_out);
}

}  // extern "C"
