// A15 host-side Adam for the offloaded SH rows, and the camera-order TSP
// heuristic.  HOST code by design (north star: "host-resident optimizer state
// keeps the reference's cpu-adam as part of the design").  The reference's
// implementations (submodules cpu-adam @ f169c36 and fast-tsp @ 069eb3d) are
// not in the mount; semantics come from the call sites
// (optimizer.py:130-142, engine.py:301-335, engine.py:179) -- parity unpinned.
#include <immintrin.h>
#include <math.h>
#include <stdint.h>
#include <string.h>

#include <pthread.h>
#include <sched.h>
#include <stdlib.h>
#include <time.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <functional>
#include <memory>
#include <mutex>
#include <new>
#include <thread>
#include <vector>

#include "../../include/clmgs.h"

void clmgs_set_error(const char *fmt, ...);
#define CLMGS_EXPORT extern "C" __attribute__((visibility("default")))

namespace {

// Rows of a sparse list are random 192-byte accesses: no hardware prefetcher helps.  Each row is requested
// kPrefetchRows ahead of its use; CLMGS_CPU_ADAM_PREFETCH=<rows>[,<level>] overrides the distance and the cache level
// the lines are requested into (0 = L1, 1 = L2) for tuning on another host.
constexpr int kPrefetchRows = 6;
struct PrefetchCfg { int rows, level; };
inline PrefetchCfg prefetch_cfg() {
    static const PrefetchCfg v = [] {
        PrefetchCfg c{kPrefetchRows, 0};
        if (const char *e = getenv("CLMGS_CPU_ADAM_PREFETCH")) {
            c.rows = std::max(0, atoi(e));
            if (const char *k = strchr(e, ',')) c.level = atoi(k + 1);
        }
        return c;
    }();
    return v;
}

struct AdamConst {
    float b1, b2, eps, scale, inv_bc1, bc2_rsqrt;
};

// One row of `cols` floats.  Two instruction sets, one arithmetic: every operation below is a correctly rounded
// IEEE mul / add / sqrt / div on independent lanes, and the translation unit is built with -ffp-contract=off, so the
// 8-lane and the 16-lane variant give identical bits.
// GMODE: where the gradient row comes from
//   G_TABLE  row of the pinned gradient table; consumed rows are zeroed in place (the reference's "version 3")
//   G_ZERO   the gradient row is known to be zero -- it is neither read nor re-zeroed
//   G_PACKED row of a packed staging buffer (one row per list entry, written by a DMA copy); read once, not zeroed
enum { G_TABLE = 0, G_ZERO = 1, G_PACKED = 2 };
template <int GMODE>
__attribute__((target("avx2"), always_inline)) inline void adam_row_avx2(float *p, float *g, float *m, float *v, const float *col_lr,
                                                         int cols, const AdamConst &c) {
    int k = 0;
    const __m256 b1 = _mm256_set1_ps(c.b1), omb1 = _mm256_set1_ps(1.0f - c.b1);
    const __m256 b2 = _mm256_set1_ps(c.b2), omb2 = _mm256_set1_ps(1.0f - c.b2);
    const __m256 eps = _mm256_set1_ps(c.eps), sc = _mm256_set1_ps(c.scale);
    const __m256 ibc1 = _mm256_set1_ps(-c.inv_bc1), bc2r = _mm256_set1_ps(c.bc2_rsqrt);
    for (; k + 8 <= cols; k += 8) {
        __m256 gg = GMODE == G_ZERO ? _mm256_setzero_ps() : _mm256_mul_ps(_mm256_loadu_ps(g + k), sc);
        __m256 mm = _mm256_add_ps(_mm256_mul_ps(b1, _mm256_loadu_ps(m + k)), _mm256_mul_ps(omb1, gg));
        __m256 vv = _mm256_add_ps(_mm256_mul_ps(b2, _mm256_loadu_ps(v + k)), _mm256_mul_ps(omb2, _mm256_mul_ps(gg, gg)));
        __m256 denom = _mm256_add_ps(_mm256_mul_ps(_mm256_sqrt_ps(vv), bc2r), eps);
        __m256 step = _mm256_mul_ps(_mm256_mul_ps(_mm256_loadu_ps(col_lr + k), ibc1), _mm256_div_ps(mm, denom));
        _mm256_storeu_ps(p + k, _mm256_add_ps(_mm256_loadu_ps(p + k), step));
        _mm256_storeu_ps(m + k, mm);
        _mm256_storeu_ps(v + k, vv);
        if (GMODE == G_TABLE) _mm256_storeu_ps(g + k, _mm256_setzero_ps());  // version 3: zero the consumed gradient in place
    }
    for (; k < cols; ++k) {
        float gg = GMODE == G_ZERO ? 0.f : g[k] * c.scale;
        float mm = c.b1 * m[k] + (1.0f - c.b1) * gg;
        float vv = c.b2 * v[k] + (1.0f - c.b2) * gg * gg;
        float denom = sqrtf(vv) * c.bc2_rsqrt + c.eps;
        p[k] += -(col_lr[k] * c.inv_bc1) * (mm / denom);
        m[k] = mm; v[k] = vv;
        if (GMODE == G_TABLE) g[k] = 0.f;
    }
}

template <int GMODE>
__attribute__((target("avx2,avx512f"), always_inline)) inline void adam_row_avx512(float *p, float *g, float *m, float *v,
                                                              const float *col_lr, int cols, const AdamConst &c) {
    int k = 0;
    const __m512 b1 = _mm512_set1_ps(c.b1), omb1 = _mm512_set1_ps(1.0f - c.b1);
    const __m512 b2 = _mm512_set1_ps(c.b2), omb2 = _mm512_set1_ps(1.0f - c.b2);
    const __m512 eps = _mm512_set1_ps(c.eps), sc = _mm512_set1_ps(c.scale);
    const __m512 ibc1 = _mm512_set1_ps(-c.inv_bc1), bc2r = _mm512_set1_ps(c.bc2_rsqrt);
    for (; k + 16 <= cols; k += 16) {  // a 48-float SH row = 3 x 64-byte lines = 3 iterations
        __m512 gg = GMODE == G_ZERO ? _mm512_setzero_ps() : _mm512_mul_ps(_mm512_loadu_ps(g + k), sc);
        __m512 mm = _mm512_add_ps(_mm512_mul_ps(b1, _mm512_loadu_ps(m + k)), _mm512_mul_ps(omb1, gg));
        __m512 vv = _mm512_add_ps(_mm512_mul_ps(b2, _mm512_loadu_ps(v + k)), _mm512_mul_ps(omb2, _mm512_mul_ps(gg, gg)));
        __m512 denom = _mm512_add_ps(_mm512_mul_ps(_mm512_sqrt_ps(vv), bc2r), eps);
        __m512 step = _mm512_mul_ps(_mm512_mul_ps(_mm512_loadu_ps(col_lr + k), ibc1), _mm512_div_ps(mm, denom));
        _mm512_storeu_ps(p + k, _mm512_add_ps(_mm512_loadu_ps(p + k), step));
        _mm512_storeu_ps(m + k, mm);
        _mm512_storeu_ps(v + k, vv);
        if (GMODE == G_TABLE) _mm512_storeu_ps(g + k, _mm512_setzero_ps());
    }
    for (; k < cols; ++k) {
        float gg = GMODE == G_ZERO ? 0.f : g[k] * c.scale;
        float mm = c.b1 * m[k] + (1.0f - c.b1) * gg;
        float vv = c.b2 * v[k] + (1.0f - c.b2) * gg * gg;
        float denom = sqrtf(vv) * c.bc2_rsqrt + c.eps;
        p[k] += -(col_lr[k] * c.inv_bc1) * (mm / denom);
        m[k] = mm; v[k] = vv;
        if (GMODE == G_TABLE) g[k] = 0.f;
    }
}

inline bool use_avx512() {
    static const bool v = [] {
        const char *e = getenv("CLMGS_CPU_ADAM_ISA");  // "avx2" forces the 8-lane path
        if (e && !strcmp(e, "avx2")) return false;
        return (bool)__builtin_cpu_supports("avx512f");
    }();
    return v;
}

// Rows [a,b) of a row list (or of the identity when rows == nullptr).  One function per instruction set, so that
// the 8-lane build never contains a 512-bit instruction.
#define CLMGS_ADAM_ROWS(NAME, TARGET, ROWFN)                                                                          \
    template <int GMODE>                                                                                             \
    __attribute__((target(TARGET))) void NAME(float *param, float *grad, float *exp_avg, float *exp_avg_sq,          \
                                              const float *col_lr, int cols, const int32_t *rows, int64_t a,         \
                                              int64_t b, int64_t packed_base, const AdamConst &c) {                  \
        const PrefetchCfg pf = prefetch_cfg();                                                                       \
        for (int64_t r = a; r < b; ++r) {                                                                            \
            const int64_t base = (int64_t)(rows ? rows[r] : r) * cols;                                               \
            if (rows && pf.rows && r + pf.rows < b) { /* every 64-byte line of the row that comes up pf.rows later */ \
                const int64_t nb = (int64_t)rows[r + pf.rows] * cols;                                                \
                if (pf.level == 0) {                                                                                 \
                    for (int l = 0; l < cols; l += 16) {                                                             \
                        _mm_prefetch((const char *)(param + nb + l), _MM_HINT_T0);                                   \
                        if (GMODE == G_TABLE) _mm_prefetch((const char *)(grad + nb + l), _MM_HINT_T0);              \
                        _mm_prefetch((const char *)(exp_avg + nb + l), _MM_HINT_T0);                                 \
                        _mm_prefetch((const char *)(exp_avg_sq + nb + l), _MM_HINT_T0);                              \
                    }                                                                                                \
                } else {                                                                                             \
                    for (int l = 0; l < cols; l += 16) {                                                             \
                        _mm_prefetch((const char *)(param + nb + l), _MM_HINT_T1);                                   \
                        if (GMODE == G_TABLE) _mm_prefetch((const char *)(grad + nb + l), _MM_HINT_T1);              \
                        _mm_prefetch((const char *)(exp_avg + nb + l), _MM_HINT_T1);                                 \
                        _mm_prefetch((const char *)(exp_avg_sq + nb + l), _MM_HINT_T1);                              \
                    }                                                                                                \
                }                                                                                                    \
            }                                                                                                        \
            float *g = GMODE == G_PACKED ? grad + (r - packed_base) * cols : grad + base;                            \
            ROWFN<GMODE>(param + base, g, exp_avg + base, exp_avg_sq + base, col_lr, cols, c);                       \
        }                                                                                                            \
    }
CLMGS_ADAM_ROWS(adam_rows_avx2, "avx2", adam_row_avx2)
CLMGS_ADAM_ROWS(adam_rows_avx512, "avx2,avx512f", adam_row_avx512)
#undef CLMGS_ADAM_ROWS

template <int GMODE>
inline void adam_rows(float *param, float *grad, float *exp_avg, float *exp_avg_sq, const float *col_lr, int cols,
                      const int32_t *rows, int64_t a, int64_t b, int64_t packed_base, const AdamConst &c) {
    if (use_avx512()) adam_rows_avx512<GMODE>(param, grad, exp_avg, exp_avg_sq, col_lr, cols, rows, a, b, packed_base, c);
    else adam_rows_avx2<GMODE>(param, grad, exp_avg, exp_avg_sq, col_lr, cols, rows, a, b, packed_base, c);
}

template <bool ZERO_GRAD>
inline void adam_row(float *p, float *g, float *m, float *v, const float *col_lr, int cols, const AdamConst &c) {
    adam_rows<ZERO_GRAD ? G_ZERO : G_TABLE>(p, g, m, v, col_lr, cols, nullptr, 0, 1, 0, c);
}

inline int64_t now_ns() {
    return std::chrono::duration_cast<std::chrono::nanoseconds>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

// returns the nanoseconds spent waiting
inline int64_t wait_flag(const volatile int32_t *flag) {
    if (*flag != 0) { __atomic_thread_fence(__ATOMIC_ACQUIRE); return 0; }
    // Short spin, then sleep in 20 us naps: a gradient group takes milliseconds to arrive, and every host thread
    // that spins here is a core the kernel-launching thread and the CUDA driver threads cannot have (measured at
    // C4: with all cores spinning, the GPU ran dry between launches).
    const int64_t t0 = now_ns();
    int spins = 0;
    while (*flag == 0) {
        if (++spins < 256) { _mm_pause(); continue; }
        struct timespec ts = {0, 20000};
        nanosleep(&ts, nullptr);
    }
    __atomic_thread_fence(__ATOMIC_ACQUIRE);
    return now_ns() - t0;
}

// Persistent host workers.  The host Adam runs once per training step and, in the densify / lazy paths, a few more
// times; creating and joining one std::thread per worker per call (what round 1 did) costs a clone() + an exit per
// thread per step and lands the new threads on whatever cores the scheduler picks.  The pool keeps its threads
// parked on a condition variable between calls; CLMGS_CPU_ADAM_PIN=<first cpu> pins worker t to cpu (first + t)
// modulo the cpu count (multi-process launches give each rank its own range).  One job at a time (run_mu).
class WorkerPool {
  public:
    static WorkerPool &get() {
        static WorkerPool *p = new WorkerPool();  // leaked on purpose: workers are detached and parked at exit
        return *p;
    }
    template <class F>
    void run(int n, F &&f) {
        if (n <= 1) { f(0); return; }
        std::lock_guard<std::mutex> serial(run_mu_);
        std::function<void(int)> job = std::forward<F>(f);
        {
            std::unique_lock<std::mutex> lk(mu_);
            while ((int)n_threads_ < n - 1) spawn(lk);
            job_ = &job;
            n_active_ = n - 1;   // workers 0..n-2 run tids 1..n-1; the caller runs tid 0
            remaining_ = n - 1;
            ++epoch_;
        }
        cv_start_.notify_all();
        f(0);
        std::unique_lock<std::mutex> lk(mu_);
        cv_done_.wait(lk, [&] { return remaining_ == 0; });
        job_ = nullptr;
    }

  private:
    void spawn(std::unique_lock<std::mutex> &) {
        const int id = (int)n_threads_++;
        const uint64_t born = epoch_;
        std::thread([this, id, born] { loop(id, born); }).detach();
    }
    void loop(int id, uint64_t seen) {
        const char *pin = getenv("CLMGS_CPU_ADAM_PIN");
        if (pin && *pin) {
            const int ncpu = (int)std::max(1u, std::thread::hardware_concurrency());
            cpu_set_t set;
            CPU_ZERO(&set);
            CPU_SET((atoi(pin) + id + 1) % ncpu, &set);
            pthread_setaffinity_np(pthread_self(), sizeof(set), &set);
        }
        for (;;) {
            std::function<void(int)> *job;
            {
                std::unique_lock<std::mutex> lk(mu_);
                cv_start_.wait(lk, [&] { return epoch_ != seen; });
                seen = epoch_;
                if (id >= n_active_) continue;
                job = job_;
            }
            (*job)(id + 1);
            {
                std::lock_guard<std::mutex> lk(mu_);
                if (--remaining_ == 0) cv_done_.notify_all();
            }
        }
    }
    std::mutex mu_, run_mu_;
    std::condition_variable cv_start_, cv_done_;
    std::function<void(int)> *job_ = nullptr;
    size_t n_threads_ = 0;
    int n_active_ = 0, remaining_ = 0;
    uint64_t epoch_ = 0;
};

template <class F>
void parallel_ranges(int64_t n, int n_threads, F f) {
    if (n <= 0) return;
    int T = (int)std::min<int64_t>(n_threads, std::max<int64_t>(1, n / 1024));
    if (T <= 1) { f((int64_t)0, n); return; }
    WorkerPool::get().run(T, [&](int t) { f(n * t / T, n * (t + 1) / T); });
}

// statistics of the last clmgs_cpu_adam_batched_sparse_step call of this process (clmgs_cpu_adam_last_stats)
std::atomic<int64_t> g_stat_total_ns(0), g_stat_wait_ns(0), g_stat_rows(0), g_stat_zero_rows(0);

}  // namespace

static int batched_sparse_step_impl(float *host_param, float *host_grad, float *host_grad_packed, int64_t packed_base,
                                    float *host_exp_avg, float *host_exp_avg_sq, int cols, const float *col_lr,
                                    const int32_t *host_rows, const int64_t *seg_begin, const int64_t *seg_end,
                                    const int32_t *seg_signal, const int32_t *seg_zero_grad, int n_segments,
                                    const volatile int32_t *h_signal, float b1, float b2, float eps, int64_t step,
                                    float scale, int n_threads, volatile int64_t *host_progress) {
    if (cols <= 0 || n_segments < 0 || step < 1) {
        clmgs_set_error("cpu_adam: bad arguments (cols=%d n_segments=%d step=%lld)", cols, n_segments, (long long)step);
        return CLMGS_ERR_INVALID;
    }
    if (n_threads <= 0) n_threads = (int)std::max(1u, std::thread::hardware_concurrency());
    n_threads = std::min(n_threads, 256);
    AdamConst c;
    c.b1 = b1; c.b2 = b2; c.eps = eps; c.scale = scale;
    c.inv_bc1 = (float)(1.0 / (1.0 - pow((double)b1, (double)step)));
    c.bc2_rsqrt = (float)(1.0 / sqrt(1.0 - pow((double)b2, (double)step)));
    std::vector<std::atomic<int>> seg_done(n_segments > 0 ? n_segments : 1);
    for (auto &d : seg_done) d.store(0);
    if (host_progress) *host_progress = 0;
    const int64_t t_start = now_ns();
    std::atomic<int64_t> max_wait(0);
    auto worker = [&](int tid) {
        int64_t waited = 0;
        for (int m = 0; m < n_segments; ++m) {
            if (seg_signal[m] >= 0 && h_signal) waited += wait_flag(h_signal + seg_signal[m]);  // these gradients have landed
            const int64_t lo = seg_begin[m], hi = seg_end[m], n = hi - lo;
            const bool zg = seg_zero_grad && seg_zero_grad[m];
            const int64_t a = lo + n * tid / n_threads, b = lo + n * (tid + 1) / n_threads;
            if (zg) adam_rows<G_ZERO>(host_param, host_grad, host_exp_avg, host_exp_avg_sq, col_lr, cols, host_rows, a, b, 0, c);
            else if (host_grad_packed) adam_rows<G_PACKED>(host_param, host_grad_packed, host_exp_avg, host_exp_avg_sq, col_lr, cols, host_rows, a, b, packed_base, c);
            else adam_rows<G_TABLE>(host_param, host_grad, host_exp_avg, host_exp_avg_sq, col_lr, cols, host_rows, a, b, 0, c);
            // the last thread to finish segment m publishes "segments 0..m are done" (threads run the
            // segments in order, so completion is monotonic)
            if (seg_done[m].fetch_add(1, std::memory_order_acq_rel) + 1 == n_threads && host_progress) {
                __atomic_thread_fence(__ATOMIC_RELEASE);
                *host_progress = m + 1;
            }
        }
        int64_t cur = max_wait.load();
        while (waited > cur && !max_wait.compare_exchange_weak(cur, waited)) {}
    };
    WorkerPool::get().run(n_threads, worker);
    int64_t rows = 0, zrows = 0;
    for (int m = 0; m < n_segments; ++m) {
        rows += seg_end[m] - seg_begin[m];
        if (seg_zero_grad && seg_zero_grad[m]) zrows += seg_end[m] - seg_begin[m];
    }
    g_stat_total_ns = now_ns() - t_start;
    g_stat_wait_ns = max_wait.load();
    g_stat_rows = rows;
    g_stat_zero_rows = zrows;
    return CLMGS_OK;
}

CLMGS_EXPORT int clmgs_cpu_adam_batched_sparse_step(float *host_param, float *host_grad, float *host_exp_avg,
                                                    float *host_exp_avg_sq, int cols, const float *col_lr,
                                                    const int32_t *host_rows, const int64_t *seg_begin,
                                                    const int64_t *seg_end, const int32_t *seg_signal,
                                                    const int32_t *seg_zero_grad, int n_segments,
                                                    const volatile int32_t *h_signal, float b1, float b2, float eps,
                                                    int64_t step, float scale, int n_threads,
                                                    volatile int64_t *host_progress) {
    return batched_sparse_step_impl(host_param, host_grad, nullptr, 0, host_exp_avg, host_exp_avg_sq, cols, col_lr,
                                    host_rows, seg_begin, seg_end, seg_signal, seg_zero_grad, n_segments, h_signal, b1,
                                    b2, eps, step, scale, n_threads, host_progress);
}

// Gradients in a PACKED staging buffer instead of the [N, cols] table: the gradient of list entry r is row
// (r - packed_base) of host_grad_packed.  The engine's flush writes that buffer with plain DMA copies (contiguous on
// both sides), the host reads each row once and does not have to zero it again; the gradient table is not touched.
CLMGS_EXPORT int clmgs_cpu_adam_batched_sparse_step_packed(float *host_param, float *host_grad_packed,
                                                           int64_t packed_base, float *host_exp_avg,
                                                           float *host_exp_avg_sq, int cols, const float *col_lr,
                                                           const int32_t *host_rows, const int64_t *seg_begin,
                                                           const int64_t *seg_end, const int32_t *seg_signal,
                                                           const int32_t *seg_zero_grad, int n_segments,
                                                           const volatile int32_t *h_signal, float b1, float b2,
                                                           float eps, int64_t step, float scale, int n_threads,
                                                           volatile int64_t *host_progress) {
    if (!host_grad_packed || packed_base < 0) {
        clmgs_set_error("cpu_adam packed: null staging buffer or negative base");
        return CLMGS_ERR_INVALID;
    }
    for (int m = 0; m < n_segments; ++m)
        if (!(seg_zero_grad && seg_zero_grad[m]) && seg_begin[m] < packed_base && seg_end[m] > seg_begin[m]) {
            clmgs_set_error("cpu_adam packed: segment %d starts before packed_base", m);
            return CLMGS_ERR_INVALID;
        }
    return batched_sparse_step_impl(host_param, nullptr, host_grad_packed, packed_base, host_exp_avg, host_exp_avg_sq,
                                    cols, col_lr, host_rows, seg_begin, seg_end, seg_signal, seg_zero_grad, n_segments,
                                    h_signal, b1, b2, eps, step, scale, n_threads, host_progress);
}

// out[4] = {wall milliseconds of the last batched_sparse_step call, milliseconds of it spent waiting for gradient
// flags (the slowest thread's), rows updated, of which rows with the zero-gradient shortcut}.
CLMGS_EXPORT int clmgs_cpu_adam_last_stats(double *host_out4) {
    if (!host_out4) { clmgs_set_error("cpu_adam_last_stats: null pointer"); return CLMGS_ERR_INVALID; }
    host_out4[0] = g_stat_total_ns.load() * 1e-6;
    host_out4[1] = g_stat_wait_ns.load() * 1e-6;
    host_out4[2] = (double)g_stat_rows.load();
    host_out4[3] = (double)g_stat_zero_rows.load();
    return CLMGS_OK;
}

// ---------------------------------------------------------------------------------------------------
// Lazy dense Adam.  With dense Adam every row is updated every step even when no camera saw it (its gradient is
// zero, but the moments decay and the parameter still moves): at 28 M rows that is 30 GB of host DRAM traffic per
// step for rows whose values nobody reads.  A zero-gradient step depends on nothing but the row itself, so it can be
// postponed: last_step[r] records the step through which row r is up to date, and the postponed steps are replayed
// -- the same adam_row<true> call with the same per-step bias corrections, hence bit-identical results -- right
// before the row is next needed (before its SH coefficients are sent to the GPU, or when the whole table is
// materialised for evaluation / densification / saving).  Rows: host_rows[0..n_rows), or all rows 0..n_rows if NULL.
// Each row is brought to `through_step` and then marked `mark_step` (= through_step + 1 when the caller is about to
// apply this batch's real step to exactly these rows, = through_step otherwise).
CLMGS_EXPORT int clmgs_cpu_adam_lazy_catch_up(float *host_param, float *host_exp_avg, float *host_exp_avg_sq, int cols,
                                              const float *col_lr, const int32_t *host_rows, int64_t n_rows,
                                              int32_t *last_step, int64_t through_step, int64_t mark_step, float b1,
                                              float b2, float eps, int n_threads) {
    if (cols <= 0 || n_rows < 0 || through_step < 0 || mark_step < through_step || !host_param || !host_exp_avg ||
        !host_exp_avg_sq || !col_lr || !last_step) {
        clmgs_set_error("cpu_adam_lazy_catch_up: bad arguments (cols=%d n_rows=%lld through=%lld mark=%lld)", cols,
                        (long long)n_rows, (long long)through_step, (long long)mark_step);
        return CLMGS_ERR_INVALID;
    }
    if (n_threads <= 0) n_threads = (int)std::max(1u, std::thread::hardware_concurrency());
    n_threads = std::min(n_threads, 256);
    // oldest state among the rows -> the range of steps whose constants are needed
    std::atomic<int64_t> oldest_a(through_step);
    parallel_ranges(n_rows, n_threads, [&](int64_t a, int64_t b) {
        int64_t lo = through_step;
        for (int64_t r = a; r < b; ++r) lo = std::min<int64_t>(lo, last_step[host_rows ? host_rows[r] : r]);
        int64_t cur = oldest_a.load();
        while (lo < cur && !oldest_a.compare_exchange_weak(cur, lo)) {}
    });
    const int64_t oldest = oldest_a.load();
    if (oldest < 0) {
        clmgs_set_error("cpu_adam_lazy_catch_up: negative last_step");
        return CLMGS_ERR_INVALID;
    }
    std::vector<AdamConst> consts((size_t)(through_step - oldest) + 1);  // consts[s - oldest], s in (oldest, through]
    for (int64_t st = oldest + 1; st <= through_step; ++st) {
        AdamConst &c = consts[(size_t)(st - oldest)];
        c.b1 = b1; c.b2 = b2; c.eps = eps; c.scale = 1.0f;
        c.inv_bc1 = (float)(1.0 / (1.0 - pow((double)b1, (double)st)));
        c.bc2_rsqrt = (float)(1.0 / sqrt(1.0 - pow((double)b2, (double)st)));
    }
    parallel_ranges(n_rows, n_threads, [&](int64_t a, int64_t b) {
        for (int64_t r = a; r < b; ++r) {
            const int64_t row = host_rows ? host_rows[r] : r;
            const int64_t base = row * cols;
            for (int64_t st = (int64_t)last_step[row] + 1; st <= through_step; ++st)
                adam_row<true>(host_param + base, nullptr, host_exp_avg + base, host_exp_avg_sq + base, col_lr, cols,
                               consts[(size_t)(st - oldest)]);
            last_step[row] = (int32_t)mark_step;
        }
    });
    return CLMGS_OK;
}

// ---------------------------------------------------------------------------------------------------
// Densify / prune on the pinned tables.  The reference compacts the pinned SH table with
// torch fancy indexing on the CPU (strategies/clm_offload/gaussian_model.py:566-570: a temporary of the whole
// kept set, then a copy back) and re-allocates both Adam moments (:534-537, :640-655).  Here the three tables
// [capacity, cols] are compacted IN PLACE with all host threads: host_src[0..nA) are the kept rows (strictly
// ascending, so src >= dst), host_src[nA..nA+n_tail) the rows whose SH coefficients the appended rows (clones,
// split children) copy; appended rows get zero moments.  Kept rows move in waves through a bounce buffer: every
// thread first reads its share of the wave, all meet at a barrier, then every thread writes -- later waves only read rows
// at or above their own destination, so nothing is overwritten before it is read.

CLMGS_EXPORT int clmgs_host_rows_densify(float *host_param, float *host_exp_avg, float *host_exp_avg_sq, int cols,
                                         int64_t n_old, int64_t capacity, const int32_t *host_src, int64_t nA,
                                         int64_t n_tail, int n_threads) {
    if (cols <= 0 || n_old < 0 || nA < 0 || n_tail < 0 || nA > n_old || nA + n_tail > capacity) {
        clmgs_set_error("host_rows_densify: bad sizes (cols=%d n_old=%lld nA=%lld n_tail=%lld capacity=%lld)", cols,
                        (long long)n_old, (long long)nA, (long long)n_tail, (long long)capacity);
        return CLMGS_ERR_INVALID;
    }
    if (!host_param || !host_exp_avg || !host_exp_avg_sq || (!host_src && nA + n_tail > 0)) {
        clmgs_set_error("host_rows_densify: null pointer");
        return CLMGS_ERR_INVALID;
    }
    if (n_threads <= 0) n_threads = (int)std::max(1u, std::thread::hardware_concurrency());
    n_threads = std::min(n_threads, 256);
    std::atomic<int> bad(0);
    parallel_ranges(nA + n_tail, n_threads, [&](int64_t a, int64_t b) {
        for (int64_t j = a; j < b; ++j) {
            int64_t s = host_src[j];
            if (s < 0 || s >= n_old || (j < nA && (s < j || (j > 0 && host_src[j - 1] >= s)))) bad.store(1);
        }
    });
    if (bad.load()) {
        clmgs_set_error("host_rows_densify: source list is not strictly ascending / out of range");
        return CLMGS_ERR_INVALID;
    }
    const size_t row_bytes = (size_t)cols * sizeof(float);
    int64_t first = 0;  // the untouched prefix (src == dst) is skipped
    while (first < nA && host_src[first] == first) ++first;
    // waves small enough that the bounce buffer stays in the last-level cache
    const int64_t wave = 65536;
    const int T = (int)std::max<int64_t>(1, std::min<int64_t>(n_threads, (nA - first + n_tail) / 4096 + 1));
    std::unique_ptr<float[]> tail(new (std::nothrow) float[(size_t)std::max<int64_t>(n_tail, 1) * cols]);
    std::unique_ptr<float[]> bounce(new (std::nothrow) float[(size_t)wave * cols]);
    if (!tail || !bounce) {
        clmgs_set_error("host_rows_densify: out of host memory for %lld staged rows", (long long)(n_tail + wave));
        return CLMGS_ERR_INVALID;
    }
    float *tables[3] = {host_param, host_exp_avg, host_exp_avg_sq};
    // persistent workers + a sense-reversing barrier: ~100 waves x 3 tables x 2 phases would otherwise pay a
    // thread spawn each
    std::atomic<int> arrived(0), sense(0);
    auto barrier = [&](int &local) {
        local ^= 1;
        if (arrived.fetch_add(1, std::memory_order_acq_rel) + 1 == T) {
            arrived.store(0, std::memory_order_relaxed);
            sense.store(local, std::memory_order_release);
        } else {
            int spins = 0;
            while (sense.load(std::memory_order_acquire) != local) {
                _mm_pause();
                if (++spins > 2000) { std::this_thread::yield(); spins = 0; }
            }
        }
    };
    auto worker = [&](int tid) {
        int local = 0;
        // 1. coefficients of the appended rows, read before anything moves
        for (int64_t r = n_tail * tid / T; r < n_tail * (tid + 1) / T; ++r)
            memcpy(tail.get() + r * cols, host_param + (int64_t)host_src[nA + r] * cols, row_bytes);
        barrier(local);
        // 2. kept rows, in place
        for (float *tab : tables) {
            for (int64_t w0 = first; w0 < nA; w0 += wave) {
                const int64_t n = std::min(wave, nA - w0), a = n * tid / T, b = n * (tid + 1) / T;
                for (int64_t r = a; r < b; ++r)
                    memcpy(bounce.get() + r * cols, tab + (int64_t)host_src[w0 + r] * cols, row_bytes);
                barrier(local);  // every read of this wave happened
                if (b > a) memcpy(tab + (w0 + a) * cols, bounce.get() + a * cols, (size_t)(b - a) * row_bytes);
                barrier(local);  // bounce buffer free again
            }
        }
        // 3. appended rows: copied coefficients, zero moments
        const int64_t a = n_tail * tid / T, b = n_tail * (tid + 1) / T;
        if (b > a) {
            memcpy(host_param + (nA + a) * cols, tail.get() + a * cols, (size_t)(b - a) * row_bytes);
            memset(host_exp_avg + (nA + a) * cols, 0, (size_t)(b - a) * row_bytes);
            memset(host_exp_avg_sq + (nA + a) * cols, 0, (size_t)(b - a) * row_bytes);
        }
    };
    if (T == 1) { worker(0); return CLMGS_OK; }
    std::vector<std::thread> th;
    th.reserve(T);
    for (int t = 0; t < T; ++t) th.emplace_back(worker, t);
    for (auto &t : th) t.join();
    return CLMGS_OK;
}

// Deterministic TSP heuristic: nearest-neighbour tour from node 0, improved by
// 2-opt and Or-opt passes until no move helps (n <= 64, so this is microseconds).
CLMGS_EXPORT int clmgs_tsp_tour(const int32_t *host_dist, int n, int32_t *host_tour) {
    if (n <= 0 || n > 4096) { clmgs_set_error("tsp: bad n=%d", n); return CLMGS_ERR_INVALID; }
    auto D = [&](int a, int b) { return (int64_t)host_dist[a * n + b]; };
    std::vector<int> tour(n);
    std::vector<char> used(n, 0);
    tour[0] = 0; used[0] = 1;
    for (int i = 1; i < n; ++i) {
        int best = -1;
        for (int j = 0; j < n; ++j)
            if (!used[j] && (best < 0 || D(tour[i - 1], j) < D(tour[i - 1], best))) best = j;
        tour[i] = best; used[best] = 1;
    }
    if (n > 3) {
        bool improved = true;
        int rounds = 0;
        while (improved && rounds++ < 100) {
            improved = false;
            for (int i = 0; i < n - 1; ++i)           // 2-opt on the closed tour
                for (int j = i + 2; j < n; ++j) {
                    if (i == 0 && j == n - 1) continue;
                    int a = tour[i], b = tour[i + 1], c = tour[j], d = tour[(j + 1) % n];
                    if (D(a, c) + D(b, d) < D(a, b) + D(c, d)) {
                        std::reverse(tour.begin() + i + 1, tour.begin() + j + 1);
                        improved = true;
                    }
                }
            for (int i = 0; i < n; ++i) {             // Or-opt: move one city elsewhere
                int prev = tour[(i + n - 1) % n], cur = tour[i], next = tour[(i + 1) % n];
                int64_t gain = D(prev, cur) + D(cur, next) - D(prev, next);
                if (gain <= 0) continue;
                for (int j = 0; j < n; ++j) {
                    if (j == i || (j + 1) % n == i) continue;
                    int a = tour[j], b = tour[(j + 1) % n];
                    if (D(a, cur) + D(cur, b) - D(a, b) < gain) {
                        std::vector<int> t2;
                        t2.reserve(n);
                        for (int k = 0; k < n; ++k) {
                            if (k == i) continue;
                            t2.push_back(tour[k]);
                            if (k == j) t2.push_back(cur);
                        }
                        tour.swap(t2);
                        improved = true;
                        break;
                    }
                }
            }
        }
        // rotate so that the tour starts at camera 0 (as a fixed, reproducible anchor)
        auto it = std::find(tour.begin(), tour.end(), 0);
        std::rotate(tour.begin(), it, tour.end());
    }
    for (int i = 0; i < n; ++i) host_tour[i] = tour[i];
    return CLMGS_OK;
}
