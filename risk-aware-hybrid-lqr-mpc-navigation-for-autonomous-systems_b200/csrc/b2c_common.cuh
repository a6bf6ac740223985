// Shared internals of libb2c.so (not part of the C ABI).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string>
#include "b2c.h"

struct b2c_ctx {
    int device = 0;
    cudaStream_t stream = nullptr;
    bool own_stream = false;
    int sm_count = 148;
    // device scratch (grown on demand, never shrunk)
    void* scratch = nullptr;
    size_t scratch_bytes = 0;
    // device staging for the *_host entry points
    void* dstage = nullptr;
    size_t dstage_bytes = 0;
    // pinned host staging
    void* hstage = nullptr;
    size_t hstage_bytes = 0;
    int64_t launches = 0;
    // chunk pipeline of the *_host entry points: two compute lanes (stream + private K3 scratch) and an output
    // stream, created on first use; events are reused
    cudaStream_t lane_stream[2] = {nullptr, nullptr};
    cudaStream_t out_stream = nullptr;
    void* lane_scratch[2] = {nullptr, nullptr};
    size_t lane_scratch_bytes[2] = {0, 0};
    cudaEvent_t ev[16] = {};
    int64_t chunk_seq = 0;                 // chunks issued so far (alternates the compute lanes across calls too)
    // third device arena (materialised table segments for the general kernel)
    void* aux = nullptr;
    size_t aux_bytes = 0;
    void* aux2 = nullptr;                  // device copies of b2c_rollout_host's arguments
    size_t aux2_bytes = 0;
    // the *_host MPC entry points double-buffer their device staging so that an asynchronous call can copy its
    // inputs while the previous call is still solving: at most two calls in flight
    struct HostJob { void* d = nullptr; size_t bytes = 0; cudaEvent_t done = nullptr; bool pending = false; } job[2];
    int job_next = 0;
};

// Every entry point runs on the context's device whatever the caller's current device is, and leaves the
// caller's current device as it found it (one process may drive several GPUs with one context each).
struct B2cDeviceGuard {
    int prev = -1; cudaError_t err = cudaSuccess;
    explicit B2cDeviceGuard(int dev) {
        if (cudaGetDevice(&prev) != cudaSuccess) prev = -1;
        if (prev != dev) err = cudaSetDevice(dev); else prev = -1;
    }
    ~B2cDeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};
#define B2C_DEVICE_GUARD(ctx)                                                   \
    B2cDeviceGuard _b2c_dg((ctx)->device);                                      \
    if (_b2c_dg.err != cudaSuccess) return b2c_cuda_fail(_b2c_dg.err, "cudaSetDevice", __FILE__, __LINE__)

void b2c_set_error(const std::string& s);
int b2c_cuda_fail(cudaError_t e, const char* what, const char* file, int line);

#define B2C_CUDA(expr)                                                          \
    do {                                                                        \
        cudaError_t _e = (expr);                                                \
        if (_e != cudaSuccess) return b2c_cuda_fail(_e, #expr, __FILE__, __LINE__); \
    } while (0)

#define B2C_CHECK_LAUNCH(ctx)                                                   \
    do {                                                                        \
        (ctx)->launches++;                                                      \
        cudaError_t _e = cudaGetLastError();                                    \
        if (_e != cudaSuccess) return b2c_cuda_fail(_e, "kernel launch", __FILE__, __LINE__); \
    } while (0)

int b2c_reserve(void** p, size_t* cur, size_t need, bool pinned_host);

static inline int b2c_blocks(int64_t n, int threads) { return (int)((n + threads - 1) / threads); }

// ---- device helpers -------------------------------------------------------------------------
// B2C_HD functions also compile for the host so that tests/ can run the *same* per-instance code
// on the CPU (tools/emul, -DB2C_HOST_EMUL, a separate test-only library; libb2c.so never calls them
// on the host and has no CPU path).
#define B2C_HD __host__ __device__ __forceinline__
#ifdef __CUDA_ARCH__
#define B2C_LDG(p) __ldg(p)
#define B2C_RSQRT(x) rsqrt(x)
#define B2C_MUL(a, b) __dmul_rn((a), (b))
// branch-free reciprocal for well-scaled arguments (2x2 determinants of the Riccati step): MUFU seed
// (2^-23) + two Newton steps -> <= 1 ulp.  The IEEE path of 1.0/x costs 80 cycles of dependent latency
// and a divergence barrier for its denormal branch; this one costs ~50 and none.
__device__ __forceinline__ double b2c_rcp_dev(double x) {
    double r;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
    double e = fma(-x, r, 1.0);
    r = fma(r, e, r);
    e = fma(-x, r, 1.0);
    return fma(r, e, r);
}
#define B2C_RCP(x) b2c_rcp_dev(x)
#else
#include <math.h>
#define B2C_LDG(p) (*(p))
#define B2C_RSQRT(x) (1.0 / sqrt(x))
#define B2C_MUL(a, b) ((a) * (b))
#define B2C_RCP(x) (1.0 / (x))
#endif
// low-speed guards of the reference (lqr_controller.py:120-122, mpc_controller.py:186,425)
B2C_HD double b2c_guard_v(double v, int guard) {
    if (guard == B2C_GUARD_LQR) return (fabs(v) < 1e-6) ? 0.01 : v;
    if (guard == B2C_GUARD_MPC) return (fabs(v) > 0.01) ? v : 0.1;
    return v;
}

// closed-interval wrap by +-2pi steps (lqr_controller.py:244-250): pi stays pi, -pi stays -pi.
B2C_HD double b2c_wrap(double a) {
    const double PI = 3.141592653589793, TWO_PI = 2.0 * 3.141592653589793;
    if (!(fabs(a) < 1e9)) return a;  // non-finite / absurd: leave (the reference would spin)
    while (a > PI) a -= TWO_PI;
    while (a < -PI) a += TWO_PI;
    return a;
}

// np.clip: compare-and-select, NaN stays NaN (as in numpy).  fmin(fmax()) costs 16 instructions and 58 cycles of
// dependent latency per call on sm_100a (no DMNMX: DSETP + FSEL/SEL + NaN quieting, twice); this form costs 6.
B2C_HD double b2c_clip(double x, double lo, double hi) {
    x = (x < lo) ? lo : x;
    return (x > hi) ? hi : x;
}
