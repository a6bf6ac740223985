// Context lifecycle, host-buffer entry points (pinned staging + on-device re-layout) and the
// FMA-peak microbenchmark of libb2c.so.
#include <stdlib.h>
#include <string.h>
#include <mutex>
#include "mpc_args.cuh"

static thread_local std::string g_err;
void b2c_set_error(const std::string& s) { g_err = s; }
const char* b2c_last_error(void) { return g_err.c_str(); }
int b2c_version(void) { return 200; }
int b2c_current_device(int* out) {
    if (!out) { b2c_set_error("null out pointer"); return -1; }
    cudaError_t e = cudaGetDevice(out);
    if (e != cudaSuccess) { *out = -1; b2c_set_error(std::string("no CUDA device: ") + cudaGetErrorString(e)); return -10; }
    return 0;
}

int b2c_cuda_fail(cudaError_t e, const char* what, const char* file, int line) {
    char buf[512];
    snprintf(buf, sizeof buf, "CUDA error %d (%s) at %s:%d: %s", (int)e, cudaGetErrorString(e), file, line, what);
    g_err = buf;
    return -100 - (int)e;
}

int b2c_reserve(void** p, size_t* cur, size_t need, bool pinned_host) {
    if (need <= *cur) return 0;
    size_t want = need + need / 4 + 4096;
    if (*p) {
        cudaError_t e = pinned_host ? cudaFreeHost(*p) : cudaFree(*p);
        *p = nullptr; *cur = 0;
        if (e != cudaSuccess) return b2c_cuda_fail(e, "free", __FILE__, __LINE__);
    }
    cudaError_t e = pinned_host ? cudaMallocHost(p, want) : cudaMalloc(p, want);
    if (e != cudaSuccess) { *p = nullptr; return b2c_cuda_fail(e, "alloc", __FILE__, __LINE__); }
    *cur = want;
    return 0;
}

int b2c_create_on_stream(b2c_ctx** out, int device, void* cuda_stream) {
    if (!out) { b2c_set_error("null out pointer"); return -1; }
    *out = nullptr;
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) {
        b2c_set_error(std::string("no CUDA device available (libb2c has no CPU fallback): ") +
                      cudaGetErrorString(e));
        return -10;
    }
    if (device < 0) B2C_CUDA(cudaGetDevice(&device));
    if (device >= ndev) { b2c_set_error("no such CUDA device"); return -11; }
    b2c_ctx* c = new b2c_ctx();
    c->device = device;                               // the caller's current device is left as it is
    int sms = 0;
    e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
    if (e != cudaSuccess) { delete c; return b2c_cuda_fail(e, "device attribute", __FILE__, __LINE__); }
    c->sm_count = sms;
    // the caller's stream is used as given -- a null handle is the legacy default stream
    c->stream = (cudaStream_t)cuda_stream; c->own_stream = false;
    *out = c;
    return 0;
}

int b2c_create(b2c_ctx** out, int device) {
    int rc = b2c_create_on_stream(out, device, nullptr);
    if (rc) return rc;
    B2cDeviceGuard guard((*out)->device);
    cudaError_t e = guard.err;
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&(*out)->stream, cudaStreamNonBlocking);
    if (e != cudaSuccess) { delete *out; *out = nullptr; return b2c_cuda_fail(e, "stream create", __FILE__, __LINE__); }
    (*out)->own_stream = true;
    return 0;
}

void b2c_destroy(b2c_ctx* c) {
    if (!c) return;
    B2cDeviceGuard guard(c->device);
    cudaStreamSynchronize(c->stream);
    if (c->scratch) cudaFree(c->scratch);
    if (c->dstage) cudaFree(c->dstage);
    if (c->hstage) cudaFreeHost(c->hstage);
    for (int k = 0; k < 2; ++k) {
        if (c->lane_scratch[k]) cudaFree(c->lane_scratch[k]);
        if (c->lane_stream[k]) { cudaStreamSynchronize(c->lane_stream[k]); cudaStreamDestroy(c->lane_stream[k]); }
    }
    if (c->out_stream) { cudaStreamSynchronize(c->out_stream); cudaStreamDestroy(c->out_stream); }
    for (int k = 0; k < 16; ++k) if (c->ev[k]) cudaEventDestroy(c->ev[k]);
    for (int k = 0; k < 2; ++k) {
        if (c->job[k].d) cudaFree(c->job[k].d);
        if (c->job[k].done) cudaEventDestroy(c->job[k].done);
    }
    if (c->aux) cudaFree(c->aux);
    if (c->aux2) cudaFree(c->aux2);
    if (c->own_stream) cudaStreamDestroy(c->stream);
    delete c;
}

int b2c_sync(b2c_ctx* c) {
    if (!c) { b2c_set_error("null context"); return -1; }
    B2C_DEVICE_GUARD(c);
    B2C_CUDA(cudaStreamSynchronize(c->stream));
    return 0;
}
void* b2c_stream(b2c_ctx* c) { return c ? (void*)c->stream : nullptr; }
int64_t b2c_launch_count(b2c_ctx* c) { return c ? c->launches : 0; }

// ---- re-layout kernels: instance-major (B,F) <-> instance-minor [F][B] ----------------------
// 32x32 tiles through shared memory so both sides are coalesced.
template <typename T>
__global__ void __launch_bounds__(256)
k_aos_to_soa(int64_t B, int F, const T* __restrict__ aos, T* __restrict__ soa) {
    __shared__ T tile[32][33];
    const int64_t i0 = (int64_t)blockIdx.x * 32;
    const int f0 = blockIdx.y * 32;
    for (int r = threadIdx.y; r < 32; r += 8) {          // r: instance within tile
        int64_t i = i0 + r; int f = f0 + threadIdx.x;
        if (i < B && f < F) tile[r][threadIdx.x] = aos[i * F + f];
    }
    __syncthreads();
    for (int r = threadIdx.y; r < 32; r += 8) {          // r: field within tile
        int f = f0 + r; int64_t i = i0 + threadIdx.x;
        if (i < B && f < F) soa[(int64_t)f * B + i] = tile[threadIdx.x][r];
    }
}
template <typename T>
__global__ void __launch_bounds__(256)
k_soa_to_aos(int64_t B, int F, const T* __restrict__ soa, T* __restrict__ aos) {
    __shared__ T tile[32][33];
    const int64_t i0 = (int64_t)blockIdx.x * 32;
    const int f0 = blockIdx.y * 32;
    for (int r = threadIdx.y; r < 32; r += 8) {          // r: field
        int f = f0 + r; int64_t i = i0 + threadIdx.x;
        if (i < B && f < F) tile[r][threadIdx.x] = soa[(int64_t)f * B + i];
    }
    __syncthreads();
    for (int r = threadIdx.y; r < 32; r += 8) {          // r: instance
        int64_t i = i0 + r; int f = f0 + threadIdx.x;
        if (i < B && f < F) aos[i * F + f] = tile[threadIdx.x][r];
    }
}
// symmetric 3x3 from upper-triangle SoA [6][B] -> (B,3,3)
__global__ void __launch_bounds__(256)
k_sym6_to_full(int64_t B, const double* __restrict__ p6, double* __restrict__ full) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= B) return;
    double xx = p6[i], xy = p6[B + i], xt = p6[2 * B + i], yy = p6[3 * B + i], yt = p6[4 * B + i], tt = p6[5 * B + i];
    double* o = full + i * 9;
    o[0] = xx; o[1] = xy; o[2] = xt; o[3] = xy; o[4] = yy; o[5] = yt; o[6] = xt; o[7] = yt; o[8] = tt;
}

static int to_soa(b2c_ctx* c, int64_t B, int F, const double* aos, double* soa) {
    if (F <= 0) return 0;
    dim3 g((unsigned)((B + 31) / 32), (unsigned)((F + 31) / 32)), t(32, 8);
    k_aos_to_soa<double><<<g, t, 0, c->stream>>>(B, F, aos, soa);
    B2C_CHECK_LAUNCH(c);
    return 0;
}
static int to_aos(b2c_ctx* c, int64_t B, int F, const double* soa, double* aos) {
    if (F <= 0) return 0;
    dim3 g((unsigned)((B + 31) / 32), (unsigned)((F + 31) / 32)), t(32, 8);
    k_soa_to_aos<double><<<g, t, 0, c->stream>>>(B, F, soa, aos);
    B2C_CHECK_LAUNCH(c);
    return 0;
}

// ---- host-buffer entry points ---------------------------------------------------------------
// cudaMemcpyAsync straight from/to the caller's buffers: pass pinned memory (b2c_host_alloc or
// torch pin_memory) for full PCIe rate; pageable memory works but is staged by the driver.
struct Bump {
    char* base; size_t off = 0;
    explicit Bump(void* b) : base((char*)b) {}
    template <typename T> T* take(size_t n) {
        T* p = (T*)(base + off);
        off += ((n * sizeof(T) + 255) / 256) * 256;
        return p;
    }
};
static size_t pad256(size_t bytes) { return ((bytes + 255) / 256) * 256; }

int b2c_host_alloc(void** p, int64_t bytes) {
    if (!p || bytes < 0) { b2c_set_error("bad host alloc request"); return -1; }
    B2C_CUDA(cudaMallocHost(p, (size_t)bytes));
    return 0;
}
int b2c_host_free(void* p) {
    if (p) B2C_CUDA(cudaFreeHost(p));
    return 0;
}

int b2c_linearize_host(b2c_ctx* c, int64_t B, int method, int guard, double dt,
                       const double* v, const double* th, double* Ad, double* Bd) {
    if (!c) { b2c_set_error("null context"); return -1; }
    if (method < 0 || method > 2) { b2c_set_error("Unknown discretization method"); return -2; }
    if (B <= 0) return 0;
    B2C_DEVICE_GUARD(c);
    size_t need = 2 * pad256(B * 8) + pad256(B * 72) + pad256(B * 48);
    int rc = b2c_reserve(&c->dstage, &c->dstage_bytes, need, false);
    if (rc) return rc;
    Bump bp(c->dstage);
    double* dv = bp.take<double>(B); double* dth = bp.take<double>(B);
    double* dA = bp.take<double>(9 * B); double* dB = bp.take<double>(6 * B);
    B2C_CUDA(cudaMemcpyAsync(dv, v, B * 8, cudaMemcpyHostToDevice, c->stream));
    B2C_CUDA(cudaMemcpyAsync(dth, th, B * 8, cudaMemcpyHostToDevice, c->stream));
    rc = b2c_linearize_packed_dev(c, B, method, guard, dt, dv, dth, dA, dB);
    if (rc) return rc;
    B2C_CUDA(cudaMemcpyAsync(Ad, dA, B * 72, cudaMemcpyDeviceToHost, c->stream));
    B2C_CUDA(cudaMemcpyAsync(Bd, dB, B * 48, cudaMemcpyDeviceToHost, c->stream));
    B2C_CUDA(cudaStreamSynchronize(c->stream));
    return 0;
}

int b2c_dare_host(b2c_ctx* c, int64_t B, int method, int guard, double dt,
                  const double Qd[3], const double Rd[2], const double* v, const double* th,
                  double* K, double* P, int32_t* iters, int32_t* status) {
    if (!c) { b2c_set_error("null context"); return -1; }
    if (B <= 0) return 0;
    B2C_DEVICE_GUARD(c);
    size_t need = 2 * pad256(B * 8) + 2 * pad256(B * 48) + pad256(B * 48) + pad256(B * 72) + 2 * pad256(B * 4);
    int rc = b2c_reserve(&c->dstage, &c->dstage_bytes, need, false);
    if (rc) return rc;
    Bump bp(c->dstage);
    double* dv = bp.take<double>(B); double* dth = bp.take<double>(B);
    double* dKs = bp.take<double>(6 * B); double* dKa = bp.take<double>(6 * B);
    double* dP6 = bp.take<double>(6 * B); double* dPf = bp.take<double>(9 * B);
    int32_t* dit = bp.take<int32_t>(B); int32_t* dst = bp.take<int32_t>(B);
    B2C_CUDA(cudaMemcpyAsync(dv, v, B * 8, cudaMemcpyHostToDevice, c->stream));
    B2C_CUDA(cudaMemcpyAsync(dth, th, B * 8, cudaMemcpyHostToDevice, c->stream));
    rc = b2c_dare_dev(c, B, method, guard, dt, Qd, Rd, dv, dth, dKs, P ? dP6 : nullptr, dit, dst);
    if (rc) return rc;
    rc = to_aos(c, B, 6, dKs, dKa);
    if (rc) return rc;
    B2C_CUDA(cudaMemcpyAsync(K, dKa, B * 48, cudaMemcpyDeviceToHost, c->stream));
    if (P) {
        k_sym6_to_full<<<b2c_blocks(B, 256), 256, 0, c->stream>>>(B, dP6, dPf);
        B2C_CHECK_LAUNCH(c);
        B2C_CUDA(cudaMemcpyAsync(P, dPf, B * 72, cudaMemcpyDeviceToHost, c->stream));
    }
    if (iters) B2C_CUDA(cudaMemcpyAsync(iters, dit, B * 4, cudaMemcpyDeviceToHost, c->stream));
    if (status) B2C_CUDA(cudaMemcpyAsync(status, dst, B * 4, cudaMemcpyDeviceToHost, c->stream));
    B2C_CUDA(cudaStreamSynchronize(c->stream));
    return 0;
}

// While alive, the context launches on compute lane k: its stream and its private K3 scratch.
struct LaneScope {
    b2c_ctx* c; int k; cudaStream_t s0; void* sc0; size_t sb0;
    LaneScope(b2c_ctx* c_, int k_) : c(c_), k(k_), s0(c_->stream), sc0(c_->scratch), sb0(c_->scratch_bytes) {
        c->stream = c->lane_stream[k]; c->scratch = c->lane_scratch[k]; c->scratch_bytes = c->lane_scratch_bytes[k];
    }
    ~LaneScope() {
        c->lane_scratch[k] = c->scratch; c->lane_scratch_bytes[k] = c->scratch_bytes;
        c->stream = s0; c->scratch = sc0; c->scratch_bytes = sb0;
    }
};

static int pipeline_init(b2c_ctx* c) {
    for (int k = 0; k < 2; ++k)
        if (!c->lane_stream[k]) B2C_CUDA(cudaStreamCreateWithFlags(&c->lane_stream[k], cudaStreamNonBlocking));
    if (!c->out_stream) B2C_CUDA(cudaStreamCreateWithFlags(&c->out_stream, cudaStreamNonBlocking));
    for (int k = 0; k < 16; ++k)
        if (!c->ev[k]) B2C_CUDA(cudaEventCreateWithFlags(&c->ev[k], cudaEventDisableTiming));
    for (int k = 0; k < 2; ++k)
        if (!c->job[k].done) B2C_CUDA(cudaEventCreateWithFlags(&c->job[k].done, cudaEventDisableTiming));
    return 0;
}

// every stream a host call may have work on: drained before an error is reported, so that no copy still targets the
// caller's buffers when the call returns
static void pipeline_drain(b2c_ctx* c) {
    cudaStreamSynchronize(c->stream);
    for (int k = 0; k < 2; ++k) if (c->lane_stream[k]) cudaStreamSynchronize(c->lane_stream[k]);
    if (c->out_stream) cudaStreamSynchronize(c->out_stream);
    c->job[0].pending = c->job[1].pending = false;
}

int b2c_wait(b2c_ctx* c) {
    if (!c) { b2c_set_error("null context"); return -1; }
    B2C_DEVICE_GUARD(c);
    for (int k = 0; k < 2; ++k)
        if (c->job[k].pending) {
            c->job[k].pending = false;
            B2C_CUDA(cudaEventSynchronize(c->job[k].done));
        }
    return 0;
}

// Host entry of K3, both forms (materialised segments / shared table + start rows), synchronous or asynchronous.
//
// Staging is double-buffered per call (job 0 / 1) and the work is spread over four streams -- H2D on the context's
// stream, solve on one of two compute lanes, D2H on an output stream -- so that (a) inside one synchronous call the
// batch can be cut into chunks whose copies overlap the previous chunk's solve, and (b) an asynchronous call s+1
// copies its inputs while call s is still solving.  The slot path reads the numpy-layout staging buffers as they are
// (k_mpc_prepare gathers; outputs are written instance-major): no re-layout kernels.  The general kernel (move
// blocking, LTI, hard rows) keeps its SoA layout and goes through the transposes, unchunked.
static int mpc_host_impl(b2c_ctx* c, int64_t B, const b2c_mpc_params* p, const bool async,
                         const double* table, const int table_rows, const int32_t* k0,
                         const double* x0, const double* xref, const double* uref, const double* obs,
                         int32_t* step_count,
                         double* u0, double* cost, int32_t* status, int32_t* iters, int32_t* slack_used,
                         double* useq, double* xpred) {
    if (!c || !p) { b2c_set_error("null context/params"); return -1; }
    if (B <= 0) return 0;
    if (p->N < 1 || p->M < 0) { b2c_set_error("bad horizon / obstacle count"); return -2; }
    if (table && (table_rows < 1 || !k0)) { b2c_set_error("table form needs rows >= 1 and k0"); return -2; }
    if (table && p->kind != B2C_MPC_LTV) { b2c_set_error("the table form is solve_with_ltv's (kind = B2C_MPC_LTV)"); return -2; }
    B2C_DEVICE_GUARD(c);
    int rc = pipeline_init(c);
    if (rc) return rc;
    const int N = p->N, M = p->M;
    const int FX = 3 * (N + 1), FU = 2 * N, FO = 3 * M;
    MpcArgs probe;
    mpc_fill_args(probe, B, p, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr,
                  nullptr, nullptr);
    const bool slots = mpc_uses_slots(probe, p->engine);
    // ---- this call's staging arena (a third call in flight waits for the first) ---------------------------------
    b2c_ctx::HostJob& job = c->job[c->job_next];
    c->job_next ^= 1;
    if (job.pending) { job.pending = false; B2C_CUDA(cudaEventSynchronize(job.done)); }
    const size_t n_tab = table ? (size_t)table_rows * 6 : 0;
    const size_t need = pad256(n_tab * 8) + pad256((size_t)B * 4) + pad256((size_t)B * 3 * 8) + 2 * pad256((size_t)B * FX * 8)
                      + 2 * pad256((size_t)B * FU * 8) + 2 * pad256(((size_t)B * FO + 1) * 8) + pad256((size_t)B * 3 * 8)
                      + 2 * pad256((size_t)B * 2 * 8) + pad256((size_t)B * 8) + 2 * pad256((size_t)B * FU * 8)
                      + 2 * pad256((size_t)B * FX * 8) + 4 * pad256((size_t)B * 4) + 64 * 256;
    rc = b2c_reserve(&job.d, &job.bytes, need, false);
    if (rc) return rc;
    Bump bp(job.d);
    double* d_tab = table ? bp.take<double>(n_tab) : nullptr;
    int32_t* d_k0 = table ? bp.take<int32_t>(B) : nullptr;
    double* a_x0 = bp.take<double>(3 * B);
    double* a_xr = table ? nullptr : bp.take<double>((size_t)FX * B);
    double* a_ur = table ? nullptr : bp.take<double>((size_t)FU * B);
    double* a_ob = bp.take<double>((size_t)FO * B + 1);
    double* a_u0 = bp.take<double>(2 * B);
    double* d_cost = bp.take<double>(B);
    double* a_us = useq ? bp.take<double>((size_t)FU * B) : nullptr;
    double* a_xp = xpred ? bp.take<double>((size_t)FX * B) : nullptr;
    int32_t* d_sc = bp.take<int32_t>(B); int32_t* d_st = bp.take<int32_t>(B);
    int32_t* d_it = bp.take<int32_t>(B); int32_t* d_sl = bp.take<int32_t>(B);
    // general kernel only: the SoA twins
    double *s_x0 = nullptr, *s_xr = nullptr, *s_ur = nullptr, *s_ob = nullptr, *s_u0 = nullptr, *s_us = nullptr, *s_xp = nullptr;
    if (!slots) {
        s_x0 = bp.take<double>(3 * B);
        s_xr = table ? nullptr : bp.take<double>((size_t)FX * B);
        s_ur = table ? nullptr : bp.take<double>((size_t)FU * B);
        s_ob = bp.take<double>((size_t)FO * B + 1); s_u0 = bp.take<double>(2 * B);
        s_us = useq ? bp.take<double>((size_t)FU * B) : nullptr; s_xp = xpred ? bp.take<double>((size_t)FX * B) : nullptr;
    }
    // ---- chunks: a multiple of 32 instances each; only a synchronous call with a large upload is cut
    const size_t h2d = ((size_t)B * (3 + FO) + (table ? n_tab : (size_t)B * (FX + FU))) * 8;
    int C = 1;
    if (slots && !async && h2d >= ((size_t)24 << 20)) { C = (int)(B / 16384); if (C > 8) C = 8; if (C < 1) C = 1; }
    if (const char* e = getenv("B2C_HOST_CHUNKS")) { int v = atoi(e); if (slots && v >= 1 && v <= 8 && B >= 64 * v) C = v; }
    const int64_t per = C > 1 ? ((B / C + 31) / 32) * 32 : B;
    cudaStream_t s_in = c->stream;
    cudaError_t ce = cudaSuccess;
#define HOSTCP(dst, src, bytes, kind, st) \
    do { if (ce == cudaSuccess && (bytes) > 0) ce = cudaMemcpyAsync((dst), (src), (bytes), (kind), (st)); } while (0)
    if (table) {
        HOSTCP(d_tab, table, n_tab * 8, cudaMemcpyHostToDevice, s_in);
        HOSTCP(d_k0, k0, (size_t)B * 4, cudaMemcpyHostToDevice, s_in);
    }
    for (int ch = 0; ch < C && rc == 0 && ce == cudaSuccess; ++ch) {
        const int64_t i0 = (int64_t)ch * per, Bc = (ch == C - 1) ? B - i0 : per;
        if (Bc <= 0) break;
        HOSTCP(a_x0 + 3 * i0, x0 + 3 * i0, (size_t)Bc * 3 * 8, cudaMemcpyHostToDevice, s_in);
        if (!table) {
            HOSTCP(a_xr + FX * i0, xref + FX * i0, (size_t)Bc * FX * 8, cudaMemcpyHostToDevice, s_in);
            HOSTCP(a_ur + FU * i0, uref + FU * i0, (size_t)Bc * FU * 8, cudaMemcpyHostToDevice, s_in);
        }
        if (FO) HOSTCP(a_ob + FO * i0, obs + FO * i0, (size_t)Bc * FO * 8, cudaMemcpyHostToDevice, s_in);
        if (step_count) HOSTCP(d_sc + i0, step_count + i0, (size_t)Bc * 4, cudaMemcpyHostToDevice, s_in);
        if (ce != cudaSuccess) break;
        const int lane = (int)(c->chunk_seq++ & 1);
        ce = cudaEventRecord(c->ev[ch], s_in);
        if (ce == cudaSuccess) ce = cudaStreamWaitEvent(c->lane_stream[lane], c->ev[ch], 0);
        if (ce != cudaSuccess) break;
        {
            LaneScope scope(c, lane);
            if (slots) {
                MpcArgs a;
                mpc_fill_args(a, Bc, p, a_x0 + 3 * i0, table ? nullptr : a_xr + FX * i0, table ? nullptr : a_ur + FU * i0,
                              a_ob + FO * i0, step_count ? d_sc + i0 : nullptr, a_u0 + 2 * i0, d_cost + i0, d_st + i0,
                              d_it + i0, d_sl + i0, useq ? a_us + FU * i0 : nullptr, xpred ? a_xp + FX * i0 : nullptr);
                a.in_aos = 1; a.out_aos = 1;
                if (table) { a.table = d_tab; a.table_rows = table_rows; a.k0 = d_k0 + i0; }
                rc = mpc_launch(c, a, p->engine);
            } else {                                  // C == 1: the whole batch through the transposes
                if (rc == 0) rc = to_soa(c, B, 3, a_x0, s_x0);
                if (rc == 0 && !table) rc = to_soa(c, B, FX, a_xr, s_xr);
                if (rc == 0 && !table) rc = to_soa(c, B, FU, a_ur, s_ur);
                if (rc == 0) rc = to_soa(c, B, FO, a_ob, s_ob);
                if (rc == 0)
                    rc = table ? b2c_mpc_table_dev(c, B, p, d_tab, table_rows, d_k0, s_x0, s_ob, step_count ? d_sc : nullptr,
                                                   s_u0, d_cost, d_st, d_it, d_sl, s_us, s_xp)
                               : b2c_mpc_dev(c, B, p, s_x0, s_xr, s_ur, s_ob, step_count ? d_sc : nullptr,
                                             s_u0, d_cost, d_st, d_it, d_sl, s_us, s_xp);
                if (rc == 0) rc = to_aos(c, B, 2, s_u0, a_u0);
                if (rc == 0 && useq) rc = to_aos(c, B, FU, s_us, a_us);
                if (rc == 0 && xpred) rc = to_aos(c, B, FX, s_xp, a_xp);
            }
            if (rc == 0) ce = cudaEventRecord(c->ev[8 + ch], c->stream);
        }
        if (rc || ce != cudaSuccess) break;
        cudaStream_t s_out = c->out_stream;
        ce = cudaStreamWaitEvent(s_out, c->ev[8 + ch], 0);
        HOSTCP(u0 + 2 * i0, a_u0 + 2 * i0, (size_t)Bc * 2 * 8, cudaMemcpyDeviceToHost, s_out);
        HOSTCP(cost + i0, d_cost + i0, (size_t)Bc * 8, cudaMemcpyDeviceToHost, s_out);
        if (useq) HOSTCP(useq + FU * i0, a_us + FU * i0, (size_t)Bc * FU * 8, cudaMemcpyDeviceToHost, s_out);
        if (xpred) HOSTCP(xpred + FX * i0, a_xp + FX * i0, (size_t)Bc * FX * 8, cudaMemcpyDeviceToHost, s_out);
        if (status) HOSTCP(status + i0, d_st + i0, (size_t)Bc * 4, cudaMemcpyDeviceToHost, s_out);
        if (iters) HOSTCP(iters + i0, d_it + i0, (size_t)Bc * 4, cudaMemcpyDeviceToHost, s_out);
        if (slack_used) HOSTCP(slack_used + i0, d_sl + i0, (size_t)Bc * 4, cudaMemcpyDeviceToHost, s_out);
        if (step_count) HOSTCP(step_count + i0, d_sc + i0, (size_t)Bc * 4, cudaMemcpyDeviceToHost, s_out);
    }
#undef HOSTCP
    if (rc == 0 && ce == cudaSuccess) ce = cudaEventRecord(job.done, c->out_stream);
    if (rc || ce != cudaSuccess) {                    // nothing may still be in flight when an error is reported
        const std::string msg = b2c_last_error();
        pipeline_drain(c);
        if (rc) { b2c_set_error(msg); return rc; }
        return b2c_cuda_fail(ce, "host pipeline", __FILE__, __LINE__);
    }
    if (async) { job.pending = true; return 0; }
    B2C_CUDA(cudaEventSynchronize(job.done));
    return 0;
}

int b2c_mpc_host(b2c_ctx* c, int64_t B, const b2c_mpc_params* p,
                 const double* x0, const double* xref, const double* uref, const double* obs,
                 int32_t* step_count,
                 double* u0, double* cost, int32_t* status, int32_t* iters, int32_t* slack_used,
                 double* useq, double* xpred) {
    return mpc_host_impl(c, B, p, false, nullptr, 0, nullptr, x0, xref, uref, obs, step_count, u0, cost, status, iters,
                         slack_used, useq, xpred);
}
int b2c_mpc_host_async(b2c_ctx* c, int64_t B, const b2c_mpc_params* p,
                       const double* x0, const double* xref, const double* uref, const double* obs,
                       int32_t* step_count,
                       double* u0, double* cost, int32_t* status, int32_t* iters, int32_t* slack_used,
                       double* useq, double* xpred) {
    return mpc_host_impl(c, B, p, true, nullptr, 0, nullptr, x0, xref, uref, obs, step_count, u0, cost, status, iters,
                         slack_used, useq, xpred);
}
int b2c_mpc_table_host(b2c_ctx* c, int64_t B, const b2c_mpc_params* p,
                       const double* table, int32_t table_rows, const int32_t* k0,
                       const double* x0, const double* obs, int32_t* step_count,
                       double* u0, double* cost, int32_t* status, int32_t* iters, int32_t* slack_used,
                       double* useq, double* xpred) {
    if (!table) { b2c_set_error("null table"); return -1; }
    return mpc_host_impl(c, B, p, false, table, table_rows, k0, x0, nullptr, nullptr, obs, step_count, u0, cost, status,
                         iters, slack_used, useq, xpred);
}
int b2c_mpc_table_host_async(b2c_ctx* c, int64_t B, const b2c_mpc_params* p,
                             const double* table, int32_t table_rows, const int32_t* k0,
                             const double* x0, const double* obs, int32_t* step_count,
                             double* u0, double* cost, int32_t* status, int32_t* iters, int32_t* slack_used,
                             double* useq, double* xpred) {
    if (!table) { b2c_set_error("null table"); return -1; }
    return mpc_host_impl(c, B, p, true, table, table_rows, k0, x0, nullptr, nullptr, obs, step_count, u0, cost, status,
                         iters, slack_used, useq, xpred);
}

// ---- FMA peak microbenchmark (FP64 / FP32 roofline denominators) ------------------------------
template <typename T>
__global__ void __launch_bounds__(256)
k_fma_peak(T* out, int iters, T a, T b) {
    T x0 = (T)threadIdx.x, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            x0 = x0 * a + b; x1 = x1 * a + b; x2 = x2 * a + b; x3 = x3 * a + b;
            x4 = x4 * a + b; x5 = x5 * a + b; x6 = x6 * a + b; x7 = x7 * a + b;
        }
    }
    T s = x0 + x1 + x2 + x3 + x4 + x5 + x6 + x7;
    if (s == (T)123456789) out[0] = s;     // never true; keeps the chain alive
}

int b2c_measure_fma_peak(b2c_ctx* c, int bits, double* tflops) {
    if (!c || !tflops) { b2c_set_error("null argument"); return -1; }
    B2C_DEVICE_GUARD(c);
    int rc = b2c_reserve(&c->dstage, &c->dstage_bytes, 4096, false);
    if (rc) return rc;
    const int iters = 4096, blocks = c->sm_count * 8, threads = 256;
    cudaEvent_t e0, e1;
    B2C_CUDA(cudaEventCreate(&e0)); B2C_CUDA(cudaEventCreate(&e1));
    float best = 1e30f;
    for (int rep = 0; rep < 5; ++rep) {
        B2C_CUDA(cudaEventRecord(e0, c->stream));
        if (bits == 64) k_fma_peak<double><<<blocks, threads, 0, c->stream>>>((double*)c->dstage, iters, 0.999999, 1e-6);
        else            k_fma_peak<float><<<blocks, threads, 0, c->stream>>>((float*)c->dstage, iters, 0.9999f, 1e-4f);
        B2C_CHECK_LAUNCH(c);
        B2C_CUDA(cudaEventRecord(e1, c->stream));
        B2C_CUDA(cudaEventSynchronize(e1));
        float ms = 0; B2C_CUDA(cudaEventElapsedTime(&ms, e0, e1));
        if (rep > 0 && ms < best) best = ms;
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    double flops = 2.0 * 64.0 * (double)iters * (double)blocks * (double)threads;
    *tflops = flops / (best * 1e-3) / 1e12;
    return 0;
}
