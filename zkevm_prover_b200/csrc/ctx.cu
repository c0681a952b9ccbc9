// ctx.cu — per-device contexts of libzkgpu.so: creation, the entry guard (Enter), scratch / stream / event helpers.
#include "gl64.cuh"
#include "zk_internal.h"
#include <cstdarg>
#include <cstdio>
#include <cstring>

namespace zk {

std::atomic<unsigned long long> g_launches{0};
static thread_local char g_err[512] = "";
const char *last_error() { return g_err; }
void set_error(const char *fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

static const int MAX_DEV = 64;
static Ctx *g_ctxs[MAX_DEV] = {};
static std::mutex g_ctxs_mu;
static Ctx *g_default = nullptr;          // context of threads that never called zkgpu_init
static thread_local Ctx *t_bound = nullptr; // context this thread named in zkgpu_init
thread_local Ctx *t_ctx = nullptr;

static int ctx_create(Ctx &C)
{
    ZK_CUDA(cudaSetDevice(C.device));
    cudaDeviceProp prop;
    ZK_CUDA(cudaGetDeviceProperties(&prop, C.device));
    if (prop.major < 10)
    {
        set_error("zkgpu: device %d is sm_%d%d; this build contains sm_100a code only", C.device, prop.major, prop.minor);
        return -3;
    }
    C.num_sms = prop.multiProcessorCount;
    ZK_CUDA(cudaStreamCreateWithFlags(&C.own, cudaStreamNonBlocking));
    C.cur = C.own;
    for (auto &ev : C.ev) ZK_CUDA(cudaEventCreate(&ev));
    C.ready = true;
    return 0;
}

Ctx *ctx_for_device(int device, bool create)
{
    std::lock_guard<std::mutex> lk(g_ctxs_mu);
    if (device < 0 || device >= MAX_DEV) return nullptr;
    if (!g_ctxs[device] && create)
    {
        Ctx *c = new Ctx();
        c->device = device;
        if (ctx_create(*c) < 0)
        {
            delete c;
            return nullptr;
        }
        g_ctxs[device] = c;
        if (!g_default) g_default = c;
    }
    return g_ctxs[device];
}

// device < 0: the thread's bound context, else the process default, else a new context on the current device
static Ctx *ctx_pick(int device)
{
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0)
    {
        cudaGetLastError();
        set_error("zkgpu: no usable CUDA device (%s); this library has no CPU fallback",
                  e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0");
        return nullptr;
    }
    if (device >= 0)
    {
        if (device >= n)
        {
            set_error("zkgpu: device %d does not exist (%d visible)", device, n);
            return nullptr;
        }
        return ctx_for_device(device, true);
    }
    if (t_bound) return t_bound;
    if (g_default) return g_default;
    int cur = 0;
    if (cudaGetDevice(&cur) != cudaSuccess)
    {
        cudaGetLastError();
        cur = 0;
    }
    return ctx_for_device(cur, true);
}

Enter::Enter(int device)
{
    c = ctx_pick(device);
    if (!c)
    {
        rc = -3;
        return;
    }
    lk = std::unique_lock<std::recursive_mutex>(c->mu);
    if (cudaSetDevice(c->device) != cudaSuccess)
    {
        set_error("zkgpu: cudaSetDevice(%d) failed: %s", c->device, cudaGetErrorString(cudaGetLastError()));
        rc = -2;
        return;
    }
    prev = t_ctx;
    t_ctx = c;
    if (device >= 0) t_bound = c;
    rc = poseidon_init_constants();
}
Enter::~Enter()
{
    if (c && rc >= 0) t_ctx = prev;
}

int ctx_ensure_scratch(Ctx &C, size_t words)
{
    if (words <= C.scratch_words) return 0;
    if (C.scratch)
    {
        ZK_CUDA(cudaStreamSynchronize(C.cur));
        ZK_CUDA(cudaFree(C.scratch));
        C.scratch = nullptr;
        C.scratch_words = 0;
    }
    cudaError_t e = cudaMalloc(&C.scratch, words * sizeof(u64));
    if (e != cudaSuccess)
    {
        cudaGetLastError();
        set_error("zkgpu: cannot allocate %llu MiB of device scratch (%s)", (unsigned long long)(words >> 17), cudaGetErrorString(e));
        return -4;
    }
    C.scratch_words = words;
    return 0;
}

int ctx_pipe_streams(Ctx &C)
{
    if (C.h2d) return 0;
    ZK_CUDA(cudaStreamCreateWithFlags(&C.h2d, cudaStreamNonBlocking));
    ZK_CUDA(cudaStreamCreateWithFlags(&C.d2h, cudaStreamNonBlocking));
    for (int b = 0; b < 2; b++)
    {
        ZK_CUDA(cudaEventCreateWithFlags(&C.in_ready[b], cudaEventDisableTiming));
        ZK_CUDA(cudaEventCreateWithFlags(&C.in_free[b], cudaEventDisableTiming));
        ZK_CUDA(cudaEventCreateWithFlags(&C.out_ready[b], cudaEventDisableTiming));
        ZK_CUDA(cudaEventCreateWithFlags(&C.out_free[b], cudaEventDisableTiming));
    }
    return 0;
}

int ctx_pipe_event(Ctx &C, int *idx, cudaStream_t st)
{
    if (idx == nullptr)
    {
        C.pool_used = 0;
        return 0;
    }
    if (C.pool_used == C.pool.size())
    {
        cudaEvent_t e;
        ZK_CUDA(cudaEventCreate(&e));
        C.pool.push_back(e);
    }
    *idx = (int)C.pool_used++;
    ZK_CUDA(cudaEventRecord(C.pool[*idx], st));
    return 0;
}

bool is_device_ptr(const void *p) { return ptr_device(p) >= 0; }
int ptr_device(const void *p)
{
    if (!p) return -1;
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess)
    {
        cudaGetLastError();
        return -1;
    }
    if (a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged) return a.device;
    return -1;
}

} // namespace zk
