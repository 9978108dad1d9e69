// Library-level state of libbmhrl_b200: last-error string, kernel-launch counter, version / device queries.
#include "common.cuh"
#include "../../include/bmhrl_b200.h"
#include <atomic>
#include <stdlib.h>
#include <string.h>

static thread_local char g_err[512] = "";
static std::atomic<long long> g_launches{0};

extern "C" void bmhrl_set_error(const char* msg) {
    strncpy(g_err, msg ? msg : "", sizeof(g_err) - 1);
    g_err[sizeof(g_err) - 1] = 0;
}
extern "C" void bmhrl_count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }
extern "C" const char* bmhrl_last_error(void) { return g_err; }
extern "C" long long bmhrl_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }
extern "C" void bmhrl_reset_launch_count(void) { g_launches.store(0, std::memory_order_relaxed); }
extern "C" int bmhrl_abi_version(void) { return BMHRL_ABI_VERSION; }

// Programmatic dependent launch of the chain kernels (common.cuh): -1 = not decided yet (BMHRL_PDL, default on)
static std::atomic<int> g_pdl{-1};
namespace bmhrl {
int pdl_enabled() {
    int v = g_pdl.load(std::memory_order_relaxed);
    if (v < 0) {
        const char* e = getenv("BMHRL_PDL");
        v = e ? (atoi(e) != 0) : 1;
        g_pdl.store(v, std::memory_order_relaxed);
    }
    return v;
}
}  // namespace bmhrl
extern "C" int bmhrl_set_pdl(int on) {
    g_pdl.store(on ? 1 : 0, std::memory_order_relaxed);
    return BMHRL_OK;
}
extern "C" int bmhrl_get_pdl(void) { return bmhrl::pdl_enabled(); }

static std::atomic<int> g_launch_priority{bmhrl::kNoPriority};
namespace bmhrl {
int launch_priority() { return g_launch_priority.load(std::memory_order_relaxed); }
}  // namespace bmhrl
extern "C" int bmhrl_set_launch_priority(int use, int priority) {
    g_launch_priority.store(use ? priority : bmhrl::kNoPriority, std::memory_order_relaxed);
    return BMHRL_OK;
}

// A new non-blocking stream on the current device with a handle no other stream shares (see the header).  Never
// destroyed: the host layer keeps its side streams for the life of the process.
extern "C" int bmhrl_stream_create(void** stream) {
    if (!stream) {
        bmhrl_set_error("bmhrl_stream_create: null pointer");
        return BMHRL_ERR_ARG;
    }
    cudaStream_t s = nullptr;
    cudaError_t e = cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking);
    if (e != cudaSuccess) {
        bmhrl_set_error(cudaGetErrorString(e));
        return BMHRL_ERR_CUDA;
    }
    *stream = (void*)s;
    return BMHRL_OK;
}

// Returns 0 when the current device is a Blackwell sm_100 part this library was built for.
extern "C" int bmhrl_check_device(void) {
    int dev = 0;
    cudaDeviceProp prop;
    if (cudaGetDevice(&dev) != cudaSuccess || cudaGetDeviceProperties(&prop, dev) != cudaSuccess) {
        bmhrl_set_error("no CUDA device");
        return BMHRL_ERR_CUDA;
    }
    if (prop.major != 10) {
        bmhrl_set_error("libbmhrl_b200 is built for sm_100a (B200) only");
        return BMHRL_ERR_UNSUPPORTED;
    }
    return BMHRL_OK;
}
